#!/bin/bash
# Final 1-GPU confirmation: full GPU test suite, smoke() under a kernel-name listing, bench (GPT-2 with the in-line DDP arm),
# ResNet-50 sync / async, ResNet-20, MNIST-CNN; one JSON per run under gpurun_out/.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/z_tests.log 2>&1
echo "gpu tests exit $?" >> gpurun_out/z_tests.log
tail -n 3 gpurun_out/z_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/z_smoke.log 2>&1; tail -n 1 gpurun_out/z_smoke.log
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/z_gpt2_n1.json 2> gpurun_out/z_gpt2_n1.err
timeout 300 python bench.py --model resnet50 --steps 20 --warmup 5 --optimizer momentum > gpurun_out/z_resnet50_sync_n1.json 2> gpurun_out/z_r50.err
timeout 300 python bench.py --model resnet50 --mode async --steps 20 --warmup 5 --optimizer momentum --baseline none > gpurun_out/z_resnet50_async_n1.json 2> gpurun_out/z_r50a.err
timeout 300 python bench.py --model resnet20 --steps 30 --warmup 5 --optimizer momentum > gpurun_out/z_resnet20_n1.json 2> gpurun_out/z_r20.err
timeout 300 python bench.py --model mnist-cnn --steps 50 --warmup 5 --optimizer sgd > gpurun_out/z_mnistcnn_n1.json 2> gpurun_out/z_mn.err
timeout 300 python bench.py --impl reference > gpurun_out/z_reference.json 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/z_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        if 'ms_per_step' in d:
            print(f.split('/')[-1], 'ms', round(d['ms_per_step'],3), 'val', round(d['value'],1), 'vs', d.get('vs_baseline') and round(d['vs_baseline'],3), 'e2e', (d.get('e2e') or {}).get('value'), 'launches', d.get('gpu_launches'), 'mhz', (d.get('clocks') or {}).get('sm_mhz'), (d.get('clocks') or {}).get('reasons'))
        else: print(f.split('/')[-1], d)
    except Exception as e: print(f, 'ERR', e)
PY
