#!/bin/bash
# Evidence run at N = $N GPUs (2 or 4): multi-GPU tests at that world size (incl. the late / leaving rank test), GPT-2 with the
# same-box DDP arm, ResNet-50 sync + async, ResNet-20 sync.  One JSON per run under gpurun_out/.
N=${N:-4}
mkdir -p gpurun_out
PORT=29700
run() { name=$1; shift; PORT=$((PORT+1))
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" \
      > gpurun_out/s${N}_$name.json 2> gpurun_out/s${N}_$name.err
  echo "$name exit $?"; }
DFB_TEST_WORLD=$N timeout 900 python -m pytest tests/test_multigpu.py -q -rP > gpurun_out/s${N}_tests.log 2>&1
echo "multigpu tests (world $N) exit $?"; grep -a "distance to the NCCL" gpurun_out/s${N}_tests.log; tail -n 3 gpurun_out/s${N}_tests.log
run gpt2_default --steps 20 --warmup 5
run resnet50_sync --model resnet50 --steps 20 --warmup 5 --optimizer momentum
run resnet50_async --model resnet50 --mode async --steps 20 --warmup 5 --optimizer momentum --baseline none
run resnet20_sync --model resnet20 --steps 30 --warmup 5 --optimizer momentum
python - <<PY
import json,glob
for f in sorted(glob.glob('gpurun_out/s${N}_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], 'ms', round(d['ms_per_step'],3), 'val', round(d['value'],1), 'vs', d.get('vs_baseline') and round(d['vs_baseline'],3), 'ident', d.get('replicas_identical'), 'mhz', (d.get('clocks') or {}).get('sm_mhz'), 'async', d.get('async'))
    except Exception as e: print(f, 'ERR', e)
PY
tail -n 2 gpurun_out/s${N}_*.err 2>/dev/null | tail -n 12
