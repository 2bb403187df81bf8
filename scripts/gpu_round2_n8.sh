#!/bin/bash
# Whole-box evidence run (N = $N GPUs, default 8): multi-GPU tests, GPT-2 bench + A/Bs + bucket sweep, gradient-path sweep,
# ResNet-50 / ResNet-20 sync + async, MNIST-CNN async through the DistriServer facade.  One JSON per run under gpurun_out/.
N=${N:-8}
mkdir -p gpurun_out
PORT=29600
run() { # name, then bench.py args
  name=$1; shift; PORT=$((PORT+1))
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" \
      > gpurun_out/n${N}_$name.json 2> gpurun_out/n${N}_$name.err
  echo "$name exit $? $(tail -c 300 gpurun_out/n${N}_$name.json | head -c 0)"
}
if [ "${TESTS:-1}" = "1" ]; then
  DFB_TEST_WORLD=$N timeout 900 python -m pytest tests/test_multigpu.py -q > gpurun_out/n${N}_tests.log 2>&1
  echo "multigpu tests (world $N) exit $?"; tail -n 4 gpurun_out/n${N}_tests.log
fi
run gpt2_default --steps 20 --warmup 5
run gpt2_trace --steps 10 --warmup 3 --trace --baseline none --no-e2e
run gpt2_nvls --steps 20 --warmup 5 --nvls on --baseline none --no-e2e
run gpt2_bucketclose --steps 20 --warmup 5 --close-barrier bucket --baseline none --no-e2e
run gpt2_nofuseag --steps 20 --warmup 5 --no-fuse-allgather --baseline none --no-e2e
for mb in 4 16 64; do run gpt2_bucket${mb} --steps 15 --warmup 3 --bucket-mb $mb --baseline none --no-e2e; done
run gpt2_bucket256 --steps 15 --warmup 3 --bucket-mb 256 --no-e2e
PORT=$((PORT+1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT scripts/bench_comm.py --opt adam --nvls --ctas 148 --out gpurun_out/n${N}_comm_sweep.jsonl > gpurun_out/n${N}_comm_sweep.log 2>&1
echo "comm sweep exit $?"
run resnet50_sync --model resnet50 --steps 20 --warmup 5 --optimizer momentum
run resnet50_async --model resnet50 --mode async --steps 20 --warmup 5 --optimizer momentum --baseline none
run resnet20_sync --model resnet20 --steps 30 --warmup 5 --optimizer momentum
PORT=$((PORT+1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT -m distriflow_b200.experiment.mnist.mnist_gpu --model cnn --mode async --max-staleness 4 --examples 16384 --out gpurun_out/n${N}_mnist_cnn_async.json > /dev/null 2> gpurun_out/n${N}_mnist_cnn_async.err
echo "mnist async exit $?"
python - <<PY
import json,glob
for f in sorted(glob.glob('gpurun_out/n${N}_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        if 'ms_per_step' in d:
            print(f.split('/')[-1], 'ms', round(d['ms_per_step'],3), 'val', round(d['value'],1), 'vs', d.get('vs_baseline') and round(d['vs_baseline'],3), 'ident', d.get('replicas_identical'), 'mhz', (d.get('clocks') or {}).get('sm_mhz'), 'trace', d.get('trace'), 'async', d.get('async'))
        else:
            print(f.split('/')[-1], {k:d.get(k) for k in ('samples_per_sec','samples_per_sec_applied','counters','loss_first','loss_last')})
    except Exception as e: print(f, 'ERR', e)
PY
cat gpurun_out/n${N}_comm_sweep.jsonl 2>/dev/null | head -12
grep -l -i "error\|Traceback" gpurun_out/n${N}_*.err 2>/dev/null | head
