#!/bin/bash
# 2-GPU trip: multi-GPU tests, bench N=2 (default, bucket-close, nvls), MNIST async through the facade
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_multigpu.py -x -q > gpurun_out/n2_tests.log 2>&1
echo "multigpu tests exit $?" >> gpurun_out/n2_tests.log
tail -n 6 gpurun_out/n2_tests.log
run() { name=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 5 "$@" > gpurun_out/n2_bench_$name.json 2> gpurun_out/n2_bench_$name.err; echo "$name exit $?"; }
run default
run nvls --baseline none --nvls
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 -m distriflow_b200.experiment.mnist.mnist_gpu --model cnn --mode async --max-staleness 4 --examples 8192 --out gpurun_out/n2_mnist_async.json > /dev/null 2> gpurun_out/n2_mnist_async.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/n2_bench_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['ms_per_step'],3), round(d['value'],1), d.get('vs_baseline'), d.get('replicas_identical'), d['clocks']['sm_mhz'])
    except Exception as e: print(f, 'ERR', e)
try:
    d=json.load(open('gpurun_out/n2_mnist_async.json')); print({k:d[k] for k in ('samples_per_sec','samples_per_sec_applied','counters','loss_first','loss_last','workerIDs')})
except Exception as e: print('mnist async ERR', e)
PY
tail -n 3 gpurun_out/n2_bench_*.err gpurun_out/n2_mnist_async.err 2>/dev/null | tail -n 20
