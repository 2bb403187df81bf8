#!/bin/bash
N=2; PORT=29700; mkdir -p gpurun_out
run() { name=$1; shift; PORT=$((PORT+1)); timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 20 --warmup 5 --baseline none --no-e2e "$@" > gpurun_out/i2_$name.json 2> gpurun_out/i2_$name.err; }
run default
run bucket256 --bucket-mb 256
run ctas32 --comm-ctas 32
run ctas16 --comm-ctas 16
run default_b
run bucket1024 --bucket-mb 1024
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29790 -m distriflow_b200.experiment.mnist.mnist_gpu --model cnn --mode async --max-staleness 4 --examples 16384 --out gpurun_out/i2_mnist.json > gpurun_out/i2_mnist.out 2> gpurun_out/i2_mnist.err; echo "mnist exit $?"; tail -c 600 gpurun_out/i2_mnist.out
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/i2_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f.split('/')[-1], round(d['ms_per_step'],3), d['clocks']['sm_mhz'], d['config']['buckets'])
    except Exception as e: print(f,'ERR',e)
PY
