#!/bin/bash
N=${N:-8}; PORT=29800; mkdir -p gpurun_out
run() { name=$1; shift; PORT=$((PORT+1)); timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > gpurun_out/m${N}_$name.json 2> gpurun_out/m${N}_$name.err; echo "$name exit $?"; }
run gpt2_auto --steps 20 --warmup 5
run gpt2_auto_nvls --steps 20 --warmup 5 --nvls on --baseline none --no-e2e
run gpt2_auto_nofuseag --steps 20 --warmup 5 --no-fuse-allgather --baseline none --no-e2e
run gpt2_auto_nvls_nofuseag --steps 20 --warmup 5 --nvls on --no-fuse-allgather --baseline none --no-e2e
python - <<PY
import json,glob
for f in sorted(glob.glob('gpurun_out/m${N}_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('/')[-1], 'ms', round(d['ms_per_step'],3), 'val', round(d['value'],1), 'vs', d.get('vs_baseline') and round(d['vs_baseline'],3), 'ident', d.get('replicas_identical'), 'mhz', (d.get('clocks') or {}).get('sm_mhz'), 'buckets', d['config']['buckets'])
    except Exception as e: print(f, 'ERR', e)
PY
