#!/bin/bash
# round-2 trip D (1 GPU): dgelu+colsum epilogue test, LN tests, A/B of the fused fc1 bias gradient in the whole step
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -k "dgelu or layer_norm or mlp" > gpurun_out/d_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/d_tests.log
tail -n 3 gpurun_out/d_tests.log
DFB_FUSE_DGELU_COLSUM=0 timeout 400 python bench.py --steps 20 --warmup 5 --baseline none > gpurun_out/d_bench_sep.json 2> gpurun_out/d_bench_sep.err
timeout 400 python bench.py --steps 20 --warmup 5 --baseline none > gpurun_out/d_bench_fused.json 2> gpurun_out/d_bench_fused.err
DFB_FUSE_DGELU_COLSUM=0 timeout 400 python bench.py --steps 20 --warmup 5 --baseline none > gpurun_out/d_bench_sep2.json 2> gpurun_out/d_bench_sep2.err
timeout 400 python bench.py --steps 20 --warmup 5 --baseline none > gpurun_out/d_bench_fused2.json 2> gpurun_out/d_bench_fused2.err
for f in sep fused sep2 fused2; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(f"gpurun_out/d_bench_{sys.argv[1]}.json").read().strip().splitlines()[-1])
    print(sys.argv[1], d["ms_per_step"], d["value"], d["clocks"].get("sm_mhz"), d["clocks"].get("reasons"), d.get("gpu_launches"))
except Exception as e:
    print(sys.argv[1], "ERR", e)
PY
done
