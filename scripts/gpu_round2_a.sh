#!/bin/bash
# first GPU trip of round 2: attention numerics, full GPU suite, attention microbench, bench line with the in-line DDP arm
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/a_gpu.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_attention.py -q > gpurun_out/a_attn_test.log 2>&1
ATT=$?
echo "attention tests exit $ATT" >> gpurun_out/a_attn_test.log
if [ $ATT -ne 0 ]; then export DFB_ATTN=sdpa; fi
timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_attention.py > gpurun_out/a_gpu_tests.log 2>&1
echo "gpu tests exit $?" >> gpurun_out/a_gpu_tests.log
if [ $ATT -eq 0 ]; then timeout 300 python scripts/bench_attention.py > gpurun_out/a_attn_bench.json 2> gpurun_out/a_attn_bench.err; fi
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err
echo "bench exit $?" >> gpurun_out/a_bench.err
DFB_ATTN=sdpa timeout 600 python bench.py --steps 10 --warmup 3 --baseline none > gpurun_out/a_bench_sdpa.json 2> gpurun_out/a_bench_sdpa.err
tail -n 3 gpurun_out/a_attn_test.log; tail -n 3 gpurun_out/a_gpu_tests.log; cat gpurun_out/a_attn_bench.json gpurun_out/a_bench.json gpurun_out/a_bench_sdpa.json
