#!/bin/bash
# attention iteration: numerics, microbench, ncu of one forward + one backward launch
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attention.py -q > gpurun_out/i_attn_test.log 2>&1
echo "attention tests exit $?" >> gpurun_out/i_attn_test.log
timeout 300 python scripts/bench_attention.py > gpurun_out/i_attn_bench.json 2> gpurun_out/i_attn_bench.err
python scripts/profile_attention.py > gpurun_out/i_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:attn_[fb]wd -s 2 -c 2 -o gpurun_out/attn_iter python scripts/profile_attention.py > gpurun_out/i_ncu.log 2>&1
tail -n 4 gpurun_out/i_attn_test.log; cat gpurun_out/i_attn_bench.json; tail -n 2 gpurun_out/i_ncu.log
