#!/bin/bash
# round-2 trip C (1 GPU): comm + attention tests (barrier time-out), attention microbench, ncu of the final attention kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_comm.py tests/test_gpu_attention.py -q > gpurun_out/c_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/c_tests.log
timeout 300 python scripts/bench_attention.py > gpurun_out/c_attn_bench.json 2> gpurun_out/c_attn_bench.err
python scripts/profile_attention.py > gpurun_out/c_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:attn_ -s 4 -c 4 -o gpurun_out/attn_final python scripts/profile_attention.py > gpurun_out/c_ncu.log 2>&1
tail -n 4 gpurun_out/c_tests.log; cat gpurun_out/c_attn_bench.json; tail -n 2 gpurun_out/c_ncu.log
