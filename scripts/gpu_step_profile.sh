#!/bin/bash
# per-kernel share of the GPT-2 step (ncu launch list over 2 eager steps) + the attention dbias test output
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_attention.py -q -k bias 2>&1 | grep -E "^E |assert|passed|failed" | head -12
python bench.py --steps 2 --warmup 3 --no-graph --no-e2e --baseline none > gpurun_out/sp_plain.log 2>&1 && \
DFB_PROFILE_RANGE=1 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-graph --no-e2e --baseline none > gpurun_out/sp_ncu.log 2>&1
python scripts/step_profile.py gpurun_out/launches.csv --steps 2 --top 30 > gpurun_out/step_profile.md; cat gpurun_out/step_profile.md
