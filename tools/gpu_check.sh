#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8
timeout 300 python tools/profile_plan.py enb0-hu 64 --top 24 > gpurun_out/prof_enb0_64.log 2>&1; head -30 gpurun_out/prof_enb0_64.log
timeout 300 python tools/profile_plan.py rn50-hu 32 --top 12 > gpurun_out/prof_rn50_32.log 2>&1; head -18 gpurun_out/prof_rn50_32.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_enb0.log 2>&1; tail -1 gpurun_out/bench_enb0.log
