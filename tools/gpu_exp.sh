#!/bin/bash
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "conv2d" 2>&1 | tail -2
timeout 900 python -m pytest tests/test_model_parity_gpu.py -m gpu -q -x 2>&1 | tail -2
for v in 1 0; do export EDE_UMMA_SHARE_PASSES=$v; echo "== SHARE_PASSES=$v"
timeout 300 python tools/profile_plan.py enb0-hu 64 --top 140 > gpurun_out/exp_plan_s$v.log 2>&1; head -2 gpurun_out/exp_plan_s$v.log | cut -c1-200
timeout 300 python tools/profile_plan.py rn50-hu 64 --top 140 > gpurun_out/exp_plan_rs$v.log 2>&1; head -2 gpurun_out/exp_plan_rs$v.log | cut -c1-200
done
