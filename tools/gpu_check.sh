#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "conv2d or upconv or dwconv" 2>&1 | tail -3
EDE_DW_NC=1 timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "dwconv" 2>&1 | tail -1
timeout 600 python -m pytest tests/test_model_parity_gpu.py -m gpu -q -x 2>&1 | tail -2
timeout 300 python tools/profile_plan.py enb0-hu 64 --top 30 > gpurun_out/prof_enb0_64.log 2>&1; head -34 gpurun_out/prof_enb0_64.log
timeout 300 python tools/profile_plan.py rn50-hu 32 --top 12 > gpurun_out/prof_rn50_32.log 2>&1; head -6 gpurun_out/prof_rn50_32.log
