#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "dwconv" 2>&1 | tail -1
timeout 600 python -m pytest tests/test_model_parity_gpu.py -m gpu -q -x 2>&1 | tail -1
timeout 300 python tools/profile_plan.py enb0-hu 64 --top 140 > gpurun_out/prof_enb0_full.log 2>&1; head -2 gpurun_out/prof_enb0_full.log; grep " dw " gpurun_out/prof_enb0_full.log | head -8
EDE_DW_NC=2 timeout 300 python tools/profile_plan.py enb0-hu 64 --top 140 > gpurun_out/prof_enb0_nc2.log 2>&1; head -2 gpurun_out/prof_enb0_nc2.log; grep " dw " gpurun_out/prof_enb0_nc2.log | head -8
