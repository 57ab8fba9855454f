#!/bin/bash
timeout 900 python -m pytest tests/test_multigpu_nccl_gpu.py tests/test_dropin_gpu.py -m gpu -q -x 2>&1 | tail -4
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_model_parity_gpu.py -m gpu -q -x -k "resize or parity or model" 2>&1 | tail -2
timeout 300 python tools/profile_plan.py enb0-hu 64 --top 140 > gpurun_out/exp_plan.log 2>&1; head -2 gpurun_out/exp_plan.log | cut -c1-200; grep " resize " gpurun_out/exp_plan.log | head -8 | cut -c1-150
timeout 300 python tools/profile_plan.py rn50-hu 64 --top 140 > gpurun_out/exp_plan.log 2>&1; head -2 gpurun_out/exp_plan.log | cut -c1-200; grep " resize " gpurun_out/exp_plan.log | head -8 | cut -c1-150
