#!/bin/bash
timeout 900 python -m pytest tests/test_model_parity_gpu.py tests/test_config_parity_gpu.py -m gpu -q -x 2>&1 | tail -2
timeout 300 python tools/profile_plan.py enb0-hu 64 --top 140 > gpurun_out/exp_plan.log 2>&1; head -2 gpurun_out/exp_plan.log | cut -c1-200; grep "#  0 \|#  1 " gpurun_out/exp_plan.log | cut -c1-130
timeout 300 python tools/profile_plan.py enb4-hu 32 --top 10 2>&1 | head -1 | cut -c1-120
