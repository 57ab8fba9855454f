#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_model_parity_gpu.py -m gpu -q -s 2>&1 | grep -E "rel_l1|passed|failed|Error|error" | head -60 > gpurun_out/parity.log
cat gpurun_out/parity.log
python tools/profile_plan.py enb0-hu 64 --top 40 > gpurun_out/prof_enb0_64.log 2>&1; head -50 gpurun_out/prof_enb0_64.log
python tools/profile_plan.py rn50-hu 32 --top 40 > gpurun_out/prof_rn50_32.log 2>&1; head -50 gpurun_out/prof_rn50_32.log
