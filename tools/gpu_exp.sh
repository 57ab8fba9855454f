#!/bin/bash
timeout 300 python tools/profile_plan.py enb4-hu 32 --top 45 > gpurun_out/exp_plan_enb4.log 2>&1
timeout 300 python tools/profile_plan.py enb4-lr 8 --top 45 > gpurun_out/exp_plan_enb4lr.log 2>&1
