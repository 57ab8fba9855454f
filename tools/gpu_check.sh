#!/bin/bash
# first GPU contact: model parity + probe; logs land in gpurun_out/
mkdir -p gpurun_out
python -m pytest tests/test_model_parity_gpu.py -m gpu -x -q -s 2>&1 | tail -60 > gpurun_out/parity.log
cat gpurun_out/parity.log
