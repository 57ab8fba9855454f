#!/bin/bash
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "upconv" 2>&1 | tail -3
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
