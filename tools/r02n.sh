#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/ab_solve.py 32,120 2>&1 | tail -1
timeout 600 python -m pytest tests/test_gpu_gn_configs.py tests/test_gpu_gn.py -x -q -m gpu 2>&1 | tail -3
