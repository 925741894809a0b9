#!/bin/bash
timeout 120 python tools/ab_solve.py 256,1000 64,300 40,160 2>&1 | grep "GN x10"
M3R_SOLVE_NOV3=1 timeout 120 python tools/ab_solve.py 256,1000 64,300 40,160 2>&1 | grep "GN x10"
timeout 600 python -m pytest tests/test_gpu_gn_configs.py tests/test_gpu_gn.py -x -q -m gpu 2>&1 | tail -3
