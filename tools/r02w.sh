#!/bin/bash
timeout 120 python tools/ab_solve.py 256,1000 128,500 64,300 40,160 2>&1 | grep "GN x10"
M3R_SOLVE_NOV4=1 timeout 120 python tools/ab_solve.py 256,1000 128,500 64,300 40,160 2>&1 | grep "GN x10"
timeout 900 python -m pytest tests/test_gpu_gn_configs.py tests/test_gpu_gn.py tests/test_gpu_solver_paths.py -x -q -m gpu 2>&1 | tail -3
