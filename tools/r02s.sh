#!/bin/bash
for v in lib_old lib lib_old lib; do M3R_LIB_DIR=$v timeout 300 python tools/ab_solve.py 32,120 64,300 256,1000 2>&1 | grep "GN x10"; done
timeout 600 python -m pytest tests/test_gpu_gn_configs.py tests/test_gpu_gn.py tests/test_gpu_solver_paths.py -x -q -m gpu 2>&1 | tail -3
