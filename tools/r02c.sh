#!/bin/bash
set -o pipefail
mkdir -p gpurun_out
M3R_LIB_DIR=lib_prof timeout 120 python tools/solve_prof2.py 32,120 256,1000 2>&1 | grep "DFPROF\|DFRAW\|DFHELP\|rror" | tail -18
timeout 120 python tools/ab_solve.py 12,20 32,120 64,160 256,1000 2>&1 | tail -4; echo "ab rc=$?"
timeout 600 python -m pytest tests/test_gpu_gn.py tests/test_gpu_gn_configs.py -m gpu -x -q -s > gpurun_out/r02c_tests.log 2>&1; echo "tests rc=$?"; grep -E "vs|passed|failed|Error|error" gpurun_out/r02c_tests.log | tail -12
