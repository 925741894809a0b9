#!/bin/bash
# solver A/B: v1 (grid-step) vs dataflow, then GN parity tests on the dataflow path
set -o pipefail
mkdir -p gpurun_out
timeout 120 python tools/ab_solve.py 32,120 256,1000 2>&1 | tail -2
M3R_SOLVE=df timeout 120 python tools/ab_solve.py 12,20 32,120 64,160 256,1000 2>&1 | tail -4; echo "ab rc=$?"
timeout 600 python -m pytest tests/test_gpu_gn.py tests/test_gpu_gn_configs.py tests/test_gpu_factor_graph.py -m gpu -x -q -s > gpurun_out/r02b_tests.log 2>&1; echo "tests rc=$?"; grep -E "vs|passed|failed|Error|error" gpurun_out/r02b_tests.log | tail -30
