#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r02r_tests_gpu.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r02r_tests_gpu.log
python bench.py --global-opt none --no-sequence --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], [ (k[:30],round(v['ms'],4)) for k,v in d['kernels'].items()])"
