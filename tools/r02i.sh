#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02i_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/r02i_noalloc_alone.log
python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_multi.py > gpurun_out/r02i_tests_gpu.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/r02i_tests_gpu.log
