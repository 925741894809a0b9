#!/bin/bash
# compute-sanitizer over the flag / LL / mbarrier kernels (one tool per gpurun call: B200_PROFILING.md).  usage: r02g_sanitize.sh racecheck|memcheck|synccheck
tool=${1:-racecheck}
mkdir -p gpurun_out
sel='test_full_loop_vs_oracle or test_cooperative_solver_sizes or test_quirks or test_bitwise_determinism or test_rays_default_thresholds_vs_oracle or test_calib_vs_oracle or test_refine_staged_path_bit_exact or test_iter_proj_padded'
timeout 2400 compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 66 python -m pytest tests/test_gpu_gn.py tests/test_gpu_tracker.py tests/test_gpu_matching.py -m gpu -x -q -k "$sel" > gpurun_out/r02g_sanitizer_$tool.log 2>&1
echo "sanitizer $tool rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|hazard|Invalid|Barrier error" gpurun_out/r02g_sanitizer_$tool.log | sort | uniq -c | head -20
tail -3 gpurun_out/r02g_sanitizer_$tool.log
