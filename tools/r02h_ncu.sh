#!/bin/bash
# round-2 ncu evidence: launch list of the tracking-step bench command (this library's kernels only)
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --global-opt none --no-cpu-baseline --no-sequence"
$B > gpurun_out/r02h_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"iter_proj|refine_|gn_|match_prep|pixel_to_lin|f32_to_f16|track_" -c 400 --csv --log-file gpurun_out/r02h_ncu_launches_raw.csv $B > gpurun_out/r02h_ncu_bench.log 2>&1
echo "launch list rc=$?"; grep -c m3r gpurun_out/r02h_ncu_launches_raw.csv
