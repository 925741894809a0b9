#!/bin/bash
# final round-2 evidence: full GPU suite, default bench, launch list + --set full of the same commands
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu > gpurun_out/r02t_tests_gpu.log 2>&1; echo "suite rc=$?"; tail -2 gpurun_out/r02t_tests_gpu.log
python bench.py > gpurun_out/r02t_bench_n1.json 2> gpurun_out/r02t_bench_n1.err; echo "bench rc=$?"
B="python bench.py --steps 2 --warmup 1 --global-opt none --no-cpu-baseline --no-sequence"
$B > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"iter_proj|refine_|gn_|match_prep|pixel_to_lin|f32_to_f16|track_" -c 400 --csv --log-file gpurun_out/r02t_ncu_launches_raw.csv $B > gpurun_out/r02t_ncu_bench.log 2>&1
echo "launch list rc=$?"
