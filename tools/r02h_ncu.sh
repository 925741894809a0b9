#!/bin/bash
# round-2 ncu evidence: (1) launch list of the tracking-step bench command, (2) --set full of the hot kernels
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --global-opt none --no-cpu-baseline --no-sequence"
$B > gpurun_out/r02h_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02h_ncu_launches_raw.csv $B > gpurun_out/r02h_ncu_bench.log 2>&1
echo "launch list rc=$?"
P="python tools/prof_step.py --reps 1 --iters 2 --graph 32,120"
$P > gpurun_out/r02h_plain_prof.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"refine_staged|iter_proj|gn_track|gn_accum|gn_solve|gn_pack|gn_prepare" -c 14 -o gpurun_out/r02h_prof -f $P > gpurun_out/r02h_ncu_prof.log 2>&1
echo "full rc=$?"; ls -la gpurun_out/r02h_prof.ncu-rep; tail -3 gpurun_out/r02h_ncu_prof.log
