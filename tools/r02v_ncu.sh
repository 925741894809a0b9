#!/bin/bash
mkdir -p gpurun_out
P="python tools/prof_step.py --reps 1 --iters 2 --graph 256,1000 --kind calib"
$P > gpurun_out/r02z_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"gn_solve_kernel|gn_accum_kernel|gn_pack_kernel" -c 5 -o gpurun_out/r02z_c4 -f $P > gpurun_out/r02z_ncu.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/r02z_c4.ncu-rep
