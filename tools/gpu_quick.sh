#!/bin/bash
# quick GPU loop: matching parity tests, short bench (tracking step only), ncu bank-conflict counters of refine
set -o pipefail
timeout 600 python -m pytest tests/test_gpu_matching.py -m gpu -x -q > gpurun_out/t.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/t.log
timeout 300 python bench.py --steps 10 --warmup 3 --global-opt none --no-cpu-baseline > gpurun_out/b.json 2> gpurun_out/b.err; echo "bench rc=$?"; tail -3 gpurun_out/b.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/b.json"))
print("ms_per_step",d["ms_per_step"],"e2e ms",d["e2e"]["ms_per_step"])
for k,v in d["kernels"].items(): print("  %-90s %.4f ms"%(k,v["ms"]))
PY
if [ -n "$NCU_K" ]; then
python tools/prof_step.py --reps 1 --iters 1 > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"$NCU_K" -c 2 python tools/prof_step.py --reps 1 --iters 1 2>&1 | grep -E "^\s+(gpu__|l1tex|smsp)|^  [a-z_:<>A-Za-z0-9, ]+\(" 
fi
