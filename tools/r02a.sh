#!/bin/bash
# round-2 call A: new full-size parity tests, whole GPU suite, default bench (c3 + c4 legs), solver microbenchmarks
set -o pipefail
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r02a_smi.txt 2>&1
timeout 1500 python -m pytest tests/test_gpu_gn_configs.py -m gpu -x -q -s > gpurun_out/r02a_tests_configs.log 2>&1; echo "configs tests rc=$?"; tail -15 gpurun_out/r02a_tests_configs.log
timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_gpu_gn_configs.py > gpurun_out/r02a_tests_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -5 gpurun_out/r02a_tests_gpu.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02a_bench_n1.json 2> gpurun_out/r02a_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/r02a_bench_n1.err
python - <<'PY'
import json
try:
    d=json.load(open("gpurun_out/r02a_bench_n1.json"))
    print("ms_per_step",d["ms_per_step"],"e2e ms",d["e2e"]["ms_per_step"])
    for k,v in d["kernels"].items(): print("  %-90s %.4f ms"%(k,v["ms"]))
    for k in ("global_opt","global_opt_c4"):
        g=d[k]; print(k, {x:g[x] for x in ("ms_per_solve","gn_edge_iters_per_s","accumulate_ms_per_iter","pack_ms_per_solve","solve_alone_ms","solve_and_comm_ms_per_iter","accumulate_frac_of_hbm_peak","info","max_abs_pose_err_vs_gt")})
except Exception as e: print("bench parse failed", e)
PY
timeout 120 tools/ubench_solve.bin > gpurun_out/r02a_ubench_solve.log 2>&1; echo "ubench rc=$?"; cat gpurun_out/r02a_ubench_solve.log
