#!/bin/bash
# 2-GPU: hardware multi-GPU correctness test + bench at N=2 (c3 + c4 legs)
set -o pipefail
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -s > gpurun_out/r02e_tests_multi.log 2>&1; echo "multi test rc=$?"; tail -8 gpurun_out/r02e_tests_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02e_bench_n2.json 2> gpurun_out/r02e_bench_n2.err; echo "bench n2 rc=$?"; tail -3 gpurun_out/r02e_bench_n2.err
python - <<'PY'
import json
try:
    d=json.load(open("gpurun_out/r02e_bench_n2.json"))
    print("value",d["value"],"e2e",d["e2e"]["value"])
    for k in ("global_opt","global_opt_c4"):
        g=d[k]; print(k, {x:g[x] for x in ("ms_per_solve","gn_edge_iters_per_s","accumulate_ms_per_iter","solve_alone_ms","solve_and_comm_ms_per_iter","poses_bit_identical_to_unsharded","poses_identical_on_all_ranks","collective")})
except Exception as e: print("bench parse failed", e)
PY
