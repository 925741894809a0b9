#!/bin/bash
# A/B of the per-iteration exchange of the sharded global optimisation: own peer-memory kernel vs NCCL all-gather
mkdir -p gpurun_out
N=$1
for X in peer nccl; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2960$N bench.py --gpus $N --steps 20 --warmup 3 --no-sequence --exchange $X > gpurun_out/r02peer_n${N}_$X.json 2> gpurun_out/r02peer_n${N}_$X.err
  echo "N=$N $X rc=$?"; tail -c 200 gpurun_out/r02peer_n${N}_$X.err
done
python - <<PY
import json
for X in ("peer","nccl"):
    d=json.loads(open("gpurun_out/r02peer_n${N}_%s.json"%X).read().strip().splitlines()[-1])
    for k in ("global_opt","global_opt_c4"):
        g=d[k]; print(X,k,"%.0f edge-it/s"%g["gn_edge_iters_per_s"],"ms/solve %.3f"%g["ms_per_solve"],"solve+comm %.4f"%g["solve_and_comm_ms_per_iter"],g["poses_bit_identical_to_unsharded"],g["poses_identical_on_all_ranks"])
PY
