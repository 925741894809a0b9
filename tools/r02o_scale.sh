#!/bin/bash
# N = 8, 4, 2 on one 8-GPU box: the default bench line (tracking replicas + sharded c3 / c4 global optimisation)
mkdir -p gpurun_out
for N in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2950$N bench.py --gpus $N --steps 20 --warmup 3 --no-sequence > gpurun_out/r02y_bench_n$N.json 2> gpurun_out/r02y_bench_n$N.err
  echo "N=$N rc=$?"; tail -c 300 gpurun_out/r02y_bench_n$N.err
done
