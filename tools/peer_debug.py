"""2-rank probe of the peer exchange: several solves in a row on one exchange, against the unsharded result."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
abi = importlib.import_module("monst3r-slam_b200.abi")
synth = importlib.import_module("monst3r-slam_b200.synth")
mdist = importlib.import_module("monst3r-slam_b200.dist")
for (N, E, n) in ((40, 90, 3000), (24, 80, 3000)):
    g = synth.make_graph_exact(N, E, n, seed=40, noise=1e-3, device=dev)
    order = mdist.pair_aligned_order(g["ii"], g["jj"])
    ii, jj = g["ii"][order].contiguous(), g["jj"][order].contiguous()
    idx, valid, Q = g["idx"][order].contiguous(), g["valid"][order].contiguous(), g["Q"][order].contiguous()
    E2 = ii.numel()
    prm = abi.make_params(abi.GN_RAYS, 0.003, 10.0, 0.0, 1.5, 10, 0.0)
    T_ref = g["Twc"].clone()
    abi.gauss_newton(prm, T_ref, g["Xs"], g["Cs"], None, ii, jj, idx, valid, Q)
    torch.cuda.synchronize()
    parts = mdist.partition_edges(E2, world)
    lo, hi = parts[rank]
    sl = (idx[lo:hi].contiguous(), valid[lo:hi].contiguous(), Q[lo:hi].contiguous())
    ws, _ = abi.gn_workspace(N, E2, n, dev, edges_local=hi - lo)
    xchg = mdist.init_peer_exchange(rank, world, E2)
    sh = mdist.make_shard(lo, hi, None, abi.EXCHANGE_PEER, xchg)
    res = []
    for k in range(5):
        T = g["Twc"].clone()
        dx, info = abi.gauss_newton(prm, T, g["Xs"], g["Cs"], None, ii, jj, *sl, shard=sh, workspace=ws)
        if k == 2:
            torch.cuda.synchronize(); dist.barrier()
        res.append(T)
    torch.cuda.synchronize()
    print(f"rank {rank} N={N}: max|T - T_unsharded| per solve:", [float((t - T_ref).abs().max()) for t in res], flush=True)
    mdist.destroy_peer_exchange(xchg)
dist.destroy_process_group()
