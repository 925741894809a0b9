"""Worker of tests/test_gpu_multi.py: one process per GPU under torchrun (NCCL).  Shards a 40-keyframe graph by edge
through the drop-in module's additive API exactly as INTEGRATION.md §3 shows, and through the C ABI, and checks on
EVERY rank that the poses are bit-identical to the unsharded single-GPU call.  Prints one JSON line per rank."""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

RAYS = (0.003, 10.0, 0.0, 1.5)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    pkg = importlib.import_module("monst3r-slam_b200")
    abi = importlib.import_module("monst3r-slam_b200.abi")
    synth = importlib.import_module("monst3r-slam_b200.synth")
    mdist = importlib.import_module("monst3r-slam_b200.dist")
    be = pkg.load_drop_in()
    out = {"rank": rank, "world": world, "cases": {}}

    g = synth.make_graph_exact(40, 90, 3000, seed=40, noise=1e-3, device=dev)  # dim 273: the multi-CTA solver
    order = mdist.pair_aligned_order(g["ii"], g["jj"])
    ii, jj = g["ii"][order].contiguous(), g["jj"][order].contiguous()
    idx, valid, Q = g["idx"][order].contiguous(), g["valid"][order].contiguous(), g["Q"][order].contiguous()
    E2 = ii.numel()

    def run_be(lo=None, hi=None, how=None):
        Twc = g["Twc"].clone()
        if lo is None:
            be.clear_edge_shard()
            sl = slice(0, E2)
        else:
            be.set_edge_shard(lo, hi, how)
            sl = slice(lo, hi)
        (dx,) = be.gauss_newton_rays(Twc, g["Xs"], g["Cs"], ii, jj, idx[sl].contiguous(), valid[sl].contiguous(),
                                     Q[sl].contiguous(), *RAYS, 10, 0.0)
        torch.cuda.synchronize()
        be.clear_edge_shard()
        return Twc, dx, list(be.last_info())

    # INTEGRATION.md §3, verbatim
    uid = [be.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    be.init_distributed(rank, world, uid[0])

    T_ref, dx_ref, info_ref = run_be()  # unsharded (a communicator exists but no shard is set: plain local call)
    out["unsharded_info"] = info_ref

    def record(name, T, dx, info):
        same = bool(torch.equal(T, T_ref) and torch.equal(dx, dx_ref))
        box = [None] * world
        dist.all_gather_object(box, T.cpu().numpy().tobytes())
        out["cases"][name] = {"bit_identical_to_unsharded": same, "identical_across_ranks": all(b == box[0] for b in box),
                              "info": info, "max_abs_diff": float((T - T_ref).abs().max())}

    parts = mdist.partition_edges(E2, world)
    assert E2 % world == 0
    how = "allgather" if mdist.exchange_for_partition(parts, E2) == abi.EXCHANGE_ALLGATHER else "allreduce"
    out["auto_exchange"] = how
    record("pybind_equal_allgather", *run_be(*parts[rank], "allgather"))
    record("pybind_equal_allreduce", *run_be(*parts[rank], "allreduce"))
    # uneven contiguous partition (rank 0 gets 7 edges more than an equal share): all-reduce is the only legal exchange
    cuts = [0] + [min(E2, (r + 1) * (E2 // world) + 7) for r in range(world - 1)] + [E2]
    uneven = [(cuts[r], cuts[r + 1]) for r in range(world)]
    assert mdist.exchange_for_partition(uneven, E2) == abi.EXCHANGE_ALLREDUCE
    record("pybind_uneven_allreduce", *run_be(*uneven[rank], "allreduce"))
    # asking for the all-gather on a range that is not the equal share must fail BEFORE any collective is enqueued
    # (every rank asks, every rank fails: nobody is left waiting)
    try:
        run_be(*uneven[rank], "allgather")
        out["cases"]["allgather_on_uneven_rejected"] = False
    except RuntimeError:
        be.clear_edge_shard()
        out["cases"]["allgather_on_uneven_rejected"] = True
    torch.cuda.synchronize()
    dist.barrier()

    # the same through the C ABI (ctypes), caller-owned sharded workspace
    comm = mdist.init_comm(rank, world, dev)
    prm = abi.make_params(abi.GN_RAYS, *RAYS, 10, 0.0)
    for name, pr in (("cabi_equal_allgather", parts), ("cabi_uneven_allreduce", uneven)):
        lo, hi = pr[rank]
        sh = mdist.make_shard(lo, hi, comm, mdist.exchange_for_partition(pr, E2))
        ws, _ = abi.gn_workspace(40, E2, 3000, dev, edges_local=hi - lo)
        Twc = g["Twc"].clone()
        dx, info = abi.gauss_newton(prm, Twc, g["Xs"], g["Cs"], None, ii, jj, idx[lo:hi].contiguous(),
                                    valid[lo:hi].contiguous(), Q[lo:hi].contiguous(), shard=sh, workspace=ws)
        torch.cuda.synchronize()
        record(name, Twc, dx, [int(v) for v in info.cpu().tolist()])
    # the library's own all-gather over NVLink peer memory (no NCCL on the data path): two solves in a row (the second
    # re-uses both parities of the exchange buffer with the next solve's epochs), and a request it must refuse
    xchg = mdist.init_peer_exchange(rank, world, E2)
    lo, hi = parts[rank]
    sh = mdist.make_shard(lo, hi, None, abi.EXCHANGE_PEER, xchg)
    ws, _ = abi.gn_workspace(40, E2, 3000, dev, edges_local=hi - lo)
    # (solves that start with both ranks aligned after a host-side sync are the ones that exposed a flag overtaking other
    # warps' NVLink stores before every storing thread fenced: csrc/peer_exchange.cu)
    for name in ("cabi_equal_peer", "cabi_equal_peer_again", "cabi_equal_peer_3", "cabi_equal_peer_4"):
        Twc = g["Twc"].clone()
        dx, info = abi.gauss_newton(prm, Twc, g["Xs"], g["Cs"], None, ii, jj, idx[lo:hi].contiguous(),
                                    valid[lo:hi].contiguous(), Q[lo:hi].contiguous(), shard=sh, workspace=ws)
        torch.cuda.synchronize()
        record(name, Twc, dx, [int(v) for v in info.cpu().tolist()])
    try:
        ulo, uhi = uneven[rank]
        abi.gauss_newton(prm, g["Twc"].clone(), g["Xs"], g["Cs"], None, ii, jj, idx[ulo:uhi].contiguous(), valid[ulo:uhi].contiguous(),
                         Q[ulo:uhi].contiguous(), shard=mdist.make_shard(ulo, uhi, None, abi.EXCHANGE_PEER, xchg),
                         workspace=abi.gn_workspace(40, E2, 3000, dev, edges_local=uhi - ulo)[0])
        out["cases"]["peer_on_uneven_rejected"] = False
    except abi.M3RError:
        out["cases"]["peer_on_uneven_rejected"] = True
    torch.cuda.synchronize()
    mdist.destroy_peer_exchange(xchg)
    # stand-alone m3r_comm_allreduce_f64
    buf = torch.arange(1000, dtype=torch.float64, device=dev) * (rank + 1)
    abi._check(abi.lib.m3r_comm_allreduce_f64(comm, abi._ptr(buf), abi.C.c_size_t(buf.numel()), abi._stream()),
               "m3r_comm_allreduce_f64")
    torch.cuda.synchronize()
    want = torch.arange(1000, dtype=torch.float64, device=dev) * (world * (world + 1) // 2)
    out["cases"]["comm_allreduce_f64_exact"] = bool(torch.equal(buf, want))
    # rank-sharded edge store: the same add_factors calls on every rank, slabs land on their owners,
    # sharded solve; against the single-device FactorGraph fed the same pairs (edge order differs -> fp64 summation order)
    fgm = importlib.import_module("monst3r-slam_b200.factor_graph")
    N, n, Eu = 40, 3000, 90
    ones = torch.ones((1, n, 1), dtype=torch.float32, device=dev)

    def feed(graph):
        for lo in range(0, Eu, 6):
            hi = min(lo + 6, Eu)
            b = hi - lo
            m = (g["idx"][lo:hi].contiguous(), g["idx"][Eu + lo:Eu + hi].contiguous(), g["valid"][lo:hi].contiguous(),
                 g["valid"][Eu + lo:Eu + hi].contiguous(), ones.expand(b, n, 1).contiguous(), ones.expand(b, n, 1).contiguous(),
                 (g["Q"][lo:hi] ** 2).contiguous(), (g["Q"][Eu + lo:Eu + hi] ** 2).contiguous())
            assert graph.add_factors_from_matches(g["ii"][lo:hi].tolist(), g["jj"][lo:hi].tolist(), m, 0.0)

    def frames_for():
        fr = fgm.KeyframeBuffer(n, capacity=N, device=dev)
        for k in range(N):
            fr.append(g["Xs"][k], g["Cs"][k], g["Twc"][k])
        return fr

    fr1, fr2 = frames_for(), frames_for()
    single = fgm.FactorGraph(frames=fr1, device=dev, capacity=8, cfg={"delta_norm": 0.0})
    feed(single)
    single.solve_GN_rays()
    sharded = fgm.ShardedFactorGraph(fr2, rank, world, comm=comm, device=dev, capacity=8, cfg={"delta_norm": 0.0})
    feed(sharded)
    sharded.solve_GN_rays()
    torch.cuda.synchronize()
    box = [None] * world
    dist.all_gather_object(box, fr2.T_WC[:N].cpu().numpy().tobytes())
    out["cases"]["sharded_store"] = {
        "local_pairs": int(sharded.store.num_und), "global_pairs": int(sharded.num_edges_global),
        "slab_rows_local": int(sharded.store.idx.shape[0]), "identical_across_ranks": all(b == box[0] for b in box),
        "max_abs_diff_vs_single_device_graph": float((fr2.T_WC[:N] - fr1.T_WC[:N]).abs().max()),
        "moved": float((fr2.T_WC[:N] - g["Twc"]).abs().max())}
    abi.comm_destroy(comm)
    be.shutdown_distributed()
    dist.barrier()
    sys.stdout.write("M3R_MULTI " + json.dumps(out) + "\n")
    sys.stdout.flush()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
