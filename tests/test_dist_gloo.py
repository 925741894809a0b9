"""world_size-2 `gloo` test of the edge-sharded Gauss-Newton protocol on CPU:
the host logic of monst3r-slam_b200/dist.py (pair-aligned ordering, contiguous partition, slab slicing,
unique-id broadcast) around the ORACLE's accumulate / solve, with the per-edge blocks exchanged as a
sum all-reduce over zero-padded disjoint slots -- exactly what the CUDA path does with NCCL.
Checks: every rank ends with bit-identical poses, identical to the unsharded single-process loop."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PRM = dict(sigma_a=0.003, sigma_b=10.0, C_thresh=0.0, Q_thresh=1.5)
ITERS = 3


def _graph():
    import importlib
    synth = importlib.import_module("monst3r-slam_b200.synth")
    g = synth.make_graph_exact(5, 7, 600, seed=4, noise=1e-3, device="cpu")
    return {k: (v.numpy() if torch.is_tensor(v) else v) for k, v in g.items()}


def _loop(g, order, lo, hi, allreduce):
    """GN loop of one rank that owns directed edges order[lo:hi]"""
    from oracle import oracle as orc
    ii, jj = g["ii"][order], g["jj"][order]
    ii_c, jj_c, nuniq = orc.compact_indices(ii, jj)  # every rank sees the FULL edge list
    E = len(ii)
    idx, valid, Q = g["idx"][order][lo:hi], g["valid"][order][lo:hi], g["Q"][order][lo:hi]  # owned slabs only
    Twc = g["Twc"].copy()
    for _ in range(ITERS):
        Hs_l, gs_l = orc.gn_accumulate("rays", Twc, g["Xs"], g["Cs"], None, ii_c[lo:hi], jj_c[lo:hi], idx, valid, Q,
                                       PRM["sigma_a"], PRM["sigma_b"], PRM["C_thresh"], PRM["Q_thresh"])
        buf = np.zeros((E, 4 * 49 + 2 * 7), np.float64)  # non-owned slots stay exact zeros
        buf[lo:hi, :196] = Hs_l.transpose(1, 0, 2, 3).reshape(hi - lo, 196)
        buf[lo:hi, 196:] = gs_l.transpose(1, 0, 2).reshape(hi - lo, 14)
        buf = allreduce(buf)
        Hs = buf[:, :196].reshape(E, 4, 7, 7).transpose(1, 0, 2, 3).astype(np.float32)
        gs = buf[:, 196:].reshape(E, 2, 7).transpose(1, 0, 2).astype(np.float32)
        dx, failed = orc.gn_solve(Hs, gs, ii_c - 1, jj_c - 1, nuniq - 1)
        assert not failed
        Twc = orc.pose_retr(Twc, dx)
    return Twc


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import importlib
        mdist = importlib.import_module("monst3r-slam_b200.dist")
        g = _graph()
        order = mdist.pair_aligned_order(torch.from_numpy(g["ii"]), torch.from_numpy(g["jj"])).numpy()
        lo, hi = mdist.partition_edges(len(order), world, rank)
        # the 128-byte communicator id travels the same way init_comm ships NCCL's
        box = [bytes(range(128)) if rank == 0 else bytes(128)]
        dist.broadcast_object_list(box, src=0)
        assert box[0] == bytes(range(128))

        def allreduce(buf):
            t = torch.from_numpy(buf)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            return t.numpy()

        Twc = _loop(g, order, lo, hi, allreduce)
        gathered = [torch.zeros(Twc.shape, dtype=torch.float32) for _ in range(world)]
        dist.all_gather(gathered, torch.from_numpy(Twc))
        if rank == 0:
            np.save(out, np.stack([t.numpy() for t in gathered]))
    finally:
        dist.destroy_process_group()


def test_edge_sharded_gn_two_ranks_gloo(tmp_path):
    from oracle import oracle as orc
    orc._lib()
    out = str(tmp_path / "twc.npy")
    port = 29600 + os.getpid() % 300
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    ranks = np.load(out)
    assert np.array_equal(ranks[0].view(np.uint32), ranks[1].view(np.uint32))  # bit-identical on every rank
    g = _graph()
    import importlib
    mdist = importlib.import_module("monst3r-slam_b200.dist")
    order = mdist.pair_aligned_order(torch.from_numpy(g["ii"]), torch.from_numpy(g["jj"])).numpy()
    single = _loop(g, order, 0, len(order), lambda b: b)
    assert np.array_equal(single.view(np.uint32), ranks[0].view(np.uint32))  # sharding changes nothing
    ref, _, iters, nfail = orc.gauss_newton("rays", g["Twc"], g["Xs"], g["Cs"], None, g["ii"], g["jj"], g["idx"], g["valid"],
                                            g["Q"], PRM["sigma_a"], PRM["sigma_b"], PRM["C_thresh"], PRM["Q_thresh"], ITERS, 0.0)
    assert iters == ITERS and nfail == 0
    assert np.abs(ref - ranks[0]).max() < 1e-5  # edge order differs from the caller's: same poses up to rounding


def test_partition_and_order_properties():
    import importlib
    mdist = importlib.import_module("monst3r-slam_b200.dist")
    for E, world in ((240, 8), (7, 2), (3, 4), (2000, 8)):
        parts = mdist.partition_edges(E, world)
        assert parts[0][0] == 0 and parts[-1][1] == E
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
        sizes = [hi - lo for lo, hi in parts]
        assert max(sizes) - min(sizes) <= 1
    ii = torch.tensor([3, 0, 2, 1, 0, 2]); jj = torch.tensor([0, 3, 1, 2, 2, 0])
    o = mdist.pair_aligned_order(ii, jj)
    pairs = [(min(int(ii[k]), int(jj[k])), max(int(ii[k]), int(jj[k]))) for k in o]
    assert pairs == sorted(pairs) and pairs[0] == pairs[1] and pairs[2] == pairs[3]


def test_exchange_choice_is_a_function_of_the_global_partition():
    """ranks must not pick the collective from their own range.  exchange_for_partition sees the whole
    partition: the mixed case of the advisor (world 4, E 8, ranges (0,2),(2,5),(5,6),(6,8)) is all-reduce for everyone."""
    import importlib
    mdist = importlib.import_module("monst3r-slam_b200.dist")
    abi = importlib.import_module("monst3r-slam_b200.abi")
    assert mdist.exchange_for_partition(mdist.partition_edges(240, 8), 240) == abi.EXCHANGE_ALLGATHER
    assert mdist.exchange_for_partition(mdist.partition_edges(2000, 8), 2000) == abi.EXCHANGE_ALLGATHER
    assert mdist.exchange_for_partition(mdist.partition_edges(14, 4), 14) == abi.EXCHANGE_ALLREDUCE
    assert mdist.exchange_for_partition([(0, 2), (2, 5), (5, 6), (6, 8)], 8) == abi.EXCHANGE_ALLREDUCE
    assert mdist.exchange_for_partition([(0, 2), (2, 4), (4, 6), (6, 8)], 8) == abi.EXCHANGE_ALLGATHER
    sh = mdist.make_shard(2, 4, None, abi.EXCHANGE_ALLGATHER)
    assert (sh.edge_lo, sh.edge_hi, sh.exchange) == (2, 4, 1)


def test_shard_plan_round_robin_and_rank_major_order():
    """host bookkeeping of the rank-sharded edge store: ownership by offered-pair counter, contiguous rank-major ranges"""
    import importlib
    fgm = importlib.import_module("monst3r-slam_b200.factor_graph")
    p = fgm.ShardPlan(3)
    own = p.owners(4)
    assert own == [0, 1, 2, 0]
    p.commit([(0, 1, own[0]), (1, 2, own[1]), (2, 3, own[3])])  # pair 2 (owner 2) was rejected
    assert p.owners(2) == [1, 2]                                   # the counter runs over OFFERED pairs
    p.commit([(3, 4, 1), (0, 4, 2)])
    parts, total = p.ranges()
    assert parts == [(0, 4), (4, 8), (8, 10)] and total == 10
    ii, jj = p.directed_lists()
    assert ii == [0, 1, 2, 3, 1, 2, 3, 4, 0, 4] and jj == [1, 0, 3, 2, 2, 1, 4, 3, 4, 0]
    mdist = importlib.import_module("monst3r-slam_b200.dist")
    abi = importlib.import_module("monst3r-slam_b200.abi")
    assert mdist.exchange_for_partition(parts, total) == abi.EXCHANGE_ALLREDUCE
