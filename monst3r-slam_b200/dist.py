"""Multi-GPU global optimisation: edge partitioner + communicator setup.

The keyframe graph is partitioned across ranks BY DIRECTED EDGE: rank r owns a contiguous range of
the (two-way) edge list and holds only those edges' slabs (idx / valid / Q, 13 B per point per edge,
96 % of the bytes); Xs / Cs / Twc are replicated.  Once per Gauss-Newton iteration the per-edge
7x7 + 7 fp64 blocks are exchanged (in-place all-gather when the GLOBAL partition has equal ranges — decided by
`exchange_for_partition`, identically on every rank — else a sum all-reduce over zero-padded disjoint slots; exact
and bit-identical on every rank either way) and every rank
runs the same deterministic solve + retraction.
One process per GPU; torch.distributed is only used to ship the 128-byte NCCL unique id.
"""
import ctypes as C


def partition_edges(num_edges, world, rank=None):
    """Balanced contiguous ranges over the directed edge list.  -> [(lo, hi)] or (lo, hi) for `rank`."""
    base, rem = divmod(int(num_edges), int(world))
    out, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < rem else 0)
        out.append((lo, hi))
        lo = hi
    return out if rank is None else out[rank]


def shard_edge_tensors(lo, hi, *tensors):
    """slice per-edge slabs (first dim = directed edge) to the owned range, contiguous"""
    return tuple(t[lo:hi].contiguous() for t in tensors)


def pair_aligned_order(ii, jj):
    """Permutation of the directed edge list that puts both directions of a pair next to each other
    and orders pairs by min(i, j): a contiguous shard then touches a compact keyframe range."""
    import torch
    a = torch.minimum(ii, jj)
    b = torch.maximum(ii, jj)
    big = int(max(int(b.max().item()) + 1, 1))
    key = (a * big + b) * 2 + (ii > jj).long()
    return torch.argsort(key, stable=True)


def init_comm(rank, world, device=None, group=None):
    """Create the backend's own NCCL communicator.  Needs torch.distributed to be initialised (any
    backend) to broadcast the unique id from rank 0.  Returns an abi.GnShard-compatible handle."""
    import torch
    import torch.distributed as dist
    from . import abi
    if device is not None:
        torch.cuda.set_device(device)
    if rank == 0:
        uid = abi.comm_unique_id()
    else:
        uid = bytes(128)
    if world > 1:
        box = [uid]
        dist.broadcast_object_list(box, src=0, group=group)
        uid = box[0]
    return abi.comm_init(rank, world, uid)


def exchange_for_partition(parts, num_edges):
    """Which collective carries the per-iteration exchange for the GLOBAL partition `parts` = [(lo, hi)] in rank
    order.  A pure function of data every rank holds, so all ranks pick the same collective (m3r_abi.h:
    m3r_gn_shard.exchange).  Equal contiguous ranges covering the list -> in-place all-gather, else all-reduce."""
    from . import abi
    world = len(parts)
    if world < 1 or num_edges % world != 0:
        return abi.EXCHANGE_ALLREDUCE
    cnt = num_edges // world
    ok = all(int(lo) == r * cnt and int(hi) == (r + 1) * cnt for r, (lo, hi) in enumerate(parts))
    return abi.EXCHANGE_ALLGATHER if ok else abi.EXCHANGE_ALLREDUCE


def make_shard(lo, hi, comm=None, exchange=0, xchg=None):
    """`exchange` must be the same on every rank: take it from exchange_for_partition(all ranks' ranges)
    (abi.EXCHANGE_PEER with `xchg` from init_peer_exchange when the partition is the equal one)."""
    from . import abi
    return abi.GnShard(int(lo), int(hi), comm if comm is not None else C.c_void_p(0), int(exchange),
                       xchg if xchg is not None else C.c_void_p(0))


def init_peer_exchange(rank, world, max_directed_edges, group=None):
    """The library's own all-gather of the edge blocks over NVLink peer memory (csrc/peer_exchange.cu): every rank
    allocates an exchange buffer for graphs of up to `max_directed_edges` edges, the 64-byte CUDA IPC handles go
    round with torch.distributed (any backend), every rank maps its peers' buffers.  One process per GPU, one node.
    -> handle for make_shard(..., exchange=abi.EXCHANGE_PEER, xchg=handle); destroy_peer_exchange() when done."""
    import torch.distributed as dist
    from . import abi
    x, mine = abi.xchg_create(int(max_directed_edges) * 56 * 8)
    handles = [None] * world
    dist.all_gather_object(handles, mine, group=group)
    abi.xchg_connect(x, rank, world, handles)
    dist.barrier(group=group)  # nobody pushes before everybody is connected
    return x


def destroy_peer_exchange(x, group=None):
    import torch
    import torch.distributed as dist
    from . import abi
    torch.cuda.synchronize()
    dist.barrier(group=group)  # a peer may still be writing into this rank's buffer otherwise
    abi.xchg_destroy(x)
