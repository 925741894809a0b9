"""Device-resident factor-graph staging: SURVEY.md §8(f) rank 2 (+ rank 4's constrain_points_to_ray).

Host-side mirror of `FactorGraph` (mast3r_slam/global_opt.py:12-215) for everything that is not the
ViT: same method names, argument meaning and return values, but the edge slabs live in an `EdgeStore`
with geometric capacity instead of growing by `torch.cat` on every `add_factors` (global_opt.py:89-96,
O(E^2) bytes over a run) and being `cat`-ed again for every solve (`prep_two_way_edges`, :104-110,
13 B/pt x 2E per solve = 5 GB at config 4).

Layout.  Undirected edge e owns two adjacent directed rows of every slab:
    row 2e   = i -> j : (ii[e], jj[e], idx_i2j, valid_match_j, Qj)
    row 2e+1 = j -> i : (jj[e], ii[e], idx_j2i, valid_match_i, Qi)
so `prep_two_way_edges()` is five zero-copy views `[:2E]` that go straight into `gauss_newton_*`
(both directions adjacent is also the order the accumulate kernel launches in, and what the
pair-aligned multi-GPU partition of dist.py wants).  The reference orders the directed edges
[all i->j, then all j->i]; the normal equations are a sum over edges, so only the fp64 summation
order of the edge blocks differs (tests compare both orders).

The CUDA work is m3r_edge_admit (Q, match counts and the append in one pass) and
m3r_constrain_points_to_ray; torch is used for storage and the O(batch) accept/reject bookkeeping.
There is no CPU path.
"""
import numpy as np
import torch

from . import abi
from . import backends

DEFAULT_CFG = {  # config/base.yaml:42-57 (local_opt)
    "pin": 1, "window_size": 1e6, "C_conf": 0.0, "Q_conf": 1.5, "min_match_frac": 0.1, "pixel_border": -10,
    "depth_eps": 1e-6, "max_iters": 10, "sigma_ray": 0.003, "sigma_dist": 10.0, "sigma_pixel": 1.0,
    "sigma_depth": 10.0, "sigma_point": 0.05, "delta_norm": 1e-8,
}


class EdgeStore:
    """Slabs [2*capacity, n] for idx (int64), valid (bool), Q (f32) + ii/jj [2*capacity]; capacity doubles."""

    def __init__(self, num_pts, capacity=64, device="cuda"):
        self.n = int(num_pts)
        self.device = torch.device(device)
        self.capacity = 0
        self.num_und = 0
        self.idx = self.valid = self.Q = self.ii2 = self.jj2 = None
        self.moved_bytes = 0  # bytes copied by capacity growth (amortised O(1) per edge)
        self.reserve(max(int(capacity), 1))

    def reserve(self, capacity):
        if capacity <= self.capacity:
            return
        cap = max(capacity, 2 * self.capacity)
        rows, n, dev = 2 * cap, self.n, self.device
        new = (torch.empty((rows, n), dtype=torch.int64, device=dev), torch.empty((rows, n), dtype=torch.bool, device=dev),
               torch.empty((rows, n), dtype=torch.float32, device=dev), torch.empty(rows, dtype=torch.int64, device=dev),
               torch.empty(rows, dtype=torch.int64, device=dev))
        used = 2 * self.num_und
        if used:
            for dst, src in zip(new, (self.idx, self.valid, self.Q, self.ii2, self.jj2)):
                dst[:used].copy_(src[:used])
            self.moved_bytes += used * n * 13
        self.idx, self.valid, self.Q, self.ii2, self.jj2 = new
        self.capacity = cap

    def move_slot(self, src, dst):
        """undirected slot src -> dst (dst < src); used only when a batch has rejected pairs"""
        for t in (self.idx, self.valid, self.Q, self.ii2, self.jj2):
            t[2 * dst:2 * dst + 2].copy_(t[2 * src:2 * src + 2])


class KeyframeBuffer:
    """Minimal device-resident stand-in for SharedKeyframes (mast3r_slam/frame.py): the three per-keyframe
    tensors the graph needs, preallocated so `get_poses_points` is a view when the graph holds keyframes 0..N-1."""

    def __init__(self, num_pts, capacity=512, device="cuda"):
        self.X_canon = torch.zeros((capacity, num_pts, 3), dtype=torch.float32, device=device)
        self.C = torch.zeros((capacity, num_pts, 1), dtype=torch.float32, device=device)
        self.T_WC = torch.zeros((capacity, 8), dtype=torch.float32, device=device)
        self.T_WC[:, 6] = 1.0
        self.T_WC[:, 7] = 1.0
        self.count = 0

    def __len__(self):
        return self.count

    def append(self, X_canon, C, T_WC):
        k = self.count
        self.X_canon[k].copy_(X_canon.reshape(-1, 3))
        self.C[k].copy_(C.reshape(-1, 1))
        self.T_WC[k].copy_(T_WC.reshape(8))
        self.count += 1
        return k

    def gather(self, unique_kf_idx, is_prefix=None):
        N = unique_kf_idx.numel()
        if is_prefix is None:  # (two scalar reads; FactorGraph passes the answer from its host-side id set)
            is_prefix = bool(N) and int(unique_kf_idx[0]) == 0 and int(unique_kf_idx[-1]) == N - 1
        if is_prefix:  # sorted unique ids 0..N-1: zero-copy views, poses updated in place
            return self.X_canon[:N], self.T_WC[:N], self.C[:N], True
        return (self.X_canon.index_select(0, unique_kf_idx), self.T_WC.index_select(0, unique_kf_idx),
                self.C.index_select(0, unique_kf_idx), False)

    def update_T_WCs(self, T_WCs, idx):
        self.T_WC.index_copy_(0, idx, T_WCs)


class _PoseData:
    """what SharedKeyframes.update_T_WCs reads of a lietorch.Sim3: `.data` [k,1,8]"""

    def __init__(self, data):
        self.data = data


class FactorGraph:
    """global_opt.py:12-215 with the model call factored out: `add_factors` takes a `match_fn(ii, jj)`
    returning the eight tensors of mast3r_match_symmetric; `add_factors_from_matches` takes them directly."""

    def __init__(self, frames, K=None, device="cuda", cfg=None, match_fn=None, num_pts=None, capacity=64):
        self.frames = frames
        self.device = torch.device(device)
        self.cfg = dict(DEFAULT_CFG)
        if cfg:
            self.cfg.update(cfg)
        self.K = K
        self.match_fn = match_fn
        self.window_size = self.cfg["window_size"]
        self._num_pts = num_pts
        self._capacity = capacity
        self.store = None
        self._ids = set()      # keyframe ids in the graph, kept on the host: no torch.unique, no sync per solve
        self._unique = None

    # -- views with the reference's attribute names (global_opt.py:17-25), zero-copy, strided over the pairs
    _EMPTY = {"ii2": (torch.int64, ()), "jj2": (torch.int64, ()), "idx": (torch.int64, (0,)), "valid": (torch.bool, (0,)),
              "Q": (torch.float32, (0,))}

    def _dir(self, name, d):
        """direction d (0: i->j, 1: j->i) of slab `name`; before the first add_factors the reference's attributes are
        EMPTY tensors (global_opt.py:17-25), so are these"""
        if self.store is None:
            dt, tail = self._EMPTY[name]
            return torch.empty((0,) + tail, dtype=dt, device=self.device)
        return getattr(self.store, name)[d:2 * self.num_edges:2]

    @property
    def num_edges(self):
        return self.store.num_und if self.store is not None else 0

    ii = property(lambda self: self._dir("ii2", 0))
    jj = property(lambda self: self._dir("jj2", 0))
    idx_ii2jj = property(lambda self: self._dir("idx", 0))
    idx_jj2ii = property(lambda self: self._dir("idx", 1))
    valid_match_j = property(lambda self: self._dir("valid", 0).unsqueeze(-1))
    valid_match_i = property(lambda self: self._dir("valid", 1).unsqueeze(-1))
    Q_ii2jj = property(lambda self: self._dir("Q", 0).unsqueeze(-1))
    Q_jj2ii = property(lambda self: self._dir("Q", 1).unsqueeze(-1))

    def add_factors(self, ii, jj, min_match_frac, is_reloc=False):
        """global_opt.py:31-99; the symmetric matcher (:40-52) is the injected match_fn"""
        if self.match_fn is None:
            raise RuntimeError("FactorGraph.add_factors needs match_fn (the ViT matcher is outside this backend); "
                               "use add_factors_from_matches with precomputed matches")
        return self.add_factors_from_matches(ii, jj, self.match_fn(ii, jj), min_match_frac, is_reloc)

    def add_factors_from_matches(self, ii, jj, matches, min_match_frac, is_reloc=False):
        """global_opt.py:54-99.  matches = (idx_i2j, idx_j2i, valid_match_j, valid_match_i, Qii, Qjj, Qji, Qij)
        shaped as mast3r_match_symmetric returns them ([b,n], [b,n,1] ...)."""
        idx_i2j, idx_j2i, valid_match_j, valid_match_i, Qii, Qjj, Qji, Qij = matches
        batch, n = idx_i2j.shape[0], idx_i2j.shape[1]
        if self.store is None:
            self.store = EdgeStore(self._num_pts or n, self._capacity, self.device)
        st = self.store
        if n != st.n:
            raise RuntimeError(f"add_factors: {n} points per map, the store holds {st.n}")
        if len(ii) != batch or len(jj) != batch:
            raise RuntimeError("add_factors: ii/jj and the match batch differ in length")
        st.reserve(st.num_und + batch)
        cont = lambda t: t.contiguous()
        counts = abi.edge_admit(cont(idx_i2j), cont(idx_j2i), cont(valid_match_j), cont(valid_match_i), cont(Qii),
                                cont(Qjj), cont(Qji), cont(Qij), float(self.cfg["Q_conf"]), st.idx, st.valid, st.Q,
                                st.num_und)
        # :66-76 — both directions must clear min_match_frac unless the pair is consecutive.  Decided on the host
        # from the 2*batch integer counts (the one sync; the reference syncs at `.any()`, :77, too) with torch's
        # arithmetic: float32(count) / float32(n), compared with float32(min_match_frac).
        c = counts.cpu().numpy().astype(np.float32) / np.float32(n)
        invalid = np.minimum(c[:, 0], c[:, 1]) < np.float32(min_match_frac)
        ii_h = np.asarray([int(v) for v in ii], dtype=np.int64)
        jj_h = np.asarray([int(v) for v in jj], dtype=np.int64)
        invalid = (~(ii_h == (jj_h - 1))) & invalid
        if invalid.any() and is_reloc:
            return False  # provisional rows past num_und are simply not committed
        keep = [b for b in range(batch) if not invalid[b]]
        base = st.num_und
        for new_b, b in enumerate(keep):
            if new_b != b:
                st.move_slot(base + b, base + new_b)
        if keep:
            pairs = np.stack((ii_h[keep], jj_h[keep]), axis=1)  # [k, 2]
            rows = slice(2 * base, 2 * (base + len(keep)))
            st.ii2[rows] = torch.from_numpy(pairs.reshape(-1).copy()).to(self.device)          # i0 j0 i1 j1 ...
            st.jj2[rows] = torch.from_numpy(pairs[:, ::-1].reshape(-1).copy()).to(self.device)  # j0 i0 j1 i1 ...
        st.num_und = base + len(keep)
        for b in keep:
            self._ids.add(int(ii[b]))
            self._ids.add(int(jj[b]))
        self._unique = None
        return len(keep) > 0

    def get_unique_kf_idx(self):
        """:101-102"""
        if self._unique is None:
            self._unique = torch.as_tensor(sorted(self._ids), dtype=torch.int64, device=self.device)
        return self._unique

    def prep_two_way_edges(self):
        """:104-110 without the five torch.cat: views in pair-interleaved order"""
        st, E2 = self.store, 2 * self.num_edges
        return st.ii2[:E2], st.jj2[:E2], st.idx[:E2], st.valid[:E2].unsqueeze(-1), st.Q[:E2].unsqueeze(-1)

    def get_poses_points(self, unique_kf_idx):
        """:112-119 -> (Xs [N,n,3], T_WCs [N,8] sim3 data, Cs [N,n,1], poses_are_views)"""
        if hasattr(self.frames, "gather"):
            N = len(self._ids)
            return self.frames.gather(unique_kf_idx, is_prefix=bool(N) and min(self._ids) == 0 and max(self._ids) == N - 1)
        kfs = [self.frames[int(i)] for i in unique_kf_idx]
        Xs = torch.stack([kf.X_canon for kf in kfs])
        T = torch.stack([getattr(kf.T_WC, "data", kf.T_WC).reshape(-1)[:8] for kf in kfs])
        Cs = torch.stack([kf.get_average_conf() for kf in kfs])
        return Xs, T, Cs, False

    def _finish(self, T_WCs, unique_kf_idx, in_place):
        pin = self.cfg["pin"]
        if in_place or not hasattr(self.frames, "update_T_WCs"):
            return
        if isinstance(self.frames, KeyframeBuffer):
            self.frames.update_T_WCs(T_WCs[pin:], unique_kf_idx[pin:])  # :152, :213
        else:
            # the reference's SharedKeyframes (frame.py:362-364) does `self.T_WC[idx] = T_WCs.data` on a [buffer,1,8]
            # tensor with a lietorch.Sim3 whose .data is [k,1,8]: hand it the same shape behind the same attribute
            self.frames.update_T_WCs(_PoseData(T_WCs[pin:].reshape(-1, 1, 8)), unique_kf_idx[pin:])

    def solve_GN_rays(self):
        """:121-153"""
        c = self.cfg
        unique_kf_idx = self.get_unique_kf_idx()
        if unique_kf_idx.numel() <= c["pin"]:
            return None
        Xs, T_WCs, Cs, in_place = self.get_poses_points(unique_kf_idx)
        ii, jj, idx, valid, Q = self.prep_two_way_edges()
        pose = T_WCs if in_place else T_WCs.contiguous()
        out = backends.gauss_newton_rays(pose, Xs, Cs, ii, jj, idx, valid, Q, c["sigma_ray"], c["sigma_dist"],
                                         c["C_conf"], c["Q_conf"], c["max_iters"], c["delta_norm"])
        self._finish(pose, unique_kf_idx, in_place)
        return out

    def solve_GN_calib(self, img_size):
        """:155-215; img_size = (height, width) (the reference reads it from frames[0].img)"""
        c = self.cfg
        unique_kf_idx = self.get_unique_kf_idx()
        if unique_kf_idx.numel() <= c["pin"]:
            return None
        Xs, T_WCs, Cs, in_place = self.get_poses_points(unique_kf_idx)
        Xs = abi.constrain_points_to_ray(img_size, Xs.contiguous(), self.K)  # :166-167
        ii, jj, idx, valid, Q = self.prep_two_way_edges()
        pose = T_WCs if in_place else T_WCs.contiguous()
        h, w = int(img_size[0]), int(img_size[1])
        out = backends.gauss_newton_calib(pose, Xs, Cs, self.K, ii, jj, idx, valid, Q, h, w, c["pixel_border"],
                                          c["depth_eps"], c["sigma_pixel"], c["sigma_depth"], c["C_conf"],
                                          c["Q_conf"], c["max_iters"], c["delta_norm"])
        self._finish(pose, unique_kf_idx, in_place)
        return out


class ShardPlan:
    """Host bookkeeping of the RANK-SHARDED edge store (multi-GPU half): which rank owns which pair and
    where its directed edges sit in the global edge list.  Pure python, identical on every rank by construction (every rank
    sees the same add_factors calls and the same accept / reject decisions).

      * ownership: pair number c (counted over every pair ever OFFERED) belongs to rank c mod world -- round-robin, so
        slabs stay balanced without ever moving between ranks;
      * global directed edge list used by a solve: rank-major -- rank 0's accepted pairs in acceptance order, then rank
        1's, ... each pair as (i->j, j->i).  A rank's edges are therefore one contiguous range [lo, hi) of the list,
        which is what m3r_gn_shard wants, and its local slabs are already in that order.
    """

    def __init__(self, world):
        self.world = int(world)
        self.offered = 0
        self.accepted = [[] for _ in range(self.world)]  # per rank: [(i, j)] in acceptance order

    def owners(self, batch):
        own = [(self.offered + b) % self.world for b in range(batch)]
        self.offered += batch
        return own

    def commit(self, pairs_with_owner):
        for i, j, r in pairs_with_owner:
            self.accepted[r].append((int(i), int(j)))

    def ranges(self):
        """[(lo, hi)] of DIRECTED edges per rank, and the total"""
        out, lo = [], 0
        for r in range(self.world):
            hi = lo + 2 * len(self.accepted[r])
            out.append((lo, hi))
            lo = hi
        return out, lo

    def directed_lists(self):
        ii, jj = [], []
        for r in range(self.world):
            for i, j in self.accepted[r]:
                ii += [i, j]
                jj += [j, i]
        return ii, jj


class ShardedFactorGraph(FactorGraph):
    """FactorGraph whose edge slabs are sharded over the ranks of a process group (one process per GPU): `add_factors` on
    rank r lands a pair's slabs on its OWNER only (13 B per point per directed edge: 96 % of the graph's bytes), keyframes
    and poses are replicated, and `solve_GN_rays / _calib` is the edge-sharded solve of DESIGN.md section 4 (one exchange of
    the fp64 edge blocks per iteration; every rank ends with bit-identical poses).  Every rank must make the same calls.
    `comm`: the backend's NCCL communicator (dist.init_comm); `group`: the torch.distributed group used for the host-side
    exchange of the 2 x batch integer match counts (the accept / reject rule needs them on every rank)."""

    def __init__(self, frames, rank, world, comm=None, group=None, **kw):
        super().__init__(frames, **kw)
        self.rank, self.world, self.comm, self.group = int(rank), int(world), comm, group
        self.plan = ShardPlan(world)
        self._edge_cache = None

    def add_factors_from_matches(self, ii, jj, matches, min_match_frac, is_reloc=False):
        """as FactorGraph.add_factors_from_matches; `matches` holds the whole batch, the rank admits the rows it owns"""
        idx_i2j = matches[0]
        batch, n = idx_i2j.shape[0], idx_i2j.shape[1]
        if self.store is None:
            self.store = EdgeStore(self._num_pts or n, self._capacity, self.device)
        st = self.store
        own = self.plan.owners(batch)
        mine = [b for b in range(batch) if own[b] == self.rank]
        local = {}
        if mine:
            st.reserve(st.num_und + len(mine))
            sel = torch.as_tensor(mine, dtype=torch.int64, device=self.device)
            sub = tuple(t.index_select(0, sel).contiguous() for t in matches)
            c = abi.edge_admit(*sub, float(self.cfg["Q_conf"]), st.idx, st.valid, st.Q, st.num_und).cpu().numpy()
            local = {b: (int(c[k, 0]), int(c[k, 1])) for k, b in enumerate(mine)}
        counts = np.zeros((batch, 2), np.int64)
        if self.world > 1:
            import torch.distributed as tdist
            box = [None] * self.world
            tdist.all_gather_object(box, local, group=self.group)
            for d in box:
                for b, v in d.items():
                    counts[b] = v
        else:
            for b, v in local.items():
                counts[b] = v
        c32 = counts.astype(np.float32) / np.float32(n)
        invalid = np.minimum(c32[:, 0], c32[:, 1]) < np.float32(min_match_frac)
        ii_h = np.asarray([int(v) for v in ii], dtype=np.int64)
        jj_h = np.asarray([int(v) for v in jj], dtype=np.int64)
        invalid = (~(ii_h == (jj_h - 1))) & invalid
        if invalid.any() and is_reloc:
            return False
        keep = [b for b in range(batch) if not invalid[b]]
        base, new_local = st.num_und, 0
        for k, b in enumerate(mine):  # compact the owned, accepted rows
            if invalid[b]:
                continue
            if new_local != k:
                st.move_slot(base + k, base + new_local)
            rows = slice(2 * (base + new_local), 2 * (base + new_local) + 2)
            st.ii2[rows] = torch.tensor([ii_h[b], jj_h[b]], dtype=torch.int64, device=self.device)
            st.jj2[rows] = torch.tensor([jj_h[b], ii_h[b]], dtype=torch.int64, device=self.device)
            new_local += 1
        st.num_und = base + new_local
        self.plan.commit([(ii_h[b], jj_h[b], own[b]) for b in keep])
        for b in keep:
            self._ids.add(int(ii_h[b]))
            self._ids.add(int(jj_h[b]))
        self._unique = None
        self._edge_cache = None
        return len(keep) > 0

    @property
    def num_edges_global(self):
        return sum(len(a) for a in self.plan.accepted)

    def _sharded_args(self):
        from . import dist as mdist
        if self._edge_cache is None:
            ii, jj = self.plan.directed_lists()
            parts, total = self.plan.ranges()
            if any(hi == lo for lo, hi in parts):
                raise RuntimeError("sharded solve: every rank must own at least one accepted pair")
            self._edge_cache = (torch.as_tensor(ii, dtype=torch.int64, device=self.device),
                                torch.as_tensor(jj, dtype=torch.int64, device=self.device), parts, total)
        ii, jj, parts, total = self._edge_cache
        lo, hi = parts[self.rank]
        shard = None
        if self.world > 1:
            shard = mdist.make_shard(lo, hi, self.comm, mdist.exchange_for_partition(parts, total))
        st, E2 = self.store, hi - lo
        return ii, jj, st.idx[:E2], st.valid[:E2].unsqueeze(-1), st.Q[:E2].unsqueeze(-1), shard

    def _solve(self, kind, K, img_size):
        c = self.cfg
        unique_kf_idx = self.get_unique_kf_idx()
        if unique_kf_idx.numel() <= c["pin"]:
            return None
        Xs, T_WCs, Cs, in_place = self.get_poses_points(unique_kf_idx)
        if kind == abi.GN_CALIB:
            Xs = abi.constrain_points_to_ray(img_size, Xs.contiguous(), K)
            prm = abi.make_params(abi.GN_CALIB, c["sigma_pixel"], c["sigma_depth"], c["C_conf"], c["Q_conf"], c["max_iters"],
                                  c["delta_norm"], int(img_size[0]), int(img_size[1]), c["pixel_border"], c["depth_eps"])
        else:
            prm = abi.make_params(abi.GN_RAYS, c["sigma_ray"], c["sigma_dist"], c["C_conf"], c["Q_conf"], c["max_iters"],
                                  c["delta_norm"])
        ii, jj, idx, valid, Q, shard = self._sharded_args()
        pose = T_WCs if in_place else T_WCs.contiguous()
        ws, _ = abi.gn_workspace(Xs.shape[0], ii.numel(), Xs.shape[1], self.device, edges_local=idx.shape[0])
        out = abi.gauss_newton(prm, pose, Xs.contiguous(), Cs.contiguous(), K, ii, jj, idx, valid, Q, shard=shard, workspace=ws)
        self._finish(pose, unique_kf_idx, in_place)
        return out

    def solve_GN_rays(self):
        return self._solve(abi.GN_RAYS, None, None)

    def solve_GN_calib(self, img_size):
        return self._solve(abi.GN_CALIB, self.K, img_size)
