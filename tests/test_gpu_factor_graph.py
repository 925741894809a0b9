"""GPU parity of the factor-graph staging: m3r_edge_admit,
m3r_constrain_points_to_ray and the FactorGraph host mirror, against (a) the numpy oracle, (b) the golden
fixture produced by the reference's own python, (c) the reference's torch expressions run on the GPU.
Integer / index / bool results and the float32 Q are bit-exact; solves meet the GN pose tolerance."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, need_gpu, np_, pkg, pose_diff

pytestmark = pytest.mark.gpu
REL_TOL, ANG_TOL = 1e-5, 1e-4
MATCH_KEYS = ("idx_i2j", "idx_j2i", "valid_j", "valid_i", "Qii", "Qjj", "Qji", "Qij")


@pytest.fixture(scope="module")
def fgo():
    from oracle import factor_graph_oracle
    return factor_graph_oracle


@pytest.fixture(scope="module")
def fgm():
    return pkg("factor_graph")


def _matches(batch, n, seed, valid_prob=0.85, q_off=1.0):
    g = torch.Generator().manual_seed(seed)
    idx = [torch.randint(0, n, (batch, n), generator=g, dtype=torch.int64) for _ in range(2)]
    valid = [torch.rand((batch, n, 1), generator=g) < valid_prob for _ in range(2)]
    Q = [torch.exp(torch.randn((batch, n, 1), generator=g) * 0.5 + q_off).float() for _ in range(4)]
    return (idx[0], idx[1], valid[0], valid[1], Q[0], Q[1], Q[2], Q[3])


@pytest.mark.parametrize("batch,n", [(1, 768), (3, 1001), (2, 4098), (2, 384 * 512)])
def test_edge_admit_bit_exact(abi, fgo, batch, n):
    need_gpu()
    m = _matches(batch, n, seed=50 + batch + n % 97)
    mc = tuple(t.cuda() for t in m)
    cap, slot0 = batch + 3, 2
    s_idx = torch.full((2 * cap, n), -7, dtype=torch.int64, device="cuda")
    s_valid = torch.zeros((2 * cap, n), dtype=torch.bool, device="cuda")
    s_Q = torch.full((2 * cap, n), -1.0, dtype=torch.float32, device="cuda")
    counts = abi.edge_admit(*mc, 1.5, s_idx, s_valid, s_Q, slot0)
    # (c) the reference's expressions on the GPU (global_opt.py:54-67)
    idx_i2j, idx_j2i, vj, vi, Qii, Qjj, Qji, Qij = mc
    binds = torch.arange(batch, device="cuda")[:, None].repeat(1, n)
    Qj = torch.sqrt(Qii[binds, idx_i2j] * Qji)
    Qi = torch.sqrt(Qjj[binds, idx_j2i] * Qij)
    cj = (vj & (Qj > 1.5)).sum(dim=(1, 2))
    ci = (vi & (Qi > 1.5)).sum(dim=(1, 2))
    rows = 2 * slot0 + 2 * torch.arange(batch, device="cuda")
    assert torch.equal(s_Q[rows], Qj[..., 0]) and torch.equal(s_Q[rows + 1], Qi[..., 0])  # float32 bit-exact
    assert torch.equal(s_idx[rows], idx_i2j) and torch.equal(s_idx[rows + 1], idx_j2i)
    assert torch.equal(s_valid[rows], vj[..., 0]) and torch.equal(s_valid[rows + 1], vi[..., 0])
    assert torch.equal(counts[:, 0].long(), cj) and torch.equal(counts[:, 1].long(), ci)
    # (a) the oracle
    assert np.array_equal(np_(s_Q[rows]), fgo.edge_q(np_(m[4]), np_(m[0]), np_(m[6]))[..., 0])
    # rows outside the batch are untouched
    other = torch.ones(2 * cap, dtype=torch.bool, device="cuda")
    other[2 * slot0:2 * (slot0 + batch)] = False
    assert bool((s_idx[other] == -7).all()) and bool((s_Q[other] == -1.0).all())


def _replay_golden(fgm, g, capacity):
    fg = fgm.FactorGraph(frames=None, device="cuda", capacity=capacity)
    rets = []
    for t in range(int(g["num_batches"])):
        m = tuple(torch.from_numpy(g[f"b{t}_{k}"]).cuda() for k in MATCH_KEYS)
        r = fg.add_factors_from_matches(g[f"b{t}_ii"].tolist(), g[f"b{t}_jj"].tolist(), m,
                                        float(g[f"b{t}_min_match_frac"]), bool(g[f"b{t}_is_reloc"]))
        rets.append(bool(r))
        assert fg.num_edges == int(g[f"b{t}_num_edges_after"]), f"batch {t}"
    return fg, rets


@pytest.mark.parametrize("capacity", [1, 64])
def test_factor_graph_replays_reference_fixture(fgm, fgo, capacity):
    """capacity=1 forces geometric growth in between; rejected pairs force slot moves"""
    need_gpu()
    g = np.load(os.path.join(GOLDEN, "factor_graph.npz"))
    fg, rets = _replay_golden(fgm, g, capacity)
    assert rets == [bool(g[f"b{t}_ret"]) for t in range(int(g["num_batches"]))]
    # the oracle replays the same batches: bit-exact state
    st = fgo.GraphState(int(g["n"]))
    for t in range(int(g["num_batches"])):
        fgo.add_factors_tail(st, g[f"b{t}_ii"], g[f"b{t}_jj"], tuple(g[f"b{t}_{k}"] for k in MATCH_KEYS), 1.5,
                             float(g[f"b{t}_min_match_frac"]), bool(g[f"b{t}_is_reloc"]))
    for name in ("ii", "jj", "idx_ii2jj", "idx_jj2ii", "valid_match_j", "valid_match_i", "Q_ii2jj", "Q_jj2ii"):
        mine, ora, ref = np_(getattr(fg, name)), getattr(st, name), g[f"state_{name}"]
        assert mine.shape == ref.shape and mine.dtype == ref.dtype, name
        assert np.array_equal(mine, ora), name
        if name.startswith("Q_"):  # the fixture's CPU torch.sqrt is MKL VML (<= 1 ulp), see test_factor_graph_oracle
            np.testing.assert_array_max_ulp(mine, ref, maxulp=1)
        else:
            assert np.array_equal(mine, ref), name
    assert np.array_equal(np_(fg.get_unique_kf_idx()), g["unique_kf_idx"])
    two_ref = fgo.interleave_two_way(fgo.prep_two_way_edges(st))
    for name, a, b in zip(("ii", "jj", "idx", "valid", "Q"), fg.prep_two_way_edges(), two_ref):
        assert np.array_equal(np_(a), b), name
    if capacity == 1:
        assert fg.store.capacity >= fg.num_edges and fg.store.moved_bytes > 0


def test_views_are_zero_copy(fgm):
    need_gpu()
    g = np.load(os.path.join(GOLDEN, "factor_graph.npz"))
    fg, _ = _replay_golden(fgm, g, 64)
    ii, jj, idx, valid, Q = fg.prep_two_way_edges()
    st = fg.store
    assert idx.data_ptr() == st.idx.data_ptr() and Q.data_ptr() == st.Q.data_ptr()
    assert valid.data_ptr() == st.valid.data_ptr() and ii.data_ptr() == st.ii2.data_ptr()
    assert idx.is_contiguous() and valid.is_contiguous() and Q.is_contiguous()
    assert valid.shape == (2 * fg.num_edges, st.n, 1) and Q.shape == (2 * fg.num_edges, st.n, 1)


def test_constrain_points_to_ray_bit_exact(abi, fgo):
    need_gpu()
    g = np.load(os.path.join(GOLDEN, "factor_graph.npz"))
    h, w = int(g["ray_h"]), int(g["ray_w"])
    out = abi.constrain_points_to_ray((h, w), torch.from_numpy(g["ray_Xs"]).cuda(), torch.from_numpy(g["ray_K"]).cuda())
    assert np.array_equal(np_(out), g["ray_out"])  # the reference's python (sub, div, mul are IEEE on the CPU too)
    # full size, against the reference's expressions on the GPU (geometry.py:107-123) and the oracle, in place
    h, w = 384, 512
    gen = torch.Generator().manual_seed(9)
    Xs = (torch.randn((3, h * w, 3), generator=gen) + torch.tensor([0.0, 0.0, 3.0])).cuda()
    K = torch.tensor([[435.2, 0.0, 255.5], [0.0, 433.1, 191.5], [0.0, 0.0, 1.0]]).cuda()
    u, v = torch.meshgrid(torch.arange(w), torch.arange(h), indexing="xy")
    uv = torch.stack((u, v), dim=-1).unsqueeze(0).repeat(3, 1, 1, 1).to(device="cuda", dtype=torch.float32).view(3, h * w, 2)
    tmp1 = (uv[..., 0] - K[0, 2]) / K[0, 0]
    tmp2 = (uv[..., 1] - K[1, 2]) / K[1, 1]
    ref = torch.stack((Xs[..., 2] * tmp1, Xs[..., 2] * tmp2, Xs[..., 2] * 1.0), dim=-1)
    ora = fgo.constrain_points_to_ray((h, w), np_(Xs), np_(K))
    buf = Xs.clone()
    abi.constrain_points_to_ray((h, w), buf, K, out=buf)
    assert torch.equal(buf, ref) and np.array_equal(np_(buf), ora)


def _store_from_graph(fgm, g, frames, batch=4, capacity=2, K=None):
    """feeds a synth graph (reference two-way order) to the FactorGraph pair by pair"""
    E = g["ii"].shape[0] // 2
    n = g["idx"].shape[1]
    fg = fgm.FactorGraph(frames=frames, K=K, device="cuda", capacity=capacity, cfg={"delta_norm": 0.0})
    ones = torch.ones((batch, n, 1), dtype=torch.float32, device="cuda")
    for lo in range(0, E, batch):
        hi = min(lo + batch, E)
        b = hi - lo
        m = (g["idx"][lo:hi].contiguous(), g["idx"][E + lo:E + hi].contiguous(), g["valid"][lo:hi].contiguous(),
             g["valid"][E + lo:E + hi].contiguous(), ones[:b], ones[:b], (g["Q"][lo:hi] ** 2).contiguous(),
             (g["Q"][E + lo:E + hi] ** 2).contiguous())
        assert fg.add_factors_from_matches(g["ii"][lo:hi].tolist(), g["jj"][lo:hi].tolist(), m, 0.0)
    assert fg.num_edges == E
    return fg


@pytest.mark.parametrize("kind", ["rays", "calib"])
def test_solve_through_store_matches_reference_order(fgm, fgo, oracle, synth, kind):
    """FactorGraph.solve_GN_* on the interleaved zero-copy views == the op called on the reference's
    [fwd..., bwd...] tensors == the CPU oracle GN, within the GN pose tolerance."""
    need_gpu()
    ops = pkg("backends")
    if kind == "calib":
        g = synth.make_graph(6, 9, 48, 64, seed=77, calib=False)  # unconstrained points: the solve must constrain them
    else:
        g = synth.make_graph_exact(6, 9, 3000, seed=78, noise=1e-3)
    g = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in g.items()}
    N, n = g["Xs"].shape[0], g["Xs"].shape[1]
    frames = fgm.KeyframeBuffer(n, capacity=8)
    for k in range(N):
        frames.append(g["Xs"][k], g["Cs"][k], g["Twc"][k])
    fg = _store_from_graph(fgm, g, frames, K=g["K"] if kind == "calib" else None)
    # reference-order tensors with the Q the store holds
    ii, jj = torch.cat((fg.ii, fg.jj)), torch.cat((fg.jj, fg.ii))
    idx = torch.cat((fg.idx_ii2jj, fg.idx_jj2ii)).contiguous()
    valid = torch.cat((fg.valid_match_j, fg.valid_match_i)).contiguous()
    Q = torch.cat((fg.Q_ii2jj, fg.Q_jj2ii)).contiguous()
    assert torch.equal(ii, g["ii"]) and torch.equal(idx, g["idx"]) and torch.equal(valid, g["valid"])
    np.testing.assert_array_max_ulp(np_(Q), np_(g["Q"]), maxulp=1)
    Twc_ref = g["Twc"].clone()
    c = fg.cfg
    if kind == "rays":
        ops.gauss_newton_rays(Twc_ref, g["Xs"], g["Cs"], ii, jj, idx, valid, Q, c["sigma_ray"], c["sigma_dist"],
                              c["C_conf"], c["Q_conf"], c["max_iters"], 0.0)
        Twc_or = oracle.gauss_newton_rays(np_(g["Twc"]), np_(g["Xs"]), np_(g["Cs"]), np_(ii), np_(jj), np_(idx),
                                             np_(valid), np_(Q), c["sigma_ray"], c["sigma_dist"], c["C_conf"],
                                             c["Q_conf"], c["max_iters"], 0.0)[0]
        fg.solve_GN_rays()
    else:
        h, w = g["h"], g["w"]
        Xc = torch.from_numpy(fgo.constrain_points_to_ray((h, w), np_(g["Xs"]), np_(g["K"]))).cuda()
        ops.gauss_newton_calib(Twc_ref, Xc, g["Cs"], g["K"], ii, jj, idx, valid, Q, h, w, c["pixel_border"],
                               c["depth_eps"], c["sigma_pixel"], c["sigma_depth"], c["C_conf"], c["Q_conf"],
                               c["max_iters"], 0.0)
        Twc_or = oracle.gauss_newton_calib(np_(g["Twc"]), np_(Xc), np_(g["Cs"]), np_(g["K"]), np_(ii), np_(jj),
                                              np_(idx), np_(valid), np_(Q), h, w, c["pixel_border"], c["depth_eps"],
                                              c["sigma_pixel"], c["sigma_depth"], c["C_conf"], c["Q_conf"],
                                              c["max_iters"], 0.0)[0]
        fg.solve_GN_calib((h, w))
    mine = np_(frames.T_WC[:N])  # prefix ids -> poses were updated in place
    assert not np.array_equal(mine, np_(g["Twc"]))
    for other, what in ((np_(Twc_ref), "reference-order call"), (np.asarray(Twc_or), "oracle")):
        rel, ang = pose_diff(mine, other)
        assert rel < REL_TOL and ang < ANG_TOL, (what, rel, ang)


def test_non_prefix_keyframe_ids_gather_and_write_back(fgm, synth):
    """graph over keyframes {0, 2, 3, 5}: get_poses_points gathers, update_T_WCs writes back pin: only"""
    need_gpu()
    ops = pkg("backends")
    g = synth.make_graph_exact(4, 5, 2000, seed=79, noise=1e-3)
    g = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in g.items()}
    ids = torch.tensor([0, 2, 3, 5], device="cuda")
    n = g["Xs"].shape[1]
    frames = fgm.KeyframeBuffer(n, capacity=8)
    frames.count = 6
    frames.X_canon[ids] = g["Xs"]
    frames.C[ids] = g["Cs"]
    frames.T_WC[ids] = g["Twc"]
    before = frames.T_WC.clone()
    g2 = dict(g)
    g2["ii"], g2["jj"] = ids[g["ii"]], ids[g["jj"]]
    fg = _store_from_graph(fgm, g2, frames)
    assert np_(fg.get_unique_kf_idx()).tolist() == [0, 2, 3, 5]
    fg.solve_GN_rays()
    Twc_ref = g["Twc"].clone()
    ii, jj, idx, valid, Q = fg.prep_two_way_edges()
    c = fg.cfg
    ops.gauss_newton_rays(Twc_ref, g["Xs"], g["Cs"], ii, jj, idx, valid, Q, c["sigma_ray"], c["sigma_dist"],
                          c["C_conf"], c["Q_conf"], c["max_iters"], 0.0)
    assert torch.equal(frames.T_WC[ids[1:]], Twc_ref[1:])          # same kernel, same inputs: identical
    untouched = torch.tensor([0, 1, 4, 6, 7], device="cuda")
    assert torch.equal(frames.T_WC[untouched], before[untouched])   # pinned pose and foreign slots untouched


def test_is_reloc_rejection_commits_nothing(fgm):
    need_gpu()
    n = 1024
    fg = fgm.FactorGraph(frames=None, device="cuda", capacity=4)
    assert fg.add_factors_from_matches([0], [1], tuple(t.cuda() for t in _matches(1, n, 5)), 0.1)
    snap = [t.clone() for t in fg.prep_two_way_edges()]
    bad = list(_matches(2, n, 6))
    bad[2] = torch.zeros_like(bad[2])  # no valid matches i->j for both pairs
    assert fg.add_factors_from_matches([0, 1], [3, 4], tuple(t.cuda() for t in bad), 0.1, is_reloc=True) is False
    assert fg.num_edges == 1 and np_(fg.get_unique_kf_idx()).tolist() == [0, 1]
    for a, b in zip(snap, fg.prep_two_way_edges()):
        assert torch.equal(a, b)
    # not a relocalisation: the bad pairs are dropped, the call reports that nothing was added
    assert fg.add_factors_from_matches([0, 1], [3, 4], tuple(t.cuda() for t in bad), 0.1) is False
    assert fg.num_edges == 1


class _SharedKeyframesStub:
    """The three things FactorGraph touches of the reference's SharedKeyframes (mast3r_slam/frame.py:245-364):
    T_WC [buffer,1,8], `frames[idx]` -> an object with X_canon / T_WC.data [1,8] / get_average_conf(), and
    update_T_WCs doing `self.T_WC[idx] = T_WCs.data` (:362-364) — shapes exactly as in the reference."""

    class _Pose:
        def __init__(self, data):
            self.data = data

    class _Kf:
        def __init__(self, X, C, T):
            self.X_canon, self._C, self.T_WC = X, C, _SharedKeyframesStub._Pose(T)

        def get_average_conf(self):
            return self._C

    def __init__(self, Xs, Cs, Twc, buffer=8):
        self.X, self.C = Xs, Cs
        self.T_WC = torch.zeros((buffer, 1, 8), dtype=torch.float32, device=Xs.device)
        self.T_WC[:Twc.shape[0], 0] = Twc

    def __getitem__(self, idx):
        k = int(idx)
        return self._Kf(self.X[k], self.C[k], self.T_WC[k])

    def update_T_WCs(self, T_WCs, idx):
        self.T_WC[idx] = T_WCs.data


def test_adopts_shared_keyframes_shaped_frames_and_empty_attributes(fgm, synth):
    """the non-`gather` frames branch (the reference's SharedKeyframes) with more than two keyframes, and the
    reference's attribute names before the first add_factors (empty tensors, global_opt.py:17-25)."""
    need_gpu()
    ops = pkg("backends")
    g = synth.make_graph_exact(5, 7, 2000, seed=81, noise=1e-3)
    g = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in g.items()}
    empty = fgm.FactorGraph(frames=None, device="cuda")
    for name in ("ii", "jj", "idx_ii2jj", "idx_jj2ii", "valid_match_j", "valid_match_i", "Q_ii2jj", "Q_jj2ii"):
        assert getattr(empty, name).numel() == 0
    assert empty.num_edges == 0 and empty.get_unique_kf_idx().numel() == 0
    frames = _SharedKeyframesStub(g["Xs"], g["Cs"], g["Twc"])
    before = frames.T_WC.clone()
    fg = _store_from_graph(fgm, g, frames)
    fg.solve_GN_rays()
    Twc_ref = g["Twc"].clone()
    ii, jj, idx, valid, Q = fg.prep_two_way_edges()
    c = fg.cfg
    ops.gauss_newton_rays(Twc_ref, g["Xs"], g["Cs"], ii, jj, idx, valid, Q, c["sigma_ray"], c["sigma_dist"],
                          c["C_conf"], c["Q_conf"], c["max_iters"], 0.0)
    assert torch.equal(frames.T_WC[1:5, 0], Twc_ref[1:])       # k = 4 > 1 poses written back through [k,1,8]
    assert torch.equal(frames.T_WC[0], before[0]) and torch.equal(frames.T_WC[5:], before[5:])
    assert not torch.equal(frames.T_WC[1:5], before[1:5])
