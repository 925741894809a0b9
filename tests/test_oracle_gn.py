"""Analytic known-answer tests of the GN oracle (SURVEY §8c iii-v) — CPU only."""
import numpy as np
import pytest
import torch

from conftest import np_, pkg, pose_diff

RAYS = dict(sigma_a=0.003, sigma_b=10.0, C_thresh=0.0, Q_thresh=1.5)
CALIB = dict(sigma_a=1.0, sigma_b=10.0, C_thresh=0.0, Q_thresh=1.5, height=96, width=128, pixel_border=-10, z_eps=1e-6)
POINTS = dict(sigma_a=0.05, sigma_b=0.0, C_thresh=0.0, Q_thresh=1.5)


def _graph(synth, N=4, E=5, h=24, w=32, seed=3, calib=False):
    if calib:  # pixel targets need a real image grid: height-field scene at moderate resolution
        g = synth.make_graph(N, E, 96, 128, seed=seed, calib=True)
    else:      # exact correspondences through random permutations (+1e-3 noise)
        g = synth.make_graph_exact(N, E, h * w, seed=seed, noise=1e-3)
    return {k: (np_(v) if torch.is_tensor(v) else v) for k, v in g.items()}


@pytest.mark.parametrize("kind,prm", [("rays", RAYS), ("points", POINTS), ("calib", CALIB)])
def test_gn_recovers_ground_truth(oracle, synth, kind, prm):
    g = _graph(synth, calib=(kind == "calib"))
    K = g["K"] if kind == "calib" else None
    err0 = pose_diff(g["Twc"], g["Twc_gt"])
    Twc, dx, iters, nfail = oracle.gauss_newton(kind, g["Twc"], g["Xs"], g["Cs"], K, g["ii"], g["jj"], g["idx"],
                                                g["valid"], g["Q"], max_iter=10, delta_thresh=0.0, **prm)
    err1 = pose_diff(Twc, g["Twc_gt"])
    assert iters == 10 and nfail == 0
    assert err1[0] < 0.3 * err0[0] and err1[1] < 0.3 * err0[1], (err0, err1)
    assert np.allclose(Twc[0], g["Twc"][0])  # pose 0 is fixed


def test_block_structure_justifies_restructure(oracle, synth):
    """H_ii = H_jj = -H_ij = -H_ji^T and g_i = -g_j up to fp32 summation noise (Appendix A)."""
    g = _graph(synth)
    ii_c, jj_c, nu = oracle.compact_indices(g["ii"], g["jj"])
    assert nu == g["Xs"].shape[0]
    Hs, gs = oracle.gn_accumulate("rays", g["Twc"], g["Xs"], g["Cs"], None, ii_c, jj_c, g["idx"], g["valid"],
                                  g["Q"], **RAYS)
    scale = np.abs(Hs[0]).max()
    assert np.abs(Hs[0] - Hs[3]).max() < 1e-4 * scale
    assert np.abs(Hs[0] + Hs[1]).max() < 1e-4 * scale
    assert np.abs(Hs[1] - np.transpose(Hs[2], (0, 2, 1))).max() < 1e-4 * scale
    assert np.abs(gs[0] + gs[1]).max() < 1e-4 * np.abs(gs).max()


def test_early_stop_and_zero_iter(oracle, synth):
    g = _graph(synth)
    Twc, dx, iters, _ = oracle.gauss_newton("rays", g["Twc"], g["Xs"], g["Cs"], None, g["ii"], g["jj"], g["idx"],
                                            g["valid"], g["Q"], max_iter=10, delta_thresh=1e9, **RAYS)
    assert iters == 1  # the test happens AFTER the first update was applied
    assert not np.allclose(Twc, g["Twc"])
    Twc0, dx0, iters0, _ = oracle.gauss_newton("rays", g["Twc"], g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                                               g["idx"], g["valid"], g["Q"], max_iter=0, delta_thresh=0.0, **RAYS)
    assert iters0 == 0 and np.array_equal(Twc0, g["Twc"])


def test_non_spd_gives_zero_dx(oracle, synth):
    """all matches invalid -> H = 0 -> Cholesky fails -> dx = 0, poses untouched (gn_kernels.cu:147-150)"""
    g = _graph(synth)
    Twc, dx, iters, nfail = oracle.gauss_newton("rays", g["Twc"], g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                                                g["idx"], np.zeros_like(g["valid"]), g["Q"], max_iter=3,
                                                delta_thresh=0.0, **RAYS)
    assert nfail == 3 and iters == 3
    assert np.array_equal(dx, np.zeros_like(dx))
    assert np.array_equal(Twc, g["Twc"])


def test_exp_and_retraction(oracle):
    rng = np.random.default_rng(0)
    assert np.allclose(oracle.sim3_exp(np.zeros(7)), [0, 0, 0, 0, 0, 0, 1, 1])
    for scale in (1e-4, 1e-2, 0.3):
        xi = (rng.standard_normal(7) * scale).astype(np.float32)
        T = oracle.sim3_exp(xi)
        assert abs(np.linalg.norm(T[3:7]) - 1) < 1e-6
        assert abs(T[7] - np.exp(xi[6])) < 1e-6 * max(1, T[7])
        # exp(xi) * exp(-xi) = identity
        P = oracle.pose_retr(np.stack([np.zeros(8, np.float32), oracle.sim3_exp(-xi)]), xi[None], num_fix=1)
        assert np.allclose(P[1], [0, 0, 0, 0, 0, 0, 1, 1], atol=5e-6)


def test_solver_matches_numpy(oracle):
    rng = np.random.default_rng(1)
    E, nf = 6, 3
    ii_o = np.array([-1, 0, 1, 0, 1, 2]); jj_o = np.array([0, 1, 2, -1, 0, 1])
    A = rng.standard_normal((E, 7, 7)).astype(np.float32)
    A = A @ np.transpose(A, (0, 2, 1)) + 7 * np.eye(7, dtype=np.float32)
    b = rng.standard_normal((E, 7)).astype(np.float32)
    Hs = np.stack([A, -A, -A, A]); gs = np.stack([-b, b])
    dx, failed = oracle.gn_solve(Hs, gs, ii_o, jj_o, nf)
    H = np.zeros((21, 21)); g = np.zeros(21)
    for e in range(E):
        i, j = ii_o[e], jj_o[e]
        for (r, c, s) in ((i, i, 1), (i, j, -1), (j, i, -1), (j, j, 1)):
            if r >= 0 and c >= 0:
                H[7 * r:7 * r + 7, 7 * c:7 * c + 7] += s * A[e].astype(np.float64)
        if i >= 0: g[7 * i:7 * i + 7] -= b[e]
        if j >= 0: g[7 * j:7 * j + 7] += b[e]
    want = -np.linalg.solve(H, g)
    assert not failed and np.allclose(dx.reshape(-1), want, rtol=1e-5, atol=1e-6)
