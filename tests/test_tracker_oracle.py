"""Pins oracle/tracker_oracle.py (restatement of mast3r_slam/tracker.py:141-270) through the Gauss-Newton oracle,
which is itself pinned bit-for-bit / to 1e-7 against the reference's CUDA build (tests/test_oracle_pin.py):
the tracker's problem is the two-keyframe, one-directed-edge graph {k: fixed identity, f: T_CkCf} with the frame
points pre-gathered, so with the cost test disabled both oracles must walk the same Gauss-Newton path.  CPU only."""
import numpy as np
import pytest

from conftest import pkg, pose_diff


@pytest.fixture(scope="module")
def trk():
    from oracle import tracker_oracle
    return tracker_oracle


def tracking_problem(synth, n=3000, seed=91, noise=1e-3, calib=False, h=None, w=None):
    """keyframe = kf 0, frame = kf 1 of a 2-keyframe synthetic graph; Xf gathered by idx_f2k"""
    if calib:
        g = synth.make_graph(2, 1, h, w, seed=seed, calib=True)
    else:
        g = synth.make_graph_exact(2, 1, n, seed=seed, noise=noise)
    e = 1  # directed edge ii=1 (frame, gathered side), jj=0 (keyframe): idx[k] = frame pixel matched to keyframe pixel k
    assert int(g["ii"][e]) == 1 and int(g["jj"][e]) == 0
    idx = g["idx"][e].numpy()
    P = dict(Xf=g["Xs"][1].numpy()[idx], Xk=g["Xs"][0].numpy(), T_WCf=g["Twc"][1].numpy(), T_WCk=g["Twc"][0].numpy(),
             Q=g["Q"][e].numpy().reshape(-1), valid=g["valid"][e].numpy().reshape(-1) & (g["Q"][e].numpy().reshape(-1) > 1.5),
             K=g["K"].numpy(), T_gt=g["Twc_gt"][1].numpy(), h=g["h"], w=g["w"])
    return P


def gn_equivalent(oracle, P, kind, sig_a, sig_b, iters, delta, **kw):
    """the same problem as a gauss_newton_* call: Xs = [Xk, Xf_gathered], identity index"""
    n = P["Xk"].shape[0]
    Xs = np.stack([P["Xk"], P["Xf"]]).astype(np.float32)
    Cs = np.ones((2, n, 1), np.float32)
    T_rel0 = None
    from oracle import tracker_oracle as t
    T_rel0 = t.sim3_mul(t.sim3_inv(P["T_WCk"]), P["T_WCf"]).astype(np.float32)
    Twc = np.stack([np.array([0, 0, 0, 0, 0, 0, 1, 1], np.float32), T_rel0])
    args = (Twc, Xs, Cs) + ((P["K"],) if kind == "calib" else ()) + (
        np.array([0]), np.array([1]), np.arange(n)[None], P["valid"][None, :, None], P["Q"][None, :, None].astype(np.float32))
    if kind == "rays":
        out = oracle.gauss_newton_rays(*args, sig_a, sig_b, -1e30, -1e30, iters, delta)
    else:
        out = oracle.gauss_newton_calib(*args, P["h"], P["w"], kw["border"], kw["z_eps"], sig_a, sig_b, -1e30, -1e30, iters,
                                        delta)
    return out[0][1], out[2]


def test_rays_matches_pinned_gn_oracle(trk, oracle):
    P = tracking_problem(pkg("synth"))
    for iters in (1, 4):
        T_WCf, T_rel, steps = trk.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Q"], P["valid"], 0.003,
                                                         10.0, iters, rel_error=0.0, delta_norm=0.0)
        T_gn, it_gn = gn_equivalent(oracle, P, "rays", 0.003, 10.0, iters, 0.0)
        assert steps == iters == it_gn
        rel, ang = pose_diff(T_rel[None], T_gn[None])
        assert rel < 2e-6 and ang < 2e-6, (iters, rel, ang)


def test_calib_matches_pinned_gn_oracle(trk, oracle):
    P = tracking_problem(pkg("synth"), calib=True, h=48, w=64, seed=92)
    T_WCf, T_rel, steps = trk.opt_pose_calib_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Q"], P["valid"], P["K"],
                                                  (P["h"], P["w"]), 1.0, 10.0, -10, 1e-6, 4, rel_error=0.0, delta_norm=0.0)
    T_gn, _ = gn_equivalent(oracle, P, "calib", 1.0, 10.0, 4, 0.0, border=-10, z_eps=1e-6)
    rel, ang = pose_diff(T_rel[None], T_gn[None])
    assert rel < 2e-6 and ang < 2e-6, (rel, ang)


def test_converges_to_ground_truth_and_stops_on_cost(trk):
    P = tracking_problem(pkg("synth"), noise=0.0)
    T_WCf, T_rel, steps = trk.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Q"], P["valid"], 0.003,
                                                     10.0, 50, rel_error=1e-3, delta_norm=1e-3)
    assert 2 <= steps < 50  # base.yaml tracking thresholds stop long before max_iters
    rel, ang = pose_diff(T_WCf[None], P["T_gt"][None])
    assert rel < 1e-3 and ang < 1e-3
    # group identities used by the mirror
    I = trk.sim3_mul(trk.sim3_inv(P["T_WCf"]), P["T_WCf"])
    assert np.allclose(I, [0, 0, 0, 0, 0, 0, 1, 1], atol=1e-12)
    X = np.random.default_rng(0).normal(size=(5, 3))
    assert np.allclose(trk.sim3_act(trk.sim3_mul(P["T_WCk"], T_rel), X), trk.sim3_act(P["T_WCk"], trk.sim3_act(T_rel, X)))


def test_cholesky_failure_raises_like_torch(trk):
    P = tracking_problem(pkg("synth"))
    with pytest.raises(np.linalg.LinAlgError):
        trk.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Q"], np.zeros_like(P["valid"]), 0.003, 10.0,
                                   5, 1e-3, 1e-3)


# ---- pinned to the REFERENCE's python: tests/golden/tracker_{rays,calib}.npz were produced by running the reference's own
# FrameTracker.opt_pose_* / solve / check_convergence / geometry code (tests/golden/make_golden_tracker.py; only lietorch's
# Sim3 container is a stand-in, its retraction being the bit-pinned orc.pose_retr).  The reference runs in float32, the
# oracle in float64: per-iteration costs agree to 1e-4 relative, poses to 2e-5, iteration counts exactly.
def _golden(kind):
    import os
    from conftest import GOLDEN
    return np.load(os.path.join(GOLDEN, f"tracker_{kind}.npz"))


def test_rays_matches_reference_python_fixture(trk):
    d = _golden("rays")
    T_WCf, T_rel, steps = trk.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Q"].reshape(-1), d["valid"].reshape(-1),
                                                     float(d["cfg_sigma_ray"]), float(d["cfg_sigma_dist"]), int(d["cfg_max_iters"]),
                                                     rel_error=float(d["cfg_rel_error"]), delta_norm=float(d["cfg_delta_norm"]))
    assert steps == int(d["ref_iterations"]) and steps >= 2
    for T, ref in ((T_WCf, d["ref_T_WCf"]), (T_rel, d["ref_T_CkCf"])):
        rel, ang = pose_diff(np.asarray(T)[None], ref[None])
        assert rel < 2e-5 and ang < 2e-5, (rel, ang)


def test_calib_matches_reference_python_fixture(trk):
    d = _golden("calib")
    T_WCf, T_rel, steps = trk.opt_pose_calib_sim3(d["Xf_on_ray"], d["Xk_on_ray"], d["T_WCf"], d["T_WCk"], d["Q"].reshape(-1),
                                                  d["valid"].reshape(-1), d["K"], (int(d["h"]), int(d["w"])),
                                                  float(d["cfg_sigma_pixel"]), float(d["cfg_sigma_depth"]), int(d["cfg_pixel_border"]),
                                                  float(d["cfg_depth_eps"]), int(d["cfg_max_iters"]),
                                                  rel_error=float(d["cfg_rel_error"]), delta_norm=float(d["cfg_delta_norm"]))
    assert steps == int(d["ref_iterations"]) and steps >= 2
    for T, ref in ((T_WCf, d["ref_T_WCf"]), (T_rel, d["ref_T_CkCf"])):
        rel, ang = pose_diff(np.asarray(T)[None], ref[None])
        assert rel < 2e-5 and ang < 2e-5, (rel, ang)


def test_solve_costs_match_reference_python_fixture(trk):
    """first-iteration cost and update of the reference's FrameTracker.solve (tracker.py:141-154) on the fixture inputs"""
    d = _golden("rays")
    Xf, Xk = d["Xf"].astype(np.float64), d["Xk"].astype(np.float64)
    T = trk.sim3_mul(trk.sim3_inv(d["T_WCk"]), d["T_WCf"])
    valid, Q = d["valid"].reshape(-1, 1).astype(np.float64), d["Q"].reshape(-1, 1).astype(np.float64)
    si = np.concatenate([np.repeat(valid * np.sqrt(Q) / float(d["cfg_sigma_ray"]), 3, 1), valid * np.sqrt(Q) / float(d["cfg_sigma_dist"])], 1)
    pW, dX = trk.act_sim3_jac(T, Xf)
    rd_f, drd = trk.point_to_ray_dist(pW, jacobian=True)
    r = trk.point_to_ray_dist(Xk) - rd_f
    tau, cost = trk.solve(si, r, -drd @ dX)
    assert abs(cost - d["ref_costs"][0]) <= 1e-4 * d["ref_costs"][0]
    assert np.abs(np.asarray(tau).reshape(7) - d["ref_taus"][0]).max() <= 2e-5 * max(1.0, np.abs(d["ref_taus"][0]).max())
