"""GPU parity of the frame-tracker pose optimisation (m3r_track_frame through the host mirror
monst3r-slam_b200/tracker.py) against oracle/tracker_oracle.py — the restatement of mast3r_slam/tracker.py:141-270 —
and against gauss_newton_* on the equivalent two-keyframe graph.  Bar: the GN pose tolerance of BASELINE.json
(1e-5 relative, 1e-4 rad) and the SAME number of iterations as the reference's stop tests give."""
import numpy as np
import pytest
import torch

from conftest import need_gpu, np_, pkg, pose_diff
from test_tracker_oracle import tracking_problem

pytestmark = pytest.mark.gpu
REL_TOL, ANG_TOL = 1e-5, 1e-4


@pytest.fixture(scope="module")
def trk():
    from oracle import tracker_oracle
    return tracker_oracle


def _cu(P):
    t = lambda a, dt=None: torch.from_numpy(np.ascontiguousarray(a)).to("cuda", dtype=dt)
    return dict(Xf=t(P["Xf"], torch.float32), Xk=t(P["Xk"], torch.float32), T_WCf=t(P["T_WCf"], torch.float32).view(1, 8),
                T_WCk=t(P["T_WCk"], torch.float32).view(1, 8), Q=t(P["Q"], torch.float32).view(-1, 1),
                valid=t(P["valid"]).view(-1, 1), K=t(P["K"], torch.float32))


def _check(T_a, T_b, what):
    rel, ang = pose_diff(np.asarray(T_a, np.float64).reshape(1, 8), np.asarray(T_b, np.float64).reshape(1, 8))
    assert rel < REL_TOL and ang < ANG_TOL, (what, rel, ang)


@pytest.mark.parametrize("n,noise,seed", [(3000, 1e-3, 91), (4096, 0.0, 93), (12288, 5e-3, 94)])
def test_rays_default_thresholds_vs_oracle(trk, n, noise, seed):
    need_gpu()
    P = tracking_problem(pkg("synth"), n=n, seed=seed, noise=noise)
    opt = pkg("tracker").PoseOptimizer()
    c = opt.cfg
    T_ref, Trel_ref, steps = trk.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Q"], P["valid"],
                                                        c["sigma_ray"], c["sigma_dist"], c["max_iters"], c["rel_error"],
                                                        c["delta_norm"])
    d = _cu(P)
    T_WCf, T_CkCf = opt.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Q"], d["valid"])
    it, fails, early, _ = (int(v) for v in opt.last_info.cpu())
    assert T_WCf.shape == (1, 8) and T_CkCf.shape == (1, 8)
    assert it == steps and fails == 0 and early == (1 if steps < c["max_iters"] else 0)
    _check(np_(T_CkCf), Trel_ref, "T_CkCf")
    _check(np_(T_WCf), T_ref, "T_WCf")
    assert torch.equal(d["T_WCf"], torch.from_numpy(P["T_WCf"]).cuda().view(1, 8))  # the input pose is not modified


def test_rays_full_resolution_vs_oracle(trk):
    """512x384: the configuration the persistent kernel is sized for (148 CTAs x 1329 points)"""
    need_gpu()
    synth = pkg("synth")
    g = synth.make_graph(2, 1, 384, 512, seed=1002, calib=False)
    e = 1
    assert int(g["ii"][e]) == 1 and int(g["jj"][e]) == 0
    idx = g["idx"][e].numpy()
    Q = g["Q"][e].numpy().reshape(-1)
    P = dict(Xf=g["Xs"][1].numpy()[idx], Xk=g["Xs"][0].numpy(), T_WCf=g["Twc"][1].numpy(), T_WCk=g["Twc"][0].numpy(), Q=Q,
             valid=g["valid"][e].numpy().reshape(-1) & (Q > 1.5), K=g["K"].numpy())
    opt = pkg("tracker").PoseOptimizer()
    c = opt.cfg
    T_ref, Trel_ref, steps = trk.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Q"], P["valid"],
                                                        c["sigma_ray"], c["sigma_dist"], c["max_iters"], c["rel_error"],
                                                        c["delta_norm"])
    d = _cu(P)
    T_WCf, T_CkCf = opt.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Q"], d["valid"])
    assert int(opt.last_info[0]) == steps
    _check(np_(T_WCf), T_ref, "T_WCf")
    # a second frame against a non-identity keyframe pose: T_WCk composes on both sides
    Tk = np.array([0.3, -0.2, 0.1, 0.0, np.sin(0.2), 0.0, np.cos(0.2), 1.3], np.float32)
    P2 = dict(P, T_WCk=Tk, T_WCf=trk.sim3_mul(Tk, trk.sim3_mul(trk.sim3_inv(P["T_WCk"]), P["T_WCf"])).astype(np.float32))
    T_ref2, Trel_ref2, steps2 = trk.opt_pose_ray_dist_sim3(P2["Xf"], P2["Xk"], P2["T_WCf"], P2["T_WCk"], P2["Q"],
                                                           P2["valid"], c["sigma_ray"], c["sigma_dist"], c["max_iters"],
                                                           c["rel_error"], c["delta_norm"])
    d2 = _cu(P2)
    T2, Trel2 = opt.opt_pose_ray_dist_sim3(d2["Xf"], d2["Xk"], d2["T_WCf"], d2["T_WCk"], d2["Q"], d2["valid"])
    assert int(opt.last_info[0]) == steps2 == steps
    _check(np_(T2), T_ref2, "T_WCf, non-identity keyframe")
    _check(np_(Trel2), np_(T_CkCf), "relative pose independent of the keyframe pose")


@pytest.mark.parametrize("h,w", [(48, 64), (96, 128)])
def test_calib_vs_oracle(trk, h, w):
    need_gpu()
    P = tracking_problem(pkg("synth"), calib=True, h=h, w=w, seed=95)
    opt = pkg("tracker").PoseOptimizer()
    c = opt.cfg
    T_ref, Trel_ref, steps = trk.opt_pose_calib_sim3(P["Xf"], P["Xk"], P["T_WCf"], P["T_WCk"], P["Q"], P["valid"], P["K"],
                                                     (h, w), c["sigma_pixel"], c["sigma_depth"], c["pixel_border"],
                                                     c["depth_eps"], c["max_iters"], c["rel_error"], c["delta_norm"])
    d = _cu(P)
    T_WCf, T_CkCf = opt.opt_pose_calib_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Q"], d["valid"], None, None, d["K"],
                                            (h, w))
    assert int(opt.last_info[0]) == steps
    _check(np_(T_CkCf), Trel_ref, "T_CkCf")
    _check(np_(T_WCf), T_ref, "T_WCf")


def test_equals_gauss_newton_rays_when_cost_test_is_off(trk):
    """rel_error = 0: the tracker mode is the GN op on {k: identity (fixed), f: T_CkCf} with one directed edge"""
    need_gpu()
    ops = pkg("backends")
    P = tracking_problem(pkg("synth"), n=4000, seed=96)
    d = _cu(P)
    opt = pkg("tracker").PoseOptimizer(cfg={"rel_error": 0.0, "delta_norm": 0.0, "max_iters": 6})
    T_WCf, T_CkCf = opt.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Q"], d["valid"])
    assert int(opt.last_info[0]) == 6
    n = d["Xk"].shape[0]
    T0 = torch.from_numpy(trk.sim3_mul(trk.sim3_inv(P["T_WCk"]), P["T_WCf"]).astype(np.float32)).cuda()
    Twc = torch.stack((torch.tensor([0, 0, 0, 0, 0, 0, 1, 1.0], device="cuda"), T0)).contiguous()
    Xs = torch.stack((d["Xk"], d["Xf"])).contiguous()
    Cs = torch.ones((2, n, 1), device="cuda")
    ops.gauss_newton_rays(Twc, Xs, Cs, torch.tensor([0], device="cuda"), torch.tensor([1], device="cuda"),
                          torch.arange(n, device="cuda").view(1, n), d["valid"].view(1, n, 1).contiguous(),
                          d["Q"].view(1, n, 1).contiguous(), 0.003, 10.0, -1e30, -1e30, 6, 0.0)
    _check(np_(T_CkCf), np_(Twc[1]), "vs gauss_newton_rays")
    assert float((T_CkCf.view(-1) - Twc[1]).abs().max()) < 1e-6


def test_cholesky_failure_raises_and_unsupported_sizes_fail_loudly():
    need_gpu()
    abi = pkg("abi")
    P = tracking_problem(pkg("synth"), n=3000, seed=97)
    d = _cu(P)
    opt = pkg("tracker").PoseOptimizer()
    with pytest.raises(RuntimeError, match="Cholesky"):
        opt.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], d["Q"], torch.zeros_like(d["valid"]))
    with pytest.raises(abi.M3RError):  # 3001 points: not a multiple of 8, no staged path -> error, not a slow fallback
        opt.opt_pose_ray_dist_sim3(torch.zeros((3001, 3), device="cuda"), torch.ones((3001, 3), device="cuda"), d["T_WCf"],
                                   d["T_WCk"], torch.ones((3001, 1), device="cuda"),
                                   torch.ones((3001, 1), dtype=torch.bool, device="cuda"))
    with pytest.raises(abi.M3RError):  # another Huber threshold than the built 1.345
        pkg("tracker").PoseOptimizer(cfg={"huber": 2.0}).opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"],
                                                                              d["Q"], d["valid"])


@pytest.mark.parametrize("kind", ["rays", "calib"])
def test_vs_reference_python_fixture(kind):
    """m3r_track_frame against tests/golden/tracker_{rays,calib}.npz: inputs, final poses and iteration count recorded
    from the reference's own FrameTracker.opt_pose_* python (tests/golden/make_golden_tracker.py; float32 torch on the
    CPU).  Same number of iterations, poses within the BASELINE pose tolerance."""
    need_gpu()
    import os
    from conftest import GOLDEN
    d = np.load(os.path.join(GOLDEN, f"tracker_{kind}.npz"))
    t = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to("cuda", dtype=dt)
    opt = pkg("tracker").PoseOptimizer()
    for k in ("max_iters", "rel_error", "delta_norm", "huber", "sigma_ray", "sigma_dist", "sigma_pixel", "sigma_depth",
              "pixel_border", "depth_eps"):
        assert float(opt.cfg[k]) == float(d["cfg_" + k]), k  # the mirror's defaults are config/base.yaml's
    T_WCf, T_WCk = t(d["T_WCf"]).view(1, 8), t(d["T_WCk"]).view(1, 8)
    Q, valid = t(d["Q"]).view(-1, 1), t(d["valid"], torch.bool).view(-1, 1)
    if kind == "rays":
        T_new, T_rel = opt.opt_pose_ray_dist_sim3(t(d["Xf"]), t(d["Xk"]), T_WCf, T_WCk, Q, valid)
    else:
        T_new, T_rel = opt.opt_pose_calib_sim3(t(d["Xf_on_ray"]), t(d["Xk_on_ray"]), T_WCf, T_WCk, Q, valid, None, None,
                                               t(d["K"]), (int(d["h"]), int(d["w"])))
    it, fails, early, _ = (int(v) for v in opt.last_info.cpu())
    print(kind, "iterations", it, "reference", int(d["ref_iterations"]))
    assert it == int(d["ref_iterations"]) and fails == 0 and early == 1
    _check(np_(T_rel), d["ref_T_CkCf"], "T_CkCf vs reference python")
    _check(np_(T_new), d["ref_T_WCf"], "T_WCf vs reference python")
