"""Generates tests/golden/tracker_{rays,calib}.npz by running the REFERENCE's own python:
`FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 / solve` (mast3r_slam/tracker.py:141-270),
`check_convergence / huber` (nonlinear_optimizer.py:5-34) and `act_Sim3 / point_to_ray_dist / project_calib /
constrain_points_to_ray / get_pixel_coords` (geometry.py) imported from /root/reference, in float32 on the CPU.

What is stubbed, and only that:
  * `lietorch` — absent from /root/reference and from the image (unpinned git dependency, pyproject.toml:15).  The stand-in
    `Sim3` below implements exactly the five members the tracker touches: `.data` ([1,8] = t, q_xyzw, s), `inv()`,
    `__mul__`, `act(X)` = s R X + t, and `retr(tau)` = Exp(tau) * T.  inv / mul / act are the textbook Sim(3) group
    operations written with torch float32 ops; `retr` calls the oracle's host retraction `orc.pose_retr`, which is pinned
    BIT FOR BIT to the reference's own `pose_retr_kernel` / `expSim3` (gn_kernels.cu:299-453, lietorch's closed form) by
    tests/test_oracle_pin.py and the golden GN fixtures.
  * `mast3r_slam.frame`, `mast3r_slam.mast3r_utils` (ViT side; not on this path), `mast3r_slam.config.config["tracking"]`
    filled from config/base.yaml:23-38.
Everything else that runs is the reference's unmodified code.  Runs in the build container only:
    python tests/golden/make_golden_tracker.py
"""
import importlib
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/MASt3R-SLAM"
sys.path.insert(0, ROOT)

TRACKING_CFG = {  # config/base.yaml:23-38
    "min_match_frac": 0.05, "max_iters": 50, "C_conf": 0.0, "Q_conf": 1.5, "rel_error": 1e-3, "delta_norm": 1e-3,
    "huber": 1.345, "match_frac_thresh": 0.333, "sigma_ray": 0.003, "sigma_dist": 1e+1, "sigma_pixel": 1.0,
    "sigma_depth": 1e+1, "sigma_point": 0.05, "pixel_border": -10, "depth_eps": 1e-6,
}


def _quat_R(q):
    x, y, z, w = q.unbind(-1)
    return torch.stack((1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                        2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                        2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)), dim=-1).view(3, 3)


class Sim3:
    """Minimal lietorch.Sim3 stand-in (see the module docstring).  data: [1,8] float32."""
    embedded_dim = 8

    def __init__(self, data):
        self.data = data.reshape(1, 8).to(torch.float32)

    def _parts(self):
        d = self.data[0]
        return d[0:3], d[3:7], d[7]

    def inv(self):
        t, q, s = self._parts()
        qi = q * torch.tensor([-1.0, -1.0, -1.0, 1.0])
        return Sim3(torch.cat((-(_quat_R(qi) @ t) / s, qi, (1.0 / s).reshape(1))))

    def __mul__(self, other):
        t1, q1, s1 = self._parts()
        t2, q2, s2 = other._parts()
        x1, y1, z1, w1 = q1
        x2, y2, z2, w2 = q2
        q = torch.stack((w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                         w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2))
        return Sim3(torch.cat((s1 * (_quat_R(q1) @ t2) + t1, q, (s1 * s2).reshape(1))))

    def act(self, X):
        t, q, s = self._parts()
        return s * (X @ _quat_R(q).T) + t

    def retr(self, tau):
        from oracle import oracle as orc
        T = np.concatenate([np.array([[0, 0, 0, 0, 0, 0, 1, 1]], np.float32), self.data.numpy().astype(np.float32)])
        out = orc.pose_retr(T, tau.detach().numpy().astype(np.float32).reshape(1, 7))
        return Sim3(torch.from_numpy(np.asarray(out[1], np.float32)))


def import_reference():
    sys.path.insert(0, REF)
    lt = types.ModuleType("lietorch")
    lt.Sim3 = Sim3
    sys.modules["lietorch"] = lt
    fr = types.ModuleType("mast3r_slam.frame")
    fr.Frame = object
    sys.modules["mast3r_slam.frame"] = fr
    mu = types.ModuleType("mast3r_slam.mast3r_utils")
    mu.mast3r_match_asymmetric = None
    sys.modules["mast3r_slam.mast3r_utils"] = mu
    cfg = importlib.import_module("mast3r_slam.config")
    cfg.config["tracking"] = dict(TRACKING_CFG)
    geo = importlib.import_module("mast3r_slam.geometry")
    trk = importlib.import_module("mast3r_slam.tracker")
    return cfg, geo, trk


def make_tracker(trk, log):
    t = trk.FrameTracker.__new__(trk.FrameTracker)
    t.cfg = dict(TRACKING_CFG)
    inner = trk.FrameTracker.solve

    def solve(self, sqrt_info, r, J):  # the reference's solve, with the cost / update recorded
        tau, cost = inner(self, sqrt_info, r, J)
        log.append((float(cost), tau.detach().numpy().reshape(7).copy()))
        return tau, cost

    t.solve = types.MethodType(solve, t)
    return t


def problem(synth, calib, seed):
    """keyframe = kf 0, frame = kf 1 of a 2-keyframe synthetic graph; Xf gathered by idx_f2k (tracker.py:139)"""
    if calib:
        g = synth.make_graph(2, 1, 48, 64, seed=seed, calib=True, perturb=(0.05, 0.03, 0.03))
    else:
        g = synth.make_graph_exact(2, 1, 3000, seed=seed, noise=2e-3, perturb=(0.05, 0.03, 0.03))
    e = 1  # directed edge ii = 1 (frame, gathered side), jj = 0 (keyframe)
    assert int(g["ii"][e]) == 1 and int(g["jj"][e]) == 0
    idx = g["idx"][e]
    Q = g["Q"][e].reshape(-1, 1).float()
    valid = g["valid"][e].reshape(-1, 1) & (Q > TRACKING_CFG["Q_conf"])  # tracker.py:62-67 (C_conf = 0)
    return dict(Xf=g["Xs"][1][idx].float().contiguous(), Xf_full=g["Xs"][1].float().contiguous(), idx_f2k=idx,
                Xk=g["Xs"][0].float().contiguous(), T_WCf=g["Twc"][1].float(),
                T_WCk=g["Twc"][0].float(), Q=Q, valid=valid, K=g["K"].float(), h=g["h"], w=g["w"], T_gt=g["Twc_gt"][1].float())


def main():
    cfg, geo, trk = import_reference()
    synth = importlib.import_module("monst3r-slam_b200.synth")
    torch.manual_seed(0)
    for kind, seed in (("rays", 91), ("calib", 92)):
        P = problem(synth, kind == "calib", seed)
        log = []
        tracker = make_tracker(trk, log)
        T_WCf, T_WCk = Sim3(P["T_WCf"]), Sim3(P["T_WCk"])
        out = {k: (v.numpy() if torch.is_tensor(v) else v) for k, v in P.items() if v is not None}
        if kind == "rays":
            T_new, T_rel = tracker.opt_pose_ray_dist_sim3(P["Xf"], P["Xk"], T_WCf, T_WCk, P["Q"], P["valid"])
        else:
            # tracker.py:116-139: both maps are snapped to their own pixel rays, the measurement is the keyframe's pixel
            # grid + log depth, and only then is the frame map gathered by idx_f2k
            img_size = (P["h"], P["w"])
            Xf = geo.constrain_points_to_ray(img_size, P["Xf_full"][None], P["K"]).squeeze(0)[P["idx_f2k"]]
            Xk = geo.constrain_points_to_ray(img_size, P["Xk"][None], P["K"]).squeeze(0)
            uv_k = geo.get_pixel_coords(1, img_size, device=Xf.device, dtype=Xf.dtype).view(-1, 2)
            meas_k = torch.cat((uv_k, torch.log(Xk[..., 2:3])), dim=-1)
            valid_meas_k = Xk[..., 2:3] > TRACKING_CFG["depth_eps"]
            meas_k[~valid_meas_k.repeat(1, 3)] = 0.0
            out["Xf_on_ray"], out["Xk_on_ray"] = Xf.numpy(), Xk.numpy()
            T_new, T_rel = tracker.opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, P["Q"], P["valid"], meas_k, valid_meas_k,
                                                       P["K"], img_size)
        out["ref_T_WCf"] = T_new.data.numpy().reshape(8)
        out["ref_T_CkCf"] = T_rel.data.numpy().reshape(8)
        out["ref_iterations"] = len(log)
        out["ref_costs"] = np.asarray([c for c, _ in log], np.float64)
        out["ref_taus"] = np.stack([t for _, t in log]).astype(np.float32)
        for k, v in TRACKING_CFG.items():
            out["cfg_" + k] = v
        path = os.path.join(HERE, f"tracker_{kind}.npz")
        np.savez_compressed(path, **out)
        err0 = float(np.abs(P["T_WCf"].numpy() - P["T_gt"].numpy()).max())
        err1 = float(np.abs(out["ref_T_WCf"] - P["T_gt"].numpy()).max())
        print(f"{kind}: {len(log)} iterations, cost {out['ref_costs'][0]:.4g} -> {out['ref_costs'][-1]:.4g}, "
              f"|T - gt| {err0:.3e} -> {err1:.3e}, {os.path.getsize(path) / 1024:.0f} KB -> {path}")


if __name__ == "__main__":
    main()
