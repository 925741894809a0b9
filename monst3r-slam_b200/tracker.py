"""Frame-tracker pose optimisation on the B200 backend: SURVEY.md §8(f) rank 3.

Mirror of `FrameTracker.opt_pose_ray_dist_sim3` / `opt_pose_calib_sim3` (mast3r_slam/tracker.py:156-270): same
names, argument order and return values `(T_WCf, T_CkCf)`, but the whole loop (up to `max_iters` Gauss-Newton
iterations with both stop tests of nonlinear_optimizer.py:5-26) is ONE cooperative kernel launch
(`m3r_track_frame`, csrc/frame_track.cu) instead of ~30 torch kernels and one `.item()` sync per iteration.

Poses are Sim(3) data tensors `[t(3), q_xyzw(4), s(1)]` of shape [8] or [1,8] (what `lietorch.Sim3.data` holds);
objects with a `.data` attribute are accepted.  A Cholesky failure raises, like `torch.linalg.cholesky` does in
the reference (caught by `FrameTracker.track`, tracker.py:92-94).  There is no CPU path.
"""
import torch

from . import abi

DEFAULT_CFG = {  # config/base.yaml:23-40 (tracking)
    "max_iters": 50, "C_conf": 0.0, "Q_conf": 1.5, "rel_error": 1e-3, "delta_norm": 1e-3, "huber": 1.345,
    "sigma_ray": 0.003, "sigma_dist": 10.0, "sigma_pixel": 1.0, "sigma_depth": 10.0, "pixel_border": -10,
    "depth_eps": 1e-6,
}


def _pose(T):
    t = getattr(T, "data", T)
    return t.reshape(-1)[:8].to(torch.float32).contiguous()


class PoseOptimizer:
    """holds the config and a reusable device workspace (one per image size)"""

    def __init__(self, cfg=None, sync_check=True):
        self.cfg = dict(DEFAULT_CFG)
        if cfg:
            self.cfg.update(cfg)
        self.sync_check = sync_check  # read info back (one sync per frame) and raise on a Cholesky failure
        self._ws = {}
        self.last_info = None

    def _run(self, kind, Xf, Xk, T_WCf, T_WCk, Qk, valid, K, sig_a, sig_b, h=0, w=0):
        c = self.cfg
        n = Xk.shape[0]
        prm = abi.TrackParams(kind, sig_a, sig_b, int(c["max_iters"]), c["rel_error"], c["delta_norm"], c["huber"], int(h),
                              int(w), int(c["pixel_border"]), c["depth_eps"])
        key = (n, Xk.device)
        if key not in self._ws:
            self._ws[key] = torch.empty(int(abi.lib.m3r_track_workspace_bytes(n)), dtype=torch.uint8, device=Xk.device)
        T_new = _pose(T_WCf).clone()
        T_rel, dx, info = abi.track_frame(T_new, _pose(T_WCk), Xf.contiguous(), Xk.contiguous(),
                                          valid.reshape(-1).contiguous(), Qk.reshape(-1).contiguous(), K, prm,
                                          workspace=self._ws[key])
        self.last_info = info
        if self.sync_check:
            it, fails, early, _ = (int(v) for v in info.cpu())
            if fails:
                raise RuntimeError("track_frame: Cholesky failed (normal matrix not positive definite)")
        return T_new.view(1, 8), T_rel.view(1, 8)

    def opt_pose_ray_dist_sim3(self, Xf, Xk, T_WCf, T_WCk, Qk, valid):
        """tracker.py:156-214.  Xf [n,3]: frame points already gathered by idx_f2k; valid [n,1] bool; Qk [n,1]"""
        c = self.cfg
        return self._run(abi.GN_RAYS, Xf, Xk, T_WCf, T_WCk, Qk, valid, None, c["sigma_ray"], c["sigma_dist"])

    def opt_pose_calib_sim3(self, Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size):
        """tracker.py:216-270.  meas_k / valid_meas_k (pixel grid, log depth of Xk, Xk.z > depth_eps, tracker.py:128-137)
        are recomputed in the kernel from Xk; the arguments are accepted for signature parity and ignored."""
        c = self.cfg
        h, w = int(img_size[0]), int(img_size[1])
        return self._run(abi.GN_CALIB, Xf, Xk, T_WCf, T_WCk, Qk, valid, K.contiguous(), c["sigma_pixel"], c["sigma_depth"], h, w)
