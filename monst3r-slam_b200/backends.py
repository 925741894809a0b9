"""The five ops of the reference extension (mast3r_slam/backend/src/gn.cpp:116-123), same names and
positional signatures, implemented by calling the C ABI (abi.py) — i.e. exactly what the pybind
module `mast3r_slam_backends` does, but without a compiled torch shim in between.  Used by the
parity tests so that they exercise the C ABI and read like calls to the reference.

Differences from the reference, all supersets: inputs must be CUDA tensors of the
documented dtype (the reference would crash in packed_accessor32), the current torch stream is used.
"""
import torch

from . import abi


def iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh):
    """gn.cpp:84-99 -> [p_new f32 [b,n,2], converged bool [b,n]]"""
    return abi.iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh)


def refine_matches(D11, D21, p1, radius, dilation_max):
    """gn.cpp:101-114 -> [p1_new i64 [b,n,2]]"""
    return abi.refine_matches(D11, D21, p1, radius, dilation_max)


def _gn(params, Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, shard=None):
    if Xs.shape[0] < 2 or ii.shape[0] < 1:
        return [torch.zeros((max(Xs.shape[0] - 1, 0), 7), dtype=torch.float32, device=Twc.device)]
    dx, info = abi.gauss_newton(params, Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, shard=shard)
    flags = info.cpu()
    if int(flags[3]) != 0:
        raise RuntimeError("gauss_newton: keyframe ids in ii/jj must lie in [0, 2^20) and have exactly "
                           f"Xs.shape[0] distinct values (flag {int(flags[3])})")
    _gn.last_info = [int(x) for x in flags]
    return [dx]


_gn.last_info = [0, 0, 0, 0]


def last_info():
    """[iterations executed, cholesky failures, stopped early, bad-index flag] of the last GN call"""
    return list(_gn.last_info)


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_point, C_thresh, Q_thresh,
                        max_iter, delta_thresh, shard=None):
    """gn.cpp:3-26; Twc [N,8] is updated in place; -> [dx f32 [N-1,7]]"""
    p = abi.make_params(abi.GN_POINTS, sigma_point, 1.0, C_thresh, Q_thresh, max_iter, delta_thresh)
    return _gn(p, Twc, Xs, Cs, None, ii, jj, idx_ii2jj, valid_match, Q, shard)


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_ray, sigma_dist, C_thresh,
                      Q_thresh, max_iter, delta_thresh, shard=None):
    """gn.cpp:28-52"""
    p = abi.make_params(abi.GN_RAYS, sigma_ray, sigma_dist, C_thresh, Q_thresh, max_iter, delta_thresh)
    return _gn(p, Twc, Xs, Cs, None, ii, jj, idx_ii2jj, valid_match, Q, shard)


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, height, width, pixel_border,
                       z_eps, sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter, delta_thresh,
                       shard=None):
    """gn.cpp:54-82"""
    p = abi.make_params(abi.GN_CALIB, sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter, delta_thresh,
                        height, width, pixel_border, z_eps)
    return _gn(p, Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, shard)
