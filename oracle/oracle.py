"""TEST INFRASTRUCTURE — numpy front-end of the CPU oracle (oracle/cpu_oracle.cpp) and of the
reference's own host build (oracle/_ref/libref_matching_host.so, see build_ref.sh).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product path never does.

Function names mirror the reference ops (mast3r_slam/backend/src/gn.cpp:116-123) and take / return
numpy arrays with the reference's shapes and dtypes.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBS = {}

KIND = {"points": 0, "rays": 1, "calib": 2}


def build(force=False):
    """Compile the oracle (and, if /root/reference exists, oracle/_ref)."""
    if force or not os.path.exists(os.path.join(_HERE, "libm3r_oracle.so")):
        subprocess.check_call(["make", "-C", _HERE, "all"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["bash", os.path.join(_HERE, "build_ref.sh"), "all"])


def _lib(name="libm3r_oracle.so"):
    if name not in _LIBS:
        path = os.path.join(_HERE, name)
        if not os.path.exists(path) and not name.startswith("_ref"):
            subprocess.check_call(["make", "-C", _HERE, "all"], stdout=subprocess.DEVNULL)
        _LIBS[name] = C.CDLL(path)
    return _LIBS[name]


def have_ref_host():
    return os.path.exists(os.path.join(_HERE, "_ref", "libref_matching_host.so"))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    a = np.ascontiguousarray(a, dtype=dt)
    return a


# ------------------------------------------------------------------------------------ matching
def iter_proj(rays_img, pts3d_norm, p_init, max_iter, lambda_init, cost_thresh, impl="port",
              fma=False):
    """-> (p_new f32 [b,n,2], converged bool [b,n]).  impl: 'port' (cpu_oracle.cpp) or
    'reference' (the reference kernel body compiled for the host)."""
    rays = _c(rays_img, np.float32)
    pts = _c(pts3d_norm, np.float32)
    p0 = _c(p_init, np.float32)
    b, h, w, c = rays.shape
    assert c == 9
    n = pts.shape[1]
    p_new = np.zeros((b, n, 2), np.float32)
    conv = np.zeros((b, n), np.uint8)
    args = (_p(rays), _p(pts), _p(p0), _p(p_new), _p(conv), b, h, w, n, int(max_iter),
            C.c_float(lambda_init), C.c_float(cost_thresh))
    if impl == "reference":
        _lib("_ref/libref_matching_host.so").ref_host_iter_proj(*args)
    else:
        _lib("libm3r_oracle_fma.so" if fma else "libm3r_oracle.so").orc_iter_proj(*args)
    return p_new, conv.astype(bool)


def refine_matches(D11, D21, p1, radius, dilation_max, impl="port"):
    """-> p1_new int64 [b,n,2].  dtype of D11 picks the instantiation (float16/32/64)."""
    D11 = np.ascontiguousarray(D11)
    D21 = np.ascontiguousarray(D21, dtype=D11.dtype)
    p1 = _c(p1, np.int64)
    b, h, w, fdim = D11.shape
    n = p1.shape[1]
    out = np.zeros((b, n, 2), np.int64)
    suffix = {np.dtype(np.float16): "f16", np.dtype(np.float32): "f32",
              np.dtype(np.float64): "f64"}[D11.dtype]
    args = (_p(D11), _p(D21), _p(p1), _p(out), b, h, w, n, fdim, int(radius), int(dilation_max))
    if impl == "reference":
        getattr(_lib("_ref/libref_matching_host.so"), "ref_host_refine_matches_" + suffix)(*args)
    else:
        getattr(_lib(), "orc_refine_matches_" + suffix)(*args)
    return out


# ------------------------------------------------------------------------------------------ GN
def compact_indices(ii, jj):
    ii = _c(ii, np.int64)
    jj = _c(jj, np.int64)
    ii_c = np.zeros_like(ii)
    jj_c = np.zeros_like(jj)
    nuniq = _lib().orc_compact_indices(_p(ii), _p(jj), len(ii), _p(ii_c), _p(jj_c))
    return ii_c, jj_c, nuniq


def gn_accumulate(kind, Twc, Xs, Cs, K, ii_c, jj_c, idx, valid, Q, sigma_a, sigma_b, C_thresh,
                  Q_thresh, height=0, width=0, pixel_border=0, z_eps=0.0, fma=False):
    """One accumulate pass in the reference layout -> (Hs f32 [4,E,7,7], gs f32 [2,E,7])."""
    Twc = _c(Twc, np.float32)
    Xs = _c(Xs, np.float32)
    Cs = _c(Cs, np.float32)
    Kc = _c(K if K is not None else np.eye(3), np.float32)
    ii_c = _c(ii_c, np.int64)
    jj_c = _c(jj_c, np.int64)
    idx = _c(idx, np.int64)
    valid = _c(valid, np.uint8)
    Q = _c(Q, np.float32)
    N, n = Xs.shape[0], Xs.shape[1]
    E = len(ii_c)
    Hs = np.zeros((4, E, 7, 7), np.float32)
    gs = np.zeros((2, E, 7), np.float32)
    lib = _lib("libm3r_oracle_fma.so" if fma else "libm3r_oracle.so")
    lib.orc_gn_accumulate(KIND[kind], _p(Twc), _p(Xs), _p(Cs), _p(Kc), _p(ii_c), _p(jj_c), _p(idx),
                          _p(valid), _p(Q), N, n, E, C.c_float(sigma_a), C.c_float(sigma_b),
                          C.c_float(C_thresh), C.c_float(Q_thresh), int(height), int(width),
                          int(pixel_border), C.c_float(z_eps), _p(Hs), _p(gs))
    return Hs, gs


def gn_solve(Hs, gs, ii_opt, jj_opt, num_free):
    Hs = _c(Hs, np.float32)
    gs = _c(gs, np.float32)
    ii_opt = _c(ii_opt, np.int64)
    jj_opt = _c(jj_opt, np.int64)
    dx = np.zeros((num_free, 7), np.float32)
    failed = C.c_int(0)
    _lib().orc_gn_solve(_p(Hs), _p(gs), _p(ii_opt), _p(jj_opt), len(ii_opt), int(num_free), _p(dx),
                        C.byref(failed))
    return dx, bool(failed.value)


def pose_retr(Twc, dx, num_fix=1):
    Twc = np.array(Twc, dtype=np.float32, order="C", copy=True)
    dx = _c(dx, np.float32)
    _lib().orc_pose_retr(_p(Twc), _p(dx), Twc.shape[0], int(num_fix))
    return Twc


def sim3_exp(xi):
    xi = _c(xi, np.float32)
    out = np.zeros(8, np.float32)
    _lib().orc_sim3_exp(_p(xi), _p(out))
    return out


def gauss_newton(kind, Twc, Xs, Cs, K, ii, jj, idx, valid, Q, sigma_a, sigma_b, C_thresh, Q_thresh,
                 max_iter, delta_thresh, height=0, width=0, pixel_border=0, z_eps=0.0):
    """Full loop.  -> (Twc_new [N,8] f32, dx [N-1,7] f32, iterations, cholesky_failures)."""
    Twc = np.array(Twc, dtype=np.float32, order="C", copy=True)
    Xs = _c(Xs, np.float32)
    Cs = _c(Cs, np.float32)
    Kc = _c(K if K is not None else np.eye(3), np.float32)
    ii = _c(ii, np.int64)
    jj = _c(jj, np.int64)
    idx = _c(idx, np.int64)
    valid = _c(valid, np.uint8)
    Q = _c(Q, np.float32)
    N, n = Xs.shape[0], Xs.shape[1]
    E = len(ii)
    dx = np.zeros((N - 1, 7), np.float32)
    nfail = C.c_int(0)
    iters = _lib().orc_gauss_newton(KIND[kind], _p(Twc), _p(Xs), _p(Cs), _p(Kc), _p(ii), _p(jj),
                                    _p(idx), _p(valid), _p(Q), N, n, E, C.c_float(sigma_a),
                                    C.c_float(sigma_b), C.c_float(C_thresh), C.c_float(Q_thresh),
                                    int(height), int(width), int(pixel_border), C.c_float(z_eps),
                                    int(max_iter), C.c_float(delta_thresh), _p(dx), C.byref(nfail))
    return Twc, dx, iters, nfail.value


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx, valid, Q, sigma_ray, sigma_dist, C_thresh, Q_thresh,
                      max_iter, delta_thresh):
    return gauss_newton("rays", Twc, Xs, Cs, None, ii, jj, idx, valid, Q, sigma_ray, sigma_dist,
                        C_thresh, Q_thresh, max_iter, delta_thresh)


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx, valid, Q, sigma_point, C_thresh, Q_thresh,
                        max_iter, delta_thresh):
    return gauss_newton("points", Twc, Xs, Cs, None, ii, jj, idx, valid, Q, sigma_point, 0.0,
                        C_thresh, Q_thresh, max_iter, delta_thresh)


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx, valid, Q, height, width, pixel_border, z_eps,
                       sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter, delta_thresh):
    return gauss_newton("calib", Twc, Xs, Cs, K, ii, jj, idx, valid, Q, sigma_pixel, sigma_depth,
                        C_thresh, Q_thresh, max_iter, delta_thresh, height, width, pixel_border,
                        z_eps)
