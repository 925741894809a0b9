"""TEST INFRASTRUCTURE ONLY — numpy (float64) restatement of the frame tracker's pose optimisation, the checker
for csrc/frame_track.cu and monst3r-slam_b200/tracker.py.  Nothing in the product imports this file.

Follows mast3r_slam/tracker.py:141-154 (solve), :156-214 (opt_pose_ray_dist_sim3), :216-270
(opt_pose_calib_sim3), mast3r_slam/nonlinear_optimizer.py:5-34 (check_convergence, huber) and
mast3r_slam/geometry.py:12-57 (point_to_ray_dist, act_Sim3), :60-105 (project_calib).

Pinning.  **Pinned against the reference's own python** (round 2): tests/golden/tracker_{rays,calib}.npz hold inputs,
per-iteration costs / updates, iteration counts and final poses recorded by running the reference's unmodified
`FrameTracker.opt_pose_ray_dist_sim3 / opt_pose_calib_sim3 / solve`, `check_convergence`, `huber`, `act_Sim3`,
`point_to_ray_dist`, `project_calib`, `constrain_points_to_ray` (tests/golden/make_golden_tracker.py imports them from
/root/reference; float32 torch on the CPU).  Only `lietorch` -- absent from /root/reference and from the image -- is a
stand-in there: a Sim3 container with `.data`, `inv`, `*`, `act` (textbook Sim(3) group operations) and `retr`, which
is the oracle's host retraction `orc.pose_retr`, itself bit-pinned to the reference's `pose_retr_kernel`
(gn_kernels.cu:299-453, lietorch's closed form).  tests/test_tracker_oracle.py checks this restatement against the
fixtures (same iteration count, poses within 2e-5, first cost / update within 1e-4) and, independently, against the
pinned Gauss-Newton oracle on the equivalent two-keyframe graph.
"""
import math

import numpy as np

from . import oracle as orc

HUBER_K = 1.345


def quat_R(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def quat_mul(a, b):
    x1, y1, z1, w1 = a
    x2, y2, z2, w2 = b
    return np.array([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                     w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2])


def sim3_inv(T):
    T = np.asarray(T, np.float64).reshape(8)
    qi = np.array([-T[3], -T[4], -T[5], T[6]])
    si = 1.0 / T[7]
    return np.concatenate([-si * (quat_R(qi) @ T[:3]), qi, [si]])


def sim3_mul(A, B):
    A = np.asarray(A, np.float64).reshape(8)
    B = np.asarray(B, np.float64).reshape(8)
    return np.concatenate([A[7] * (quat_R(A[3:7]) @ B[:3]) + A[:3], quat_mul(A[3:7], B[3:7]), [A[7] * B[7]]])


def sim3_act(T, X):
    T = np.asarray(T, np.float64).reshape(8)
    return T[7] * (X @ quat_R(T[3:7]).T) + T[:3]


def sim3_retr(T, tau):
    """lietorch retr: exp(tau) * T — through the pinned host retraction (fp32, like lietorch)"""
    return orc.pose_retr(np.asarray(T, np.float32).reshape(1, 8), np.asarray(tau, np.float32).reshape(1, 7),
                         num_fix=0)[0].astype(np.float64)


def skew(X):
    o = np.zeros(X.shape[0])
    x, y, z = X[:, 0], X[:, 1], X[:, 2]
    return np.stack([o, -z, y, z, o, -x, -y, x, o], axis=-1).reshape(-1, 3, 3)


def act_sim3_jac(T, X):
    """geometry.py:45-57: (pW, [I | -skew(pW) | pW])"""
    pW = sim3_act(T, X)
    I = np.broadcast_to(np.eye(3), (X.shape[0], 3, 3))
    return pW, np.concatenate([I, -skew(pW), pW[:, :, None]], axis=-1)


def point_to_ray_dist(X, jacobian=False):
    """geometry.py:17-34"""
    d = np.linalg.norm(X, axis=-1, keepdims=True)
    d_inv = 1.0 / d
    r = d_inv * X
    rd = np.concatenate([r, d], axis=-1)
    if not jacobian:
        return rd
    dr = d_inv[:, :, None] * (np.eye(3)[None] - (d_inv ** 2)[:, :, None] * (X[:, :, None] @ X[:, None, :]))
    return rd, np.concatenate([dr, r[:, None, :]], axis=-2)


def project_calib(P, K, img_size, border, z_eps):
    """geometry.py:60-105 with jacobian=True"""
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    x, y, z = P[:, 0], P[:, 1], P[:, 2]
    with np.errstate(divide="ignore", invalid="ignore"):
        p = (P @ K.T)
        p = p / p[:, 2:3]
        u, v = p[:, 0], p[:, 1]
        valid = (u > border) & (u < img_size[1] - 1 - border) & (v > border) & (v < img_size[0] - 1 - border) & (z > z_eps)
        logz = np.where(z > z_eps, np.log(np.where(z > z_eps, z, 1.0)), 0.0)
        z_inv = 1.0 / z
    pz = np.stack([u, v, logz], axis=-1)
    J = np.zeros((P.shape[0], 3, 3))
    J[:, 0, 0] = fx
    J[:, 1, 1] = fy
    J[:, 0, 2] = -fx * x * z_inv
    J[:, 1, 2] = -fy * y * z_inv
    J *= z_inv[:, None, None]
    J[:, 2, 2] = z_inv
    return pz, J, valid


def huber(r, k=HUBER_K):
    """nonlinear_optimizer.py:29-34"""
    a = np.abs(r)
    with np.errstate(divide="ignore"):
        return np.where(a < k, 1.0, k / a)


def solve(sqrt_info, r, J):
    """tracker.py:141-154 -> (tau [7], cost); raises LinAlgError like torch.linalg.cholesky"""
    robust = sqrt_info * np.sqrt(huber(sqrt_info * r))
    A = (robust[..., None] * J).reshape(-1, J.shape[-1])
    b = (robust * r).reshape(-1, 1)
    H = A.T @ A
    g = -A.T @ b
    cost = 0.5 * float((b.T @ b)[0, 0])
    L = np.linalg.cholesky(H)
    tau = np.linalg.solve(L.T, np.linalg.solve(L, g)).reshape(-1)
    return tau, cost


def check_convergence(rel_error_threshold, delta_norm_threshold, old_cost, new_cost, delta):
    """nonlinear_optimizer.py:5-26"""
    cost_diff = old_cost - new_cost
    rel_dec = math.fabs(cost_diff / old_cost) if not (math.isinf(old_cost) and math.isinf(cost_diff)) else float("nan")
    return bool(rel_dec < rel_error_threshold or np.linalg.norm(delta) < delta_norm_threshold)


def opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, sigma_ray, sigma_dist, max_iters, rel_error, delta_norm):
    """tracker.py:156-214 -> (T_WCf [8], T_CkCf [8], iterations executed)"""
    Xf, Xk = np.asarray(Xf, np.float64), np.asarray(Xk, np.float64)
    Qk = np.asarray(Qk, np.float64).reshape(-1, 1)
    valid = np.asarray(valid).reshape(-1, 1).astype(np.float64)
    si_ray = 1.0 / sigma_ray * valid * np.sqrt(Qk)
    si_dist = 1.0 / sigma_dist * valid * np.sqrt(Qk)
    sqrt_info = np.concatenate([si_ray, si_ray, si_ray, si_dist], axis=1)
    T = sim3_mul(sim3_inv(T_WCk), T_WCf)
    rd_k = point_to_ray_dist(Xk)
    old_cost = float("inf")
    steps = 0
    for step in range(max_iters):
        Xf_Ck, dX = act_sim3_jac(T, Xf)
        rd_f, drd = point_to_ray_dist(Xf_Ck, jacobian=True)
        r = rd_k - rd_f
        J = -drd @ dX
        tau, new_cost = solve(sqrt_info, r, J)
        T = sim3_retr(T, tau)
        steps = step + 1
        if check_convergence(rel_error, delta_norm, old_cost, new_cost, tau):
            break
        old_cost = new_cost
    return sim3_mul(T_WCk, T), T, steps


def opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, K, img_size, sigma_pixel, sigma_depth, pixel_border, depth_eps,
                        max_iters, rel_error, delta_norm):
    """tracker.py:216-270 with meas_k / valid_meas_k built as in get_points_poses (:128-137)"""
    Xf, Xk, K = np.asarray(Xf, np.float64), np.asarray(Xk, np.float64), np.asarray(K, np.float64)
    h, w = img_size
    Qk = np.asarray(Qk, np.float64).reshape(-1, 1)
    valid = np.asarray(valid).reshape(-1, 1).astype(np.float64)
    k = np.arange(h * w)
    valid_meas = Xk[:, 2:3] > depth_eps
    meas = np.stack([(k % w).astype(np.float64), (k // w).astype(np.float64),
                     np.log(np.where(valid_meas[:, 0], Xk[:, 2], 1.0))], axis=-1)
    meas[~valid_meas[:, 0]] = 0.0
    si_p = 1.0 / sigma_pixel * valid * np.sqrt(Qk)
    si_d = 1.0 / sigma_depth * valid * np.sqrt(Qk)
    sqrt_info = np.concatenate([si_p, si_p, si_d], axis=1)
    T = sim3_mul(sim3_inv(T_WCk), T_WCf)
    old_cost = float("inf")
    steps = 0
    for step in range(max_iters):
        Xf_Ck, dX = act_sim3_jac(T, Xf)
        pz, dpz, valid_proj = project_calib(Xf_Ck, K, img_size, pixel_border, depth_eps)
        v2 = (valid_proj[:, None] & valid_meas).astype(np.float64)
        si2 = v2 * sqrt_info
        r = np.where(v2 > 0, meas - np.nan_to_num(pz), 0.0)
        J = np.where(v2[:, :, None] > 0, -np.nan_to_num(dpz) @ dX, 0.0)
        tau, new_cost = solve(si2, r, J)
        T = sim3_retr(T, tau)
        steps = step + 1
        if check_convergence(rel_error, delta_norm, old_cost, new_cost, tau):
            break
        old_cost = new_cost
    return sim3_mul(T_WCk, T), T, steps
