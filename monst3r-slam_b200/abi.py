"""ctypes binding of libm3r_b200.so (include/m3r_abi.h) operating on torch CUDA tensors.

This is the thinnest possible host layer: it passes `tensor.data_ptr()` and sizes straight to the
C ABI and never touches the data.  It exists so that tests and bench.py can drive the C ABI
directly (the torch-free boundary a reference maintainer would bind), next to the pybind module
`mast3r_slam_backends` that the reference's python code imports.

There is no CPU fallback: if the library is missing, importing this module raises.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, os.environ.get("M3R_LIB_DIR", "lib"), "libm3r_b200.so")  # (M3R_LIB_DIR: A/B variant builds)

GN_POINTS, GN_RAYS, GN_CALIB = 0, 1, 2
MAX_KF_ID = 1 << 20

EXPORTS = [
    "m3r_abi_version", "m3r_last_error", "m3r_iter_proj", "m3r_iter_proj_ws", "m3r_iter_proj_workspace_bytes",
    "m3r_refine_matches_f16", "m3r_match_workspace_bytes", "m3r_match_fused",
    "m3r_refine_matches_f32", "m3r_refine_matches_f64", "m3r_gn_workspace_bytes", "m3r_gn_workspace_bytes_sharded",
    "m3r_gauss_newton", "m3r_gn_prepare", "m3r_gn_pack", "m3r_gn_accumulate", "m3r_gn_edge_blocks", "m3r_gn_solve_retract",
    "m3r_gn_export_hs_gs", "m3r_pose_retr", "m3r_comm_unique_id", "m3r_comm_init", "m3r_comm_destroy",
    "m3r_comm_allreduce_f64", "m3r_edge_admit", "m3r_constrain_points_to_ray",
    "m3r_track_frame", "m3r_track_workspace_bytes", "m3r_iter_proj_static", "m3r_iter_proj_static_max_texels",
    "m3r_xchg_create", "m3r_xchg_connect", "m3r_xchg_destroy",
]


class GnParams(C.Structure):
    _fields_ = [("kind", C.c_int), ("sigma_a", C.c_float), ("sigma_b", C.c_float),
                ("C_thresh", C.c_float), ("Q_thresh", C.c_float), ("max_iter", C.c_int),
                ("delta_thresh", C.c_float), ("height", C.c_int), ("width", C.c_int),
                ("pixel_border", C.c_int), ("z_eps", C.c_float)]


class TrackParams(C.Structure):
    _fields_ = [("kind", C.c_int), ("sigma_a", C.c_float), ("sigma_b", C.c_float), ("max_iter", C.c_int),
                ("rel_error", C.c_float), ("delta_norm", C.c_float), ("huber", C.c_float), ("height", C.c_int),
                ("width", C.c_int), ("pixel_border", C.c_int), ("z_eps", C.c_float)]


class GnShard(C.Structure):
    _fields_ = [("edge_lo", C.c_int), ("edge_hi", C.c_int), ("comm", C.c_void_p), ("exchange", C.c_int),
                ("xchg", C.c_void_p)]


EXCHANGE_ALLREDUCE, EXCHANGE_ALLGATHER, EXCHANGE_PEER = 0, 1, 2
IPC_HANDLE_BYTES = 64


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python monst3r-slam_b200/build.py` "
            "(there is no CPU or PyTorch fallback for this backend)")
    lib = C.CDLL(LIB_PATH)
    lib.m3r_last_error.restype = C.c_char_p
    lib.m3r_iter_proj_workspace_bytes.restype = C.c_size_t
    lib.m3r_iter_proj_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.m3r_match_workspace_bytes.restype = C.c_size_t
    lib.m3r_match_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.m3r_gn_workspace_bytes.restype = C.c_size_t
    lib.m3r_gn_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.m3r_gn_workspace_bytes_sharded.restype = C.c_size_t
    lib.m3r_gn_workspace_bytes_sharded.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
    lib.m3r_track_workspace_bytes.restype = C.c_size_t
    lib.m3r_track_workspace_bytes.argtypes = [C.c_int]
    lib.m3r_iter_proj_static_max_texels.restype = C.c_size_t
    lib.m3r_iter_proj_static_max_texels.argtypes = []
    return lib


lib = _load()


class M3RError(RuntimeError):
    pass


def _check(rc, what):
    if rc != 0:
        raise M3RError(f"{what} failed: code {rc} ({lib.m3r_last_error().decode()})")


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _req(t, name, dtype=None):
    if not t.is_cuda:
        raise M3RError(f"{name} must be a CUDA tensor (this backend has no CPU path)")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if dtype is not None and t.dtype != dtype:
        raise M3RError(f"{name} must be {dtype}, got {t.dtype}")


def iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh, padded=True):
    """padded=True: m3r_iter_proj_ws (texel planes in a scratch buffer); False: the plain kernel."""
    for t, nm in ((rays_img_with_grad, "rays_img_with_grad"), (pts_3d_norm, "pts_3d_norm"), (p_init, "p_init")):
        _req(t, nm, torch.float32)
    b, h, w, c = rays_img_with_grad.shape
    assert c == 9
    n = p_init.shape[1]
    p_new = torch.empty((b, n, 2), dtype=torch.float32, device=p_init.device)
    conv = torch.empty((b, n), dtype=torch.bool, device=p_init.device)
    if padded:
        nbytes = int(lib.m3r_iter_proj_workspace_bytes(b, h, w))
        ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=p_init.device)
        _check(lib.m3r_iter_proj_ws(_ptr(rays_img_with_grad), _ptr(pts_3d_norm), _ptr(p_init), _ptr(p_new),
                                    _ptr(conv), b, h, w, n, int(max_iter), C.c_float(lambda_init),
                                    C.c_float(cost_thresh), _ptr(ws), C.c_size_t(nbytes), _stream()),
               "m3r_iter_proj_ws")
    else:
        _check(lib.m3r_iter_proj(_ptr(rays_img_with_grad), _ptr(pts_3d_norm), _ptr(p_init), _ptr(p_new),
                                 _ptr(conv), b, h, w, n, int(max_iter), C.c_float(lambda_init),
                                 C.c_float(cost_thresh), _stream()), "m3r_iter_proj")
    return [p_new, conv]


def iter_proj_static(rays_img_with_grad, pts_3d_norm, p_init, p_new, converged, max_iter, lambda_init, cost_thresh):
    """m3r_iter_proj_static: the library's static device scratch, caller-provided outputs -- nothing is allocated, here or
    in the library.  p_new [b,n,2] f32 and converged [b,n] bool are written in place.  Calls share the scratch: keep them
    on one stream."""
    for t, nm in ((rays_img_with_grad, "rays_img_with_grad"), (pts_3d_norm, "pts_3d_norm"), (p_init, "p_init"),
                  (p_new, "p_new")):
        _req(t, nm, torch.float32)
    _req(converged, "converged", torch.bool)
    b, h, w, c = rays_img_with_grad.shape
    assert c == 9
    n = p_init.shape[1]
    assert tuple(p_new.shape) == (b, n, 2) and tuple(converged.shape) == (b, n)
    _check(lib.m3r_iter_proj_static(_ptr(rays_img_with_grad), _ptr(pts_3d_norm), _ptr(p_init), _ptr(p_new), _ptr(converged),
                                    b, h, w, n, int(max_iter), C.c_float(lambda_init), C.c_float(cost_thresh), _stream()),
           "m3r_iter_proj_static")


def match_fused(X11, X21, D11, D21, idx_init, max_iter, lambda_init, cost_thresh, dist_thresh, radius,
                dilation_max):
    """matching.match() in five launches (m3r_match_fused) -> (idx_1_to_2 int64 [b,n], valid bool [b,n,1])"""
    _req(X11, "X11", torch.float32)
    _req(X21, "X21", torch.float32)
    b, h, w, _ = X11.shape
    n = h * w
    D21 = D21.reshape(b, n, -1)
    fdim = D11.shape[-1]
    assert D11.dtype == D21.dtype and D11.dtype in (torch.float32, torch.float16)
    _req(D11, "D11")
    _req(D21, "D21")
    is_f32 = 1 if D11.dtype == torch.float32 else 0
    if idx_init is not None:
        _req(idx_init, "idx_1_to_2_init", torch.int64)
    idx = torch.empty((b, n), dtype=torch.int64, device=X11.device)
    valid = torch.empty((b, n, 1), dtype=torch.bool, device=X11.device)
    nbytes = int(lib.m3r_match_workspace_bytes(b, h, w, fdim, is_f32))
    ws = torch.empty(nbytes, dtype=torch.uint8, device=X11.device)
    _check(lib.m3r_match_fused(_ptr(X11), _ptr(X21), _ptr(D11), _ptr(D21), fdim, is_f32,
                               _ptr(idx_init) if idx_init is not None else None, b, h, w, int(max_iter),
                               C.c_float(lambda_init), C.c_float(cost_thresh), C.c_float(dist_thresh), int(radius),
                               int(dilation_max), _ptr(idx), _ptr(valid), _ptr(ws), C.c_size_t(nbytes), _stream()),
           "m3r_match_fused")
    return idx, valid


def refine_matches(D11, D21, p1, radius, dilation_max):
    _req(D11, "D11")
    _req(D21, "D21", D11.dtype)
    _req(p1, "p1", torch.int64)
    b, h, w, fdim = D11.shape
    n = p1.shape[1]
    out = torch.empty((b, n, 2), dtype=torch.int64, device=p1.device)
    fn = {torch.float16: lib.m3r_refine_matches_f16, torch.float32: lib.m3r_refine_matches_f32,
          torch.float64: lib.m3r_refine_matches_f64}[D11.dtype]
    _check(fn(_ptr(D11), _ptr(D21), _ptr(p1), _ptr(out), b, h, w, n, fdim, int(radius), int(dilation_max),
              _stream()), "m3r_refine_matches")
    return [out]


def gn_workspace(num_kf, num_edges, num_pts, device, edges_local=None):
    """device workspace; a rank that owns only `edges_local` edges keeps packed slabs for those only"""
    if edges_local is None:
        nbytes = lib.m3r_gn_workspace_bytes(num_kf, num_edges, num_pts)
    else:
        nbytes = lib.m3r_gn_workspace_bytes_sharded(num_kf, num_edges, num_pts, int(edges_local))
    if nbytes == 0:
        raise M3RError("m3r_gn_workspace_bytes: bad problem size")
    return torch.empty(nbytes, dtype=torch.uint8, device=device), nbytes


def make_params(kind, sigma_a, sigma_b, C_thresh, Q_thresh, max_iter, delta_thresh, height=0, width=0,
                pixel_border=0, z_eps=0.0):
    return GnParams(kind, sigma_a, sigma_b, C_thresh, Q_thresh, int(max_iter), delta_thresh, int(height),
                    int(width), int(pixel_border), z_eps)


def gauss_newton(params, Twc, Xs, Cs, K, ii, jj, idx, valid, Q, shard=None, workspace=None, info=None,
                 dx=None):
    """Full asynchronous GN loop (m3r_gauss_newton).  Twc is updated in place.
    Returns (dx [N-1,7] f32, info int32[4] device tensor)."""
    for t, nm, dt in ((Twc, "Twc", torch.float32), (Xs, "Xs", torch.float32), (Cs, "Cs", torch.float32),
                      (ii, "ii", torch.int64), (jj, "jj", torch.int64), (idx, "idx_ii2jj", torch.int64),
                      (valid, "valid_match", torch.bool), (Q, "Q", torch.float32)):
        _req(t, nm, dt)
    if K is not None:
        _req(K, "K", torch.float32)
    N, n = Xs.shape[0], Xs.shape[1]
    E = ii.shape[0]
    if workspace is None:
        workspace, nbytes = gn_workspace(N, E, n, Twc.device)
    else:
        nbytes = workspace.numel()
    if dx is None:
        dx = torch.empty((N - 1, 7), dtype=torch.float32, device=Twc.device)
    if info is None:
        info = torch.empty(4, dtype=torch.int32, device=Twc.device)
    sh = C.byref(shard) if shard is not None else None
    _check(lib.m3r_gauss_newton(_ptr(Twc), _ptr(Xs), _ptr(Cs), _ptr(K), _ptr(ii), _ptr(jj), _ptr(idx),
                                _ptr(valid), _ptr(Q), N, n, E, C.byref(params), _ptr(dx), _ptr(info),
                                _ptr(workspace), C.c_size_t(nbytes), sh, _stream()), "m3r_gauss_newton")
    return dx, info


def gn_prepare(ii, jj, num_kf, num_pts, workspace):
    _check(lib.m3r_gn_prepare(_ptr(ii), _ptr(jj), num_kf, num_pts, ii.shape[0], _ptr(workspace),
                              C.c_size_t(workspace.numel()), _stream()), "m3r_gn_prepare")


def gn_pack(params, Cs, idx, valid, Q, num_kf, num_edges, workspace, shard=None):
    """fold the pose-independent per-point data into the 8-byte records (once per problem)"""
    n = Cs.shape[1]
    sh = C.byref(shard) if shard is not None else None
    _check(lib.m3r_gn_pack(_ptr(Cs), _ptr(idx), _ptr(valid), _ptr(Q), num_kf, n, num_edges, C.byref(params),
                           _ptr(workspace), C.c_size_t(workspace.numel()), sh, _stream()), "m3r_gn_pack")


def gn_accumulate_packed(params, Twc, Xs, K, num_edges, workspace, shard=None):
    N, n = Xs.shape[0], Xs.shape[1]
    sh = C.byref(shard) if shard is not None else None
    _check(lib.m3r_gn_accumulate(_ptr(Twc), _ptr(Xs), _ptr(K), N, n, num_edges, C.byref(params),
                                 _ptr(workspace), C.c_size_t(workspace.numel()), sh, _stream()),
           "m3r_gn_accumulate")


def gn_accumulate(params, Twc, Xs, Cs, K, idx, valid, Q, num_edges, workspace, shard=None):
    """pack + accumulate (the two building blocks of one GN iteration before the solve)"""
    gn_pack(params, Cs, idx, valid, Q, Xs.shape[0], num_edges, workspace, shard)
    gn_accumulate_packed(params, Twc, Xs, K, num_edges, workspace, shard)


def gn_export_hs_gs(workspace, num_kf, num_edges, num_pts):
    Hs = torch.empty((4, num_edges, 7, 7), dtype=torch.float32, device=workspace.device)
    gs = torch.empty((2, num_edges, 7), dtype=torch.float32, device=workspace.device)
    _check(lib.m3r_gn_export_hs_gs(_ptr(workspace), num_kf, num_edges, num_pts, _ptr(Hs), _ptr(gs),
                                   _stream()), "m3r_gn_export_hs_gs")
    return Hs, gs


def gn_edge_blocks(workspace, num_kf, num_edges, num_pts):
    """fp64 per-edge blocks [E,56] = (A_e 7x7 | b_e 7), copied out of the workspace."""
    p = C.c_void_p(0)
    _check(lib.m3r_gn_edge_blocks(_ptr(workspace), num_kf, num_edges, num_pts, C.byref(p)),
           "m3r_gn_edge_blocks")
    off = p.value - workspace.data_ptr()
    return workspace[off: off + num_edges * 56 * 8].view(torch.float64).view(num_edges, 56).clone()


def gn_solve_retract(Twc, num_edges, num_pts, delta_thresh, workspace):
    N = Twc.shape[0]
    dx = torch.zeros((N - 1, 7), dtype=torch.float32, device=Twc.device)
    info = torch.zeros(4, dtype=torch.int32, device=Twc.device)
    _check(lib.m3r_gn_solve_retract(_ptr(Twc), N, num_pts, num_edges, C.c_float(delta_thresh), _ptr(dx),
                                    _ptr(info), _ptr(workspace), C.c_size_t(workspace.numel()), _stream()),
           "m3r_gn_solve_retract")
    return dx, info


def pose_retr(Twc, dx, num_fix=1):
    _req(Twc, "Twc", torch.float32)
    _req(dx, "dx", torch.float32)
    _check(lib.m3r_pose_retr(_ptr(Twc), _ptr(dx), Twc.shape[0], int(num_fix), _stream()), "m3r_pose_retr")


# ---- NCCL plumbing ------------------------------------------------------------------------------
def comm_unique_id():
    buf = C.create_string_buffer(128)
    _check(lib.m3r_comm_unique_id(buf), "m3r_comm_unique_id")
    return buf.raw


def comm_init(rank, world, uid):
    comm = C.c_void_p(0)
    _check(lib.m3r_comm_init(int(rank), int(world), C.c_char_p(uid), C.byref(comm)), "m3r_comm_init")
    return comm


def comm_destroy(comm):
    _check(lib.m3r_comm_destroy(comm), "m3r_comm_destroy")


def edge_admit(idx_i2j, idx_j2i, valid_j, valid_i, Qii, Qjj, Qji, Qij, Q_conf, store_idx, store_valid, store_Q, slot0):
    """m3r_edge_admit: Q of both directions of `batch` new pairs written into the store slabs at undirected
    slots slot0.., returns int32 counts [batch, 2] of (valid & Q > Q_conf).  global_opt.py:54-67, :89-96."""
    batch, n = idx_i2j.shape[0], idx_i2j.shape[1]
    for t, nm in ((idx_i2j, "idx_i2j"), (idx_j2i, "idx_j2i")):
        _req(t, nm, torch.int64)
    for t, nm in ((valid_j, "valid_match_j"), (valid_i, "valid_match_i")):
        _req(t, nm, torch.bool)
    for t, nm in ((Qii, "Qii"), (Qjj, "Qjj"), (Qji, "Qji"), (Qij, "Qij"), (store_Q, "store_Q")):
        _req(t, nm, torch.float32)
    _req(store_idx, "store_idx", torch.int64)
    _req(store_valid, "store_valid", torch.bool)
    for t in (idx_j2i, valid_j, valid_i, Qii, Qjj, Qji, Qij):
        if t.numel() != batch * n:
            raise M3RError("edge_admit: every input must hold batch x num_pts elements")
    if store_idx.shape[-1] != n or 2 * (slot0 + batch) > store_idx.shape[0]:
        raise M3RError("edge_admit: store slabs too small for this batch")
    counts = torch.empty((batch, 2), dtype=torch.int32, device=idx_i2j.device)
    _check(lib.m3r_edge_admit(_ptr(idx_i2j), _ptr(idx_j2i), _ptr(valid_j), _ptr(valid_i), _ptr(Qii), _ptr(Qjj),
                              _ptr(Qji), _ptr(Qij), int(batch), int(n), C.c_float(Q_conf), _ptr(store_idx),
                              _ptr(store_valid), _ptr(store_Q), C.c_int64(slot0), _ptr(counts), _stream()),
           "m3r_edge_admit")
    return counts


def constrain_points_to_ray(img_size, Xs, K, out=None):
    """geometry.py:37-42 on [N, h*w, 3] (or [N, h, w, 3]) float32 CUDA points."""
    h, w = int(img_size[0]), int(img_size[1])
    _req(Xs, "Xs", torch.float32)
    _req(K, "K", torch.float32)
    if Xs.numel() % (h * w * 3) != 0:
        raise M3RError("constrain_points_to_ray: Xs does not hold whole h x w x 3 maps")
    num = Xs.numel() // (h * w * 3)
    if out is None:
        out = torch.empty_like(Xs)
    _check(lib.m3r_constrain_points_to_ray(_ptr(Xs), _ptr(out), _ptr(K), int(num), h, w, _stream()),
           "m3r_constrain_points_to_ray")
    return out


def track_frame(T_WCf, T_WCk, Xf, Xk, valid, Q, K, params, workspace=None):
    """m3r_track_frame (tracker.py:156-270 in one launch).  T_WCf [8] / [1,8] f32 is updated in place.
    Xf: frame points already gathered by idx_f2k.  -> (T_CkCf [8], dx [7], info int32[4] device tensors)"""
    for t, nm, dt in ((T_WCf, "T_WCf", torch.float32), (T_WCk, "T_WCk", torch.float32), (Xf, "Xf", torch.float32),
                      (Xk, "Xk", torch.float32), (valid, "valid", torch.bool), (Q, "Q", torch.float32)):
        _req(t, nm, dt)
    if K is not None:
        _req(K, "K", torch.float32)
    n = Xk.numel() // 3
    if Xf.numel() != 3 * n or valid.numel() != n or Q.numel() != n or T_WCf.numel() != 8 or T_WCk.numel() != 8:
        raise M3RError("track_frame: Xf/Xk [n,3], valid/Q [n], poses [8] expected")
    dev = Xk.device
    if workspace is None:
        nbytes = int(lib.m3r_track_workspace_bytes(n))
        workspace = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    T_rel = torch.empty(8, dtype=torch.float32, device=dev)
    dx = torch.empty(7, dtype=torch.float32, device=dev)
    info = torch.empty(4, dtype=torch.int32, device=dev)
    _check(lib.m3r_track_frame(_ptr(T_WCf), _ptr(T_rel), _ptr(T_WCk), _ptr(Xf), _ptr(Xk), _ptr(valid), _ptr(Q), _ptr(K),
                               int(n), C.byref(params), _ptr(dx), _ptr(info), _ptr(workspace),
                               C.c_size_t(workspace.numel()), _stream()), "m3r_track_frame")
    return T_rel, dx, info


# ---- peer-memory exchange of the sharded Gauss-Newton loop (m3r_abi.h: m3r_xchg_*) -----------------------------
def xchg_create(capacity_bytes):
    """-> (handle c_void_p, 64-byte CUDA IPC handle of this rank's exchange buffer on the current device)"""
    x = C.c_void_p(0)
    buf = C.create_string_buffer(IPC_HANDLE_BYTES)
    _check(lib.m3r_xchg_create(C.c_size_t(int(capacity_bytes)), C.byref(x), buf), "m3r_xchg_create")
    return x, bytes(buf.raw)


def xchg_connect(x, rank, world, handles):
    """handles: the `world` 64-byte IPC handles in rank order"""
    assert len(handles) == world and all(len(h) == IPC_HANDLE_BYTES for h in handles)
    _check(lib.m3r_xchg_connect(x, int(rank), int(world), b"".join(handles)), "m3r_xchg_connect")


def xchg_destroy(x):
    _check(lib.m3r_xchg_destroy(x), "m3r_xchg_destroy")
