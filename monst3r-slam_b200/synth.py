"""Synthetic inputs for the five backend ops (SURVEY.md §8d generators, seeds 1001-1004).

There is no network and no dataset on the box, so tests and bench.py use these: a smooth pinhole
depth scene + low-frequency flow for the matching pair, and a height-field world ray-cast into N
Sim(3) cameras for the Gauss-Newton graphs.  Everything is torch and device-agnostic; all
randomness comes from one seeded torch.Generator on the CPU so CPU and GPU runs see the same data.
"""
import math

import torch
import torch.nn.functional as F


def _gen(seed):
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return g


def _smooth_field(h, w, g, terms, amp_lo, amp_hi, f_lo=0.5, f_hi=3.0):
    """sum of `terms` random low-frequency sinusoids on the pixel grid -> [h,w] float64"""
    v, u = torch.meshgrid(torch.arange(h, dtype=torch.float64), torch.arange(w, dtype=torch.float64),
                          indexing="ij")
    out = torch.zeros(h, w, dtype=torch.float64)
    for _ in range(terms):
        a = amp_lo + (amp_hi - amp_lo) * torch.rand(1, generator=g).item()
        fu = f_lo + (f_hi - f_lo) * torch.rand(1, generator=g).item()
        fv = f_lo + (f_hi - f_lo) * torch.rand(1, generator=g).item()
        ph = 2 * math.pi * torch.rand(1, generator=g).item()
        out += a * torch.sin(2 * math.pi * (fu * u / w + fv * v / h) + ph)
    return out


def _bilinear(img_bhwc, pu, pv):
    """sample img [b,h,w,c] at float pixel coords pu,pv [b,h,w] (align_corners pixel centres)"""
    b, h, w, _ = img_bhwc.shape
    gx = 2.0 * pu / (w - 1) - 1.0
    gy = 2.0 * pv / (h - 1) - 1.0
    grid = torch.stack((gx, gy), dim=-1).to(img_bhwc.dtype)
    out = F.grid_sample(img_bhwc.permute(0, 3, 1, 2), grid, mode="bilinear", padding_mode="border",
                        align_corners=True)
    return out.permute(0, 2, 3, 1)


def intrinsics(h, w):
    fx = fy = 0.85 * w
    return torch.tensor([[fx, 0.0, w / 2.0], [0.0, fy, h / 2.0], [0.0, 0.0, 1.0]], dtype=torch.float32)


def make_matching_pair(h=384, w=512, b=1, fdim=24, seed=1001, flow_amp=12.0, mask_frac=0.10,
                       device="cpu"):
    """Config 1.  Returns dict with X11,X21 [b,h,w,3] f32, D11,D21 [b,h,w,fdim] f32, the true
    float match (pu,pv) and an 'idx_init' previous-frame style initialisation (true match + integer
    jitter <= 3 px)."""
    g = _gen(seed)
    K = intrinsics(h, w).double()
    v, u = torch.meshgrid(torch.arange(h, dtype=torch.float64), torch.arange(w, dtype=torch.float64),
                          indexing="ij")
    out = {k: [] for k in ("X11", "X21", "D11", "D21", "pu", "pv")}
    for _ in range(b):
        z = 2.5 + _smooth_field(h, w, g, 6, 0.05, 0.25)
        X11 = torch.stack(((u - K[0, 2]) / K[0, 0] * z, (v - K[1, 2]) / K[1, 1] * z, z), dim=-1)
        fu = _smooth_field(h, w, g, 4, 0.0, flow_amp / 4, 0.3, 1.5)
        fv = _smooth_field(h, w, g, 4, 0.0, flow_amp / 4, 0.3, 1.5)
        pu = (u + fu).clamp(2.0, w - 3.0)
        pv = (v + fv).clamp(2.0, h - 3.0)
        X21 = _bilinear(X11[None], pu[None], pv[None])[0]
        X21 = X21 + 2e-3 * torch.randn(h, w, 3, generator=g, dtype=torch.float64)
        G = torch.randn(1, fdim, h, w, generator=g, dtype=torch.float64)
        G = F.avg_pool2d(F.pad(G, (4, 4, 4, 4), mode="reflect"), 9, stride=1)
        D11 = F.normalize(G[0].permute(1, 2, 0), dim=-1)
        D21 = _bilinear(D11[None], pu[None], pv[None])[0]
        D21 = F.normalize(D21 + 0.05 * torch.randn(h, w, fdim, generator=g, dtype=torch.float64), dim=-1)
        if mask_frac > 0:  # dynamic-mask case: zeroed descriptors never move (refine score > 0 rule)
            m = torch.rand(h, w, generator=g) < mask_frac
            D11 = D11.clone()
            D11[m] = 0.0
        for k_, t in (("X11", X11), ("X21", X21), ("D11", D11), ("D21", D21), ("pu", pu), ("pv", pv)):
            out[k_].append(t)
    res = {k_: torch.stack(t).float().to(device) for k_, t in out.items()}
    jit = torch.randint(-3, 4, (b, h, w, 2), generator=g)
    ui = (res["pu"].cpu().round().long() + jit[..., 0]).clamp(0, w - 1)
    vi = (res["pv"].cpu().round().long() + jit[..., 1]).clamp(0, h - 1)
    res["idx_init"] = (ui + w * vi).view(b, -1).to(device)
    res["K"] = intrinsics(h, w).to(device)
    return res


# ------------------------------------------------------------------------------------- Sim(3)
def quat_to_R(q):
    """q [...,4] (x,y,z,w) -> R [...,3,3]"""
    x, y, z, w = q.unbind(-1)
    R = torch.stack((1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                     2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                     2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)), dim=-1)
    return R.view(q.shape[:-1] + (3, 3))


def sim3_act(T, X):
    """T [...,8] = (t, q, s) ; X [...,n,3] -> s R X + t"""
    R = quat_to_R(T[..., 3:7])
    return T[..., None, 7:8] * (X @ R.transpose(-1, -2)) + T[..., None, 0:3]


def sim3_inv(T):
    R = quat_to_R(T[..., 3:7])
    s = T[..., 7:8]
    qinv = T[..., 3:7] * torch.tensor([-1.0, -1.0, -1.0, 1.0], dtype=T.dtype, device=T.device)
    tinv = -(R.transpose(-1, -2) @ T[..., 0:3, None])[..., 0] / s
    return torch.cat((tinv, qinv, 1.0 / s), dim=-1)


def _rand_pose(g, trans, max_deg, s_lo, s_hi):
    t = (torch.rand(3, generator=g, dtype=torch.float64) * 2 - 1) * torch.tensor(trans, dtype=torch.float64)
    axis = F.normalize(torch.randn(3, generator=g, dtype=torch.float64), dim=0)
    ang = math.radians(max_deg) * torch.rand(1, generator=g, dtype=torch.float64)
    q = torch.cat((axis * torch.sin(ang / 2), torch.cos(ang / 2)))
    s = s_lo + (s_hi - s_lo) * torch.rand(1, generator=g, dtype=torch.float64)
    return torch.cat((t, q, s))


def _surface(coef, x, y):
    z = torch.full_like(x, 3.0)
    for a, fx, fy, ph in coef:
        z = z + a * torch.sin(fx * x + fy * y + ph)
    return z


def make_graph(num_kf=32, num_und_edges=120, h=384, w=512, seed=1003, device="cpu", calib=False,
               valid_prob=0.9, perturb=(0.02, 0.01, 0.01), edges_chunk=16):
    """Configs 2-4.  Height-field world seen by num_kf Sim(3) cameras.  Returns dict:
    Twc_gt, Twc [N,8] f32 (init = gt perturbed, pose 0 = identity & fixed), Xs [N,n,3], Cs [N,n,1],
    K [3,3], ii,jj [2E] i64 (two-way, global ids = 0..N-1), idx [2E,n] i64, valid [2E,n,1] bool,
    Q [2E,n,1] f32."""
    g = _gen(seed)
    dev = torch.device(device)
    n = h * w
    K = intrinsics(h, w).double()
    coef = [(0.05 + 0.05 * torch.rand(1, generator=g).item(), 0.5 + 1.5 * torch.rand(1, generator=g).item(),
             0.5 + 1.5 * torch.rand(1, generator=g).item(), 2 * math.pi * torch.rand(1, generator=g).item())
            for _ in range(6)]
    ident = torch.tensor([0, 0, 0, 0, 0, 0, 1, 1], dtype=torch.float64)
    poses = [ident] + [_rand_pose(g, (0.5, 0.5, 0.2), 10.0, 0.9, 1.1) for _ in range(num_kf - 1)]
    Twc_gt = torch.stack(poses).to(dev)
    # ray-cast every pixel of every camera onto the surface (fixed-point on the depth along the ray)
    v, u = torch.meshgrid(torch.arange(h, dtype=torch.float64), torch.arange(w, dtype=torch.float64),
                          indexing="ij")
    rays = torch.stack(((u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], torch.ones_like(u)), -1).view(n, 3).to(dev)
    Xs = torch.empty(num_kf, n, 3, dtype=torch.float64, device=dev)
    for k in range(num_kf):
        T = Twc_gt[k]
        d = torch.full((n, 1), 3.0 / T[7].item(), dtype=torch.float64, device=dev)
        Rz = quat_to_R(T[3:7])[2]  # world z row
        dzdd = (T[7] * (rays @ Rz))[:, None]
        for _ in range(8):
            Xw = sim3_act(T, (rays * d))
            zs = _surface(coef, Xw[:, 0], Xw[:, 1])
            d = d + (zs - Xw[:, 2])[:, None] / dzdd
        Xs[k] = rays * d
    # edges: chain + random pairs, undirected then two-way (global_opt.py:104-110)
    und = [(k, k + 1) for k in range(min(num_kf - 1, num_und_edges))]
    seen = set(und)
    while len(und) < num_und_edges:
        a, b_ = sorted(torch.randint(0, num_kf, (2,), generator=g).tolist())
        if a != b_ and (a, b_) not in seen:
            seen.add((a, b_))
            und.append((a, b_))
    ii = torch.tensor([e[0] for e in und] + [e[1] for e in und], dtype=torch.int64)
    jj = torch.tensor([e[1] for e in und] + [e[0] for e in und], dtype=torch.int64)
    E2 = ii.numel()
    idx = torch.empty(E2, n, dtype=torch.int64, device=dev)
    valid = torch.empty(E2, n, 1, dtype=torch.bool, device=dev)
    Kd = K.to(dev)
    for e0 in range(0, E2, edges_chunk):
        sl = slice(e0, min(E2, e0 + edges_chunk))
        Ti, Tj = Twc_gt[ii[sl].to(dev)], Twc_gt[jj[sl].to(dev)]
        Pw = sim3_act(Tj, Xs[jj[sl].to(dev)])
        Pi = sim3_act(sim3_inv(Ti), Pw)
        uu = Kd[0, 0] * Pi[..., 0] / Pi[..., 2] + Kd[0, 2]
        vv = Kd[1, 1] * Pi[..., 1] / Pi[..., 2] + Kd[1, 2]
        ur, vr = uu.round().long(), vv.round().long()
        inside = (ur >= 0) & (ur < w) & (vr >= 0) & (vr < h) & (Pi[..., 2] > 0.1)
        keep = (torch.rand(ur.shape, generator=g) < valid_prob).to(dev)
        idx[sl] = (ur.clamp(0, w - 1) + w * vr.clamp(0, h - 1))
        valid[sl] = (inside & keep)[..., None]
    Q = torch.exp(1.0 + 0.5 * torch.randn(E2, n, 1, generator=g)).float().to(dev)
    Cs = (1.0 + torch.exp(torch.randn(num_kf, n, 1, generator=g))).float().to(dev)
    # initial poses: gt composed with a small left perturbation (first order is enough here)
    Twc = Twc_gt.clone()
    st, sr, ss = perturb
    for k in range(1, num_kf):
        dt = st * torch.randn(3, generator=g, dtype=torch.float64)
        dw = sr * torch.randn(3, generator=g, dtype=torch.float64)
        ds = ss * torch.randn(1, generator=g, dtype=torch.float64)
        ang = dw.norm()
        dq = torch.cat((dw / ang * torch.sin(ang / 2), torch.cos(ang / 2)[None])).to(dev)
        Tk = Twc_gt[k]
        R = quat_to_R(dq)
        es = torch.exp(ds).to(dev)
        t_new = es * (R @ Tk[0:3]) + dt.to(dev)
        x1, y1, z1, w1 = dq
        x2, y2, z2, w2 = Tk[3:7]
        q_new = torch.stack((w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                             w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2))
        Twc[k] = torch.cat((t_new, q_new, es * Tk[7:8]))
    Xs32 = Xs.float()
    if calib:  # constrain_points_to_ray semantics (geometry.py:37-42): keep depth, snap x,y to the ray
        Xs32 = (rays.float() * Xs32[..., 2:3]).contiguous()
    return dict(Twc_gt=Twc_gt.float(), Twc=Twc.float().contiguous(), Xs=Xs32.contiguous(), Cs=Cs, K=K.float().to(dev),
                ii=ii.to(dev), jj=jj.to(dev), idx=idx, valid=valid, Q=Q, h=h, w=w)


def make_graph_exact(num_kf=4, num_und_edges=5, n=768, seed=7, device="cpu", noise=0.0, valid_prob=0.9,
                     perturb=(0.02, 0.01, 0.01), h=None, w=None, calib=False):
    """Small well-conditioned graph for parity tests: every keyframe sees the same n world points
    under its own random pixel permutation, so correspondences are exact (plus optional gaussian
    noise) and the gather index is non-trivial.  With `calib`, n must be h*w: points are the
    back-projection of the pixel grid of each camera at random depth, matched through world space
    is impossible exactly, so calib uses camera 0's grid re-observed by nearby cameras and keeps the
    identity permutation (targets are pixel centres)."""
    g = _gen(seed)
    dev = torch.device(device)
    ident = torch.tensor([0, 0, 0, 0, 0, 0, 1, 1], dtype=torch.float64)
    poses = [ident] + [_rand_pose(g, (0.3, 0.3, 0.1), 6.0, 0.9, 1.1) for _ in range(num_kf - 1)]
    Twc_gt = torch.stack(poses)
    W = torch.randn(n, 3, generator=g, dtype=torch.float64) * torch.tensor([0.8, 0.6, 0.4], dtype=torch.float64)
    W = W + torch.tensor([0.0, 0.0, 3.0], dtype=torch.float64)
    perms = [torch.arange(n)] + [torch.randperm(n, generator=g) for _ in range(num_kf - 1)]
    if calib:
        perms = [torch.arange(n) for _ in range(num_kf)]
    Xs = torch.empty(num_kf, n, 3, dtype=torch.float64)
    for k in range(num_kf):
        Xc = sim3_act(sim3_inv(Twc_gt[k]), W)
        Xs[k, perms[k]] = Xc + noise * torch.randn(n, 3, generator=g, dtype=torch.float64)
    und = [(k, k + 1) for k in range(min(num_kf - 1, num_und_edges))]
    seen = set(und)
    while len(und) < num_und_edges:
        a, b_ = sorted(torch.randint(0, num_kf, (2,), generator=g).tolist())
        if a != b_ and (a, b_) not in seen:
            seen.add((a, b_))
            und.append((a, b_))
    ii = torch.tensor([e[0] for e in und] + [e[1] for e in und], dtype=torch.int64)
    jj = torch.tensor([e[1] for e in und] + [e[0] for e in und], dtype=torch.int64)
    E2 = ii.numel()
    inv = [torch.empty(n, dtype=torch.int64) for _ in range(num_kf)]
    for k in range(num_kf):
        inv[k][perms[k]] = torch.arange(n)
    idx = torch.stack([perms[ii[e]][inv[jj[e]]] for e in range(E2)])  # pixel k of j -> pixel of i
    valid = (torch.rand(E2, n, 1, generator=g) < valid_prob)
    Q = torch.exp(1.0 + 0.5 * torch.randn(E2, n, 1, generator=g)).float()
    Cs = (1.0 + torch.exp(torch.randn(num_kf, n, 1, generator=g))).float()
    Twc = Twc_gt.clone()
    st, sr, ss = perturb
    for k in range(1, num_kf):
        d = torch.cat((st * torch.randn(3, generator=g, dtype=torch.float64),
                       torch.zeros(4, dtype=torch.float64), torch.zeros(1, dtype=torch.float64)))
        dw = sr * torch.randn(3, generator=g, dtype=torch.float64)
        ang = dw.norm()
        dq = torch.cat((dw / ang * torch.sin(ang / 2), torch.cos(ang / 2)[None]))
        es = torch.exp(ss * torch.randn(1, generator=g, dtype=torch.float64))
        Tk = Twc_gt[k]
        x1, y1, z1, w1 = dq
        x2, y2, z2, w2 = Tk[3:7]
        q_new = torch.stack((w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                             w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2))
        Twc[k] = torch.cat((es * (quat_to_R(dq) @ Tk[0:3]) + d[0:3], q_new, es * Tk[7:8]))
    K = intrinsics(h, w) if (h and w) else intrinsics(24, 32)
    out = dict(Twc_gt=Twc_gt.float(), Twc=Twc.float().contiguous(), Xs=Xs.float().contiguous(), Cs=Cs, K=K,
               ii=ii, jj=jj, idx=idx.contiguous(), valid=valid, Q=Q, h=h, w=w)
    return {k_: (v.to(dev) if torch.is_tensor(v) else v) for k_, v in out.items()}


# ------------------------------------------------------------------------------------- sequences
def make_sequence(num_frames=60, h=384, w=512, fdim=24, seed=1005, device="cpu", noise=1e-3):
    """Config 5 stand-in: a smooth camera trajectory over the height-field world of make_graph.  For every frame: its
    ground-truth Sim(3) pose (scale 1), the canonical pointmap X [n,3] in its own camera frame (exact ray-cast + gaussian
    noise), a per-point confidence C and matching score Q, and a 24-d descriptor map that is a fixed function of the WORLD
    point (random Fourier features), so descriptors of the same surface point agree across frames like a trained
    descriptor head's would.  Returns dict(Twc_gt [F,8], X [F,n,3], C [F,n,1], Q [F,n,1], D [F,n,fdim], K, h, w)."""
    g = _gen(seed)
    dev = torch.device(device)
    n = h * w
    K = intrinsics(h, w).double()
    coef = [(0.05 + 0.05 * torch.rand(1, generator=g).item(), 0.5 + 1.5 * torch.rand(1, generator=g).item(),
             0.5 + 1.5 * torch.rand(1, generator=g).item(), 2 * math.pi * torch.rand(1, generator=g).item())
            for _ in range(6)]
    v, u = torch.meshgrid(torch.arange(h, dtype=torch.float64), torch.arange(w, dtype=torch.float64), indexing="ij")
    rays = torch.stack(((u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], torch.ones_like(u)), -1).view(n, 3).to(dev)
    fa = ((torch.rand(fdim, generator=g, dtype=torch.float64) * 2 - 1) * 12.0).to(dev)
    fb = ((torch.rand(fdim, generator=g, dtype=torch.float64) * 2 - 1) * 12.0).to(dev)
    fc = (torch.rand(fdim, generator=g, dtype=torch.float64) * 2 * math.pi).to(dev)
    poses, Xs, Ds = [], [], []
    for f in range(num_frames):
        s = f / max(num_frames - 1, 1)
        t = torch.tensor([0.6 * math.sin(math.pi * s), 0.25 * math.sin(2 * math.pi * s), 0.1 * math.sin(3 * math.pi * s)],
                         dtype=torch.float64)
        ang = math.radians(8.0) * math.sin(2 * math.pi * s)
        q = torch.tensor([0.0, math.sin(ang / 2), 0.0, math.cos(ang / 2)], dtype=torch.float64)
        T = torch.cat((t, q, torch.ones(1, dtype=torch.float64))).to(dev)
        d = torch.full((n, 1), 3.0, dtype=torch.float64, device=dev)
        Rz = quat_to_R(T[3:7])[2]
        dzdd = (rays @ Rz)[:, None]
        for _ in range(8):
            Xw = sim3_act(T, rays * d)
            d = d + (_surface(coef, Xw[:, 0], Xw[:, 1]) - Xw[:, 2])[:, None] / dzdd
        Xw = sim3_act(T, rays * d)
        D = torch.sin(Xw[:, 0:1] * fa + Xw[:, 1:2] * fb + fc)
        poses.append(T)
        Xs.append((rays * d + noise * torch.randn(n, 3, generator=g, dtype=torch.float64).to(dev)).float())
        Ds.append(F.normalize(D, dim=-1).float())
    C = (1.0 + torch.exp(torch.randn(num_frames, n, 1, generator=g))).float().to(dev)
    Q = torch.exp(1.0 + 0.5 * torch.randn(num_frames, n, 1, generator=g)).float().to(dev)
    return dict(Twc_gt=torch.stack(poses).float(), X=torch.stack(Xs), C=C, Q=Q, D=torch.stack(Ds), K=K.float().to(dev), h=h, w=w)
