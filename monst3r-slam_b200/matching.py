"""Mirror of the reference's matching glue (mast3r_slam/matching.py:8-90, image.py:5-38) on top of
the B200 ops.  Same function names, arguments and return values; the config dict is passed in
instead of read from the reference's global (defaults = config/base.yaml:15-21).  The torch glue
(normalise, Scharr-like gradients, occlusion test) is unchanged reference behaviour; `match(...,
fused=True)` is the additive fused path of SURVEY §8f-1 (csrc/iter_proj.cu: m3r_match_fused).
"""
import torch
import torch.nn.functional as F

DEFAULT_CFG = dict(max_iter=10, lambda_init=1e-8, convergence_thresh=1e-6, dist_thresh=1e-1, radius=3,
                   dilation_max=5)


def img_gradient(img):
    """image.py:5-38: 3x3 Scharr-like gradients (1/32 * [3,10,3]) with reflect padding, per channel."""
    c = img.shape[1]
    k = torch.tensor([[-3.0, 0.0, 3.0], [-10.0, 0.0, 10.0], [-3.0, 0.0, 3.0]], device=img.device,
                     dtype=img.dtype) / 32.0
    gx_k = k.repeat(c, 1, 1, 1)
    gy_k = k.t().contiguous().repeat(c, 1, 1, 1)
    pad = F.pad(img, (1, 1, 1, 1), mode="reflect")
    return F.conv2d(pad, gx_k, groups=c), F.conv2d(pad, gy_k, groups=c)


def pixel_to_lin(p1, w):
    return p1[..., 0] + (w * p1[..., 1])


def lin_to_pixel(idx_1_to_2, w):
    return torch.stack((idx_1_to_2 % w, idx_1_to_2 // w), dim=-1)


def prep_for_iter_proj(X11, X21, idx_1_to_2_init):
    """matching.py:25-49"""
    b, h, w, _ = X11.shape
    rays_img = F.normalize(X11, dim=-1).permute(0, 3, 1, 2)
    gx_img, gy_img = img_gradient(rays_img)
    rays_with_grad_img = torch.cat((rays_img, gx_img, gy_img), dim=1).permute(0, 2, 3, 1).contiguous()
    pts3d_norm = F.normalize(X21.view(b, -1, 3), dim=-1)
    if idx_1_to_2_init is None:
        idx_1_to_2_init = torch.arange(h * w, device=X11.device)[None, :].repeat(b, 1)
    p_init = lin_to_pixel(idx_1_to_2_init, w).float()
    return rays_with_grad_img, pts3d_norm, p_init


def match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init=None, cfg=None, ops=None):
    """matching.py:52-90.  `ops` is any module exposing iter_proj / refine_matches (default: the C-ABI
    mirror in backends.py; the pybind drop-in or the reference's own build work too)."""
    if ops is None:
        from . import backends as ops
    cfg = dict(DEFAULT_CFG, **(cfg or {}))
    b, h, w = X21.shape[:3]
    rays_with_grad_img, pts3d_norm, p_init = prep_for_iter_proj(X11, X21, idx_1_to_2_init)
    p1, valid_proj2 = ops.iter_proj(rays_with_grad_img, pts3d_norm, p_init, cfg["max_iter"],
                                    cfg["lambda_init"], cfg["convergence_thresh"])
    p1 = p1.long()
    batch_inds = torch.arange(b, device=X11.device)[:, None].repeat(1, h * w)
    dists2 = torch.linalg.norm(X11[batch_inds, p1[..., 1], p1[..., 0], :].reshape(b, h, w, 3) - X21, dim=-1)
    valid_proj2 = valid_proj2 & (dists2 < cfg["dist_thresh"]).view(b, -1)
    if cfg["radius"] > 0:
        (p1,) = ops.refine_matches(D11.half(), D21.view(b, h * w, -1).half(), p1, cfg["radius"],
                                   cfg["dilation_max"])
    return pixel_to_lin(p1, w), valid_proj2.unsqueeze(-1)


def match(X11, X21, D11, D21, idx_1_to_2_init=None, cfg=None, ops=None, fused=False):
    """matching.py:8-10.  fused=True (additive): the whole function runs as five launches of
    m3r_match_fused instead of ~20 torch kernels around the two ops; same outputs within the parity bar
    (the torch glue's normalise / conv arithmetic is not reproduced bit for bit)."""
    if fused:
        from . import abi
        cfg = dict(DEFAULT_CFG, **(cfg or {}))
        b, h, w = X21.shape[:3]
        return abi.match_fused(X11.contiguous(), X21.contiguous(), D11.contiguous(), D21.reshape(b, h * w, -1).contiguous(),
                               idx_1_to_2_init, cfg["max_iter"], cfg["lambda_init"], cfg["convergence_thresh"],
                               cfg["dist_thresh"], cfg["radius"], cfg["dilation_max"])
    return match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init, cfg, ops)
