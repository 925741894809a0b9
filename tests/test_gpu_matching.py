"""GPU parity of iter_proj / refine_matches (through the C ABI) against the CPU oracle and, when
oracle/_ref/ref_backends.so is present, against the reference's own CUDA kernels on the same box.
Bar (BASELINE.json): indices / masks identical on >= 99.9 % of pixels, every disagreement <= 1 px;
fp16 refine is integer/fp16 arithmetic and must be bit exact.
"""
import numpy as np
import pytest
import torch

from conftest import match_agreement, need_gpu, np_, pkg

pytestmark = pytest.mark.gpu
CFG = pkg("matching").DEFAULT_CFG


def _prep(synth, h, w, seed, b=1, init="identity"):
    d = synth.make_matching_pair(h, w, b=b, seed=seed, device="cuda")
    m = pkg("matching")
    rays, pts, p0 = m.prep_for_iter_proj(d["X11"], d["X21"], None if init == "identity" else d["idx_init"])
    return d, rays, pts, p0


@pytest.mark.parametrize("init", ["identity", "previous"])
def test_iter_proj_vs_oracle(backends, oracle, synth, init):
    need_gpu()
    d, rays, pts, p0 = _prep(synth, 96, 128, 21, b=2, init=init)
    p, c = backends.iter_proj(rays, pts, p0, CFG["max_iter"], CFG["lambda_init"], CFG["convergence_thresh"])
    po, co = oracle.iter_proj(np_(rays), np_(pts), np_(p0), CFG["max_iter"], CFG["lambda_init"], CFG["convergence_thresh"])
    same, worst = match_agreement(np_(p.long()), po.astype(np.int64))
    assert same >= 0.999 and worst <= 1, (same, worst)
    assert (np_(c) == co).mean() >= 0.999
    assert np.abs(np_(p) - po).max() < 0.05


@pytest.mark.parametrize("b", [1, 2])
def test_iter_proj_vs_reference_cuda_full_size(backends, ref_backends, synth, b):
    need_gpu()
    d, rays, pts, p0 = _prep(synth, 384, 512, 1001, b=b)
    args = (rays, pts, p0, CFG["max_iter"], CFG["lambda_init"], CFG["convergence_thresh"])
    p, c = backends.iter_proj(*args)
    pr, cr = ref_backends.iter_proj(*args)
    same, worst = match_agreement(np_(p.long()), np_(pr.long()))
    bitwise = (p == pr).all(-1).float().mean().item()
    print(f"iter_proj vs reference CUDA: long-identical {same:.6f}, worst {worst}px, bitwise {bitwise:.6f}")
    assert same >= 0.999 and worst <= 1
    assert (c == cr).float().mean().item() >= 0.999


def test_iter_proj_padded_equals_plain_kernel(abi, synth):
    """the texel-plane path (m3r_iter_proj_ws) and the plain kernel run the same arithmetic on the same
    values (only the compiler's FMA contraction may differ): identical pixels and flags, sub-1e-4 px floats;
    identity and previous-frame inits, batch 2, ragged n (scalar tail of the re-pack)"""
    need_gpu()
    for init, n, hw in (("identity", None, (96, 128)), ("previous", None, (96, 128)), ("identity", 1001, (96, 128)),
                        ("identity", None, (37, 53))):
        d, rays, pts, p0 = _prep(synth, hw[0], hw[1], 33, b=2, init=init)
        if n:
            pts, p0 = pts[:, :n].contiguous(), p0[:, :n].contiguous()
        a = abi.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6, padded=True)
        c = abi.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6, padded=False)
        assert torch.equal(a[0].long(), c[0].long()) and torch.equal(a[1], c[1])
        assert (a[0] - c[0]).abs().max().item() < 1e-4
        assert (a[0] != c[0]).any(-1).float().mean().item() < 1e-2


def test_iter_proj_ragged_n_and_zero_iter(backends, oracle, synth):
    need_gpu()
    d, rays, pts, p0 = _prep(synth, 48, 64, 5)
    n = 1001  # not a multiple of any block size; the reference would read out of bounds here
    p, c = backends.iter_proj(rays, pts[:, :n].contiguous(), p0[:, :n].contiguous(), 10, 1e-8, 1e-6)
    po, co = oracle.iter_proj(np_(rays), np_(pts[:, :n]), np_(p0[:, :n]), 10, 1e-8, 1e-6)
    assert p.shape == (1, n, 2)
    assert match_agreement(np_(p.long()), po.astype(np.int64))[0] >= 0.999
    p, c = backends.iter_proj(rays, pts, p0 - 5.0, 0, 1e-8, 1e-6)
    assert not c.any() and p.min() >= 1.0


def test_iter_proj_identity_known_answer(backends, synth):
    """X21 = X11 and identity init -> identity indices, converged."""
    need_gpu()
    d = synth.make_matching_pair(96, 128, seed=3, device="cuda", flow_amp=0.0)
    m = pkg("matching")
    rays, pts, p0 = m.prep_for_iter_proj(d["X11"], d["X11"], None)
    p, c = backends.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
    inner = (p0[..., 0] >= 1) & (p0[..., 0] <= 126) & (p0[..., 1] >= 1) & (p0[..., 1] <= 94)
    assert (p.round().long() == p0.long())[inner].all()
    assert c[inner].float().mean() > 0.999


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32, torch.float64])
def test_refine_vs_oracle(backends, oracle, synth, dtype):
    need_gpu()
    h, w = 96, 128
    d, rays, pts, p0 = _prep(synth, h, w, 22, b=2)
    p1 = pkg("matching").lin_to_pixel(d["idx_init"], w)
    D11 = d["D11"].to(dtype)
    D21 = d["D21"].view(2, h * w, -1).to(dtype)
    (out,) = backends.refine_matches(D11, D21, p1, 3, 5)
    want = oracle.refine_matches(np_(D11), np_(D21), np_(p1), 3, 5)
    if dtype == torch.float16:
        assert np.array_equal(np_(out), want)
    else:
        same, worst = match_agreement(np_(out), want)
        assert same >= 0.999


def test_refine_borders_zero_descriptors_generic_fdim(backends, oracle):
    need_gpu()
    g = torch.Generator().manual_seed(5)
    h, w = 40, 56
    for fdim in (24, 16):
        D11 = torch.randn(1, h, w, fdim, generator=g).half().cuda()
        D21 = torch.randn(1, 500, fdim, generator=g).half().cuda()
        # starts on and next to every border
        u = torch.randint(0, w, (500,), generator=g)
        v = torch.randint(0, h, (500,), generator=g)
        u[:50] = 0; u[50:100] = w - 1; v[100:150] = 0; v[150:200] = h - 1
        p1 = torch.stack((u, v), -1)[None].cuda()
        (out,) = backends.refine_matches(D11, D21, p1, 3, 5)
        want = oracle.refine_matches(np_(D11), np_(D21), np_(p1), 3, 5)
        assert np.array_equal(np_(out), want), fdim
        assert out[..., 0].min() >= 0 and out[..., 0].max() < w and out[..., 1].max() < h
        (out0,) = backends.refine_matches(torch.zeros_like(D11), D21, p1, 3, 5)
        assert torch.equal(out0, p1)
    for radius, dil in ((1, 1), (2, 3), (3, 0), (4, 2)):
        (out,) = backends.refine_matches(D11, D21, p1, radius, dil)
        want = oracle.refine_matches(np_(D11), np_(D21), np_(p1), radius, dil)
        assert np.array_equal(np_(out), want), (radius, dil)


def _staged_case(h, w, kind, seed, b=1, quant=False):
    """n == h*w queries (the staged shared-memory path) with start pixels of a given structure"""
    g = torch.Generator().manual_seed(seed)
    D11 = torch.randn(b, h, w, 24, generator=g)
    D11 = torch.nn.functional.avg_pool2d(D11.permute(0, 3, 1, 2), 5, 1, 2).permute(0, 2, 3, 1).contiguous()
    D11 = D11 / D11.norm(dim=-1, keepdim=True)
    if quant:  # few distinct values -> many exactly tied scores: the scan order decides
        D11 = (D11 * 2).round() / 2
    vv, uu = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    uu, vv = uu.float(), vv.float()
    if kind == "shift":
        pu, pv = uu + 7, vv - 4
    elif kind == "zoom":  # 1.7x magnification about the centre: boxes exceed one pass
        pu, pv = (uu - w / 2) * 1.7 + w / 2, (vv - h / 2) * 1.7 + h / 2
    elif kind == "shrink":
        pu, pv = (uu - w / 2) * 0.3 + w / 2, (vv - h / 2) * 0.3 + h / 2
    elif kind == "random":
        pu, pv = torch.rand(h, w, generator=g) * w, torch.rand(h, w, generator=g) * h
    elif kind == "mixed":  # smooth flow + 8 % far outliers + a few out-of-image / absurd starts
        pu = uu + 5 * torch.sin(vv / 9) + 3
        pv = vv + 4 * torch.cos(uu / 11) - 2
        out = torch.rand(h, w, generator=g) < 0.08
        pu = torch.where(out, torch.rand(h, w, generator=g) * w, pu)
        pv = torch.where(out, torch.rand(h, w, generator=g) * h, pv)
    p1 = torch.stack((pu, pv), -1).round().long().view(1, h * w, 2).repeat(b, 1, 1)
    if kind == "mixed" and h * w > 303:
        p1[:, 5] = torch.tensor([-3, 2]); p1[:, 77] = torch.tensor([w + 2, h + 9]); p1[:, 300] = torch.tensor([-100000, 5])
        p1[:, 301] = torch.tensor([3, 2 ** 20]); p1[:, 302] = torch.tensor([w - 1, h - 1]); p1[:, 303] = torch.tensor([0, 0])
    elif kind != "mixed":
        p1[..., 0].clamp_(0, w - 1); p1[..., 1].clamp_(0, h - 1)
    qsrc = p1.clone(); qsrc[..., 0].clamp_(0, w - 1); qsrc[..., 1].clamp_(0, h - 1)
    jit = torch.randint(-6, 7, qsrc.shape, generator=g)
    qsrc = qsrc + jit; qsrc[..., 0].clamp_(0, w - 1); qsrc[..., 1].clamp_(0, h - 1)
    bi = torch.arange(b).view(b, 1).expand(b, h * w)
    D21 = D11[bi, qsrc[..., 1], qsrc[..., 0]] + 0.02 * torch.randn(b, h * w, 24, generator=g)
    return D11.half().cuda(), D21.half().cuda().contiguous(), p1.cuda().contiguous()


@pytest.mark.parametrize("kind,h,w,dil,quant", [
    ("shift", 64, 96, 5, False), ("zoom", 64, 96, 5, False), ("shrink", 64, 96, 5, False),
    ("random", 64, 96, 5, False), ("mixed", 64, 96, 5, False), ("mixed", 50, 70, 5, True),
    ("shift", 33, 47, 8, True), ("zoom", 50, 70, 1, True), ("mixed", 16, 16, 3, False), ("shift", 7, 5, 5, False)])
def test_refine_staged_path_bit_exact(backends, oracle, kind, h, w, dil, quant):
    """every route through refine_staged_kernel (single pass, column passes, robust inlier box,
    global-gather outliers, border checks, partial tiles) against the oracle, bit for bit"""
    need_gpu()
    D11, D21, p1 = _staged_case(h, w, kind, seed=31 + h + dil, b=2, quant=quant)
    (out,) = backends.refine_matches(D11, D21, p1, 3, dil)
    want = oracle.refine_matches(np_(D11), np_(D21), np_(p1), 3, dil)
    bad = (np_(out) != want).any(-1)
    assert not bad.any(), (kind, int(bad.sum()), np.argwhere(bad)[:5], np_(out)[bad][:5], want[bad][:5])


def test_refine_staged_equals_gather_full_size(backends):
    """512x384: the staged kernel (n == h*w) and the global-gather kernel (same queries presented as
    two half-length lists, n != h*w) must agree exactly"""
    need_gpu()
    h, w = 384, 512
    D11, D21, p1 = _staged_case(h, w, "mixed", seed=99)
    (a,) = backends.refine_matches(D11, D21, p1, 3, 5)
    n2 = h * w // 2
    (b0,) = backends.refine_matches(D11, D21[:, :n2].contiguous(), p1[:, :n2].contiguous(), 3, 5)
    (b1,) = backends.refine_matches(D11, D21[:, n2:].contiguous(), p1[:, n2:].contiguous(), 3, 5)
    assert torch.equal(a, torch.cat((b0, b1), 1))


def test_refine_vs_reference_cuda_full_size(backends, ref_backends, synth):
    need_gpu()
    h, w = 384, 512
    d, rays, pts, p0 = _prep(synth, h, w, 1001)
    p1 = pkg("matching").lin_to_pixel(d["idx_init"], w)
    D11 = d["D11"].half()
    D21 = d["D21"].view(1, h * w, -1).half()
    (out,) = backends.refine_matches(D11, D21, p1, 3, 5)
    (want,) = ref_backends.refine_matches(D11, D21, p1, 3, 5)
    assert torch.equal(out, want)


def test_match_pipeline_vs_reference_cuda(backends, ref_backends, synth):
    """matching.match end to end (torch glue + both ops), ours vs the reference's kernels."""
    need_gpu()
    m = pkg("matching")
    d = synth.make_matching_pair(384, 512, seed=1001, device="cuda")
    for init in (None, d["idx_init"]):
        idx_a, val_a = m.match(d["X11"], d["X21"], d["D11"], d["D21"], init, ops=backends)
        idx_b, val_b = m.match(d["X11"], d["X21"], d["D11"], d["D21"], init, ops=ref_backends)
        pa, pb = m.lin_to_pixel(idx_a, 512), m.lin_to_pixel(idx_b, 512)
        same, worst = match_agreement(np_(pa), np_(pb))
        assert same >= 0.999, same
        assert (val_a == val_b).float().mean().item() >= 0.999
        # sanity: the matches are close to the planted flow where valid
        gt_u, gt_v = d["pu"].view(1, -1), d["pv"].view(1, -1)
        err = torch.maximum((pa[..., 0] - gt_u).abs(), (pa[..., 1] - gt_v).abs())[val_a[..., 0]]
        assert err.median() <= 1.5


@pytest.mark.parametrize("h,w,b,half", [(96, 128, 2, False), (384, 512, 1, False), (50, 70, 1, True)])
def test_match_fused_vs_torch_glue(backends, synth, h, w, b, half):
    """m3r_match_fused (prep + LM with fused occlusion test + fp16 conversion + refine + pixel_to_lin) against
    matching.match with the torch glue around the same two ops: BASELINE bar (>= 99.9 % identical, <= 1 px),
    identity and previous-frame initialisation, partial tiles, fp32 and fp16 descriptors, radius 0"""
    need_gpu()
    m = pkg("matching")
    d = synth.make_matching_pair(h, w, b=b, seed=11 + h, device="cuda")
    D11, D21 = (d["D11"].half(), d["D21"].half()) if half else (d["D11"], d["D21"])
    for init in (None, d["idx_init"]):
        for cfg in (None, {"radius": 0}):
            idx_a, val_a = m.match(d["X11"], d["X21"], D11, D21, init, cfg=cfg, fused=True)
            idx_b, val_b = m.match(d["X11"], d["X21"], D11, D21, init, cfg=cfg, ops=backends)
            assert idx_a.shape == idx_b.shape and val_a.shape == val_b.shape and val_a.dtype == torch.bool
            same, worst = match_agreement(np_(m.lin_to_pixel(idx_a, w)), np_(m.lin_to_pixel(idx_b, w)))
            agree_valid = (val_a == val_b).float().mean().item()
            assert same >= 0.999 and worst <= 1, (same, worst, init is None, cfg)
            assert agree_valid >= 0.999, agree_valid


def test_drop_in_module_equals_c_abi(backends, synth):
    need_gpu()
    mod = pkg().load_drop_in()
    d, rays, pts, p0 = _prep(synth, 96, 128, 9)
    a = backends.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
    b = mod.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1]) and b[1].dtype == torch.bool
    p1 = a[0].long()
    D11, D21 = d["D11"].half(), d["D21"].view(1, 96 * 128, -1).half()
    assert torch.equal(backends.refine_matches(D11, D21, p1, 3, 5)[0], mod.refine_matches(D11, D21, p1, 3, 5)[0])
    fa = pkg("matching").match(d["X11"], d["X21"], d["D11"], d["D21"], None, fused=True)
    fb = mod.match_fused(d["X11"], d["X21"], d["D11"], d["D21"].view(1, 96 * 128, -1))
    assert torch.equal(fa[0], fb[0]) and torch.equal(fa[1], fb[1])
