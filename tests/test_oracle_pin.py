"""Pin the CPU oracle (oracle/cpu_oracle.cpp) against the REFERENCE's own code:
  * bit-for-bit against oracle/_ref/libref_matching_host.so (the reference's iter_proj /
    refine_matches kernel bodies compiled for the host by oracle/build_ref.sh), and
  * against tests/golden/*.npz produced by the reference's CUDA build on a B200
    (tests/golden/make_golden.py), within FMA-contraction noise.
"""
import glob
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, match_agreement, np_, pkg, pose_diff

CFG = pkg("matching").DEFAULT_CFG


def _pair(synth, h, w, seed, b=1):
    d = synth.make_matching_pair(h, w, b=b, seed=seed)
    m = pkg("matching")
    rays, pts, p0 = m.prep_for_iter_proj(d["X11"], d["X21"], None)
    return d, rays, pts, p0


def test_fp16_conversions_match_numpy(oracle):
    lib = oracle._lib()
    import ctypes as C
    lib.orc_h2f.restype = C.c_float
    lib.orc_f2h.restype = C.c_uint16
    lib.orc_f2h.argtypes = [C.c_float]
    lib.orc_h2f.argtypes = [C.c_uint16]
    bits = np.arange(0, 65536, 7, dtype=np.uint16)
    for hb in bits[:: 13]:
        f = np.array([hb], dtype=np.uint16).view(np.float16)[0]
        if np.isnan(f):
            continue
        assert lib.orc_h2f(int(hb)) == float(f)
    rng = np.random.default_rng(0)
    xs = np.concatenate([rng.standard_normal(2000) * s for s in (1e-8, 1e-5, 1e-3, 1.0, 100.0, 7e4)]).astype(np.float32)
    for x in xs:
        want = np.array([x], dtype=np.float32).astype(np.float16).view(np.uint16)[0]
        assert lib.orc_f2h(float(x)) == int(want), x


@pytest.mark.parametrize("init", ["identity", "previous"])
def test_iter_proj_port_equals_reference_host_build(oracle, synth, init):
    if not oracle.have_ref_host():
        pytest.skip("oracle/_ref/libref_matching_host.so not built (no /root/reference here)")
    d, rays, pts, p0 = _pair(synth, 48, 64, 11)
    if init == "previous":
        p0 = pkg("matching").lin_to_pixel(d["idx_init"], 64).float()
    args = (np_(rays), np_(pts), np_(p0), CFG["max_iter"], CFG["lambda_init"], CFG["convergence_thresh"])
    p_port, c_port = oracle.iter_proj(*args, impl="port")
    p_ref, c_ref = oracle.iter_proj(*args, impl="reference")
    # same compiler, same flags class: the port must reproduce the reference bit for bit up to the
    # host compiler's contraction choices; after .long() they must be identical everywhere
    same, worst = match_agreement(p_port.astype(np.int64), p_ref.astype(np.int64))
    assert same >= 0.999 and worst <= 1, (same, worst)
    assert (c_port == c_ref).mean() >= 0.999
    assert np.abs(p_port - p_ref).max() < 1e-2


def test_iter_proj_zero_iterations(oracle, synth):
    d, rays, pts, p0 = _pair(synth, 24, 32, 5)
    p, c = oracle.iter_proj(np_(rays), np_(pts), np_(p0) - 3.0, 0, 1e-8, 1e-6)
    assert not c.any()
    assert p[..., 0].min() >= 1.0 and p[..., 1].min() >= 1.0  # clamped even without iterations


@pytest.mark.parametrize("dtype", [np.float16, np.float32])
def test_refine_port_equals_reference_host_build(oracle, synth, dtype):
    if not oracle.have_ref_host():
        pytest.skip("oracle/_ref/libref_matching_host.so not built")
    d, rays, pts, p0 = _pair(synth, 48, 64, 12)
    h, w = 48, 64
    p1 = pkg("matching").lin_to_pixel(d["idx_init"], w)
    D11 = np_(d["D11"]).astype(dtype)
    D21 = np_(d["D21"]).reshape(1, h * w, -1).astype(dtype)
    a = oracle.refine_matches(D11, D21, np_(p1), 3, 5, impl="port")
    b = oracle.refine_matches(D11, D21, np_(p1), 3, 5, impl="reference")
    if dtype == np.float16:
        assert np.array_equal(a, b)  # integer/fp16 arithmetic: bit exact
    else:
        same, worst = match_agreement(a, b)
        assert same >= 0.999


def test_refine_known_answers(oracle):
    """planted descriptor is found; all-zero D11 never moves (score > 0 rule)."""
    rng = np.random.default_rng(3)
    h, w, f = 40, 48, 24
    D11 = rng.standard_normal((1, h, w, f)).astype(np.float32)
    D11 /= np.linalg.norm(D11, axis=-1, keepdims=True)
    start = np.stack([rng.integers(8, w - 8, 200), rng.integers(8, h - 8, 200)], -1)[None].astype(np.int64)
    off = rng.integers(-1, 2, (1, 200, 2)) * 5  # reachable at the first (d=5) level
    tgt = start + off
    D21 = D11[0, tgt[0, :, 1], tgt[0, :, 0]][None] * 4.0
    out = oracle.refine_matches(D11.astype(np.float16), D21.astype(np.float16), start, 3, 5)
    assert (np.abs(out - tgt).max(-1) == 0).mean() > 0.95
    out0 = oracle.refine_matches(np.zeros_like(D11, dtype=np.float16), D21.astype(np.float16), start, 3, 5)
    assert np.array_equal(out0, start)


def _golden_files(prefix):
    return sorted(glob.glob(os.path.join(GOLDEN, prefix + "*.npz")))


def test_golden_matching(oracle):
    files = _golden_files("matching_")
    if not files:
        pytest.skip("no golden matching fixtures yet (generate on the GPU box: tests/golden/make_golden.py)")
    for f in files:
        g = np.load(f)
        p, c = oracle.iter_proj(g["rays"], g["pts"], g["p_init"], int(g["max_iter"]), float(g["lambda_init"]),
                                float(g["cost_thresh"]))
        same, worst = match_agreement(p.astype(np.int64), g["p_new"].astype(np.int64))
        assert same >= 0.999 and worst <= 1, (f, same, worst)
        assert (c == g["converged"]).mean() >= 0.999
        r = oracle.refine_matches(g["D11"], g["D21"], g["p1"], int(g["radius"]), int(g["dilation_max"]))
        assert np.array_equal(r, g["p1_new"]), f


def test_golden_gn(oracle):
    files = _golden_files("gn_")
    if not files:
        pytest.skip("no golden GN fixtures yet")
    for f in files:
        g = np.load(f)
        kind = str(g["kind"])
        K = g["K"] if kind == "calib" else None
        Twc, dx, iters, nfail = oracle.gauss_newton(
            kind, g["Twc0"], g["Xs"], g["Cs"], K, g["ii"], g["jj"], g["idx"], g["valid"], g["Q"],
            float(g["sigma_a"]), float(g["sigma_b"]), float(g["C_thresh"]), float(g["Q_thresh"]),
            int(g["max_iter"]), float(g["delta_thresh"]), int(g["height"]), int(g["width"]),
            int(g["pixel_border"]), float(g["z_eps"]))
        rel, ang = pose_diff(Twc, g["Twc_ref"])
        assert rel < 1e-5 and ang < 1e-4, (f, rel, ang)


def test_double_reciprocal_rounded_to_float_equals_float_division():
    """iter_proj.cu replaces the reference's `float y = 1.0 / x` (double divide, rounded to float) by the IEEE
    single-precision division: double rounding is innocuous for division when the wide format has >= 2p+2
    digits.  Checked here on 20 M random bit patterns plus the special values."""
    rng = np.random.default_rng(7)
    bits = rng.integers(0, 2 ** 32, size=20_000_000, dtype=np.uint64).astype(np.uint32)
    x = np.concatenate([bits.view(np.float32), np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 1e-45, -1e-45, 1.1754944e-38,
                                                         3.4028235e38, 1.0, 3.0, 1e-6, 0.99999994], dtype=np.float32)])
    with np.errstate(all="ignore"):
        a = (1.0 / x.astype(np.float64)).astype(np.float32)
        b = np.float32(1.0) / x
    same = (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))
    assert same.all(), x[~same][:10]
