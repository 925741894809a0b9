"""m3r_iter_proj_static: the allocation-free entry point (library-owned `static __device__` scratch), through the C ABI.

No counterpart in the reference (include/m3r_abi.h).  Checked against
the existing entry points on the same inputs: m3r_iter_proj_ws runs the same two kernels on a caller-provided scratch
(bit-identical expected) and the plain m3r_iter_proj within the tolerance test_gpu_matching.py uses for the padded
path (identical pixels and flags, floats within 1e-4 px).  Two calls in a row share the array; the second one is a
smaller problem, so stale contents of the first call lie behind it.  Collected last on purpose (pytest -x).
"""
import ctypes as C

import pytest
import torch

from conftest import need_gpu, pkg

pytestmark = pytest.mark.gpu


def _prep(synth, h, w, seed, b, init):
    d = synth.make_matching_pair(h, w, b=b, seed=seed, device="cuda")
    return pkg("matching").prep_for_iter_proj(d["X11"], d["X21"], None if init == "identity" else d["idx_init"])


def _check(abi, rays, pts, p0):
    b, n = p0.shape[0], p0.shape[1]
    p_new = torch.full((b, n, 2), -7.0, dtype=torch.float32, device="cuda")
    conv = torch.zeros((b, n), dtype=torch.bool, device="cuda")
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    abi.iter_proj_static(rays, pts, p0, p_new, conv, 10, 1e-8, 1e-6)
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before  # no torch allocation on the call path
    torch.cuda.synchronize()
    ws = abi.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6, padded=True)
    plain = abi.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6, padded=False)
    assert torch.equal(p_new, ws[0]) and torch.equal(conv, ws[1])
    assert torch.equal(p_new.long(), plain[0].long()) and torch.equal(conv, plain[1])
    assert (p_new - plain[0]).abs().max().item() < 1e-4


def test_iter_proj_static_two_calls_in_a_row(abi, synth):
    need_gpu()
    assert int(abi.lib.m3r_iter_proj_static_max_texels()) == 512 * 384
    rays, pts, p0 = _prep(synth, 384, 512, 1001, 1, "identity")        # exactly the maximum: one 512x384 frame
    _check(abi, rays, pts, p0)
    rays2, pts2, p02 = _prep(synth, 96, 128, 33, 2, "previous")       # smaller, batch 2, ragged n, right after
    _check(abi, rays2, pts2[:, :1001].contiguous(), p02[:, :1001].contiguous())
    _check(abi, rays, pts, p0)                                          # and the first one again


def test_iter_proj_static_rejects_oversize(abi, synth):
    need_gpu()
    rays, pts, p0 = _prep(synth, 96, 128, 5, 1, "identity")
    n = p0.shape[1]
    p_new = torch.empty((1, n, 2), dtype=torch.float32, device="cuda")
    conv = torch.empty((1, n), dtype=torch.bool, device="cuda")
    # claim a 513x384 image: one row of texels more than the array holds; rejected before anything is enqueued
    rc = abi.lib.m3r_iter_proj_static(abi._ptr(rays), abi._ptr(pts), abi._ptr(p0), abi._ptr(p_new), abi._ptr(conv),
                                      1, 384, 513, n, 10, C.c_float(1e-8), C.c_float(1e-6), abi._stream())
    assert rc == -2
    msg = abi.lib.m3r_last_error().decode()
    assert "exceeds the static scratch" in msg and str(384 * 513) in msg, msg
    torch.cuda.synchronize()
