"""The C-ABI library loads without a GPU and exports every symbol include/m3r_abi.h declares; the
pybind drop-in imports under the reference's module name and carries the five op names."""
import ctypes
import os
import re
import sys

import pytest
import torch

from conftest import ROOT, pkg


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "m3r_abi.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(m3r_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(abi):
    syms = _declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(abi.lib, s), s
    assert set(syms) == set(abi.EXPORTS)
    assert abi.lib.m3r_abi_version() == 4


def test_workspace_size_is_monotone(abi):
    a = abi.lib.m3r_gn_workspace_bytes(2, 2, 196608)
    b = abi.lib.m3r_gn_workspace_bytes(32, 240, 196608)
    c = abi.lib.m3r_gn_workspace_bytes(256, 2000, 196608)
    # c4 on ONE rank: 2000 edges x 196608 pts x 8 B packed records = 3.1 GB of the 180 GB
    assert 0 < a < b < c < 4 * 2 ** 30
    # a rank that owns 1/8 of the edges keeps 1/8 of the packed slabs
    d = abi.lib.m3r_gn_workspace_bytes_sharded(256, 2000, 196608, 250)
    assert 0 < d < c / 4
    assert abi.lib.m3r_gn_workspace_bytes(1, 2, 10) == 0


def test_bad_arguments_return_error_codes(abi):
    assert abi.lib.m3r_iter_proj(None, None, None, None, None, 1, 8, 8, 4, 1, ctypes.c_float(0), ctypes.c_float(0), None) == -1
    assert abi.lib.m3r_refine_matches_f16(None, None, None, None, 1, 8, 8, 4, 24, 3, 5, None) == -1


def test_drop_in_module_has_reference_surface():
    mod = pkg().load_drop_in()
    for name in ("iter_proj", "refine_matches", "gauss_newton_points", "gauss_newton_rays", "gauss_newton_calib"):
        assert callable(getattr(mod, name))
    assert mod.__name__ == "mast3r_slam_backends"


def test_drop_in_rejects_cpu_and_noncontiguous_tensors():
    mod = pkg().load_drop_in()
    rays = torch.zeros(1, 8, 8, 9)
    with pytest.raises(RuntimeError, match="CUDA"):
        mod.iter_proj(rays, torch.zeros(1, 4, 3), torch.zeros(1, 4, 2), 1, 1e-8, 1e-6)
    with pytest.raises(RuntimeError, match="must be contiguous"):
        mod.iter_proj(rays.permute(0, 2, 1, 3), torch.zeros(1, 4, 3), torch.zeros(1, 4, 2), 1, 1e-8, 1e-6)


def test_reference_matching_py_imports_with_our_module():
    """matching.py of the reference is importable unmodified once our module is on sys.path (only
    in the build container: /root/reference does not exist on the GPU box)."""
    ref = "/root/reference/MASt3R-SLAM"
    if not os.path.isdir(ref):
        pytest.skip("reference not mounted")
    pkg().load_drop_in()
    sys.path.insert(0, ref)
    try:
        import importlib
        m = importlib.import_module("mast3r_slam.matching")
        assert m.mast3r_slam_backends.__m3r_b200__
    finally:
        sys.path.remove(ref)
