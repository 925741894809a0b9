"""Shared test plumbing.  `-m "not gpu"`: oracle vs reference host build / golden fixtures, host
logic, C-ABI load + symbol checks (no GPU needed).  `-m gpu`: parity tests proper, through the C ABI.
Only tests (and smoke / bench's cpu_baseline leg) may import oracle/.
"""
import importlib
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pkg(name=None):
    return importlib.import_module("monst3r-slam_b200" + ("." + name if name else ""))


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as orc
    orc._lib()
    return orc


@pytest.fixture(scope="session")
def synth():
    return pkg("synth")


@pytest.fixture(scope="session")
def abi():
    return pkg("abi")


@pytest.fixture(scope="session")
def backends():
    return pkg("backends")


@pytest.fixture(scope="session")
def ref_backends():
    """The reference's own CUDA build (oracle/_ref/ref_backends.so, built by oracle/build_ref.sh)."""
    path = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.exists(os.path.join(path, "ref_backends.so")):
        pytest.skip("oracle/_ref/ref_backends.so not built")
    if path not in sys.path:
        sys.path.insert(0, path)
    return importlib.import_module("ref_backends")


def need_gpu():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")


def np_(t):
    return t.detach().cpu().numpy()


def match_agreement(p_a, p_b):
    """fraction of identical integer pixel pairs and the max L-inf disagreement"""
    a = np.asarray(p_a).reshape(-1, 2).astype(np.int64)
    b = np.asarray(p_b).reshape(-1, 2).astype(np.int64)
    d = np.abs(a - b).max(axis=1)
    return float((d == 0).mean()), int(d.max()) if d.size else 0


def pose_diff(Ta, Tb):
    """(max relative translation/scale error, max rotation angle [rad]) between [N,8] pose arrays"""
    Ta = np.asarray(Ta, dtype=np.float64)
    Tb = np.asarray(Tb, dtype=np.float64)
    scale = np.maximum(1.0, np.linalg.norm(Tb[:, :3], axis=1))
    dt = np.linalg.norm(Ta[:, :3] - Tb[:, :3], axis=1) / scale
    ds = np.abs(Ta[:, 7] - Tb[:, 7]) / np.abs(Tb[:, 7])
    qa = Ta[:, 3:7] / np.linalg.norm(Ta[:, 3:7], axis=1, keepdims=True)
    qb = Tb[:, 3:7] / np.linalg.norm(Tb[:, 3:7], axis=1, keepdims=True)
    dot = np.clip(np.abs((qa * qb).sum(1)), 0.0, 1.0)
    ang = 2.0 * np.arccos(dot)
    return float(max(dt.max(), ds.max())), float(ang.max())
