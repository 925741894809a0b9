"""monst3r-slam_b200 — B200 (sm_100a) geometric backend for MASt3R-/MonST3R-SLAM.

The directory name carries a hyphen (it mirrors the reference repo's name), so import it with
`importlib.import_module("monst3r-slam_b200")` or put this directory's `lib/` on sys.path and
`import mast3r_slam_backends` — the drop-in pybind module the reference's python code expects.

Submodules:
  abi       ctypes binding of lib/libm3r_b200.so (C ABI, include/m3r_abi.h)
  backends  the five reference ops (same names / positional signatures) through the C ABI
  matching  mirror of mast3r_slam/matching.py on top of the ops
  dist      edge partitioner + NCCL communicator setup for multi-GPU global optimisation
  synth     synthetic inputs
  build     in-tree nvcc build

Nothing here imports oracle/ — the CPU oracle is test infrastructure only.
"""
import os

PACKAGE_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_DIR = os.path.join(PACKAGE_DIR, "lib")


def load_drop_in():
    """Import and return the pybind module `mast3r_slam_backends` built in lib/."""
    import importlib
    import sys
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    if LIB_DIR not in sys.path:
        sys.path.insert(0, LIB_DIR)
    mod = importlib.import_module("mast3r_slam_backends")
    if not getattr(mod, "__m3r_b200__", False):
        raise ImportError("a different mast3r_slam_backends shadows the B200 build: " + str(mod.__file__))
    return mod
