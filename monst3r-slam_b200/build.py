"""In-tree build of the B200 backend (nvcc cross-compiles sm_100a without a GPU).

  lib/libm3r_b200.so            CUDA kernels + C ABI (include/m3r_abi.h); no torch, static cudart
  lib/mast3r_slam_backends.so   torch/pybind shim: the drop-in python module

Run as a script (`python monst3r-slam_b200/build.py [--force]`) or through
__graft_entry__.build().  Only sm_100a SASS is emitted — there is no other target.
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, os.environ.get("M3R_LIB_DIR", "lib"))  # M3R_LIB_DIR: side-by-side variant builds for A/B timing
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

CU_SOURCES = ["iter_proj.cu", "refine_matches.cu", "gn_accum.cu", "gn_solve.cu", "gn_solve_resident.cu", "gn_driver.cu", "edge_store.cu", "frame_track.cu", "peer_exchange.cu"]
HEADERS = ["common.cuh", "gn_common.cuh", "tma.cuh", "gn_device.cuh", "gn_tile.cuh"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr"]
NVCC_FLAGS += os.environ.get("M3R_NVCC_EXTRA", "").split()  # e.g. -DM3R_TRACK_PROF for tools/track_prof.py (then rebuild with force)
HOST_CXX = "/usr/bin/g++"


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd, log=None):
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log:
        with open(log, "w") as f:
            f.write(" ".join(cmd) + "\n" + res.stdout)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError("build failed: " + " ".join(cmd[:3]))
    return res.stdout


def build_lib(force=False, verbose=False):
    os.makedirs(LIB, exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in HEADERS] + [os.path.join(INCLUDE, "m3r_abi.h")]
    objs = []
    procs = []
    for src in CU_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(LIB, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _newer(o, [s] + hdrs):
            cmd = ["nvcc", "-ccbin", HOST_CXX] + NVCC_FLAGS + ["-c", s, "-o", o]
            log = open(os.path.join(LIB, src + ".ptxas.log"), "w")
            procs.append((src, subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT), log))
    for src, p, log in procs:
        rc = p.wait()
        log.close()
        if rc != 0:
            sys.stderr.write(open(os.path.join(LIB, src + ".ptxas.log")).read())
            raise RuntimeError("nvcc failed on " + src)
        if verbose:
            print(open(os.path.join(LIB, src + ".ptxas.log")).read())
    out = os.path.join(LIB, "libm3r_b200.so")
    if force or procs or _newer(out, objs):
        _run(["nvcc", "-ccbin", HOST_CXX, "-shared", "-o", out] + objs + ["-cudart", "static", "-ldl", "-lpthread"])
    return out


def build_bindings(force=False):
    import torch.utils.cpp_extension as cpp
    import torch
    os.makedirs(LIB, exist_ok=True)
    src = os.path.join(CSRC, "bindings.cpp")
    out = os.path.join(LIB, "mast3r_slam_backends.so")
    if not (force or _newer(out, [src, os.path.join(INCLUDE, "m3r_abi.h")])):
        return out
    inc = []
    for p in cpp.include_paths("cuda") + [sysconfig.get_paths()["include"]]:
        inc += ["-I", p]
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = [HOST_CXX, "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=mast3r_slam_backends",
           "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=1"] + inc + [
        "-I", "/usr/local/cuda/include", src, "-o", out, "-L", LIB, "-lm3r_b200", "-L", torch_lib,
        "-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-lc10", "-lc10_cuda", "-ltorch_python",
        "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + torch_lib]
    _run(cmd, log=os.path.join(LIB, "bindings.build.log"))
    return out


def build_all(force=False, verbose=False):
    a = build_lib(force, verbose)
    b = build_bindings(force)
    return a, b


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
