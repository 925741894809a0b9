"""A/B timing of iter_proj (re-pack + LM kernel, workspace path) at 512x384 on one box: median of 50 cold-L2 calls.
usage: M3R_LIB_DIR=lib|lib_x python tools/ab_iterproj.py"""
import importlib, os, statistics, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
abi = importlib.import_module("monst3r-slam_b200.abi")
synth = importlib.import_module("monst3r-slam_b200.synth")
matching = importlib.import_module("monst3r-slam_b200.matching")
d = synth.make_matching_pair(384, 512, seed=1001, device="cuda")
rays, pts, p0 = matching.prep_for_iter_proj(d["X11"], d["X21"], None)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
ts = []
ref = None
for r in range(60):
    flush.fill_(1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); out = abi.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6, padded=True); b.record(); b.synchronize()
    if r >= 10: ts.append(a.elapsed_time(b))
print(os.environ.get("M3R_LIB_DIR", "lib"), "iter_proj: median %.4f min %.4f ms  checksum %.6f %d" % (
    statistics.median(ts), min(ts), float(out[0].double().sum()), int(out[1].sum())))
