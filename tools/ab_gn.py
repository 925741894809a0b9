"""A/B timing of the tracking-sized gauss_newton_rays (2 keyframes, 2 x 196 608 points, 10 iterations) through the C ABI:
median of 40 cold-L2 calls.  Run with M3R_LIB_DIR=lib / lib_b for two builds on the same box."""
import importlib, os, statistics, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
pkg, abi, synth, matching, mdist = bench.load_pkg()
dev = torch.device("cuda", 0)
inp = bench.make_tracking_inputs(synth, matching, dev, 1001)
prm = abi.make_params(abi.GN_RAYS, 0.003, 10.0, 0.0, 1.5, 10, 0.0)
ws, _ = abi.gn_workspace(2, 2, bench.N_PX, dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
Twc = inp["Twc0"].clone()
dx = torch.empty((1, 7), device=dev); info = torch.empty(4, dtype=torch.int32, device=dev)
ts = []
for r in range(50):
    Twc.copy_(inp["Twc0"]); flush.fill_(1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    abi.gauss_newton(prm, Twc, inp["Xs"], inp["Cs"], None, inp["ii"], inp["jj"], inp["idx"], inp["valid"], inp["Q"], workspace=ws, dx=dx, info=info)
    b.record(); b.synchronize()
    if r >= 10: ts.append(a.elapsed_time(b))
print(os.environ.get("M3R_LIB_DIR", "lib"), "GN x10: median %.4f min %.4f ms" % (statistics.median(ts), min(ts)), Twc[1].tolist()[:3])
