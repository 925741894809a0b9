"""A/B timing of the tracking-sized gauss_newton_rays x10 call (prepare + pack + persistent kernel), cold L2."""
import os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
pkg, abi, synth, matching, mdist = bench.load_pkg()
dev = torch.device("cuda", 0)
inp = bench.make_tracking_inputs(synth, matching, dev, 1001)
N, n, E2 = 2, bench.N_PX, 2
prm = abi.make_params(abi.GN_RAYS, 0.003, 10.0, 0.0, 1.5, 10, 0.0)
ws, _ = abi.gn_workspace(N, E2, n, dev)
Twc = inp["Twc0"].clone()
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
ts = []
for r in range(80):
    flush.fill_(1)
    Twc.copy_(inp["Twc0"])
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    abi.gauss_newton(prm, Twc, inp["Xs"], inp["Cs"], None, inp["ii"], inp["jj"], inp["idx"], inp["valid"], inp["Q"], workspace=ws)
    b.record(); b.synchronize()
    if r >= 10: ts.append(a.elapsed_time(b))
print(os.environ.get("M3R_LIB_DIR", "lib"), "gauss_newton_rays x10: median %.4f min %.4f ms" % (statistics.median(ts), min(ts)))
