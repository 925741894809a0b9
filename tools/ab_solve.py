"""A/B timing of the multi-keyframe solve: gauss_newton_rays x10 on a graph with few points per keyframe (96x128), so
that gn_solve_kernel dominates.  usage: M3R_LIB_DIR=lib|lib_b python tools/ab_solve.py 32,120 [256,1000]"""
import importlib, os, statistics, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
pkg, abi, synth, matching, mdist = bench.load_pkg()
dev = torch.device("cuda", 0)
for spec in sys.argv[1:] or ["32,120"]:
    Ng, Eg = (int(x) for x in spec.split(","))
    g = synth.make_graph(Ng, Eg, 96, 128, seed=1003, device=dev, calib=False)
    n = 96 * 128
    prm = abi.make_params(abi.GN_RAYS, 0.003, 10.0, 0.0, 1.5, 10, 0.0)
    ws, _ = abi.gn_workspace(Ng, 2 * Eg, n, dev)
    Twc = g["Twc"].clone()
    ts = []
    for r in range(14):
        Twc.copy_(g["Twc"])
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        abi.gauss_newton(prm, Twc, g["Xs"], g["Cs"], None, g["ii"], g["jj"], g["idx"], g["valid"], g["Q"], workspace=ws)
        b.record(); b.synchronize()
        if r >= 4: ts.append(a.elapsed_time(b))
    err = float((Twc - g["Twc_gt"]).abs().max())
    print(os.environ.get("M3R_LIB_DIR", "lib"), spec, "GN x10: median %.3f ms (%.1f us / iteration)  max|T-gt| %.2e" % (statistics.median(ts), statistics.median(ts) * 100, err))
