#!/usr/bin/env python
"""Phase breakdown of gn_solve_kernel<coop> (needs a -DM3R_TRACK_PROF build): cycles/16 of CTA 0 per phase."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
dev = torch.device("cuda", 0)
pkg, abi, synth, matching, mdist = bench.load_pkg()
Ng, Eg = (int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "256,1000").split(","))
g = synth.make_graph(Ng, Eg, 96, 128, seed=1003, device=dev, calib=False)
n = 96 * 128
prm = abi.make_params(abi.GN_RAYS, 0.003, 10.0, 0.0, 1.5, 1, 0.0)
ws, _ = abi.gn_workspace(Ng, 2 * Eg, n, dev)
for rep in range(2):
    Twc = g["Twc"].clone()
    abi.gauss_newton(prm, Twc, g["Xs"], g["Cs"], None, g["ii"], g["jj"], g["idx"], g["valid"], g["Q"], workspace=ws)
    torch.cuda.synchronize()
ctrl = ws[:64].view(torch.int32).cpu().tolist()
names = ["diag load+potrf+invert", "panel TRSM", "barrier 1", "trailing update", "barrier 2", "back-substitution"]
tot = sum(ctrl[6:12])
for k, nm in enumerate(names):
    print("%-24s %8.1f us  %5.1f%%" % (nm, ctrl[6 + k] * 16 / 1.965e3, 100.0 * ctrl[6 + k] / max(tot, 1)))
print("total %.1f us" % (tot * 16 / 1.965e3))
