#!/usr/bin/env python
"""Phase breakdown of gn_solve_df_kernel as seen by the chain CTA (needs a -DM3R_SOLVE_PROF build in lib_prof):
M3R_LIB_DIR=lib_prof python tools/solve_prof2.py 32,120 256,1000"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
dev = torch.device("cuda", 0)
pkg, abi, synth, matching, mdist = bench.load_pkg()
for spec in sys.argv[1:] or ["256,1000"]:
    Ng, Eg = (int(x) for x in spec.split(","))
    g = synth.make_graph(Ng, Eg, 96, 128, seed=1003, device=dev, calib=False)
    n = 96 * 128
    prm = abi.make_params(abi.GN_RAYS, 0.003, 10.0, 0.0, 1.5, 1, 0.0)
    ws, _ = abi.gn_workspace(Ng, 2 * Eg, n, dev)
    for rep in range(3):
        Twc = g["Twc"].clone()
        abi.gauss_newton(prm, Twc, g["Xs"], g["Cs"], None, g["ii"], g["jj"], g["idx"], g["valid"], g["Q"], workspace=ws)
        torch.cuda.synchronize()
