#!/usr/bin/env python
"""Small, fixed launch sequence for ncu: `reps` tracking steps (iter_proj, refine_matches, GN x`iters`)
at 512x384 through the C ABI, optionally one config-3-shaped accumulate+solve.  No timing, no L2
flush: under ncu every launch is serialised and cold anyway.  Usage:
    python tools/prof_step.py [--reps 2] [--iters 2] [--graph N,E] [--kind rays|calib|points]
"""
import argparse
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--iters", type=int, default=2)
    ap.add_argument("--graph", default="")
    ap.add_argument("--kind", default="rays")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    pkg, abi, synth, matching, mdist = bench.load_pkg()
    inp = bench.make_tracking_inputs(synth, matching, dev, 1001)
    N, n, E2 = 2, bench.N_PX, 2
    prm = abi.make_params(abi.GN_RAYS, 0.003, 10.0, 0.0, 1.5, a.iters, 0.0)
    ws, _ = abi.gn_workspace(N, E2, n, dev)
    for _ in range(a.reps):
        p, conv = abi.iter_proj(inp["rays"], inp["pts"], inp["p0"], 10, 1e-8, 1e-6)
        abi.refine_matches(inp["D11"], inp["D21"], p.long(), 3, 5)
        Twc = inp["Twc0"].clone()
        abi.gauss_newton(prm, Twc, inp["Xs"], inp["Cs"], None, inp["ii"], inp["jj"], inp["idx"], inp["valid"],
                         inp["Q"], workspace=ws)
    torch.cuda.synchronize()
    if a.graph:
        Ng, Eg = (int(x) for x in a.graph.split(","))
        calib = a.kind == "calib"
        g = synth.make_graph(Ng, Eg, bench.H, bench.W, seed=1003, device=dev, calib=calib)
        if calib:
            prm = abi.make_params(abi.GN_CALIB, 1.0, 10.0, 0.0, 1.5, a.iters, 0.0, bench.H, bench.W, -10, 1e-6)
        elif a.kind == "points":
            prm = abi.make_params(abi.GN_POINTS, 0.05, 1.0, 0.0, 1.5, a.iters, 0.0)
        ws, _ = abi.gn_workspace(Ng, 2 * Eg, n, dev)
        Twc = g["Twc"].clone()
        for _ in range(a.reps):
            abi.gauss_newton(prm, Twc, g["Xs"], g["Cs"], g["K"] if calib else None, g["ii"], g["jj"], g["idx"],
                             g["valid"], g["Q"], workspace=ws)
        torch.cuda.synchronize()
    print("prof_step ok")


if __name__ == "__main__":
    main()
