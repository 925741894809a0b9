#!/usr/bin/env python
"""Phase breakdown of gn_track_kernel (needs a build with -DM3R_TRACK_PROF): cycles of CTA 0 per phase."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
dev = torch.device("cuda", 0)
pkg, abi, synth, matching, mdist = bench.load_pkg()
inp = bench.make_tracking_inputs(synth, matching, dev, 1001)
N, n, E2 = 2, bench.N_PX, 2
prm = abi.make_params(abi.GN_RAYS, 0.003, 10.0, 0.0, 1.5, 10, 0.0)
ws, _ = abi.gn_workspace(N, E2, n, dev)
for rep in range(1):
    Twc = inp["Twc0"].clone()
    ws[:64].zero_()
    abi.gauss_newton(prm, Twc, inp["Xs"], inp["Cs"], None, inp["ii"], inp["jj"], inp["idx"], inp["valid"], inp["Q"], workspace=ws)
    torch.cuda.synchronize()
    ctrl = ws[:64].view(torch.int32).cpu().tolist()
names = ["stage+gather", "edge_xform", "accumulate", "cta reduce+store", "grid.sync", "partial sums", "adjoints", "solve+retract"]
tot = sum(ctrl[6:14]) or 1
for k, nm in enumerate(names):
    print("%-18s %8d cycles  %6.2f us  %5.1f%%" % (nm, ctrl[6 + k], ctrl[6 + k] / 1.965e3, 100.0 * ctrl[6 + k] / tot))
print("total %.1f us over %d iterations" % (tot / 1.965e3, ctrl[1]))

import ctypes, numpy as np, statistics
buf = np.zeros(1024, dtype=np.uint64)
if hasattr(abi.lib, "m3r_debug_read"):
    abi.lib.m3r_debug_read(buf.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(buf.nbytes))
    st = [int(v) for v in buf[700:708]]
    print("iteration 5 stamps in CTA 0 (ns after the xform stamp):", [v - st[1] for v in st[1:]],
          "= xform, accumulate, reduce+store, gather, sums, adjoints, solve")
    k0 = int(buf[720])
    print("kernel start -> staged:", int(buf[721]) - k0, "ns; pose published at (ns after kernel start):",
          [int(v) - k0 for v in buf[730:740]])
    t0 = int(buf[601])
    pub = [int(v) - t0 for v in buf[:148]]
    got = [int(v) - t0 for v in buf[301:448]]
    print("pose published (it 4) at 0 ns; CTA0 gathered it-5 partials at", int(buf[600]) - t0, "ns")
    print("other CTAs got pose: min/med/max", min(got), statistics.median(got), max(got))
    print("partials (it 5) published: cta0", pub[0], "others min/med/max", min(pub[1:]), statistics.median(pub[1:]), max(pub[1:]))
    print("latest 8 CTAs:", sorted(range(148), key=lambda c: -pub[c])[:8])
