#!/usr/bin/env python
"""Timing probe (not a bench line): e2e step with inputs copied on the compute stream vs on a copy stream,
plus the H2D copy alone.  Usage: python tools/e2e_probe.py"""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
pkg, abi, synth, matching, mdist = bench.load_pkg()
ops = pkg.load_drop_in()
inp = bench.make_tracking_inputs(synth, matching, dev, 1001)
host = {k: v.cpu().pin_memory() for k, v in inp.items()}
M, G = bench.MATCH_CFG, bench.GN_RAYS
groups = (("rays", "pts", "p0"), ("D11", "D21"), ("Twc0", "Xs", "Cs", "ii", "jj", "idx", "valid", "Q"))
s_up = torch.cuda.Stream(dev)

def compute(dv, waits=None):
    cur = torch.cuda.current_stream(dev)
    if waits: cur.wait_event(waits[0])
    p, conv = ops.iter_proj(dv["rays"], dv["pts"], dv["p0"], M["max_iter"], M["lambda_init"], M["convergence_thresh"])
    if waits: cur.wait_event(waits[1])
    (r,) = ops.refine_matches(dv["D11"], dv["D21"], p.long(), M["radius"], M["dilation_max"])
    if waits: cur.wait_event(waits[2])
    ops.gauss_newton_rays(dv["Twc0"], dv["Xs"], dv["Cs"], dv["ii"], dv["jj"], dv["idx"], dv["valid"], dv["Q"],
                          G["sigma_ray"], G["sigma_dist"], G["C_thresh"], G["Q_thresh"], G["max_iter"], G["delta_thresh"])
    return r, conv

def copy_only():
    return {k: v.to(dev, non_blocking=True) for k, v in host.items()}

def serial():
    dv = copy_only(); compute(dv)

def piped():
    cur = torch.cuda.current_stream(dev)
    s_up.wait_stream(cur)
    dv, ready = {}, []
    with torch.cuda.stream(s_up):
        for g_ in groups:
            for k_ in g_: dv[k_] = host[k_].to(dev, non_blocking=True)
            ready.append(s_up.record_event())
    for t_ in dv.values(): t_.record_stream(cur)
    compute(dv, ready)

def compute_only():
    compute(inp_dev)

inp_dev = {k: v.clone() for k, v in inp.items()}
def timed(fn, reps=20, warm=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    return statistics.median(ts), min(ts)
for name, fn in (("copy_only", copy_only), ("compute_only(warm L2)", compute_only), ("serial", serial), ("piped", piped)):
    print("%-24s median %.3f ms  min %.3f ms" % ((name,) + timed(fn)))
import time
t0 = time.perf_counter()
for _ in range(20): compute_only()
t1 = time.perf_counter(); torch.cuda.synchronize()
print("host enqueue time per compute step: %.3f ms" % ((t1 - t0) / 20 * 1e3))
