"""A/B timing of refine_matches at 512x384 on one box (run twice with different env): median of 50 cold-L2 launches."""
import importlib, os, statistics, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
abi = importlib.import_module("monst3r-slam_b200.abi")
synth = importlib.import_module("monst3r-slam_b200.synth")
matching = importlib.import_module("monst3r-slam_b200.matching")
d = synth.make_matching_pair(384, 512, seed=1001, device="cuda")
rays, pts, p0 = matching.prep_for_iter_proj(d["X11"], d["X21"], None)
p, _ = abi.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
p1 = p.long()
D11, D21 = d["D11"].half().contiguous(), d["D21"].view(1, 384 * 512, -1).half().contiguous()
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
ts = []
for r in range(60):
    flush.fill_(1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); abi.refine_matches(D11, D21, p1, 3, 5); b.record(); b.synchronize()
    if r >= 10: ts.append(a.elapsed_time(b))
print(os.environ.get("M3R_REFINE_REUSE", "1"), "median %.4f min %.4f ms" % (statistics.median(ts), min(ts)))
# ideal-coherence probe: every query centred on its own pixel and D21 == D11 (no centre ever moves, lanes stay 48 B apart)
if os.environ.get("M3R_AB_IDEAL"):
    v, u = torch.meshgrid(torch.arange(384, device="cuda"), torch.arange(512, device="cuda"), indexing="ij")
    pg = torch.stack((u, v), -1).view(1, -1, 2).long().contiguous()
    Dq = D11.view(1, 384 * 512, -1).contiguous()
    ts = []
    for r in range(60):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); abi.refine_matches(D11, Dq, pg, 3, 5); b.record(); b.synchronize()
        if r >= 10: ts.append(a.elapsed_time(b))
    print("ideal coherent: median %.4f min %.4f ms" % (statistics.median(ts), min(ts)))
