"""Per-CTA timeline of refine_staged_kernel (needs a -DM3R_REFINE_PROF build: M3R_LIB_DIR=lib_prof)."""
import ctypes as C, importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
abi = importlib.import_module("monst3r-slam_b200.abi")
synth = importlib.import_module("monst3r-slam_b200.synth")
matching = importlib.import_module("monst3r-slam_b200.matching")
d = synth.make_matching_pair(384, 512, seed=1001, device="cuda")
rays, pts, p0 = matching.prep_for_iter_proj(d["X11"], d["X21"], None)
p, _ = abi.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
p1 = p.long()
D11, D21 = d["D11"].half().contiguous(), d["D21"].view(1, 384 * 512, -1).half().contiguous()
for _ in range(3):
    abi.refine_matches(D11, D21, p1, 3, 5)
torch.cuda.synchronize()
abi.lib.m3r_debug_refine_prof(None, 1)
abi.refine_matches(D11, D21, p1, 3, 5)
torch.cuda.synchronize()
buf = np.zeros(4096 * 40, np.int64)
abi.lib.m3r_debug_refine_prof(buf.ctypes.data_as(C.c_void_p), 0)
P = buf.reshape(4096, 40)
P = P[P[:, 0] > 0]
t0 = P[:, 0].min()
print("ctas", len(P), "span us", (P[:, 1].max() - t0) / 1e3, "mean cta us", (P[:, 1] - P[:, 0]).mean() / 1e3,
      "min/max cta us", (P[:, 1] - P[:, 0]).min() / 1e3, (P[:, 1] - P[:, 0]).max() / 1e3, "passes mean", P[:, 3].mean())
L = P[:, 4:24].reshape(-1, 5, 4)
for l in range(5):
    print(f"level d={5 - l}: wait {L[:, l, 0].mean():.0f} score {L[:, l, 1].mean():.0f} gather {L[:, l, 2].mean():.0f} cycles; "
          f"outliers/cta {L[:, l, 3].mean():.2f} max {L[:, l, 3].max()}")
tot = L[:, :, :3].sum(axis=(0, 1))
print("sum over levels (mean cycles per cta): wait %.0f score %.0f gather %.0f" % tuple(tot / len(P)))
# per SM: busy time = union of CTA intervals
sm = P[:, 2]
ends = []
for s_ in np.unique(sm):
    q = P[sm == s_]
    ends.append((q[:, 1].max() - t0) / 1e3)
    if s_ < 3:
        print("sm", s_, "ctas", len(q), "start/end us", np.round((q[:, 0] - t0) / 1e3, 1), np.round((q[:, 1] - t0) / 1e3, 1))
ends = np.array(ends)
print("per-SM last end us: min %.1f mean %.1f max %.1f; ctas per SM min %d max %d" % (ends.min(), ends.mean(), ends.max(),
      np.bincount(sm).min(), np.bincount(sm).max()))
starts = np.sort(P[:, 0] - t0) / 1e3
print("cta start times us (every 74th):", np.round(starts[::74], 1))
dur = (P[:, 1] - P[:, 0]) / 1e3
order = np.argsort(-dur)
ids = np.nonzero(buf.reshape(4096, 40)[:, 0] > 0)[0]
print("slowest CTAs: id (tx,ty) dur_us start_us passes | per-level wait/score/gather kcycles, outliers")
for k in list(order[:8]) + list(order[-3:]):
    cid = ids[k]
    print(cid, (cid % 32, cid // 32), round(dur[k], 1), round((P[k, 0] - t0) / 1e3, 1), P[k, 3], "|",
          " ".join(f"{L[k, l, 0] / 1e3:.1f}/{L[k, l, 1] / 1e3:.1f}/{L[k, l, 2] / 1e3:.1f}/{L[k, l, 3]}" for l in range(5)))
print("duration percentiles us:", np.round(np.percentile(dur, [5, 25, 50, 75, 95, 99]), 1))
print("sum of score cycles per level-cta percentiles:", np.round(np.percentile(L[:, :, 1].sum(1), [5, 50, 95, 99]) / 1e3, 1))
