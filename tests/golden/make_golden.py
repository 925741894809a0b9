"""Generates tests/golden/*.npz from the REFERENCE's own CUDA build (oracle/_ref/ref_backends.so).
Runs on a B200 box only (`gpurun -- python tests/golden/make_golden.py`); outputs land in
gpurun_out/golden/ and are copied into tests/golden/ by hand and committed.  Small sizes keep the
fixtures a few hundred kB.  The CPU oracle is then pinned against them in tests/test_oracle_pin.py.
"""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))


def main():
    out = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out, exist_ok=True)
    ref = importlib.import_module("ref_backends")
    synth = importlib.import_module("monst3r-slam_b200.synth")
    m = importlib.import_module("monst3r-slam_b200.matching")
    np_ = lambda t: t.detach().cpu().numpy()
    # matching: 48x64, identity and previous-frame init
    for tag, seed, use_init in (("identity", 31, False), ("previous", 32, True)):
        h, w = 48, 64
        d = synth.make_matching_pair(h, w, seed=seed, device="cuda")
        rays, pts, p0 = m.prep_for_iter_proj(d["X11"], d["X21"], d["idx_init"] if use_init else None)
        p_new, conv = ref.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
        p1 = p_new.long()
        D11 = d["D11"].half()
        D21 = d["D21"].view(1, h * w, -1).half()
        (p1_new,) = ref.refine_matches(D11, D21, p1, 3, 5)
        np.savez_compressed(os.path.join(out, f"matching_{tag}.npz"), rays=np_(rays), pts=np_(pts), p_init=np_(p0),
                            p_new=np_(p_new), converged=np_(conv), D11=np_(D11), D21=np_(D21), p1=np_(p1),
                            p1_new=np_(p1_new), max_iter=10, lambda_init=1e-8, cost_thresh=1e-6, radius=3,
                            dilation_max=5)
    # GN: three kinds on small graphs, 10 iterations
    for kind in ("rays", "points", "calib"):
        if kind == "calib":
            g = synth.make_graph(4, 5, 48, 64, seed=41, calib=True, device="cuda")
        else:
            g = synth.make_graph_exact(4, 5, 2000, seed=42, noise=1e-3, device="cuda")
        Twc = g["Twc"].clone()
        a = (Twc, g["Xs"], g["Cs"])
        e = (g["ii"], g["jj"], g["idx"], g["valid"], g["Q"])
        meta = dict(height=0, width=0, pixel_border=0, z_eps=0.0)
        if kind == "rays":
            ref.gauss_newton_rays(*a, *e, 0.003, 10.0, 0.0, 1.5, 10, 0.0)
            sa, sb = 0.003, 10.0
        elif kind == "points":
            ref.gauss_newton_points(*a, *e, 0.05, 0.0, 1.5, 10, 0.0)
            sa, sb = 0.05, 0.0
        else:
            ref.gauss_newton_calib(*a, g["K"], *e, 48, 64, -10, 1e-6, 1.0, 10.0, 0.0, 1.5, 10, 0.0)
            sa, sb = 1.0, 10.0
            meta = dict(height=48, width=64, pixel_border=-10, z_eps=1e-6)
        torch.cuda.synchronize()
        np.savez_compressed(os.path.join(out, f"gn_{kind}.npz"), kind=kind, Twc0=np_(g["Twc"]), Twc_ref=np_(Twc),
                            Xs=np_(g["Xs"]), Cs=np_(g["Cs"]), K=np_(g["K"]), ii=np_(g["ii"]), jj=np_(g["jj"]),
                            idx=np_(g["idx"]), valid=np_(g["valid"]), Q=np_(g["Q"]), sigma_a=sa, sigma_b=sb,
                            C_thresh=0.0, Q_thresh=1.5, max_iter=10, delta_thresh=0.0, **meta)
    print("golden fixtures written to", out, os.listdir(out))


if __name__ == "__main__":
    main()
