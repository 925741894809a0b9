"""Generates tests/golden/factor_graph.npz by running the REFERENCE's own python:
`FactorGraph.add_factors / get_unique_kf_idx / prep_two_way_edges` (mast3r_slam/global_opt.py) and
`constrain_points_to_ray` (mast3r_slam/geometry.py), imported from /root/reference on the CPU.

Only the modules that cannot exist here are stubbed before the import: `lietorch` (type annotation only on
this path), `mast3r_slam_backends` (not called), `mast3r_slam.frame` (SharedKeyframes is an annotation) and
`mast3r_slam.mast3r_utils`, whose `mast3r_match_symmetric` (the ViT) is replaced by a function that returns
the seeded synthetic match tensors saved in the fixture.  Everything after the matcher is the reference's
unmodified code.  Runs in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_golden_factor_graph.py
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/MASt3R-SLAM"


def import_reference():
    sys.path.insert(0, REF)
    lt = types.ModuleType("lietorch")
    lt.Sim3 = object
    sys.modules["lietorch"] = lt
    sys.modules["mast3r_slam_backends"] = types.ModuleType("mast3r_slam_backends")
    fr = types.ModuleType("mast3r_slam.frame")
    fr.SharedKeyframes = object
    sys.modules["mast3r_slam.frame"] = fr
    mu = types.ModuleType("mast3r_slam.mast3r_utils")
    mu.mast3r_match_symmetric = None
    sys.modules["mast3r_slam.mast3r_utils"] = mu
    import mast3r_slam.config as cfg
    import mast3r_slam.geometry as geo
    import mast3r_slam.global_opt as go
    return cfg, geo, go


def synth_matches(g, batch, n, valid_prob, q_scale):
    """shapes of mast3r_match_symmetric's outputs: idx [b,n] i64, valid [b,n,1] bool, Q* [b,n,1] f32"""
    idx = [torch.randint(0, n, (batch, n), generator=g, dtype=torch.int64) for _ in range(2)]
    valid = [torch.rand((batch, n, 1), generator=g) < valid_prob.view(batch, 1, 1) for _ in range(2)]
    Q = [torch.exp(torch.randn((batch, n, 1), generator=g) * 0.5 + q_scale.view(batch, 1, 1)).float() for _ in range(4)]
    return idx[0], idx[1], valid[0], valid[1], Q[0], Q[1], Q[2], Q[3]


class _KF:  # what FactorGraph.add_factors reads from a keyframe before the matcher
    feat = torch.zeros(1, 1)
    pos = torch.zeros(1, 1)
    img_true_shape = torch.zeros(1, 2)


def main():
    cfg, geo, go = import_reference()
    cfg.config["local_opt"] = {"window_size": 1e6, "Q_conf": 1.5, "C_conf": 0.0, "pin": 1}
    n = 24 * 32
    g = torch.Generator().manual_seed(2201)
    frames = [_KF() for _ in range(16)]
    fg = go.FactorGraph(model=None, frames=frames, K=None, device="cpu")
    # batches: (ii, jj, valid_prob per pair, log-Q offset per pair, min_match_frac, is_reloc)
    plan = [
        ([0], [1], [0.9], [1.0], 0.1, False),                                 # first edge
        ([1, 0, 2], [2, 2, 3], [0.9, 0.02, 0.03], [1.0, 1.0, 1.0], 0.1, False),  # (0,2) rejected, (2,3) consecutive kept
        ([3, 1], [4, 4], [0.8, 0.9], [0.0, -2.0], 0.1, False),               # (1,4) fails through Q only -> rejected
        ([0, 2], [5, 5], [0.9, 0.01], [1.0, 1.0], 0.1, True),                # reloc with a bad pair: nothing added
        ([4, 0], [5, 5], [0.9, 0.7], [1.0, 0.8], 0.1, True),                 # reloc, all good
        ([7], [6], [0.5], [0.6], 0.3, False),                                # ii > jj, not consecutive by the reference's test
    ]
    out = {"n": n, "num_batches": len(plan)}
    for t, (ii, jj, vp, qs, mmf, reloc) in enumerate(plan):
        m = synth_matches(g, len(ii), n, torch.tensor(vp), torch.tensor(qs))
        go.mast3r_match_symmetric = lambda *a, _m=m: _m
        ret = fg.add_factors(ii, jj, mmf, reloc)
        out[f"b{t}_ii"] = np.asarray(ii, dtype=np.int64)
        out[f"b{t}_jj"] = np.asarray(jj, dtype=np.int64)
        out[f"b{t}_min_match_frac"] = mmf
        out[f"b{t}_is_reloc"] = reloc
        out[f"b{t}_ret"] = bool(ret)
        for k, name in enumerate(("idx_i2j", "idx_j2i", "valid_j", "valid_i", "Qii", "Qjj", "Qji", "Qij")):
            out[f"b{t}_{name}"] = m[k].numpy()
        out[f"b{t}_num_edges_after"] = int(fg.ii.numel())
    for name in ("ii", "jj", "idx_ii2jj", "idx_jj2ii", "valid_match_j", "valid_match_i", "Q_ii2jj", "Q_jj2ii"):
        out[f"state_{name}"] = getattr(fg, name).numpy()
    out["unique_kf_idx"] = fg.get_unique_kf_idx().numpy()
    for name, t in zip(("ii", "jj", "idx", "valid", "Q"), fg.prep_two_way_edges()):
        out[f"twoway_{name}"] = t.numpy()
    # constrain_points_to_ray
    h, w = 24, 32
    K = torch.tensor([[27.2, 0.0, 15.5], [0.0, 26.9, 11.7], [0.0, 0.0, 1.0]])
    Xs = (torch.randn((3, h * w, 3), generator=g) * 0.7 + torch.tensor([0.0, 0.0, 2.5])).float()
    out["ray_h"], out["ray_w"], out["ray_K"], out["ray_Xs"] = h, w, K.numpy(), Xs.numpy()
    out["ray_out"] = geo.constrain_points_to_ray((h, w), Xs, K).numpy()
    path = os.path.join(HERE, "factor_graph.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; edges", int(fg.ii.numel()), "rets",
          [out[f"b{t}_ret"] for t in range(len(plan))])


if __name__ == "__main__":
    main()
