"""One invocation of the section-8(f) kernels for ncu: edge_admit (120 pairs x 196 608 points, one launch),
ray_constrain (32 maps) and the frame tracker (512x384).  `ncu --set full -k regex:"edge_admit|ray_constrain|gn_track"`."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
abi = importlib.import_module("monst3r-slam_b200.abi")
synth = importlib.import_module("monst3r-slam_b200.synth")
trk = importlib.import_module("monst3r-slam_b200.tracker")

dev = "cuda"
H, W = 384, 512
n, E = H * W, 120
g = torch.Generator(device=dev).manual_seed(5)
base = torch.arange(n, device=dev).view(1, n)
idx = [(base + torch.randint(-3, 4, (E, n), device=dev, generator=g)).clamp_(0, n - 1) for _ in range(2)]
valid = [torch.rand((E, n, 1), device=dev, generator=g) < 0.9 for _ in range(2)]
Qs = [torch.exp(torch.randn((E, n, 1), device=dev, generator=g) * 0.5 + 1.0) for _ in range(4)]
s_idx = torch.empty((2 * E, n), dtype=torch.int64, device=dev)
s_valid = torch.empty((2 * E, n), dtype=torch.bool, device=dev)
s_Q = torch.empty((2 * E, n), dtype=torch.float32, device=dev)
for _ in range(2):
    abi.edge_admit(idx[0], idx[1], valid[0], valid[1], Qs[0], Qs[1], Qs[2], Qs[3], 1.5, s_idx, s_valid, s_Q, 0)
Xs = torch.randn((32, n, 3), device=dev, generator=g) + torch.tensor([0.0, 0.0, 3.0], device=dev)
K = torch.tensor([[435.2, 0.0, 255.5], [0.0, 433.1, 191.5], [0.0, 0.0, 1.0]], device=dev)
for _ in range(2):
    abi.constrain_points_to_ray((H, W), Xs, K)
gr = synth.make_graph(2, 1, H, W, seed=1002, device=dev, calib=False)
Xf = gr["Xs"][1][gr["idx"][1]].contiguous()
Q = gr["Q"][1].reshape(-1, 1).contiguous()
val = (gr["valid"][1].reshape(-1, 1) & (Q > 1.5)).contiguous()
opt = trk.PoseOptimizer(cfg={"rel_error": 0.0, "delta_norm": 0.0, "max_iters": 10})
for _ in range(2):
    opt.opt_pose_ray_dist_sim3(Xf, gr["Xs"][0].contiguous(), gr["Twc"][1].view(1, 8), gr["Twc"][0].view(1, 8), Q, val)
torch.cuda.synchronize()
print("done", opt.last_info.tolist())
