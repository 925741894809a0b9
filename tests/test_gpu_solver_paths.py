"""Every solve kernel that `launch_gn_solve` can dispatch to gives the same poses.

The dispatch is by system size (resident kernel up to 7 tiles, grid kernel with the single-barrier v4 tile step from 8
tiles), with environment knobs that force the other kernels for A/B timing: M3R_SOLVE=v1 (grid kernel for small
systems), M3R_SOLVE_NOV4=1 (the two-barrier v2 step), M3R_SOLVE_NOPIPE=1 (round-1 tile step: what systems too large
for the cp.async ring fall back to), M3R_SOLVE=df (the experimental dataflow kernel).  The knobs are read once per process, so each variant runs in its own interpreter and
leaves its poses in a file; all are compared with the default path's (different summation orders: 1e-5 / 1e-4 rad,
the bar of BASELINE.json) and the default with the CPU oracle.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, need_gpu, pose_diff

pytestmark = pytest.mark.gpu

SCRIPT = r"""
import importlib, sys
import numpy as np, torch
sys.path.insert(0, {root!r})
pkg = importlib.import_module("monst3r-slam_b200")
synth = importlib.import_module("monst3r-slam_b200.synth")
mod = pkg.load_drop_in()
N, E = {N}, {E}
g = synth.make_graph_exact(N, E, 1200, seed=300 + N, noise=1e-3, device="cuda")
Twc = g["Twc"].clone()
mod.gauss_newton_rays(Twc, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"], 0.003, 10.0, 0.0, 1.5, 5, 0.0)
torch.cuda.synchronize()
np.save({out!r}, np.concatenate([Twc.cpu().numpy().ravel(), np.asarray(list(mod.last_info()), np.float32)]))
"""


def _run(tmp_path, tag, env, N, E):
    out = str(tmp_path / f"{tag}_{N}.npy")
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, "-c", SCRIPT.format(root=ROOT, N=N, E=E, out=out)], env=e, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    v = np.load(out)
    return v[:-4].reshape(N, 8), v[-4:]


@pytest.mark.parametrize("N,E,variants", [
    (24, 80, {"grid_v1": {"M3R_SOLVE": "v1"}}),                                      # 6 tiles: resident by default
    (35, 120, {"two_barrier_v2": {"M3R_SOLVE_NOV4": "1"}}),                          # 8 tiles: the smallest v4 system
    (72, 300, {"two_barrier_v2": {"M3R_SOLVE_NOV4": "1"}, "round1_step": {"M3R_SOLVE_NOPIPE": "1"},
               "dataflow": {"M3R_SOLVE": "df"}}),                                     # 16 tiles: v4 by default
])
def test_solver_variants_agree(tmp_path, oracle, synth, N, E, variants):
    need_gpu()
    base, info = _run(tmp_path, "default", {}, N, E)
    assert list(info[:2]) == [5, 0], info
    g = synth.make_graph_exact(N, E, 1200, seed=300 + N, noise=1e-3)
    c = {k: (v.numpy() if hasattr(v, "numpy") else v) for k, v in g.items()}
    To, _, _, nfail = oracle.gauss_newton_rays(c["Twc"], c["Xs"], c["Cs"], c["ii"], c["jj"], c["idx"], c["valid"], c["Q"],
                                               0.003, 10.0, 0.0, 1.5, 5, 0.0)
    rel, ang = pose_diff(base, To)
    print(f"N={N}: default path vs oracle rel {rel:.2e} ang {ang:.2e}")
    assert nfail == 0 and rel < 1e-5 and ang < 1e-4
    for tag, env in variants.items():
        T, inf = _run(tmp_path, tag, env, N, E)
        rel, ang = pose_diff(T, base)
        print(f"N={N}: {tag} vs default rel {rel:.2e} ang {ang:.2e} info {inf}")
        assert list(inf[:2]) == [5, 0], (tag, inf)
        assert rel < 1e-5 and ang < 1e-4, (tag, rel, ang)
