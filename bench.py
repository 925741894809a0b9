#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 geometric backend (contract: see DESIGN.md §Measurement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--global-opt c3|c4|none]

One "step" = one frame-to-keyframe tracking pass on synthetic 512x384 data (BASELINE.json configs[1]):
iter_proj + refine_matches on one pointmap pair (24-d descriptors) and gauss_newton_rays on the
2-keyframe / 2-directed-edge graph, 10 iterations.  metric = matched px/s = h*w / step time.
  value        inputs resident in HBM, C ABI, CUDA events, L2 flushed between steps
  e2e          same step through the drop-in python module `mast3r_slam_backends` with pinned HOST
               inputs: H2D of every input + D2H of matches / mask / poses inside the timed region
  roofline     dominant kernel of the step vs the measured HBM peak (MEASURED_PEAKS.json)
  cpu_baseline the reference's CPU path (oracle/_ref host build of its matching kernels + the GN
               port) on this box's cores, bounded sample; `--impl reference` times only that
  global_opt   multi-keyframe GN (config 3: 32 kf / 240 directed edges; c4: 256 kf / 2000 edges),
               edge-sharded over the N ranks with one NCCL all-reduce of the edge blocks per iteration
At N>1 (torchrun) every rank tracks its own frame (replicas, weak scaling; the path has no exchange
step for single-frame tracking) and the ranks jointly solve the global-opt graph (strong scaling).
"""
import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H, W, FDIM = 384, 512, 24
N_PX = H * W
MATCH_CFG = dict(max_iter=10, lambda_init=1e-8, convergence_thresh=1e-6, radius=3, dilation_max=5)
GN_RAYS = dict(sigma_ray=0.003, sigma_dist=10.0, C_thresh=0.0, Q_thresh=1.5, max_iter=10, delta_thresh=0.0)
BYTES_ITER_PROJ = 65      # per px   
BYTES_REFINE = 128        # per px
METRIC = "matched px/s (tracking step: iter_proj + refine_matches + gauss_newton_rays x10, 512x384)"


def gn_bytes(E_local, N_touched, n=N_PX):
    return 13 * E_local * n + 16 * N_touched * n


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def load_pkg():
    pkg = importlib.import_module("monst3r-slam_b200")
    return (pkg, importlib.import_module("monst3r-slam_b200.abi"), importlib.import_module("monst3r-slam_b200.synth"),
            importlib.import_module("monst3r-slam_b200.matching"), importlib.import_module("monst3r-slam_b200.dist"))


def make_tracking_inputs(synth, matching, device, seed):
    """config 2: one pair for matching + the N=2 / E2=2 rays graph, all on `device`."""
    import torch
    d = synth.make_matching_pair(H, W, b=1, fdim=FDIM, seed=seed, device=device)
    rays, pts, p0 = matching.prep_for_iter_proj(d["X11"], d["X21"], None)
    g = synth.make_graph(2, 1, H, W, seed=seed + 1, device=device)
    inp = dict(rays=rays.contiguous(), pts=pts.contiguous(), p0=p0.contiguous(), D11=d["D11"].half().contiguous(),
               D21=d["D21"].view(1, N_PX, FDIM).half().contiguous(), Twc0=g["Twc"].contiguous(),
               Xs=g["Xs"], Cs=g["Cs"], ii=g["ii"], jj=g["jj"], idx=g["idx"].contiguous(),
               valid=g["valid"].contiguous(), Q=g["Q"].contiguous())
    return inp


# ------------------------------------------------------------------------------------------ GPU arm
def run_gpu(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this backend has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pkg, abi, synth, matching, mdist = load_pkg()
    ops = pkg.load_drop_in()
    K, Wm = args.steps, max(args.warmup, 3)

    inp = make_tracking_inputs(synth, matching, dev, 1001 + 10 * rank)
    N, n, E2 = 2, N_PX, 2
    prm = abi.make_params(abi.GN_RAYS, GN_RAYS["sigma_ray"], GN_RAYS["sigma_dist"], GN_RAYS["C_thresh"],
                          GN_RAYS["Q_thresh"], GN_RAYS["max_iter"], GN_RAYS["delta_thresh"])
    ws, _ = abi.gn_workspace(N, E2, n, dev)
    Twc = inp["Twc0"].clone()
    dx = torch.empty((N - 1, 7), dtype=torch.float32, device=dev)
    info = torch.empty(4, dtype=torch.int32, device=dev)
    flush = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def flush_l2():
        flush.fill_(1)

    def step_resident():
        p, conv = abi.iter_proj(inp["rays"], inp["pts"], inp["p0"], MATCH_CFG["max_iter"], MATCH_CFG["lambda_init"],
                                MATCH_CFG["convergence_thresh"])
        p1 = p.long()
        (r,) = abi.refine_matches(inp["D11"], inp["D21"], p1, MATCH_CFG["radius"], MATCH_CFG["dilation_max"])
        Twc.copy_(inp["Twc0"])
        abi.gauss_newton(prm, Twc, inp["Xs"], inp["Cs"], None, inp["ii"], inp["jj"], inp["idx"], inp["valid"],
                         inp["Q"], workspace=ws, info=info, dx=dx)
        return r, conv

    def timed(fn, reps, warm):
        """per-rep CUDA events on the current stream, L2 flushed (outside the events) before each rep"""
        for _ in range(warm):
            flush_l2()
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            flush_l2()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            ts.append(a.elapsed_time(b))
        return ts

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- headline: K resident steps -------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    ts = timed(step_resident, K, Wm)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    total_ms = max_over_ranks(sum(ts))
    ms_per_step = total_ms / K
    value = world * N_PX / (ms_per_step * 1e-3)
    launches_per_step = 1 + 1 + 1 + 2 * GN_RAYS["max_iter"]  # iter_proj, refine, gn_prepare, 10x(accum+solve)

    # ---- per-kernel timings for the roofline (rank 0 numbers; kernels timed alone, cold L2) -------
    peak, peak_kind = peaks()
    kern = {}
    p_fixed, _ = abi.iter_proj(inp["rays"], inp["pts"], inp["p0"], 10, 1e-8, 1e-6)
    p1_fixed = p_fixed.long()
    abi.gn_prepare(inp["ii"], inp["jj"], N, n, ws)

    def k_iter():
        abi.iter_proj(inp["rays"], inp["pts"], inp["p0"], 10, 1e-8, 1e-6)

    def k_refine():
        abi.refine_matches(inp["D11"], inp["D21"], p1_fixed, 3, 5)

    def k_accum():
        abi.gn_accumulate(prm, inp["Twc0"], inp["Xs"], inp["Cs"], None, inp["idx"], inp["valid"], inp["Q"], E2, ws)

    Ttmp = inp["Twc0"].clone()

    def k_solve():
        abi.lib.m3r_gn_solve_retract(abi._ptr(Ttmp), N, n, E2, abi.C.c_float(0.0), abi._ptr(dx), abi._ptr(info),
                                     abi._ptr(ws), abi.C.c_size_t(ws.numel()), abi._stream())

    reps = max(K, 10)
    for name, fn, nbytes, per_step in (("iter_proj_kernel", k_iter, BYTES_ITER_PROJ * N_PX, 1),
                                       ("refine_f16_fd24_kernel", k_refine, BYTES_REFINE * N_PX, 1),
                                       ("gn_accum_kernel<rays>", k_accum, gn_bytes(E2, N), GN_RAYS["max_iter"]),
                                       ("gn_solve_kernel", k_solve, 0, GN_RAYS["max_iter"])):
        t = timed(fn, reps, 3)
        ms = statistics.median(t)
        kern[name] = {"ms": ms, "launches_per_step": per_step, "bytes": nbytes,
                      "GBps": (nbytes / (ms * 1e-3) / 1e9) if nbytes else None,
                      "frac_of_hbm_peak": (nbytes / (ms * 1e-3) / 1e9 / peak) if nbytes else None,
                      "share_of_step": ms * per_step / ms_per_step}
    abi.gn_prepare(inp["ii"], inp["jj"], N, n, ws)
    dom = max(kern, key=lambda k_: kern[k_]["share_of_step"] if kern[k_]["bytes"] else -1)
    roofline = {"kernel": dom, "bound": "hbm", "achieved": kern[dom]["GBps"], "peak": peak, "unit": "GB/s",
                "frac": kern[dom]["frac_of_hbm_peak"], "traffic": None, "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs)",
                "note": "algorithmic bytes / CUDA-event time of the kernel launched alone with a cold L2; "
                        "time-bound by instruction issue / L1 gathers, not HBM (see DESIGN.md, profiles/)"}

    # ---- e2e: python drop-in module, pinned host inputs, H2D + D2H inside the timed region --------
    host = {k: v.cpu().pin_memory() for k, v in inp.items() if k not in ("Twc0",)}
    host["Twc0"] = inp["Twc0"].cpu().pin_memory()
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    out_host = {"idx": torch.empty((1, n, 2), dtype=torch.int64).pin_memory(),
                "conv": torch.empty((1, n), dtype=torch.bool).pin_memory(),
                "Twc": torch.empty((N, 8), dtype=torch.float32).pin_memory()}
    d2h = sum(v.numel() * v.element_size() for v in out_host.values())

    def step_e2e():
        dv = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
        p, conv = ops.iter_proj(dv["rays"], dv["pts"], dv["p0"], MATCH_CFG["max_iter"], MATCH_CFG["lambda_init"],
                                MATCH_CFG["convergence_thresh"])
        (r,) = ops.refine_matches(dv["D11"], dv["D21"], p.long(), MATCH_CFG["radius"], MATCH_CFG["dilation_max"])
        T = dv["Twc0"]
        ops.gauss_newton_rays(T, dv["Xs"], dv["Cs"], dv["ii"], dv["jj"], dv["idx"], dv["valid"], dv["Q"],
                              GN_RAYS["sigma_ray"], GN_RAYS["sigma_dist"], GN_RAYS["C_thresh"], GN_RAYS["Q_thresh"],
                              GN_RAYS["max_iter"], GN_RAYS["delta_thresh"])
        out_host["idx"].copy_(r, non_blocking=True)
        out_host["conv"].copy_(conv, non_blocking=True)
        out_host["Twc"].copy_(T, non_blocking=True)

    barrier()
    te = timed(step_e2e, K, 3)
    barrier()
    e2e_ms = max_over_ranks(sum(te)) / K
    e2e = {"value": world * N_PX / (e2e_ms * 1e-3), "unit": "px/s", "ms_per_step": e2e_ms,
           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "api": "mast3r_slam_backends.iter_proj / refine_matches / gauss_newton_rays (pybind drop-in), pinned host buffers"}

    # ---- global optimisation, edge-sharded -------------------------------------------------------
    gopt = None
    if args.global_opt != "none":
        gopt = run_global_opt(args, abi, synth, mdist, dev, rank, world, barrier, max_over_ranks, peak)

    # ---- CPU baseline on rank 0 at N=1 --------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference_step(inp, steps=1, warmup=0, budget_s=25.0)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "px/s", "n_gpus": world, "steps": K, "warmup": Wm,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32 (iter_proj, GN accumulate) / f16 (refine_matches) / f64 (assemble + Cholesky)",
                "data": "synthetic",
                "config": {"workload": "configs[1]: frame-to-keyframe tracking, 512x384, 24-d desc, matching + "
                                       "gauss_newton_rays N=2/E2=2 x10 iters; one replica per GPU",
                           "l2": "L2 flushed (512 MB fill) before every timed step and kernel rep",
                           "parallelism": f"replicas x{world}" if world > 1 else "1 GPU"},
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches_per_step * K,
                "roofline": roofline, "kernels": kern, "cpu_baseline": cpu, "global_opt": gopt}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
