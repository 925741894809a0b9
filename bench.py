#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 geometric backend (contract: see DESIGN.md §Measurement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--global-opt both|c3|c4|none]

One "step" = one frame-to-keyframe tracking pass on synthetic 512x384 data (BASELINE.json configs[1]):
iter_proj + refine_matches on one pointmap pair (24-d descriptors) and gauss_newton_rays on the
2-keyframe / 2-directed-edge graph, 10 iterations.  metric = matched px/s = h*w / step time.
  value        inputs resident in HBM, C ABI, CUDA events, L2 flushed between steps
  e2e          same step through the drop-in python module `mast3r_slam_backends` with pinned HOST
               inputs: H2D of every input + D2H of matches / mask / poses inside the timed region
  roofline     dominant kernel of the step vs the measured HBM peak (MEASURED_PEAKS.json)
  cpu_baseline the reference's CPU path (oracle/_ref host build of its matching kernels + the GN
               port) on this box's cores, bounded sample; `--impl reference` times only that
  global_opt   multi-keyframe GN, BOTH graphs by default at every N: `global_opt` = configs[2] (32 kf / 240
               directed edges, rays), `global_opt_c4` = configs[3] (256 kf / 2000 edges, calib); edge-sharded over
               the N ranks with one NCCL exchange of the edge blocks per iteration; each leg reports
               `poses_bit_identical_to_unsharded` (rank 0 repeats the solve as a world-size-1 call)
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
WORKLOAD = ("configs[1]: frame-to-keyframe tracking, 512x384, 24-d desc, matching + gauss_newton_rays N=2/E2=2 x10 iters")


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


class GpuLocalHostMemory:
    """While active, the calling thread runs on the CPUs of the GPU's NUMA node, so pinned host buffers allocated
    inside are first-touched on the memory closest to that GPU's PCIe root (matters when 8 ranks upload at once).
    The previous affinity is restored on exit.  Silently a no-op where sysfs does not say (single node, no /sys)."""

    def __init__(self, device_index):
        self.node, self.cpus, self.prev = None, None, None
        try:
            import torch
            p = torch.cuda.get_device_properties(device_index)
            bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
            node = int(open("/sys/bus/pci/devices/%s/numa_node" % bdf).read())
            if node < 0:
                return
            cpus = set()
            for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
            cpus &= os.sched_getaffinity(0)
            if cpus:
                self.node, self.cpus = node, cpus
        except (OSError, ValueError, AttributeError, RuntimeError):
            pass

    def __enter__(self):
        if self.cpus:
            self.prev = os.sched_getaffinity(0)
            os.sched_setaffinity(0, self.cpus)
        return self

    def __exit__(self, *exc):
        if self.prev:
            os.sched_setaffinity(0, self.prev)
        return False


_REAL_STDOUT = None


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def ncu_traffic(kernel_name):
    """dram__bytes_read + dram__bytes_write per launch of `kernel_name`, from the committed `ncu --set full`
    summary of the same step (profiles/*_ncu_full_summary.csv, newest first); None if not captured."""
    import csv
    import glob
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_full_summary.csv")), reverse=True):
        try:
            rows = list(csv.reader(open(path)))
            hdr, units = rows[0], rows[1]
            ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            for r in rows[2:]:
                if kernel_name.split("(")[0].strip() in r[0]:
                    return float(r[ir]) * scale.get(units[ir], 1.0) + float(r[iw]) * scale.get(units[iw], 1.0)
        except (OSError, ValueError, IndexError):
            continue
    return None


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
    # iter_proj_repack, iter_proj_padded, refine_staged, gn_prepare, gn_pack, gn_track (one cooperative launch
    # runs all GN iterations)
    launches_per_step = 2 + 1 + 1 + 1 + 1

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

    abi.gn_pack(prm, inp["Cs"], inp["idx"], inp["valid"], inp["Q"], N, E2, ws)

    def k_pack():
        abi.gn_pack(prm, inp["Cs"], inp["idx"], inp["valid"], inp["Q"], N, E2, ws)

    def k_accum():
        abi.gn_accumulate_packed(prm, inp["Twc0"], inp["Xs"], None, E2, ws)

    def k_gn():
        Twc.copy_(inp["Twc0"])
        abi.gauss_newton(prm, Twc, inp["Xs"], inp["Cs"], None, inp["ii"], inp["jj"], inp["idx"], inp["valid"],
                         inp["Q"], workspace=ws, info=info, dx=dx)

    Ttmp = inp["Twc0"].clone()

    def k_solve():
        abi.lib.m3r_gn_solve_retract(abi._ptr(Ttmp), N, n, E2, abi.C.c_float(0.0), abi._ptr(dx), abi._ptr(info),
                                     abi._ptr(ws), abi.C.c_size_t(ws.numel()), abi._stream())

    reps = max(K, 10)
    for name, fn, nbytes, per_step in (("iter_proj (iter_proj_repack_kernel + iter_proj_padded_kernel)", k_iter, BYTES_ITER_PROJ * N_PX, 1),
                                       ("refine_staged_kernel", k_refine, BYTES_REFINE * N_PX, 1),
                                       ("gn_pack_kernel (inside gauss_newton)", k_pack, (13 + 8) * E2 * n + 8 * N * n, 0),
                                       ("gn_accum_kernel<rays> (accumulate only, cold L2)", k_accum, gn_bytes(E2, N), 0),
                                       ("gn_small_solve_kernel (stand-alone)", k_solve, 0, 0),
                                       ("gauss_newton_rays x10 (gn_prepare + gn_pack + persistent gn_track_kernel)",
                                        k_gn, gn_bytes(E2, N) * GN_RAYS["max_iter"], 1)):
        t = timed(fn, reps, 3)
        ms = statistics.median(t)
        kern[name] = {"ms": ms, "launches_per_step": per_step, "bytes": nbytes,
                      "GBps": (nbytes / (ms * 1e-3) / 1e9) if nbytes else None,
                      "frac_of_hbm_peak": (nbytes / (ms * 1e-3) / 1e9 / peak) if nbytes else None,
                      "frac_of_nominal_8TBps": (nbytes / (ms * 1e-3) / 1e9 / 8000.0) if nbytes else None,
                      "share_of_step": ms * per_step / ms_per_step}
    abi.gn_prepare(inp["ii"], inp["jj"], N, n, ws)
    abi.gn_pack(prm, inp["Cs"], inp["idx"], inp["valid"], inp["Q"], N, E2, ws)
    dom = max(kern, key=lambda k_: kern[k_]["share_of_step"] if kern[k_]["bytes"] else -1)
    # "bound" names the roofline the fraction is taken against (the contract's hbm | tensor); "limiter" is what ncu says
    # actually bounds the kernel
    roofline = {"kernel": dom, "bound": "hbm", "limiter": "fp16x2 + shared-memory pipes at 16 warps/SM (ncu: profiles/r02j_refine_experiments.md), not HBM"
                if dom.startswith("refine") else "instruction issue / latency, not HBM", "achieved": kern[dom]["GBps"], "peak": peak, "unit": "GB/s",
                "frac": kern[dom]["frac_of_hbm_peak"], "frac_of_nominal_8TBps": kern[dom]["frac_of_nominal_8TBps"],
                "traffic": ncu_traffic(dom), "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs)",
                "note": "algorithmic bytes / CUDA-event time of the kernel launched alone with a cold L2; "
                        "time-bound by instruction issue / L1 gathers, not HBM (see DESIGN.md, profiles/)"}
    if dom.startswith("refine"):
        # the limits that do bind this kernel: bit-exact c10::Half arithmetic needs one HMUL2 and one
        # HADD2 per two rounded MACs (0.5 warp-inst/clk/SMSP, profiles/r01b_ubench_pipes.log), and every
        # (query, candidate) pair reads 48 B of descriptors from shared memory (128 B/clk/SM)
        radius, dil_max, feat = 3, 5, inp["D11"].shape[-1]
        cand = dil_max * (2 * radius + 1) ** 2
        sm_hz = 148 * 1965.0e6  # B200 boost clock (clocks.max.sm)
        fp16_us = N_PX * cand * feat / 32 / (4 * 0.5 * sm_hz) * 1e6
        lds_us = N_PX * cand * feat * 2 / (128 * sm_hz) * 1e6
        roofline["pipe_bounds"] = {"fp16x2_issue_floor_us": fp16_us, "shared_memory_floor_us": lds_us,
                                   "frac_of_fp16x2_floor": fp16_us / (kern[dom]["ms"] * 1e3),
                                   "note": "the two co-equal binding limits of the bit-exact half-precision scoring"}

    # ---- e2e: python drop-in module, pinned host inputs, H2D + D2H inside the timed region --------
    # The caller keeps a step's inputs in ONE pinned arena (tensors are 256-byte aligned views, laid out in the order
    # the ops consume them) and mirrors it on the device, so a group of inputs goes up as one contiguous copy.
    groups = (("rays", "pts", "p0"), ("D11", "D21"), ("Twc0", "Xs", "Cs", "ii", "jj", "idx", "valid", "Q"))
    assert sorted(k for g_ in groups for k in g_) == sorted(inp.keys())
    offs, pos, spans = {}, 0, []
    for g_ in groups:
        lo = pos
        for k_ in g_:
            offs[k_] = pos
            pos = (pos + inp[k_].numel() * inp[k_].element_size() + 255) // 256 * 256
        spans.append((lo, pos))
    arena_bytes = pos

    def arena_views(arena):
        return {k_: arena[offs[k_]:offs[k_] + inp[k_].numel() * inp[k_].element_size()].view(inp[k_].dtype).view(inp[k_].shape)
                for k_ in inp}

    numa = GpuLocalHostMemory(local_rank)
    with numa:
        host_arena = torch.empty(arena_bytes, dtype=torch.uint8, pin_memory=True)
        out_host = {"idx": torch.empty((1, n, 2), dtype=torch.int64, pin_memory=True),
                    "conv": torch.empty((1, n), dtype=torch.bool, pin_memory=True),
                    "Twc": torch.empty((N, 8), dtype=torch.float32, pin_memory=True)}
        out_host_b = {k: torch.empty_like(v, pin_memory=True) for k, v in out_host.items()}
    host = arena_views(host_arena)
    for k_, v in inp.items():
        host[k_].copy_(v)
    h2d = arena_bytes  # what is copied per step (tensor bytes + < 3 KB of alignment padding)
    d2h = sum(v.numel() * v.element_size() for v in out_host.values())

    # The caller's streams: inputs go up on a copy stream in the order the ops need them (matching inputs,
    # descriptors, graph slabs) while the compute stream runs each op as soon as its inputs have landed;
    # results come back on a third stream.  Everything is fenced into the timed events of the compute stream.
    s_up, s_down = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    dev_arena = torch.empty(arena_bytes, dtype=torch.uint8, device=dev)  # device staging buffers, allocated once
    dv = arena_views(dev_arena)

    def step_e2e():
        cur = torch.cuda.current_stream(dev)
        s_up.wait_stream(cur)  # nothing starts before the timed region opens
        ready = []
        with torch.cuda.stream(s_up):
            for lo_, hi_ in spans:
                dev_arena[lo_:hi_].copy_(host_arena[lo_:hi_], non_blocking=True)
                ready.append(s_up.record_event())
        cur.wait_event(ready[0])
        p, conv = ops.iter_proj(dv["rays"], dv["pts"], dv["p0"], MATCH_CFG["max_iter"], MATCH_CFG["lambda_init"],
                                MATCH_CFG["convergence_thresh"])
        cur.wait_event(ready[1])
        (r,) = ops.refine_matches(dv["D11"], dv["D21"], p.long(), MATCH_CFG["radius"], MATCH_CFG["dilation_max"])
        s_down.wait_stream(cur)
        with torch.cuda.stream(s_down):
            out_host["idx"].copy_(r, non_blocking=True)
            out_host["conv"].copy_(conv, non_blocking=True)
        r.record_stream(s_down)
        conv.record_stream(s_down)
        cur.wait_event(ready[2])
        T = dv["Twc0"]
        ops.gauss_newton_rays(T, dv["Xs"], dv["Cs"], dv["ii"], dv["jj"], dv["idx"], dv["valid"], dv["Q"],
                              GN_RAYS["sigma_ray"], GN_RAYS["sigma_dist"], GN_RAYS["C_thresh"], GN_RAYS["Q_thresh"],
                              GN_RAYS["max_iter"], GN_RAYS["delta_thresh"])
        out_host["Twc"].copy_(T, non_blocking=True)
        cur.wait_stream(s_down)  # the step ends when every result is back in host memory

    barrier()
    te = timed(step_e2e, K, 3)
    barrier()
    lat_ms = max_over_ranks(sum(te)) / K

    # Throughput form of the same loop: two sets of device / pinned result buffers, step i+1's uploads are
    # enqueued before step i's ops, so the copy engine works while the SMs do (frames arrive as a stream; a
    # frame's inputs do not depend on the previous frame's result).  Every step still uploads all of its
    # inputs and reads back all of its results inside the timed region; K steps are timed as one region.
    dev_arena2 = (dev_arena, torch.empty_like(dev_arena))
    dv2 = (dv, arena_views(dev_arena2[1]))
    out2 = (out_host, out_host_b)
    done = [None, None]

    # What goes up per step: `full` = every input of the step (the headline e2e); `frame_only` = the keyframe side stays
    # resident on the device between frames tracked against the same keyframe -- its ray image, descriptors, points and
    # confidences do not change (what FrameTracker effectively does, tracker.py:31-75) -- and only the frame side, the
    # matches' initial guess and the graph slabs are uploaded.  Byte ranges of the arena, grouped by the op that waits for them.
    def rng(k_, row=None):
        nb_ = inp[k_].numel() * inp[k_].element_size()
        if row is None:
            return (offs[k_], offs[k_] + nb_)
        per = nb_ // inp[k_].shape[0]
        return (offs[k_] + row * per, offs[k_] + (row + 1) * per)

    def merged(rs):
        out_ = []
        for lo_, hi_ in sorted(rs):
            if out_ and lo_ - out_[-1][1] < 4096:
                out_[-1] = (out_[-1][0], hi_)
            else:
                out_.append((lo_, hi_))
        return out_

    upload_plan = {"full": [[sp] for sp in spans],
                   "frame_only": [merged([rng("pts"), rng("p0")]), merged([rng("D21")]),
                                  merged([rng("Twc0"), rng("Xs", 1), rng("Cs", 1), rng("ii"), rng("jj"), rng("idx"),
                                          rng("valid"), rng("Q")])]}
    mode = ["full"]

    def upload(slot):
        ready = []
        if done[slot] is not None:
            s_up.wait_event(done[slot])  # the step that last used this buffer set has finished
        with torch.cuda.stream(s_up):
            for group in upload_plan[mode[0]]:
                for lo_, hi_ in group:
                    dev_arena2[slot][lo_:hi_].copy_(host_arena[lo_:hi_], non_blocking=True)
                ready.append(s_up.record_event())
        return ready

    def compute(slot, ready):
        cur = torch.cuda.current_stream(dev)
        d_, o_ = dv2[slot], out2[slot]
        cur.wait_event(ready[0])
        p, conv = ops.iter_proj(d_["rays"], d_["pts"], d_["p0"], MATCH_CFG["max_iter"], MATCH_CFG["lambda_init"],
                                MATCH_CFG["convergence_thresh"])
        cur.wait_event(ready[1])
        (r,) = ops.refine_matches(d_["D11"], d_["D21"], p.long(), MATCH_CFG["radius"], MATCH_CFG["dilation_max"])
        s_down.wait_stream(cur)
        with torch.cuda.stream(s_down):
            o_["idx"].copy_(r, non_blocking=True)
            o_["conv"].copy_(conv, non_blocking=True)
        r.record_stream(s_down)
        conv.record_stream(s_down)
        cur.wait_event(ready[2])
        T = d_["Twc0"]
        ops.gauss_newton_rays(T, d_["Xs"], d_["Cs"], d_["ii"], d_["jj"], d_["idx"], d_["valid"], d_["Q"],
                              GN_RAYS["sigma_ray"], GN_RAYS["sigma_dist"], GN_RAYS["C_thresh"], GN_RAYS["Q_thresh"],
                              GN_RAYS["max_iter"], GN_RAYS["delta_thresh"])
        o_["Twc"].copy_(T, non_blocking=True)
        cur.wait_stream(s_down)
        done[slot] = cur.record_event()

    def pipelined(steps):
        cur = torch.cuda.current_stream(dev)
        s_up.wait_stream(cur)
        nxt = upload(0)
        for i in range(steps):
            ready = nxt
            if i + 1 < steps:
                nxt = upload((i + 1) & 1)
            compute(i & 1, ready)

    pipelined(3)
    torch.cuda.synchronize()
    barrier()
    flush_l2()
    ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ea.record()
    pipelined(K)
    eb.record()
    eb.synchronize()
    barrier()
    e2e_ms = max_over_ranks(ea.elapsed_time(eb)) / K
    full_out = {k_: v.clone() for k_, v in out2[0].items()}
    # same loop, keyframe side resident (both device buffer sets hold it from the full uploads above)
    mode[0] = "frame_only"
    done[0] = done[1] = None
    pipelined(3)
    torch.cuda.synchronize()
    barrier()
    flush_l2()
    ea2, eb2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ea2.record()
    pipelined(K)
    eb2.record()
    eb2.synchronize()
    barrier()
    e2e_res_ms = max_over_ranks(ea2.elapsed_time(eb2)) / K
    h2d_res = sum(hi_ - lo_ for group in upload_plan["frame_only"] for lo_, hi_ in group)
    res_ok = bool(all(torch.equal(out2[0][k_], full_out[k_]) for k_ in full_out))
    mode[0] = "full"
    pipe_ok = bool(torch.equal(out2[0]["Twc"], out2[1]["Twc"]) and torch.equal(out2[0]["idx"], out2[1]["idx"])
                   and torch.equal(out2[0]["Twc"], out_host["Twc"]))
    e2e = {"value": world * N_PX / (e2e_ms * 1e-3), "unit": "px/s", "ms_per_step": e2e_ms,
           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "latency_ms_per_step": lat_ms, "latency_px_per_s": world * N_PX / (lat_ms * 1e-3),
           "pipeline": "2 buffer sets: step i+1's H2D is enqueued before step i's ops; K steps timed as one region "
                       "(pipeline fill included); every step uploads all its inputs and reads back all its results (no L2 flush "
                       "inside the region: each step's inputs arrive fresh over PCIe into the other buffer set); "
                       "`latency_*` is the same loop without overlap between steps (each step timed alone, L2 flushed)",
           "results_identical_across_buffer_sets": pipe_ok,
           "keyframe_resident": {"value": world * N_PX / (e2e_res_ms * 1e-3), "unit": "px/s", "ms_per_step": e2e_res_ms,
                                 "h2d_bytes_per_step": h2d_res, "d2h_bytes_per_step": d2h,
                                 "results_identical_to_full_upload": res_ok,
                                 "what": "same pipelined loop; the keyframe's ray image, descriptors, points and confidences "
                                         "(rays, D11, Xs[0], Cs[0]) stay on the device between frames tracked against the same "
                                         "keyframe, every step uploads the frame side + match initialisation + graph slabs"},
           "host_numa_node": numa.node,  # pinned buffers first-touched on the GPU's node (None: sysfs does not say)
           "api": "mast3r_slam_backends.iter_proj / refine_matches / gauss_newton_rays (pybind drop-in), pinned host "
                  "buffers (one arena, three contiguous copies); H2D on a copy stream in op order, each op waits only for its own inputs, D2H on a third stream"}

    # ---- matching.match(): torch glue around the two ops vs the fused path, rank 0 numbers ---
    d_raw = synth.make_matching_pair(H, W, seed=1001 + 10 * rank, device=dev)

    def m_glue():
        matching.match(d_raw["X11"], d_raw["X21"], d_raw["D11"], d_raw["D21"], None, ops=ops)

    def m_fused():
        ops.match_fused(d_raw["X11"], d_raw["X21"], d_raw["D11"], d_raw["D21"].view(1, N_PX, -1))

    t_glue = statistics.median(timed(m_glue, reps, 3))
    t_fused = statistics.median(timed(m_fused, reps, 3))
    match_pipeline = {"what": "matching.match (prep + iter_proj + occlusion test + .half() + refine_matches + pixel_to_lin), "
                              "fp32 descriptors in, identity init, 512x384",
                      "torch_glue_ms": t_glue, "fused_ms": t_fused, "torch_glue_px_per_s": N_PX / (t_glue * 1e-3),
                      "fused_px_per_s": N_PX / (t_fused * 1e-3), "fused_launches": 5}

    # ---- global optimisation, edge-sharded: configs[2] (c3) AND configs[3] (c4) at every N ----------
    gopt = {}
    if args.global_opt != "none":
        comm = mdist.init_comm(rank, world, dev) if world > 1 else None
        for name in (("c3", "c4") if args.global_opt == "both" else (args.global_opt,)):
            gopt[name] = run_global_opt(args, name, abi, synth, mdist, dev, rank, world, comm, barrier, max_over_ranks, peak)
        if comm is not None:
            torch.cuda.synchronize()
            abi.comm_destroy(comm)

    # ---- factor-graph staging, one GPU ---------------------------------------
    fgraph = None
    if args.global_opt in ("c3", "both") and world == 1:
        torch.cuda.empty_cache()
        fgraph = run_factor_graph(abi, synth, dev, peak)

    # ---- frame tracker pose optimisation, one GPU ------------------------------
    ftrack = None
    if world == 1:
        ftrack = run_frame_tracker(synth, dev)

    # ---- sequence-level backend frames/s (BASELINE configs[4] stand-in), one GPU ---------------------
    sequence = None
    if world == 1 and not args.no_sequence:
        torch.cuda.empty_cache()
        sequence = run_sequence(synth, ops, dev)

    # ---- like-for-like baseline: the reference's own CUDA kernels on THIS GPU (rank 0, N=1) -------------------
    # oracle/_ref/ref_backends.so is the reference's backend built for sm_100 by oracle/build_ref.sh (test
    # infrastructure, also what the GPU parity tests compare against).  It is timed here exactly like the resident
    # step above -- same inputs, CUDA events, L2 flushed before every step -- and only reported; nothing of ours
    # runs through it.  Its gauss_newton_rays includes the reference's host Eigen solve and two syncs per iteration.
    # (part of the baseline leg: reported inside `cpu_baseline`, the one place bench.py may execute oracle/)
    gpu_ref = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        gpu_ref = reference_cuda_step(inp, timed, K, ms_per_step)

    # ---- CPU baseline on rank 0 at N=1 --------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference_step(inp, steps=1, warmup=0, budget_s=25.0)
        cpu["reference_cuda_same_gpu"] = gpu_ref

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "px/s", "n_gpus": world, "steps": K, "warmup": Wm,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32 (iter_proj, GN accumulate) / f16 (refine_matches) / f64 (assemble + Cholesky)",
                "data": "synthetic",
                "config": {"workload": WORKLOAD, "arm": "B200 kernels, one replica per GPU",
                           "l2": "L2 flushed (512 MB fill) before every timed step and kernel rep",
                           "parallelism": f"replicas x{world}" if world > 1 else "1 GPU"},
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches_per_step * K,
                "roofline": roofline, "kernels": kern, "cpu_baseline": cpu, "global_opt": gopt.get("c3"), "global_opt_c4": gopt.get("c4"),
                "match_pipeline": match_pipeline, "factor_graph": fgraph, "frame_tracker": ftrack,
                "sequence": sequence}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def reference_cuda_step(inp, timed, reps, ours_ms_per_step):
    import importlib
    import torch
    path = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.exists(os.path.join(path, "ref_backends.so")):
        return {"unavailable": "oracle/_ref/ref_backends.so not built (oracle/build_ref.sh needs /root/reference)"}
    if path not in sys.path:
        sys.path.insert(0, path)
    try:
        ref = importlib.import_module("ref_backends")
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"ref_backends import failed: {e}"}
    Twc = inp["Twc0"].clone()
    out = {}

    def r_iter():
        out["p"], out["c"] = ref.iter_proj(inp["rays"], inp["pts"], inp["p0"], MATCH_CFG["max_iter"], MATCH_CFG["lambda_init"],
                                           MATCH_CFG["convergence_thresh"])

    r_iter()
    p1 = out["p"].long()

    def r_refine():
        ref.refine_matches(inp["D11"], inp["D21"], p1, MATCH_CFG["radius"], MATCH_CFG["dilation_max"])

    def r_gn():
        Twc.copy_(inp["Twc0"])
        ref.gauss_newton_rays(Twc, inp["Xs"], inp["Cs"], inp["ii"], inp["jj"], inp["idx"], inp["valid"], inp["Q"],
                              GN_RAYS["sigma_ray"], GN_RAYS["sigma_dist"], GN_RAYS["C_thresh"], GN_RAYS["Q_thresh"],
                              GN_RAYS["max_iter"], GN_RAYS["delta_thresh"])

    def r_step():
        r_iter()
        ref.refine_matches(inp["D11"], inp["D21"], out["p"].long(), MATCH_CFG["radius"], MATCH_CFG["dilation_max"])
        r_gn()

    res = {"what": "the reference's own CUDA kernels + host Eigen solve (oracle/_ref/ref_backends.so, sm_100 build of its "
                   "unmodified backend sources) on the same GPU, same inputs, same timing as `value`"}
    for name, fn in (("iter_proj_ms", r_iter), ("refine_matches_ms", r_refine), ("gauss_newton_rays_x10_ms", r_gn), ("step_ms", r_step)):
        res[name] = statistics.median(timed(fn, max(reps, 10), 3))
    res["px_per_s"] = N_PX / (res["step_ms"] * 1e-3)
    res["ours_step_ms"] = ours_ms_per_step
    res["speedup_same_gpu"] = res["step_ms"] / ours_ms_per_step
    return res


GOPT_CFG = {"c3": dict(N=32, E=120, calib=False, seed=1003, configs_index=2),
            "c4": dict(N=256, E=1000, calib=True, seed=1004, configs_index=3)}


def run_global_opt(args, name, abi, synth, mdist, dev, rank, world, comm, barrier, max_over_ranks, peak):
    """Multi-keyframe gauss_newton_rays (config 3) / gauss_newton_calib (config 4), directed edges
    sharded over the ranks: every rank holds Xs/Cs/Twc and ONLY its own edges' slabs;
    one NCCL exchange of the fp64 edge blocks per iteration; strong scaling (the graph is fixed).
    At N > 1 rank 0 also keeps the full slabs and repeats the solve UNSHARDED (world-size-1 call, no
    communicator): `poses_bit_identical_to_unsharded` compares the two results bit for bit."""
    import hashlib
    import torch
    cfg = GOPT_CFG[name]
    g = synth.make_graph(cfg["N"], cfg["E"], H, W, seed=cfg["seed"], device=dev, calib=cfg["calib"])
    N, n, E2 = cfg["N"], N_PX, 2 * cfg["E"]
    order = mdist.pair_aligned_order(g["ii"], g["jj"])
    ii, jj = g["ii"][order].contiguous(), g["jj"][order].contiguous()
    parts = mdist.partition_edges(E2, world)
    lo, hi = parts[rank]
    exchange = mdist.exchange_for_partition(parts, E2) if args.exchange != "allreduce" else abi.EXCHANGE_ALLREDUCE
    # --exchange peer: the library's own all-gather kernel over NVLink peer memory instead of ncclAllGather (A/B on 2 and 8
    # GPUs: a tie to 3 % slower, profiles/r02peer_*; the NCCL call stays the default)
    xchg = None
    if world > 1 and exchange == abi.EXCHANGE_ALLGATHER and args.exchange == "peer":
        xchg = mdist.init_peer_exchange(rank, world, E2)
        exchange = abi.EXCHANGE_PEER
    own = order[lo:hi]
    idx, valid, Q = g["idx"][own].contiguous(), g["valid"][own].contiguous(), g["Q"][own].contiguous()
    full = None
    if world > 1 and rank == 0:
        full = (g["idx"][order].contiguous(), g["valid"][order].contiguous(), g["Q"][order].contiguous())
    del g["idx"], g["valid"], g["Q"]
    torch.cuda.empty_cache()
    shard = mdist.make_shard(lo, hi, comm, exchange, xchg) if world > 1 else None
    iters = 10
    if cfg["calib"]:
        prm = abi.make_params(abi.GN_CALIB, 1.0, 10.0, 0.0, 1.5, iters, 0.0, H, W, -10, 1e-6)
        K = g["K"].contiguous()
    else:
        prm = abi.make_params(abi.GN_RAYS, 0.003, 10.0, 0.0, 1.5, iters, 0.0)
        K = None
    ws, _ = abi.gn_workspace(N, E2, n, dev, edges_local=hi - lo)
    Twc = g["Twc"].clone()
    dx = torch.empty((N - 1, 7), dtype=torch.float32, device=dev)
    info = torch.empty(4, dtype=torch.int32, device=dev)

    def solve():
        Twc.copy_(g["Twc"])
        abi.gauss_newton(prm, Twc, g["Xs"], g["Cs"], K, ii, jj, idx, valid, Q, shard=shard, workspace=ws, info=info, dx=dx)

    reps = max(3, min(args.steps, 10))
    for _ in range(3):
        solve()
    barrier()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        a.record()
        solve()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b))
    ms = max_over_ranks(statistics.median(ts))
    info_h = [int(v) for v in info.cpu().tolist()]
    T_sharded = Twc.clone()
    # every rank must hold the same poses bit for bit (replicated deterministic solve on identical blocks)
    sha = hashlib.sha1(T_sharded.cpu().numpy().tobytes()).hexdigest()
    same_on_all_ranks = True
    if world > 1:
        import torch.distributed as dist
        box = [None] * world
        dist.all_gather_object(box, sha)
        same_on_all_ranks = all(b_ == box[0] for b_ in box)
    # accumulate alone (rank-local, no collective) for the roofline of the sharded kernel
    abi.gn_prepare(ii, jj, N, n, ws)
    sh_local = mdist.make_shard(lo, hi, None) if world > 1 else None
    ta, tp = [], []
    for r in range(3 + reps):
        a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        a.record()
        abi.gn_pack(prm, g["Cs"], idx, valid, Q, N, E2, ws, shard=sh_local)
        b.record()
        abi.gn_accumulate_packed(prm, g["Twc"], g["Xs"], K, E2, ws, shard=sh_local)
        c.record()
        c.synchronize()
        if r >= 3:
            tp.append(a.elapsed_time(b))
            ta.append(b.elapsed_time(c))
    acc_ms = max_over_ranks(statistics.median(ta))
    pack_ms = max_over_ranks(statistics.median(tp))
    # the replicated solve alone (assemble + Cholesky + substitutions + retraction from the blocks in the workspace)
    Ttmp = g["Twc"].clone()
    tsv = []
    for r in range(3 + reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        Ttmp.copy_(g["Twc"])
        abi.gn_prepare(ii, jj, N, n, ws)  # resets the iteration / stop state; the edge blocks stay
        a.record()
        abi.lib.m3r_gn_solve_retract(abi._ptr(Ttmp), N, n, E2, abi.C.c_float(0.0), abi._ptr(dx), abi._ptr(info),
                                     abi._ptr(ws), abi.C.c_size_t(ws.numel()), abi._stream())
        b.record()
        b.synchronize()
        if r >= 3:
            tsv.append(a.elapsed_time(b))
    solve_ms = max_over_ranks(statistics.median(tsv))
    touched = int(torch.unique(torch.cat((ii[lo:hi], jj[lo:hi]))).numel())
    nbytes = gn_bytes(hi - lo, touched)
    gerr = float((T_sharded - g["Twc_gt"]).abs().max().item())
    # unsharded repeat on rank 0 (world-size-1 call on the full slabs): bit-identity of the sharded result
    identical = True if world == 1 else None
    if full is not None:
        del ws
        torch.cuda.empty_cache()
        ws1, _ = abi.gn_workspace(N, E2, n, dev)
        T1 = g["Twc"].clone()
        abi.gauss_newton(prm, T1, g["Xs"], g["Cs"], K, ii, jj, full[0], full[1], full[2], shard=None, workspace=ws1,
                         info=info, dx=dx)
        torch.cuda.synchronize()
        identical = bool(torch.equal(T1, T_sharded))
        del ws1, full
    barrier()
    if xchg is not None:
        mdist.destroy_peer_exchange(xchg)
    torch.cuda.empty_cache()
    return {"config": f"configs[{cfg['configs_index']}]: {N} keyframes / {E2} directed edges, 512x384, "
                      f"{'gauss_newton_calib' if cfg['calib'] else 'gauss_newton_rays'} x{iters} iters",
            "scaling": "strong", "n_gpus": world, "ms_per_solve": ms, "iters": iters, "info": info_h,
            "gn_edge_iters_per_s": E2 * iters / (ms * 1e-3), "accumulate_ms_per_iter": acc_ms,
            "accumulate_GBps_per_gpu": nbytes / (acc_ms * 1e-3) / 1e9,
            "accumulate_frac_of_hbm_peak": nbytes / (acc_ms * 1e-3) / 1e9 / peak,
            "accumulate_frac_of_nominal_8TBps": nbytes / (acc_ms * 1e-3) / 1e9 / 8000.0,
            "whole_op_frac_of_hbm_peak": gn_bytes(E2, N) / (ms / iters * 1e-3) / 1e9 / (peak * world),
            "accumulate_bytes_per_rank": nbytes, "pack_ms_per_solve": pack_ms,
            "accumulate_plus_amortised_pack_frac_of_hbm_peak": nbytes / ((acc_ms + pack_ms / iters) * 1e-3) / 1e9 / peak,
            "accumulate_streamed_bytes_per_rank": 8 * (hi - lo) * n + 12 * touched * n,
            "solve_alone_ms": solve_ms, "solve_dim": 7 * (N - 1),
            "solve_and_comm_ms_per_iter": (ms - pack_ms) / iters - acc_ms,
            "collective": (({abi.EXCHANGE_PEER: "own kernel: all-gather of the [E2,56] fp64 edge blocks by 16-byte stores into the peers' buffers over "
                                                "NVLink (CUDA IPC) + release/acquire flags, once per iteration, no NCCL on the data path",
                             abi.EXCHANGE_ALLGATHER: "in-place ncclAllGather(f64) of the [E2,56] edge blocks (equal contiguous edge ranges), once per iteration",
                             abi.EXCHANGE_ALLREDUCE: "ncclAllReduce(sum, f64) of the zero-padded [E2,56] edge blocks, once per iteration"}[exchange])
                           if world > 1 else None),
            "poses_bit_identical_to_unsharded": identical, "poses_identical_on_all_ranks": same_on_all_ranks,
            "poses_sha1": sha, "max_abs_pose_err_vs_gt": gerr,
            "working_set": "13 B/pt/edge slabs (>= 0.6 GB per rank at c3/8) exceed the 126 MB L2; not flushed"}


def run_sequence(synth, ops, dev, num_frames=60, kf_every=6):
    """BASELINE configs[4] stand-in ("SLAM frames/s"): the reference's main loop (main.py:237-326) cannot be imported here
    (lietorch, thirdparty/monst3r and the ViT weights are absent), so this strings the repo's own mirrors of its backend
    calls over a synthetic 512x384 sequence, in the reference's order:
        per frame    mast3r_match_asymmetric -> matching.match (tracker.py:31-37; fused op) -> valid / Q (:41-67)
                     -> FrameTracker.opt_pose_ray_dist_sim3 (:156-214; PoseOptimizer)
        per keyframe FactorGraph.add_factors against the previous keyframes (main.py:133, global_opt.py:31-99; symmetric
                     matching = two fused match calls per pair) -> solve_GN_rays (main.py:142, global_opt.py:121-153)
    The ViT is EXCLUDED: its outputs (pointmaps in the other view's frame, confidences, descriptors) are synthesised from
    ground truth + noise, which is what the two-view network hands the backend.  A frame becomes a keyframe when the
    reference's test fires (match fraction < match_frac_thresh, main.py:300-305) or every `kf_every` frames (added so that the
    60-frame sequence exercises the global optimisation).  Reported: backend frames/s (wall, one sync per frame like the
    reference's .item()), per-stage times, and the trajectory error against ground truth."""
    import torch
    fgm = importlib.import_module("monst3r-slam_b200.factor_graph")
    trk = importlib.import_module("monst3r-slam_b200.tracker")
    seq = synth.make_sequence(num_frames, H, W, fdim=FDIM, seed=1005, device=dev)
    n = N_PX
    gt = seq["Twc_gt"].to(dev)
    gen = torch.Generator(device="cpu").manual_seed(7)

    def in_frame_of(a, b):
        """points of frame b expressed in frame a's coordinates (what the two-view network predicts), + noise"""
        Tab = synth.sim3_act(synth.sim3_inv(gt[a].double()), synth.sim3_act(gt[b].double(), seq["X"][b].double()))
        return (Tab + 1e-3 * torch.randn(n, 3, generator=gen, dtype=torch.float64).to(dev)).float().view(1, H, W, 3).contiguous()

    t_vit = [0.0]  # time spent synthesising the two-view network's outputs: NOT backend time, subtracted below

    def fused_match(a, b):
        """matching.match(X_aa, X_ba, D_aa, D_ba): for every pixel of b the matching pixel of a (idx_a2b), valid mask"""
        torch.cuda.synchronize()
        ts = time.perf_counter()
        X_ba = in_frame_of(a, b)
        torch.cuda.synchronize()
        t_vit[0] += time.perf_counter() - ts
        return ops.match_fused(seq["X"][a].view(1, H, W, 3), X_ba, seq["D"][a].view(1, H, W, FDIM), seq["D"][b].view(1, n, FDIM))

    frames = fgm.KeyframeBuffer(n, capacity=num_frames, device=dev)
    graph = fgm.FactorGraph(frames=frames, device=dev, capacity=64)
    opt = trk.PoseOptimizer()
    kf_frame = []  # sequence index of every keyframe

    def add_keyframe(f, T):
        kf_frame.append(f)
        return frames.append(seq["X"][f], seq["C"][f], T)

    add_keyframe(0, gt[0])
    T_prev = gt[0].view(1, 8).clone()
    est = [gt[0].clone()]
    t_track, t_kf, iters = [], [], []
    since_kf = 0
    torch.cuda.synchronize()
    t_all = time.perf_counter()
    for f in range(1, num_frames):
        t0 = time.perf_counter()
        v0 = t_vit[0]
        k = len(kf_frame) - 1
        fk = kf_frame[k]
        idx_f2k, valid_k = fused_match(f, fk)                       # tracker.py:31-37
        idx = idx_f2k[0]
        Qk = torch.sqrt(seq["Q"][f][idx] * seq["Q"][fk])            # :41
        valid_opt = valid_k[0] & (Qk > opt.cfg["Q_conf"])           # :62-67 (C_conf = 0: confidences are > 1)
        if os.environ.get("M3R_SEQ_VERBOSE"):
            print(f"  track {f}: valid {float(valid_opt.float().mean()):.3f} T_prev {[round(x, 4) for x in T_prev.view(8).tolist()]} "
                  f"T_WCk {[round(x, 4) for x in frames.T_WC[k].tolist()]}", file=sys.stderr)
        T_WCf, T_CkCf = opt.opt_pose_ray_dist_sim3(seq["X"][f][idx].contiguous(), seq["X"][fk], T_prev, frames.T_WC[k].view(1, 8),
                                                   Qk, valid_opt)   # :85-88
        iters.append(int(opt.last_info[0]))
        if os.environ.get("M3R_SEQ_VERBOSE"):
            print(f"frame {f} vs kf {k} (frame {fk}): valid {float(valid_opt.float().mean()):.3f} iters {iters[-1]} "
                  f"|t - gt| {float((T_WCf.view(8)[:3] - gt[f, :3]).norm()):.2e}", file=sys.stderr)
        T_prev = T_WCf.clone()
        est.append(T_WCf.view(8).clone())
        since_kf += 1
        match_frac = float(valid_opt.float().mean())
        new_kf = match_frac < 0.333 or since_kf >= kf_every        # main.py:300-305 (+ the periodic rule, see docstring)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        t_track.append((t1 - t0 - (t_vit[0] - v0)) * 1e3)
        v1 = t_vit[0]
        if new_kf:
            since_kf = 0
            j = add_keyframe(f, T_WCf.view(8))
            prev = list(range(max(0, j - 3), j))                    # the last keyframes
            if j >= 5:  # retrieval stand-in (main.py:117-131): the nearest older keyframe, if the views overlap
                dpos = (gt[kf_frame[:j - 3], :3] - gt[f, :3]).norm(dim=1)
                if float(dpos.min()) < 0.2:
                    prev = [int(dpos.argmin())] + prev
            m = [[], [], [], [], [], [], [], []]
            for i in prev:
                fi = kf_frame[i]
                idx_i2j, valid_j = fused_match(fi, f)               # mast3r_match_symmetric: both orders (global_opt.py:40-52)
                idx_j2i, valid_i = fused_match(f, fi)
                for lst, t in zip(m, (idx_i2j, idx_j2i, valid_j, valid_i, seq["Q"][fi][None], seq["Q"][f][None],
                                      seq["Q"][f][None], seq["Q"][fi][None])):
                    lst.append(t)
            graph.add_factors_from_matches(prev, [j] * len(prev), tuple(torch.cat(x).contiguous() for x in m), 0.1)
            graph.solve_GN_rays()
            if os.environ.get("M3R_SEQ_VERBOSE"):
                print(f"keyframe {j} (frame {f}): pairs {prev} -> edges {graph.num_edges}, |t - gt| "
                      f"{float((frames.T_WC[j, :3] - gt[f, :3]).norm()):.2e}", file=sys.stderr)
            T_prev = frames.T_WC[j].view(1, 8).clone()
            torch.cuda.synchronize()
            t_kf.append((time.perf_counter() - t1 - (t_vit[0] - v1)) * 1e3)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t_all - t_vit[0]
    est = torch.stack(est)
    # Sim(3) gauge: the first pose is ground truth and keyframe 0 is fixed, so poses are directly comparable
    terr = (est[:, :3] - gt[:, :3]).norm(dim=1)
    kf_err = (frames.T_WC[:len(kf_frame), :3] - gt[kf_frame, :3]).norm(dim=1)
    return {"what": "BASELINE configs[4] stand-in: reference main-loop order over a synthetic 512x384 sequence, ViT excluded "
                    "(its outputs synthesised from ground truth + noise); fused match -> PoseOptimizer per frame, "
                    "FactorGraph.add_factors + solve_GN_rays per keyframe",
            "frames": num_frames, "keyframes": len(kf_frame), "graph_edges_undirected": graph.num_edges,
            "backend_frames_per_s": (num_frames - 1) / wall, "backend_wall_s": wall, "excluded_network_synthesis_s": t_vit[0],
            "tracking_ms_per_frame_median": statistics.median(t_track),
            "keyframe_ms_median": statistics.median(t_kf) if t_kf else None,
            "tracker_iterations_median": statistics.median(iters),
            "ate_rmse_m": float(torch.sqrt((terr ** 2).mean())), "max_translation_err_m": float(terr.max()),
            "keyframe_translation_err_after_global_opt_m": float(kf_err.max())}


def run_frame_tracker(synth, dev):
    """FrameTracker.opt_pose_ray_dist_sim3 (tracker.py:156-214) at 512x384 with config/base.yaml's tracking thresholds:
    one persistent launch (m3r_track_frame) vs a torch restatement of ONE iteration of the reference loop
    (act_Sim3 + point_to_ray_dist Jacobians + FrameTracker.solve, tracker.py:141-154,169-176; lietorch is absent,
    so the Sim(3) action is written out and the retraction is left out) on the same GPU."""
    import torch
    trk = importlib.import_module("monst3r-slam_b200.tracker")
    g = synth.make_graph(2, 1, H, W, seed=1002, device=dev, calib=False)
    e = 1  # ii = 1 (frame, gathered side), jj = 0 (keyframe)
    Xf = g["Xs"][1][g["idx"][e]].contiguous()
    Xk = g["Xs"][0].contiguous()
    Q = g["Q"][e].reshape(-1, 1).contiguous()
    valid = (g["valid"][e].reshape(-1, 1) & (Q > 1.5)).contiguous()
    T_WCf, T_WCk = g["Twc"][1].reshape(1, 8).contiguous(), g["Twc"][0].reshape(1, 8).contiguous()
    opt = trk.PoseOptimizer()

    def ev():
        return torch.cuda.Event(enable_timing=True)

    def timed(fn, reps=10, warm=3):
        ts, ws = [], []
        for r in range(reps + warm):
            a, b = ev(), ev()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            a.record()
            fn()
            b.record()
            b.synchronize()
            if r >= warm:
                ws.append((time.perf_counter() - t0) * 1e3)
                ts.append(a.elapsed_time(b))
        return statistics.median(ts), statistics.median(ws)

    out = {}

    def ours():
        out["T"], out["Trel"] = opt.opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Q, valid)

    dev_ms, wall_ms = timed(ours)
    iters = int(opt.last_info[0])
    opt_async = trk.PoseOptimizer(sync_check=False)
    dev_async_ms, _ = timed(lambda: opt_async.opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Q, valid))
    gerr = float((out["T"].reshape(-1) - g["Twc_gt"][1]).abs().max().item())

    # torch restatement of one reference iteration (fp32, the reference's statements)
    c = opt.cfg
    t_, q_, s_ = T_WCf[0, :3], T_WCf[0, 3:7], T_WCf[0, 7]
    x, y, z, w = q_
    R = torch.stack((1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w), 2 * (x * y + z * w),
                     1 - 2 * (x * x + z * z), 2 * (y * z - x * w), 2 * (x * z - y * w), 2 * (y * z + x * w),
                     1 - 2 * (x * x + y * y))).view(3, 3)
    sqrt_info_ray = 1 / c["sigma_ray"] * valid * torch.sqrt(Q)
    sqrt_info_dist = 1 / c["sigma_dist"] * valid * torch.sqrt(Q)
    sqrt_info = torch.cat((sqrt_info_ray.repeat(1, 3), sqrt_info_dist), dim=1)

    def rd(X, jac):
        d = torch.linalg.norm(X, dim=-1, keepdim=True)
        d_inv = 1.0 / d
        r = d_inv * X
        out_ = torch.cat((r, d), dim=-1)
        if not jac:
            return out_
        I = torch.eye(3, device=X.device, dtype=X.dtype).repeat(X.shape[0], 1, 1)
        dr = d_inv.unsqueeze(-1) * (I - (d_inv ** 2).unsqueeze(-1) * (X.unsqueeze(-1) @ X.unsqueeze(-2)))
        return out_, torch.cat((dr, r.unsqueeze(-2)), dim=-2)

    rd_k = rd(Xk, False)

    def ref_iteration():
        pW = s_ * (Xf @ R.T) + t_
        o = torch.zeros_like(pW[:, 0])
        sk = torch.stack([o, -pW[:, 2], pW[:, 1], pW[:, 2], o, -pW[:, 0], -pW[:, 1], pW[:, 0], o], dim=-1).view(-1, 3, 3)
        dX = torch.cat([torch.eye(3, device=dev).repeat(pW.shape[0], 1, 1), -sk, pW.reshape(-1, 3, 1)], dim=-1)
        rd_f, drd = rd(pW, True)
        r = rd_k - rd_f
        J = -drd @ dX
        whitened = sqrt_info * r
        r_abs = torch.abs(whitened)
        wgt = torch.where(r_abs < c["huber"], torch.ones(1, device=dev), c["huber"] / r_abs)
        robust = sqrt_info * torch.sqrt(wgt)
        A = (robust[..., None] * J).view(-1, 7)
        b = (robust * r).view(-1, 1)
        Hm = A.T @ A
        gv = -A.T @ b
        cost = 0.5 * (b.T @ b).item()
        L = torch.linalg.cholesky(Hm, upper=False)
        return torch.cholesky_solve(gv, L, upper=False).view(1, -1), cost

    ref_dev_ms, ref_wall_ms = timed(ref_iteration, reps=5, warm=2)
    return {"what": "SURVEY 8(f) rank 3: FrameTracker.opt_pose_ray_dist_sim3 at 512x384, config/base.yaml tracking thresholds "
                    "(rel_error 1e-3, delta_norm 1e-3, max_iters 50)",
            "iterations": iters, "ms_per_frame": wall_ms, "device_ms_per_frame": dev_ms,
            "device_ms_per_frame_without_info_readback": dev_async_ms, "launches_per_frame": 5,
            "px_per_s": N_PX / (wall_ms * 1e-3),
            "reference_torch_ms_per_iteration": ref_wall_ms, "reference_torch_device_ms_per_iteration": ref_dev_ms,
            "reference_torch_ms_per_frame_estimate": ref_wall_ms * iters,
            "reference_note": "torch restatement of ONE iteration of tracker.py:141-154,169-176 (no retraction; lietorch is "
                              "absent) x the iterations our kernel executed",
            "max_abs_pose_err_vs_gt": gerr}


def run_factor_graph(abi, synth, dev, peak):
    """Config-3 graph fed pair by pair into the device-resident FactorGraph (factor_graph.py), against a torch
    restatement of the reference's staging (global_opt.py:54-58, 89-96, 104-119) on the same GPU."""
    import torch
    fgm = importlib.import_module("monst3r-slam_b200.factor_graph")
    N, E, batch = 32, 120, 8
    g = synth.make_graph(N, E, H, W, seed=1003, device=dev, calib=False)
    n = N_PX
    frames = fgm.KeyframeBuffer(n, capacity=N, device=dev)
    for k in range(N):
        frames.append(g["Xs"][k], g["Cs"][k], g["Twc"][k])
    ones = torch.ones((batch, n, 1), dtype=torch.float32, device=dev)
    Q2 = g["Q"] ** 2
    feeds = []
    for lo in range(0, E, batch):
        hi = min(lo + batch, E)
        feeds.append((g["ii"][lo:hi].tolist(), g["jj"][lo:hi].tolist(),
                      (g["idx"][lo:hi], g["idx"][E + lo:E + hi], g["valid"][lo:hi], g["valid"][E + lo:E + hi],
                       ones[:hi - lo], ones[:hi - lo], Q2[lo:hi], Q2[E + lo:E + hi])))

    def ev():
        return torch.cuda.Event(enable_timing=True)

    # (1) admit kernel alone: all 15 batches back to back into a scratch store (1.4 GB of cold inputs)
    st = fgm.EdgeStore(n, E, dev)
    t_admit = []
    for rep in range(4):
        a, b = ev(), ev()
        a.record()
        for k, (_, _, m) in enumerate(feeds):
            abi.edge_admit(*m, 1.5, st.idx, st.valid, st.Q, k * batch)
        b.record()
        b.synchronize()
        if rep:
            t_admit.append(a.elapsed_time(b))
    admit_ms = statistics.median(t_admit)
    admit_bytes = 2 * E * n * 30
    # ... and as ONE launch over all 120 pairs (no launch gaps: the kernel's own bandwidth)
    one = (g["idx"][:E], g["idx"][E:], g["valid"][:E], g["valid"][E:], torch.ones((E, n, 1), dtype=torch.float32, device=dev),
           torch.ones((E, n, 1), dtype=torch.float32, device=dev), Q2[:E], Q2[E:])
    t_one = []
    for rep in range(4):
        a, b = ev(), ev()
        a.record()
        abi.edge_admit(*one, 1.5, st.idx, st.valid, st.Q, 0)
        b.record()
        b.synchronize()
        if rep:
            t_one.append(a.elapsed_time(b))
    admit_one_ms = statistics.median(t_one)
    del st, one
    # (2) add_factors end to end (kernel + accept/reject bookkeeping + its one host sync), growth from capacity 8
    add_ms = {}
    for cap in (8, E, E):  # growth from 8 pairs (cudaMalloc + copies); preallocated (second run: warm)
        fg = fgm.FactorGraph(frames=frames, device=dev, capacity=cap, cfg={"delta_norm": 0.0})
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for f_ii, f_jj, m in feeds:
            fg.add_factors_from_matches(f_ii, f_jj, m, 0.1)
        torch.cuda.synchronize()
        add_ms[cap] = (time.perf_counter() - t0) * 1e3
        if cap == 8:
            grown = fg.store.moved_bytes
            del fg
            torch.cuda.empty_cache()
    E_acc = fg.num_edges
    # (3) per-solve staging: zero-copy views vs the reference's cat/stack statements
    ref_parts = [fg.ii.clone(), fg.jj.clone(), fg.idx_ii2jj.contiguous(), fg.idx_jj2ii.contiguous(),
                 fg.valid_match_j.contiguous(), fg.valid_match_i.contiguous(), fg.Q_ii2jj.contiguous(),
                 fg.Q_jj2ii.contiguous()]
    kf_X = [frames.X_canon[k].clone() for k in range(N)]
    kf_C = [frames.C[k].clone() for k in range(N)]
    kf_T = [frames.T_WC[k].clone() for k in range(N)]

    def stage_reference():
        r_ii, r_jj, i1, i2, v1, v2, q1, q2 = ref_parts
        torch.unique(torch.cat([r_ii, r_jj]), sorted=True)
        out = (torch.stack(kf_X), torch.stack(kf_T), torch.stack(kf_C), torch.cat((r_ii, r_jj)), torch.cat((r_jj, r_ii)),
               torch.cat((i1, i2)), torch.cat((v1, v2)), torch.cat((q1, q2)))
        return out

    def stage_store():
        u = fg.get_unique_kf_idx()
        return fg.get_poses_points(u), fg.prep_two_way_edges()

    def timed(fn, reps=5):
        ts = []
        for r in range(reps + 2):
            a, b = ev(), ev()
            torch.cuda.synchronize()
            t = time.perf_counter()
            a.record()
            out = fn()
            b.record()
            b.synchronize()
            wall = (time.perf_counter() - t) * 1e3
            del out
            if r >= 2:
                ts.append(max(a.elapsed_time(b), wall))
        return statistics.median(ts)

    stage_ref_ms = timed(stage_reference)
    stage_store_ms = timed(stage_store)
    pose0 = frames.T_WC.clone()

    def solve():
        frames.T_WC.copy_(pose0)
        fg.solve_GN_rays()

    solve_ms = timed(solve)
    gerr = float((frames.T_WC[:N] - g["Twc_gt"]).abs().max().item())
    return {"what": "SURVEY 8(f) rank 2: config-3 graph (32 kf) fed in batches of 8 pairs to the device-resident FactorGraph",
            "pairs_offered": E, "pairs_accepted": E_acc, "edge_admit_ms_total": admit_ms,
            "edge_admit_GBps": admit_bytes / (admit_ms * 1e-3) / 1e9,
            "edge_admit_frac_of_hbm_peak": admit_bytes / (admit_ms * 1e-3) / 1e9 / peak,
            "edge_admit_bytes": admit_bytes, "edge_admit_bytes_per_pt": 30,
            "edge_admit_single_launch_ms": admit_one_ms,
            "edge_admit_single_launch_GBps": admit_bytes / (admit_one_ms * 1e-3) / 1e9,
            "edge_admit_single_launch_frac_of_hbm_peak": admit_bytes / (admit_one_ms * 1e-3) / 1e9 / peak,
            "add_factors_wall_ms_15_batches_preallocated": add_ms[E], "add_factors_wall_ms_15_batches_growing_from_8": add_ms[8],
            "store_growth_bytes_moved": grown,
            "per_solve_staging_ms_reference_cat_stack": stage_ref_ms, "per_solve_staging_ms_store_views": stage_store_ms,
            "solve_GN_rays_ms_through_store": solve_ms, "max_abs_pose_err_vs_gt": gerr}


# ------------------------------------------------------------------------------------------ CPU arm
def _np(t):
    return t.detach().cpu().numpy()


def cpu_reference_step(inp, steps=1, warmup=0, budget_s=25.0):
    """The reference's CPU path for one tracking step, timed on this box's host cores: the reference's
    own iter_proj / refine_matches kernel source built for the host (oracle/_ref/libref_matching_host.so,
    kind "reference") and the line-by-line host port of ray_align_kernel + Eigen-equivalent fp64 solve
    (oracle/cpu_oracle.cpp, kind "port": the reference's GN has no host build — Eigen + CUDA kernel).
    Bounded sample: the first `n_s` of the 196 608 query pixels for matching and `it_s` of the 10 GN
    iterations, sized from a calibration pass so one step fits `budget_s`; scaled to whole steps."""
    import numpy as np
    from oracle import oracle as orc
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    # torchrun exports OMP_NUM_THREADS=1 and torch has initialised the OpenMP runtime by now, so the environment no
    # longer decides: set the thread count of the runtime the oracle libraries share (libgomp) explicitly
    os.environ["OMP_NUM_THREADS"] = str(cores)
    try:
        import ctypes
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(int(cores))
    except OSError:
        pass
    impl = "reference" if orc.have_ref_host() else "port"
    rays, pts, p0 = _np(inp["rays"]), _np(inp["pts"]), _np(inp["p0"])
    D11, D21 = _np(inp["D11"]), _np(inp["D21"])
    gn = [_np(inp[k]) for k in ("Twc0", "Xs", "Cs", "ii", "jj", "idx", "valid", "Q")]

    def matching(n_s):
        t0 = time.perf_counter()
        p, _ = orc.iter_proj(rays, pts[:, :n_s], p0[:, :n_s], MATCH_CFG["max_iter"], MATCH_CFG["lambda_init"],
                             MATCH_CFG["convergence_thresh"], impl=impl)
        t1 = time.perf_counter()
        orc.refine_matches(D11, D21[:, :n_s], p.astype(np.int64), MATCH_CFG["radius"], MATCH_CFG["dilation_max"], impl=impl)
        t2 = time.perf_counter()
        return t1 - t0, t2 - t1

    def gn_run(it_s):
        t0 = time.perf_counter()
        orc.gauss_newton_rays(*gn, GN_RAYS["sigma_ray"], GN_RAYS["sigma_dist"], GN_RAYS["C_thresh"], GN_RAYS["Q_thresh"],
                              it_s, GN_RAYS["delta_thresh"])
        return time.perf_counter() - t0

    # calibration: 1/16 of the pixels, 1 GN iteration
    n_c = N_PX // 16
    ti, tr = matching(n_c)
    tg = gn_run(1)
    full = (ti + tr) * 16 + tg * GN_RAYS["max_iter"]
    frac = min(1.0, budget_s / full) if full > 0 else 1.0
    n_s = max(4096, min(N_PX, int(N_PX * frac) // 1024 * 1024))
    it_s = max(1, min(GN_RAYS["max_iter"], int(round(GN_RAYS["max_iter"] * frac))))
    est = []
    parts = None
    for s in range(warmup + steps):
        ti, tr = matching(n_s)
        tg = gn_run(it_s)
        step_s = (ti + tr) * (N_PX / n_s) + tg * (GN_RAYS["max_iter"] / it_s)
        if s >= warmup:
            est.append(step_s)
            parts = {"iter_proj_s": ti * N_PX / n_s, "refine_matches_s": tr * N_PX / n_s,
                     "gauss_newton_rays_s": tg * GN_RAYS["max_iter"] / it_s}
    step_s = statistics.median(est)
    return {"value": N_PX / step_s, "unit": "px/s", "cores": cores,
            # the GN port parallelises over directed edges like the reference's one-block-per-edge grid: E2 = 2 here
            "gn_threads_used": min(cores, int(gn[3].shape[0])), "matching_threads_used": cores,
            "kind": "reference" if impl == "reference" else "port",
            "sample": f"{n_s} of {N_PX} query px through the reference's matching kernels built for the host "
                      f"({impl}) + {it_s} of {GN_RAYS['max_iter']} GN iterations (N=2, E2=2, all {N_PX} points) "
                      f"through the host port of ray_align_kernel + fp64 LLT (edge-parallel: 2 edges -> 2 of the {cores} threads busy in that leg); "
                      f"scaled to one whole step",
            "s_per_step": step_s, "parts_s_per_step": parts}


def run_reference(args):
    """--impl reference: the reference's CPU path for the same step, on the host cores of this box."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    import torch
    # synth + torch glue ONLY: the reference arm must not map libm3r_b200.so (abi / dist / the drop-in stay unimported)
    synth = importlib.import_module("monst3r-slam_b200.synth")
    matching = importlib.import_module("monst3r-slam_b200.matching")
    inp = make_tracking_inputs(synth, matching, torch.device("cpu"), 1001)
    K, Wm = args.steps, args.warmup
    budget = max(4.0, min(25.0, 150.0 / max(1, K + Wm)))
    cpu = cpu_reference_step(inp, steps=K, warmup=Wm, budget_s=budget)
    line = {"impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": "px/s", "n_gpus": world,
            "steps": K, "warmup": Wm, "ms_per_step": cpu["s_per_step"] * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 / f16 (c10::Half on host) / f64 solve", "data": "synthetic",
            "config": {"workload": WORKLOAD, "arm": "reference CPU path, rank 0 only",
                       "parallelism": f"host cores x{cpu['cores']} (OpenMP)"},
            "cpu_baseline": cpu,
            "e2e": {"value": cpu["value"], "unit": "px/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--global-opt", default="both", choices=["both", "c3", "c4", "none"])
    ap.add_argument("--exchange", default="auto", choices=["auto", "peer", "nccl", "allreduce"],
                    help="edge-block exchange at N > 1: auto = all-gather when the partition is equal, else all-reduce")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sequence", action="store_true")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: anything a library prints on fd 1 meanwhile (NCCL's version banner,
    # torchrun notices) is sent to stderr, and the line is written to the real stdout at the end.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
