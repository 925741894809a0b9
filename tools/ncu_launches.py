#!/usr/bin/env python
"""Filter an `ncu --metrics gpu__time_duration.sum --csv` log down to this library's kernels and add per-kernel totals.
Usage: python tools/ncu_launches.py gpurun_out/raw.csv profiles/rXX_ncu_launches.csv "<the profiled command>"
"""
import csv
import re
import sys

OURS = re.compile(r"iter_proj|refine_|gn_|match_prep|pixel_to_lin|f32_to_f16|edge_admit|ray_constrain|track_|pose_retr|export_hs")


def main():
    raw, out, cmd = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "")
    lines = [l for l in open(raw, errors="replace") if l.startswith('"')]
    rows = list(csv.reader(lines))
    hdr = rows[0]
    ik, iv, ii = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("ID")
    iu = hdr.index("Metric Unit")
    keep, tot = [], {}
    for r in rows[1:]:
        if len(r) <= iv or not OURS.search(r[ik]):
            continue
        ns = float(r[iv].replace(",", "")) * {"ns": 1.0, "us": 1e3, "ms": 1e6}.get(r[iu], 1.0)
        name = r[ik].split("(")[0][:80]
        keep.append((r[ii], name, ns))
        t = tot.setdefault(name, [0, 0.0])
        t[0] += 1
        t[1] += ns
    total = sum(t[1] for t in tot.values()) or 1.0
    with open(out, "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none: {cmd} (our kernels only; ns, cold-cache, serialised)\n")
        f.write("id,kernel,gpu__time_duration_ns\n")
        for i, n, ns in keep:
            f.write(f"{i},{n},{ns:.0f}\n")
        f.write("# totals per kernel: name,launches,total_ns,share\n")
        for n, (c, ns) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
            f.write(f"# {n},{c},{ns:.0f},{ns / total:.4f}\n")
    print("wrote", out, len(keep), "launches of", len(tot), "kernels")


if __name__ == "__main__":
    main()
