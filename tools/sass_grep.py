#!/usr/bin/env python
"""Per-kernel SASS evidence for lib/libm3r_b200.so (cuobjdump -sass): counts of the mnemonics that show what each kernel
is made of -- UBLKCP (cp.async.bulk, TMA bulk copies), SYNCS (mbarrier), HMUL2/HADD2 (packed fp16), DFMA / DMMA (fp64 SIMT /
tensor), REDUX, LDGSTS, MEMBAR, BAR.  Usage: python tools/sass_grep.py > profiles/r02_sass_grep.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "monst3r-slam_b200", "lib", "libm3r_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], stdout=subprocess.PIPE, text=True).stdout
pats = ["UBLKCP", "UTMALDG", "SYNCS", "HMUL2", "HADD2", "HFMA2", "DFMA", "DMMA", "FFMA", "REDUX", "LDGSTS", "MEMBAR", "BAR.", "SHFL", "LDS", "STS", "LDG", "STG", "MUFU"]
cur, counts, order = None, collections.defaultdict(collections.Counter), []
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
        cur = re.sub(r"\(anonymous namespace\)::", "", cur)
        cur = re.sub(r"\(.*", "", cur)[:80]
        order.append(cur)
        continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        counts[cur]["_total"] += 1
        for p_ in pats:
            if p_ in ln:
                counts[cur][p_] += 1
print("# cuobjdump -sass monst3r-slam_b200/lib/libm3r_b200.so : instruction counts per kernel (sm_100a)")
print("%-82s %7s " % ("kernel", "instrs") + " ".join("%7s" % p_ for p_ in pats))
for k in order:
    c = counts[k]
    print("%-82s %7d " % (k, c["_total"]) + " ".join("%7d" % c[p_] for p_ in pats))
