#!/usr/bin/env python
"""SASS evidence for profiles/: per kernel the count of the mnemonics that prove the
bulk-copy (UBLKCP), mbarrier (SYNCS), cp.async (LDGSTS), vector reduction (REDG ... F32x4)
and warp-collective instructions and the cluster barrier (UCGABAR_*), plus an excerpt of the fast march loop of
integrate_keyframes_kernel<true, true, true>.
usage: python scripts/sass_counts.py [lib] > profiles/sass_counts.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(
    ROOT, "mast3r_slam_quality_dualtsdf_b200", "libb200tsdf.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)[1:]
KEEP = ("UBLKCP", "SYNCS", "LDGSTS", "RED", "ATOM", "REDUX", "MUFU", "BAR", "STL", "LDL", "UTMA",
        "MATCH", "VOTE", "SHFL", "LDGDEPBAR", "DEPBAR", "UCGABAR", "MEMBAR")
INS = re.compile(r"/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Za-z0-9_.]+)")


def demangle(n):
    return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()


print("# cuobjdump -sass of", os.path.relpath(lib, ROOT), "(sm_100a); written by scripts/sass_counts.py")
excerpt = None
for f in funcs:
    name = demangle(f.split("\n", 1)[0].strip())
    lines = [l for l in f.split("\n") if re.search(r"/\*[0-9a-f]{4,6}\*/", l)]
    ops = collections.Counter()
    for l in lines:
        m = INS.search(l)
        if m:
            ops[m.group(1)] += 1
    kept = {k: v for k, v in ops.items() if k.startswith(KEEP)}
    print(name[:160])
    print(f"  instructions {len(lines)}; " + ", ".join(f"{k} {v}" for k, v in sorted(kept.items())))
    if name.startswith("void integrate_keyframes_kernel<true, true, true>"):
        idx = [i for i, l in enumerate(lines) if "F32x4" in l]
        if idx:
            lo, hi = max(0, idx[0] - 70), min(len(lines), idx[0] + 30)
            excerpt = (name, lines[lo:hi])
if excerpt:
    print("\n# excerpt: fast march loop of", excerpt[0][:80])
    print("# (one-FMA keys: FFMA + FADD magic + FADD back; LOP3 0xbe change test; fold = REDG F32x4 + REDG 64)")
    for l in excerpt[1]:
        print(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/\s*$", "", l.rstrip()))
