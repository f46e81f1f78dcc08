#!/usr/bin/env python
"""Opcode histogram per kernel of libasdmatch_cuda.so (cuobjdump -sass): the evidence that the hot kernels are what
DESIGN.md says they are -- UTCHMMA / UTMALDG / LDTM / UTCBAR / SYNCS for the tcgen05 + TMA + TMEM contraction, FADD /
FMUL without FFMA for the exact-distance kernels, ATOMS / MATCH / VOTE for the selection stages.
usage: python tools/sass_histogram.py [path to .so] > profiles/<round>_sass_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "asd-slam_b200", "libasdmatch_cuda.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
demangle = lambda s: subprocess.run(["c++filt", s], capture_output=True, text=True).stdout.strip()
KEY = ("UTCHMMA", "UTCBAR", "UTMALDG", "UTMAPF", "LDTM", "STTM", "SYNCS", "HMMA", "FFMA", "FADD", "FMUL", "DADD", "DFMA", "ATOMS", "ATOMG",
       "ATOM", "RED", "MATCH", "VOTE", "SHFL", "LDG", "STG", "LDS", "STS", "LDGSTS", "BAR", "MEMBAR", "NANOSLEEP", "MULTIMEM")
cur, hist = None, {}
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        hist[cur] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        hist[cur][m.group(1)] += 1
print(f"# {os.path.relpath(so, ROOT)}: {len(hist)} kernels, sm_100a SASS (cuobjdump -sass), opcode counts by mnemonic prefix")
for fn in sorted(hist, key=lambda f: -sum(hist[f].values())):
    h = hist[fn]
    name = demangle(fn)
    name = name.replace("(anonymous namespace)::", "")
    name = re.sub(r"\((?:[^()]|\([^()]*\))*\)\s*$", "", name)[:90]
    picks = [(k, sum(v for op, v in h.items() if op == k or op.startswith(k + "."))) for k in KEY]
    picks = [(k, v) for k, v in picks if v]
    print(f"{name:90s} total {sum(h.values()):6d}  " + " ".join(f"{k}={v}" for k, v in picks))
