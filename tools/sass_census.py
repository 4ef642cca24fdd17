#!/usr/bin/env python
"""SASS census of libeegspec.so: per kernel, the mnemonics that prove what the code runs on
(UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UBLKCP = cp.async.bulk, UTMALDG / UTMASTG = tensor-map TMA,
SYNCS = mbarrier, FFMA2 / FMUL2 / FADD2 = packed fp32, DFMA / DMUL / DADD = fp64, MUFU, HMMA = mma.sync).

    python tools/sass_census.py > profiles/sass_census_<tag>.txt     (build container, no GPU needed)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "multimodal-brain-pattern-identification_xai_b200", "lib", "libeegspec.so")
WATCH = ["UTCHMMA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "FFMA2", "FMUL2", "FADD2", "FFMA", "FMUL",
         "FADD", "FMNMX", "FMNMX3", "MUFU", "DFMA", "DMUL", "DADD", "HMMA", "IMAD", "LOP3", "LDS", "STS", "LDG", "STG",
         "LDL", "STL", "SHFL", "BAR", "ATOMS", "RED", "NANOSLEEP"]
txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
kern, counts = None, {}
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and kern:
        counts[kern][m.group(1)] += 1
        counts[kern]["_total"] += 1
print(f"# SASS census of {os.path.relpath(LIB, ROOT)} (static instruction counts per kernel)\n")
for k in sorted(counts):
    c = counts[k]
    print(f"{k}\n    total {c['_total']}: " + ", ".join(f"{w} {c[w]}" for w in WATCH if c[w]))
