#!/usr/bin/env python
"""Counts the SASS mnemonics that prove a Blackwell-native kernel (B200_PROFILING.md: UTC*MMA = tcgen05.mma, LDTM / STTM =
tcgen05.ld / st, UTMALDG / UTMASTG / UBLKCP / UBLKRED = TMA and bulk copies / reductions, HMMA = the legacy mma.sync path) per
kernel of libmstk.so.  No GPU needed: python tools/sass_mnemonics.py [path/to/libmstk.so]  (prints a markdown table)"""
import collections, os, re, subprocess, sys
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "music_style_transfer_ldm_b200", "libmstk.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
pat = re.compile(r"\b(UTC[A-Z]*MMA|UTMALDG|UTMASTG|UBLKCP|UBLKRED|UTCBAR|LDTM|STTM|HMMA|SYNCS)\b")
cur, counts = None, collections.OrderedDict()
for ln in sass.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
    elif cur:
        for t in pat.findall(ln):
            counts[cur][t] += 1
names = list(counts)
dem = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
cols = ["UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "UBLKCP", "UBLKRED", "UTCBAR", "SYNCS", "HMMA"]
print(f"{len(names)} kernels in {os.path.basename(so)}; those with tensor-core or TMA instructions:\n")
print("| kernel | " + " | ".join(cols) + " |\n|---|" + "---|" * len(cols))
for n, d in zip(names, dem):
    c = counts[n]
    if any(c[k] for k in cols if k != "SYNCS"):
        print("| `" + re.sub(r"\(.*", "", d)[:80] + "` | " + " | ".join(str(c[k]) if c[k] else "" for k in cols) + " |")
