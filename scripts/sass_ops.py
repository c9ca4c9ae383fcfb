#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove a Blackwell-native kernel (B200_PROFILING.md):
UTC*MMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UBLKCP (cp.async.bulk), UTMALDG / UTMASTG (tensor-map TMA),
HMMA (legacy mma.sync), from `cuobjdump -sass haet_b200/libhaet_b200.so`.   python scripts/sass_ops.py > profiles/r2_sass_ops.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "haet_b200", "libhaet_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
pat = {"UTC*MMA": r"\bUTC[A-Z]*MMA\b", "LDTM": r"\bLDTM\b", "STTM": r"\bSTTM\b", "UBLKCP": r"\bUBLKCP\b",
       "UTMALDG/STG": r"\bUTMA(LDG|STG)\b", "HMMA": r"\bHMMA\b", "MUFU.EX2": r"\bMUFU\.EX2\b", "F2FP": r"\bF2FP\b",
       "FFMA2/FADD2": r"\bF(FMA|ADD)2\b",
       "SYNCS (mbarrier)": r"\bSYNCS\b", "total": r"^\s+/\*[0-9a-f]{4,}\*/"}
counts, cur = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = demangle(m.group(1))
        cur = re.sub(r"\(.*", "", cur).replace("void ", "").replace("haet::", "")
        counts[cur] = collections.Counter()
        continue
    if cur:
        for k, p in pat.items():
            if re.search(p, line):
                counts[cur][k] += 1
cols = list(pat)
print("# SASS evidence (round 2): `cuobjdump -sass haet_b200/libhaet_b200.so`, counted by scripts/sass_ops.py\n")
print("sm_100a only (`-gencode arch=compute_100a,code=sm_100a`).  tcgen05.mma -> `UTC*MMA` (UTCHMMA = kind::f16), "
      "tcgen05.ld / st -> `LDTM` / `STTM`, `cp.async.bulk` -> `UBLKCP`.  No kernel uses tensor-map TMA (`UTMALDG`): every "
      "operand tile is a contiguous pre-swizzled image, so a 1-D bulk copy moves it; none uses the legacy `HMMA` path.  `FFMA2` / `FADD2`: packed fp32 pairs (fma.rn.f32x2) in the softmax "
      "loops of the two passes.\n")
print("| kernel | " + " | ".join(cols) + " |")
print("|---|" + "---|" * len(cols))
for k, c in counts.items():
    print(f"| `{k}` | " + " | ".join(str(c[x]) for x in cols) + " |")
