#!/usr/bin/env python
"""SASS evidence for the Blackwell-native path (B200_PROFILING.md "What proves a Blackwell-native kernel"):

    python tools/sass_histogram.py [frn_b200/libfrn_b200.so] > profiles/r2_sass_histogram.txt

Counts, per kernel of the shipped library, the mnemonics that only the tcgen05 / TMEM / TMA path produces
(UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG = cp.async.bulk.tensor, UBLKCP = cp.async.bulk,
UTCBAR = tcgen05.commit), the packed fp32 pair arithmetic (FADD2 / FMUL2 / FFMA2), and the legacy tensor path
(HMMA: must be absent)."""
import collections
import re
import subprocess
import sys

MNEMONICS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "UTCCP", "FADD2", "FMUL2", "FFMA2",
             "HMMA", "LDGSTS", "SYNCS", "RED.E"]
lib = sys.argv[1] if len(sys.argv) > 1 else "frn_b200/libfrn_b200.so"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
per, total, kern = collections.defaultdict(collections.Counter), collections.Counter(), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\(anonymous namespace\)::|frn::", "", kern).split("(")[0].replace("void ", "")
        continue
    for mn in MNEMONICS:
        if re.search(r"\b" + re.escape(mn) + r"\b", line) or (mn.endswith(".E") and mn in line):
            per[kern][mn] += 1
            total[mn] += 1
print(f"# {lib}: SASS mnemonic counts (sm_100a cubins)")
print("total: " + ", ".join(f"{k} {v}" for k, v in sorted(total.items(), key=lambda kv: -kv[1])))
print()
for k in sorted(per, key=lambda k: -sum(per[k].values())):
    if any(per[k][m] for m in ("UTCHMMA", "LDTM", "STTM", "UTMALDG", "UBLKCP", "FFMA2", "HMMA")):
        print(f"{k}: " + ", ".join(f"{m} {n}" for m, n in per[k].most_common()))
