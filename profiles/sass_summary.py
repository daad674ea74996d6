"""SASS evidence of the Blackwell-native paths: per kernel of librnnt_b200.so, the count of tensor-core / TMEM / TMA
mnemonics (UTCIMMA = tcgen05.mma kind::i8, UTCHMMA = kind::f16, HMMA = mma.sync, LDTM = tcgen05.ld, UTMALDG / UTMASTG =
TMA load / store, LDSM = ldmatrix).  Usage: python profiles/sass_summary.py > profiles/r02/sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "rnnt_inference_b200", "lib", "librnnt_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
pat = re.compile(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)")
kern, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[kern] = collections.Counter()
        continue
    m = pat.match(line)
    if m and kern:
        counts[kern][m.group(1)] += 1
keys = ["UTCIMMA", "UTCHMMA", "HMMA", "LDTM", "UTMALDG", "UTMASTG", "UTCBAR", "LDSM", "SYNCS", "FFMA", "total"]
print(f"{'kernel':70s} " + " ".join(f"{k:>8s}" for k in keys))
for k, c in counts.items():
    c["total"] = sum(v for kk, v in c.items() if kk != "total")
    print(f"{k[:70]:70s} " + " ".join(f"{c.get(x, 0):8d}" for x in keys))
