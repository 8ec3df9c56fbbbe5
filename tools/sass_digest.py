#!/usr/bin/env python3
"""Per-kernel digest of the SASS in libfctc_b200.so: registers, static shared memory and the counts of the instructions
that show what the kernels are built on (bulk async copies UBLKCP, mbarrier SYNCS, warp reductions REDUX/CREDUX, MATCH,
votes, shuffles, 128-bit shared loads, FP64 and MUFU work, divisions' out-of-line calls).  Runs without a GPU.
Usage: sass_digest.py [libfctc_b200.so] > profiles/sass_digest_rNN.txt"""
import os
import re
import subprocess
import sys
from collections import Counter

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "fast-ctc-decode_b200", "libfctc_b200.so")
WATCH = ["UBLKCP", "SYNCS", "REDUX", "CREDUX", "MATCH", "VOTE", "SHFL", "LDS.128", "LDS.64", "LDG", "STG", "ATOM", "RED",
         "DFMA", "DMUL", "DADD", "F2F", "MUFU", "FCHK", "CALL", "BSSY", "BRA", "FFMA", "FADD", "FMUL"]

res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True).stdout
usage = {}
name = None
for line in res.splitlines():
    m = re.search(r"Function (\S+):", line)
    if m:
        name = m.group(1)
    m = re.search(r"REG:(\d+).*SHARED:(\d+)", line)
    if m and name:
        usage[name] = (int(m.group(1)), int(m.group(2)))

sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
counts, total = {}, Counter()
name = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1)
        counts[name] = Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and name:
        op = m.group(1)
        total[name] += 1
        for w in WATCH:
            if op == w or op.startswith(w + ".") or (("." in w) and op.startswith(w)):
                counts[name][w] += 1


def demangle(n):
    return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()


print(f"# SASS digest of {os.path.relpath(lib, ROOT)} (cuobjdump -sass / -res-usage, sm_100a); counts are static instructions")
for n in sorted(counts, key=lambda k: -total[k]):
    if total[n] < 40:
        continue
    reg, sh = usage.get(n, (0, 0))
    d = demangle(n)
    d = re.sub(r"\((?!anonymous namespace).*$", "", d)
    print(f"\n{d}\n  instructions {total[n]}, registers {reg}, static shared {sh} B")
    print("  " + ", ".join(f"{w} {counts[n][w]}" for w in WATCH if counts[n][w]))
