#!/usr/bin/env python3
"""Stall reasons and issue figures of one profiled launch.  Usage: ncu_stalls.py report.ncu-rep"""
import csv
import io
import subprocess
import sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, v = rows[0], rows[2]
for i, name in enumerate(h):
    if "issue_stalled" in name and name.endswith(".pct") or "average_warp" in name or "warps_eligible" in name or \
            "issue_active" in name or "hit_rate" in name or "inst_executed_pipe" in name and name.endswith("sum"):
        try:
            x = float(v[i].replace(",", ""))
        except ValueError:
            continue
        if x >= 1:
            print(f"{name:95s} {x:14.2f}")
