#!/usr/bin/env python3
"""Executed warp-instructions per source line, in file order, of one profiled launch (needs
-lineinfo and --import-source on).  Usage: ncu_lines_all.py report.ncu-rep units [min_per_unit]
The first line prints the launch's true smsp__inst_executed.sum per unit (inlined functions are
listed under both the callee's and the caller's line, so the per-line sum exceeds it)."""
import csv
import io
import subprocess
import sys

rep, units = sys.argv[1], float(sys.argv[2])
floor = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h = rows[0]
for name in ('smsp__inst_executed.sum', 'gpu__time_duration.sum', 'launch__registers_per_thread',
             'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers'):
    if name in h:
        v = rows[2][h.index(name)]
        print(name, v, f'{float(v.replace(",", "")) / units:.1f}/unit' if 'inst' in name else '')
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass,cuda'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, fname, out = None, '', []
for r in rows:
    if r and r[0] == 'File Path':
        fname = r[1].split('/')[-1]
    elif r and r[0] == 'Line No':
        ie = r.index('Instructions Executed')
        hdr = 1
    elif hdr and r and r[0].isdigit():
        try:
            out.append((fname, int(r[0]), int(r[ie]), r[1].strip()))
        except ValueError:
            pass
for f, l, n, src in sorted(out):
    if n / units >= floor:
        print(f'{n / units:7.1f} {f}:{l}: {src[:110]}')
