#!/usr/bin/env python3
"""Per-source-line executed-instruction counts of one profiled launch (needs -lineinfo and
--import-source on).  Usage: ncu_lines.py report.ncu-rep units [top]"""
import csv
import io
import subprocess
import sys

rep, units = sys.argv[1], float(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass,cuda'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
out, hdr, fname = [], None, ''
for r in rows:
    if r and r[0] == 'File Path':
        fname = r[1].split('/')[-1]
    elif r and r[0] == 'Line No':
        hdr = {h: i for i, h in enumerate(r)}
        ie = r.index('Instructions Executed')
        ist = r.index('Warp Stall Sampling (All Samples)')
    elif hdr and r and r[0].isdigit():
        try:
            out.append((int(r[ie]), int(r[ist] or 0), fname, r[0], r[1].strip()))
        except ValueError:
            pass
tot = sum(o[0] for o in out)
print(f'total per unit {tot / units:.1f}')
for n, st, f, l, src in sorted(out, key=lambda o: -o[0])[:top]:
    print(f'{n / units:8.1f} {st:8d}  {f}:{l}: {src[:105]}')
