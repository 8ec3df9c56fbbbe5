#!/usr/bin/env python3
"""Summarise an .ncu-rep (one kernel launch) into the handful of numbers we track; optional
opcode histogram per work unit.  Usage: ncu_summary.py report.ncu-rep [units_for_normalisation]"""
import csv
import io
import re
import subprocess
import sys
from collections import Counter

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed.sum', 'sm__inst_executed.avg.per_cycle_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'launch__grid_size', 'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__cycles_elapsed.max',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'lts__t_bytes.sum', 'l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum',
        'smsp__cycles_active.avg', 'sm__cycles_active.avg']


def main():
    rep = sys.argv[1]
    units = float(sys.argv[2]) if len(sys.argv) > 2 else None
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, unit, vals = rows[0], rows[1], rows[2]
    for w in WANT:
        for i, h in enumerate(hdr):
            if h == w:
                print(f'{w:80s} {vals[i]:>18s} {unit[i]}')
    for i, h in enumerate(hdr):
        if 'warp_issue_stalled' in h and h.endswith('per_warp_active.pct'):
            try:
                v = float(vals[i])
            except ValueError:
                continue
            if v >= 3:
                print(f'  stall {h[len("smsp__warp_issue_stalled_"):-len("_per_warp_active.pct")]:30s} {v:6.1f} %')
    src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    start = None
    for i, r in enumerate(rows):
        if r and r[0] == 'Address':
            hdr, start = r, i + 1
            break
    if start is None:
        return
    ix = {h: i for i, h in enumerate(hdr)}
    c, st, tot = Counter(), Counter(), 0
    lines = []
    for r in rows[start:]:
        if len(r) < len(hdr):
            continue
        try:
            n = int(r[ix['Instructions Executed']])
        except ValueError:
            continue
        op = re.sub(r'^@!?U?P\d+\s+', '', r[ix['Source']].strip()).split()[0].split('.')[0]
        c[op] += n
        st[op] += int(r[ix['Warp Stall Sampling (All Samples)']] or 0)
        tot += n
        lines.append((n, r[ix['Source']].strip(), int(r[ix['Warp Stall Sampling (All Samples)']] or 0), r[ix['Avg. Threads Executed']]))
    norm = units or 1.0
    print(f'total warp instructions {tot}  per unit {tot / norm:.2f}')
    for op, n in c.most_common(25):
        print(f'  {op:10s} {n / norm:9.2f}   stall-samples {st[op]}')
    if '--listing' in sys.argv:
        for n, s, stl, thr in lines:
            if n / norm >= 0.3:
                print(f'{n / norm:6.2f} {stl:5d} {thr:>4s} {s[:110]}')


if __name__ == '__main__':
    main()
