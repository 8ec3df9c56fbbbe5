#!/usr/bin/env python3
"""Turn the .ncu-rep files a round brought back in gpurun_out/ into the tracked text summaries under
profiles/ (summary + opcode histogram, the raw metric table, per-source-line instruction counts) and
refresh profiles/traffic.json.  Usage: make_profiles.py rNN"""
import csv
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
GO = os.path.join(ROOT, "gpurun_out")
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"

# report file, profile name, work units of the profiled launch (for per-unit instruction counts), what a unit is
# the captures of tools/profile_round.sh
REPORTS = [
    ("r02_viterbi.ncu-rep", "viterbi_stream", 65536 * 4000 / 128, "128-row tile (65,536 reads x 4000 rows, qstring)"),
    ("r02_beam32.ncu-rep", "beam_wide_b32", 3552 * 2000, "posterior row (3,552 reads x 2000 rows, beam 32)"),
    ("r02_beam5.ncu-rep", "beam_small_b5", 18944 * 2000, "posterior row (18,944 reads x 2000 rows, beam 5, threshold 0.1)"),
    ("r02_duplex.ncu-rep", "duplex", 9472 * 500, "network-1 row (9,472 pairs x 500 rows, band +-32, beam 5, exact exp; four pairs per warp)"),
]
# profile name -> (bench.py workload, posterior rows of the profiled launch)
WORKLOAD_OF = {"viterbi_stream": ("viterbi", 65536 * 4000), "beam_wide_b32": ("beam32", 3552 * 2000),
               "beam_small_b5": ("beam5", 18944 * 2000), "duplex": ("duplex", 9472 * 500)}


def run(cmd):
    return subprocess.run(cmd, capture_output=True, text=True).stdout


for rep, name, units, what in REPORTS:
    path = os.path.join(GO, rep)
    if not os.path.exists(path):
        print("missing", rep)
        continue
    head = f"# {rep}: ncu --set full --clock-control none --import-source on, one launch; unit = {what}\n"
    with open(os.path.join(OUT, f"{tag}_{name}_ncu_full_summary.txt"), "w") as f:
        f.write(head + run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), path, str(units)]))
    with open(os.path.join(OUT, f"{tag}_{name}_ncu_source_lines.txt"), "w") as f:
        f.write(head + "# executed warp instructions per unit and source line (inlined callees are listed under both lines)\n" +
                run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines_all.py"), path, str(units), "1.0"]))
    rows = list(csv.reader(io.StringIO(run(["ncu", "-i", path, "--page", "raw", "--csv"]))))
    with open(os.path.join(OUT, f"{tag}_{name}_ncu_raw_metrics.csv"), "w") as f:
        for h, u, v in sorted(zip(rows[0], rows[1], rows[2])):
            if "__" in h and v not in ("", "n/a"):
                f.write(f"{h},{u},{v}\n")
    h = rows[0]
    val = lambda k: float(rows[2][h.index(k)].replace(",", "")) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(rows[1][h.index(k)], 1)
    traffic = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
    tj = os.path.join(OUT, "traffic.json")
    d = json.load(open(tj))
    wl, wrows = WORKLOAD_OF[name]
    e = d.setdefault(wl, {})
    e["issue_slots_busy"] = val("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100
    e["warp_instructions_per_row"] = val("smsp__inst_executed.sum") / wrows
    if name != "viterbi_stream":
        e["dram_bytes_per_row"] = traffic / wrows
        e["source"] = f"profiles/{tag}_{name}_ncu_raw_metrics.csv: one launch, {what} (ncu --set full)"
    json.dump(d, open(tj, "w"), indent=1)
    if name == "viterbi_stream":
        d["viterbi"]["dram_bytes_per_read"] = traffic / 65536
        d["viterbi"]["source"] = (f"profiles/{tag}_viterbi_stream_ncu_raw_metrics.csv: dram__bytes_read.sum + dram__bytes_write.sum of one "
                                  "launch over 65,536 reads x 4000 x 5 (ncu --set full, bench.py --reads 65536); traffic is linear in "
                                  "reads, bench.py scales it to the launch's read count")
        json.dump(d, open(tj, "w"), indent=1)
        print("traffic per read", traffic / 65536)

launches = os.path.join(GO, f"launches_{tag}.csv")
if os.path.exists(launches):
    shutil.copy(launches, os.path.join(OUT, f"{tag}_bench_launches.csv"))
