#!/usr/bin/env python3
"""Summarise an .ncu-rep: key counters per kernel (raw page) and, with --src, the hottest source lines."""
import csv, io, re, subprocess, sys, json
from collections import Counter
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h = rows[0]
pat = re.compile(r'^(dram__bytes_(read|write)\.sum|gpu__time_duration.sum|sm__warps_active.avg.pct_of_peak_sustained_active|launch__registers_per_thread|smsp__issue_active.avg.pct_of_peak_sustained_active|smsp__inst_executed.sum|l1tex__t_sector_hit_rate.pct|lts__t_sector_hit_rate.pct|smsp__average_warps_issue_stalled.*_per_issue_active.ratio|launch__grid_size|launch__block_size|sass__inst_executed_local_loads|sass__inst_executed_local_stores|smsp__thread_inst_executed_per_inst_executed.ratio|launch__shared_mem_per_block_static|sm__throughput.avg.pct_of_peak_sustained_elapsed|gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed|launch__occupancy_limit_.*|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum|smsp__inst_executed_op_shared.*sum|sm__inst_executed_pipe_.*pct_of_peak_sustained_active)$')
out = []
for v in rows[2:]:
    d = {"kernel": v[h.index("Kernel Name")].split("(")[0]}
    for a, u, c in zip(h, rows[1], v):
        if pat.match(a):
            try: x = float(c.replace(",", ""))
            except ValueError: continue
            if ("stalled" in a and x < 0.2) or ("pipe_" in a and x < 2): continue
            d[a + (" [" + u + "]" if u else "")] = x
    out.append(d)
print(json.dumps(out, indent=1))
if "--src" in sys.argv:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
    cur = None; lines = []; tot = 0
    for r in csv.reader(io.StringIO(src)):
        if r and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
        if r and r[0].isdigit():
            try: s = int(r[4]); ie = int(r[7])
            except ValueError: continue
            lines.append((s, ie, cur, int(r[0]), r[1][:100])); tot += s
    lines.sort(reverse=True)
    print("total samples", tot, "total warp-inst", sum(l[1] for l in lines))
    for l in lines[:int(sys.argv[sys.argv.index("--src") + 1]) if len(sys.argv) > sys.argv.index("--src") + 1 else 40]: print(l)
