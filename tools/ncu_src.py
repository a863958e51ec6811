#!/usr/bin/env python3
"""Hot source lines of one kernel of an .ncu-rep (source page, CUDA lines aggregated over their SASS):
   tools/ncu_src.py rep.ncu-rep kernel_name [top_n]  -> samples, warp-instructions, avg active threads per source line."""
import csv, io, subprocess, sys
from collections import defaultdict
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda", "-k", kern], capture_output=True, text=True).stdout
cur = None
agg = defaultdict(lambda: [0, 0, 0, ""])
for r in csv.reader(io.StringIO(src)):
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0].isdigit():
        try: s = int(r[6]); ie = int(r[7]); te = int(r[8])
        except (ValueError, IndexError): continue
        a = agg[(cur, int(r[0]))]
        a[0] += s; a[1] += ie; a[2] += te; a[3] = r[1][:110]
tot_s = sum(a[0] for a in agg.values()); tot_i = sum(a[1] for a in agg.values())
print(f"kernel {kern}: {tot_s} samples, {tot_i} warp-instructions")
byfile = defaultdict(lambda: [0, 0])
for (f, l), a in agg.items(): byfile[f][0] += a[0]; byfile[f][1] += a[1]
for f, (s, i) in sorted(byfile.items(), key=lambda kv: -kv[1][1]): print(f"  {f:28s} samples {100*s/max(tot_s,1):5.1f}%  instr {100*i/max(tot_i,1):5.1f}%")
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{100*a[1]/tot_i:5.1f}% instr {100*a[0]/max(tot_s,1):5.1f}% smp  thr {a[2]/max(a[1],1):4.1f}  {f}:{l}  {a[3]}")
