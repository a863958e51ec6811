#!/bin/bash
# Sweep of the pass size (frames per pass) for the headline workload; prints value / ms per step / e2e per setting.
for m in "$@"; do
  MOUSIKI_B200_MAX_FRAMES_PER_PASS=$m python bench.py --steps 5 --warmup 3 2>/dev/null > gpurun_out/sweep_$m.json
  python - "$m" <<'P'
import json, sys
d = json.load(open(f"gpurun_out/sweep_{sys.argv[1]}.json"))
print("pass", sys.argv[1], round(d["value"]), round(d["ms_per_step"], 2), round(d["e2e"]["value"]), {k: round(v, 2) for k, v in d["kernel_ms_per_step"].items()})
P
done
