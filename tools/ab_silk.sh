#!/bin/bash
# A/B of a kernel change on the SILK-heavy workloads: C3, C4 and C5 bench lines, value and k_silk ms per step.
tag=$1
python tools/bench_silk.py > gpurun_out/ab_${tag}_silk.json 2>gpurun_out/ab_${tag}.err
python tools/bench_silk.py --kind hybrid > gpurun_out/ab_${tag}_hybrid.json 2>>gpurun_out/ab_${tag}.err
python tools/bench_mixed.py > gpurun_out/ab_${tag}_mixed.json 2>>gpurun_out/ab_${tag}.err
python - "$tag" <<'P'
import json, sys
for k in ("silk", "hybrid", "mixed"):
    try:
        d = json.load(open(f"gpurun_out/ab_{sys.argv[1]}_{k}.json"))
        print(sys.argv[1], k, round(d["value"]), round(d["ms_per_step"], 2), "k_silk", round(d["kernel_ms_per_step"].get("k_silk", 0), 2), d["parity"])
    except Exception as e:
        print(sys.argv[1], k, "failed", e)
P
