#!/usr/bin/env python3
"""Golden per-stream final_range checksums of bench.py's strong-scaling job (ONE fixed corpus of 10^5 CELT stereo streams x
50 frames), from the ORACLE on the CPU: sha256 over the uint32 vector of per-stream sums of final_range. final_range of a
CELT-only packet is a function of the packet alone, so the vector does not depend on how often the corpus was decoded.
Writes tests/golden/strong_scaling_checksums.json (the pcm_bits hash, which depends on the kernels' float rounding and on
the number of repeats, is added by tools/record_strong_pcm_hash.py from a 1-GPU run)."""
import hashlib
import json
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import bench  # noqa: E402
import helpers  # noqa: E402

total = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 50
repeats = [int(x) for x in sys.argv[3:]] or [8]
sums = np.zeros(total, dtype=np.uint64)
step = 5000
for lo in range(0, total, step):
    hi = min(lo + step, total)
    arena, offsets = bench.make_strong_corpus(total, ticks, lo, hi)
    _, samples, rng = helpers.oracle_decode_grid(arena, offsets, hi - lo, ticks, 2, 960, i16=True)
    assert (samples == 960).all()
    sums[lo:hi] = rng.reshape(ticks, hi - lo).astype(np.uint64).sum(axis=0) & np.uint64(0xFFFFFFFF)
    print(f"streams {lo}..{hi}", file=sys.stderr, flush=True)
h = hashlib.sha256(sums.astype(np.uint32).tobytes()).hexdigest()
path = ROOT / "tests" / "golden" / "strong_scaling_checksums.json"
d = json.loads(path.read_text()) if path.exists() else {}
for r in repeats:
    d.setdefault(f"{total}x{ticks}x{r}", {})["final_range_sha256"] = h
path.write_text(json.dumps(d, indent=1, sort_keys=True) + "\n")
print(h)
