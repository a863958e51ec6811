#!/usr/bin/env python3
"""Per-kernel device time of the end-to-end host call (bd_decode_i16, pinned host buffers) on bench.py's workload."""
import json, sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
import torch  # noqa: E402
from mousiki_b200 import BatchDecoder  # noqa: E402

streams, ticks = 10000, 50
n = streams * ticks
arena, offsets = bench.make_corpus(streams, ticks, 1000)
dec = BatchDecoder(streams, 48000, bench.CH, device=0)
h_arena = torch.from_numpy(arena).pin_memory().numpy()
h_offsets = torch.from_numpy(offsets.view(np.int32)).pin_memory().numpy().view(np.uint32)
h_ids = torch.from_numpy(np.tile(np.arange(streams, dtype=np.int32), ticks)).pin_memory().numpy().view(np.uint32)
h_pcm = torch.empty((n, bench.FRAME * bench.CH), dtype=torch.int16).pin_memory().numpy()
h_samples = torch.empty(n, dtype=torch.int32).pin_memory().numpy()
h_range = torch.empty(n, dtype=torch.int32).pin_memory().numpy().view(np.uint32)
for _ in range(2):
    dec.decode_into(h_arena, h_offsets, h_ids, h_pcm, h_samples, h_range)
for prof in (False, True):
    dec.profile(prof)
    t0 = time.perf_counter()
    for _ in range(3):
        dec.decode_into(h_arena, h_offsets, h_ids, h_pcm, h_samples, h_range)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 3 * 1e3
    out = {"profiling": prof, "ms_per_call": dt}
    if prof:
        out["kernel_ms_per_call"] = {k: round(v[0] / 3, 2) for k, v in dec.profile_read().items() if v[1]}
    print(json.dumps(out))
