#!/usr/bin/env python3
"""Does the decode path gain from running G independent lane groups side by side (one BatchDecoder + CUDA stream per group,
streams split evenly)? Device-resident C2 workload, CUDA events across all streams. Prints one JSON line per G."""
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import bench  # noqa: E402
from mousiki_b200 import BatchDecoder  # noqa: E402

streams, ticks, steps = 10_000, 50, 4
dev = torch.device("cuda", 0)
for G in (1, 2, 3, 4):
    per = streams // G
    groups = []
    for g in range(G):
        arena, offsets = bench.make_corpus(per, ticks, 1000 + g)
        n = per * ticks
        groups.append(dict(dec=BatchDecoder(per, 48000, 2), s=torch.cuda.Stream(), a=torch.from_numpy(arena).to(dev),
                           o=torch.from_numpy(offsets.view(np.int32)).to(dev), pcm=torch.empty((n, 1920), dtype=torch.float32, device=dev),
                           smp=torch.zeros(n, dtype=torch.int32, device=dev), rng=torch.zeros(n, dtype=torch.int32, device=dev)))

    def step():
        for x in groups:
            x["dec"].decode_grid_device(x["a"].data_ptr(), x["o"].data_ptr(), per, ticks, x["pcm"].data_ptr(), 1920, x["smp"].data_ptr(),
                                        x["rng"].data_ptr(), stream=x["s"].cuda_stream)

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for x in groups:
        x["s"].wait_event(e0)
    for _ in range(steps):
        step()
    for x in groups:
        torch.cuda.current_stream().wait_stream(x["s"])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print(json.dumps({"groups": G, "streams_per_group": per, "ms_per_step": ms, "audio_s_per_s": per * G * ticks * 0.02 / (ms * 1e-3)}), flush=True)
    del groups
    torch.cuda.empty_cache()
