#!/usr/bin/env python3
"""Per-kernel CUDA-event times of the device-resident C2 workload (no parity, no CPU arm): quick A/B of kernel variants.
   MOUSIKI_B200_LIB=<variant .so> python tools/kernel_times.py [streams] [ticks] [steps]"""
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import bench  # noqa: E402
from mousiki_b200 import BatchDecoder  # noqa: E402

streams = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000
ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 50
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda", 0)
arena, offsets = bench.make_corpus(streams, ticks, 1000)
n = streams * ticks
dec = BatchDecoder(streams, 48000, 2)
a = torch.from_numpy(arena).to(dev)
o = torch.from_numpy(offsets.view(np.int32)).to(dev)
pcm = torch.empty((n, 1920), dtype=torch.float32, device=dev)
smp = torch.zeros(n, dtype=torch.int32, device=dev)
rng = torch.zeros(n, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream()


def step():
    dec.decode_grid_device(a.data_ptr(), o.data_ptr(), streams, ticks, pcm.data_ptr(), 1920, smp.data_ptr(), rng.data_ptr(), stream=st.cuda_stream)


for _ in range(2):
    step()
torch.cuda.synchronize()
dec.profile(True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for _ in range(steps):
    step()
e1.record(st)
torch.cuda.synchronize()
prof = dec.profile_read()
ms = e0.elapsed_time(e1) / steps
chk = int(rng.to(torch.int64).bitwise_and(0xFFFFFFFF).sum().item()) & 0xFFFFFFFF
pchk = int(pcm.view(torch.int32).to(torch.int64).sum().item()) & 0xFFFFFFFFFFFF
print(json.dumps({"lib": os.environ.get("MOUSIKI_B200_LIB", "default"), "ms_per_step": round(ms, 3),
                  "kernels": {k: round(v[0] / steps, 3) for k, v in prof.items() if v[1]}, "range_sum": chk, "pcm_bits_sum": pchk}), flush=True)
