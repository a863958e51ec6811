#!/usr/bin/env python3
"""Work balance of k_pvq on the C2 corpus: dumps every leaf's (N, K) through the host emulation and models the SIMT cost of
thread-per-leaf schedules (steps of a leaf ~ N + K, a warp pays the maximum over its lanes)."""
import ctypes as C
import sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers  # noqa: E402
import bench  # noqa: E402

streams, ticks = 64, 4
arena, offsets = bench.make_corpus(streams, ticks, 1000)
lib = helpers.emul_lib()
buf = np.zeros(3 * 200 * streams * ticks, dtype=np.uint32)
lib.emu_leaf_dump.argtypes = [C.c_void_p, C.c_uint32]
lib.emu_leaf_count.restype = C.c_uint32
lib.emu_leaf_dump(helpers.ptr(buf), len(buf))
helpers.emul_decode_grid(arena, offsets, streams, ticks, 2, 960)
n = lib.emu_leaf_count()
lib.emu_leaf_dump(None, 0)
d = buf[:n].reshape(-1, 3).astype(np.int64)
frames = np.unique(d[:, 0])
print("frames", len(frames), "leaves/frame", len(d) / len(frames), "mean N", d[:, 1].mean(), "mean K", d[:, 2].mean(), "max N", d[:, 1].max(), "max K", d[:, 2].max())
def cost(nk, early_exit):
    N, K = nk
    return (N + K) if not early_exit else (min(N, max(K, 1) * 6 + 2) + K)   # crude: with the early exit a sparse leaf stops soon after its last pulse
for name, ee, sort in (("now", False, False), ("early exit", True, False), ("sorted", False, True), ("early exit + sorted", True, True)):
    tot_warp = tot_thread = 0
    for f in frames:
        L = d[d[:, 0] == f][:, 1:]
        c = np.array([cost(x, ee) for x in L], dtype=np.float64)
        if sort:
            c = np.sort(c)[::-1]
        pad = (-len(c)) % 32
        cw = np.concatenate([c, np.zeros(pad)]).reshape(-1, 32)
        # 64 threads: leaves l = tid, tid + 64, ...: warp w of round r takes leaves [64 r + 32 w, +32)
        tot_warp += cw.max(axis=1).sum()
        tot_thread += c.sum()
    print(f"{name:22s} warp-steps/frame {tot_warp / len(frames):8.1f}  thread-steps/frame {tot_thread / len(frames):8.1f}  lane use {tot_thread / (32 * tot_warp):.2f}")
