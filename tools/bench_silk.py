#!/usr/bin/env python3
"""Secondary measurement (not the driver's bench line): BASELINE config 3 — SILK-only WB 16 kHz mono 20 ms streams,
48 kHz mono int16/f32 out — on one GPU, with the CPU oracle timed beside it. Prints one JSON line.

Packets: `TOC 0x48 || random bytes` (the reference's fuzz construction) restricted to packets without a redundancy
frame; a pool of POOL such packets is drawn with replacement per (stream, tick), so every stream sees its own sequence.
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import helpers  # noqa: E402

FRAME = 960
POOL = 2048


KINDS = {
    # BASELINE config 3: SILK-only WB mono 20 ms @ ~16 kb/s
    "silk": {"toc": helpers.silk_toc(2, 1, False), "payload": (8, 40), "ch": 1, "streams": 100_000,
             "workload": "silk_only_wb_mono_20ms_1e5_streams"},
    # BASELINE config 4: hybrid FB stereo 20 ms @ ~32 kb/s (SILK WB mid/side + CELT bands 17..20)
    "hybrid": {"toc": ((15 << 3) | 4), "payload": (60, 100), "ch": 2, "streams": 50_000,
               "workload": "hybrid_fb_stereo_20ms_5e4_streams"},
}
KIND = KINDS["silk"]


def corpus(streams, ticks, seed):
    pa, po = helpers.silk_fuzz_grid(POOL, 1, KIND["payload"], lambda k, r: KIND["toc"], seed=seed, channels=KIND["ch"])
    lens = np.diff(po.astype(np.int64))
    rng = np.random.default_rng(seed + 1)
    pick = rng.integers(0, POOL, size=streams * ticks)
    offsets = np.zeros(streams * ticks + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum(lens[pick])
    arena = np.empty(int(offsets[-1]), dtype=np.uint8)
    maxlen = int(lens.max())
    padded = np.zeros((POOL, maxlen), dtype=np.uint8)
    for k in range(POOL):
        padded[k, :lens[k]] = pa[po[k]:po[k + 1]]
    col = np.arange(maxlen)[None, :]
    mask = col < lens[pick][:, None]
    arena[:] = padded[pick][mask]
    return arena, offsets


def main():
    import torch

    from mousiki_b200 import BatchDecoder

    ap = argparse.ArgumentParser()
    ap.add_argument("--kind", choices=sorted(KINDS), default="silk")
    ap.add_argument("--streams", type=int, default=0)
    ap.add_argument("--ticks", type=int, default=10)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    global KIND
    KIND = KINDS[args.kind]
    CH = KIND["ch"]
    streams, ticks = args.streams or KIND["streams"], args.ticks
    n = streams * ticks
    arena, offsets = corpus(streams, ticks, 5)
    dec = BatchDecoder(streams, 48000, CH, device=0, max_frames_per_pass=max(1 << 18, streams * ticks))  # one pass per step
    # parity gate: first 64 streams
    cs = 64
    sel = (np.arange(ticks)[:, None] * streams + np.arange(cs)[None, :]).reshape(-1)
    lens = np.diff(offsets.astype(np.int64))
    so = np.zeros(len(sel) + 1, dtype=np.uint32)
    so[1:] = np.cumsum(lens[sel])
    sa = np.concatenate([arena[offsets[p]:offsets[p + 1]] for p in sel])
    ids = np.tile(np.arange(cs, dtype=np.uint32), ticks)
    pcm, samples, rng = dec.decode(sa, so, ids, FRAME)
    rp, rs, rr = helpers.oracle_decode_grid(sa, so, cs, ticks, CH, FRAME, i16=True)
    parity = {"packets_checked": int(len(sel)), "final_range_mismatches": int((rng != rr).sum()),
              "sample_count_mismatches": int((samples != rs).sum()), "pcm_i16_mismatches": int((pcm != rp).sum()),
              "pcm_i16_max_abs_diff": int(np.abs(pcm.astype(np.int32) - rp.astype(np.int32)).max())}
    dec.reset()
    d_arena = torch.from_numpy(arena).to(dev)
    d_offsets = torch.from_numpy(offsets.view(np.int32)).to(dev)
    d_pcm = torch.empty((n, FRAME * CH), dtype=torch.float32, device=dev)
    d_samples = torch.zeros(n, dtype=torch.int32, device=dev)
    d_range = torch.zeros(n, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream()

    def step():
        dec.decode_grid_device(d_arena.data_ptr(), d_offsets.data_ptr(), streams, ticks, d_pcm.data_ptr(), FRAME * CH, d_samples.data_ptr(),
                               d_range.data_ptr(), stream=stream.cuda_stream)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    dec.profile(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    prof = dec.profile_read()
    assert int((d_samples != FRAME).sum().item()) == 0
    audio_s = n * FRAME / 48000.0
    # CPU oracle on a bounded sample, all cores
    cores = os.cpu_count() or 1
    cst = 256 * cores
    ca, co = corpus(cst, ticks, 9)
    lib = helpers.oracle_lib()
    cp = np.zeros((cst * ticks, FRAME * CH), dtype=np.int16)
    csm = np.zeros(cst * ticks, dtype=np.int32)
    crg = np.zeros(cst * ticks, dtype=np.uint32)
    tot, passes = 0.0, 0
    while tot < 8.0:
        t0 = time.perf_counter()
        lib.orc_decode_grid(48000, CH, helpers.QUIRKS_MOUSIKI, helpers.ptr(ca), helpers.ptr(co), cst, ticks, FRAME, None, helpers.ptr(cp),
                            helpers.ptr(csm), helpers.ptr(crg), cores)
        tot += time.perf_counter() - t0
        passes += 1
    cpu_value = passes * cst * ticks * FRAME / 48000.0 / tot
    kms = {k: v[0] / args.steps for k, v in prof.items() if v[1]}
    # algorithmic bytes per 20 ms WB mono frame: packet ~27 B + record (9 + 160 words) written and read + state 2 x 1.1 KB
    # read and written once per pass (amortised over `ticks`) + PCM out 960 x 4 B
    alg = 27 + 2 * 169 * 4 + 2 * 2 * 1100 / ticks + 960 * 4
    if args.kind == "hybrid":  # stereo SILK record + state, and the f32 stereo row is written by k_silk then re-read and re-written by k_deemph
        alg = 80 + 2 * (3 + 2 * 169) * 4 + 2 * 2 * 2200 / ticks + 960 * 2 * 4
    peak = 6535.4
    try:
        peak = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
    except Exception:
        pass
    silk_ms = kms.get("k_silk", 0.0)
    line = {"metric": "decoded_audio_seconds_per_second", "value": audio_s * args.steps / (ms * 1e-3), "unit": "audio-s/s", "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "dtype": "int32" if args.kind == "silk" else "int32+f32",
            "config": {"workload": KIND["workload"], "streams": streams, "frames_per_stream_per_step": ticks,
                       "mean_payload_bytes": float(lens.mean()), "output": "f32 PCM left in HBM"},
            "kernel_ms_per_step": kms,
            "roofline": {"bound": "hbm", "kernel": "k_silk", "alg_bytes_per_frame": alg, "achieved": alg * n / (silk_ms * 1e-3) / 1e9 if silk_ms else None,
                         "peak": peak, "unit": "GB/s", "frac": (alg * n / (silk_ms * 1e-3) / 1e9 / peak) if silk_ms else None},
            "cpu_baseline": {"value": cpu_value, "unit": "audio-s/s", "cores": cores, "kind": "port",
                             "sample": f"{passes} passes of {cst} streams x {ticks} frames ({tot:.1f} s)"},
            "parity": parity}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
