#!/usr/bin/env python3
"""Secondary measurement (not the driver's bench line): BASELINE config 5 — a mixed-shape batch on one GPU, with the CPU
oracle timed beside it. Prints one JSON line.

Even streams are CELT-only FB stereo and draw a new frame size (2.5 / 5 / 10 / 20 ms) for every packet, so transient
and non-transient frames of all four MDCT sizes sit side by side in one launch; odd streams are SILK-only WB and
alternate freely between 40 ms and 60 ms packets, mono or stereo coded (two or three SILK frames per packet per
channel). Packets are `TOC || random bytes` (the reference's fuzz construction, tests/test_opus_decode.rs:372-407;
SURVEY.md section 8d corpus P-fuzz, lengths U[20,400] B for CELT); the SILK pool keeps packets without a redundancy
frame (40 / 60 ms packets with one are not built yet). A stream keeps its mode: a switch without
redundancy conceals a frame of the old mode, which is row F2.
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

FRAME = 2880  # 60 ms at 48 kHz: the row stride every packet gets
CH = 2
POOL = 2048


def pools(seed):
    ca, co = helpers.fuzz_grid(POOL, 1, (20, 400), lambda t, j, r: helpers.celt_toc(int(r.integers(0, 4)), True), seed=seed)
    sa, so = helpers.silk_fuzz_grid(POOL, 1, (16, 90), lambda k, r: helpers.silk_toc(2, 2 + int(r.integers(0, 2)), bool(r.integers(0, 2))),
                                    seed=seed + 1, channels=CH)
    return (ca, co), (sa, so)


def _padded(arena, offsets):
    lens = np.diff(offsets.astype(np.int64))
    pad = np.zeros((len(lens), int(lens.max())), dtype=np.uint8)
    for k in range(len(lens)):
        pad[k, :lens[k]] = arena[offsets[k]:offsets[k + 1]]
    return pad, lens


def corpus(streams, ticks, pool, seed):
    """Tick-major grid: packet (t, j) comes from the CELT pool when j is even, from the SILK pool when j is odd."""
    (cp, cl), (sp, sl) = (_padded(*pool[0]), _padded(*pool[1]))
    width = max(cp.shape[1], sp.shape[1])
    both = np.zeros((2 * POOL, width), dtype=np.uint8)
    both[:POOL, :cp.shape[1]] = cp
    both[POOL:, :sp.shape[1]] = sp
    lens = np.concatenate([cl, sl])
    rng = np.random.default_rng(seed)
    pick = rng.integers(0, POOL, size=streams * ticks) + POOL * (np.arange(streams * ticks) % streams % 2)
    offsets = np.zeros(streams * ticks + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum(lens[pick])
    arena = np.empty(int(offsets[-1]), dtype=np.uint8)
    step = 1 << 16
    for a in range(0, len(pick), step):  # bounded temporaries
        p = pick[a:a + step]
        arena[offsets[a]:offsets[min(a + step, len(pick))]] = both[p][np.arange(width)[None, :] < lens[p][:, None]]
    return arena, offsets


def main():
    import torch

    from mousiki_b200 import BatchDecoder

    ap = argparse.ArgumentParser()
    ap.add_argument("--streams", type=int, default=100_000, help="streams of the whole job; a torchrun launch shards them by rank")
    ap.add_argument("--ticks", type=int, default=4)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    sys.stdout.flush()
    result_fd = os.dup(1)  # stdout carries the JSON line only (NCCL prints its version banner on descriptor 1)
    os.dup2(2, 1)
    import torch.distributed as dist
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(local)
    if world > 1:  # streams are independent: rank r owns streams r, r + world, ... and there is no data-path collective
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    streams, ticks = len(range(rank, args.streams, world)), args.ticks
    streams -= streams % 2  # CELT / SILK pairs
    n = streams * ticks
    pool = pools(11)
    arena, offsets = corpus(streams, ticks, pool, 5 + 100 * rank)
    dec = BatchDecoder(streams, 48000, CH, device=local)
    # parity gate: the first 128 streams (64 CELT, 64 SILK), every tick, f32 and the sample counts
    cs = 128
    sel = (np.arange(ticks)[:, None] * streams + np.arange(cs)[None, :]).reshape(-1)
    lens = np.diff(offsets.astype(np.int64))
    so = np.zeros(len(sel) + 1, dtype=np.uint32)
    so[1:] = np.cumsum(lens[sel])
    sa = np.concatenate([arena[offsets[p]:offsets[p + 1]] for p in sel])
    ids = np.tile(np.arange(cs, dtype=np.uint32), ticks)
    pcm, samples, rng = dec.decode(sa, so, ids, FRAME)
    rp, rs, rr = helpers.oracle_decode_grid(sa, so, cs, ticks, CH, FRAME, i16=True)
    parity = {"packets_checked": int(len(sel)), "final_range_mismatches": int((rng != rr).sum()),
              "sample_count_mismatches": int((samples != rs).sum()), "rejected_packets": int((rs <= 0).sum()),
              "pcm_i16_mismatches": int((pcm != rp).sum()),
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
    if world > 1:
        dist.barrier()
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
    job = torch.tensor([ms, 0.0, float(streams)], dtype=torch.float64, device=dev)
    assert int((d_samples <= 0).sum().item()) == 0, "a packet of the corpus was rejected"
    per_ch = d_samples.to(torch.int64)
    audio_s = float(per_ch.sum().item()) / 48000.0
    celt_s = float(per_ch.view(ticks, streams)[:, 0::2].sum().item()) / 48000.0
    if world > 1:  # max over ranks of the device time, sum of the audio and of the streams
        job[1] = audio_s
        t = job[:1].clone()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(job, op=dist.ReduceOp.SUM)
        ms, audio_s, celt_s = float(t.item()), float(job[1].item()), celt_s * float(job[1].item()) / audio_s
        if rank:
            dist.destroy_process_group()
            return
    # CPU oracle on a bounded sample, all cores
    cores = os.cpu_count() or 1
    cst = 128 * cores
    ca, co = corpus(cst, ticks, pool, 9)
    lib = helpers.oracle_lib()
    cp = np.zeros((cst * ticks, FRAME * CH), dtype=np.int16)
    csm = np.zeros(cst * ticks, dtype=np.int32)
    crg = np.zeros(cst * ticks, dtype=np.uint32)
    tot, passes, cpu_audio = 0.0, 0, 0.0
    while tot < 8.0:
        t0 = time.perf_counter()
        lib.orc_decode_grid(48000, CH, helpers.QUIRKS_MOUSIKI, helpers.ptr(ca), helpers.ptr(co), cst, ticks, FRAME, None, helpers.ptr(cp),
                            helpers.ptr(csm), helpers.ptr(crg), cores)
        tot += time.perf_counter() - t0
        passes += 1
        cpu_audio += float(csm.clip(min=0).sum()) / 48000.0
    kms = {k: v[0] / args.steps for k, v in prof.items() if v[1]}
    line = {"metric": "decoded_audio_seconds_per_second", "value": audio_s * args.steps / (ms * 1e-3), "unit": "audio-s/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "dtype": "int32+f32",
            "config": {"workload": "mixed_celt_2.5_5_10_20ms_fb_stereo_and_silk_wb_40_60ms", "streams": int(job[2].item()) if world > 1 else streams,
                       "streams_per_gpu": streams, "sharding": f"streams by rank, {world} rank(s), no data-path collective",
                       "packets_per_stream_per_step": ticks, "audio_s_per_step": audio_s, "celt_share_of_audio": celt_s / audio_s,
                       "mean_payload_bytes": float(lens.mean()), "output": "f32 PCM left in HBM, row stride 2880 x 2"},
            "kernel_ms_per_step": kms,
            "cpu_baseline": {"value": cpu_audio / tot, "unit": "audio-s/s", "cores": cores, "kind": "port",
                             "sample": f"{passes} passes of {cst} streams x {ticks} packets ({tot:.1f} s)"},
            "parity": parity}
    os.write(result_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
