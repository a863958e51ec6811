#!/usr/bin/env python3
"""bench.py — decoded audio-seconds per second for 48 kHz stereo CELT streams (BASELINE.json metric).

Workload (BASELINE.json configs[1]): CELT-only FB 48 kHz stereo, 20 ms frames at 64 kb/s (160-byte payloads),
10^4 streams x 50 consecutive frames (1 s of audio per stream) per GPU. One step = one pass of the whole decode
path over that batch = 10^4 audio-seconds per GPU. Packets are synthetic (`TOC || seeded random bytes`, the
reference's own fuzz construction, tests/test_opus_decode.rs:372-407): neither mousiki's encoder (Rust) nor
libopus exists on these boxes, see DESIGN.md "Corpus".

  value  device-resident: packets already in HBM, PCM left in HBM (bd_decode_grid_dev), CUDA events on the
         launch stream, max over ranks.
  e2e    through the reference-facing host call (BatchDecoder.decode / bd_decode_i16 = mousiki Decoder::decode):
         pinned host packets -> H2D -> kernels -> D2H i16 PCM + sample counts + final_range, wall time of the call.
  --impl reference: the CPU restatement of mousiki's float decoder (oracle/, the reference itself cannot be built
         here: no Rust toolchain), all host threads, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

STREAMS = 10_000
TICKS = 50
PAYLOAD = 160
FRAME = 960
CH = 2
AUDIO_S_PER_FRAME = FRAME / 48000.0
METRIC = "decoded_audio_seconds_per_second"
UNIT = "audio-s/s"
WORKLOAD = "celt_only_48k_stereo_20ms_64kbps_1e4_streams_x50_frames"


def make_corpus(streams, ticks, seed):
    """[ticks][streams] grid of TOC(0xFC: CELT FB 20 ms stereo code 0) + 160 random bytes."""
    rng = np.random.default_rng(seed)
    n = streams * ticks
    arena = rng.integers(0, 256, size=(n, PAYLOAD + 1), dtype=np.uint8)
    arena[:, 0] = 0xFC
    offsets = (np.arange(n + 1, dtype=np.uint64) * (PAYLOAD + 1)).astype(np.uint32)
    return arena.reshape(-1), offsets


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    FIELDS = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            parts = [p.strip() for p in r.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "samples": len(sm),
                "reasons": sorted(reasons)}


def cpu_reference_rate(streams, ticks, threads, min_seconds=0.0, seed=7):
    """The CPU restatement of mousiki's float decoder (oracle/, kind "port") on `streams` x `ticks` frames of the same
    workload, `threads` std::threads, i16 output like the reference's own codec_bench (src/bin/codec_bench.rs:679-702:
    fresh decoder per stream, timer around the decode loop only). The pass is repeated until `min_seconds` of CPU
    wall time have been spent. Returns (audio-s/s over all passes, seconds, passes)."""
    import helpers

    arena, offsets = make_corpus(streams, ticks, seed)
    lib = helpers.oracle_lib()
    n = streams * ticks
    pcm = np.zeros((n, FRAME * CH), dtype=np.int16)
    samples = np.zeros(n, dtype=np.int32)
    rng = np.zeros(n, dtype=np.uint32)
    total, passes = 0.0, 0
    while passes == 0 or total < min_seconds:
        t0 = time.perf_counter()
        lib.orc_decode_grid(48000, CH, helpers.QUIRKS_MOUSIKI, helpers.ptr(arena), helpers.ptr(offsets), streams, ticks, FRAME, None,
                            helpers.ptr(pcm), helpers.ptr(samples), helpers.ptr(rng), threads)
        total += time.perf_counter() - t0
        passes += 1
    assert (samples == FRAME).all()
    return passes * streams * ticks * AUDIO_S_PER_FRAME / total, total, passes


def run_reference(args):
    """--impl reference: rank 0 only; each step decodes a bounded sample (cores x 32 streams x 50 frames)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    streams = 128 * cores
    for _ in range(args.warmup):
        cpu_reference_rate(streams, TICKS, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_reference_rate(streams, TICKS, cores)
    dt = time.perf_counter() - t0
    value = args.steps * streams * TICKS * AUDIO_S_PER_FRAME / dt
    sample = f"{streams} streams x {TICKS} frames per step ({cores} threads, one decoder per stream, i16 out)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": WORKLOAD, "note": "CPU restatement of mousiki's float decoder (oracle/); the Rust "
                                        "reference cannot be built on this box (no cargo/rustc)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)
    return 0


# Algorithmic HBM bytes per 20 ms stereo frame, per kernel (DESIGN.md "Roofline"): what each kernel must read and
# write once if nothing were re-fetched. The whole-path figure is SURVEY.md §8(d)'s 30.2 KB per stream-frame.
ALG_BYTES = {
    "k_parse": 161 + 8 + 1800,          # packet + offsets in, ~450 record words out
    "k_energy": 42 * 4 + 126 * 4 + 16,  # energy symbols in, band energies + frame meta out
    "k_pvq": 600 + 7680,                # leaf records in, spectrum row (pulse leaves) out
    "k_plan": 1200 + 1200,              # node stream + leaf masks in, operation list out
    "k_bands": 1200 + 504 + 7680 + 7680,  # operation list + energies + spectrum row in, freq row out (in place)
    "k_imdct": 7680 + 7680,             # freq row in, 2 x 960 f32 time-domain samples out
    "k_comb": 7680 + 3840,              # time-domain samples in, post-filtered samples out (about half the frames filter)
    "k_deemph": 7680 + 7680,            # post-filtered samples in, interleaved f32 PCM out
}
PATH_BYTES = 30200


def hbm_peak():
    try:
        return float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def verify_against_oracle(dec, arena, offsets, streams, ticks, check_streams=64):
    """Parity gate printed with the number: first `check_streams` streams, all ticks, against the oracle."""
    import helpers

    sel = (np.arange(ticks)[:, None] * streams + np.arange(check_streams)[None, :]).reshape(-1)
    lens = np.diff(offsets.astype(np.int64))
    sub_off = np.zeros(len(sel) + 1, dtype=np.uint32)
    sub_off[1:] = np.cumsum(lens[sel])
    sub_arena = np.concatenate([arena[offsets[p]:offsets[p + 1]] for p in sel])
    ids = np.tile(np.arange(check_streams, dtype=np.uint32), ticks)
    dec.reset()
    pcm, samples, rng = dec.decode_float(sub_arena, sub_off, ids, FRAME)
    ref_pcm, ref_samples, ref_rng = helpers.oracle_decode_grid(sub_arena, sub_off, check_streams, ticks, CH, FRAME)
    dec.reset()
    i16a = np.clip(np.rint(pcm * 32768.0), -32768, 32767)
    i16b = np.clip(np.rint(ref_pcm * 32768.0), -32768, 32767)
    num = float((ref_pcm.astype(np.float64) ** 2).sum())
    den = float(((ref_pcm.astype(np.float64) - pcm) ** 2).sum())
    return {
        "packets_checked": int(len(sel)),
        "final_range_mismatches": int((rng != ref_rng).sum()),
        "sample_count_mismatches": int((samples != ref_samples).sum()),
        "pcm_max_abs_err_i16_lsb": float(np.abs(i16a - i16b).max()),
        "pcm_max_abs_err_f32": float(np.abs(pcm - ref_pcm).max()),
        "snr_db": None if den == 0 else 10 * float(np.log10(max(num, 1e-30) / den)),
    }


def run_gpu(args):
    import torch
    import torch.distributed as dist

    from mousiki_b200 import BatchDecoder

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — mousiki_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    streams, ticks = args.streams, args.ticks
    n = streams * ticks
    arena, offsets = make_corpus(streams, ticks, 1000 + rank)  # every rank owns its own streams: no data-path collective
    dec = BatchDecoder(streams, 48000, CH, device=local_rank)

    parity = verify_against_oracle(dec, arena, offsets, streams, ticks) if rank == 0 else None

    # ---- device-resident arm ("value")
    d_arena = torch.from_numpy(arena).to(dev)
    d_offsets = torch.from_numpy(offsets.view(np.int32)).to(dev)
    d_pcm = torch.empty((n, FRAME * CH), dtype=torch.float32, device=dev)
    d_samples = torch.zeros(n, dtype=torch.int32, device=dev)
    d_range = torch.zeros(n, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream()

    def step_dev():
        dec.decode_grid_device(d_arena.data_ptr(), d_offsets.data_ptr(), streams, ticks, d_pcm.data_ptr(), FRAME * CH, d_samples.data_ptr(),
                               d_range.data_ptr(), stream=stream.cuda_stream)

    for _ in range(args.warmup):
        step_dev()
    torch.cuda.synchronize()
    launches_per_step = dec.launch_count()
    dec.profile(True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        step_dev()
    ev1.record(stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    prof = dec.profile_read()
    dec.profile(False)
    assert int((d_samples != FRAME).sum().item()) == 0, "a packet failed to decode in the timed region"
    checksum = int(d_range.to(torch.int64).bitwise_and(0xFFFFFFFF).sum().item()) & 0xFFFFFFFF

    # ---- end-to-end arm: the host call a user makes, pinned host buffers, copies inside the timed region
    h_arena = torch.from_numpy(arena).pin_memory().numpy()
    h_offsets = torch.from_numpy(offsets.view(np.int32)).pin_memory().numpy().view(np.uint32)
    h_ids = torch.from_numpy(np.tile(np.arange(streams, dtype=np.int32), ticks)).pin_memory().numpy().view(np.uint32)
    h_pcm = torch.empty((n, FRAME * CH), dtype=torch.int16).pin_memory().numpy()
    h_samples = torch.empty(n, dtype=torch.int32).pin_memory().numpy()
    h_range = torch.empty(n, dtype=torch.int32).pin_memory().numpy().view(np.uint32)
    del d_pcm
    torch.cuda.empty_cache()
    e2e_steps = max(1, min(args.steps, 3))
    for _ in range(2):
        dec.decode_into(h_arena, h_offsets, h_ids, h_pcm, h_samples, h_range)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        dec.decode_into(h_arena, h_offsets, h_ids, h_pcm, h_samples, h_range)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert (h_samples == FRAME).all()
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    sums = torch.tensor([checksum], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)  # max over ranks
        gathered = [torch.zeros_like(sums) for _ in range(world)]
        dist.all_gather(gathered, sums)           # NCCL carries only the per-rank checksums
        sums = torch.stack(gathered).reshape(-1)
    ms, e2e_ms = float(t[0].item()), float(t[1].item())
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    audio_s_per_step = world * n * AUDIO_S_PER_FRAME
    value = audio_s_per_step * args.steps / (ms * 1e-3)
    e2e_value = audio_s_per_step * e2e_steps / (e2e_ms * 1e-3)
    peak, peak_kind = hbm_peak()
    top = max((k for k in ALG_BYTES if prof[k][1] > 0), key=lambda k: prof[k][0])
    top_ms, top_launches = prof[top]
    frames_per_launch = n * args.steps / top_launches
    achieved = ALG_BYTES[top] * frames_per_launch / (top_ms / top_launches * 1e-3) / 1e9
    traffic = None
    tf = ROOT / "profiles" / "roofline_traffic.json"
    if tf.exists():
        traffic = json.loads(tf.read_text()).get(top)
    cores = os.cpu_count() or 1
    cpu_streams = 128 * cores
    cpu_value, cpu_s, cpu_passes = cpu_reference_rate(cpu_streams, TICKS, cores, min_seconds=12.0)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "streams_per_gpu": streams, "frames_per_stream_per_step": ticks, "payload_bytes": PAYLOAD,
                   "packets": "TOC 0xFC + seeded random bytes (reference fuzz construction)", "output": "f32 PCM left in HBM",
                   "l2": "per-step working set (scratch + PCM, several GB) exceeds the 126 MB L2", "sharding": f"streams by rank, {world} rank(s), no data-path collective"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h_arena.nbytes + h_offsets.nbytes + h_ids.nbytes + n),
                "d2h_bytes_per_step": int(h_pcm.nbytes + h_samples.nbytes + h_range.nbytes), "ms_per_step": e2e_ms / e2e_steps,
                "steps": e2e_steps, "api": "BatchDecoder.decode_into -> bd_decode_i16 (mousiki Decoder::decode), pinned host buffers"},
        "gpu_launches": launches_per_step * args.steps,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "peak_kind": peak_kind, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "alg_bytes_per_frame": ALG_BYTES[top],
                     "avg_launch_ms": top_ms / top_launches,
                     "path": {"alg_bytes_per_frame": PATH_BYTES, "achieved": PATH_BYTES * n * args.steps / (ms * 1e-3) / 1e9 / world,
                              "frac": PATH_BYTES * n * args.steps / (ms * 1e-3) / 1e9 / world / peak}},
        "kernel_ms_per_step": {k: v[0] / args.steps for k, v in prof.items() if v[1]},
        "cpu_baseline": {"value": cpu_value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{cpu_passes} passes of {cpu_streams} streams x {TICKS} frames ({cpu_s:.1f} s of CPU wall time), {cores} threads, "
                                   "fresh decoder per stream, i16 out"},
        "parity": parity,
        "checksums": [int(x) for x in sums.tolist()],
    }
    _emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


_RESULT_FD = None


def _claim_stdout():
    """Keep stdout for the one JSON line: anything a library writes to file descriptor 1 meanwhile (NCCL prints its version
    there when the first communicator comes up) goes to stderr instead."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def _emit(line):
    sys.stdout.flush()
    os.write(_RESULT_FD if _RESULT_FD is not None else 1, (json.dumps(line) + "\n").encode())


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--streams", type=int, default=STREAMS, help="streams per GPU (default: the BASELINE workload)")
    ap.add_argument("--ticks", type=int, default=TICKS, help="frames per stream per step")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
