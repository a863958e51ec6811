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
         here: no Rust toolchain), all host threads, on a bounded sample of the same workload; built on the box it runs
         on with -O3 -march=native -ffp-contract=off.
  --scaling strong: ONE fixed corpus of --streams streams in total (default 10^5, the north star's job) sharded over the
         ranks by stream id; the per-stream checksums are all-gathered with NCCL and must not depend on the rank count.
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


CORPUS = os.environ.get("MOUSIKI_BENCH_CORPUS", "fuzz")   # set by --corpus
CORPUS_NOTE = {
    "fuzz": "P-fuzz: TOC 0xFC + seeded random bytes (reference fuzz construction)",
    "struct": "P-struct: TOC 0xFC + 160 bytes written through the restated decoder (oracle/gen_corpus.cpp): band energies along a "
              "per-stream envelope, ~10 % transients, post-filter on ~30 % of the frames, PCM inside +-1 on 99.8 % of the packets",
}


def make_corpus(streams, ticks, seed):
    """[ticks][streams] grid of TOC(0xFC: CELT FB 20 ms stereo code 0) + 160 bytes: random (P-fuzz) or structured (P-struct)."""
    if CORPUS == "struct":
        import helpers
        return helpers.struct_grid(streams, ticks, PAYLOAD, 0xFC, CH, seed=seed)
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


_NATIVE = None


def oracle_native_lib():
    """The oracle built for THIS host's CPU (-O3 -march=native -ffp-contract=off), for the timed
    CPU arm only; the file name carries a signature of the CPU so that a library built elsewhere is never loaded.
    Falls back to the portable build (-O2) if the compiler is missing. Returns (lib, "native" | "portable")."""
    global _NATIVE
    if _NATIVE is not None:
        return _NATIVE
    import ctypes as C
    import hashlib

    import helpers
    try:
        flags = next(l for l in open("/proc/cpuinfo") if l.startswith("flags"))
        model = next(l for l in open("/proc/cpuinfo") if l.startswith("model name"))
        sig = hashlib.sha1((flags + model).encode()).hexdigest()[:10]
        target = f"_build/liboracle_native_{sig}.so"
        subprocess.run(["make", "-C", str(ROOT / "oracle"), target], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        lib = C.CDLL(str(ROOT / "oracle" / target))
        lib.orc_decode_grid.argtypes = helpers.oracle_lib().orc_decode_grid.argtypes
        _NATIVE = (lib, "native")
    except Exception:
        _NATIVE = (helpers.oracle_lib(), "portable")
    return _NATIVE


def cpu_reference_rate(streams, ticks, threads, min_seconds=0.0, seed=7):
    """The CPU restatement of mousiki's float decoder (oracle/, kind "port") on `streams` x `ticks` frames of the same
    workload, `threads` std::threads, i16 output like the reference's own codec_bench (src/bin/codec_bench.rs:679-702:
    fresh decoder per stream, timer around the decode loop only). The pass is repeated until `min_seconds` of CPU
    wall time have been spent. Returns (audio-s/s over all passes, seconds, passes)."""
    import helpers

    arena, offsets = make_corpus(streams, ticks, seed)
    lib, _ = oracle_native_lib()
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
    dt = 0.0
    for _ in range(args.steps):
        dt += cpu_reference_rate(streams, TICKS, cores)[1]     # the decode loop only (src/bin/codec_bench.rs:679-702), not the corpus set-up
    value = args.steps * streams * TICKS * AUDIO_S_PER_FRAME / dt
    sample = f"{streams} streams x {TICKS} frames per step ({cores} threads, one decoder per stream, i16 out)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": WORKLOAD, "packets": CORPUS_NOTE[CORPUS], "note": "CPU restatement of mousiki's float decoder (oracle/); the Rust "
                                        "reference cannot be built on this box (no cargo/rustc)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "build": oracle_native_lib()[1]},
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


def oracle_tail(arena, offsets, streams, ticks, check, repeats, i16):
    """What the reference path leaves for the first `check` streams of a [ticks][streams] grid after the grid has been
    decoded `repeats` times in a row by the same decoders (the bench never resets between steps): the outputs of the LAST
    repeat, from the oracle (one decoder per stream, all `ticks * repeats` packets in sequence)."""
    import helpers

    sel = (np.arange(ticks)[:, None] * streams + np.arange(check)[None, :]).reshape(-1)
    lens = np.diff(offsets.astype(np.int64))
    one = np.concatenate([arena[offsets[p]:offsets[p + 1]] for p in sel])
    sub_arena = np.tile(one, repeats)
    sub_off = np.zeros(len(sel) * repeats + 1, dtype=np.uint32)
    sub_off[1:] = np.cumsum(np.tile(lens[sel], repeats))
    pcm, samples, rng = helpers.oracle_decode_grid(sub_arena, sub_off, check, ticks * repeats, CH, FRAME, i16=i16)
    k = len(sel) * (repeats - 1)
    return sel, pcm[k:], samples[k:], rng[k:]


def parity_report(pcm, samples, rng, ref_pcm, ref_samples, ref_rng, what):
    """Gates printed with the number: final_range and sample counts exact, PCM in i16 LSB / f32 / SNR."""
    if pcm.dtype == np.int16:
        a, b = pcm.astype(np.int64), ref_pcm.astype(np.int64)
        fa, fb = pcm.astype(np.float64) / 32768.0, ref_pcm.astype(np.float64) / 32768.0
    else:
        a = np.clip(np.rint(pcm.astype(np.float64) * 32768.0), -32768, 32767)
        b = np.clip(np.rint(ref_pcm.astype(np.float64) * 32768.0), -32768, 32767)
        fa, fb = pcm.astype(np.float64), ref_pcm.astype(np.float64)
    num, den = float((fb ** 2).sum()), float(((fb - fa) ** 2).sum())
    return {
        "checked": what, "packets_checked": int(len(samples)),
        "final_range_mismatches": int((rng != ref_rng).sum()),
        "sample_count_mismatches": int((samples != ref_samples).sum()),
        "pcm_max_abs_err_i16_lsb": float(np.abs(a - b).max()),
        "pcm_max_abs_err_f32": float(np.abs(fa - fb).max()),
        "snr_db": None if den == 0 else 10 * float(np.log10(max(num, 1e-30) / den)),
    }


def make_strong_corpus(total_streams, ticks, lo, hi, seed=4242):
    """Streams [lo, hi) of ONE fixed corpus of `total_streams` streams (bytes of stream s do not depend on the sharding):
    generated stream-major in blocks of 1000 streams, returned as this rank's [ticks][hi - lo] grid."""
    blocks = []
    for b0 in range(lo - lo % 1000, hi, 1000):
        g = np.random.default_rng([seed, b0 // 1000]).integers(0, 256, size=(1000, ticks, PAYLOAD + 1), dtype=np.uint8)
        blocks.append(g[max(lo - b0, 0):max(min(hi - b0, 1000), 0)])
    g = np.concatenate(blocks) if blocks else np.zeros((0, ticks, PAYLOAD + 1), dtype=np.uint8)
    g[:, :, 0] = 0xFC
    arena = np.ascontiguousarray(g.transpose(1, 0, 2)).reshape(-1)
    offsets = (np.arange((hi - lo) * ticks + 1, dtype=np.uint64) * (PAYLOAD + 1)).astype(np.uint32)
    return arena, offsets


def d2h_probe(torch, dev, mbytes=256, reps=4):
    """Device -> pinned host copy rate of this rank while every rank copies at once (explains the e2e arm at N GPUs: the box's
    host fabric, not the decoder, bounds the PCM going back). GB/s."""
    n = mbytes << 20
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    h.copy_(d, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        h.copy_(d, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    return reps * n / (e0.elapsed_time(e1) * 1e-3) / 1e9


def run_gpu(args):
    import hashlib

    import torch
    import torch.distributed as dist

    from mousiki_b200 import BatchDecoder
    from mousiki_b200.sharding import gather_checksums, shard_range

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
    ticks = args.ticks
    strong = args.scaling == "strong"
    if strong:
        total_streams = args.streams if args.streams != STREAMS else 100_000
        lo, hi = shard_range(total_streams, world, rank)
        streams = hi - lo
        arena, offsets = make_strong_corpus(total_streams, ticks, lo, hi)
    else:
        streams = args.streams
        total_streams = streams * world
        arena, offsets = make_corpus(streams, ticks, 1000 + rank)  # every rank owns its own streams: no data-path collective
    n = streams * ticks
    dec = BatchDecoder(streams, 48000, CH, device=local_rank)
    check = min(64, streams)

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
    # per-stream checksums of the LAST timed step: sum of final_range over the stream's packets (low word) and a sum over the
    # PCM bit patterns (high word); both computed on the device, 8 bytes per stream leave it
    fr = d_range.view(ticks, streams).to(torch.int64).bitwise_and_(0xFFFFFFFF).sum(dim=0).bitwise_and_(0xFFFFFFFF)
    pc = torch.zeros(streams, dtype=torch.int64, device=dev)
    bits = d_pcm.view(torch.int32).view(ticks, streams, FRAME * CH)
    for t in range(ticks):
        pc += bits[t].sum(dim=1, dtype=torch.int64)
    local_sums = ((pc.bitwise_and_(0xFFFFFFFF) << 32) | fr).cpu().numpy().view(np.uint64)
    # parity gate on the TIMED configuration: the first streams of the last timed call, straight from its output buffers
    parity = None
    if rank == 0:
        sel, ref_pcm, ref_samples, ref_rng = oracle_tail(arena, offsets, streams, ticks, check, args.warmup + args.steps, False)
        idx = torch.from_numpy(sel.astype(np.int64)).to(dev)
        parity = parity_report(d_pcm[idx].cpu().numpy(), d_samples[idx].cpu().numpy(), d_range[idx].cpu().numpy().view(np.uint32),
                               ref_pcm, ref_samples, ref_rng,
                               f"first {check} streams x {ticks} frames of the last timed bd_decode_grid_dev call ({streams} lanes), f32")

    # ---- end-to-end arm: the host call a user makes, pinned host buffers, copies inside the timed region
    h_arena = torch.from_numpy(arena).pin_memory().numpy()
    h_offsets = torch.from_numpy(offsets.view(np.int32)).pin_memory().numpy().view(np.uint32)
    h_ids = torch.from_numpy(np.tile(np.arange(streams, dtype=np.int32), ticks)).pin_memory().numpy().view(np.uint32)
    h_pcm = torch.empty((1 if args.no_e2e else n, FRAME * CH), dtype=torch.int16).pin_memory().numpy()
    h_samples = torch.empty(n, dtype=torch.int32).pin_memory().numpy()
    h_range = torch.empty(n, dtype=torch.int32).pin_memory().numpy().view(np.uint32)
    del d_pcm, bits
    torch.cuda.empty_cache()
    e2e_steps = max(1, min(args.steps, 5))
    e2e_warm = max(3, min(args.warmup, 5))   # (W >= 3 warm-up calls here too: the first calls of a fresh buffer set run ~5 ms long)
    dec.reset()
    for _ in range(0 if args.no_e2e else e2e_warm):
        dec.decode_into(h_arena, h_offsets, h_ids, h_pcm, h_samples, h_range)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(0 if args.no_e2e else e2e_steps):
        dec.decode_into(h_arena, h_offsets, h_ids, h_pcm, h_samples, h_range)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert args.no_e2e or (h_samples == FRAME).all()
    clocks = sampler.stop() if rank == 0 else None
    e2e_parity = None
    if rank == 0 and not args.no_e2e:
        sel, ref_pcm, ref_samples, ref_rng = oracle_tail(arena, offsets, streams, ticks, check, e2e_warm + e2e_steps, True)
        e2e_parity = parity_report(h_pcm[sel], h_samples[sel], h_range[sel], ref_pcm, ref_samples, ref_rng,
                                   f"first {check} streams x {ticks} frames of the last timed bd_decode_i16 call ({streams} lanes), i16")
    if world > 1:
        dist.barrier()
    d2h_gbs = d2h_probe(torch, dev)

    t = torch.tensor([ms, e2e_s * 1e3, -d2h_gbs], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)  # max over ranks (slowest copy rate)
    if strong:
        sums = gather_checksums(local_sums, total_streams, world, rank, device=dev)   # NCCL carries only the per-stream checksums
    else:  # weak scaling: every rank has `streams` streams of its own
        buf = torch.from_numpy(local_sums.view(np.int64)).to(dev)
        if world > 1:
            out = [torch.zeros_like(buf) for _ in range(world)]
            dist.all_gather(out, buf)
            buf = torch.cat(out)
        sums = buf.cpu().numpy().view(np.uint64)
    ms, e2e_ms, d2h_min = float(t[0].item()), float(t[1].item()), -float(t[2].item())
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    audio_s_per_step = total_streams * ticks * AUDIO_S_PER_FRAME
    value = audio_s_per_step * args.steps / (ms * 1e-3)
    e2e_value = audio_s_per_step * e2e_steps / (e2e_ms * 1e-3)
    peak, peak_kind = hbm_peak()
    top = max((k for k in ALG_BYTES if prof[k][1] > 0), key=lambda k: prof[k][0])
    top_ms, top_launches = prof[top]
    frames_per_launch = n * args.steps / top_launches
    achieved = ALG_BYTES[top] * frames_per_launch / (top_ms / top_launches * 1e-3) / 1e9
    # the band pass has two forms; calls with >= 3552 lanes run every frame through the per-stream one (mb_api.cu)
    top_name = "k_bands_hyb" if top == "k_bands" and streams >= 148 * 6 * 4 else top
    traffic = None
    tf = ROOT / "profiles" / "roofline_traffic.json"
    if tf.exists():
        table = json.loads(tf.read_text())
        traffic = table.get(top_name)
        if traffic is not None and table.get("_per_frame"):   # per-frame figures: scaled to the frames one launch processes
            traffic = int(traffic * frames_per_launch)
    cpu_baseline = None
    if world == 1:
        cores = os.cpu_count() or 1
        cpu_streams = 128 * cores
        cpu_value, cpu_s, cpu_passes = cpu_reference_rate(cpu_streams, TICKS, cores, min_seconds=12.0)
        cpu_baseline = {"value": cpu_value, "unit": UNIT, "cores": cores, "kind": "port", "build": oracle_native_lib()[1],
                        "sample": f"{cpu_passes} passes of {cpu_streams} streams x {TICKS} frames ({cpu_s:.1f} s of CPU wall time), {cores} threads, "
                                  "fresh decoder per stream, i16 out"}
    fr_hash = hashlib.sha256((sums & np.uint64(0xFFFFFFFF)).astype(np.uint32).tobytes()).hexdigest()
    pcm_hash = hashlib.sha256((sums >> np.uint64(32)).astype(np.uint32).tobytes()).hexdigest()
    checks = {"streams": int(len(sums)), "final_range_sha256": fr_hash, "pcm_bits_sha256": pcm_hash,
              "what": "per-stream sums over the last timed step, all-gathered with NCCL" if world > 1 else "per-stream sums over the last timed step"}
    if strong:
        gold = ROOT / "tests" / "golden" / "strong_scaling_checksums.json"
        if gold.exists():
            table = json.loads(gold.read_text())
            want = table.get(f"{total_streams}x{ticks}x{args.warmup + args.steps}")
            if not want:   # another number of decodes of the corpus: a CELT-only packet's final_range does not depend on the stream's
                           # state, so any entry of the same corpus pins it (the PCM hash does depend on it and is left out)
                same = [v for k, v in table.items() if k.startswith(f"{total_streams}x{ticks}x")]
                want = {"final_range_sha256": same[0]["final_range_sha256"]} if same else None
            if want:
                checks["final_range_matches_golden"] = want["final_range_sha256"] == fr_hash
                if "pcm_bits_sha256" in want:
                    checks["pcm_matches_1gpu_run"] = want["pcm_bits_sha256"] == pcm_hash
                assert checks["final_range_matches_golden"], "per-stream final_range checksums differ from the committed 1-GPU / oracle vector"
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD if not strong else f"celt_only_48k_stereo_20ms_64kbps_{total_streams}_streams_x{ticks}_frames_one_job",
                   "streams_per_gpu": streams, "streams_total": total_streams, "frames_per_stream_per_step": ticks, "payload_bytes": PAYLOAD,
                   "packets": CORPUS_NOTE[CORPUS], "output": "f32 PCM left in HBM",
                   "l2": "per-step working set (scratch + PCM, several GB) exceeds the 126 MB L2", "sharding": f"streams by rank, {world} rank(s), no data-path collective"},
        "e2e": None if args.no_e2e else {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h_arena.nbytes + h_offsets.nbytes + h_ids.nbytes + n),
                "d2h_bytes_per_step": int(h_pcm.nbytes + h_samples.nbytes + h_range.nbytes), "ms_per_step": e2e_ms / e2e_steps,
                "steps": e2e_steps, "api": "BatchDecoder.decode_into -> bd_decode_i16 (mousiki Decoder::decode), pinned host buffers",
                "pcie_d2h_gbs": d2h_min, "pcie_d2h_note": "device -> pinned host rate of the slowest rank with all ranks copying at once; "
                "d2h_bytes_per_step / this = the floor of the e2e step"},
        "gpu_launches": launches_per_step * args.steps,
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": top_name, "achieved": achieved, "peak": peak, "peak_kind": peak_kind, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "alg_bytes_per_frame": ALG_BYTES[top],
                     "avg_launch_ms": top_ms / top_launches,
                     "path": {"alg_bytes_per_frame": PATH_BYTES, "achieved": PATH_BYTES * n * args.steps / (ms * 1e-3) / 1e9,
                              "frac": PATH_BYTES * n * args.steps / (ms * 1e-3) / 1e9 / peak}},
        "kernel_ms_per_step": {k: v[0] / args.steps for k, v in prof.items() if v[1]},
        "cpu_baseline": cpu_baseline,
        "parity": parity,
        "parity_e2e": e2e_parity,
        "checksums": checks,
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
    global CORPUS
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--streams", type=int, default=STREAMS, help="streams per GPU (default: the BASELINE workload)")
    ap.add_argument("--ticks", type=int, default=TICKS, help="frames per stream per step")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer arm (builder runs of very large jobs on one GPU)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --streams streams per GPU (the driver's contract); strong: one fixed corpus of --streams streams "
                         "(default 10^5) sharded over the ranks, per-stream checksums all-gathered and compared with the golden vector")
    ap.add_argument("--corpus", default=CORPUS, choices=["fuzz", "struct"],
                    help="packet construction of the weak-scaling workload: fuzz (TOC + random bytes) or struct (oracle/gen_corpus.cpp)")
    args = ap.parse_args()
    CORPUS = args.corpus
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
