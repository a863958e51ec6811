"""Host-side mirror of the reference's decoder API for a batch of streams, over the C ABI.

Reference interface mirrored (cijiugechu/mousiki):
  mousiki::Decoder::{new, decode, decode_float, reset, final_range, last_packet_duration}
      src/codec.rs:792-977
  c_style_api::opus_decoder::{opus_decode, opus_decode_float}   src/opus_decoder.rs:1632-1826
One `BatchDecoder` owns n_streams decoder states on one GPU; `decode_float` / `decode` take a batch of
packets tagged with stream ids (packets of a stream in order) and return interleaved PCM plus, per
packet, the sample count or Opus error code and `final_range`.

This module is only a ctypes binding: all work happens in libmousiki_b200.so (CUDA, sm_100a).
There is no CPU fallback; constructing a decoder without the library or without a GPU raises.
"""
import ctypes as C
from pathlib import Path

import numpy as np

import os

# MOUSIKI_B200_LIB: another build of the same library (A/B runs of kernel variants, tools/ab_build.sh)
_LIB_PATH = Path(os.environ.get("MOUSIKI_B200_LIB") or Path(__file__).resolve().parent / "libmousiki_b200.so")
QUIRKS_MOUSIKI = 0xF7F

OPUS_BAD_ARG, OPUS_BUFFER_TOO_SMALL, OPUS_INTERNAL_ERROR, OPUS_INVALID_PACKET, OPUS_UNIMPLEMENTED = -1, -2, -3, -4, -5

EXPORTS = [
    "bd_create", "bd_destroy", "bd_reset_streams", "bd_last_error", "bd_decode_f32", "bd_decode_i16", "bd_decode_grid_dev",
    "bd_get_stream_info", "bd_alloc_pinned", "bd_free_pinned", "bd_last_launch_count", "bd_debug_capture", "bd_profile_enable", "bd_profile_read",
    "bd_packet_get_mode", "bd_packet_get_bandwidth", "bd_packet_get_nb_channels", "bd_packet_get_samples_per_frame", "bd_packet_get_nb_frames",
    "bd_packet_get_nb_samples", "bd_debug_kat", "bd_decode_i24", "bd_parse_noalloc",
    "bd_lite_create", "bd_lite_destroy", "bd_lite_reset", "bd_lite_decode_f32", "bd_lite_decode_i16", "bd_lite_decode_i24",
]


class BdConfig(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("n_streams", C.c_uint32), ("sample_rate_hz", C.c_int32), ("channels", C.c_int32),
                ("gain_q8", C.c_int32), ("disable_phase_inversion", C.c_int32), ("quirks", C.c_uint32), ("device", C.c_int32),
                ("max_frames_per_pass", C.c_uint32), ("reserved", C.c_uint32)]


class BdStreamInfo(C.Structure):
    _fields_ = [("final_range", C.c_uint32), ("last_packet_duration", C.c_int32), ("prev_mode", C.c_int32), ("celt_error", C.c_int32),
                ("bandwidth", C.c_int32), ("pitch", C.c_int32)]


class DecoderError(RuntimeError):
    pass


_lib = None


def load_library():
    """dlopen the in-tree CUDA library and declare the C ABI. Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise DecoderError(f"{_LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(nvcc, sm_100a). mousiki_b200 has no CPU fallback.")
    lib = C.CDLL(str(_LIB_PATH))
    vp, u32p = C.c_void_p, C.c_void_p
    lib.bd_create.argtypes = [C.POINTER(BdConfig), C.POINTER(vp)]
    lib.bd_destroy.argtypes = [vp]
    lib.bd_destroy.restype = None
    lib.bd_reset_streams.argtypes = [vp, u32p, C.c_uint32]
    lib.bd_last_error.argtypes = [vp]
    lib.bd_last_error.restype = C.c_char_p
    for fn in (lib.bd_decode_f32, lib.bd_decode_i16, lib.bd_decode_i24):
        fn.argtypes = [vp, vp, vp, vp, C.c_uint32, vp, vp, C.c_uint32, vp, vp]
    lib.bd_decode_grid_dev.argtypes = [vp, vp, vp, C.c_uint32, C.c_uint32, vp, vp, vp, C.c_uint32, vp, vp, vp]
    lib.bd_get_stream_info.argtypes = [vp, C.c_uint32, C.POINTER(BdStreamInfo)]
    lib.bd_alloc_pinned.argtypes = [C.c_uint64]
    lib.bd_alloc_pinned.restype = vp
    lib.bd_free_pinned.argtypes = [vp]
    lib.bd_free_pinned.restype = None
    lib.bd_last_launch_count.argtypes = [vp]
    lib.bd_last_launch_count.restype = C.c_uint32
    lib.bd_debug_capture.argtypes = [vp, vp, vp, C.c_uint32]
    lib.bd_profile_enable.argtypes = [vp, C.c_int]
    lib.bd_profile_read.argtypes = [vp, vp, vp]
    for name in ("bd_packet_get_mode", "bd_packet_get_bandwidth", "bd_packet_get_nb_channels"):
        getattr(lib, name).argtypes = [C.c_char_p, C.c_uint32]
    lib.bd_packet_get_samples_per_frame.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32]
    lib.bd_packet_get_nb_frames.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32]
    lib.bd_packet_get_nb_samples.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32]
    _lib = lib
    return lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def pack_packets(packets):
    """list of bytes-like -> (arena uint8, offsets uint32[n+1])"""
    lens = np.fromiter((len(p) for p in packets), dtype=np.int64, count=len(packets))
    offsets = np.zeros(len(packets) + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum(lens)
    arena = np.frombuffer(b"".join(bytes(p) for p in packets), dtype=np.uint8).copy() if len(packets) else np.zeros(0, np.uint8)
    return arena, offsets


class BatchDecoder:
    """n_streams independent Opus decoders resident on one B200 (mirror of mousiki::Decoder per stream)."""

    def __init__(self, n_streams, sample_rate=48000, channels=2, gain_q8=0, disable_phase_inversion=None, quirks=QUIRKS_MOUSIKI,
                 device=0, max_frames_per_pass=0):
        self._lib = load_library()
        self._h = C.c_void_p()
        self.n_streams, self.sample_rate, self.channels = int(n_streams), int(sample_rate), int(channels)
        cfg = BdConfig(C.sizeof(BdConfig), self.n_streams, self.sample_rate, self.channels, int(gain_q8),
                       -1 if disable_phase_inversion is None else int(bool(disable_phase_inversion)), int(quirks), int(device),
                       int(max_frames_per_pass), 0)
        rc = self._lib.bd_create(C.byref(cfg), C.byref(self._h))
        if rc != 0:
            self._h = C.c_void_p()
            what = {-1: "bad argument (OpusDecoderInitError::BadArgument)", -5: "configuration not implemented yet",
                    -100: "no usable CUDA device (mousiki_b200 has no CPU fallback)", -101: "device allocation failed"}.get(rc, "error")
            raise DecoderError(f"bd_create failed ({rc}): {what}")

    # -- lifetime
    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.bd_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc, what):
        if rc != 0:
            raise DecoderError(f"{what} failed ({rc}): {self._lib.bd_last_error(self._h).decode(errors='replace')}")

    # -- Decoder::reset
    def reset(self, stream_ids=None):
        if stream_ids is None:
            self._check(self._lib.bd_reset_streams(self._h, None, 0), "bd_reset_streams")
        else:
            ids = np.ascontiguousarray(stream_ids, dtype=np.uint32)
            self._check(self._lib.bd_reset_streams(self._h, _p(ids), len(ids)), "bd_reset_streams")

    # -- Decoder::decode_float / Decoder::decode on a batch
    def _decode(self, arena, offsets, stream_ids, frame_size, flags, i16, capture=False, i24=False):
        arena = np.ascontiguousarray(arena, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
        stream_ids = np.ascontiguousarray(stream_ids, dtype=np.uint32)
        n = len(stream_ids)
        if len(offsets) != n + 1:
            raise ValueError("offsets must have n_packets + 1 entries")
        stride = int(frame_size) * self.channels
        pcm = np.zeros((n, stride), dtype=np.int16 if i16 else (np.int32 if i24 else np.float32))
        samples = np.zeros(n, dtype=np.int32)
        rng = np.zeros(n, dtype=np.uint32)
        fl = None if flags is None else np.ascontiguousarray(flags, dtype=np.uint8)
        iy = xs = None
        if capture:
            iy = np.zeros((n, 2, 960), dtype=np.int32)
            xs = np.zeros((n, 2, 960), dtype=np.float32)
            self._lib.bd_debug_capture(self._h, _p(iy), _p(xs), n)
        fn = self._lib.bd_decode_i16 if i16 else (self._lib.bd_decode_i24 if i24 else self._lib.bd_decode_f32)
        rc = fn(self._h, _p(arena), _p(offsets), _p(stream_ids), n, _p(fl), _p(pcm), stride, _p(samples), _p(rng))
        if capture:
            self._lib.bd_debug_capture(self._h, None, None, 0)
        self._check(rc, "bd_decode")
        return (pcm, samples, rng, iy, xs) if capture else (pcm, samples, rng)

    def decode_into(self, arena, offsets, stream_ids, pcm_out, samples_out, final_range_out, flags=None):
        """Zero-allocation form of decode/decode_float: caller-owned (ideally pinned) numpy buffers.
        pcm_out: [n, stride] float32 (decode_float) or int16 (decode); stride = frame_size * channels."""
        n = len(stream_ids)
        assert pcm_out.dtype in (np.float32, np.int16) and pcm_out.flags.c_contiguous and pcm_out.shape[0] == n
        fn = self._lib.bd_decode_i16 if pcm_out.dtype == np.int16 else self._lib.bd_decode_f32
        rc = fn(self._h, _p(arena), _p(offsets), _p(stream_ids), n, _p(flags), _p(pcm_out), pcm_out.shape[1], _p(samples_out),
                _p(final_range_out))
        self._check(rc, "bd_decode")

    def decode_float(self, arena, offsets, stream_ids, frame_size=960, flags=None, capture=False):
        """-> (pcm float32 [n, frame_size*channels], samples int32 [n], final_range uint32 [n])"""
        return self._decode(arena, offsets, stream_ids, frame_size, flags, False, capture)

    def decode(self, arena, offsets, stream_ids, frame_size=960, flags=None):
        """i16 output with the reference's soft clip (opus_decode)."""
        return self._decode(arena, offsets, stream_ids, frame_size, flags, True)

    def decode_24bit(self, arena, offsets, stream_ids, frame_size=960, flags=None):
        """24-bit samples in int32 containers (Decoder::decode_24bit, src/codec.rs:867; opus_decode24): no soft clip."""
        return self._decode(arena, offsets, stream_ids, frame_size, flags, False, i24=True)

    def decode_float_packets(self, packets, stream_ids, frame_size=960):
        arena, offsets = pack_packets(packets)
        return self.decode_float(arena, offsets, stream_ids, frame_size)

    # -- device-resident grid (torch tensors or raw device pointers as ints)
    def decode_grid_device(self, d_arena, d_offsets, n_lanes, n_ticks, d_pcm_f32, pcm_stride, d_samples, d_final_range,
                           d_lane_stream=0, d_pcm_i16=0, stream=0):
        rc = self._lib.bd_decode_grid_dev(self._h, C.c_void_p(d_arena), C.c_void_p(d_offsets), n_lanes, n_ticks,
                                          C.c_void_p(d_lane_stream or None), C.c_void_p(d_pcm_f32 or None), C.c_void_p(d_pcm_i16 or None),
                                          pcm_stride, C.c_void_p(d_samples), C.c_void_p(d_final_range), C.c_void_p(stream or None))
        self._check(rc, "bd_decode_grid_dev")

    KERNELS = ("k_parse", "k_energy", "k_pvq", "k_plan", "k_bands", "k_imdct", "k_comb", "k_deemph", "k_clip_i16", "k_silk")

    def profile(self, on=True):
        self._check(self._lib.bd_profile_enable(self._h, int(on)), "bd_profile_enable")

    def profile_read(self):
        """-> {kernel: (total_ms, launches)} accumulated since the last read (CUDA-event timed on the launch stream)."""
        ms = np.zeros(len(self.KERNELS), dtype=np.float64)
        cnt = np.zeros(len(self.KERNELS), dtype=np.uint32)
        self._check(self._lib.bd_profile_read(self._h, _p(ms), _p(cnt)), "bd_profile_read")
        return {k: (float(ms[i]), int(cnt[i])) for i, k in enumerate(self.KERNELS)}

    def launch_count(self):
        return int(self._lib.bd_last_launch_count(self._h))

    def stream_info(self, stream_id):
        info = BdStreamInfo()
        self._check(self._lib.bd_get_stream_info(self._h, int(stream_id), C.byref(info)), "bd_get_stream_info")
        return {"final_range": info.final_range, "last_packet_duration": info.last_packet_duration, "prev_mode": info.prev_mode,
                "celt_error": info.celt_error, "bandwidth": info.bandwidth, "pitch": info.pitch}


# ---- packet queries without a decoder (src/packet.rs:30-56, 177-280; Decoder::packet_samples, src/codec.rs:838) --------------
class PacketError(ValueError):
    """PacketError of the reference: .code is the Opus error code (-1 BadArgument, -4 InvalidPacket)."""

    def __init__(self, code):
        super().__init__({-1: "BadArgument", -4: "InvalidPacket"}.get(code, f"error {code}"))
        self.code = code


def _query(name, packet, *args):
    packet = bytes(packet)
    r = getattr(load_library(), name)(packet, len(packet), *args)
    if r < 0:
        raise PacketError(r)
    return r


def packet_get_mode(packet):
    """-> 1000 SILK, 1001 hybrid, 1002 CELT (opus_packet_get_mode)."""
    return _query("bd_packet_get_mode", packet)


def packet_get_bandwidth(packet):
    """-> 1101 narrow .. 1105 full (opus_packet_get_bandwidth, Bandwidth::to_opus_int)."""
    return _query("bd_packet_get_bandwidth", packet)


def packet_get_nb_channels(packet):
    return _query("bd_packet_get_nb_channels", packet)


def packet_get_samples_per_frame(packet, fs_hz=48000):
    return _query("bd_packet_get_samples_per_frame", packet, int(fs_hz))


def packet_get_nb_frames(packet, length=None):
    return _query("bd_packet_get_nb_frames", packet, len(packet) if length is None else int(length))


def packet_get_nb_samples(packet, length=None, fs_hz=48000):
    """Samples per channel the packet decodes to (opus_packet_get_nb_samples; `Decoder::packet_samples` at the decoder's rate)."""
    return _query("bd_packet_get_nb_samples", packet, len(packet) if length is None else int(length), int(fs_hz))


class LiteBatchDecoder:
    """mousiki's lightweight `decoder::Decoder` (src/decoder.rs:142-262), one per stream, batched on the GPU: SILK-only, mono,
    20 ms packets, 960 samples out per packet (row A2; include/mousiki_b200.h bd_lite_*). `decode_float32`, `decode` and
    `decode_int24` take a tick-major grid of packets (packet p = tick * n_lanes + lane, lane j = stream j) and return
    (pcm[n_packets, 960], status[n_packets]) with status 960 or the reference's error code for that packet."""

    def __init__(self, n_streams: int, device: int = 0):
        self._lib = load_library()
        lib = self._lib
        lib.bd_lite_create.argtypes = [C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]
        lib.bd_lite_destroy.argtypes = [C.c_void_p]
        lib.bd_lite_reset.argtypes = [C.c_void_p]
        for f in (lib.bd_lite_decode_f32, lib.bd_lite_decode_i16, lib.bd_lite_decode_i24):
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        h = C.c_void_p()
        rc = lib.bd_lite_create(device, n_streams, C.byref(h))
        if rc != 0:
            raise DecoderError(f"bd_lite_create failed: {rc}")
        self._h = h
        self.n_streams = n_streams

    def close(self):
        if getattr(self, "_h", None):
            self._lib.bd_lite_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        self.close()

    def reset(self):
        rc = self._lib.bd_lite_reset(self._h)
        if rc != 0:
            raise DecoderError(f"bd_lite_reset failed: {rc}")

    def _run(self, fn, dtype, arena, offsets, n_lanes):
        arena = np.ascontiguousarray(arena, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
        n = len(offsets) - 1
        if n % n_lanes:
            raise ValueError("the packet grid must hold the same number of packets for every lane")
        pcm = np.zeros((n, 960), dtype=dtype)
        status = np.zeros(n, dtype=np.int32)
        rc = fn(self._h, arena.ctypes.data_as(C.c_void_p), offsets.ctypes.data_as(C.c_void_p), n_lanes, n // n_lanes,
                pcm.ctypes.data_as(C.c_void_p), status.ctypes.data_as(C.c_void_p))
        if rc != 0:
            raise DecoderError(f"lightweight decode failed: {rc}")
        return pcm, status

    def decode_float32(self, arena, offsets, n_lanes):
        return self._run(self._lib.bd_lite_decode_f32, np.float32, arena, offsets, n_lanes)

    def decode(self, arena, offsets, n_lanes):
        return self._run(self._lib.bd_lite_decode_i16, np.int16, arena, offsets, n_lanes)

    def decode_int24(self, arena, offsets, n_lanes):
        return self._run(self._lib.bd_lite_decode_i24, np.int32, arena, offsets, n_lanes)
