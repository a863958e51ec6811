"""Shared test helpers: seeded packet corpora, ctypes bindings for the oracle (CPU restatement of the
reference, the checker) and for the host emulation of the kernel bodies (CPU tests only)."""
import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
QUIRKS_MOUSIKI = 0xF7F  # CELT Q1..Q9 (bits 0-6) + SILK S1..S4 (bits 8-11): mousiki as written


def _make(dirpath: Path, target: str) -> Path:
    subprocess.run(["make", "-C", str(dirpath), target], check=True, stdout=subprocess.DEVNULL)
    return dirpath / target


_ORACLE = None
_EMUL = None


def oracle_lib():
    global _ORACLE
    if _ORACLE is None:
        lib = C.CDLL(str(_make(ROOT / "oracle", "_build/liboracle.so")))
        lib.orc_create.restype = C.c_void_p
        lib.orc_create.argtypes = [C.c_int, C.c_int, C.c_uint32]
        lib.orc_destroy.argtypes = [C.c_void_p]
        lib.orc_reset.argtypes = [C.c_void_p]
        lib.orc_set_gain.argtypes = [C.c_void_p, C.c_int]
        lib.orc_final_range.restype = C.c_uint32
        lib.orc_final_range.argtypes = [C.c_void_p]
        lib.orc_decode_float.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int]
        lib.orc_decode_i16.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int]
        lib.orc_trace_enable.argtypes = [C.c_void_p, C.c_int]
        lib.orc_last_redundancy.argtypes = [C.c_void_p]
        lib.orc_matrix_out_float.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        lib.orc_matrix_out_short.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        lib.orc_trace_info.argtypes = [C.c_void_p, C.c_void_p]
        lib.orc_trace_get.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        lib.orc_decode_grid.argtypes = [C.c_int, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        _ORACLE = lib
    return _ORACLE


def emul_lib():
    global _EMUL
    if _EMUL is None:
        lib = C.CDLL(str(_make(ROOT / "tests" / "emul", "_build/libemul.so")))
        lib.emu_decode_grid.argtypes = [C.c_int, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _EMUL = lib
    return _EMUL


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def lcg_bytes(seed: int, n: int) -> np.ndarray:
    """Deterministic byte stream (numpy PCG is fine: the generator only has to be seeded and committed)."""
    return np.random.default_rng(seed).integers(0, 256, size=n, dtype=np.uint8)


def celt_toc(lm: int, stereo: bool, bandwidth: int = 3) -> int:
    """CELT-only TOC, code 0: config = 16 + 4*bandwidth(0 NB,1 WB,2 SWB,3 FB) + lm."""
    return ((16 + 4 * bandwidth + lm) << 3) | (4 if stereo else 0)


def fuzz_grid(n_lanes: int, n_ticks: int, payload: int, toc_fn, seed: int = 1):
    """[n_ticks][n_lanes] grid of `TOC || random bytes` packets (the construction of the reference's own
    decode fuzz, tests/test_opus_decode.rs:372-407). toc_fn(t, j, rng) -> TOC byte; payload may be an int or
    a (lo, hi) range. Returns (arena uint8, offsets uint32[n+1])."""
    rng = np.random.default_rng(seed)
    n = n_lanes * n_ticks
    if isinstance(payload, tuple):
        lens = rng.integers(payload[0], payload[1] + 1, size=n)
    else:
        lens = np.full(n, payload)
    offsets = np.zeros(n + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum(lens + 1)
    arena = rng.integers(0, 256, size=int(offsets[-1]), dtype=np.uint8)
    for t in range(n_ticks):
        for j in range(n_lanes):
            p = t * n_lanes + j
            arena[offsets[p]] = toc_fn(t, j, rng)
    return arena, offsets


def oracle_decode_grid(arena, offsets, n_lanes, n_ticks, channels, frame_size, i16=False, quirks=QUIRKS_MOUSIKI, threads=None, fs=48000):
    lib = oracle_lib()
    n = n_lanes * n_ticks
    pcm = np.zeros((n, frame_size * channels), dtype=np.int16 if i16 else np.float32)
    samples = np.zeros(n, dtype=np.int32)
    rng = np.zeros(n, dtype=np.uint32)
    lib.orc_decode_grid(fs, channels, quirks, ptr(arena), ptr(offsets), n_lanes, n_ticks, frame_size,
                        None if i16 else ptr(pcm), ptr(pcm) if i16 else None, ptr(samples), ptr(rng),
                        threads or os.cpu_count() or 1)
    return pcm, samples, rng


def emul_decode_grid(arena, offsets, n_lanes, n_ticks, channels, frame_size, i16=False, quirks=QUIRKS_MOUSIKI, debug=False, fs=48000):
    lib = emul_lib()
    lib.emu_set_downsample(48000 // fs)
    n = n_lanes * n_ticks
    stride = frame_size * channels
    pcm = np.zeros((n, stride), dtype=np.float32)
    pcm16 = np.zeros((n, stride), dtype=np.int16) if i16 else None
    samples = np.zeros(n, dtype=np.int32)
    rng = np.zeros(n, dtype=np.uint32)
    iy = np.zeros((n, 2, 960), dtype=np.int32) if debug else None
    xs = np.zeros((n, 2, 960), dtype=np.float32) if debug else None
    lib.emu_decode_grid(channels, quirks, ptr(arena), ptr(offsets), n_lanes, n_ticks, stride, ptr(pcm), ptr(pcm16), ptr(samples),
                        ptr(rng), ptr(iy), ptr(xs))
    lib.emu_set_downsample(1)
    return (pcm16 if i16 else pcm), samples, rng, iy, xs


def silk_toc(bandwidth: int, size_code: int, stereo: bool) -> int:
    """SILK-only TOC, code 0: config = 4*bandwidth(0 NB, 1 MB, 2 WB) + size_code(0: 10 ms, 1: 20, 2: 40, 3: 60)."""
    return ((4 * bandwidth + size_code) << 3) | (4 if stereo else 0)


def silk_fuzz_grid(n_lanes: int, n_ticks: int, payload, toc_fn, seed: int = 1, channels: int = 2, redundancy: bool = False):
    """Like fuzz_grid, for SILK-only TOCs, keeping only packets WITHOUT a redundancy frame: in SILK-only mode any
    packet with >= 17 unread bits after the SILK payload carries a CELT redundancy frame (opus_decoder.rs:746-759),
    an arm the CUDA path reports as unimplemented. The entropy decode is state-free, so candidates are probed one by
    one with a fresh oracle decoder. toc_fn(k, rng) -> TOC byte for the k-th candidate. redundancy=True keeps those
    packets too (whatever decodes on a fresh decoder)."""
    lib = oracle_lib()
    rng = np.random.default_rng(seed)
    need = n_lanes * n_ticks
    pkts = []
    pcm = np.zeros(2880 * 2, dtype=np.float32)
    k = 0
    while len(pkts) < need:
        ln = int(rng.integers(payload[0], payload[1] + 1)) if isinstance(payload, tuple) else int(payload)
        pkt = rng.integers(0, 256, size=ln + 1, dtype=np.uint8)
        pkt[0] = toc_fn(k, rng)
        k += 1
        h = lib.orc_create(48000, channels, QUIRKS_MOUSIKI)
        r = lib.orc_decode_float(h, ptr(pkt), len(pkt), ptr(pcm), 2880, 0)
        red = lib.orc_last_redundancy(h)
        lib.orc_destroy(h)
        if r > 0 and (redundancy or not red):
            pkts.append(pkt)
        assert k < 200 * need, "payload range leaves too many packets with redundancy"
    offsets = np.zeros(need + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in pkts])
    return np.concatenate(pkts), offsets


# ---------------------------------------------------------------------------------------------- multi-frame packets
def _frame_len_bytes(n: int) -> bytes:
    """RFC 6716 section 3.2.1 frame length coding."""
    if n < 252:
        return bytes([n])
    return bytes([252 + (n & 3), (n - (252 + (n & 3))) // 4])


def multiframe_packet(toc: int, frames, code: int, vbr: bool = False, pad: int = 0) -> np.ndarray:
    """Packs `frames` (byte strings) behind a TOC with frame-count code 1, 2 or 3 (RFC 6716 section 3.2). code 1 needs two
    equal-sized frames, code 2 two frames, code 3 any count (CBR unless vbr), optionally with `pad` bytes of padding (plus its length bytes)."""
    toc = (toc & ~3) | code
    frames = [bytes(f) for f in frames]
    if code == 1:
        assert len(frames) == 2 and len(frames[0]) == len(frames[1])
        body = frames[0] + frames[1]
    elif code == 2:
        assert len(frames) == 2
        body = _frame_len_bytes(len(frames[0])) + frames[0] + frames[1]
    else:
        fc = len(frames) | (0x80 if vbr else 0) | (0x40 if pad else 0)
        body = bytes([fc])
        padding = b""
        if pad:  # `pad` padding bytes: each 255 length byte stands for 254 of them and announces another length byte
            body += bytes([255] * (pad // 254)) + bytes([pad % 254])
            padding = bytes(pad)
        if vbr:
            for f in frames[:-1]:
                body += _frame_len_bytes(len(f))
        else:
            assert len({len(f) for f in frames}) == 1
        body += b"".join(frames) + padding
    return np.frombuffer(bytes([toc]) + body, dtype=np.uint8).copy()


def split_batch(arena, offsets, stream_ids, cap_per_ch, channels, flags=None):
    """The product's host front end (mousiki_b200/csrc/mb_host_packets.h) through the emulation library."""
    lib = emul_lib()
    lib.emu_split_batch.restype = C.c_uint32
    n = len(offsets) - 1
    arena = np.ascontiguousarray(arena, dtype=np.uint8)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
    stream_ids = np.ascontiguousarray(stream_ids, dtype=np.uint32)
    nbytes = C.c_uint32(0)
    nv = lib.emu_split_batch(ptr(arena), ptr(offsets), ptr(stream_ids), ptr(flags) if flags is not None else None, n, int(cap_per_ch),
                             int(channels), C.byref(nbytes))
    out = {"arena": np.zeros(max(nbytes.value, 1), dtype=np.uint8), "offsets": np.zeros(nv + 1, dtype=np.uint32),
           "stream_ids": np.zeros(nv, dtype=np.uint32), "flags": np.zeros(nv, dtype=np.uint8), "out_map": np.zeros(2 * nv, dtype=np.uint32),
           "first": np.zeros(n + 1, dtype=np.uint32)}
    lib.emu_split_fetch(ptr(out["arena"]), ptr(out["offsets"]), ptr(out["stream_ids"]), ptr(out["flags"]), ptr(out["out_map"]),
                        ptr(out["first"]))
    out["arena"] = out["arena"][:nbytes.value]
    return out


_ORCGEN = None


def orcgen_lib():
    global _ORCGEN
    if _ORCGEN is None:
        lib = C.CDLL(str(_make(ROOT / "oracle", "_build/liborcgen.so")))
        lib.orcgen_celt_grid.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_uint32,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        _ORCGEN = lib
    return _ORCGEN


STRUCT_LEVEL = 0.0   # mean band energy of a P-struct stream relative to the codec's eMeans (log2 units); see DESIGN.md section 6


def struct_grid(n_lanes: int, n_ticks: int, payload: int, toc: int, channels: int, seed: int = 1, level: float = STRUCT_LEVEL,
                want_pcm: bool = False, quirks: int = QUIRKS_MOUSIKI, threads=None):
    """P-struct: [n_ticks][n_lanes] grid of structured CELT-only packets (oracle/gen_corpus.cpp: written through the restated
    decoder, band energies steered along a per-stream envelope, ~10 % transients, post-filter on ~30 % of the frames, PCM inside
    +-1). Same layout as fuzz_grid. -> (arena, offsets[, pcm, final_range])."""
    lib = orcgen_lib()
    n = n_lanes * n_ticks
    arena = np.zeros(n * (payload + 1), dtype=np.uint8)
    offsets = np.zeros(n + 1, dtype=np.uint32)
    frame = 120 << ((toc >> 3) & 3)
    pcm = np.zeros((n, frame * channels), dtype=np.float32) if want_pcm else None
    rng = np.zeros(n, dtype=np.uint32) if want_pcm else None
    rc = lib.orcgen_celt_grid(seed, n_lanes, n_ticks, payload, toc, channels, level, quirks & 0x7F, ptr(arena), ptr(offsets), ptr(pcm), ptr(rng),
                              threads or os.cpu_count() or 1)
    assert rc == 0, "the corpus writer could not finish a frame"
    return (arena, offsets, pcm, rng) if want_pcm else (arena, offsets)
