"""Row A2: the batched lightweight decoder (k_lite through the C-ABI bd_lite_*; its kernel body also through the host
emulation) against the oracle's restatement of mousiki's `decoder::Decoder` (oracle/lite_oracle.cpp) on seeded streams:
NB / MB / WB 20 ms mono SILK packets of random payload (voiced, unvoiced and inactive frames, bandwidth switches inside a
stream) mixed with every packet kind the reference refuses. Bit-exact: status, f32 samples, i16 and 24-bit integers."""
import ctypes as C
import json
from pathlib import Path

import numpy as np
import pytest

import helpers

VEC = json.loads((Path(__file__).parent / "golden" / "lite_vectors.json").read_text())


def oracle():
    lib = helpers.oracle_lib()
    lib.orl_create.restype = C.c_void_p
    lib.orl_destroy.argtypes = [C.c_void_p]
    for f in (lib.orl_decode_f32, lib.orl_decode_i16, lib.orl_decode_i24):
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    return lib


def oracle_grid(arena, offsets, n_lanes, mode="f32"):
    lib = oracle()
    n = len(offsets) - 1
    dtype, fn = {"f32": (np.float32, lib.orl_decode_f32), "i16": (np.int16, lib.orl_decode_i16), "i24": (np.int32, lib.orl_decode_i24)}[mode]
    pcm = np.zeros((n, 960), dtype=dtype)
    status = np.zeros(n, dtype=np.int32)
    for j in range(n_lanes):
        h = lib.orl_create()
        for p in range(j, n, n_lanes):
            pkt = np.ascontiguousarray(arena[offsets[p]:offsets[p + 1]])
            row = np.zeros(960, dtype=dtype)
            status[p] = fn(h, helpers.ptr(pkt) if len(pkt) else None, len(pkt), helpers.ptr(row))
            if status[p] == 960:
                pcm[p] = row
        lib.orl_destroy(h)
    return pcm, status


def emul_grid(arena, offsets, n_lanes, mode="f32"):
    lib = helpers.emul_lib()
    lib.emu_lite_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int]
    n = len(offsets) - 1
    dtype, m = {"f32": (np.float32, 0), "i16": (np.int16, 1), "i24": (np.int32, 2)}[mode]
    pcm = np.zeros((n, 960), dtype=dtype)
    status = np.zeros(n, dtype=np.int32)
    arena = np.ascontiguousarray(np.concatenate([arena, np.zeros(8, dtype=np.uint8)]))
    offsets = np.ascontiguousarray(offsets, dtype=np.uint32)
    assert lib.emu_lite_decode(helpers.ptr(arena), helpers.ptr(offsets), n_lanes, n // n_lanes, helpers.ptr(pcm), helpers.ptr(status), m) == 0
    return pcm, status


def first_bytes_without_lbrr():
    """Payload first bytes whose two header bits decode to LBRR = 0 (half of all values), found with the oracle."""
    lib = oracle()
    ok = []
    row = np.zeros(960, dtype=np.float32)
    for b in range(256):
        h = lib.orl_create()
        pkt = np.asarray([0x48, b, 0x55, 0xAA, 0x33, 0xCC, 0x0F, 0xF0, 0x5A], dtype=np.uint8)
        if lib.orl_decode_f32(h, helpers.ptr(pkt), len(pkt), helpers.ptr(row)) == 960:
            ok.append(b)
        lib.orl_destroy(h)
    assert 64 <= len(ok) <= 192
    return np.asarray(ok, dtype=np.uint8)


def lite_grid(n_lanes, n_ticks, seed, refuse=0.08):
    rng = np.random.default_rng(seed)
    ok_first = first_bytes_without_lbrr()
    lane_bw = rng.integers(0, 3, n_lanes)
    pkts = []
    for t in range(n_ticks):
        for j in range(n_lanes):
            bw = int(lane_bw[j]) if rng.random() > 0.15 else int(rng.integers(0, 3))   # some streams switch bandwidth
            toc = ((bw * 4 + 1) << 3)
            body = rng.integers(0, 256, int(rng.integers(6, 90)), dtype=np.uint8)
            body[0] = ok_first[rng.integers(0, len(ok_first))]
            r = rng.random()
            if r < refuse:
                kind = int(rng.integers(0, 6))
                if kind == 0: pkt = np.zeros(0, dtype=np.uint8)                                    # too short
                elif kind == 1: pkt = np.concatenate([[toc | int(rng.integers(1, 4))], body])      # frame codes 1..3
                elif kind == 2: pkt = np.concatenate([[int(rng.integers(96, 256)) & 0xF8], body])  # hybrid / CELT
                elif kind == 3: pkt = np.concatenate([[((bw * 4 + int(rng.choice([0, 2, 3]))) << 3)], body])   # 10 / 40 / 60 ms
                elif kind == 4: pkt = np.concatenate([[toc | 4], body])                            # stereo
                else: pkt = np.concatenate([[toc], rng.integers(0, 256, 12, dtype=np.uint8)])      # any header bits (LBRR refusals)
            else:
                pkt = np.concatenate([[toc], body])
            pkts.append(pkt.astype(np.uint8))
    offsets = np.zeros(len(pkts) + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in pkts])
    return np.concatenate(pkts).astype(np.uint8), offsets


def assert_same(got, want, what):
    gp, gs = got
    wp, ws = want
    assert np.array_equal(gs, ws), (what, np.flatnonzero(gs != ws)[:8], gs[gs != ws][:8], ws[gs != ws][:8])
    if gp.dtype == np.float32:
        bad = np.flatnonzero((gp != wp).any(axis=1))
        assert len(bad) == 0, (what, bad[:8], float(np.abs(gp - wp).max()))
    else:
        assert np.array_equal(gp, wp), what


def test_emulation_matches_the_oracle():
    arena, offsets = lite_grid(24, 12, seed=5)
    want = oracle_grid(arena, offsets, 24)
    assert (want[1] == 960).mean() > 0.8 and len(set(want[1].tolist())) >= 6       # every refusal kind is in the corpus
    assert_same(emul_grid(arena, offsets, 24), want, "f32")
    assert_same(emul_grid(arena, offsets, 24, "i16"), oracle_grid(arena, offsets, 24, "i16"), "i16")
    assert_same(emul_grid(arena, offsets, 24, "i24"), oracle_grid(arena, offsets, 24, "i24"), "i24")


def test_emulation_reference_vectors():
    """The reference's own vectors through the kernel body: src/decoder.rs:298-316 and the Go fixtures
    src/silk/decoder.rs:2732-2780 (two consecutive frames of one stream), 1e-6."""
    pk = [np.asarray(VEC["packet"], dtype=np.uint8), np.asarray([0x48] + VEC["silk_frame_second"], dtype=np.uint8)]
    offsets = np.asarray([0, len(pk[0]), len(pk[0]) + len(pk[1])], dtype=np.uint32)
    pcm, status = emul_grid(np.concatenate(pk), offsets, 1)
    assert status.tolist() == [960, 960]
    assert np.abs(pcm[0, 0:36:3] - np.asarray(VEC["packet_prefix"], dtype=np.float32)).max() < 1e-6
    assert np.abs(pcm[0, ::3] - np.asarray(VEC["decode_out_0"], dtype=np.float32)).max() < 1e-6
    assert np.abs(pcm[1, ::3] - np.asarray(VEC["decode_out_1"], dtype=np.float32)).max() < 1e-6


@pytest.mark.gpu
def test_gpu_matches_the_oracle():
    from mousiki_b200 import LiteBatchDecoder
    arena, offsets = lite_grid(200, 8, seed=6)
    with LiteBatchDecoder(200) as dec:
        assert_same(dec.decode_float32(arena, offsets, 200), oracle_grid(arena, offsets, 200), "f32")
        dec.reset()
        assert_same(dec.decode(arena, offsets, 200), oracle_grid(arena, offsets, 200, "i16"), "i16")
        dec.reset()
        assert_same(dec.decode_int24(arena, offsets, 200), oracle_grid(arena, offsets, 200, "i24"), "i24")


@pytest.mark.gpu
def test_gpu_state_carries_across_calls():
    """Two calls of 3 ticks decode like one call of 6: the stream state lives in the handle."""
    from mousiki_b200 import LiteBatchDecoder
    arena, offsets = lite_grid(70, 6, seed=7)
    want = oracle_grid(arena, offsets, 70)
    half = 70 * 3
    with LiteBatchDecoder(80) as dec:
        a = dec.decode_float32(arena[:offsets[half]], offsets[:half + 1], 70)
        b = dec.decode_float32(arena[offsets[half]:], offsets[half:] - offsets[half], 70)
    assert_same((np.concatenate([a[0], b[0]]), np.concatenate([a[1], b[1]])), want, "split")


@pytest.mark.gpu
def test_gpu_reference_vectors_and_tiny_ogg():
    """src/decoder.rs:298-316, the Go fixtures, and BASELINE config 1: the packet of testdata/tiny.ogg (tests/golden/tiny.ogg)
    through `decode_float32` like examples/decode.rs does."""
    from mousiki_b200 import LiteBatchDecoder, loaders
    pk = [np.asarray(VEC["packet"], dtype=np.uint8), np.asarray([0x48] + VEC["silk_frame_second"], dtype=np.uint8)]
    offsets = np.asarray([0, len(pk[0]), len(pk[0]) + len(pk[1])], dtype=np.uint32)
    with LiteBatchDecoder(1) as dec:
        pcm, status = dec.decode_float32(np.concatenate(pk), offsets, 1)
    assert status.tolist() == [960, 960]
    assert np.abs(pcm[0, 0:36:3] - np.asarray(VEC["packet_prefix"], dtype=np.float32)).max() < 1e-6
    assert np.abs(pcm[0, ::3] - np.asarray(VEC["decode_out_0"], dtype=np.float32)).max() < 1e-6
    assert np.abs(pcm[1, ::3] - np.asarray(VEC["decode_out_1"], dtype=np.float32)).max() < 1e-6
    _, packets = loaders.read_ogg_opus((Path(__file__).parent / "golden" / "tiny.ogg").read_bytes(), by_segment=True)   # examples/decode.rs:68-93
    assert bytes(packets[0]) == bytes.fromhex("4883cade8ae567d51caca254faffbf")
    arena = np.concatenate([np.frombuffer(bytes(p), dtype=np.uint8) for p in packets])
    offsets = np.zeros(len(packets) + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in packets])
    with LiteBatchDecoder(1) as dec:
        got = dec.decode_float32(arena, offsets, 1)
    assert_same(got, oracle_grid(arena, offsets, 1), "tiny.ogg")
    assert (got[1] == 960).all()
