"""Pins the SILK / hybrid part of the oracle against the reference's own known answers (CPU only).

Sources (paths in the reference tree):
  tests/test_opus_decode.rs:344-370      lres: sum of final_range over 65536 four-byte SILK packets (NB/MB/WB 10 ms)
  tests/hybrid_decode.rs + tests/fixtures/hybrid_decode_vectors.rs
                                         SILK WB packet then hybrid FB packet: final_range and 960 f32 samples (tol 1e-4)
  src/silk/nlsf_decode.rs:197-223, lpc_inv_pred_gain.rs:222-262, gain_quant.rs:120-178, decode_pitch.rs:96-122,
  log2lin.rs:41-70, resampler_private_up2_hq.rs:133-163  unit known answers

Two of the reference's vectors are only reproduced with one quirk switched to the libopus behaviour; that is
asserted here in both directions so the finding stays visible:
  * lres[1] (12 kHz, 10 ms) needs QUIRK_S3 off — the reference itself only checks lres under
    TEST_OPUS_STRICT_FINAL_RANGE (tests/test_opus_decode.rs:74-76,368);
  * the hybrid PCM vector (produced by opus-c) needs QUIRK_S2 off.
"""
import ctypes as C

import numpy as np
import pytest

import helpers

S1, S2, S3, S4 = 1 << 8, 1 << 9, 1 << 10, 1 << 11
ALL = helpers.QUIRKS_MOUSIKI


def _i8(v):
    return np.array(v, dtype=np.int8)


def _i16(v):
    return np.array(v, dtype=np.int16)


def test_log2lin_table():
    lib = helpers.oracle_lib()
    lib.orc_silk_log2lin.restype = C.c_int32
    cases = [(-128, 0), (-1, 0), (0, 1), (1, 1), (2, 1), (127, 1), (128, 2), (203, 3), (256, 4), (296, 4), (384, 8), (512, 16),
             (640, 32), (896, 128), (1024, 256), (1919, 32512), (2048, 65536), (3966, 2122317824), (3967, 2147483647), (2264, 210944)]
    for x, y in cases:
        assert lib.orc_silk_log2lin(x) == y, x


def test_inverse_pred_gain_known_answers():
    lib = helpers.oracle_lib()
    lib.orc_silk_inverse_pred_gain.restype = C.c_int32
    for coeffs, want in [([2048], 805306368), ([1024, -512, 256, -128], 1006430076), ([4096], 0),
                         ([3000, -2000, 1000, -500], 691862120), ([0] * 24, 1 << 30)]:
        a = _i16(coeffs)
        assert lib.orc_silk_inverse_pred_gain(helpers.ptr(a), len(coeffs)) == want


def test_gains_dequant_known_answers():
    lib = helpers.oracle_lib()
    g = np.zeros(4, dtype=np.int32)
    ind = _i8([6, 0, 5, 7])
    assert lib.orc_silk_gains_dequant(helpers.ptr(ind), 10, 0, 4, helpers.ptr(g)) == 6
    assert g.tolist() == [210944, 112640, 131072, 210944]
    ind = _i8([0, 4, 8, 0])
    assert lib.orc_silk_gains_dequant(helpers.ptr(ind), 4, 1, 4, helpers.ptr(g)) == 0
    assert g.tolist() == [81920, 81920, 153600, 81920]


def test_decode_pitch_known_answers():
    lib = helpers.oracle_lib()
    lags = np.zeros(4, dtype=np.int32)
    for args, want in [((0, 1, 16, 4), [32, 32, 33, 33]), ((5, 2, 8, 4), [20, 21, 22, 23]), ((3, 5, 16, 2), [34, 37, 0, 0]),
                       ((2, 1, 8, 2), [19, 18, 0, 0])]:
        lib.orc_silk_decode_pitch(*args, helpers.ptr(lags))
        assert lags.tolist() == want


def test_nlsf_decode_known_answer():
    lib = helpers.oracle_lib()
    idx = _i8([9, -1, 2, 0, -3, 1, -2, 3, -1, 0, 2, -2, 1, -3, 2, 0, -1])
    out = np.zeros(16, dtype=np.int16)
    lib.orc_silk_nlsf_decode(1, helpers.ptr(idx), helpers.ptr(out), ALL)
    assert out.tolist() == [480, 5399, 5689, 5692, 10036, 10039, 14962, 14976, 17231, 20376, 20387, 22638, 22646, 28263, 28458, 30080]
    idx = _i8([0] * 17)
    lib.orc_silk_nlsf_decode(1, helpers.ptr(idx), helpers.ptr(out), ALL)
    assert out[0] > 0 and np.all(np.diff(out.astype(np.int32)) > 0)  # residual 0 -> the (sorted) stage-1 vector


def test_nlsf2a_is_stable_for_wideband_vector():
    lib = helpers.oracle_lib()
    lib.orc_silk_inverse_pred_gain.restype = C.c_int32
    idx = _i8([5] + [1] * 16)
    nlsf = np.zeros(16, dtype=np.int16)
    a = np.zeros(16, dtype=np.int16)
    for q in (ALL, ALL & ~S2):
        lib.orc_silk_nlsf_decode(1, helpers.ptr(idx), helpers.ptr(nlsf), q)
        lib.orc_silk_nlsf2a(helpers.ptr(nlsf), 16, helpers.ptr(a), q)
        assert lib.orc_silk_inverse_pred_gain(helpers.ptr(a), 16) > 0


def test_up2_hq_known_answers():
    lib = helpers.oracle_lib()
    out = np.zeros(64, dtype=np.int16)
    st = np.zeros(6, dtype=np.int32)
    # 8 kHz -> 16 kHz uses the up2_hq kernel directly; the API wrapper delays the input by DELAY_MATRIX_DEC[0][2] = 2
    # samples, so feed [x, 0, 0] * and read the kernel state after the 4 real samples have gone through.
    x = _i16([1000, -1000, 2000, -2000] + [0] * 4)
    n = lib.orc_silk_resample(8000, 16000, helpers.ptr(x), 8, helpers.ptr(out), helpers.ptr(st))
    assert n == 16
    # output = two leading delay zeros (4 output samples) then the reference's expected sequence
    assert out[:4].tolist() == [0, 0, 0, 0]
    assert out[4:12].tolist() == [4, 35, 152, 381, 571, 423, -52, -236]


def _lres(mode, quirks):
    n = 65536
    i = np.arange(n)
    arena = np.zeros((n, 4), dtype=np.uint8)
    arena[:, 0] = mode << 3
    arena[:, 1] = i >> 8
    arena[:, 2] = i & 255
    arena[:, 3] = 255
    offsets = (np.arange(n + 1) * 4).astype(np.uint32)
    _, samples, rng = helpers.oracle_decode_grid(arena.reshape(-1), offsets, 1, n, 1, 480, i16=True, quirks=quirks, threads=1)
    assert np.all(samples == 480)
    return int(rng.astype(np.uint64).sum() & 0xFFFFFFFF)


@pytest.mark.parametrize("mode,want", [(0, 3285687739), (4, 1481572662), (8, 694350475)])
def test_lres_known_answer(mode, want):
    """tests/test_opus_decode.rs:344-370. NB and WB hold for mousiki as written; MB (120-sample frames, padded shell
    block) only with the sign-decode quirk S3 switched to libopus behaviour."""
    if mode == 4:
        assert _lres(mode, ALL & ~S3) == want
        assert _lres(mode, ALL) != want
    else:
        assert _lres(mode, ALL) == want


def _decode(lib, h, pkt, frame=960, ch=1):
    pcm = np.zeros(frame * ch, dtype=np.float32)
    pkt = np.ascontiguousarray(pkt)
    r = lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(pcm), frame, 0)
    return r, pcm, lib.orc_final_range(h)


def test_hybrid_transition_vectors():
    """tests/hybrid_decode.rs:55-124: SILK WB 20 ms then hybrid FB 20 ms on a mono 48 kHz decoder."""
    lib = helpers.oracle_lib()
    g = np.load(helpers.ROOT / "tests" / "golden" / "ref_hybrid_vectors.npz")
    errs = {}
    for q in (ALL, ALL & ~S2):
        h = lib.orc_create(48000, 1, q)
        r1, _, _ = _decode(lib, h, g["TRANSITION_SILK_PACKET"])
        r2, pcm, rng = _decode(lib, h, g["TRANSITION_HYBRID_PACKET"])
        lib.orc_destroy(h)
        assert r1 == 960 and r2 == 960
        assert rng == int(g["TRANSITION_HYBRID_RANGE"][0])  # entropy path: pinned for both settings
        errs[q] = float(np.abs(pcm - g["TRANSITION_HYBRID_PCM"]).max())
    assert errs[ALL & ~S2] <= 1e-4, errs  # the reference's tolerance, libopus NLSF2A ordering
    assert errs[ALL] > 1e-4, errs         # mousiki's nlsf2a.rs as written misses its own vector (documented)


def test_hybrid_encode_packets_decode():
    """Three real libopus hybrid packets (tests/fixtures/hybrid_encode_vectors.rs). Decoding them is a probe, not a
    pin (the ranges are asserted encoder-side by the reference); the decode must succeed with 960 samples."""
    lib = helpers.oracle_lib()
    g = np.load(helpers.ROOT / "tests" / "golden" / "ref_hybrid_vectors.npz")
    for name, ch in (("HP_DELAY_PACKET0", 1), ("HP_DELAY_PACKET1", 1), ("STEREO_WIDTH_PACKET", 2)):
        h = lib.orc_create(48000, 2, ALL)
        r, pcm, rng = _decode(lib, h, g[name], 960, 2)
        lib.orc_destroy(h)
        assert r == 960, (name, r)
        assert np.isfinite(pcm).all()
