"""Row A2: the oracle's restatement of mousiki's lightweight decoder (oracle/lite_oracle.cpp) pinned on every vector the
reference holds for that path (tests/golden/lite_vectors.json, copied from the reference's own unit tests by
tools/gen_lite_golden.py): src/decoder.rs:275-325, src/silk/decoder.rs:2479-2860, :2925-2970, src/bitdepth.rs:101-130."""
import ctypes as C
import json
from pathlib import Path

import numpy as np
import pytest

import helpers

VEC = json.loads((Path(__file__).parent / "golden" / "lite_vectors.json").read_text())
WIDE, NS20 = 2, 20_000_000


def lite():
    lib = helpers.oracle_lib()
    lib.orl_create.restype = C.c_void_p
    lib.orl_destroy.argtypes = [C.c_void_p]
    for f in (lib.orl_decode_f32, lib.orl_decode_i16, lib.orl_decode_i24):
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.orl_silk_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint32, C.c_int, C.c_void_p]
    lib.orl_last_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.orl_lsf_coefficients.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_uint32, C.c_void_p]
    lib.orl_lsf_stabilize.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.orl_lsf_to_lpc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return lib


def u8(x):
    return np.asarray(x, dtype=np.uint8)


def last_frame(lib, h):
    ints = np.zeros(350, dtype=np.int32)
    fl = np.zeros(38, dtype=np.float32)
    lib.orl_last_frame(h, helpers.ptr(ints), helpers.ptr(fl))
    return ints, fl


def test_decode_float32_matches_the_reference_prefix():
    """src/decoder.rs:298-316: packet 48 0b e4 c1 36 ec c5 80 -> first sample of each of the first 12 triplets, 1e-6."""
    lib = lite()
    h = lib.orl_create()
    out = np.zeros(960, dtype=np.float32)
    pkt = u8(VEC["packet"])
    assert lib.orl_decode_f32(h, helpers.ptr(pkt), len(pkt), helpers.ptr(out)) == 960
    want = np.asarray(VEC["packet_prefix"], dtype=np.float32)
    assert np.abs(out[0:36:3] - want).max() < 1.0e-6
    assert np.array_equal(out[0::3], out[1::3]) and np.array_equal(out[0::3], out[2::3])   # src/resample.rs:3-15
    lib.orl_destroy(h)


def test_go_fixtures_two_consecutive_frames():
    """src/silk/decoder.rs:2732-2780: EXPECTED_DECODE_OUT_0 (fresh decoder) and _1 (the frame after it), 1e-6 each."""
    lib = lite()
    h = lib.orl_create()
    out = np.zeros(320, dtype=np.float32)
    f0, f1 = u8(VEC["silk_frame"]), u8(VEC["silk_frame_second"])
    assert lib.orl_silk_decode(h, helpers.ptr(f0), len(f0), 0, NS20, WIDE, helpers.ptr(out)) == 0
    assert np.abs(out - np.asarray(VEC["decode_out_0"], dtype=np.float32)).max() < 1.0e-6
    ints, fl = last_frame(lib, h)
    assert (ints[0], ints[1]) == (0, 1)                                  # :2479-2503 Inactive / High
    assert fl[:4].tolist() == VEC["gains_q16"]                            # :2505-2527
    assert lib.orl_silk_decode(h, helpers.ptr(f1), len(f1), 0, NS20, WIDE, helpers.ptr(out)) == 0
    assert np.abs(out - np.asarray(VEC["decode_out_1"], dtype=np.float32)).max() < 1.0e-6
    lib.orl_destroy(h)


def test_error_contract():
    """src/decoder.rs:285-296, src/silk/decoder.rs:2696-2716 and the LBRR refusal (:1013-1015)."""
    lib = lite()
    h = lib.orl_create()
    out = np.zeros(960, dtype=np.float32)
    assert lib.orl_decode_f32(h, None, 0, helpers.ptr(out)) == -1                      # TooShortForTableOfContents
    p = u8([0x49, 0x00])
    assert lib.orl_decode_f32(h, helpers.ptr(p), 2, helpers.ptr(out)) == -2            # UnsupportedFrameCode(1)
    p = u8([0x80, 0x00])
    assert lib.orl_decode_f32(h, helpers.ptr(p), 2, helpers.ptr(out)) == -3            # UnsupportedConfigurationMode
    p = u8(VEC["packet"]); p[0] = 0x4C
    assert lib.orl_decode_f32(h, helpers.ptr(p), len(p), helpers.ptr(out)) == -5       # stereo (src/decoder.rs STEREO_PACKET)
    p = u8(VEC["packet"]); p[0] = 0x40                                                 # WB 10 ms
    assert lib.orl_decode_f32(h, helpers.ptr(p), len(p), helpers.ptr(out)) == -4       # UnsupportedFrameDuration
    p = u8([0x48, 0x7f, 0xff, 0xff])                                                   # VAD 0, LBRR flag 1
    assert lib.orl_decode_f32(h, helpers.ptr(p), len(p), helpers.ptr(out)) == -7
    lib.orl_destroy(h)


def test_nlsf_and_lpc_unit_vectors():
    """src/silk/decoder.rs:2925-2970 (+ :2560-2587 stabilisation)."""
    lib = lite()
    res = np.asarray(VEC["res_q10"], dtype=np.int16)
    nlsf = np.zeros(16, dtype=np.int16)
    lib.orl_lsf_coefficients(16, WIDE, helpers.ptr(res), 9, helpers.ptr(nlsf))
    assert nlsf.tolist() == VEC["nlsf_q15"]
    for a, b in zip(VEC["stab_in"], VEC["stab_out"]):
        x = np.asarray(a, dtype=np.int16)
        lib.orl_lsf_stabilize(helpers.ptr(x), 16, WIDE)
        assert x.tolist() == b
    n = np.asarray(VEC["convert_nlsf_q15"], dtype=np.int16)
    a32 = np.zeros(16, dtype=np.int32)
    aq = np.zeros(16, dtype=np.float32)
    lib.orl_lsf_to_lpc(helpers.ptr(n), 16, WIDE, helpers.ptr(a32), helpers.ptr(aq))
    assert a32.tolist() == VEC["a32_q17"]
    assert aq.tolist() == VEC["aq12"]


def test_voiced_frame_pitch_and_ltp_vectors():
    """src/silk/decoder.rs:2609-2652: TEST_PITCH_LAG_FRAME parsed from its first bit reaches the pitch lags and LTP taps the
    reference's mid-stream range-decoder snapshots expect."""
    lib = lite()
    h = lib.orl_create()
    out = np.zeros(320, dtype=np.float32)
    f = u8(VEC["pitch_lag_frame"])
    assert lib.orl_silk_decode(h, helpers.ptr(f), len(f), 0, NS20, WIDE, helpers.ptr(out)) == 0
    ints, fl = last_frame(lib, h)
    assert ints[0] == 2 and ints[4] == 288
    assert ints[5:9].tolist() == VEC["pitch_lags"]
    assert ints[9:29].reshape(4, 5).tolist() == VEC["ltp_taps"]
    assert fl[4] == 15565.0                                               # :2665-2678
    lib.orl_destroy(h)


def test_excitation_go_fixture():
    """src/silk/decoder.rs:2806-2857: the 320 Q23 excitation values of TEST_LCG_FRAME (unvoiced, low offset, seed 0)."""
    lib = lite()
    h = lib.orl_create()
    out = np.zeros(320, dtype=np.float32)
    f = u8(VEC["lcg_frame"])
    assert lib.orl_silk_decode(h, helpers.ptr(f), len(f), 0, NS20, WIDE, helpers.ptr(out)) == 0
    ints, _ = last_frame(lib, h)
    assert ints[29] == 320
    assert ints[30:350].tolist() == VEC["excitation"]
    lib.orl_destroy(h)


def test_bitdepth_vectors():
    """src/bitdepth.rs:101-130 through the oracle's output stages: floor(x * 32767) and rint(x * 2^23), each held 3x."""
    lib = lite()
    h = lib.orl_create()
    pkt = u8(VEC["packet"])
    f = np.zeros(960, dtype=np.float32)
    lib.orl_decode_f32(h, helpers.ptr(pkt), len(pkt), helpers.ptr(f))
    lib.orl_destroy(h)
    h = lib.orl_create()
    i16 = np.zeros(960, dtype=np.int16)
    assert lib.orl_decode_i16(h, helpers.ptr(pkt), len(pkt), helpers.ptr(i16)) == 960
    lib.orl_destroy(h)
    assert np.array_equal(i16, np.floor(f.astype(np.float32) * np.float32(32767.0)).astype(np.int16))
    h = lib.orl_create()
    i24 = np.zeros(960, dtype=np.int32)
    assert lib.orl_decode_i24(h, helpers.ptr(pkt), len(pkt), helpers.ptr(i24)) == 960
    lib.orl_destroy(h)
    assert np.array_equal(i24, np.rint(f.astype(np.float64) * 8388608.0).astype(np.int32))
    x = np.asarray([0.3, 0.0, 0.55, 0.72, -0.05], dtype=np.float32)
    assert np.floor(x * np.float32(32767.0)).astype("<i2").tobytes() == bytes([0x66, 0x26, 0x00, 0x00, 0x65, 0x46, 0x28, 0x5c, 0x99, 0xf9])
