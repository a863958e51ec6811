"""Packet-loss concealment, in-band FEC (LBRR) and mode switches without redundancy on the ORACLE (SURVEY.md section 8
rows F1 / F2, oracle row (c)): the reference-held vectors nothing else pins —
  tests/hybrid_decode.rs:89-124   TRANSITION_HYBRID_PCM then HYBRID_PLC_PCM (decode(None) after a hybrid frame), 1e-4
  tests/hybrid_decode.rs:126-157  FEC_RANGE = 36081568 (decode_fec on a hybrid packet after a hybrid packet)
  tests/hybrid_decode.rs:159-186  FEC_PCM, 1e-4
and the API contract of tests/test_opus_decode.rs:135-308 (PLC sizes, error codes, every code-0 TOC followed by losses).
The three PCM vectors come from opus-c: like the transition vector they are met with quirk S2 off (libopus' LSF ordering
in NLSF2A) and missed by 4e-3 .. 4e-2 with mousiki's identity ordering; final_range does not depend on S2."""
import numpy as np
import pytest

import helpers

VEC = np.load(helpers.ROOT / "tests" / "golden" / "ref_hybrid_vectors.npz")
S2 = 0x200


def _dec(quirks, ch=1):
    lib = helpers.oracle_lib()
    return lib, lib.orc_create(48000, ch, quirks)


def _decode(lib, h, pkt, frame_size, fec=0, ch=1, i16=False):
    pcm = np.zeros(frame_size * ch, dtype=np.int16 if i16 else np.float32)
    fn = lib.orc_decode_i16 if i16 else lib.orc_decode_float
    r = fn(h, None if pkt is None else helpers.ptr(pkt), 0 if pkt is None else len(pkt), helpers.ptr(pcm), frame_size, fec)
    return r, pcm


@pytest.mark.parametrize("s2_on", [False, True])
def test_hybrid_plc_vector(s2_on):
    lib, h = _dec(helpers.QUIRKS_MOUSIKI if s2_on else helpers.QUIRKS_MOUSIKI & ~S2)
    for name in ("TRANSITION_SILK_PACKET", "TRANSITION_HYBRID_PACKET"):
        r, pcm = _decode(lib, h, VEC[name], 960)
        assert r == 960
    assert lib.orc_final_range(h) == int(VEC["TRANSITION_HYBRID_RANGE"][0])
    r, plc = _decode(lib, h, None, 960)
    assert r == 960 and lib.orc_final_range(h) == 0          # a concealed frame leaves final_range 0 (opus_decoder.rs:1123-1128)
    err = float(np.abs(plc - VEC["HYBRID_PLC_PCM"]).max())
    lib.orc_destroy(h)
    if s2_on:
        assert 1e-4 < err < 2e-2, err      # mousiki as written misses the opus-c vector
    else:
        assert err <= 1e-4, err


@pytest.mark.parametrize("s2_on", [False, True])
def test_hybrid_fec_vectors(s2_on):
    lib, h = _dec(helpers.QUIRKS_MOUSIKI if s2_on else helpers.QUIRKS_MOUSIKI & ~S2)
    assert _decode(lib, h, VEC["FEC_PREV_PACKET"], 960)[0] == 960
    r, pcm = _decode(lib, h, VEC["FEC_PACKET"], 960, fec=1)
    assert r == 960
    assert lib.orc_final_range(h) == int(VEC["FEC_RANGE"][0]) == 36081568
    err = float(np.abs(pcm - VEC["FEC_PCM"]).max())
    lib.orc_destroy(h)
    if s2_on:
        assert 1e-4 < err < 1e-1, err
    else:
        assert err <= 1e-4, err


@pytest.mark.parametrize("ch", [1, 2])
def test_plc_sizes_and_error_codes(ch):
    """tests/test_opus_decode.rs:135-218 at 48 kHz."""
    lib, h = _dec(helpers.QUIRKS_MOUSIKI, ch)
    for fec in (0, 1):
        assert _decode(lib, h, None, 120, fec, ch, i16=True)[0] == 120
        assert lib.orc_last_packet_duration(h) == 120
        assert _decode(lib, h, None, 122, fec, ch, i16=True)[0] == -1          # not a multiple of 2.5 ms: BadArgument
        empty = np.zeros(0, dtype=np.uint8)
        assert _decode(lib, h, empty, 120, fec, ch, i16=True)[0] == 120        # zero-length packet = loss
        assert _decode(lib, h, empty, 0, fec, ch, i16=True)[0] == -1           # zero frame size: BadArgument
        lib.orc_reset(h)
    lib.orc_destroy(h)


def test_every_code0_toc_then_losses():
    """tests/test_opus_decode.rs:220-308: every code-0 TOC x second byte decodes to the packet's sample count on mono and
    stereo decoders with the same final_range; six lost packets of that size, then a 2.5 ms one, conceal to the sizes asked
    for; a buffer one sample short is rejected."""
    lib = helpers.oracle_lib()
    decs = [lib.orc_create(48000, 1, helpers.QUIRKS_MOUSIKI), lib.orc_create(48000, 2, helpers.QUIRKS_MOUSIKI)]
    for i in range(64):
        toc = i << 2
        cfg = toc >> 3
        if cfg >= 16:
            expect = 120 << (cfg & 3)
        elif cfg >= 12:
            expect = 960 if cfg & 1 else 480
        else:
            expect = [480, 960, 1920, 2880][cfg & 3]
        for j in (0, 1, 37, 128, 255):
            pkt = np.array([toc, j, 255], dtype=np.uint8)
            rngs = []
            for t, h in enumerate(decs):
                r, _ = _decode(lib, h, pkt, 5760, 0, t + 1, i16=True)
                assert r == expect, (i, j, t, r)
                assert lib.orc_last_packet_duration(h) == expect
                rngs.append(lib.orc_final_range(h))
            assert rngs[0] == rngs[1]
        for t, h in enumerate(decs):
            for _ in range(6):
                assert _decode(lib, h, None, expect, 0, t + 1, i16=True)[0] == expect
                assert lib.orc_last_packet_duration(h) == expect
            if expect != 120:
                assert _decode(lib, h, None, 120, 0, t + 1, i16=True)[0] == 120
                assert lib.orc_last_packet_duration(h) == 120
            assert _decode(lib, h, np.array([toc, 255], dtype=np.uint8), expect - 1, 0, t + 1, i16=True)[0] < 0
    for h in decs:
        lib.orc_destroy(h)


def test_streams_with_losses_decode_without_errors():
    """tests/test_opus_decode.rs:372-407 style fuzz with losses and FEC requests mixed in: every call returns the size asked
    for or the packet's size, never an error, whatever the mode sequence (this is what turns the -5 arms into audio)."""
    lib = helpers.oracle_lib()
    rng = np.random.default_rng(5)
    for ch in (1, 2):
        h = lib.orc_create(48000, ch, helpers.QUIRKS_MOUSIKI)
        decoded = lost = fec_n = 0
        for k in range(600):
            kind = int(rng.integers(0, 10))
            toc = int(rng.integers(0, 64)) << 2 | 0
            pkt = np.concatenate([[toc], rng.integers(0, 256, size=int(rng.integers(2, 120)), dtype=np.uint8)]).astype(np.uint8)
            if kind < 6:
                r, _ = _decode(lib, h, pkt, 5760, 0, ch)
                if r == -4:      # random bytes can exceed the frame's bit budget: the reference reports InvalidPacket too
                    continue
                assert r > 0, (k, r, hex(toc))
                decoded += 1
            elif kind < 8:
                n = lib.orc_last_packet_duration(h) or 960
                r, _ = _decode(lib, h, None, n, 0, ch)
                assert r == n, (k, r, n)
                lost += 1
            else:
                n = 960
                r, _ = _decode(lib, h, pkt, n, 1, ch)
                assert r == n or r == -4 or r == -2, (k, r)
                fec_n += 1
        lib.orc_destroy(h)
        assert decoded > 200 and lost > 60 and fec_n > 60
