"""Packet queries of the C-ABI (bd_packet_get_*) against the reference's own expectations: tests/packet_helpers.rs:10-165
(the libopus bit-trick formulas for bandwidth and frame duration over all 256 TOC bytes, frame counts for every
TOC x second byte, the sample-count cases and their error codes). Host-only: no GPU involved."""
import pytest

import mousiki_b200 as mb

SAMPLE_RATES = [8000, 12000, 16000, 24000, 48000]


def _reference_bandwidth(toc):  # tests/packet_helpers.rs:10-19
    bw = toc >> 4
    return 1101 + (((((bw & 7) * 9) & (63 - (bw & 8))) + 2 + 12 * (1 if bw & 8 else 0)) >> 4)


def _reference_fp3s(toc):  # tests/packet_helpers.rs:21-27 (frames per 3 seconds)
    f = toc >> 3
    return ((((3 - (f & 3)) * 13 & 119) + 9) >> 2) * ((int(f > 13) * (3 - int((f & 3) == 3))) + 1) * 25


def _err(fn, *a):
    with pytest.raises(mb.PacketError) as e:
        fn(*a)
    return e.value.code


def test_bandwidth_for_all_toc_values():
    for toc in range(256):
        assert mb.packet_get_bandwidth(bytes([toc])) == _reference_bandwidth(toc), hex(toc)
    assert _err(mb.packet_get_bandwidth, b"") == -1


def test_samples_per_frame_for_all_toc_values_and_rates():
    for toc in range(256):
        fp3s = _reference_fp3s(toc)
        assert fp3s != 0
        for rate in SAMPLE_RATES:
            assert mb.packet_get_samples_per_frame(bytes([toc]), rate) == rate * 3 // fp3s, (hex(toc), rate)
    assert _err(mb.packet_get_samples_per_frame, b"", 8000) == -1


def test_frame_count_cases():
    assert _err(mb.packet_get_nb_frames, bytes(2), 0) == -1
    assert _err(mb.packet_get_nb_frames, bytes(1), 2) == -1
    for toc in range(256):
        code = toc & 3
        if code == 3:
            assert _err(mb.packet_get_nb_frames, bytes([toc, 0]), 1) == -4
        else:
            assert mb.packet_get_nb_frames(bytes([toc, 0]), 1) == (1 if code == 0 else 2)
        for second in range(256):
            want = (second & 0x3F) if code == 3 else (1 if code == 0 else 2)
            assert mb.packet_get_nb_frames(bytes([toc, second]), 2) == want


def test_channels_and_mode():
    assert mb.packet_get_nb_channels(b"\x00") == 1 and mb.packet_get_nb_channels(b"\x04") == 2
    assert _err(mb.packet_get_nb_channels, b"") == -1
    assert (mb.packet_get_mode(b"\x00"), mb.packet_get_mode(b"\x60"), mb.packet_get_mode(b"\x80")) == (1000, 1001, 1002)
    assert _err(mb.packet_get_mode, b"") == -1


def test_sample_count_cases():
    p = bytes(2)
    assert [mb.packet_get_nb_samples(p, 1, r) for r in (48000, 96000, 32000, 8000)] == [480, 960, 320, 80]
    assert _err(mb.packet_get_nb_samples, bytes([3, 0]), 1, 24000) == -4
    p = bytes([(63 << 2) | 3, 63])
    assert _err(mb.packet_get_nb_samples, p, 0, 24000) == -1
    assert _err(mb.packet_get_nb_samples, p, 2, 48000) == -4          # 63 frames of 20 ms: more than 120 ms


def test_agrees_with_the_decoded_sample_counts_of_the_fixtures():
    """Every TOC: the query equals what the oracle decoder returns for a one-frame packet (code 0)."""
    import numpy as np

    import helpers
    lib = helpers.oracle_lib()
    pcm = np.zeros(2880 * 2, dtype=np.float32)
    rng = np.random.default_rng(3)
    for toc in range(0, 256, 4):
        pkt = rng.integers(0, 256, size=40, dtype=np.uint8)
        pkt[0] = toc
        h = lib.orc_create(48000, 2, helpers.QUIRKS_MOUSIKI)
        r = lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(pcm), 2880, 0)
        lib.orc_destroy(h)
        if r > 0:
            assert r == mb.packet_get_nb_samples(pkt.tobytes()), hex(toc)
