"""Multistream decode: `opus_multistream_decode / _decode_float` (src/opus_multistream.rs:717-1000) —
N substream packets per packet, the first N-1 in self-delimited framing (src/packet.rs:316-498, src/opus_decoder.rs:1139-1160),
coupled substreams decoded as stereo and the rest as mono, channels copied out by the mapping table.

CPU: the oracle's multistream wrapper against N single-stream oracle decoders and a hand-made channel mapping; the host-side framing
(`mousiki_b200.multistream.parse_packet`) against the oracle's parser on every framing code, valid and broken. GPU (-m gpu):
`MultistreamBatchDecoder` — the substreams as lanes of a stereo and a mono BatchDecoder, through the C ABI — against the oracle,
f32 and int16, with lost packets, packets of several frames, mixed modes, and the reference's error codes."""
import ctypes as C

import numpy as np
import pytest

import helpers
from mousiki_b200.multistream import PacketParseError, pack_multistream, parse_packet, regular_packet
from test_hybrid_parity import hyb_toc

SURROUND_51 = dict(channels=6, streams=4, coupled=2, mapping=[0, 4, 1, 2, 3, 5])     # Vorbis order 5.1 (RFC 7845 section 5.1.1.2)
LAYOUTS = [
    ("stereo", dict(channels=2, streams=1, coupled=1, mapping=[0, 1])),
    ("5.1", SURROUND_51),
    ("three monos, one output muted, one stream played twice", dict(channels=5, streams=3, coupled=0, mapping=[0, 255, 1, 2, 1])),
    ("mono + stereo swapped", dict(channels=3, streams=2, coupled=1, mapping=[2, 1, 0])),
]


def _ms_lib():
    lib = helpers.oracle_lib()
    lib.orc_ms_create.restype = C.c_void_p
    lib.orc_ms_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_uint32]
    lib.orc_ms_destroy.argtypes = [C.c_void_p]
    lib.orc_ms_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    lib.orc_ms_final_range.restype = C.c_uint32
    lib.orc_ms_final_range.argtypes = [C.c_void_p]
    return lib


def _sub_packets(layout, ticks, seed, kinds=("celt", "silk", "hybrid"), frame_ms=20):
    """[tick][stream] regular packets of one duration: substream s is stereo when coupled; the mode may differ per stream and tick"""
    rng = np.random.default_rng(seed)
    out = []
    for t in range(ticks):
        row = []
        for s in range(layout["streams"]):
            stereo = s < layout["coupled"]
            kind = kinds[int(rng.integers(0, len(kinds)))]
            if kind == "celt":
                toc, lo, hi = helpers.celt_toc(3 if frame_ms == 20 else 2, stereo), 20, 120
            elif kind == "silk":
                toc, lo, hi = helpers.silk_toc(int(rng.integers(0, 3)), 1 if frame_ms == 20 else 0, stereo), 20, 70
            else:
                toc, lo, hi = hyb_toc(int(rng.integers(0, 2)), 1 if frame_ms == 20 else 0, stereo), 40, 160
            a, o = helpers.silk_fuzz_grid(1, 1, (lo, hi), lambda k, r: toc, seed=int(rng.integers(1, 1 << 30)), channels=2 if stereo else 1)
            row.append(bytes(a[o[0]:o[1]]))
        out.append(row)
    return out


def _oracle_ms(layout, packets, frame_size, i16, fs=48000, fec=False):
    """packets: per tick bytes or None -> (pcm [ticks, frame_size * channels], samples, final ranges)"""
    lib = _ms_lib()
    mp = np.array(layout["mapping"], dtype=np.uint8)
    h = lib.orc_ms_create(fs, layout["channels"], layout["streams"], layout["coupled"], helpers.ptr(mp), helpers.QUIRKS_MOUSIKI)
    assert h
    pcm = np.zeros((len(packets), frame_size * layout["channels"]), dtype=np.int16 if i16 else np.float32)
    samples = np.zeros(len(packets), dtype=np.int32)
    rng = np.zeros(len(packets), dtype=np.uint32)
    for t, pkt in enumerate(packets):
        buf = np.frombuffer(pkt, dtype=np.uint8).copy() if pkt else None
        samples[t] = lib.orc_ms_decode(h, helpers.ptr(buf), len(pkt) if pkt else 0, None if i16 else helpers.ptr(pcm[t]),
                                       helpers.ptr(pcm[t]) if i16 else None, frame_size, int(fec))
        rng[t] = lib.orc_ms_final_range(h)
    lib.orc_ms_destroy(h)
    return pcm, samples, rng


@pytest.mark.parametrize("name,layout", LAYOUTS, ids=[l[0] for l in LAYOUTS])
def test_oracle_multistream_is_n_decoders_plus_a_channel_map(name, layout):
    ticks = 6
    subs = _sub_packets(layout, ticks, seed=5)
    lost = [False, False, True, False, True, False]
    packets = [None if lost[t] else pack_multistream(subs[t]) for t in range(ticks)]
    pcm, samples, rng = _oracle_ms(layout, packets, 960, False)
    assert (samples == 960).all()
    lib = helpers.oracle_lib()
    want = np.zeros_like(pcm).reshape(ticks, 960, layout["channels"])
    xr = np.zeros(ticks, dtype=np.uint32)
    for s in range(layout["streams"]):
        ch = 2 if s < layout["coupled"] else 1
        h = lib.orc_create(48000, ch, helpers.QUIRKS_MOUSIKI)
        out = np.zeros(960 * ch, dtype=np.float32)
        for t in range(ticks):
            p = None if lost[t] else np.frombuffer(subs[t][s], dtype=np.uint8).copy()
            assert lib.orc_decode_float(h, helpers.ptr(p), 0 if p is None else len(p), helpers.ptr(out), 960, 0) == 960
            xr[t] ^= lib.orc_final_range(h)
            o2 = out.reshape(960, ch)
            for c, m in enumerate(layout["mapping"]):
                if s < layout["coupled"] and m in (2 * s, 2 * s + 1):
                    want[t, :, c] = o2[:, m & 1]
                elif s >= layout["coupled"] and m == s + layout["coupled"]:
                    want[t, :, c] = o2[:, 0]
        lib.orc_destroy(h)
    assert np.array_equal(pcm.reshape(want.shape), want) and np.array_equal(rng, xr)


def test_host_framing_matches_the_oracle_parser():
    """Every framing code, regular and self-delimited, and broken packets: parse_packet raises where the oracle's multistream
    validation fails, and the re-framed substreams decode like the originals."""
    rng = np.random.default_rng(9)
    fa, fo = helpers.fuzz_grid(1, 6, 30, lambda t, j, r: helpers.celt_toc(1, False), seed=3)
    fr = [bytes(fa[fo[i] + 1:fo[i + 1]]) for i in range(6)]
    toc = helpers.celt_toc(1, False)
    regular = [
        bytes([toc]) + fr[0],
        bytes(helpers.multiframe_packet(toc, fr[:2], 1)),
        bytes(helpers.multiframe_packet(toc, [fr[0], fr[1][:17]], 2)),
        bytes(helpers.multiframe_packet(toc, fr[:4], 3)),
        bytes(helpers.multiframe_packet(toc, [fr[0], fr[1][:9], fr[2][:20]], 3, vbr=True, pad=5)),
    ]
    layout = dict(channels=2, streams=2, coupled=0, mapping=[0, 1])
    for first in regular:
        n = len(parse_packet(first, False)[1])
        last = bytes(helpers.multiframe_packet(toc, fr[:n], 3)) if n > 1 else bytes([toc]) + fr[5]
        ms = pack_multistream([first, last])
        t0, f0, off = parse_packet(ms, True)
        assert f0 == parse_packet(first, False)[1] and parse_packet(regular_packet(t0, f0), False)[1] == f0
        assert parse_packet(ms[off:], False)[1] == parse_packet(last, False)[1]
        _, samples, _ = _oracle_ms(layout, [ms], 120 * 2 * n, False)
        assert samples[0] == 240 * n
        # broken: cut anywhere, or a random byte flipped in the framing bytes
        for cut in range(1, min(len(ms), 12)):
            bad = ms[:cut]
            _, s_or, _ = _oracle_ms(layout, [bad], 5760, False)
            try:
                pos, ok = 0, len(bad) >= 3
                if ok:
                    _, _, o1 = parse_packet(bad, True)
                    ok = o1 < len(bad)
                    if ok:
                        parse_packet(bad[o1:], False)
            except PacketParseError:
                ok = False
            assert ok == (s_or[0] > 0 or s_or[0] == -2), (cut, s_or[0])


@pytest.mark.gpu
@pytest.mark.parametrize("name,layout", LAYOUTS, ids=[l[0] for l in LAYOUTS])
def test_gpu_multistream_decode(name, layout):
    from mousiki_b200.multistream import MultistreamBatchDecoder

    n_dec, ticks = 6, 7
    subs = [_sub_packets(layout, ticks, seed=11 + d) for d in range(n_dec)]
    lost = np.random.default_rng(3).random((n_dec, ticks)) < 0.2
    lost[:, 0] = False
    packets = [[None if lost[d, t] else pack_multistream(subs[d][t]) for t in range(ticks)] for d in range(n_dec)]
    for i16 in (False, True):
        want = [_oracle_ms(layout, packets[d], 960, i16) for d in range(n_dec)]
        with MultistreamBatchDecoder(n_dec, 48000, layout["channels"], layout["streams"], layout["coupled"], layout["mapping"]) as ms:
            for t in range(ticks):
                pcm, samples = ms.decode([packets[d][t] for d in range(n_dec)], 960) if i16 else ms.decode_float([packets[d][t] for d in range(n_dec)], 960)
                fr = ms.final_range()
                for d in range(n_dec):
                    assert samples[d] == want[d][1][t] == 960, (name, d, t)
                    assert np.array_equal(pcm[d], want[d][0][t]), (name, i16, d, t)
                    assert fr[d] == want[d][2][t]


@pytest.mark.gpu
def test_gpu_multistream_multiframe_substreams_other_rate_and_errors():
    from mousiki_b200.multistream import MultistreamBatchDecoder

    layout = SURROUND_51
    # substream packets of two 10 ms frames (code 1), decoded at 24 kHz
    ticks, fs = 4, 24000
    pk = []
    for t in range(ticks):
        a = _sub_packets(layout, 1, seed=31 + t, kinds=("celt", "silk"), frame_ms=10)[0]
        b = _sub_packets(layout, 1, seed=61 + t, kinds=("celt", "silk"), frame_ms=10)[0]
        row = []
        for s in range(layout["streams"]):
            toc = a[s][0]
            fa, fb = a[s][1:], b[s][1:] if b[s][0] == toc else a[s][1:]
            row.append(bytes(helpers.multiframe_packet(toc, [fa, fb], 2)))
        pk.append(pack_multistream(row))
    want_pcm, want_s, want_r = _oracle_ms(layout, pk, 480, False, fs=fs)
    assert (want_s == 480).all()
    with MultistreamBatchDecoder(1, fs, 6, 4, 2, layout["mapping"]) as ms:
        for t in range(ticks):
            pcm, samples = ms.decode_float([pk[t]], 480)
            assert samples[0] == 480 and np.array_equal(pcm[0], want_pcm[t]) and ms.final_range()[0] == want_r[t]
        # the reference's error codes; a refused packet leaves the decoders alone
        good = pk[0]
        short = good[:5]                                  # fewer than 2 * streams - 1 bytes
        mixed = pack_multistream([bytes([helpers.celt_toc(3, True)]) + bytes(30), bytes([helpers.celt_toc(2, True)]) + bytes(30),
                                  bytes([helpers.celt_toc(3, False)]) + bytes(30), bytes([helpers.celt_toc(3, False)]) + bytes(30)])
        for bad, code in ((short, -4), (mixed, -4), (good[:-200] if len(good) > 260 else good[:20], None)):
            _, s_or, _ = _oracle_ms(layout, [bad], 480, False, fs=fs)
            _, s_gpu = ms.decode_float([bad], 480)
            assert s_gpu[0] == s_or[0] and (code is None or s_or[0] == code), (s_gpu, s_or)
        _, s_gpu = ms.decode_float([good], 240)            # two 10 ms frames do not fit 240 samples at 24 kHz
        assert s_gpu[0] == -2


@pytest.mark.parametrize("ch,seed", [(2, 11), (2, 40), (1, 77)])
def test_emul_streams_that_switch_mode_and_lose_packets(ch, seed):
    """What the substreams of a multistream decoder look like to the batch decoder: every packet may change mode, any packet may be
    lost — among them a lost packet of a SILK-only stream that decoded CELT before (the row's zero fill is for streams that have
    decoded nothing, not for it) and a SILK-only packet right after a LOST packet of a hybrid stream (the parser cannot see the
    mode across the loss and makes room for the fade-out frame). Host emulation against the oracle, bit for bit."""
    layout = dict(channels=ch, streams=1, coupled=1 if ch == 2 else 0, mapping=list(range(ch)))
    lanes, ticks = 8, 9
    subs = [_sub_packets(layout, ticks, seed=seed + d) for d in range(lanes)]
    lost = np.random.default_rng(seed).random((lanes, ticks)) < 0.25
    lost[:, 0] = False
    pk, fl = [], []
    for t in range(ticks):
        for d in range(lanes):
            pk.append(np.array([960 & 255, 960 >> 8], dtype=np.uint8) if lost[d, t] else np.frombuffer(subs[d][t][0], dtype=np.uint8))
            fl.append(2 if lost[d, t] else 0)
    off = np.zeros(len(pk) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in pk])
    arena, flags = np.concatenate(pk).astype(np.uint8), np.array(fl, dtype=np.uint8)
    lib = helpers.emul_lib()
    lib.emu_set_flags(helpers.ptr(flags))
    try:
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, off, lanes, ticks, ch, 960)
    finally:
        lib.emu_set_flags(None)
    ol = helpers.oracle_lib()
    po, so, ro = np.zeros_like(pe), np.zeros_like(se), np.zeros_like(re)
    for d in range(lanes):
        h = ol.orc_create(48000, ch, helpers.QUIRKS_MOUSIKI)
        for t in range(ticks):
            p = t * lanes + d
            b = None if lost[d, t] else np.frombuffer(subs[d][t][0], dtype=np.uint8).copy()
            so[p] = ol.orc_decode_float(h, helpers.ptr(b), 0 if b is None else len(b), helpers.ptr(po[p]), 960, 0)
            ro[p] = ol.orc_final_range(h)
        ol.orc_destroy(h)
    assert (so == 960).all() and np.array_equal(so, se) and np.array_equal(ro, re)
    assert np.array_equal(po, pe)


def test_reference_multistream_known_answers():
    """The vectors the reference holds for this path (src/opus_multistream.rs:2548-2578): a two-stream packet whose first stream is
    self-delimited validates to 480 samples; a second stream of another duration is an invalid packet; a one-byte-payload SILK
    packet decodes through a stereo multistream decoder. On the oracle and on the host-side framing."""
    two = dict(channels=2, streams=2, coupled=0, mapping=[0, 1])
    ok, bad = bytes([0x00, 0x01, 0xAA, 0x00, 0xBB]), bytes([0x00, 0x01, 0xAA, 0x18, 0xBB])
    _, s, _ = _oracle_ms(two, [ok], 960, False)
    assert s[0] == 480
    _, s, _ = _oracle_ms(two, [bad], 2880, False)
    assert s[0] == -4
    toc, frames, off = parse_packet(ok, True)
    assert (toc, frames, off) == (0x00, [b"\xaa"], 3) and parse_packet(ok[off:], False)[1] == [b"\xbb"]

    from mousiki_b200.multistream import MultistreamBatchDecoder
    ms = MultistreamBatchDecoder.__new__(MultistreamBatchDecoder)      # (framing only: no device behind it)
    ms.fs, ms.streams = 48000, 2
    assert ms._validate(ok, 960)[1] == 0 and ms._validate(bad, 2880)[1] == -4 and ms._validate(ok, 240)[1] == -2
    stereo = dict(channels=2, streams=1, coupled=1, mapping=[0, 1])
    _, s, _ = _oracle_ms(stereo, [bytes([0x00, 0xAA])], 960, True)
    assert s[0] > 0


# ---------------------------------------------------------------------------------------------- projection (demixing matrix)
SIMPLE_MATRIX_DATA = [0, 32767, 0, 0, 32767, 0, 0, 0, 0, 0, 0, 32767]     # src/mapping_matrix.rs:370 (4 rows x 3 columns, column-major)
SIMPLE_MATRIX_INPUT = [32767, 0, -32768, 29491, -3277, -29491, 26214, -6554, -26214, 22938, -9830, -22938, 19661, -13107, -19661,
                       16384, -16384, -16384, 13107, -19661, -13107, 9830, -22938, -9830, 6554, -26214, -6554, 3277, -29491, -3277]
SIMPLE_MATRIX_EXPECTED = [0, 32767, 0, -32768, -3277, 29491, 0, -29491, -6554, 26214, 0, -26214, -9830, 22938, 0, -22938, -13107,
                          19661, 0, -19661, -16384, 16384, 0, -16384, -19661, 13107, 0, -13107, -22938, 9830, 0, -9830, -26214, 6554, 0,
                          -6554, -29491, 3277, 0, -3277]


def test_matrix_mixing_known_answer():
    """simple_matrix_operations_follow_reference (src/mapping_matrix.rs:415-500): the _out_short and _out_float forms, within the
    reference's own tolerance of 1 LSB — on the host-side mixer and on the oracle's."""
    from mousiki_b200.multistream import matrix_out_float, matrix_out_short
    m = np.array(SIMPLE_MATRIX_DATA, dtype=np.int16).reshape(3, 4).T.copy()
    x16 = np.array(SIMPLE_MATRIX_INPUT, dtype=np.int16).reshape(10, 3)
    want = np.array(SIMPLE_MATRIX_EXPECTED, dtype=np.int32).reshape(10, 4)
    assert np.abs(matrix_out_short(m, x16).astype(np.int32) - want).max() <= 1
    xf = x16.astype(np.float32) / np.float32(32768.0)
    yf = matrix_out_float(m, xf)
    y16 = np.clip(np.rint(yf * 32768.0), -32768, 32767).astype(np.int32)
    assert np.abs(y16 - want).max() <= 1
    lib = helpers.oracle_lib()
    md = np.array(SIMPLE_MATRIX_DATA, dtype=np.int16)
    of, o16 = np.zeros((10, 4), dtype=np.float32), np.zeros((10, 4), dtype=np.int16)
    lib.orc_matrix_out_float(helpers.ptr(md), 4, 3, helpers.ptr(xf), 3, helpers.ptr(of), 10)
    lib.orc_matrix_out_short(helpers.ptr(md), 4, 3, helpers.ptr(xf), 3, helpers.ptr(o16), 10)
    assert np.array_equal(of, yf) and np.array_equal(o16, matrix_out_short(m, x16))


def _oracle_proj(layout, matrix_i16, packets, frame_size, i16):
    lib = _ms_lib()
    lib.orc_proj_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    mp = np.arange(layout["channels"], dtype=np.uint8)
    h = lib.orc_ms_create(48000, layout["channels"], layout["streams"], layout["coupled"], helpers.ptr(mp), helpers.QUIRKS_MOUSIKI)
    assert h
    pcm = np.zeros((len(packets), frame_size * layout["channels"]), dtype=np.int16 if i16 else np.float32)
    samples = np.zeros(len(packets), dtype=np.int32)
    for t, pkt in enumerate(packets):
        buf = np.frombuffer(pkt, dtype=np.uint8).copy() if pkt else None
        samples[t] = lib.orc_proj_decode(h, helpers.ptr(matrix_i16), helpers.ptr(buf), len(pkt) if pkt else 0,
                                         None if i16 else helpers.ptr(pcm[t]), helpers.ptr(pcm[t]) if i16 else None, frame_size, 0)
    lib.orc_ms_destroy(h)
    return pcm, samples


@pytest.mark.gpu
def test_gpu_projection_decode():
    """First-order ambisonics shape: 4 output channels from 2 coupled streams (4 decoded channels), a dense random Q15 matrix."""
    from mousiki_b200.multistream import ProjectionBatchDecoder

    layout = dict(channels=4, streams=2, coupled=2)
    rng = np.random.default_rng(5)
    matrix = rng.integers(-20000, 20000, size=16).astype("<i2")     # column-major [rows = 4][cols = 4]
    n_dec, ticks = 3, 5
    subs = [_sub_packets(dict(streams=2, coupled=2), ticks, seed=71 + d) for d in range(n_dec)]
    packets = [[None if (t == 3 and d == 1) else pack_multistream(subs[d][t]) for t in range(ticks)] for d in range(n_dec)]
    for i16 in (False, True):
        want = [_oracle_proj(layout, matrix, packets[d], 960, i16) for d in range(n_dec)]
        with ProjectionBatchDecoder(n_dec, 48000, 4, 2, 2, matrix.tobytes()) as pd:
            for t in range(ticks):
                pcm, samples = (pd.decode if i16 else pd.decode_float)([packets[d][t] for d in range(n_dec)], 960)
                for d in range(n_dec):
                    assert samples[d] == want[d][1][t] == 960
                    assert np.array_equal(pcm[d], want[d][0][t]), (i16, d, t)


def test_emul_fec_on_a_silk_packet_right_after_a_hybrid_one():
    """decode(.., fec = true) taken on the hybrid -> SILK-only switch: the cell holds the LBRR-decoded SILK record AND the 2.5 ms
    CELT fade-out frame (opus_decoder.rs:868-893); the status word must keep saying "decoded from LBRR data" (stereo: the side
    channel's presence and the mid-only flag follow the LBRR rules, dec_api.rs:240-290). Then the same packet decoded normally."""
    lanes, ticks = 48, 4
    ha, ho = helpers.silk_fuzz_grid(lanes, 1, (40, 160), lambda k, r: hyb_toc(1, 1, bool(k & 1)), seed=5, channels=2)
    sa, so_ = helpers.silk_fuzz_grid(lanes, 2, (30, 90), lambda k, r: helpers.silk_toc(2, 1, True), seed=6, channels=2)
    pk, fl = [], []
    for j in range(lanes):
        pk.append(ha[ho[j]:ho[j + 1]]); fl.append(0)
    for j in range(lanes):
        pk.append(sa[so_[j]:so_[j + 1]]); fl.append(1)
    for j in range(lanes):
        pk.append(sa[so_[j]:so_[j + 1]]); fl.append(0)
    for j in range(lanes):
        pk.append(sa[so_[lanes + j]:so_[lanes + j + 1]]); fl.append(0)
    off = np.zeros(len(pk) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in pk])
    arena, ff = np.concatenate(pk).astype(np.uint8), np.array(fl, dtype=np.uint8)
    lib = helpers.emul_lib()
    lib.emu_set_flags(helpers.ptr(ff))
    try:
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, off, lanes, ticks, 2, 960)
    finally:
        lib.emu_set_flags(None)
    ol = helpers.oracle_lib()
    po, so, ro = np.zeros_like(pe), np.zeros_like(se), np.zeros_like(re)
    for j in range(lanes):
        h = ol.orc_create(48000, 2, helpers.QUIRKS_MOUSIKI)
        for t in range(ticks):
            p = t * lanes + j
            b = np.ascontiguousarray(arena[off[p]:off[p + 1]])
            so[p] = ol.orc_decode_float(h, helpers.ptr(b), len(b), helpers.ptr(po[p]), 960, int(ff[p] & 1))
            ro[p] = ol.orc_final_range(h)
        ol.orc_destroy(h)
    assert (so == 960).all() and np.array_equal(so, se) and np.array_equal(ro, re)
    assert np.array_equal(po, pe)
