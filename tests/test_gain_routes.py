"""Decode gain and the output routes of the i16 API.

apply_decode_gain_and_soft_clip (src/opus_decoder.rs:1376-1418) scales EVERY packet by 2^(gain_q8 / (256 * 20 / 3.32..))
before the soft clipper: SILK-only, hybrid and CELT-only alike. The product lets SILK-only packets leave the SILK lane as
int16 directly on the i16 API — only when no decode gain applies and no frame of the packet shares a row with a CELT frame
(one route per packet, decided in k_energy); everything else goes through the float row and k_clip_i16, which also
carries the clipper's memory into a directly-written packet when the previous packet left one.

CPU: host emulation of the kernel bodies against the oracle, bit for bit. GPU (-m gpu): the CUDA path through the C ABI."""
import ctypes as C

import numpy as np
import pytest

import helpers
from test_hybrid_parity import hyb_toc
from test_mode_switch import _any_toc, _compare_until_first_rejection

TOL = 1.0 / 32768.0


def oracle_streams(arena, offsets, lanes, ticks, ch, cap, i16, gain_q8=0):
    """Like helpers.oracle_decode_grid with a decode gain (Decoder builder option, src/opus_decoder.rs:411-463)."""
    lib = helpers.oracle_lib()
    n = lanes * ticks
    pcm = np.zeros((n, cap * ch), dtype=np.int16 if i16 else np.float32)
    samples = np.zeros(n, dtype=np.int32)
    rng = np.zeros(n, dtype=np.uint32)
    for j in range(lanes):
        h = lib.orc_create(48000, ch, helpers.QUIRKS_MOUSIKI)
        lib.orc_set_gain(h, gain_q8)
        for t in range(ticks):
            p = t * lanes + j
            pkt = np.ascontiguousarray(arena[offsets[p]:offsets[p + 1]])
            fn = lib.orc_decode_i16 if i16 else lib.orc_decode_float
            samples[p] = fn(h, helpers.ptr(pkt), len(pkt), helpers.ptr(pcm[p]), cap, 0)
            rng[p] = lib.orc_final_range(h)
        lib.orc_destroy(h)
    return pcm, samples, rng


def emul_with_gain(arena, offsets, lanes, ticks, ch, cap, i16, gain_q8):
    lib = helpers.emul_lib()
    lib.emu_set_gain(gain_q8)
    try:
        return helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, cap, i16=i16)[:3]
    finally:
        lib.emu_set_gain(0)


CORPORA = [
    ("celt fb 20 ms stereo", lambda l, t, s: helpers.fuzz_grid(l, t, (30, 160), lambda a, b, r: helpers.celt_toc(3, True), seed=s), 2),
    ("silk wb 20 ms stereo", lambda l, t, s: helpers.silk_fuzz_grid(l, t, (20, 60), lambda k, r: helpers.silk_toc(2, 1, True), seed=s, channels=2), 2),
    ("silk nb 10 ms mono", lambda l, t, s: helpers.silk_fuzz_grid(l, t, (8, 40), lambda k, r: helpers.silk_toc(0, 0, False), seed=s, channels=1), 1),
    ("hybrid fb 20 ms stereo", lambda l, t, s: helpers.silk_fuzz_grid(l, t, (40, 160), lambda k, r: hyb_toc(1, 1, True), seed=s, channels=2), 2),
    ("silk with redundancy frames", lambda l, t, s: helpers.silk_fuzz_grid(l, t, (20, 120), lambda k, r: helpers.silk_toc(2, 1, False), seed=s, channels=1,
                                                                         redundancy=True), 1),
]


@pytest.mark.parametrize("name,build,ch", CORPORA, ids=[c[0] for c in CORPORA])
@pytest.mark.parametrize("gain_q8", [768, -2560])
def test_emul_decode_gain(name, build, ch, gain_q8):
    lanes, ticks = 6, 5
    arena, offsets = build(lanes, ticks, 31)
    for i16 in (False, True):
        po, so, ro = oracle_streams(arena, offsets, lanes, ticks, ch, 960, i16, gain_q8)
        pe, se, re = emul_with_gain(arena, offsets, lanes, ticks, ch, 960, i16, gain_q8)
        assert (so > 0).all()
        assert np.array_equal(so, se) and np.array_equal(ro, re)
        assert np.array_equal(po, pe), (name, i16)
    # the gain really was applied: +3 dB is a factor of 1.4125
    p0, _, _ = oracle_streams(arena, offsets, lanes, ticks, ch, 960, False, 0)
    pg, _, _ = oracle_streams(arena, offsets, lanes, ticks, ch, 960, False, gain_q8)
    k = np.abs(p0) > 1e-3
    ratio = np.median(np.abs(pg[k]) / np.abs(p0[k]))
    assert abs(ratio - 10 ** (gain_q8 / 256 / 20)) < 1e-3 * ratio


def test_emul_mixed_streams_i16_with_and_without_gain():
    """Streams that move between CELT-only, SILK-only (with and without redundancy frames) and hybrid: the i16 route is
    chosen per packet, the clipper memory crosses from float-row packets into directly written ones."""
    lanes, ticks = 12, 10
    checked = 0
    for gain in (0, 1280):
        for seed in (5, 6):
            arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (20, 140), _any_toc, seed=seed, channels=2, redundancy=True)
            po, so, ro = oracle_streams(arena, offsets, lanes, ticks, 2, 960, True, gain)
            pe, se, re = emul_with_gain(arena, offsets, lanes, ticks, 2, 960, True, gain)
            checked += _compare_until_first_rejection(so, se, ro, re, po, pe, lanes, ticks, 0.0)
    assert checked > 100


def test_code0_frame_longer_than_1275_bytes_is_invalid():
    """src/packet.rs:465: a code-0 packet whose frame exceeds 1275 bytes is OPUS_INVALID_PACKET (the oracle agrees); the
    device-side parser reports the same code instead of decoding it."""
    pkt = np.concatenate([[helpers.celt_toc(3, True)], helpers.lcg_bytes(3, 1276)]).astype(np.uint8)
    ok = np.concatenate([[helpers.celt_toc(3, True)], helpers.lcg_bytes(3, 1275)]).astype(np.uint8)
    arena = np.concatenate([pkt, ok])
    offsets = np.array([0, len(pkt), len(pkt) + len(ok)], dtype=np.uint32)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, 2, 1, 2, 960)
    pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, 2, 1, 2, 960)
    assert so.tolist() == [-4, 960] and se.tolist() == [-4, 960]
    assert np.array_equal(po[1], pe[1]) and ro[1] == re[1]


# ---------------------------------------------------------------------------------------------- GPU
def _gpu(arena, offsets, ids, n_streams, ch, cap, i16, **kw):
    from mousiki_b200 import BatchDecoder

    with BatchDecoder(n_streams, 48000, ch, **kw) as dec:
        return dec.decode(arena, offsets, ids, cap) if i16 else dec.decode_float(arena, offsets, ids, cap)


@pytest.mark.gpu
@pytest.mark.parametrize("name,build,ch", CORPORA, ids=[c[0] for c in CORPORA])
def test_gpu_decode_gain(name, build, ch):
    lanes, ticks = 40, 5
    arena, offsets = build(lanes, ticks, 41)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    for gain in (768, -2560):
        for i16 in (False, True):
            po, so, ro = oracle_streams(arena, offsets, lanes, ticks, ch, 960, i16, gain)
            pg, sg, rg = _gpu(arena, offsets, ids, lanes, ch, 960, i16, gain_q8=gain)
            assert np.array_equal(so, sg) and np.array_equal(ro, rg)
            if i16:
                assert int(np.abs(po.astype(np.int32) - pg.astype(np.int32)).max()) <= 1, (name, gain)
            else:
                assert float(np.abs(po - pg).max()) <= TOL * max(1.0, 10 ** (gain / 256 / 20)), (name, gain)


@pytest.mark.gpu
def test_gpu_mixed_streams_i16_with_and_without_gain(monkeypatch):
    lanes, ticks = 64, 10
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    checked = 0
    for gain in (0, 1280):
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (20, 140), _any_toc, seed=50 + gain, channels=2, redundancy=True)
        po, so, ro = oracle_streams(arena, offsets, lanes, ticks, 2, 960, True, gain)
        # several passes: the clipper of pass k runs beside pass k + 1 and the frame meta halves are reused (pass k + 2)
        pg, sg, rg = _gpu(arena, offsets, ids, lanes, 2, 960, True, gain_q8=gain, max_frames_per_pass=2 * lanes)
        checked += _compare_until_first_rejection(so, sg, ro, rg, po, pg, lanes, ticks, 1.0)
    assert checked > 250


@pytest.mark.gpu
def test_gpu_packets_whose_frames_take_different_routes():
    """A packet is routed as a whole: (1) a hybrid packet followed by a two-frame SILK-only packet — its first frame shares
    the row with the CELT fade-out frame, its second frame is plain SILK; (2) three-frame SILK packets whose frames may
    end in redundancy frames. i16 output, against the oracle decoding the original packets."""
    lanes = 32
    # (1) tick 0 hybrid SWB 20 ms mono, tick 1 code-1 SILK WB 2 x 20 ms
    ha, ho = helpers.silk_fuzz_grid(lanes, 1, (40, 120), lambda k, r: hyb_toc(0, 1, False), seed=71, channels=1)
    sa, so_ = helpers.silk_fuzz_grid(lanes, 2, 40, lambda k, r: helpers.silk_toc(2, 1, False), seed=72, channels=1)
    pkts = [ha[ho[j]:ho[j + 1]] for j in range(lanes)]
    for j in range(lanes):
        fr = [bytes(sa[so_[k * lanes + j] + 1:so_[k * lanes + j + 1]]) for k in range(2)]
        pkts.append(helpers.multiframe_packet(helpers.silk_toc(2, 1, False), fr, 1))
    offsets = np.zeros(len(pkts) + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in pkts])
    arena = np.concatenate(pkts).astype(np.uint8)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), 2)
    po, so, ro = oracle_streams(arena, offsets, lanes, 2, 1, 1920, True)
    pg, sg, rg = _gpu(arena, offsets, ids, lanes, 1, 1920, True)
    assert np.array_equal(so, sg) and (sg[lanes:] == 1920).all() and np.array_equal(ro, rg)
    assert int(np.abs(po.astype(np.int32) - pg.astype(np.int32)).max()) <= 1
    # (2) SILK WB 20 ms frames, some ending in redundancy frames, three per packet (VBR)
    sa, so_ = helpers.silk_fuzz_grid(lanes, 6, (25, 110), lambda k, r: helpers.silk_toc(2, 1, True), seed=73, channels=2, redundancy=True)
    pkts = []
    for t in range(2):
        for j in range(lanes):
            fr = [bytes(sa[so_[(3 * t + k) * lanes + j] + 1:so_[(3 * t + k) * lanes + j + 1]]) for k in range(3)]
            pkts.append(helpers.multiframe_packet(helpers.silk_toc(2, 1, True), fr, 3, vbr=True))
    offsets = np.zeros(len(pkts) + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in pkts])
    arena = np.concatenate(pkts).astype(np.uint8)
    po, so, ro = oracle_streams(arena, offsets, lanes, 2, 2, 2880, True)
    pg, sg, rg = _gpu(arena, offsets, ids, lanes, 2, 2880, True)
    both = (so > 0) & (sg > 0)
    assert both.sum() >= lanes, "most packets must decode on both sides"
    for j in range(lanes):          # per stream, up to the first packet either side rejects
        for t in range(2):
            p = t * lanes + j
            if so[p] < 0 or sg[p] < 0:
                assert sg[p] < 0
                break
            assert so[p] == sg[p] and ro[p] == rg[p]
            assert int(np.abs(po[p].astype(np.int32) - pg[p].astype(np.int32)).max()) <= 1, (j, t)


@pytest.mark.gpu
def test_gpu_code0_frame_longer_than_1275_bytes_is_invalid():
    pkt = np.concatenate([[helpers.celt_toc(3, True)], helpers.lcg_bytes(3, 1276)]).astype(np.uint8)
    ok = np.concatenate([[helpers.celt_toc(3, True)], helpers.lcg_bytes(3, 1275)]).astype(np.uint8)
    arena = np.concatenate([pkt, ok])
    offsets = np.array([0, len(pkt), len(pkt) + len(ok)], dtype=np.uint32)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, 2, 1, 2, 960)
    pg, sg, rg = _gpu(arena, offsets, np.arange(2, dtype=np.uint32), 2, 2, 960, False)
    assert sg.tolist() == [-4, 960] and rg[1] == ro[1] and float(np.abs(pg[1] - po[1]).max()) <= TOL
