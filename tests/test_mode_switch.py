"""Mode switches with redundancy: SILK-only packets that end in a 5 ms CELT
redundancy frame (opus_decoder.rs:746-797, 1012-1080, smooth_fade :100-135), which is how an Opus stream moves between
CELT-only and SILK-only without concealment: CELT -> SILK plays the redundancy frame in front of the SILK audio and
cross-fades, SILK -> CELT resets the CELT decoder, decodes the frame and cross-fades the end of the row into it; the
next CELT-only packet then needs no transition. final_range = SILK range ^ redundancy range.

Corpus: `TOC || random bytes` with redundancy-carrying packets kept (in SILK-only mode every packet with >= 17 spare
bits carries one). CPU: host emulation of the kernel bodies against the oracle, bit for bit. GPU (-m gpu): the CUDA path
through the C ABI. Switches that still need concealment (no redundancy) are reported as -5 by both."""
import os

import numpy as np
import pytest

import helpers
from test_hybrid_parity import hyb_toc

TOL = 1.0 / 32768.0


def _n_redundant(arena, offsets, ch):
    lib = helpers.oracle_lib()
    n = 0
    pcm = np.zeros(2880 * 2, dtype=np.float32)
    for p in range(len(offsets) - 1):
        pkt = arena[offsets[p]:offsets[p + 1]]
        h = lib.orc_create(48000, ch, helpers.QUIRKS_MOUSIKI)
        lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(pcm), 2880, 0)
        n += int(lib.orc_last_redundancy(h))
        lib.orc_destroy(h)
    return n


SILK_CASES = [
    ("wb 20 ms mono", lambda k, r: helpers.silk_toc(2, 1, False), (20, 120), 1, 1),
    ("wb 20 ms stereo", lambda k, r: helpers.silk_toc(2, 1, True), (30, 150), 2, 2),
    ("nb 10 ms mixed", lambda k, r: helpers.silk_toc(0, 0, bool(r.integers(0, 2))), (8, 80), 2, 3),
    ("mb 20 ms stereo into mono", lambda k, r: helpers.silk_toc(1, 1, True), (20, 100), 1, 4),
]


@pytest.mark.parametrize("name,toc,payload,ch,seed", SILK_CASES, ids=[c[0] for c in SILK_CASES])
def test_emul_silk_streams_with_redundancy_frames(name, toc, payload, ch, seed):
    lanes, ticks = 8, 7
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, toc, seed=seed, channels=ch, redundancy=True)
    assert _n_redundant(arena, offsets, ch) >= 5, "corpus must carry redundancy frames"
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=i16)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=i16)
        assert (so > 0).all()
        assert np.array_equal(so, se) and np.array_equal(ro, re) and np.array_equal(po, pe)


def _any_toc(k, r):
    m = int(r.integers(0, 10))
    if m < 4:
        return helpers.celt_toc(int(r.integers(1, 4)), bool(r.integers(0, 2)), int(r.integers(0, 4)))
    if m < 9:
        return helpers.silk_toc(int(r.integers(0, 3)), int(r.integers(0, 2)), bool(r.integers(0, 2)))
    return hyb_toc(r.integers(0, 2), r.integers(0, 2), r.integers(0, 2))


def _compare_until_first_rejection(so, se, ro, re, po, pe, lanes, ticks, tol):
    """Per stream: everything equal up to the first packet either side rejects. The oracle rejects exactly the switches
    that need concealment (and may have touched its SILK state by then, opus_decoder.rs:694-809); the product rejects
    those plus the combinations it has no room for (hybrid packets with redundancy, redundancy right after hybrid)."""
    so2, se2 = so.reshape(ticks, lanes), se.reshape(ticks, lanes)
    checked = switches = 0
    for j in range(lanes):
        for t in range(ticks):
            f = t * lanes + j
            if so2[t, j] < 0 or se2[t, j] < 0:
                assert se2[t, j] < 0, "the product decoded a packet the reference path rejects"
                break
            assert so2[t, j] == se2[t, j] and ro[f] == re[f], (j, t)
            err = float(np.abs(po[f].astype(np.float64) - pe[f].astype(np.float64)).max())
            assert err <= tol, (j, t, err)
            checked += 1
    return checked


def test_emul_streams_switching_between_celt_silk_and_hybrid():
    lanes, ticks = 16, 12
    total = 0
    for ch in (1, 2):
        for seed in (1, 2, 3):
            arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (20, 140), _any_toc, seed=seed, channels=ch, redundancy=True)
            po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
            pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 960)
            total += _compare_until_first_rejection(so, se, ro, re, po, pe, lanes, ticks, 0.0)
    assert total > 200


def test_emul_parallel_band_pass_is_exact_for_celt_only_streams():
    """Small calls run CELT-only frames through the all-at-once band pass (MB_EMUL_PARALLEL_BANDS selects it in the
    emulation; the default there is the per-stream pass of large calls)."""
    arena, offsets = helpers.fuzz_grid(6, 5, (20, 200), lambda t, j, r: helpers.celt_toc(int(r.integers(0, 4)), bool(r.integers(0, 2))), seed=3)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, 6, 5, 2, 960)
    os.environ["MB_EMUL_PARALLEL_BANDS"] = "1"
    try:
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, 6, 5, 2, 960)
    finally:
        del os.environ["MB_EMUL_PARALLEL_BANDS"]
    assert np.array_equal(so, se) and np.array_equal(ro, re) and np.array_equal(po, pe)


# ---------------------------------------------------------------------------------------------- GPU
def _decode_gpu(arena, offsets, lanes, ticks, ch, frame_size, i16, **kw):
    from mousiki_b200 import BatchDecoder

    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    with BatchDecoder(lanes, 48000, ch, **kw) as dec:
        if i16:
            return dec.decode(arena, offsets, ids, frame_size)
        return dec.decode_float(arena, offsets, ids, frame_size)


@pytest.mark.gpu
@pytest.mark.parametrize("name,toc,payload,ch,seed", SILK_CASES, ids=[c[0] for c in SILK_CASES])
def test_gpu_silk_streams_with_redundancy_frames(name, toc, payload, ch, seed):
    lanes, ticks = 48, 6
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, toc, seed=seed + 10, channels=ch, redundancy=True)
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=i16)
        pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, ch, 960, i16)
        assert np.array_equal(so, sg) and np.array_equal(ro, rg)
        if i16:
            assert int(np.abs(po.astype(np.int32) - pg.astype(np.int32)).max()) <= 1
        else:
            assert float(np.abs(po - pg).max()) <= TOL


@pytest.mark.gpu
def test_gpu_streams_switching_between_celt_silk_and_hybrid(monkeypatch):
    """The per-stream band pass (what large calls use; forced here for a small one) keeps the fold buffer exact across
    CELT-only, redundancy and hybrid frames of one stream."""
    lanes, ticks = 64, 12
    total = 0
    for ch in (1, 2):
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (20, 140), _any_toc, seed=20 + ch, channels=ch, redundancy=True)
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
        pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, ch, 960, False)
        total += _compare_until_first_rejection(so, sg, ro, rg, po, pg, lanes, ticks, TOL)
    assert total > 300


# ---------------------------------------------------------------------------------------------- hybrid packets with redundancy
def _hyb_red_streams(lanes, ticks, ch, seed, with_celt):
    """Streams that mix hybrid packets carrying a redundancy frame (tests/golden/hybrid_redundancy_packets.npz, found by
    tools/gen_hybrid_red_fixture.py: the flag has probability 2^-12 in random bytes) with plain hybrid, SILK-only WB and, if
    asked, CELT-only packets. Returns (arena, offsets, is_red[ticks * lanes])."""
    g = np.load(helpers.ROOT / "tests" / "golden" / "hybrid_redundancy_packets.npz")
    ra, ro = g["arena"], g["offsets"]
    n_red = len(ro) - 1
    rng = np.random.default_rng(seed)
    n_plain = lanes * ticks

    def plain_toc(k, r):
        if r.integers(0, 3):
            return hyb_toc(r.integers(0, 2), r.integers(0, 2), r.integers(0, 2))
        return helpers.silk_toc(2, int(r.integers(0, 2)), bool(r.integers(0, 2)))

    pa, po = helpers.silk_fuzz_grid(n_plain, 1, (30, 140), plain_toc, seed=seed + 1, channels=ch)
    ca, co = helpers.fuzz_grid(n_plain, 1, (30, 140), lambda t, j, r: helpers.celt_toc(int(r.integers(1, 4)), bool(r.integers(0, 2))), seed=seed + 2)
    pk, red = [], []
    for t in range(ticks):
        for j in range(lanes):
            k = t * lanes + j
            u = int(rng.integers(0, 10))
            if t > 0 and u < 4:
                i = int(rng.integers(0, n_red))
                pk.append(ra[ro[i]:ro[i + 1]]); red.append(True)
            elif with_celt and u < 6:
                pk.append(ca[co[k]:co[k + 1]]); red.append(False)
            else:
                pk.append(pa[po[k]:po[k + 1]]); red.append(False)
    off = np.zeros(len(pk) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in pk])
    return np.concatenate(pk).astype(np.uint8), off, np.array(red)


def _first_diff(po, pe, lanes, red):
    bad = np.nonzero((po != pe).any(axis=1))[0]
    p = int(bad[0])
    d = np.nonzero(po[p] != pe[p])[0]
    return (f"first differing packet: tick {p // lanes} lane {p % lanes} red={red[p]}, first sample {d[0]} of {len(d)}, "
            f"max|d|={np.abs(po[p].astype(np.float64) - pe[p]).max()} ({len(bad)} differ: {bad[:12]})")


@pytest.mark.parametrize("ch,seed", [(1, 11), (2, 12), (2, 13)])
def test_emul_hybrid_packets_with_redundancy_frames(ch, seed):
    """A hybrid packet may end in a 5 ms CELT redundancy frame (opus_decoder.rs:746-797): two CELT frames in one packet, the
    second one in an auxiliary cell (mb_common.cuh). CELT -> SILK frames run before the packet's own CELT part on the old
    state, SILK -> CELT ones after a decoder reset; the next packet sees prev_redundancy. Streams of hybrid and SILK-only WB
    packets: every packet decodes, bit for bit."""
    lanes, ticks = 8, 9
    arena, offsets, red = _hyb_red_streams(lanes, ticks, ch, seed, with_celt=False)
    assert red.sum() >= 15
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=i16)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=i16)
        assert (so > 0).all()
        assert np.array_equal(so, se), (so.reshape(ticks, lanes), se.reshape(ticks, lanes))
        assert np.array_equal(ro, re), np.nonzero(ro != re)
        assert np.array_equal(po, pe), _first_diff(po, pe, lanes, red)


def test_emul_hybrid_redundancy_next_to_celt_only_packets():
    """The same with CELT-only packets in the mix: the switches a redundancy frame covers decode; a stream is followed up to the
    first switch the product still reports as unimplemented (those that conceal a frame of the old mode)."""
    lanes, ticks = 16, 10
    total = red_checked = 0
    for ch, seed in ((1, 21), (2, 22)):
        arena, offsets, red = _hyb_red_streams(lanes, ticks, ch, seed, with_celt=True)
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 960)
        total += _compare_until_first_rejection(so, se, ro, re, po, pe, lanes, ticks, 0.0)
        se2 = se.reshape(ticks, lanes)
        for j in range(lanes):
            for t in range(ticks):
                if se2[t, j] < 0:
                    break
                red_checked += int(red[t * lanes + j])
    assert total > 100 and red_checked >= 10, (total, red_checked)


@pytest.mark.gpu
@pytest.mark.parametrize("ch,seed", [(1, 31), (2, 32)])
def test_gpu_hybrid_packets_with_redundancy_frames(ch, seed):
    """Through the C ABI, in passes of 3 ticks (the auxiliary pool is per pass)."""
    lanes, ticks = 48, 9
    arena, offsets, red = _hyb_red_streams(lanes, ticks, ch, seed, with_celt=False)
    assert red.sum() >= 100
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=i16)
        pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, ch, 960, i16, max_frames_per_pass=3 * lanes)
        assert (so > 0).all() and np.array_equal(so, sg) and np.array_equal(ro, rg)
        if i16:
            assert int(np.abs(po.astype(np.int32) - pg.astype(np.int32)).max()) <= 1
        else:
            assert float(np.abs(po - pg).max()) <= TOL


@pytest.mark.gpu
def test_gpu_hybrid_redundancy_next_to_celt_only_packets(monkeypatch):
    lanes, ticks = 64, 10
    total = 0
    for ch, seed in ((1, 41), (2, 42)):
        arena, offsets, red = _hyb_red_streams(lanes, ticks, ch, seed, with_celt=True)
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
        pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, ch, 960, False)
        total += _compare_until_first_rejection(so, sg, ro, rg, po, pg, lanes, ticks, TOL)
    assert total > 300


# ---------------------------------------------------------------------------------------------- redundancy frames in auxiliary cells
LONG_SILK_CASES = [
    ("wb 40 ms mono", lambda k, r: helpers.silk_toc(2, 2, False), (40, 200), 1, 51),
    ("wb 60 ms stereo", lambda k, r: helpers.silk_toc(2, 3, True), (200, 500), 2, 52),
    ("nb/mb 10-60 ms mixed", lambda k, r: helpers.silk_toc(int(r.integers(0, 2)), int(r.integers(0, 4)), bool(r.integers(0, 2))), (10, 200), 2, 53),
]


@pytest.mark.parametrize("name,toc,payload,ch,seed", LONG_SILK_CASES, ids=[c[0] for c in LONG_SILK_CASES])
def test_emul_40_and_60_ms_silk_packets_with_redundancy_frames(name, toc, payload, ch, seed):
    lanes, ticks = 8, 6
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, toc, seed=seed, channels=ch, redundancy=True)
    assert _n_redundant(arena, offsets, ch) >= 5, "corpus must carry redundancy frames"
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 2880, i16=i16)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 2880, i16=i16)
        assert (so > 0).all()
        assert np.array_equal(so, se) and np.array_equal(ro, re)
        assert np.array_equal(po, pe), _first_diff(po, pe, lanes, np.zeros(len(so), dtype=bool))


def _silk_red_or_hybrid(k, r):
    """SILK-only WB packets (most of them with a redundancy frame: any packet with 17 spare bits has one) next to hybrid ones:
    SILK-only right after hybrid needs the fade-out frame AND the redundancy frame."""
    if r.integers(0, 2):
        return helpers.silk_toc(2, int(r.integers(0, 4)), bool(r.integers(0, 2)))
    return hyb_toc(r.integers(0, 2), r.integers(0, 2), r.integers(0, 2))


def test_emul_silk_redundancy_right_after_hybrid():
    lanes, ticks = 12, 10
    total = 0
    for ch, seed in ((1, 61), (2, 62)):
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (30, 200), _silk_red_or_hybrid, seed=seed, channels=ch, redundancy=True)
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 2880)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 2880)
        total += _compare_until_first_rejection(so, se, ro, re, po, pe, lanes, ticks, 0.0)
        print(ch, "rejected by the product:", int((se < 0).sum()), "of", len(se))
    assert total > 200


@pytest.mark.gpu
@pytest.mark.parametrize("name,toc,payload,ch,seed", LONG_SILK_CASES, ids=[c[0] for c in LONG_SILK_CASES])
def test_gpu_40_and_60_ms_silk_packets_with_redundancy_frames(name, toc, payload, ch, seed):
    lanes, ticks = 40, 6
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, toc, seed=seed + 100, channels=ch, redundancy=True)
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 2880, i16=i16)
        pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, ch, 2880, i16, max_frames_per_pass=4 * lanes)
        assert (so > 0).all() and np.array_equal(so, sg) and np.array_equal(ro, rg)
        if i16:
            assert int(np.abs(po.astype(np.int32) - pg.astype(np.int32)).max()) <= 1
        else:
            assert float(np.abs(po - pg).max()) <= TOL


@pytest.mark.gpu
def test_gpu_silk_redundancy_right_after_hybrid(monkeypatch):
    lanes, ticks = 48, 10
    total = 0
    for ch, seed in ((1, 161), (2, 162)):
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (30, 200), _silk_red_or_hybrid, seed=seed, channels=ch, redundancy=True)
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 2880)
        pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, ch, 2880, False, max_frames_per_pass=3 * lanes)
        assert (sg > 0).all()
        total += _compare_until_first_rejection(so, sg, ro, rg, po, pg, lanes, ticks, TOL)
    assert total == 2 * lanes * ticks


# ---------------------------------------------------------------------------------------------- SILK / hybrid -> CELT-only, no redundancy
def _old_mode_then_celt(lanes, ticks_old, old, ch, seed):
    """`ticks_old` packets of the old mode (SILK-only of random size / bandwidth, or hybrid), then CELT-only packets of every frame
    size, none of them with a redundancy frame: the first CELT packet opens with the cross-fade from 5 ms (2.5 ms for a 2.5 ms
    frame) concealed in the old mode (opus_decoder.rs:575-593, 1083-1112)."""
    n = [0]

    def toc(k, r):
        t = n[0] // lanes
        n[0] += 1
        if t < ticks_old:
            if old == "silk":
                return helpers.silk_toc(int(r.integers(0, 3)), int(r.integers(0, 2)), bool(r.integers(0, 2)))
            return hyb_toc(r.integers(0, 2), r.integers(0, 2), r.integers(0, 2))
        return helpers.celt_toc(int(r.integers(0, 4)), bool(r.integers(0, 2)), int(r.integers(0, 4)))
    return helpers.silk_fuzz_grid(lanes, ticks_old + 2, (25, 140), toc, seed=seed, channels=ch, redundancy=False)


@pytest.mark.parametrize("old,ch,seed", [("silk", 1, 41), ("silk", 2, 42), ("hybrid", 1, 43), ("hybrid", 2, 44)])
def test_emul_switch_to_celt_without_redundancy(old, ch, seed):
    lanes = 24
    arena, offsets = _old_mode_then_celt(lanes, 3, old, ch, seed)
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, 5, ch, 960, i16=i16)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, 5, ch, 960, i16=i16)
        assert (so > 0).all() and np.array_equal(so, se) and np.array_equal(ro, re)
        assert np.array_equal(po, pe), _first_diff(po, pe, lanes, np.ones(len(so), dtype=bool))


def test_emul_every_mode_switch_decodes():
    """Random walks over CELT-only, SILK-only and hybrid packets of every size, with and without redundancy frames: nothing is
    refused any more, and everything is bit-identical to the oracle."""
    lanes, ticks = 16, 12
    for ch in (1, 2):
        for seed in (1, 2, 3, 4):
            arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (20, 140), _any_toc, seed=seed, channels=ch, redundancy=True)
            po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
            pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 960)
            assert (so > 0).all() and np.array_equal(so, se) and np.array_equal(ro, re)
            assert np.array_equal(po, pe)


@pytest.mark.gpu
@pytest.mark.parametrize("old,ch,seed", [("silk", 1, 51), ("silk", 2, 52), ("hybrid", 1, 53), ("hybrid", 2, 54)])
def test_gpu_switch_to_celt_without_redundancy(old, ch, seed):
    lanes = 96
    arena, offsets = _old_mode_then_celt(lanes, 3, old, ch, seed)
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, 5, ch, 960, i16=i16)
        pg, sg, rg = _decode_gpu(arena, offsets, lanes, 5, ch, 960, i16)
        assert (so > 0).all() and np.array_equal(so, sg) and np.array_equal(ro, rg)
        err = np.abs(po.astype(np.float64) - pg.astype(np.float64)).max()
        assert err <= (1 if i16 else TOL), err


@pytest.mark.gpu
def test_gpu_every_mode_switch_decodes():
    lanes, ticks = 128, 12
    for ch in (1, 2):
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (20, 140), _any_toc, seed=60 + ch, channels=ch, redundancy=True)
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
        pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, ch, 960, False)
        assert (so > 0).all() and np.array_equal(so, sg) and np.array_equal(ro, rg)
        assert np.abs(po.astype(np.float64) - pg.astype(np.float64)).max() <= TOL
