"""Hybrid packets (SILK WB 16 kHz + CELT bands 17..): the product against the oracle and
against the reference's own hybrid decode vector.

CPU: host build of the kernel bodies (tests/emul). GPU (-m gpu): the CUDA path through the C ABI.
What is compared: sample counts and final_range bit for bit (one range decoder runs through the SILK payload, the
redundancy flag and the CELT bands); PCM bit for bit on the CPU emulation (same float expression order as the oracle)
and <= 1 LSB at int16 scale on the GPU (the tolerance BASELINE.json's north_star states for float PCM).
Corpus: `TOC || random bytes` (the reference's fuzz construction, tests/test_opus_decode.rs:372-407) restricted to
packets without a redundancy frame (helpers.silk_fuzz_grid); the reference's golden vector is
tests/hybrid_decode.rs:55-124 transcribed into tests/golden/ref_hybrid_vectors.npz (tools/gen_ref_vectors.py).

The FB 20 ms cases matter most: the reference's fold buffer (`norm_storage`, src/celt/types.rs:400,416) outlives the
frame and its last band is transformed in place (src/celt/bands.rs:2176-2200), so a hybrid frame depends on the fold
buffer the previous frames of the stream left behind."""
import numpy as np
import pytest

import helpers

S2 = 1 << 9
TOL = 1.0 / 32768.0


def hyb_toc(fb, ms20, stereo):
    """config 12/13 = SWB 10/20 ms, 14/15 = FB 10/20 ms (RFC 6716 table 2, src/packet.rs:30-78)."""
    return ((12 + 2 * int(fb) + int(ms20)) << 3) | (4 if stereo else 0)


# fullband, 20 ms, packet stereo, decoder channels, payload range, seed
CASES = [
    (1, 1, False, 1, (30, 160), 1),
    (1, 1, True, 2, (40, 200), 2),
    (0, 1, False, 2, (30, 120), 3),
    (1, 0, True, 1, (20, 100), 4),
    (0, 0, True, 2, (20, 100), 5),
    (1, 0, False, 1, (20, 80), 6),
]


def _corpus(fb, ms20, stereo, ch, payload, seed, lanes, ticks):
    return helpers.silk_fuzz_grid(lanes, ticks, payload, lambda k, r: hyb_toc(fb, ms20, stereo), seed=seed, channels=ch)


def _mixed_toc(k, rng):
    """SILK-only WB and hybrid packets drawn per packet; SILK -> hybrid resets the CELT half (opus_decoder.rs:744-751),
    hybrid -> SILK-only lets the CELT overlap fade out through a 2.5 ms silence frame added onto the first 120 samples
    (opus_decoder.rs:868-893)."""
    if rng.integers(0, 4) == 0:
        return helpers.silk_toc(2, int(rng.integers(0, 2)), bool(rng.integers(0, 2)))
    return hyb_toc(rng.integers(0, 2), rng.integers(0, 2), rng.integers(0, 2))


def _golden_grid():
    g = np.load(helpers.ROOT / "tests" / "golden" / "ref_hybrid_vectors.npz")
    pk = [g["TRANSITION_SILK_PACKET"].astype(np.uint8), g["TRANSITION_HYBRID_PACKET"].astype(np.uint8)]
    offsets = np.zeros(3, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in pk])
    return g, np.concatenate(pk), offsets


@pytest.mark.parametrize("fb,ms20,stereo,ch,payload,seed", CASES)
def test_emul_hybrid_matches_oracle(fb, ms20, stereo, ch, payload, seed):
    lanes, ticks = 6, 7
    arena, offsets = _corpus(fb, ms20, stereo, ch, payload, seed, lanes, ticks)
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=i16)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=i16)
        assert np.all(so > 0)
        assert np.array_equal(so, se)
        assert np.array_equal(ro, re)
        assert np.array_equal(po, pe)


def test_emul_hybrid_reference_vector():
    """SILK WB packet then hybrid FB packet on a mono decoder: final_range 1329918976, 960 samples within 1e-4 of the
    reference's PCM with quirk S2 off (libopus NLSF2A ordering; see test_silk_oracle_kat.test_hybrid_transition_vectors)."""
    g, arena, offsets = _golden_grid()
    for q, within in ((helpers.QUIRKS_MOUSIKI, False), (helpers.QUIRKS_MOUSIKI & ~S2, True)):
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, 1, 2, 1, 960, quirks=q)
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, 1, 2, 1, 960, quirks=q)
        assert se.tolist() == [960, 960]
        assert int(re[1]) == int(g["TRANSITION_HYBRID_RANGE"][0])
        assert np.array_equal(pe, po)
        err = float(np.abs(pe[1, :960] - g["TRANSITION_HYBRID_PCM"]).max())
        assert (err <= 1e-4) == within, (q, err)


def test_emul_hybrid_silk_switches_both_ways():
    lanes, ticks = 8, 10
    for ch in (1, 2):
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (30, 160), _mixed_toc, seed=60 + ch, channels=ch)
        tocs = arena[offsets[:-1]].reshape(ticks, lanes)
        hyb = (tocs & 0x60) == 0x60
        assert (hyb[:-1] & ~hyb[1:]).any() and (~hyb[:-1] & hyb[1:]).any(), "corpus must switch in both directions"
        for i16 in (False, True):
            pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=i16)
            po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=i16)
            assert (so > 0).all()
            assert np.array_equal(se, so) and np.array_equal(re, ro) and np.array_equal(pe, po)


def test_emul_40ms_silk_after_hybrid_decodes():
    """A 40 ms SILK-only packet right after a hybrid packet: the cell has no room for the fade-out record next to two SILK
    frames, so it goes to an auxiliary cell (mb_common.cuh). Bit for bit with the oracle, and so is the hybrid packet behind."""
    good, goff = helpers.silk_fuzz_grid(1, 2, (40, 160), lambda k, r: hyb_toc(1, 1, False), seed=5, channels=1)
    long_a, long_o = helpers.silk_fuzz_grid(1, 1, (60, 120), lambda k, r: helpers.silk_toc(2, 2, False), seed=6, channels=1)
    pk = [good[goff[0]:goff[1]], long_a[long_o[0]:long_o[1]], good[goff[1]:goff[2]]]
    offsets = np.zeros(4, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in pk])
    arena = np.concatenate(pk)
    for i16 in (False, True):
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, 1, 3, 1, 1920, i16=i16)
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, 1, 3, 1, 1920, i16=i16)
        assert se.tolist() == [960, 1920, 960] == so.tolist()
        assert np.array_equal(pe, po) and np.array_equal(re, ro)


# ---------------------------------------------------------------------------------------------- GPU
def _decode_gpu(arena, offsets, lanes, ticks, ch, frame_size, i16, **kw):
    from mousiki_b200 import BatchDecoder

    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    with BatchDecoder(lanes, 48000, ch, **kw) as dec:
        if i16:
            return dec.decode(arena, offsets, ids, frame_size)
        return dec.decode_float(arena, offsets, ids, frame_size)


@pytest.mark.gpu
@pytest.mark.parametrize("fb,ms20,stereo,ch,payload,seed", CASES)
def test_gpu_hybrid_matches_oracle(fb, ms20, stereo, ch, payload, seed):
    lanes, ticks = 70, 6
    arena, offsets = _corpus(fb, ms20, stereo, ch, payload, seed, lanes, ticks)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
    pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, ch, 960, False)
    assert np.array_equal(so, sg)
    assert np.array_equal(ro, rg), "final_range must be bit-exact"
    assert float(np.abs(po - pg).max()) <= TOL
    p16o, _, _ = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=True)
    p16g, s16, r16 = _decode_gpu(arena, offsets, lanes, ticks, ch, 960, True)
    assert np.array_equal(so, s16) and np.array_equal(ro, r16)
    assert int(np.abs(p16o.astype(np.int32) - p16g.astype(np.int32)).max()) <= 1


@pytest.mark.gpu
def test_gpu_hybrid_reference_vector():
    g, arena, offsets = _golden_grid()
    pg, sg, rg = _decode_gpu(arena, offsets, 1, 2, 1, 960, False, quirks=helpers.QUIRKS_MOUSIKI & ~S2)
    assert sg.tolist() == [960, 960]
    assert int(rg[1]) == int(g["TRANSITION_HYBRID_RANGE"][0])
    assert float(np.abs(pg[1, :960] - g["TRANSITION_HYBRID_PCM"]).max()) <= 1e-4


@pytest.mark.gpu
def test_gpu_hybrid_state_across_calls():
    """One packet per stream per call == all ticks in one call: SILK state, CELT state and the fold buffer all live
    in HBM between calls."""
    from mousiki_b200 import BatchDecoder

    lanes, ticks = 48, 6
    arena, offsets = _corpus(1, 1, True, 2, (40, 200), 71, lanes, ticks)
    pa, sa, ra = _decode_gpu(arena, offsets, lanes, ticks, 2, 960, False)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 960)
    assert np.array_equal(sa, so) and np.array_equal(ra, ro) and float(np.abs(pa - po).max()) <= TOL
    ids = np.arange(lanes, dtype=np.uint32)
    with BatchDecoder(lanes, 48000, 2) as dec:
        for t in range(ticks):
            a = arena[offsets[t * lanes]:offsets[(t + 1) * lanes]]
            o = (offsets[t * lanes:(t + 1) * lanes + 1] - offsets[t * lanes]).astype(np.uint32)
            p, s, r = dec.decode_float(a, o, ids, 960)
            assert np.array_equal(s, sa[t * lanes:(t + 1) * lanes])
            assert np.array_equal(r, ra[t * lanes:(t + 1) * lanes])
            assert np.array_equal(p[:, :1920], pa[t * lanes:(t + 1) * lanes, :1920]), "same kernels, same order: bit-identical"


@pytest.mark.gpu
def test_gpu_hybrid_silk_mix():
    lanes, ticks = 64, 10
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (30, 160), _mixed_toc, seed=81, channels=2)
    pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, 2, 960, False)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 960)
    assert (so > 0).all()
    assert np.array_equal(sg, so) and np.array_equal(rg, ro)
    assert float(np.abs(pg - po).max()) <= TOL
    p16g, s16, r16 = _decode_gpu(arena, offsets, lanes, ticks, 2, 960, True)
    p16o, _, _ = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 960, i16=True)
    assert np.array_equal(s16, so) and np.array_equal(r16, ro)
    assert int(np.abs(p16o.astype(np.int32) - p16g.astype(np.int32)).max()) <= 1
    # one packet per stream per call: the hint for the fade-out record then comes from the stream state
    from mousiki_b200 import BatchDecoder

    ids = np.arange(lanes, dtype=np.uint32)
    with BatchDecoder(lanes, 48000, 2) as dec:
        for t in range(ticks):
            a = arena[offsets[t * lanes]:offsets[(t + 1) * lanes]]
            o = (offsets[t * lanes:(t + 1) * lanes + 1] - offsets[t * lanes]).astype(np.uint32)
            p, s_, r = dec.decode_float(a, o, ids, 960)
            assert np.array_equal(s_, so[t * lanes:(t + 1) * lanes]) and np.array_equal(r, ro[t * lanes:(t + 1) * lanes])
            for j in range(lanes):
                m = int(s_[j]) * 2
                assert np.array_equal(p[j, :m], pg[t * lanes + j, :m])
