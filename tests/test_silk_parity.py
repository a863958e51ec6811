"""SILK-only packets (BASELINE config 3): the product's SILK path against the oracle.

CPU: the host build of the kernel bodies (tests/emul). GPU (-m gpu): the CUDA path through the C ABI. SILK is integer
work end to end, so everything is compared bit for bit: final_range, sample counts, int16 PCM, and the float PCM
(= int16 * 2^-15). Corpus: the reference's fuzz construction `TOC || random bytes` (tests/test_opus_decode.rs:372-407)
restricted to packets without a redundancy frame (helpers.silk_fuzz_grid)."""
import numpy as np
import pytest

import helpers

# bandwidth (0 NB, 1 MB, 2 WB), size code (0: 10 ms .. 3: 60 ms), packet stereo, decoder channels, payload bytes, seed
CASES = [
    (2, 1, False, 1, (8, 40), 1),     # BASELINE config 3: WB mono 20 ms
    (0, 0, False, 1, (3, 20), 2),
    (1, 0, True, 2, (6, 30), 3),      # MB 10 ms: 120-sample frames, padded shell block (quirk S3)
    (2, 3, True, 2, (30, 150), 4),    # 60 ms stereo: three SILK frames per packet, conditional coding
    (1, 2, False, 2, (20, 70), 5),    # mono packets into a stereo decoder
    (2, 1, True, 1, (20, 60), 6),     # stereo packets into a mono decoder
    (0, 1, True, 2, (10, 50), 7),
]


def _corpus(bw, sc, stereo, ch, payload, seed, lanes, ticks):
    return helpers.silk_fuzz_grid(lanes, ticks, payload, lambda k, r: helpers.silk_toc(bw, sc, stereo), seed=seed, channels=ch)


@pytest.mark.parametrize("bw,sc,stereo,ch,payload,seed", CASES)
def test_emul_silk_matches_oracle(bw, sc, stereo, ch, payload, seed):
    lanes, ticks = 6, 6
    arena, offsets = _corpus(bw, sc, stereo, ch, payload, seed, lanes, ticks)
    for i16 in (True, False):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 2880, i16=i16)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 2880, i16=i16)
        assert np.all(so > 0)
        assert np.array_equal(so, se)
        assert np.array_equal(ro, re)
        assert np.array_equal(po, pe)


def test_emul_silk_mixed_rates_and_channel_switches():
    """Every SILK-only configuration drawn per packet: exercises decoder_set_fs re-initialisation, the mono<->stereo
    switches of dec_api.rs:150-191 (quirk S4) and 10/20/40/60 ms framing inside one stream."""
    lanes, ticks = 8, 8

    def toc(k, rng):
        return helpers.silk_toc(int(rng.integers(0, 3)), int(rng.integers(0, 4)), bool(rng.integers(0, 2)))

    for ch in (1, 2):
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (4, 80), toc, seed=11 + ch, channels=ch)
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 2880, i16=True)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 2880, i16=True)
        assert np.array_equal(so, se) and np.array_equal(ro, re)
        assert np.array_equal(po, pe)


def test_emul_silk_quirk_switches_follow_the_oracle():
    """Each SILK quirk bit cleared in turn (libopus behaviour) still matches the oracle run with the same bits."""
    lanes, ticks = 6, 5
    arena, offsets = _corpus(1, 0, True, 2, (6, 30), 21, lanes, ticks)
    for bit in (1 << 8, 1 << 9, 1 << 10, 1 << 11):
        q = helpers.QUIRKS_MOUSIKI & ~bit
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 2880, i16=True, quirks=q)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, 2, 2880, i16=True, quirks=q)
        assert np.array_equal(so, se) and np.array_equal(ro, re) and np.array_equal(po, pe)


def test_emul_40ms_redundancy_decodes_and_a_bare_mode_switch_decodes():
    """A 40 ms SILK-only packet with a CELT redundancy frame decodes (its cell has no room for the CELT record: the frame goes
    to an auxiliary cell, mb_common.cuh); so does a SILK -> CELT switch without redundancy (the reference conceals 5 ms of the old
    mode and fades from it, opus_decoder.rs:575-593), and the SILK packet after it (CELT -> SILK, again without redundancy)."""
    lanes = 2
    good, goff = _corpus(2, 1, False, 1, (8, 40), 31, lanes, 2)
    long_pkt = np.concatenate([[helpers.silk_toc(2, 2, False)], np.zeros(300, dtype=np.uint8)]).astype(np.uint8)  # 40 ms, zeros: few bits used
    celt_pkt = np.concatenate([[helpers.celt_toc(3, False)], helpers.lcg_bytes(5, 60)]).astype(np.uint8)
    pk = lambda p: good[goff[p]:goff[p + 1]]
    rows = [[pk(0), pk(1)], [long_pkt, celt_pkt], [pk(2), pk(3)]]
    flat = [p for row in rows for p in row]
    offsets = np.zeros(len(flat) + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in flat])
    arena = np.concatenate(flat)
    pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, 3, 1, 1920, i16=True)
    assert se.tolist() == [960, 960, 1920, 960, 960, 960]
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, 3, 1, 1920, i16=True)
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
@pytest.mark.parametrize("bw,sc,stereo,ch,payload,seed", CASES)
def test_gpu_silk_matches_oracle(bw, sc, stereo, ch, payload, seed):
    lanes, ticks = 70, 5
    arena, offsets = _corpus(bw, sc, stereo, ch, payload, seed, lanes, ticks)
    for i16 in (True, False):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 2880, i16=i16)
        pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, ch, 2880, i16)
        assert np.array_equal(so, sg)
        assert np.array_equal(ro, rg), "final_range must be bit-exact"
        assert np.array_equal(po, pg), "SILK PCM is integer work: bit-exact"


@pytest.mark.gpu
def test_gpu_silk_mixed_rates_state_across_calls():
    """One packet per stream per call (live-server shape) == all ticks in one call == oracle."""
    from mousiki_b200 import BatchDecoder

    lanes, ticks = 64, 6

    def toc(k, rng):
        return helpers.silk_toc(int(rng.integers(0, 3)), int(rng.integers(0, 4)), bool(rng.integers(0, 2)))

    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (4, 80), toc, seed=41, channels=2)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 2880, i16=True)
    pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, 2, 2880, True)
    assert np.array_equal(so, sg) and np.array_equal(ro, rg) and np.array_equal(po, pg)
    ids = np.arange(lanes, dtype=np.uint32)
    with BatchDecoder(lanes, 48000, 2) as dec:
        for t in range(ticks):
            a = arena[offsets[t * lanes]:offsets[(t + 1) * lanes]]
            o = (offsets[t * lanes:(t + 1) * lanes + 1] - offsets[t * lanes]).astype(np.uint32)
            p, s, r = dec.decode(a, o, ids, 2880)
            assert np.array_equal(s, so[t * lanes:(t + 1) * lanes])
            assert np.array_equal(r, ro[t * lanes:(t + 1) * lanes])
            ref = po[t * lanes:(t + 1) * lanes]
            for j in range(lanes):  # rows are only defined up to the packet's own duration
                m = int(s[j]) * 2
                assert np.array_equal(p[j, :m], ref[j, :m])


@pytest.mark.gpu
def test_gpu_silk_and_celt_streams_in_one_batch():
    """Half the streams carry CELT packets, half SILK packets, in the same call: each kernel only touches its own cells."""
    lanes, ticks = 64, 4
    sa, so_ = helpers.silk_fuzz_grid(lanes // 2, ticks, (8, 40), lambda k, r: helpers.silk_toc(2, 1, False), seed=51, channels=2)
    ca, co = helpers.fuzz_grid(lanes // 2, ticks, 100, lambda t, j, r: helpers.celt_toc(3, True), seed=52)
    pk = []
    for t in range(ticks):
        for j in range(lanes):
            if j % 2 == 0:
                k = t * (lanes // 2) + j // 2
                pk.append(sa[so_[k]:so_[k + 1]])
            else:
                k = t * (lanes // 2) + j // 2
                pk.append(ca[co[k]:co[k + 1]])
    offsets = np.zeros(len(pk) + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in pk])
    arena = np.concatenate(pk)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 960)
    pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, 2, 960, False)
    assert np.array_equal(so, sg) and np.array_equal(ro, rg)
    assert float(np.abs(po - pg).max()) <= 1.0 / 32768.0
