"""Packet-loss concealment and in-band FEC of SILK-only and hybrid streams on the product path (SURVEY.md section 8 rows F2,
A16): lost packets run silk_plc_conceal + comfort noise (src/silk/plc.rs:186-345, src/silk/cng.rs:235-370), the first decoded
frame after a loss is glued on (plc.rs:74-115) and starts from expanded LPCs / a quarter-gain pitch predictor
(decode_parameters.rs:156-160, decode_core.rs:140-152); a hybrid stream also conceals its CELT part from band 17 with
band-shaped noise (src/celt/celt_decoder.rs:1511-1548); `decode(fec=true)` rebuilds the previous frame from the packet's LBRR
data (src/opus_decoder.rs:1240-1316, src/silk/dec_api.rs:200-330). Against the oracle, which is pinned on the reference's
HYBRID_PLC_PCM / FEC_RANGE / FEC_PCM vectors (tests/test_plc_oracle.py).
CPU: host emulation of the kernel bodies, bit for bit. GPU (-m gpu): the CUDA path through the C ABI."""
import numpy as np
import pytest

import helpers
from test_hybrid_parity import hyb_toc
from test_plc_parity import oracle_lossy, virtual_grid

TOL = 1.0 / 32768.0


def loss_pattern(lanes, ticks, seed, rate=0.25, burst=4):
    rng = np.random.default_rng(seed)
    lost = rng.random(lanes * ticks) < rate
    lost[:lanes] = False
    for j in range(0, lanes, 3):
        t0 = 1 + int(rng.integers(0, max(ticks - burst - 1, 1)))
        for t in range(t0, min(t0 + burst, ticks)):
            lost[t * lanes + j] = True
    return lost


# bandwidth (0 NB, 1 MB, 2 WB), size code (0: 10 ms, 1: 20 ms), packet stereo, decoder channels, payload range
SILK_CASES = [(2, 1, False, 1, (20, 70)), (2, 1, True, 2, (40, 120)), (0, 1, False, 2, (15, 50)), (1, 0, True, 2, (20, 60)),
              (2, 0, False, 1, (10, 40)), (1, 1, True, 1, (30, 90))]


def _check(arena, offsets, lost, lanes, ticks, ch, n):
    va, vo, vf = virtual_grid(arena, offsets, lost, n)
    lib = helpers.emul_lib()
    for i16 in (False, True):
        po, so, ro = oracle_lossy(arena, offsets, lost, lanes, ticks, ch, n, i16)
        lib.emu_set_flags(helpers.ptr(vf))
        try:
            pe, se, re, _, _ = helpers.emul_decode_grid(va, vo, lanes, ticks, ch, n, i16=i16)
        finally:
            lib.emu_set_flags(None)
        assert (so == n).all(), so
        assert np.array_equal(so, se), (so, se)
        assert np.array_equal(ro, re) and (re[lost] == 0).all()
        if not np.array_equal(po, pe):
            bad = np.nonzero((po != pe).any(axis=1))[0]
            p = int(bad[0])
            d = np.nonzero(po[p] != pe[p])[0]
            raise AssertionError(f"i16={i16}: first differing packet: tick {p // lanes} lane {p % lanes} lost={bool(lost[p])} "
                                 f"first sample {d[0]} of {len(d)}, max|d|={np.abs(po[p].astype(np.float64) - pe[p]).max()} "
                                 f"({len(bad)} packets differ: {bad[:12]})")


@pytest.mark.parametrize("bw,sc,stereo,ch,payload", SILK_CASES)
def test_emul_silk_streams_with_lost_packets(bw, sc, stereo, ch, payload):
    lanes, ticks, n = 6, 12, 480 << sc
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, lambda k, r: helpers.silk_toc(bw, sc, stereo), seed=300 + 7 * bw + sc, channels=ch)
    lost = loss_pattern(lanes, ticks, seed=17 + bw + 3 * sc)
    _check(arena, offsets, lost, lanes, ticks, ch, n)
    assert lost.sum() > 10


HYB_CASES = [(1, 1, False, 1, (30, 160)), (1, 1, True, 2, (40, 200)), (0, 0, True, 2, (20, 100)), (0, 1, False, 2, (30, 120))]


@pytest.mark.parametrize("fb,ms20,stereo,ch,payload", HYB_CASES)
def test_emul_hybrid_streams_with_lost_packets(fb, ms20, stereo, ch, payload):
    lanes, ticks, n = 6, 10, 960 if ms20 else 480
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, lambda k, r: hyb_toc(fb, ms20, stereo), seed=400 + fb + 2 * ms20, channels=ch)
    lost = loss_pattern(lanes, ticks, seed=23 + fb + ms20)
    _check(arena, offsets, lost, lanes, ticks, ch, n)


# ---------------------------------------------------------------------------------------------- in-band FEC
def fec_grid(arena, offsets, lost, lanes, ticks):
    """The receiver's view of a lossy stream with FEC: where packet t is lost and packet t + 1 arrived, the row of tick t is
    `decode(packet t + 1, fec=true)`; where t + 1 is lost too (or t is the last tick) it is `decode(None)`.
    Returns (arena, offsets, flags) of the grid handed to the decoder: flag bit0 FEC, bit1 lost."""
    pk, fl = [], []
    for t in range(ticks):
        for j in range(lanes):
            p = t * lanes + j
            if not lost[p]:
                pk.append(arena[offsets[p]:offsets[p + 1]]); fl.append(0)
            elif t + 1 < ticks and not lost[p + lanes]:
                pk.append(arena[offsets[p + lanes]:offsets[p + lanes + 1]]); fl.append(1)
            else:
                pk.append(np.zeros(0, dtype=np.uint8)); fl.append(2)
    off = np.zeros(len(pk) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in pk])
    return np.concatenate(pk).astype(np.uint8), off, np.array(fl, dtype=np.uint8)


def oracle_fec(arena, offsets, flags, lanes, ticks, ch, n, i16=False):
    lib = helpers.oracle_lib()
    pcm = np.zeros((lanes * ticks, n * ch), dtype=np.int16 if i16 else np.float32)
    samples = np.zeros(lanes * ticks, dtype=np.int32)
    rng = np.zeros(lanes * ticks, dtype=np.uint32)
    fn = lib.orc_decode_i16 if i16 else lib.orc_decode_float
    for j in range(lanes):
        h = lib.orc_create(48000, ch, helpers.QUIRKS_MOUSIKI)
        for t in range(ticks):
            p = t * lanes + j
            if flags[p] & 2:
                samples[p] = fn(h, None, 0, helpers.ptr(pcm[p]), n, 0)
            else:
                pkt = np.ascontiguousarray(arena[offsets[p]:offsets[p + 1]])
                samples[p] = fn(h, helpers.ptr(pkt), len(pkt), helpers.ptr(pcm[p]), n, int(flags[p] & 1))
            rng[p] = lib.orc_final_range(h)
        lib.orc_destroy(h)
    return pcm, samples, rng


def _check_fec(arena, offsets, lost, lanes, ticks, ch, n):
    fa, fo, ff = fec_grid(arena, offsets, lost, lanes, ticks)
    # the device's grid: a lost packet is the two-byte record of its sample count (mb_host_packets.h)
    pk = [np.array([n & 255, n >> 8], dtype=np.uint8) if ff[p] & 2 else fa[fo[p]:fo[p + 1]] for p in range(len(ff))]
    vo = np.zeros(len(pk) + 1, dtype=np.uint32)
    vo[1:] = np.cumsum([len(x) for x in pk])
    va = np.concatenate(pk).astype(np.uint8)
    lib = helpers.emul_lib()
    for i16 in (False, True):
        po, so, ro = oracle_fec(fa, fo, ff, lanes, ticks, ch, n, i16)
        lib.emu_set_flags(helpers.ptr(ff))
        try:
            pe, se, re, _, _ = helpers.emul_decode_grid(va, vo, lanes, ticks, ch, n, i16=i16)
        finally:
            lib.emu_set_flags(None)
        assert (so == n).all(), so
        assert np.array_equal(so, se), (so, se)
        assert np.array_equal(ro, re)
        if not np.array_equal(po, pe):
            bad = np.nonzero((po != pe).any(axis=1))[0]
            p = int(bad[0])
            d = np.nonzero(po[p] != pe[p])[0]
            raise AssertionError(f"i16={i16}: first differing packet: tick {p // lanes} lane {p % lanes} flag={ff[p]} "
                                 f"first sample {d[0]} of {len(d)}, max|d|={np.abs(po[p].astype(np.float64) - pe[p]).max()} "
                                 f"({len(bad)} packets differ: {bad[:12]})")
    return ff


@pytest.mark.parametrize("bw,sc,stereo,ch,payload", SILK_CASES)
def test_emul_silk_streams_with_fec(bw, sc, stereo, ch, payload):
    lanes, ticks, n = 6, 12, 480 << sc
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, lambda k, r: helpers.silk_toc(bw, sc, stereo), seed=500 + 7 * bw + sc, channels=ch)
    lost = loss_pattern(lanes, ticks, seed=31 + bw + 3 * sc, rate=0.3, burst=2)
    ff = _check_fec(arena, offsets, lost, lanes, ticks, ch, n)
    assert (ff == 1).sum() >= 6


@pytest.mark.parametrize("fb,ms20,stereo,ch,payload", HYB_CASES)
def test_emul_hybrid_streams_with_fec(fb, ms20, stereo, ch, payload):
    lanes, ticks, n = 6, 10, 960 if ms20 else 480
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, lambda k, r: hyb_toc(fb, ms20, stereo), seed=600 + fb + 2 * ms20, channels=ch)
    lost = loss_pattern(lanes, ticks, seed=41 + fb + ms20, rate=0.3, burst=2)
    ff = _check_fec(arena, offsets, lost, lanes, ticks, ch, n)
    assert (ff == 1).sum() >= 6


def test_host_front_end_cuts_a_fec_request_into_concealment_and_the_lbrr_frame():
    """decode(packet, fec=true) into a row of 40 ms after 20 ms packets: 20 ms of concealment, then the packet's first frame with
    flag bit0, as a continuation (src/opus_decoder.rs:1268-1316); a CELT-only packet, or a decoder whose last packet was
    CELT-only, conceals the whole row (:1247-1250)."""
    lib = helpers.emul_lib()
    silk = np.array([helpers.silk_toc(2, 1, False)] + [7] * 30, dtype=np.uint8)
    celt = np.array([helpers.celt_toc(3, False)] + [9] * 30, dtype=np.uint8)
    two = helpers.multiframe_packet(helpers.silk_toc(2, 1, False), [bytes([1] * 20), bytes([2] * 25)], 2)
    pk = [silk, two, celt, silk, celt]
    flags = np.array([0, 1, 0, 1, 1], dtype=np.uint8)
    off = np.zeros(len(pk) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in pk])
    lib.emu_split_set_streams(1)
    try:
        S = helpers.split_batch(np.concatenate(pk), off, np.zeros(len(pk), dtype=np.uint32), 1920, 1, flags)
    finally:
        lib.emu_split_set_streams(0)
    first, vf, vo, om = S["first"], S["flags"], S["offsets"], S["out_map"].reshape(-1, 2)
    assert list(np.diff(first)) == [1, 2, 1, 2, 2]
    # packet 1: 960 concealed at element 0, then the first frame of the two-frame packet at element 960
    v = first[1]
    assert vf[v] == 2 and bytes(S["arena"][vo[v]:vo[v + 1]]) == bytes([960 & 255, 960 >> 8]) and om[v][1] == 0
    assert vf[v + 1] == (1 | 4) and om[v + 1][1] == 960
    assert bytes(S["arena"][vo[v + 1]:vo[v + 2]]) == bytes([int(two[0]) & 0xFC]) + bytes([1] * 20)
    # packet 3: the decoder's last packet was CELT-only -> the whole row is concealed, in frames of the CELT packet's size
    v = first[3]
    assert list(vf[v:first[4]]) == [2, 2 | 4] and [int(x) for x in om[v:first[4], 1]] == [0, 960]
    # packet 4: a CELT-only packet carries no FEC data
    v = first[4]
    assert list(vf[v:first[5]]) == [2, 2 | 4]


# ---------------------------------------------------------------------------------------------- GPU
def _gpu(arena, offsets, ids, flags, n_streams, ch, cap, i16, **kw):
    from mousiki_b200 import BatchDecoder

    with BatchDecoder(n_streams, 48000, ch, **kw) as dec:
        return dec.decode(arena, offsets, ids, cap, flags=flags) if i16 else dec.decode_float(arena, offsets, ids, cap, flags=flags)


def _gpu_check(fa, fo, ff, lanes, ticks, ch, n, **kw):
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    for i16 in (False, True):
        po, so, ro = oracle_fec(fa, fo, ff, lanes, ticks, ch, n, i16)
        pg, sg, rg = _gpu(fa, fo, ids, ff, lanes, ch, n, i16, **kw)
        assert np.array_equal(so, sg), (so, sg)
        assert np.array_equal(ro, rg)
        if i16:
            assert int(np.abs(po.astype(np.int32) - pg.astype(np.int32)).max()) <= 1
        else:
            err = float(np.abs(po - pg).max())
            print(f"f32 max err {err:.3g}, identical {np.array_equal(po, pg)}")
            assert err <= TOL


def _lossy(arena, offsets, lost):
    """Empty packets with flag bit1 where `lost`."""
    pk = [np.zeros(0, dtype=np.uint8) if lost[p] else arena[offsets[p]:offsets[p + 1]] for p in range(len(lost))]
    off = np.zeros(len(pk) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in pk])
    return np.concatenate(pk).astype(np.uint8), off, (lost.astype(np.uint8) << 1)


@pytest.mark.gpu
@pytest.mark.parametrize("bw,sc,stereo,ch,payload", SILK_CASES)
def test_gpu_silk_streams_with_lost_packets_and_fec(bw, sc, stereo, ch, payload):
    lanes, ticks, n = 40, 12, 480 << sc
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, lambda k, r: helpers.silk_toc(bw, sc, stereo), seed=700 + 7 * bw + sc, channels=ch)
    lost = loss_pattern(lanes, ticks, seed=51 + bw + 3 * sc)
    _gpu_check(*_lossy(arena, offsets, lost), lanes, ticks, ch, n, max_frames_per_pass=5 * lanes)   # losses straddle pass boundaries
    _gpu_check(*fec_grid(arena, offsets, lost, lanes, ticks), lanes, ticks, ch, n, max_frames_per_pass=5 * lanes)


@pytest.mark.gpu
@pytest.mark.parametrize("fb,ms20,stereo,ch,payload", HYB_CASES)
def test_gpu_hybrid_streams_with_lost_packets_and_fec(fb, ms20, stereo, ch, payload):
    lanes, ticks, n = 40, 10, 960 if ms20 else 480
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, lambda k, r: hyb_toc(fb, ms20, stereo), seed=800 + fb + 2 * ms20, channels=ch)
    lost = loss_pattern(lanes, ticks, seed=61 + fb + ms20)
    _gpu_check(*_lossy(arena, offsets, lost), lanes, ticks, ch, n, max_frames_per_pass=4 * lanes)
    _gpu_check(*fec_grid(arena, offsets, lost, lanes, ticks), lanes, ticks, ch, n, max_frames_per_pass=4 * lanes)


@pytest.mark.gpu
def test_gpu_fec_into_a_row_longer_than_the_frame():
    """decode(packet, fec=true) with 40 ms of room after 20 ms packets: 20 ms of concealment, then the LBRR frame."""
    lanes, ticks, n = 24, 8, 960
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (30, 90), lambda k, r: helpers.silk_toc(2, 1, bool(k & 1)), seed=900, channels=2)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    # ticks 3 and 4 are lost; tick 3's row is skipped by the receiver, tick 4's row asks packet 5 for 40 ms with FEC
    keep = [t for t in range(ticks) if t != 3]
    pk, fl = [], []
    for t in keep:
        for j in range(lanes):
            src = (5 if t == 4 else t) * lanes + j
            pk.append(arena[offsets[src]:offsets[src + 1]]); fl.append(1 if t == 4 else 0)
    off = np.zeros(len(pk) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in pk])
    fa, ff = np.concatenate(pk).astype(np.uint8), np.array(fl, dtype=np.uint8)
    idk = np.tile(np.arange(lanes, dtype=np.uint32), len(keep))
    lib = helpers.oracle_lib()
    cap = 1920
    want = np.zeros((len(pk), cap * 2), dtype=np.float32)
    ws, wr = np.zeros(len(pk), dtype=np.int32), np.zeros(len(pk), dtype=np.uint32)
    for j in range(lanes):
        h = lib.orc_create(48000, 2, helpers.QUIRKS_MOUSIKI)
        for k in range(len(keep)):
            p = k * lanes + j
            pkt = np.ascontiguousarray(fa[off[p]:off[p + 1]])
            ws[p] = lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(want[p]), cap if ff[p] else n, int(ff[p]))
            wr[p] = lib.orc_final_range(h)
        lib.orc_destroy(h)
    # the rows have room for 40 ms; regular packets fill 20 ms of it, the FEC request all of it
    pg, sg, rg = _gpu(fa, off, idk, ff, lanes, 2, cap, False)
    fec_rows = ff == 1
    assert (ws[fec_rows] == cap).all() and np.array_equal(ws, sg) and np.array_equal(wr, rg)
    for p in range(len(pk)):
        assert float(np.abs(want[p, :ws[p] * 2] - pg[p, :ws[p] * 2]).max()) <= TOL, p
