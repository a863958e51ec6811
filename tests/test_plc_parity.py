"""Packet-loss concealment of CELT-only streams on the product path: lost packets
(`Decoder::decode(None)` / `conceal`, src/codec.rs:861-914) and frames of at most one byte are concealed by
celt_decode_lost (src/celt/celt_decoder.rs:1429-1855): pitch repetition for the first losses of a burst, band-shaped noise
after 40 lost blocks, after a reset, and (later) for hybrid frames; the first coded frame after a loss starts from safer
energies (:2943-2969). Against the oracle, which is pinned on the reference's own PLC / FEC vectors
(tests/test_plc_oracle.py). CPU: host emulation of the kernel bodies, bit for bit. GPU (-m gpu): the CUDA path through the C ABI."""
import numpy as np
import pytest

import helpers

TOL = 1.0 / 32768.0


def lossy_grid(lanes, ticks, lm, stereo, seed, loss_rate=0.25, burst=6, payload=(20, 160)):
    """[ticks][lanes] CELT grid plus a loss pattern: isolated losses, and for a few lanes a burst long enough to move from pitch
    repetition to noise (5 x 20 ms). Returns (arena, offsets, lost[ticks*lanes] bool)."""
    arena, offsets = helpers.fuzz_grid(lanes, ticks, payload, lambda t, j, r: helpers.celt_toc(lm, stereo), seed=seed)
    rng = np.random.default_rng(seed + 1)
    lost = rng.random(lanes * ticks) < loss_rate
    lost[:lanes] = False                       # tick 0 decodes (a loss before any packet is its own test)
    for j in range(0, lanes, 3):               # bursts
        t0 = 2 + int(rng.integers(0, max(ticks - burst - 2, 1)))
        for t in range(t0, min(t0 + burst, ticks)):
            lost[t * lanes + j] = True
    return arena, offsets, lost


def oracle_lossy(arena, offsets, lost, lanes, ticks, ch, n, i16=False):
    lib = helpers.oracle_lib()
    pcm = np.zeros((lanes * ticks, n * ch), dtype=np.int16 if i16 else np.float32)
    samples = np.zeros(lanes * ticks, dtype=np.int32)
    rng = np.zeros(lanes * ticks, dtype=np.uint32)
    fn = lib.orc_decode_i16 if i16 else lib.orc_decode_float
    for j in range(lanes):
        h = lib.orc_create(48000, ch, helpers.QUIRKS_MOUSIKI)
        for t in range(ticks):
            p = t * lanes + j
            if lost[p]:
                samples[p] = fn(h, None, 0, helpers.ptr(pcm[p]), n, 0)
            else:
                pkt = np.ascontiguousarray(arena[offsets[p]:offsets[p + 1]])
                samples[p] = fn(h, helpers.ptr(pkt), len(pkt), helpers.ptr(pcm[p]), n, 0)
            rng[p] = lib.orc_final_range(h)
        lib.orc_destroy(h)
    return pcm, samples, rng


def virtual_grid(arena, offsets, lost, n):
    """What the host front end hands the device for a regular grid: a lost packet becomes a two-byte record of its sample
    count with flag bit 1 (mb_host_packets.h)."""
    pk = []
    for p in range(len(offsets) - 1):
        pk.append(np.array([n & 255, n >> 8], dtype=np.uint8) if lost[p] else arena[offsets[p]:offsets[p + 1]])
    off = np.zeros(len(pk) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(x) for x in pk])
    return np.concatenate(pk).astype(np.uint8), off, (lost.astype(np.uint8) << 1)


CASES = [(3, True, 2), (3, False, 1), (2, True, 2), (1, False, 2), (0, True, 2)]


@pytest.mark.parametrize("lm,stereo,ch", CASES)
def test_emul_celt_streams_with_lost_packets(lm, stereo, ch):
    lanes, ticks, n = 6, 14, 120 << lm
    arena, offsets, lost = lossy_grid(lanes, ticks, lm, stereo, seed=90 + lm + ch, burst=5 * (8 >> lm) + 2 if lm >= 2 else 12)
    va, vo, vf = virtual_grid(arena, offsets, lost, n)
    lib = helpers.emul_lib()
    for i16 in (False, True):
        po, so, ro = oracle_lossy(arena, offsets, lost, lanes, ticks, ch, n, i16)
        lib.emu_set_flags(helpers.ptr(vf))
        try:
            pe, se, re, _, _ = helpers.emul_decode_grid(va, vo, lanes, ticks, ch, n, i16=i16)
        finally:
            lib.emu_set_flags(None)
        assert (so == n).all() and np.array_equal(so, se)
        assert np.array_equal(ro, re) and (re[lost] == 0).all()
        if not np.array_equal(po, pe):
            bad = np.nonzero((po != pe).any(axis=1))[0]
            p = int(bad[0])
            raise AssertionError(f"first differing packet: tick {p // lanes} lane {p % lanes} lost={bool(lost[p])} "
                                 f"max|d|={np.abs(po[p].astype(np.float64) - pe[p]).max()} ({len(bad)} packets differ)")
    assert lost.sum() > 10


def test_emul_loss_before_anything_was_decoded_is_silence():
    lanes, ticks, n = 3, 3, 960
    arena, offsets = helpers.fuzz_grid(lanes, ticks, 80, lambda t, j, r: helpers.celt_toc(3, True), seed=5)
    lost = np.zeros(lanes * ticks, dtype=bool)
    lost[:lanes] = True
    va, vo, vf = virtual_grid(arena, offsets, lost, n)
    po, so, ro = oracle_lossy(arena, offsets, lost, lanes, ticks, 2, n)
    lib = helpers.emul_lib()
    lib.emu_set_flags(helpers.ptr(vf))
    try:
        pe, se, re, _, _ = helpers.emul_decode_grid(va, vo, lanes, ticks, 2, n)
    finally:
        lib.emu_set_flags(None)
    assert np.array_equal(so, se) and np.array_equal(ro, re) and np.array_equal(po, pe)
    assert (pe[:lanes] == 0).all()


# ---------------------------------------------------------------------------------------------- GPU
def _gpu(arena, offsets, ids, flags, n_streams, ch, cap, i16, **kw):
    from mousiki_b200 import BatchDecoder

    with BatchDecoder(n_streams, 48000, ch, **kw) as dec:
        return dec.decode(arena, offsets, ids, cap, flags=flags) if i16 else dec.decode_float(arena, offsets, ids, cap, flags=flags)


@pytest.mark.gpu
@pytest.mark.parametrize("lm,stereo,ch", CASES)
def test_gpu_celt_streams_with_lost_packets(lm, stereo, ch):
    """Through the C ABI: a lost packet is an empty packet with flag bit 1; the row's capacity says how much to conceal."""
    lanes, ticks, n = 48, 14, 120 << lm
    arena, offsets, lost = lossy_grid(lanes, ticks, lm, stereo, seed=190 + lm + ch, burst=5 * (8 >> lm) + 2 if lm >= 2 else 12)
    lens = np.diff(offsets.astype(np.int64))
    lens[lost] = 0
    keep = np.concatenate([arena[offsets[p]:offsets[p + 1]] for p in range(len(lens)) if not lost[p]])
    off2 = np.zeros(len(lens) + 1, dtype=np.uint32)
    off2[1:] = np.cumsum(lens)
    flags = (lost.astype(np.uint8) << 1)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    for i16 in (False, True):
        po, so, ro = oracle_lossy(arena, offsets, lost, lanes, ticks, ch, n, i16)
        pg, sg, rg = _gpu(keep, off2, ids, flags, lanes, ch, n, i16, max_frames_per_pass=5 * lanes)   # losses straddle pass boundaries
        assert np.array_equal(so, sg) and np.array_equal(ro, rg)
        if i16:
            assert int(np.abs(po.astype(np.int32) - pg.astype(np.int32)).max()) <= 1
        else:
            err = float(np.abs(po - pg).max())
            print(f"lm={lm}: f32 max err {err:.3g}, identical {np.array_equal(po, pg)}")
            assert err <= TOL


@pytest.mark.gpu
def test_gpu_one_byte_frames_are_concealed():
    """A code-0 packet with a payload of at most one byte is a concealed frame of the packet's size (src/opus_decoder.rs:491-496);
    inside a multi-frame packet the other frames decode."""
    lanes = 16
    toc = helpers.celt_toc(2, True)
    a0, o0 = helpers.fuzz_grid(lanes, 6, 70, lambda t, j, r: toc, seed=77)
    pk = [[a0[o0[t * lanes + j]:o0[t * lanes + j + 1]] for j in range(lanes)] for t in range(6)]
    seq = []
    for j in range(lanes):
        fr = lambda t: bytes(pk[t][j][1:])
        seq.append([pk[0][j], pk[1][j], np.array([toc, 0x55], dtype=np.uint8),
                    helpers.multiframe_packet(toc, [fr(3), b"\x01", fr(4)], 3, vbr=True), pk[5][j]])
    lib = helpers.oracle_lib()
    want = []
    for j in range(lanes):
        h = lib.orc_create(48000, 2, helpers.QUIRKS_MOUSIKI)
        for pkt in seq[j]:
            out = np.zeros(1440 * 2, dtype=np.float32)
            r = lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(out), 1440, 0)
            want.append((r, out, lib.orc_final_range(h)))
        lib.orc_destroy(h)
    pkts = [seq[j][k] for k in range(5) for j in range(lanes)]
    off = np.zeros(len(pkts) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(p) for p in pkts])
    ids = np.tile(np.arange(lanes, dtype=np.uint32), 5)
    pg, sg, rg = _gpu(np.concatenate(pkts).astype(np.uint8), off, ids, None, lanes, 2, 1440, False)
    for k in range(5):
        for j in range(lanes):
            r, out, rng = want[j * 5 + k]
            slot = k * lanes + j
            assert sg[slot] == r and r in (480, 1440), (k, j, sg[slot], r)
            assert rg[slot] == rng
            assert float(np.abs(pg[slot, :r * 2] - out[:r * 2]).max()) <= TOL, (k, j)
