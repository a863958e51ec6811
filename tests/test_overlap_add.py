"""The overlap-add form of the inverse MDCT: quirk Q2 cleared ("what's weak" 4).

mousiki's clt_mdct_backward zero-fills its output slice before writing (src/celt/mdct.rs:373), which removes the
time-domain aliasing cancellation between consecutive blocks; libopus does not, and the day upstream drops the fill the
product has to carry each block's 60 unwindowed tail samples into the next block. The oracle restates both forms
(QUIRK_Q2_MDCT_ZERO_FILL). With the bit cleared the product decodes coded CELT-only frames in the libopus form: between the
short blocks of a frame inside the IMDCT unit, across frames in the stream's post-filter pass from a per-unit tail store, across
passes and calls from the persisted overlap region; lost packets, FEC and packets with a SILK part are refused per packet (-5).

CPU: host emulation of the kernel bodies against the oracle, bit for bit. GPU (-m gpu): the CUDA path through the C ABI."""
import numpy as np
import pytest

import helpers
from helpers import QUIRKS_MOUSIKI

Q_OLA = QUIRKS_MOUSIKI & ~0x2
TOL = 1.0 / 32768.0

SHAPES = [
    ("fb 20 ms stereo", helpers.celt_toc(3, True), 2, 960, (40, 160)),
    ("fb 10 ms stereo", helpers.celt_toc(2, True), 2, 480, (30, 100)),
    ("swb 5 ms mono", helpers.celt_toc(1, False, bandwidth=2), 1, 240, (20, 60)),
    ("fb 2.5 ms stereo", helpers.celt_toc(0, True), 2, 120, (10, 40)),
    ("fb 20 ms mono in a stereo decoder", helpers.celt_toc(3, False), 2, 960, (30, 120)),
    ("fb 20 ms stereo in a mono decoder", helpers.celt_toc(3, True), 1, 960, (40, 160)),
]


@pytest.mark.parametrize("name,toc,ch,frame,payload", SHAPES, ids=[s[0] for s in SHAPES])
def test_emul_overlap_add_form(name, toc, ch, frame, payload):
    lanes, ticks = 10, 12
    arena, offsets = helpers.fuzz_grid(lanes, ticks, payload, lambda t, j, r: toc, seed=5)
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=i16, quirks=Q_OLA)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=i16, quirks=Q_OLA)
        assert (so == frame).all() and np.array_equal(so, se) and np.array_equal(ro, re)
        assert np.array_equal(po, pe), (name, i16)
    # and it is a different decoder: the zero fill changes the first 120 samples of every block
    pq, _, _ = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame)
    po, _, _ = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame, quirks=Q_OLA)
    assert not np.array_equal(pq, po)


def test_emul_overlap_add_mixed_frame_sizes_and_struct_packets():
    lanes, ticks = 8, 16
    arena, offsets = helpers.fuzz_grid(lanes, ticks, (20, 120), lambda t, j, r: helpers.celt_toc(int(r.integers(0, 4)), True), seed=9)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 960, quirks=Q_OLA)
    pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, 2, 960, quirks=Q_OLA)
    assert np.array_equal(so, se) and np.array_equal(ro, re) and np.array_equal(po, pe)
    arena, offsets = helpers.struct_grid(lanes, ticks, 160, 0xFC, 2, seed=13)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 960, quirks=Q_OLA)
    pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, 2, 960, quirks=Q_OLA)
    assert np.array_equal(so, se) and np.array_equal(ro, re) and np.array_equal(po, pe)


@pytest.mark.gpu
@pytest.mark.parametrize("name,toc,ch,frame,payload", SHAPES, ids=[s[0] for s in SHAPES])
def test_gpu_overlap_add_form(name, toc, ch, frame, payload):
    from mousiki_b200 import BatchDecoder

    lanes, ticks = 160, 12
    arena, offsets = helpers.fuzz_grid(lanes, ticks, payload, lambda t, j, r: toc, seed=15)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame, quirks=Q_OLA)
    p16, _, _ = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=True, quirks=Q_OLA)
    # several passes per call (the tail crosses pass boundaries through the persisted overlap region) ...
    with BatchDecoder(lanes, 48000, ch, quirks=Q_OLA, max_frames_per_pass=5 * lanes) as dec:
        pg, sg, rg = dec.decode_float(arena, offsets, ids, frame)
    assert np.array_equal(so, sg) and np.array_equal(ro, rg)
    assert float(np.abs(po - pg).max()) <= TOL, name
    assert np.array_equal(po, pg), name
    # ... several calls (it crosses calls the same way), int16 output
    with BatchDecoder(lanes, 48000, ch, quirks=Q_OLA) as dec:
        parts = []
        for t0 in range(0, ticks, 4):
            sl = slice(t0 * lanes, (t0 + 4) * lanes)
            a = arena[offsets[sl.start]:offsets[sl.stop]]
            o = (offsets[sl.start:sl.stop + 1] - offsets[sl.start]).astype(np.uint32)
            parts.append(dec.decode(a, o, ids[sl], frame)[0])
    assert np.array_equal(np.concatenate(parts), p16), name


@pytest.mark.gpu
def test_gpu_overlap_add_refuses_what_it_does_not_carry():
    from mousiki_b200 import BatchDecoder

    lanes = 8
    arena, offsets = helpers.fuzz_grid(lanes, 3, 60, lambda t, j, r: helpers.celt_toc(3, True), seed=17)
    # tick 1: lanes 0-1 lost, lanes 2-3 SILK-only, lanes 4-5 hybrid, the rest coded CELT frames
    flags = np.zeros(3 * lanes, dtype=np.uint8)
    flags[lanes + 0] = flags[lanes + 1] = 2
    arena[offsets[lanes + 2]] = arena[offsets[lanes + 3]] = helpers.silk_toc(2, 1, True)
    arena[offsets[lanes + 4]] = arena[offsets[lanes + 5]] = 0x7C   # hybrid FB 20 ms stereo
    ids = np.tile(np.arange(lanes, dtype=np.uint32), 3)
    with BatchDecoder(lanes, 48000, 2, quirks=Q_OLA) as dec:
        pg, sg, rg = dec.decode_float(arena, offsets, ids, 960, flags=flags)
    assert (sg[lanes:lanes + 6] == -5).all(), sg[lanes:2 * lanes]
    assert (sg[:lanes] == 960).all() and (sg[lanes + 6:2 * lanes] == 960).all()
    # the streams that only saw coded CELT frames still agree with the oracle
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, 3, 2, 960, quirks=Q_OLA)
    for j in (6, 7):
        for t in range(3):
            assert np.array_equal(pg[t * lanes + j], po[t * lanes + j]) and rg[t * lanes + j] == ro[t * lanes + j]
