"""P-struct, the structured CELT corpus (oracle/gen_corpus.cpp).

The writer runs the restated decoder "driven": every symbol the decoder asks for is chosen and range-encoded (oracle/ec.h,
GenCtl) instead of read. These tests pin the construction on the CHECKER build of the oracle (no ORC_GEN in it): the finished
bytes decode to the writer's own PCM and final_range, the steered symbols come out as steered (intra on the first frame only,
~10 % transients, post-filter on ~30 % of the frames, PCM inside +-1), and the product decodes them like the oracle does — the
host emulation of the kernel bodies on CPU, the CUDA path through the C ABI under -m gpu (f32 and i16)."""
import numpy as np
import pytest

import helpers

TOL = 1.0 / 32768.0

SHAPES = [
    ("fb 20 ms stereo 160 B", helpers.celt_toc(3, True), 2, 160),
    ("fb 20 ms mono 80 B", helpers.celt_toc(3, False), 1, 80),
    ("fb 10 ms stereo 100 B", helpers.celt_toc(2, True), 2, 100),
    ("swb 5 ms stereo 60 B", helpers.celt_toc(1, True, bandwidth=2), 2, 60),
    ("wb 2.5 ms mono 30 B", helpers.celt_toc(0, False, bandwidth=1), 1, 30),
    ("fb 20 ms stereo 12 B (starved)", helpers.celt_toc(3, True), 2, 12),
    ("fb 20 ms mono in a stereo decoder", helpers.celt_toc(3, False), 2, 120),
]


@pytest.mark.parametrize("name,toc,ch,payload", SHAPES, ids=[s[0] for s in SHAPES])
def test_written_packets_decode_to_the_writers_pcm(name, toc, ch, payload):
    lanes, ticks = 24, 12
    frame = 120 << ((toc >> 3) & 3)
    arena, offsets, pcm, rng = helpers.struct_grid(lanes, ticks, payload, toc, ch, seed=3, want_pcm=True)
    assert (arena[offsets[:-1]] == toc).all() and int(offsets[-1]) == lanes * ticks * (payload + 1)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame)
    assert (so == frame).all()
    assert np.array_equal(ro, rng), "final_range: the decoder left the writer's path"
    assert np.array_equal(po, pcm)
    # deterministic in the seed, and independent of the number of threads that wrote it
    a2, o2 = helpers.struct_grid(lanes, ticks, payload, toc, ch, seed=3, threads=1)
    assert np.array_equal(a2, arena) and np.array_equal(o2, offsets)
    a3, _ = helpers.struct_grid(lanes, ticks, payload, toc, ch, seed=4)
    assert not np.array_equal(a3, arena)


def test_steered_symbols_and_loudness():
    lanes, ticks = 96, 40
    arena, offsets, pcm, _ = helpers.struct_grid(lanes, ticks, 160, 0xFC, 2, seed=11, want_pcm=True)
    peak = np.abs(pcm).max(axis=1)
    assert (peak > 1.0).mean() < 0.01, "P-struct keeps the soft clipper idle (P-fuzz: a quarter of the packets leave +-1)"
    assert 0.03 < float(np.median(peak)) < 0.5
    lib = helpers.oracle_lib()
    info = np.zeros(16, dtype=np.int32)
    out = np.zeros(1920, dtype=np.float32)
    transient, intra, postfilter = [], [], []
    for j in range(0, lanes, 4):
        h = lib.orc_create(48000, 2, helpers.QUIRKS_MOUSIKI)
        lib.orc_trace_enable(h, 1)
        for t in range(ticks):
            p = t * lanes + j
            pkt = np.ascontiguousarray(arena[offsets[p]:offsets[p + 1]])
            assert lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(out), 960, 0) == 960
            lib.orc_trace_info(h, helpers.ptr(info))
            assert info[3] == 0, "no silence frames"
            transient.append(int(info[4])); intra.append((t, int(info[5]))); postfilter.append(int(info[11] != 0))
        lib.orc_destroy(h)
    assert all(v == (t == 0) for t, v in intra), "intra energy on a stream's first frame only"
    assert 0.04 < np.mean(transient) < 0.2
    assert 0.15 < np.mean(postfilter) < 0.45


@pytest.mark.parametrize("name,toc,ch,payload", SHAPES[:4], ids=[s[0] for s in SHAPES[:4]])
def test_emul_decodes_struct_packets_like_the_oracle(name, toc, ch, payload):
    lanes, ticks = 12, 8
    frame = 120 << ((toc >> 3) & 3)
    arena, offsets = helpers.struct_grid(lanes, ticks, payload, toc, ch, seed=21)
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=i16)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=i16)
        assert np.array_equal(so, se) and np.array_equal(ro, re)
        assert np.array_equal(po, pe), (name, i16)


@pytest.mark.gpu
@pytest.mark.parametrize("name,toc,ch,payload", SHAPES, ids=[s[0] for s in SHAPES])
def test_gpu_decodes_struct_packets_like_the_oracle(name, toc, ch, payload):
    from mousiki_b200 import BatchDecoder

    lanes, ticks = 192, 12
    frame = 120 << ((toc >> 3) & 3)
    arena, offsets = helpers.struct_grid(lanes, ticks, payload, toc, ch, seed=31)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=i16)
        with BatchDecoder(lanes, 48000, ch, max_frames_per_pass=5 * lanes) as dec:
            pg, sg, rg = dec.decode(arena, offsets, ids, frame) if i16 else dec.decode_float(arena, offsets, ids, frame)
        assert np.array_equal(so, sg) and np.array_equal(ro, rg), name
        if i16:
            assert np.array_equal(po, pg), (name, int(np.abs(po.astype(np.int32) - pg.astype(np.int32)).max()))
        else:
            assert float(np.abs(po - pg).max()) <= TOL, name
