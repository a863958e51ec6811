"""CPU checks of the kernel bodies' logic: the single-threaded host build of mousiki_b200/csrc/mb_*.cuh
(tests/emul) against the oracle on seeded fuzz corpora. Integer results (final_range, sample counts,
PVQ pulse vectors) must be bit-exact; PCM is compared bit-for-bit as well because both sides follow the
reference's operation order (the GPU tests allow the north star's 1-LSB tolerance for libm differences)."""
import numpy as np
import pytest

import helpers

CASES = [
    # lm, packet stereo, decoder channels, payload bytes, bandwidth
    (3, True, 2, 160, 3),
    (3, False, 1, 80, 3),
    (0, True, 2, 40, 3),
    (1, False, 2, 60, 2),
    (2, True, 1, 100, 1),
    (3, True, 2, (2, 400), 3),
    (2, False, 1, (2, 60), 0),
]


@pytest.mark.parametrize("lm,stereo,ch,payload,bw", CASES)
def test_emul_matches_oracle(lm, stereo, ch, payload, bw):
    lanes, ticks = 24, 5
    arena, offsets = helpers.fuzz_grid(lanes, ticks, payload, lambda t, j, r: helpers.celt_toc(lm, stereo, bw), seed=100 + lm * 7 + ch)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
    pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 960)
    assert np.array_equal(so, se)
    assert np.array_equal(ro, re)
    assert np.array_equal(po.view(np.uint32), pe.view(np.uint32))


def test_emul_mixed_frame_sizes_and_channels():
    """Config 5 flavour: every CELT frame size / bandwidth / channel count in one batch, state carried per stream."""
    lanes, ticks = 32, 6

    def toc(t, j, rng):
        return helpers.celt_toc(int(rng.integers(0, 4)), bool(rng.integers(0, 2)), int(rng.integers(0, 4)))

    arena, offsets = helpers.fuzz_grid(lanes, ticks, (2, 300), toc, seed=5)
    for ch in (1, 2):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, 960)
        assert np.array_equal(so, se) and np.array_equal(ro, re)
        assert np.array_equal(po.view(np.uint32), pe.view(np.uint32))


def test_emul_i16_soft_clip():
    lanes, ticks = 16, 4
    arena, offsets = helpers.fuzz_grid(lanes, ticks, 200, lambda t, j, r: helpers.celt_toc(3, True), seed=9)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 960, i16=True)
    pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, 2, 960, i16=True)
    assert np.array_equal(so, se) and np.array_equal(ro, re)
    assert np.array_equal(po, pe)


def test_emul_pulse_vectors_bit_exact():
    """Decoded PVQ pulse vectors against the oracle's stage trace (north star: pulses bit-exact)."""
    import ctypes as C

    lib = helpers.oracle_lib()
    lanes = 12
    arena, offsets = helpers.fuzz_grid(lanes, 1, 180, lambda t, j, r: helpers.celt_toc(3, True), seed=21)
    _, _, _, iy, xs = helpers.emul_decode_grid(arena, offsets, lanes, 1, 2, 960, debug=True)
    for j in range(lanes):
        h = lib.orc_create(48000, 2, helpers.QUIRKS_MOUSIKI)
        lib.orc_trace_enable(h, 1)
        out = np.zeros(1920, dtype=np.float32)
        pkt = arena[offsets[j]:offsets[j + 1]]
        assert lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(out), 960, 0) == 960
        want = np.zeros(1920, dtype=np.int32)
        wx = np.zeros(1920, dtype=np.float32)
        assert lib.orc_trace_get(h, 0, helpers.ptr(want), 1920) == 1920
        assert lib.orc_trace_get(h, 1, helpers.ptr(wx), 1920) == 1920
        lib.orc_destroy(h)
        assert np.array_equal(iy[j].reshape(-1), want)
        assert np.array_equal(xs[j].reshape(-1).view(np.uint32), wx.view(np.uint32))


def test_emul_cres_kat():
    """The reference's final_range known answer through the product's parse logic (config 16, NB 2.5 ms)."""
    n = 65536
    i = np.arange(n, dtype=np.uint32)
    arena = np.empty((n, 4), dtype=np.uint8)
    arena[:, 0] = 16 << 3
    arena[:, 1] = i >> 8
    arena[:, 2] = i & 255
    arena[:, 3] = 255
    offsets = (np.arange(n + 1, dtype=np.uint32) * 4).astype(np.uint32)
    _, samples, rng, _, _ = helpers.emul_decode_grid(arena.reshape(-1), offsets, n, 1, 1, 120)
    assert (samples == 120).all()
    assert int(rng.astype(np.uint64).sum() & 0xFFFFFFFF) == 116290185
