"""The oracle against the reference's own known answers for this path. CPU only."""
import ctypes as C

import numpy as np

import helpers

# tests/test_opus_decode.rs:311-342 — sum of final_range over packets {cfg<<3, i>>8, i&255, 255}, i in [0, 65536)
CRES = {16: 116290185, 20: 2172123586, 24: 2172123586, 28: 2172123586}


def _kat_corpus(cfg):
    n = 65536
    i = np.arange(n, dtype=np.uint32)
    arena = np.empty((n, 4), dtype=np.uint8)
    arena[:, 0] = cfg << 3
    arena[:, 1] = i >> 8
    arena[:, 2] = i & 255
    arena[:, 3] = 255
    offsets = (np.arange(n + 1, dtype=np.uint32) * 4).astype(np.uint32)
    return arena.reshape(-1), offsets


def test_cres_final_range_sum_oracle():
    for cfg, want in CRES.items():
        arena, offsets = _kat_corpus(cfg)
        # final_range does not depend on decoder state: 65536 one-packet streams give the reference's serial sum
        _, samples, rng = helpers.oracle_decode_grid(arena, offsets, 65536, 1, 2, 960)
        assert (samples == 120).all()
        assert int(rng.astype(np.uint64).sum() & 0xFFFFFFFF) == want, cfg


def test_cres_serial_state_oracle():
    """Same KAT through ONE decoder in sequence (as the reference runs it), on a prefix, to show the sum is
    order/state independent."""
    arena, offsets = _kat_corpus(16)
    n = 4096
    _, s1, r1 = helpers.oracle_decode_grid(arena, offsets, 1, n, 2, 960, threads=1)
    _, s2, r2 = helpers.oracle_decode_grid(arena, offsets, n, 1, 2, 960)
    assert np.array_equal(r1, r2) and np.array_equal(s1, s2)


# (the symbol walk of src/range.rs:659-695 and the EcDec cases of src/celt/entdec.rs:327-377: tests/test_unit_kat.py)


def test_every_celt_toc_decodes():
    """tests/test_opus_decode.rs:220-267 — every TOC x small payload must decode or fail cleanly (CELT rows)."""
    lib = helpers.oracle_lib()
    h = lib.orc_create(48000, 2, helpers.QUIRKS_MOUSIKI)
    out = np.zeros(5760 * 2, dtype=np.float32)
    for cfg in range(16, 32):
        for stereo in (0, 1):
            for fill in (0, 1, 127, 255):
                pkt = np.array([(cfg << 3) | (stereo << 2), fill, fill, fill, fill], dtype=np.uint8)
                ret = lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(out), 5760, 0)
                assert ret == 120 << (cfg & 3), (cfg, stereo, fill, ret)
    lib.orc_destroy(h)


def test_soft_clip_known_answers():
    """src/opus.rs:157-208 — the four soft-clip cases of the reference (in range, a clipped peak, a carried
    non-linearity that ends at the zero crossing, a ramp before the first zero crossing)."""
    import ctypes as C
    lib = helpers.oracle_lib()
    lib.orc_pcm_soft_clip.restype = None

    def clip(pcm, n, c, mem):
        pcm, mem = np.array(pcm, dtype=np.float32), np.array(mem, dtype=np.float32)
        lib.orc_pcm_soft_clip(C.c_void_p(pcm.ctypes.data), n, c, C.c_void_p(mem.ctypes.data))
        return pcm, mem

    pcm, mem = clip([0.1, -0.6, 0.9, 0.3], 2, 2, [0.5, -0.25])
    assert np.array_equal(pcm, np.array([0.1, -0.6, 0.9, 0.3], dtype=np.float32)) and np.array_equal(mem, [0.0, 0.0])
    pcm, mem = clip([2.0], 1, 1, [0.0])
    assert abs(pcm[0] - 0.99999976) < 1e-7 and abs(mem[0] + 0.25000006) < 1e-7
    pcm, mem = clip([0.5, 0.25, -0.25], 3, 1, [-0.25])
    assert np.abs(pcm - np.array([0.4375, 0.234375, -0.25])).max() < 1e-6 and mem[0] == 0.0
    pcm, mem = clip([0.5, 1.5, 1.7, -0.4], 4, 1, [0.0])
    assert np.abs(pcm[[0, 1, 3]] - np.array([0.46972317, 0.9550162, -0.4])).max() < 1e-6 and abs(pcm[2] - 1.0) < 2e-6 and mem[0] == 0.0
