"""bd_parse_noalloc (include/mousiki_b200.h): the allocation-free entry point for callers that cannot call cudaMalloc. It has
no counterpart in the reference; its scratch is a file-scope device array.
Checked through the C ABI against the existing entry point bd_decode_f32 on the same packets, each as the first packet of a
fresh stream: sample counts and final_range are integers, so the tolerance is equality. Two calls in a row (the second one
smaller and with other packets: the shared array still holds the first call's records), then the size limit.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import helpers


def _packets(n, seed, stereo):
    """CELT-only code-0 packets of every frame size and bandwidth, a few one-byte frames, a SILK packet and a code-3 packet
    (both outside what the entry point covers)."""
    rng = np.random.default_rng(seed)
    pk = []
    for k in range(n):
        ln = int(rng.integers(2, 200))
        p = rng.integers(0, 256, size=ln + 1, dtype=np.uint8)
        p[0] = helpers.celt_toc(int(rng.integers(0, 4)), bool(stereo and rng.integers(0, 2)), int(rng.integers(0, 4)))
        if k % 29 == 7:
            p = p[:2]                                  # a frame of one byte: concealed
        if k % 37 == 11:
            p[0] = helpers.silk_toc(2, 1, False)       # not CELT-only
        if k % 41 == 13:
            p[0] |= 3                                  # more than one frame
        pk.append(p)
    off = np.zeros(n + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(p) for p in pk])
    return np.concatenate(pk).astype(np.uint8), off


def _existing(arena, offsets, ch, cap):
    """bd_decode_f32, one fresh stream per packet."""
    from mousiki_b200 import BatchDecoder

    n = len(offsets) - 1
    with BatchDecoder(n, 48000, ch) as dec:
        _, samples, rng = dec.decode_float(arena, offsets, np.arange(n, dtype=np.uint32), cap)
    return samples, rng


@pytest.mark.gpu
def test_gpu_parse_noalloc_matches_the_decoder_on_two_calls_in_a_row():
    import torch

    import mousiki_b200
    from mousiki_b200.batch_decoder import load_library

    lib = load_library()
    lib.bd_parse_noalloc.restype = C.c_int
    lib.bd_parse_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p,
                                     C.c_void_p]
    dev = torch.device("cuda", 0)
    cap = 960
    checked = 0
    for n, seed, ch in ((512, 1, 2), (200, 2, 1)):
        arena, offsets = _packets(n, seed, ch == 2)
        want_s, want_r = _existing(arena, offsets, ch, cap)
        d_arena = torch.from_numpy(arena).to(dev)
        d_off = torch.from_numpy(offsets.view(np.int32)).to(dev)
        d_s = torch.full((n,), -77, dtype=torch.int32, device=dev)
        d_r = torch.full((n,), -1, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        rc = lib.bd_parse_noalloc(0, d_arena.data_ptr(), d_off.data_ptr(), n, ch, helpers.QUIRKS_MOUSIKI, cap, d_s.data_ptr(), d_r.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        torch.cuda.synchronize()
        got_s = d_s.cpu().numpy()
        got_r = d_r.cpu().numpy().view(np.uint32)
        toc = arena[offsets[:-1]]
        covered = ((toc & 0x80) != 0) & ((toc & 3) == 0)
        assert covered.sum() > 0.9 * n
        assert (got_s[~covered] == -5).all() and (got_r[~covered] == 0).all()
        assert np.array_equal(got_s[covered], want_s[covered]), np.nonzero(got_s[covered] != want_s[covered])
        assert np.array_equal(got_r[covered], want_r[covered])
        assert (want_s[covered] > 0).sum() > 0.8 * n and (got_r[covered] != 0).sum() > 0.8 * n
        checked += int(covered.sum())
    assert checked > 600
    # beyond the static scratch: refused, nothing launched
    rc = lib.bd_parse_noalloc(0, d_arena.data_ptr(), d_off.data_ptr(), 513, 2, helpers.QUIRKS_MOUSIKI, cap, d_s.data_ptr(), d_r.data_ptr(), None)
    assert rc == -102
    assert mousiki_b200 is not None
