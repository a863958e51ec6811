"""24-bit output and the decoder queries of the boundary:
  bd_decode_i24        = opus_decode24 / Decoder::decode_24bit (src/opus_decoder.rs:1718-1800, src/codec.rs:867)
  bd_get_stream_info   = Decoder::{final_range, last_packet_duration, bandwidth, pitch} (src/codec.rs:941-977,
                         src/opus_decoder.rs:1832-1849)
against the oracle, stream by stream, on CELT-only, SILK-only and hybrid streams. GPU (-m gpu) through the C ABI."""
import numpy as np
import pytest

import helpers
from test_hybrid_parity import hyb_toc

pytestmark = pytest.mark.gpu

CORPORA = [
    ("celt", lambda l, t: helpers.fuzz_grid(l, t, (20, 200), lambda a, b, r: helpers.celt_toc(int(r.integers(0, 4)), bool(r.integers(0, 2)), int(r.integers(0, 4))), seed=61), 2),
    ("silk", lambda l, t: helpers.silk_fuzz_grid(l, t, (20, 60), lambda k, r: helpers.silk_toc(int(r.integers(0, 3)), 1, bool(r.integers(0, 2))), seed=62, channels=2), 2),
    ("hybrid", lambda l, t: helpers.silk_fuzz_grid(l, t, (40, 160), lambda k, r: hyb_toc(int(r.integers(0, 2)), 1, False), seed=63, channels=1), 1),
]


@pytest.mark.parametrize("name,build,ch", CORPORA, ids=[c[0] for c in CORPORA])
def test_gpu_i24_and_queries_match_the_oracle(name, build, ch):
    from mousiki_b200 import BatchDecoder

    lanes, ticks = 24, 4
    arena, offsets = build(lanes, ticks)
    lib = helpers.oracle_lib()
    want_pcm = np.zeros((lanes * ticks, 960 * ch), dtype=np.int32)
    want = {}
    for j in range(lanes):
        h = lib.orc_create(48000, ch, helpers.QUIRKS_MOUSIKI)
        for t in range(ticks):
            p = t * lanes + j
            pkt = np.ascontiguousarray(arena[offsets[p]:offsets[p + 1]])
            r = lib.orc_decode_i24(h, helpers.ptr(pkt), len(pkt), helpers.ptr(want_pcm[p]), 960, 0)
            assert r > 0
        lib.orc_final_range.restype = helpers.C.c_uint32
        want[j] = (lib.orc_final_range(h), lib.orc_last_packet_duration(h), lib.orc_bandwidth(h), lib.orc_pitch(h))
        lib.orc_destroy(h)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    with BatchDecoder(lanes, 48000, ch) as dec:
        info0 = dec.stream_info(0)
        assert info0["bandwidth"] == 0 and info0["pitch"] == 0          # nothing decoded yet (Decoder::bandwidth -> None)
        pcm, samples, rng = dec.decode_24bit(arena, offsets, ids, 960)
        assert (samples > 0).all()
        assert pcm.dtype == np.int32 and np.abs(pcm).max() <= 8388608
        assert int(np.abs(pcm.astype(np.int64) - want_pcm).max()) <= 256, "within 1 LSB at 16 bits (= 256 at 24 bits)"
        print(f"{name}: i24 max |err| {int(np.abs(pcm.astype(np.int64) - want_pcm).max())}, identical {np.array_equal(pcm, want_pcm)}")
        for j in range(lanes):
            info = dec.stream_info(j)
            got = (info["final_range"], info["last_packet_duration"], info["bandwidth"], info["pitch"])
            assert got == want[j], (name, j, got, want[j])
