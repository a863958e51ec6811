"""GPU parity tests proper (-m gpu, run on the B200 box): the CUDA path, called through the C ABI, against the
oracle on the same seeded inputs. final_range / sample counts / pulse vectors: bit-exact. PCM: within the north
star's tolerance (<= 1 LSB after i16 conversion); f32 max error and SNR are printed."""
import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu

TOL_F32 = 1.0 / 32768.0  # 1 LSB at i16 (BASELINE.json north_star)


def _snr_db(ref, got):
    num = float((ref.astype(np.float64) ** 2).sum())
    den = float(((ref.astype(np.float64) - got.astype(np.float64)) ** 2).sum())
    return float("inf") if den == 0 else 10 * np.log10(max(num, 1e-30) / den)


def _decode_gpu(arena, offsets, lanes, ticks, ch, frame_size=960, i16=False, capture=False, **kw):
    from mousiki_b200 import BatchDecoder

    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    with BatchDecoder(lanes, 48000, ch, **kw) as dec:
        if i16:
            return dec.decode(arena, offsets, ids, frame_size)
        return dec.decode_float(arena, offsets, ids, frame_size, capture=capture)


CASES = [
    (3, True, 2, 160, 3),
    (3, False, 1, 80, 3),
    (0, True, 2, 40, 3),
    (1, False, 2, 60, 2),
    (2, True, 1, 100, 1),
    (3, True, 2, (2, 400), 3),
    (2, False, 1, (2, 60), 0),
]


@pytest.mark.parametrize("lm,stereo,ch,payload,bw", CASES)
def test_gpu_matches_oracle(lm, stereo, ch, payload, bw):
    lanes, ticks = 96, 6
    arena, offsets = helpers.fuzz_grid(lanes, ticks, payload, lambda t, j, r: helpers.celt_toc(lm, stereo, bw), seed=100 + lm * 7 + ch)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
    pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, ch)
    assert np.array_equal(so, sg)
    assert np.array_equal(ro, rg), "final_range must be bit-exact"
    err = float(np.abs(po - pg).max())
    print(f"lm={lm} stereo={stereo} ch={ch}: f32 max err {err:.3g}, SNR {_snr_db(po, pg):.1f} dB, bit-identical {np.array_equal(po, pg)}")
    assert err <= TOL_F32
    i16o = np.clip(np.rint(po * 32768.0), -32768, 32767)
    i16g = np.clip(np.rint(pg * 32768.0), -32768, 32767)
    assert np.abs(i16o - i16g).max() <= 1


@pytest.mark.parametrize("parallel", [False, True])
def test_gpu_mixed_shapes_all_ticks_in_one_call(parallel, monkeypatch):
    """`parallel`: the opt-in all-at-once band pass of small calls (k_bands, MOUSIKI_B200_PARALLEL_BANDS=1), exact for streams
    that stay CELT-only; the default is the per-stream pass (k_bands_hyb)."""
    if parallel:
        monkeypatch.setenv("MOUSIKI_B200_PARALLEL_BANDS", "1")
    lanes, ticks = 128, 8

    def toc(t, j, rng):
        return helpers.celt_toc(int(rng.integers(0, 4)), bool(rng.integers(0, 2)), int(rng.integers(0, 4)))

    arena, offsets = helpers.fuzz_grid(lanes, ticks, (2, 300), toc, seed=5)
    for ch in (1, 2):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
        pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, ch)
        assert np.array_equal(so, sg) and np.array_equal(ro, rg)
        assert float(np.abs(po - pg).max()) <= TOL_F32


def test_gpu_state_carries_across_calls():
    """One packet per stream per call (the live-server shape) must equal all ticks in one call."""
    from mousiki_b200 import BatchDecoder

    lanes, ticks = 64, 5
    arena, offsets = helpers.fuzz_grid(lanes, ticks, 120, lambda t, j, r: helpers.celt_toc(3, True), seed=77)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 960)
    ids = np.arange(lanes, dtype=np.uint32)
    with BatchDecoder(lanes, 48000, 2) as dec:
        for t in range(ticks):
            lo, hi = t * lanes, (t + 1) * lanes
            sub_off = (offsets[lo:hi + 1] - offsets[lo]).astype(np.uint32)
            sub_arena = arena[offsets[lo]:offsets[hi]]
            pg, sg, rg = dec.decode_float(sub_arena, sub_off, ids, 960)
            assert np.array_equal(sg, so[lo:hi]) and np.array_equal(rg, ro[lo:hi])
            assert float(np.abs(pg - po[lo:hi]).max()) <= TOL_F32
        info = dec.stream_info(3)
        assert info["final_range"] == int(ro[(ticks - 1) * lanes + 3])
        assert info["last_packet_duration"] == 960


def test_gpu_multi_pass_equals_single_pass():
    """Scratch-bounded passes (ticks split) give the same result as one pass."""
    lanes, ticks = 64, 6
    arena, offsets = helpers.fuzz_grid(lanes, ticks, 100, lambda t, j, r: helpers.celt_toc(2, True), seed=31)
    a = _decode_gpu(arena, offsets, lanes, ticks, 2)
    b = _decode_gpu(arena, offsets, lanes, ticks, 2, max_frames_per_pass=2 * lanes)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_gpu_shuffled_stream_order_and_subset():
    """Arbitrary packet order / stream subset through the general scheduling path."""
    from mousiki_b200 import BatchDecoder

    lanes, ticks = 40, 4
    arena, offsets = helpers.fuzz_grid(lanes, ticks, 90, lambda t, j, r: helpers.celt_toc(3, False), seed=13)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 1, 960)
    rng = np.random.default_rng(0)
    # keep per-stream order, interleave streams randomly; stream ids are a permutation inside a larger id space
    order = np.argsort(rng.random(lanes * ticks) + np.repeat(np.arange(ticks), lanes) * 1.0, kind="stable")
    perm = rng.permutation(100)[:lanes].astype(np.uint32)
    lens = np.diff(offsets.astype(np.int64))
    new_off = np.zeros(lanes * ticks + 1, dtype=np.uint32)
    new_off[1:] = np.cumsum(lens[order])
    new_arena = np.concatenate([arena[offsets[p]:offsets[p + 1]] for p in order])
    ids = perm[order % lanes]
    with BatchDecoder(100, 48000, 1) as dec:
        pg, sg, rg = dec.decode_float(new_arena, new_off, ids, 960)
    assert np.array_equal(sg, so[order]) and np.array_equal(rg, ro[order])
    assert float(np.abs(pg - po[order]).max()) <= TOL_F32


def test_gpu_i16_soft_clip():
    lanes, ticks = 64, 4
    arena, offsets = helpers.fuzz_grid(lanes, ticks, 200, lambda t, j, r: helpers.celt_toc(3, True), seed=9)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 960, i16=True)
    pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, 2, i16=True)
    assert np.array_equal(so, sg) and np.array_equal(ro, rg)
    assert np.abs(po.astype(np.int32) - pg.astype(np.int32)).max() <= 1


def test_gpu_pulse_vectors_bit_exact():
    lib = helpers.oracle_lib()
    lanes = 48
    arena, offsets = helpers.fuzz_grid(lanes, 1, 180, lambda t, j, r: helpers.celt_toc(3, True), seed=21)
    _, _, _, iy, xs = _decode_gpu(arena, offsets, lanes, 1, 2, capture=True)
    for j in range(lanes):
        h = lib.orc_create(48000, 2, helpers.QUIRKS_MOUSIKI)
        lib.orc_trace_enable(h, 1)
        out = np.zeros(1920, dtype=np.float32)
        pkt = arena[offsets[j]:offsets[j + 1]]
        assert lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(out), 960, 0) == 960
        want = np.zeros(1920, dtype=np.int32)
        wx = np.zeros(1920, dtype=np.float32)
        lib.orc_trace_get(h, 0, helpers.ptr(want), 1920)
        lib.orc_trace_get(h, 1, helpers.ptr(wx), 1920)
        lib.orc_destroy(h)
        assert np.array_equal(iy[j].reshape(-1), want), "PVQ pulse vectors must be bit-exact"
        assert float(np.abs(xs[j].reshape(-1) - wx).max()) <= 1e-6


def test_gpu_cres_known_answer():
    """tests/test_opus_decode.rs:311-342 through the CUDA path: all four CELT configs."""
    from test_oracle_kat import CRES, _kat_corpus

    for cfg, want in CRES.items():
        arena, offsets = _kat_corpus(cfg)
        _, samples, rng = _decode_gpu(arena, offsets, 65536, 1, 2, frame_size=120)
        assert (samples == 120).all()
        assert int(rng.astype(np.uint64).sum() & 0xFFFFFFFF) == want


def test_gpu_unsupported_packets_report_per_packet_codes():
    """Odd packets are handled per packet and never poison the batch: a 40 ms SILK packet with a redundancy frame decodes (the
    frame goes to an auxiliary cell), a one-byte frame is concealed."""
    from mousiki_b200 import BatchDecoder, pack_packets

    good = bytes([helpers.celt_toc(3, True)]) + bytes(helpers.lcg_bytes(1, 100))
    silk = bytes([helpers.silk_toc(2, 2, False)]) + bytes(300)  # SILK WB 40 ms with a CELT redundancy frame
    code1 = bytes([helpers.celt_toc(2, False), 0x55])  # one payload byte: a concealment frame (src/opus_decoder.rs:491)
    arena, offsets = pack_packets([good, silk, code1, good])
    with BatchDecoder(4, 48000, 2) as dec:
        pcm, samples, rng = dec.decode_float(arena, offsets, np.arange(4, dtype=np.uint32), 1920)
    assert samples[0] == 960 and samples[3] == 960
    assert samples[1] == 1920
    # the one-byte frame is concealed (row F2): on a fresh decoder that is 10 ms of silence, and final_range 0
    assert samples[2] == 480 and rng[2] == 0 and not pcm[2, :960].any()
    assert np.array_equal(pcm[0], pcm[3]) and rng[0] == rng[3]


def test_gpu_buffer_too_small():
    from mousiki_b200 import BatchDecoder, pack_packets

    pkt = bytes([helpers.celt_toc(3, True)]) + bytes(helpers.lcg_bytes(1, 100))
    arena, offsets = pack_packets([pkt])
    with BatchDecoder(1, 48000, 2) as dec:
        _, samples, _ = dec.decode_float(arena, offsets, np.zeros(1, dtype=np.uint32), 480)
    assert samples[0] == -2  # OpusDecodeError::BufferTooSmall (src/opus_decoder.rs:1318-1320)


@pytest.mark.parametrize("i16", [False, True])
def test_gpu_large_call_takes_the_per_stream_band_pass_pcm(i16):
    """Calls with >= 3552 lanes (the bench's 10^4-lane configuration among them) run every frame through the per-stream
    band pass (k_bands_hyb, mb_api.cu `serial_all`): PCM, not only final_range, of such a call against the oracle,
    stereo 20 ms frames with the bench's payload size and a mix of frame sizes behind them."""
    lanes, ticks = 3584, 3

    def toc(t, j, rng):
        return helpers.celt_toc(3 if t < 2 else int(rng.integers(0, 4)), True)

    arena, offsets = helpers.fuzz_grid(lanes, ticks, 160, toc, seed=77)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 960, i16=i16)
    pg, sg, rg = _decode_gpu(arena, offsets, lanes, ticks, 2, i16=i16)
    assert np.array_equal(so, sg) and np.array_equal(ro, rg)
    if i16:
        assert int(np.abs(po.astype(np.int32) - pg.astype(np.int32)).max()) <= 1
    else:
        err = float(np.abs(po - pg).max())
        print(f"3584 lanes: f32 max err {err:.3g}, bit-identical {np.array_equal(po, pg)}")
        assert err <= TOL_F32
