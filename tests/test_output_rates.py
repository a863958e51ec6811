"""Decoder output rates below 48 kHz: `Decoder::new(8000 | 12000 | 16000 | 24000, ..)`,
src/opus_decoder.rs:1550-1556.

Built: the CELT side — `downsample = 48000 / Fs` in synthesis and de-emphasis (src/celt/celt_decoder.rs:584-776, 2232-2297;
denormalise_bands clears the bins from N / downsample on, src/celt/bands.rs:4198-4204; the de-emphasis filter runs over every
48 kHz sample and keeps every downsample-th one) for coded CELT-only frames, single and multi-frame packets, f32 / int16 / 24-bit;
and the SILK side — every mode of the resampler in the k_silk lane (src/silk/resampler.rs:87-176, 279-369): copy, 2x, the
interpolating up-sampler with the ratio's own step, and the down-sampling kernel where the output rate lies below the internal
one (resampler_private_ar2.rs, resampler_private_down_fir.rs: a WB packet in a 12 / 8 kHz decoder). Mode switches (redundancy
frames, fades, concealed transition frames) and in-band FEC work as at 48 kHz. Hybrid streams: SILK WB through the resampler, the
CELT bands down-sampled and added onto it (nothing of them is left below 16 kHz). Lost packets
of CELT-only and SILK-only streams are concealed at those rates like at 48 kHz (the concealment runs at the codec's own rate; only
the sample accounting changes).

CPU: host emulation of the kernel bodies against the oracle, bit for bit. GPU (-m gpu): the CUDA path through the C ABI."""
import numpy as np
import pytest

import helpers

RATES = [24000, 16000, 12000, 8000]
SHAPES = [
    ("fb 20 ms stereo", helpers.celt_toc(3, True), 2, 960, (40, 160)),
    ("fb 10 ms mono", helpers.celt_toc(2, False), 1, 480, (30, 90)),
    ("swb 5 ms stereo", helpers.celt_toc(1, True, bandwidth=2), 2, 240, (20, 60)),
    ("nb 2.5 ms mono in a stereo decoder", helpers.celt_toc(0, False, bandwidth=0), 2, 120, (8, 30)),
]


@pytest.mark.parametrize("fs", RATES)
@pytest.mark.parametrize("name,toc,ch,frame48,payload", SHAPES, ids=[s[0] for s in SHAPES])
def test_emul_output_rates(fs, name, toc, ch, frame48, payload):
    lanes, ticks = 6, 10
    frame = frame48 * fs // 48000
    arena, offsets = helpers.fuzz_grid(lanes, ticks, payload, lambda t, j, r: toc, seed=7)
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=i16, fs=fs)
        pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=i16, fs=fs)
        assert (so == frame).all() and np.array_equal(so, se) and np.array_equal(ro, re)
        assert np.array_equal(po, pe), (name, fs, i16)


def _gpu(arena, offsets, ids, lanes, ch, cap, fs, i16=False, flags=None, **kw):
    from mousiki_b200 import BatchDecoder

    with BatchDecoder(lanes, fs, ch, **kw) as dec:
        return dec.decode(arena, offsets, ids, cap, flags=flags) if i16 else dec.decode_float(arena, offsets, ids, cap, flags=flags)


@pytest.mark.gpu
@pytest.mark.parametrize("fs", RATES)
def test_gpu_output_rates(fs):
    lanes, ticks = 128, 10
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    for name, toc, ch, frame48, payload in SHAPES:
        frame = frame48 * fs // 48000
        arena, offsets = helpers.fuzz_grid(lanes, ticks, payload, lambda t, j, r: toc, seed=23)
        for i16 in (False, True):
            po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=i16, fs=fs)
            pg, sg, rg = _gpu(arena, offsets, ids, lanes, ch, frame, fs, i16=i16, max_frames_per_pass=4 * lanes)
            assert np.array_equal(so, sg) and np.array_equal(ro, rg), (name, fs)
            assert np.array_equal(po, pg), (name, fs, i16)


@pytest.mark.gpu
def test_gpu_output_rate_packets_of_several_frames_and_mixed_sizes():
    """Code-3 packets of three 10 ms frames and a stream that changes its frame size every packet, at 16 kHz."""
    lanes, groups, fs = 48, 4, 16000
    fa, fo = helpers.fuzz_grid(lanes, 3 * groups, 70, lambda t, j, r: helpers.celt_toc(2, True), seed=29)
    pkts = []
    for g in range(groups):
        for j in range(lanes):
            fr = [bytes(fa[fo[(3 * g + k) * lanes + j] + 1:fo[(3 * g + k) * lanes + j + 1]]) for k in range(3)]
            pkts.append(helpers.multiframe_packet(helpers.celt_toc(2, True), fr, 3))
    offsets = np.zeros(len(pkts) + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in pkts])
    arena = np.concatenate(pkts).astype(np.uint8)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), groups)
    cap = 1440 * fs // 48000
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, groups, 2, cap, i16=i16, fs=fs)
        assert (so == cap).all()
        pg, sg, rg = _gpu(arena, offsets, ids, lanes, 2, cap, fs, i16=i16)
        assert np.array_equal(so, sg) and np.array_equal(ro, rg) and np.array_equal(po, pg), i16
    arena, offsets = helpers.fuzz_grid(lanes, 12, (20, 120), lambda t, j, r: helpers.celt_toc(int(r.integers(0, 4)), True), seed=31)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), 12)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, 12, 2, 320, fs=fs)
    pg, sg, rg = _gpu(arena, offsets, ids, lanes, 2, 320, fs)
    assert np.array_equal(so, sg) and np.array_equal(ro, rg) and np.array_equal(po, pg)


SILK_SHAPES = [
    ("wb 20 ms mono", helpers.silk_toc(2, 1, False), 1, (20, 60), 960),
    ("wb 20 ms stereo", helpers.silk_toc(2, 1, True), 2, (25, 70), 960),
    ("nb 10 ms mono", helpers.silk_toc(0, 0, False), 1, (8, 30), 480),
    ("mb 40 ms stereo", helpers.silk_toc(1, 2, True), 2, (40, 100), 1920),
    ("nb 60 ms mono in a stereo decoder", helpers.silk_toc(0, 3, False), 2, (40, 90), 2880),
]


def _silk_fits(toc, fs):
    """(every SILK bandwidth decodes at every output rate: below the internal rate the resampler's down-sampling kernel runs,
    resampler_private_down_fir.rs — 16 -> 12 kHz, 12 -> 8 kHz, 16 -> 8 kHz)"""
    return True


@pytest.mark.parametrize("fs", RATES)
def test_emul_silk_only_output_rates(fs):
    lanes, ticks = 4, 6
    ran = 0
    for name, toc, ch, payload, frame48 in SILK_SHAPES:
        if not _silk_fits(toc, fs):
            continue
        frame = frame48 * fs // 48000
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, lambda k, r: toc, seed=11, channels=ch)
        for i16 in (False, True):
            po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=i16, fs=fs)
            pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=i16, fs=fs)
            assert (so == frame).all() and np.array_equal(so, se) and np.array_equal(ro, re), (name, fs)
            assert np.array_equal(po, pe), (name, fs, i16)
        ran += 1
    assert ran >= 2
    # a stream that changes its SILK bandwidth and frame size from packet to packet (the resampler is re-initialised on a rate change)
    bws = [0, 1, 2]
    arena, offsets = helpers.silk_fuzz_grid(lanes, 10, (20, 80), lambda k, r: helpers.silk_toc(int(r.choice(bws)), int(r.integers(0, 2)), True),
                                            seed=13, channels=2)
    cap = 960 * fs // 48000
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, 10, 2, cap, fs=fs)
    pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, 10, 2, cap, fs=fs)
    assert (so > 0).all() and np.array_equal(so, se) and np.array_equal(ro, re) and np.array_equal(po, pe)


@pytest.mark.gpu
@pytest.mark.parametrize("fs", RATES)
def test_gpu_silk_only_output_rates(fs):
    lanes, ticks = 64, 6
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    for name, toc, ch, payload, frame48 in SILK_SHAPES:
        if not _silk_fits(toc, fs):
            continue
        frame = frame48 * fs // 48000
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, lambda k, r: toc, seed=17, channels=ch)
        for i16 in (False, True):
            po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, frame, i16=i16, fs=fs)
            pg, sg, rg = _gpu(arena, offsets, ids, lanes, ch, frame, fs, i16=i16, max_frames_per_pass=2 * lanes)
            assert np.array_equal(so, sg) and np.array_equal(ro, rg), (name, fs)
            assert np.array_equal(po, pg), (name, fs, i16)
    bws = [0, 1, 2]
    arena, offsets = helpers.silk_fuzz_grid(lanes, 8, (20, 80), lambda k, r: helpers.silk_toc(int(r.choice(bws)), int(r.integers(0, 2)), True),
                                            seed=19, channels=2)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), 8)
    cap = 960 * fs // 48000
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, 8, 2, cap, fs=fs)
    pg, sg, rg = _gpu(arena, offsets, ids, lanes, 2, cap, fs)
    assert np.array_equal(so, sg) and np.array_equal(ro, rg) and np.array_equal(po, pg)


def _walk_toc(fs):
    """CELT-only, SILK-only and hybrid packets of every size and bandwidth: a stream that switches mode freely"""
    from test_hybrid_parity import hyb_toc
    max_bw = 2

    def toc(k, r):
        m = int(r.integers(0, 10))
        if m < 4:
            return helpers.celt_toc(int(r.integers(1, 4)), bool(r.integers(0, 2)), int(r.integers(0, 4)))
        if m < 9:
            return helpers.silk_toc(int(r.integers(0, max_bw + 1)), int(r.integers(0, 2)), bool(r.integers(0, 2)))
        return hyb_toc(r.integers(0, 2), r.integers(0, 2), r.integers(0, 2))
    return toc


@pytest.mark.parametrize("fs", RATES)
def test_emul_every_mode_switch_decodes_below_48k(fs):
    """Random walks over the modes, with and without redundancy frames (src/opus_decoder.rs:575-640, 746-819, 868-1112): the fades
    step through the window by 48000 / Fs (smooth_fade, :100-135), side rows and transition frames hold output-rate samples."""
    lanes, ticks = 8, 14
    cap = 960 * fs // 48000
    for ch, seed in ((2, 91), (1, 92)):
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (10, 140), _walk_toc(fs), seed=seed, channels=ch, redundancy=True)
        for i16 in (False, True):
            po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, cap, i16=i16, fs=fs)
            pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, lanes, ticks, ch, cap, i16=i16, fs=fs)
            assert (so > 0).all() and np.array_equal(so, se) and np.array_equal(ro, re)
            assert np.array_equal(po, pe), (fs, ch, i16)


@pytest.mark.gpu
@pytest.mark.parametrize("fs", RATES)
def test_gpu_every_mode_switch_decodes_below_48k(fs):
    lanes, ticks = 48, 14
    cap = 960 * fs // 48000
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    for ch, seed in ((2, 97), (1, 98)):
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (10, 140), _walk_toc(fs), seed=seed, channels=ch, redundancy=True)
        for i16 in (False, True):
            po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, cap, i16=i16, fs=fs)
            pg, sg, rg = _gpu(arena, offsets, ids, lanes, ch, cap, fs, i16=i16, max_frames_per_pass=5 * lanes)
            assert (so > 0).all() and np.array_equal(so, sg) and np.array_equal(ro, rg), (fs, ch)
            assert np.array_equal(po, pg), (fs, ch, i16)


def _oracle_lossy(arena, offsets, lost, lanes, ticks, ch, n, fs, i16):
    lib = helpers.oracle_lib()
    pcm = np.zeros((lanes * ticks, n * ch), dtype=np.int16 if i16 else np.float32)
    samples = np.zeros(lanes * ticks, dtype=np.int32)
    rng = np.zeros(lanes * ticks, dtype=np.uint32)
    fn = lib.orc_decode_i16 if i16 else lib.orc_decode_float
    for j in range(lanes):
        h = lib.orc_create(fs, ch, helpers.QUIRKS_MOUSIKI)
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


def _lossy_streams(kind, lanes, ticks, fs, seed):
    """-> (arena, offsets, lost, ch, frame48): CELT FB 20 ms stereo, or SILK NB 20 ms mono in a stereo decoder, with isolated losses and bursts"""
    from test_plc_parity import lossy_grid
    from test_silk_plc_parity import loss_pattern

    if kind == "celt":
        arena, offsets, lost = lossy_grid(lanes, ticks, 3, True, seed=seed)
        return arena, offsets, lost, 2, 960
    if kind == "hybrid":     # FB 20 ms stereo: SILK WB at the output rate + CELT bands 17.. down-sampled, added in k_deemph
        from test_hybrid_parity import hyb_toc
        arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (40, 200), lambda k, r: hyb_toc(1, 1, True), seed=seed, channels=2)
        return arena, offsets, loss_pattern(lanes, ticks, seed=seed + 1), 2, 960
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, (15, 50), lambda k, r: helpers.silk_toc(0, 1, False), seed=seed, channels=2)
    return arena, offsets, loss_pattern(lanes, ticks, seed=seed + 1), 2, 960


@pytest.mark.parametrize("kind,fs", [("celt", 24000), ("celt", 8000), ("silk", 24000), ("silk", 8000), ("hybrid", 24000), ("hybrid", 16000),
                                     ("hybrid", 8000)])
def test_emul_concealment_below_48k(kind, fs):
    from test_plc_parity import virtual_grid

    lanes, ticks = 6, 14
    arena, offsets, lost, ch, frame48 = _lossy_streams(kind, lanes, ticks, fs, 51)
    n = frame48 * fs // 48000
    va, vo, vf = virtual_grid(arena, offsets, lost, frame48)      # the host front end hands the device 48 kHz sample counts
    lib = helpers.emul_lib()
    for i16 in (False, True):
        po, so, ro = _oracle_lossy(arena, offsets, lost, lanes, ticks, ch, n, fs, i16)
        lib.emu_set_flags(helpers.ptr(vf))
        try:
            pe, se, re, _, _ = helpers.emul_decode_grid(va, vo, lanes, ticks, ch, n, i16=i16, fs=fs)
        finally:
            lib.emu_set_flags(None)
        assert (so == n).all() and np.array_equal(so, se) and np.array_equal(ro, re)
        assert np.array_equal(po, pe), (kind, fs, i16)
    assert lost.sum() > 10


@pytest.mark.gpu
@pytest.mark.parametrize("kind,fs", [("celt", 24000), ("celt", 16000), ("celt", 8000), ("silk", 24000), ("silk", 16000), ("silk", 8000),
                                     ("hybrid", 24000), ("hybrid", 16000), ("hybrid", 12000)])
def test_gpu_concealment_below_48k(kind, fs):
    """Through the host front end: flags[p] & 2 = decode(None); the packet's bytes are ignored."""
    lanes, ticks = 48, 14
    arena, offsets, lost, ch, frame48 = _lossy_streams(kind, lanes, ticks, fs, 61)
    n = frame48 * fs // 48000
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    flags = (lost.astype(np.uint8) << 1)
    for i16 in (False, True):
        po, so, ro = _oracle_lossy(arena, offsets, lost, lanes, ticks, ch, n, fs, i16)
        pg, sg, rg = _gpu(arena, offsets, ids, lanes, ch, n, fs, i16=i16, flags=flags, max_frames_per_pass=5 * lanes)
        assert np.array_equal(so, sg) and np.array_equal(ro, rg), (kind, fs)
        assert np.array_equal(po, pg), (kind, fs, i16)


def _oracle_fec(arena, offsets, flags, lanes, ticks, ch, n, fs, i16):
    lib = helpers.oracle_lib()
    pcm = np.zeros((lanes * ticks, n * ch), dtype=np.int16 if i16 else np.float32)
    samples = np.zeros(lanes * ticks, dtype=np.int32)
    rng = np.zeros(lanes * ticks, dtype=np.uint32)
    fn = lib.orc_decode_i16 if i16 else lib.orc_decode_float
    for j in range(lanes):
        h = lib.orc_create(fs, ch, helpers.QUIRKS_MOUSIKI)
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


def _fec_streams(kind, lanes, ticks, seed):
    """-> (arena, offsets, flags, ch, frame48): a lost packet is recovered from the NEXT packet's LBRR data (flag 1) when that one
    arrived, else concealed (flag 2) — Decoder::decode(.., fec = true), src/opus_decoder.rs:1240-1316"""
    from test_hybrid_parity import hyb_toc
    from test_silk_plc_parity import fec_grid, loss_pattern

    toc, payload = (helpers.silk_toc(2, 1, True), (40, 120)) if kind == "silk" else (hyb_toc(1, 1, True), (40, 200))
    arena, offsets = helpers.silk_fuzz_grid(lanes, ticks, payload, lambda k, r: toc, seed=seed, channels=2)
    lost = loss_pattern(lanes, ticks, seed=seed + 1, rate=0.3, burst=2)
    fa, fo, ff = fec_grid(arena, offsets, lost, lanes, ticks)
    return fa, fo, ff, 2, 960


@pytest.mark.parametrize("kind,fs", [("silk", 24000), ("silk", 16000), ("silk", 8000), ("hybrid", 24000), ("hybrid", 12000)])
def test_emul_fec_below_48k(kind, fs):
    lanes, ticks = 6, 12
    fa, fo, ff, ch, frame48 = _fec_streams(kind, lanes, ticks, 71)
    n = frame48 * fs // 48000
    pk = [np.array([frame48 & 255, frame48 >> 8], dtype=np.uint8) if ff[p] & 2 else fa[fo[p]:fo[p + 1]] for p in range(len(ff))]
    vo = np.zeros(len(pk) + 1, dtype=np.uint32)
    vo[1:] = np.cumsum([len(x) for x in pk])
    va = np.concatenate(pk).astype(np.uint8)
    lib = helpers.emul_lib()
    for i16 in (False, True):
        po, so, ro = _oracle_fec(fa, fo, ff, lanes, ticks, ch, n, fs, i16)
        lib.emu_set_flags(helpers.ptr(ff))
        try:
            pe, se, re, _, _ = helpers.emul_decode_grid(va, vo, lanes, ticks, ch, n, i16=i16, fs=fs)
        finally:
            lib.emu_set_flags(None)
        assert (so == n).all() and np.array_equal(so, se) and np.array_equal(ro, re)
        assert np.array_equal(po, pe), (kind, fs, i16)
    assert (ff == 1).sum() >= 6


@pytest.mark.gpu
@pytest.mark.parametrize("kind,fs", [("silk", 24000), ("silk", 12000), ("hybrid", 16000), ("hybrid", 8000)])
def test_gpu_fec_below_48k(kind, fs):
    lanes, ticks = 32, 12
    fa, fo, ff, ch, frame48 = _fec_streams(kind, lanes, ticks, 81)
    n = frame48 * fs // 48000
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    for i16 in (False, True):
        po, so, ro = _oracle_fec(fa, fo, ff, lanes, ticks, ch, n, fs, i16)
        pg, sg, rg = _gpu(fa, fo, ids, lanes, ch, n, fs, i16=i16, flags=ff)
        assert np.array_equal(so, sg) and np.array_equal(ro, rg), (kind, fs)
        assert np.array_equal(po, pg), (kind, fs, i16)
