"""The int16 output stage on the GPU: opus_pcm_soft_clip + float2int16
(src/opus.rs:16-140, src/opus_decoder.rs:1679-1712, src/celt/float_cast.rs:38-41).

Three forms of the same arithmetic must give the same int16 samples, bit for bit, and equal the oracle's:
  * default: k_deemph stores the int16 form of plain CELT-only frames itself and notes which packets hold a sample beyond +-1;
    k_clip_i16 only works on those (and on packets that inherit a clipper memory), correcting all regions of a row at once;
  * MOUSIKI_B200_CLIP_SERIAL=1: the clipper walks the regions of a row one after the other, like the reference's loop;
  * MOUSIKI_B200_FUSED_I16=0: every packet goes through k_clip_i16.
P-fuzz packets are loud (a quarter of them leave +-1, some by a factor of several hundred), so every branch of the clipper runs:
the carried-over coefficient, the ramp of a region that opens the row, rows with hundreds of regions."""
import numpy as np
import pytest

import helpers


def _gpu_i16(arena, offsets, ids, lanes, ch, cap, **kw):
    from mousiki_b200 import BatchDecoder

    with BatchDecoder(lanes, 48000, ch, **kw) as dec:
        return dec.decode(arena, offsets, ids, cap)


CASES = [
    ("fb 20 ms stereo", 3, True, 960, 160),
    ("fb 20 ms mono", 3, False, 960, 100),
    ("fb 10 ms stereo", 2, True, 480, 90),
    ("fb 2.5 ms stereo", 0, True, 120, 40),
]


@pytest.mark.gpu
@pytest.mark.parametrize("name,lm,stereo,cap,payload", CASES, ids=[c[0] for c in CASES])
def test_gpu_clipper_forms_agree_with_the_oracle(name, lm, stereo, cap, payload, monkeypatch):
    lanes, ticks, ch = 256, 9, 2 if stereo else 1
    arena, offsets = helpers.fuzz_grid(lanes, ticks, payload, lambda t, j, r: helpers.celt_toc(lm, stereo), seed=77 + lm)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, cap, i16=True)
    pf, _, _ = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, cap)
    loud = np.abs(pf).max(axis=1) > 1.0
    assert 0.05 < loud.mean() < 0.95, "the corpus must hold packets on both sides of the clipper"
    outs = {}
    for label, env in (("default", {}), ("serial", {"MOUSIKI_B200_CLIP_SERIAL": "1"}), ("unfused", {"MOUSIKI_B200_FUSED_I16": "0"})):
        for k in ("MOUSIKI_B200_CLIP_SERIAL", "MOUSIKI_B200_FUSED_I16"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        # three passes: the clipper of pass k runs beside pass k + 1, its notes are double-buffered like the frame meta
        pg, sg, rg = _gpu_i16(arena, offsets, ids, lanes, ch, cap, max_frames_per_pass=3 * lanes)
        assert np.array_equal(sg, so) and np.array_equal(rg, ro), label
        outs[label] = pg
    assert np.array_equal(outs["default"], outs["serial"])
    assert np.array_equal(outs["default"], outs["unfused"])
    assert np.array_equal(outs["default"], po), int(np.abs(outs["default"].astype(np.int32) - po.astype(np.int32)).max())


@pytest.mark.gpu
def test_gpu_clipper_forms_agree_on_packets_of_several_frames(monkeypatch):
    """Code-3 packets of three 10 ms frames: the packet is clipped as a whole (src/opus_decoder.rs:1352-1366), its frames sit in
    three ticks and share one row; k_deemph's notes of all three decide whether the clipper may skip it."""
    lanes, groups = 96, 4
    fa, fo = helpers.fuzz_grid(lanes, 3 * groups, 70, lambda t, j, r: helpers.celt_toc(2, True), seed=91)
    pkts = []
    for g in range(groups):
        for j in range(lanes):
            fr = [bytes(fa[fo[(3 * g + k) * lanes + j] + 1:fo[(3 * g + k) * lanes + j + 1]]) for k in range(3)]
            pkts.append(helpers.multiframe_packet(helpers.celt_toc(2, True), fr, 3))
    offsets = np.zeros(len(pkts) + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in pkts])
    arena = np.concatenate(pkts).astype(np.uint8)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), groups)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, groups, 2, 1440, i16=True)
    assert (so == 1440).all()
    outs = []
    for env in ({}, {"MOUSIKI_B200_CLIP_SERIAL": "1", "MOUSIKI_B200_FUSED_I16": "0"}):
        for k in ("MOUSIKI_B200_CLIP_SERIAL", "MOUSIKI_B200_FUSED_I16"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        pg, sg, rg = _gpu_i16(arena, offsets, ids, lanes, 2, 1440, max_frames_per_pass=6 * lanes)
        assert np.array_equal(sg, so) and np.array_equal(rg, ro)
        outs.append(pg)
    assert np.array_equal(outs[0], outs[1])
    assert np.array_equal(outs[0], po)
