"""Packets with more than one frame (TOC codes 1-3, padding).

The product cuts them into frames on the host (mousiki_b200/csrc/mb_host_packets.h, replacing opus_packet_parse_impl,
src/packet.rs:316-498, and the frame loop of decode_native, src/opus_decoder.rs:1296-1367); every frame then runs as a
code-0 packet in the next tick of its stream's lane and writes its stretch of the packet's PCM row.

CPU: the splitter itself (frame tables, rejections, against the oracle's return codes) and split -> emulated kernels ->
merge against the oracle decoding the original packets. GPU (-m gpu): the same through bd_decode_f32 / bd_decode_i16,
including the packet-wide soft clip and "a failing frame stops the packet"."""
import numpy as np
import pytest

import helpers
from test_hybrid_parity import hyb_toc

TOL = 1.0 / 32768.0


def _pack(pkts):
    offsets = np.zeros(len(pkts) + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in pkts])
    return np.concatenate(pkts).astype(np.uint8), offsets


def _oracle_code(pkt, ch=2, cap=5760):
    lib = helpers.oracle_lib()
    h = lib.orc_create(48000, ch, helpers.QUIRKS_MOUSIKI)
    out = np.zeros(cap * ch, dtype=np.float32)
    r = lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(out), cap, 0)
    lib.orc_destroy(h)
    return r


def _frames(seed, n, size):
    return [bytes(helpers.lcg_bytes(seed * 100 + k, size if np.isscalar(size) else size[k])) for k in range(n)]


def test_split_frame_tables():
    toc = helpers.celt_toc(1, True)  # 5 ms stereo
    cases = [
        (1, False, 0, _frames(1, 2, 40)),
        (2, False, 0, _frames(2, 2, [10, 70])),
        (2, False, 0, _frames(3, 2, [300, 20])),          # two-byte length
        (3, False, 0, _frames(4, 5, 33)),
        (3, True, 0, _frames(5, 4, [5, 260, 17, 90])),
        (3, False, 7, _frames(6, 3, 50)),
        (3, True, 300, _frames(7, 3, [12, 13, 14])),       # padding longer than 254 bytes
        (3, True, 0, _frames(8, 1, [25])),                 # code 3 with a single frame
    ]
    pkts = [helpers.multiframe_packet(toc, fr, code, vbr, pad) for code, vbr, pad, fr in cases]
    single = np.concatenate([[helpers.celt_toc(3, True)], helpers.lcg_bytes(9, 60)]).astype(np.uint8)
    pkts.append(single)
    arena, offsets = _pack(pkts)
    ids = np.arange(len(pkts), dtype=np.uint32)
    S = helpers.split_batch(arena, offsets, ids, 5760, 2)
    v = 0
    for p, (code, vbr, pad, fr) in enumerate(cases):
        assert S["first"][p] == v
        for k, f in enumerate(fr):
            got = bytes(S["arena"][S["offsets"][v]:S["offsets"][v + 1]])
            assert got == bytes([toc & ~3]) + f, (p, k)
            assert S["stream_ids"][v] == p
            assert S["flags"][v] == (4 if k else 0)
            assert S["out_map"][2 * v] == p and S["out_map"][2 * v + 1] == k * 240 * 2
            v += 1
        assert _oracle_code(pkts[p]) == len(fr) * 240       # the oracle reads the same framing
    assert bytes(S["arena"][S["offsets"][v]:S["offsets"][v + 1]]) == bytes(single) and S["flags"][v] == 0
    assert S["first"][-1] == v + 1


def test_split_rejections_follow_the_reference_codes():
    toc = helpers.celt_toc(3, False)
    bad = [
        np.array([toc | 1] + [1] * 5, dtype=np.uint8),                      # code 1, odd payload
        np.array([toc | 2, 200] + [0] * 20, dtype=np.uint8),                 # code 2, first length beyond the packet
        np.array([toc | 2], dtype=np.uint8),                                 # code 2, no length byte
        np.array([toc | 3], dtype=np.uint8),                                 # code 3, no count byte
        np.array([toc | 3, 0] + [0] * 8, dtype=np.uint8),                    # zero frames
        np.array([toc | 3, 7] + [0] * 70, dtype=np.uint8),                   # 7 x 20 ms > 120 ms
        np.array([toc | 3, 3] + [0] * 10, dtype=np.uint8),                   # CBR, payload not divisible by 3
        np.array([toc | 3, 0x42, 50] + [0] * 10, dtype=np.uint8),            # padding beyond the packet
        np.array([toc | 3, 0x83, 4, 250] + [0] * 30, dtype=np.uint8),        # VBR, second length beyond the packet
    ]
    arena, offsets = _pack(bad)
    S = helpers.split_batch(arena, offsets, np.arange(len(bad), dtype=np.uint32), 5760, 2)
    assert len(S["stream_ids"]) == len(bad)          # one reporting cell each
    for p, pkt in enumerate(bad):
        assert S["flags"][p] >> 4 == 4, p            # OPUS_INVALID_PACKET
        assert _oracle_code(pkt) == -4, p
    # capacity: 3 x 20 ms into a 40 ms row (src/opus_decoder.rs:1318-1320)
    big = helpers.multiframe_packet(toc, _frames(1, 3, 30), 3)
    arena, offsets = _pack([big])
    S = helpers.split_batch(arena, offsets, np.zeros(1, dtype=np.uint32), 1920, 2)
    assert S["flags"][0] >> 4 == 2 and _oracle_code(big, cap=1920) == -2


def _multiframe_grid(lanes, ticks, count, code, vbr, pad, single_arena, single_offsets):
    """Packs `count` consecutive single-frame packets of each lane (from a [ticks*count][lanes] corpus) into one packet."""
    pkts = []
    for t in range(ticks):
        for j in range(lanes):
            fr = []
            for k in range(count):
                q = (t * count + k) * lanes + j
                fr.append(bytes(single_arena[single_offsets[q] + 1:single_offsets[q + 1]]))
            toc = int(single_arena[single_offsets[(t * count) * lanes + j]])
            if not vbr:  # CBR: cut every frame of the packet to the shortest
                m = min(len(f) for f in fr)
                fr = [f[:m] for f in fr]
            pkts.append(helpers.multiframe_packet(toc, fr, code, vbr, pad))
    return _pack(pkts)


# name, single-frame corpus builder, frames per packet, code, vbr, padding, decoder channels, samples per frame
def _celt(lm, stereo):
    return lambda lanes, ticks, seed: helpers.fuzz_grid(lanes, ticks, (8, 120), lambda t, j, r: helpers.celt_toc(lm, stereo), seed=seed)


def _silk(bw, sc, stereo, ch):
    return lambda lanes, ticks, seed: helpers.silk_fuzz_grid(lanes, ticks, (20, 90), lambda k, r: helpers.silk_toc(bw, sc, stereo), seed=seed, channels=ch)


def _hyb(fb, ms20, stereo, ch):
    return lambda lanes, ticks, seed: helpers.silk_fuzz_grid(lanes, ticks, (40, 160), lambda k, r: hyb_toc(fb, ms20, stereo), seed=seed, channels=ch)


CASES = [
    ("celt 2x20ms code1", _celt(3, True), 2, 1, False, 0, 2, 960),
    ("celt 2x10ms code2", _celt(2, False), 2, 2, True, 0, 1, 480),
    ("celt 6x2.5ms vbr", _celt(0, True), 6, 3, True, 0, 2, 120),
    ("celt 4x5ms cbr pad", _celt(1, True), 4, 3, False, 9, 2, 240),
    ("silk wb 3x20ms vbr", _silk(2, 1, False, 1), 3, 3, True, 0, 1, 960),
    ("silk nb 2x10ms code2", _silk(0, 0, True, 2), 2, 2, True, 0, 2, 480),
    ("hybrid fb 2x20ms vbr pad", _hyb(1, 1, True, 2), 2, 3, True, 300, 2, 960),
    ("hybrid swb 3x10ms", _hyb(0, 0, False, 1), 3, 3, True, 0, 1, 480),
]


def _case_corpus(case, lanes, ticks, seed):
    name, build, count, code, vbr, pad, ch, per = case
    sa, so = build(lanes, ticks * count, seed)
    arena, offsets = _multiframe_grid(lanes, ticks, count, code, vbr, pad, sa, so)
    return arena, offsets, count, ch, per


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_emul_multiframe_matches_oracle(case):
    lanes, ticks = 5, 3
    arena, offsets, count, ch, per = _case_corpus(case, lanes, ticks, 7)
    cap = count * per
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, cap)
    assert np.all(so == cap), so
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    S = helpers.split_batch(arena, offsets, ids, cap, ch)
    nv = len(S["stream_ids"])
    assert nv == lanes * ticks * count
    # virtual packets are packet-major; the emulation wants a regular tick-major grid
    order = [(t * lanes + j) * count + k for t in range(ticks) for k in range(count) for j in range(lanes)]
    va, vo = _pack([S["arena"][S["offsets"][v]:S["offsets"][v + 1]] for v in order])
    pe, se, re, _, _ = helpers.emul_decode_grid(va, vo, lanes, ticks * count, ch, per)
    assert np.all(se == per)
    for t in range(ticks):
        for j in range(lanes):
            p = t * lanes + j
            rows = [(t * count + k) * lanes + j for k in range(count)]
            got = np.concatenate([pe[r, :per * ch] for r in rows])
            assert np.array_equal(got, po[p, :cap * ch]), (t, j)
            assert re[rows[-1]] == ro[p]


# ---------------------------------------------------------------------------------------------- GPU
def _decode_gpu(arena, offsets, ids, n_streams, ch, frame_size, i16, **kw):
    from mousiki_b200 import BatchDecoder

    with BatchDecoder(n_streams, 48000, ch, **kw) as dec:
        if i16:
            return dec.decode(arena, offsets, ids, frame_size)
        return dec.decode_float(arena, offsets, ids, frame_size)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_gpu_multiframe_matches_oracle(case):
    lanes, ticks = 40, 4
    arena, offsets, count, ch, per = _case_corpus(case, lanes, ticks, 11)
    cap = count * per
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    for i16 in (False, True):
        po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, cap, i16=i16)
        pg, sg, rg = _decode_gpu(arena, offsets, ids, lanes, ch, cap, i16)
        assert np.array_equal(so, sg) and np.all(sg == cap)
        assert np.array_equal(ro, rg), "final_range of the last frame, bit-exact"
        if i16:
            assert int(np.abs(po.astype(np.int32) - pg.astype(np.int32)).max()) <= 1
        else:
            assert float(np.abs(po - pg).max()) <= TOL


@pytest.mark.gpu
def test_gpu_multiframe_loud_packets_are_clipped_as_a_whole():
    """i16 output: opus_pcm_soft_clip runs once over the whole packet (src/opus_decoder.rs:1352-1366). Long random CELT
    payloads overshoot +-1 often, so frame-wise clipping would differ at the frame joints."""
    lanes, ticks, count = 48, 4, 3
    sa, so_ = helpers.fuzz_grid(lanes, ticks * count, 250, lambda t, j, r: helpers.celt_toc(3, True), seed=5)
    arena, offsets = _multiframe_grid(lanes, ticks, count, 3, True, 0, sa, so_)
    ids = np.tile(np.arange(lanes, dtype=np.uint32), ticks)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 2880, i16=True)
    pf, _, _ = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, 2, 2880)
    assert float(np.abs(pf).max()) > 1.0, "corpus must exercise the clipper"
    pg, sg, rg = _decode_gpu(arena, offsets, ids, lanes, 2, 2880, True)
    assert np.array_equal(so, sg) and np.array_equal(ro, rg)
    assert int(np.abs(po.astype(np.int32) - pg.astype(np.int32)).max()) <= 1
    # small passes: frame groups must not straddle a pass boundary
    pg2, sg2, rg2 = _decode_gpu(arena, offsets, ids, lanes, 2, 2880, True, max_frames_per_pass=4 * lanes)
    assert np.array_equal(pg, pg2) and np.array_equal(sg, sg2) and np.array_equal(rg, rg2)


@pytest.mark.gpu
def test_gpu_multiframe_mixed_with_single_and_rejected():
    """One call: single-frame packets, multi-frame packets of different counts, a packet with bad framing, and a packet
    whose middle frame is a 1-byte (concealment) frame. The failing packet reports its code, stops at the failing
    frame like the reference, and the stream carries on with the next packet exactly like the oracle's decoder."""
    from mousiki_b200 import BatchDecoder

    toc = helpers.celt_toc(2, True)  # 10 ms stereo
    fr = lambda s, n=60: bytes(helpers.lcg_bytes(s, n))
    seqs = [
        [helpers.multiframe_packet(toc, [fr(1), fr(2)], 1), np.frombuffer(bytes([toc]) + fr(3), dtype=np.uint8),
         helpers.multiframe_packet(toc, [fr(4, 20), fr(5, 90), fr(6, 33)], 3, True)],
        [np.frombuffer(bytes([toc]) + fr(7), dtype=np.uint8), np.array([toc | 1] + [9] * 5, dtype=np.uint8),
         helpers.multiframe_packet(toc, [fr(8), fr(9)], 2)],
        [helpers.multiframe_packet(toc, [fr(10, 40), b"\x01", fr(11, 40)], 3, True), np.frombuffer(bytes([toc]) + fr(12), dtype=np.uint8),
         helpers.multiframe_packet(toc, [fr(13)] * 4, 3, False, 5)],
    ]
    lib = helpers.oracle_lib()
    want = []
    for seq in seqs:
        h = lib.orc_create(48000, 2, helpers.QUIRKS_MOUSIKI)
        for pkt in seq:
            out = np.zeros(5760 * 2, dtype=np.float32)
            r = lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(out), 5760, 0)
            want.append((r, out, lib.orc_final_range(h)))
        lib.orc_destroy(h)
    pkts = [p for seq in seqs for p in seq]
    ids = np.repeat(np.arange(3, dtype=np.uint32), 3)
    # interleave the streams so that the call is not in grid order
    order = [0, 3, 6, 1, 4, 7, 2, 5, 8]
    arena, offsets = _pack([pkts[i] for i in order])
    with BatchDecoder(3, 48000, 2) as dec:
        pg, sg, rg = dec.decode_float(arena, offsets, ids[order], 5760)
    diverged = set()
    for slot, i in sorted(enumerate(order), key=lambda si: si[1]):
        r, out, rng = want[i]
        if i == 6:   # the middle frame is a 1-byte frame = a concealed frame (src/opus_decoder.rs:491-496): the oracle conceals it
            assert r == 1440 and sg[slot] in (1440, -5), "row F2: -5 until the product conceals"
            if sg[slot] == -5:
                diverged.add(i // 3)     # the oracle's stream went on through the concealed frame: nothing left to compare
        if i // 3 in diverged:
            continue
        assert sg[slot] == r, (i, sg[slot], r)
        if r > 0:
            assert rg[slot] == rng
            assert float(np.abs(pg[slot, :r * 2] - out[:r * 2]).max()) <= TOL
    assert want[4][0] == -4
