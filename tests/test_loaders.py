"""Batch loaders: Ogg Opus pages -> packets, OPUSPKT1 / OPUSBEN1 packet files, tick-major
batch layout. Known answers are the reference's own: the 80-byte container of src/oggreader.rs:203-213 with the header
fields and segment its tests assert (:215-236), and testdata/tiny.ogg (committed under tests/golden by
tools/gen_ogg_fixture.py), whose single packet SURVEY.md section 8 row A2 quotes."""
import pathlib
import struct

import numpy as np
import pytest

import helpers
from mousiki_b200 import loaders

GOLDEN = pathlib.Path(__file__).parent / "golden"

# src/oggreader.rs:203-213 (test data)
REF_CONTAINER = bytes([
    0x4f, 0x67, 0x67, 0x53, 0x00, 0x02, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00,
    0x8e, 0x9b, 0x20, 0xaa, 0x00, 0x00, 0x00, 0x00, 0x61, 0xee, 0x61, 0x17, 0x01, 0x13,
    0x4f, 0x70, 0x75, 0x73, 0x48, 0x65, 0x61, 0x64, 0x01, 0x02, 0x00, 0x0f, 0x80, 0xbb,
    0x00, 0x00, 0x00, 0x00, 0x00, 0x4f, 0x67, 0x67, 0x53, 0x00, 0x00, 0xda, 0x93, 0xc2,
    0xd9, 0x00, 0x00, 0x00, 0x00, 0x8e, 0x9b, 0x20, 0xaa, 0x02, 0x00, 0x00, 0x00, 0x49,
    0x97, 0x03, 0x37, 0x01, 0x05, 0x98, 0x36, 0xbe, 0x88, 0x9e,
])
TINY_PACKET = bytes.fromhex("4883cade8ae567d51caca254faffbf")


def _page(body_segments, serial=7, seq=0, header_type=0, granule=0):
    lacing = bytes(len(s) for s in body_segments)
    hdr = bytearray(b"OggS\0" + bytes([header_type]) + struct.pack("<qIII", granule, serial, seq, 0) + bytes([len(lacing)]) + lacing)
    body = b"".join(body_segments)
    hdr[22:26] = struct.pack("<I", loaders.ogg_crc(body, loaders.ogg_crc(bytes(hdr))))
    return bytes(hdr) + body


def _head(channels=2, pre_skip=312):
    return b"OpusHead" + struct.pack("<BBHIHB", 1, channels, pre_skip, 48000, 0, 0)


def test_reference_container_header_and_segment():
    head, packets = loaders.read_ogg_opus(REF_CONTAINER)
    assert (head.version, head.channels, head.pre_skip, head.sample_rate, head.output_gain, head.channel_map) == (1, 2, 0x0F00, 48000, 0, 0)
    assert packets == [bytes([0x98, 0x36, 0xBE, 0x88, 0x9E])]
    pages = list(loaders.ogg_pages(REF_CONTAINER))
    assert [p.sequence for p in pages] == [0, 2] and pages[0].header_type == 2 and pages[1].granule_position == 0xD9C293DA


def test_tiny_ogg_yields_the_known_packet():
    data = (GOLDEN / "tiny.ogg").read_bytes()
    assert len(data) == 180
    head, packets = loaders.read_ogg_opus(data)
    assert head.channels == 1 and head.sample_rate == 48000 and head.pre_skip == 0x138
    assert packets == [TINY_PACKET]
    assert loaders.read_ogg_opus(data, by_segment=True)[1] == [TINY_PACKET]
    pages = list(loaders.ogg_pages(data))
    assert len(pages) == 3 and pages[2].header_type == 4 and list(pages[1].segments())[0].startswith(b"OpusTags")


def test_checksum_and_signature_errors():
    data = bytearray((GOLDEN / "tiny.ogg").read_bytes())
    data[-1] ^= 1
    with pytest.raises(loaders.LoaderError, match="checksum"):
        loaders.read_ogg_opus(bytes(data))
    with pytest.raises(loaders.LoaderError, match="capture"):
        list(loaders.ogg_pages(b"NotAnOggFileAtAllNotAnOggFileAtAll"))
    with pytest.raises(loaders.LoaderError, match="OpusHead"):
        loaders.read_ogg_opus(_page([b"OpusHeaX" + bytes(11)], header_type=2))
    with pytest.raises(loaders.LoaderError, match="beginning"):
        loaders.read_ogg_opus(_page([_head()], header_type=0))
    # a truncated last page ends the stream quietly, like the UnexpectedEof the example loop breaks on
    whole = (GOLDEN / "tiny.ogg").read_bytes()
    assert loaders.read_ogg_opus(whole[:-5])[1] == []


def test_lacing_continuation_within_and_across_pages():
    big = bytes(helpers.lcg_bytes(3, 300))       # 255 + 45
    huge = bytes(helpers.lcg_bytes(4, 600))      # 255 + 255 + 90, cut after the first page's two segments
    small = bytes(helpers.lcg_bytes(5, 40))
    exact = bytes(helpers.lcg_bytes(6, 255))     # 255 + a terminating 0
    data = (_page([_head()], header_type=2) + _page([b"OpusTags" + bytes(8)], seq=1)
            + _page([big[:255], big[255:], small, huge[:255], huge[255:510]], seq=2)
            + _page([huge[510:], exact, b""], seq=3, header_type=1)
            + _page([small], seq=4, serial=99)    # a different logical stream: ignored
            + _page([small[::-1]], seq=4, header_type=4))
    head, packets = loaders.read_ogg_opus(data)
    assert packets == [big, small, huge, exact, small[::-1]]
    by_seg = loaders.read_ogg_opus(data, by_segment=True)[1]
    assert by_seg[:3] == [big[:255], big[255:], small] and len(by_seg) == 8


def test_opuspkt1_round_trip_and_errors():
    pk = [bytes(helpers.lcg_bytes(i, 10 + 7 * i)) for i in range(5)]
    blob = loaders.write_opuspkt1(pk)
    assert blob[:8] == b"OPUSPKT1" and loaders.read_opuspkt1(blob) == pk
    assert loaders.read_opuspkt1(blob + b"\0\0" + b"junk after the terminator") == pk   # a zero length ends the list
    with pytest.raises(loaders.LoaderError, match="magic"):
        loaders.read_opuspkt1(b"OPUSPKT2" + blob[8:])
    with pytest.raises(loaders.LoaderError, match="truncated"):
        loaders.read_opuspkt1(blob[:-3])


def test_opusben1_round_trip():
    hdr = loaders.BenchCorpusHeader(48000, 2, 960, 2049, 64000, 10, 1)
    pk = [bytes(helpers.lcg_bytes(20 + i, 160)) for i in range(4)]
    blob = loaders.write_opusben1(hdr, pk)
    assert len(blob) == 28 + 4 * 162 and blob[8:12] == struct.pack("<I", 48000) and blob[26:28] == b"\0\0"
    h2, p2 = loaders.read_opusben1(blob)
    assert h2 == hdr and p2 == pk
    with pytest.raises(loaders.LoaderError):
        loaders.read_opusben1(blob[:20])


def test_batch_from_streams_is_tick_major():
    a = [b"a0", b"a1_", b"a2__"]
    b = [b"b0___"]
    c = [b"c0", b"c1"]
    arena, offsets, sid, order = loaders.batch_from_streams([a, b, c], stream_ids=[5, 6, 9])
    assert order == [(0, 0), (1, 0), (2, 0), (0, 1), (2, 1), (0, 2)]
    assert sid.tolist() == [5, 6, 9, 5, 9, 5]
    got = [bytes(arena[offsets[k]:offsets[k + 1]]) for k in range(len(order))]
    assert got == [b"a0", b"b0___", b"c0", b"a1_", b"c1", b"a2__"]
    arena, offsets, sid, order = loaders.batch_from_streams([])
    assert len(arena) == 0 and offsets.tolist() == [0] and len(sid) == 0


def _three_streams():
    """tiny.ogg's SILK packet followed by fuzz SILK packets; an OPUSPKT1 file of CELT stereo packets; an Ogg stream built
    here holding hybrid packets."""
    _, tiny = loaders.read_ogg_opus((GOLDEN / "tiny.ogg").read_bytes())
    sa, so = helpers.silk_fuzz_grid(1, 3, (8, 40), lambda k, r: helpers.silk_toc(2, 1, False), seed=71, channels=2)
    s0 = tiny + [bytes(sa[so[k]:so[k + 1]]) for k in range(3)]
    ca, co = helpers.fuzz_grid(1, 4, 100, lambda t, j, r: helpers.celt_toc(3, True), seed=72)
    s1 = loaders.read_opuspkt1(loaders.write_opuspkt1([bytes(ca[co[k]:co[k + 1]]) for k in range(4)]))
    ha, ho = helpers.silk_fuzz_grid(1, 4, (60, 120), lambda k, r: 0x7C, seed=73, channels=2)   # hybrid FB 20 ms stereo
    hp = [bytes(ha[ho[k]:ho[k + 1]]) for k in range(4)]
    ogg = _page([_head()], header_type=2) + _page([b"OpusTags" + bytes(8)], seq=1) + _page(hp[:2], seq=2) + _page(hp[2:], seq=3, header_type=4)
    s2 = loaders.read_ogg_opus(ogg)[1]
    assert s2 == hp
    return [s0, s1, s2]


def test_loaded_streams_decode_like_the_oracle_in_emulation():
    streams = _three_streams()
    arena, offsets, sid, order = loaders.batch_from_streams(streams)
    assert len(order) == 12 and sid.tolist() == [0, 1, 2] * 4
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, 3, 4, 2, 960, i16=True)
    pe, se, re, _, _ = helpers.emul_decode_grid(arena, offsets, 3, 4, 2, 960, i16=True)
    assert so[0] == 960 and np.all(so[:9] == 960)
    assert np.array_equal(so, se) and np.array_equal(ro, re) and np.array_equal(po, pe)


@pytest.mark.gpu
def test_gpu_loaded_streams_decode_like_the_oracle():
    from mousiki_b200 import BatchDecoder

    streams = _three_streams()
    arena, offsets, sid, order = loaders.batch_from_streams(streams)
    po, so, ro = helpers.oracle_decode_grid(arena, offsets, 3, 4, 2, 960, i16=True)
    with BatchDecoder(3, 48000, 2) as dec:
        pg, sg, rg = dec.decode(arena, offsets, sid, 960)
    assert np.array_equal(so, sg) and np.array_equal(ro, rg) and np.array_equal(po, pg)


def test_comment_header_spanning_several_pages_is_dropped_as_one_packet():
    """An OpusTags packet with embedded cover art spans pages: its continuation pages must not come back as the
    first audio packet. 255 * 255 + 70 bytes: one full page of 255-byte segments, then the rest on a continuation page."""
    tags = b"OpusTags" + bytes(helpers.lcg_bytes(8, 255 * 255 + 70 - 8))
    segs = [tags[i:i + 255] for i in range(0, len(tags), 255)]
    assert len(segs) == 256 and len(segs[-1]) == 70
    audio = [bytes(helpers.lcg_bytes(9, 90)), bytes(helpers.lcg_bytes(10, 33))]
    data = (_page([_head()], header_type=2) + _page(segs[:255], seq=1) + _page(segs[255:], seq=2, header_type=1)
            + _page(audio, seq=3, header_type=4))
    head, packets = loaders.read_ogg_opus(data)
    assert head.channels == 2 and packets == audio
