"""Batch loaders: the data formats on the input side of the decode path.

Everything here is host-side byte work that ends in the `(arena, offsets, stream_ids)` triple `BatchDecoder` takes:

* Ogg Opus demux — pages, CRC, lacing, OpusHead / OpusTags (reference: `src/oggreader.rs:60-180`,
  `mousiki-ogg/src/page.rs`, `mousiki-ogg/src/crc.rs`, the page loop of `examples/decode.rs:59-95`);
* `OPUSPKT1` packet files (`examples/opus_packet_tool.rs:14-130`: magic, then `u16` length + packet bytes, a zero
  length ends the list);
* `OPUSBEN1` benchmark corpora (`src/bin/codec_bench.rs:405-470`, `BENCHMARK_COMPARE.md:63-82`: magic, a 20-byte
  header, then `u16` length + packet bytes);
* `batch_from_streams` — several streams' packet lists laid out tick-major, the order `BatchDecoder` decodes in.
"""
import struct
from dataclasses import dataclass

import numpy as np

OPUSPKT1_MAGIC = b"OPUSPKT1"
OPUSBEN1_MAGIC = b"OPUSBEN1"


class LoaderError(ValueError):
    pass


# ------------------------------------------------------------------------------------------------ Ogg
def _crc_table():
    # Ogg's CRC-32: polynomial 0x04c11db7, no reflection, zero initial value and no final xor (mousiki-ogg/src/crc.rs)
    t = np.zeros(256, dtype=np.uint32)
    for i in range(256):
        r = i << 24
        for _ in range(8):
            r = ((r << 1) ^ 0x04C11DB7) & 0xFFFFFFFF if r & 0x80000000 else (r << 1) & 0xFFFFFFFF
        t[i] = r
    return [int(x) for x in t]


_CRC = _crc_table()


def ogg_crc(data: bytes, crc: int = 0) -> int:
    for b in data:
        crc = ((crc << 8) & 0xFFFFFFFF) ^ _CRC[((crc >> 24) & 0xFF) ^ b]
    return crc


@dataclass
class OggPage:
    header_type: int        # bit0 continued packet, bit1 beginning of stream, bit2 end of stream
    granule_position: int
    serial: int
    sequence: int
    lacing: list            # segment table
    body: bytes

    def segments(self):
        """The lacing segments one by one — what the reference's `Page::segment_slices` yields and what
        `examples/decode.rs` feeds its decoder with (a packet shorter than 255 bytes is exactly one segment)."""
        off = 0
        for n in self.lacing:
            yield self.body[off:off + n]
            off += n


@dataclass
class OpusHead:
    version: int
    channels: int
    pre_skip: int
    sample_rate: int
    output_gain: int
    channel_map: int


def ogg_pages(data: bytes, check_crc: bool = True):
    """Pages of a physical Ogg bitstream. A bad capture pattern or checksum raises (the reference reports
    `ChecksumMismatch` / `InvalidPage`, src/oggreader.rs:141-165); a truncated last page ends the iteration like the
    `UnexpectedEof` the example loop breaks on."""
    pos = 0
    n = len(data)
    while pos < n:
        if n - pos < 27:
            return
        if data[pos:pos + 4] != b"OggS":
            raise LoaderError(f"no OggS capture pattern at byte {pos}")
        if data[pos + 4] != 0:
            raise LoaderError("unsupported Ogg stream structure version")
        header_type = data[pos + 5]
        granule, serial, seq, crc, nseg = struct.unpack_from("<qIIIB", data, pos + 6)
        if n - pos < 27 + nseg:
            return
        lacing = list(data[pos + 27:pos + 27 + nseg])
        body_len = sum(lacing)
        end = pos + 27 + nseg + body_len
        if end > n:
            return
        if check_crc:
            hdr = bytearray(data[pos:pos + 27 + nseg])
            hdr[22:26] = b"\0\0\0\0"
            if ogg_crc(data[pos + 27 + nseg:end], ogg_crc(bytes(hdr))) != crc:
                raise LoaderError(f"Ogg page checksum mismatch at byte {pos}")
        yield OggPage(header_type, granule, serial, seq, lacing, data[pos + 27 + nseg:end])
        pos = end


def parse_opus_head(segment: bytes) -> OpusHead:
    """The identification header (src/oggreader.rs:111-139): 19 bytes, `OpusHead`."""
    if len(segment) != 19:
        raise LoaderError("bad OpusHead length")
    if segment[:8] != b"OpusHead":
        raise LoaderError("bad OpusHead signature")
    version, channels, pre_skip, rate, gain, cmap = struct.unpack_from("<BBHIHB", segment, 8)
    return OpusHead(version, channels, pre_skip, rate, gain, cmap)


def read_ogg_opus(data: bytes, by_segment: bool = False):
    """-> (OpusHead, [packet bytes]). Packets are re-assembled from the lacing (a value of 255 continues into the next
    segment, possibly on the next page); `by_segment=True` gives the reference example's view instead: every non-empty
    segment of every page after the headers is a packet (examples/decode.rs:68-93)."""
    pages = ogg_pages(data)
    try:
        first = next(pages)
    except StopIteration:
        raise LoaderError("empty Ogg stream")
    if not first.header_type & 2:
        raise LoaderError("first page is not a beginning-of-stream page")
    segs = list(first.segments())
    if not segs:
        raise LoaderError("bad OpusHead length")
    head = parse_opus_head(segs[0])
    packets = []
    partial = b""
    for page in pages:
        if page.serial != first.serial:
            continue                      # another logical stream multiplexed in: not ours
        segs = list(page.segments())
        if by_segment:
            if segs and segs[0].startswith(b"OpusTags"):
                continue                  # the comment header page (examples/decode.rs:75-79)
            packets.extend(s for s in segs if s)
            continue
        for n, s in zip(page.lacing, segs):
            partial += s
            if n < 255:
                if partial:
                    packets.append(partial)
                partial = b""
    # The comment header is the stream's second logical packet and may span several pages (embedded cover art): it is dropped as
    # a whole packet, after re-assembly — its continuation pages must not turn into the first audio packet.
    if not by_segment and packets and packets[0].startswith(b"OpusTags"):
        packets.pop(0)
    return head, packets


# ------------------------------------------------------------------------------------------------ packet files
def _length_prefixed(data: bytes, pos: int, stop_at_zero: bool):
    out = []
    n = len(data)
    while pos + 2 <= n:
        (ln,) = struct.unpack_from("<H", data, pos)
        pos += 2
        if ln == 0 and stop_at_zero:
            break
        if pos + ln > n:
            raise LoaderError("truncated packet record")
        out.append(data[pos:pos + ln])
        pos += ln
    return out


def read_opuspkt1(data: bytes):
    """`opus_packet_tool` packet file -> [packet bytes] (examples/opus_packet_tool.rs:98-130; 48 kHz stereo by
    construction, :15-17)."""
    if data[:8] != OPUSPKT1_MAGIC:
        raise LoaderError("bad OPUSPKT1 magic")
    return _length_prefixed(data, 8, stop_at_zero=True)


def write_opuspkt1(packets) -> bytes:
    out = bytearray(OPUSPKT1_MAGIC)
    for p in packets:
        if not 0 < len(p) <= 0xFFFF:
            raise LoaderError("packet length out of range")
        out += struct.pack("<H", len(p)) + bytes(p)
    return bytes(out)


@dataclass
class BenchCorpusHeader:
    sample_rate: int
    channels: int
    frame_size: int
    application: int
    bitrate: int
    complexity: int
    bitrate_mode: int


def read_opusben1(data: bytes):
    """`codec_bench packets` corpus -> (BenchCorpusHeader, [packet bytes]) (BENCHMARK_COMPARE.md:63-82)."""
    if data[:8] != OPUSBEN1_MAGIC:
        raise LoaderError("bad OPUSBEN1 magic")
    if len(data) < 28:
        raise LoaderError("truncated OPUSBEN1 header")
    rate, ch, fs, app, br, cx, mode, _reserved = struct.unpack_from("<IHHIiBBH", data, 8)
    return BenchCorpusHeader(rate, ch, fs, app, br, cx, mode), _length_prefixed(data, 28, stop_at_zero=False)


def write_opusben1(header: BenchCorpusHeader, packets) -> bytes:
    out = bytearray(OPUSBEN1_MAGIC)
    out += struct.pack("<IHHIiBBH", header.sample_rate, header.channels, header.frame_size, header.application, header.bitrate,
                       header.complexity, header.bitrate_mode, 0)
    for p in packets:
        out += struct.pack("<H", len(p)) + bytes(p)
    return bytes(out)


# ------------------------------------------------------------------------------------------------ batch layout
def batch_from_streams(streams, stream_ids=None):
    """[[packet, ...] per stream] -> (arena u8, offsets u32[n+1], stream_ids u32[n], index) in tick-major order: first
    every stream's packet 0, then every stream's packet 1, ... Streams may have different lengths. `index[k]` =
    (stream position, packet position) of batch entry k, to route the PCM rows back."""
    ids = list(range(len(streams))) if stream_ids is None else list(stream_ids)
    if len(ids) != len(streams):
        raise LoaderError("stream_ids must name every stream")
    order = []
    depth = max((len(s) for s in streams), default=0)
    for t in range(depth):
        for j, s in enumerate(streams):
            if t < len(s):
                order.append((j, t))
    lens = np.fromiter((len(streams[j][t]) for j, t in order), dtype=np.int64, count=len(order))
    offsets = np.zeros(len(order) + 1, dtype=np.uint32)
    if len(order):
        if int(lens.sum()) >= 1 << 32:
            raise LoaderError("arena larger than 4 GiB")
        offsets[1:] = np.cumsum(lens)
    arena = np.frombuffer(b"".join(bytes(streams[j][t]) for j, t in order), dtype=np.uint8).copy() if len(order) else np.zeros(0, np.uint8)
    sid = np.fromiter((ids[j] for j, _ in order), dtype=np.uint32, count=len(order))
    return arena, offsets, sid, order
