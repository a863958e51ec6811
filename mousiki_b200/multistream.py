"""Multistream decode over the batch decoder.

Host-side mirror of mousiki's `OpusMultistreamDecoder` (`src/opus_multistream.rs:151-260, 717-990`): a multistream packet is
N substream packets back to back, the first N-1 in self-delimited framing (`opus_packet_parse_impl(.., self_delimited)`,
`src/packet.rs:316-498`; `opus_decode_native(.., self_delimited)`, `src/opus_decoder.rs:1139-1160`); substream s is decoded by its
own Opus decoder — stereo for the first `coupled_streams`, mono for the rest — and the decoded channels are copied to the output
channels the mapping table names (`get_left_channel / get_right_channel / get_mono_channel`, `:564-600`; 255 = muted).

Here the substreams of many multistream decoders become lanes of two `BatchDecoder`s (one stereo, one mono: "extra batch lanes",
SURVEY section 8 F4): `MultistreamBatchDecoder` holds M independent multistream decoders, one call decodes one packet of each.
The framing work is byte work on the host; all decoding goes through the C ABI (`bd_decode_f32 / bd_decode_i16`)."""
import numpy as np

from .batch_decoder import BatchDecoder

OPUS_BAD_ARG, OPUS_BUFFER_TOO_SMALL, OPUS_INTERNAL_ERROR, OPUS_INVALID_PACKET = -1, -2, -3, -4


class PacketParseError(ValueError):
    def __init__(self, code):
        super().__init__(f"opus packet parse error {code}")
        self.code = code


def _samples_per_frame(toc, fs):
    """opus_packet_get_samples_per_frame (src/packet.rs:206-236)"""
    if toc & 0x80:
        return (fs << ((toc >> 3) & 3)) // 400
    if (toc & 0x60) == 0x60:
        return fs // 50 if toc & 0x08 else fs // 100
    size = (toc >> 3) & 3
    return fs * 60 // 1000 if size == 3 else (fs << size) // 100


def _parse_size(data, pos, end):
    """parse_size (src/packet.rs:285-314): 1 or 2 bytes -> (bytes used, size)"""
    if pos >= end:
        raise PacketParseError(OPUS_INVALID_PACKET)
    b0 = data[pos]
    if b0 < 252:
        return 1, b0
    if pos + 1 >= end:
        raise PacketParseError(OPUS_INVALID_PACKET)
    return 2, 4 * data[pos + 1] + b0


def parse_packet(data, self_delimited):
    """opus_packet_parse_impl (src/packet.rs:316-498) -> (toc, [frame bytes], packet_offset)."""
    n = len(data)
    if n == 0:
        raise PacketParseError(OPUS_INVALID_PACKET)
    toc = data[0]
    idx, remaining = 1, n - 1
    framesize = _samples_per_frame(toc, 48000)
    cbr, pad, last_size = False, 0, remaining
    code = toc & 3
    if code == 0:
        count, sizes = 1, [0]
    elif code == 1:
        count, cbr, sizes = 2, True, [0, 0]
        if not self_delimited:
            if remaining & 1:
                raise PacketParseError(OPUS_INVALID_PACKET)
            last_size = remaining // 2
            sizes[0] = last_size
    elif code == 2:
        count, sizes = 2, [0, 0]
        used, size = _parse_size(data, idx, n)
        if size > max(remaining - used, 0):
            raise PacketParseError(OPUS_INVALID_PACKET)
        idx += used
        remaining -= used
        sizes[0] = size
        last_size = remaining - size
    else:
        if remaining == 0:
            raise PacketParseError(OPUS_INVALID_PACKET)
        ch = data[idx]
        idx += 1
        remaining -= 1
        count = ch & 0x3F
        if count == 0 or count > 48 or framesize * count > 5760:
            raise PacketParseError(OPUS_INVALID_PACKET)
        sizes = [0] * count
        if ch & 0x40:
            while True:
                if remaining == 0:
                    raise PacketParseError(OPUS_INVALID_PACKET)
                p = data[idx]
                idx += 1
                remaining -= 1
                tmp = 254 if p == 255 else p
                pad += tmp
                if remaining < tmp:
                    raise PacketParseError(OPUS_INVALID_PACKET)
                remaining -= tmp
                if p != 255:
                    break
        cbr = (ch & 0x80) == 0
        if not cbr:
            last_size = remaining
            for i in range(count - 1):
                used, size = _parse_size(data, idx, n)
                if size > max(remaining - used, 0):
                    raise PacketParseError(OPUS_INVALID_PACKET)
                idx += used
                remaining -= used
                sizes[i] = size
                last_size -= used + size
                if last_size < 0:
                    raise PacketParseError(OPUS_INVALID_PACKET)
        elif not self_delimited:
            per = remaining // count
            if per * count != remaining:
                raise PacketParseError(OPUS_INVALID_PACKET)
            last_size = per
            for i in range(count - 1):
                sizes[i] = per
    if self_delimited:
        used, size = _parse_size(data, idx, n)
        if size > max(remaining - used, 0):
            raise PacketParseError(OPUS_INVALID_PACKET)
        idx += used
        remaining -= used
        sizes[count - 1] = size
        if cbr:
            if size * count > remaining:
                raise PacketParseError(OPUS_INVALID_PACKET)
            for i in range(count - 1):
                sizes[i] = size
        elif used + size > last_size:
            raise PacketParseError(OPUS_INVALID_PACKET)
    else:
        if last_size < 0 or last_size > 1275:
            raise PacketParseError(OPUS_INVALID_PACKET)
        sizes[count - 1] = last_size
    frames, cur = [], idx
    for s in sizes:
        if cur + s > n:
            raise PacketParseError(OPUS_INVALID_PACKET)
        frames.append(bytes(data[cur:cur + s]))
        cur += s
    if cur + pad > n:
        raise PacketParseError(OPUS_INVALID_PACKET)
    return toc, frames, cur + pad


def _size_bytes(n):
    return bytes([n]) if n < 252 else bytes([252 + (n & 3), (n - (252 + (n & 3))) >> 2])


def regular_packet(toc, frames):
    """The frames of a (self-delimited) substream packet as a packet in regular framing: code 0 for one frame, code 3 VBR else —
    what the batch decoder's front end takes (it decodes frames, not framings: same PCM, same final_range)."""
    if len(frames) == 1:
        return bytes([toc & 0xFC]) + frames[0]
    body = bytes([toc | 3, 0x80 | len(frames)])
    for f in frames[:-1]:
        body += _size_bytes(len(f))
    return body + b"".join(frames)


class MultistreamBatchDecoder:
    """M independent `OpusMultistreamDecoder`s (`opus_multistream_decoder_create`, src/opus_multistream.rs:217-260): `channels` output
    channels, `streams` substreams of which the first `coupled_streams` are stereo, `mapping[c]` = the decoded channel output channel
    c plays (2s / 2s+1: left / right of coupled stream s; 2*coupled + k: mono stream k; 255: silence)."""

    def __init__(self, n_decoders, sample_rate, channels, streams, coupled_streams, mapping, device=0, **kw):
        mapping = [int(m) for m in mapping]
        if not (1 <= channels <= 255 and streams >= 1 and 0 <= coupled_streams <= streams and streams + coupled_streams <= 255
                and len(mapping) == channels and all(m == 255 or m < streams + coupled_streams for m in mapping)):
            raise ValueError("bad multistream layout")          # validate_layout (:262-300): OPUS_BAD_ARG
        self.n, self.fs, self.channels, self.streams, self.coupled, self.mapping = n_decoders, sample_rate, channels, streams, coupled_streams, mapping
        self.mono = streams - coupled_streams
        self._st = BatchDecoder(n_decoders * coupled_streams, sample_rate, 2, device=device, **kw) if coupled_streams else None
        self._mo = BatchDecoder(n_decoders * self.mono, sample_rate, 1, device=device, **kw) if self.mono else None
        self._ranges = np.zeros((n_decoders, streams), dtype=np.uint32)

    def close(self):
        for d in (self._st, self._mo):
            if d is not None:
                d.close()
        self._st = self._mo = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def final_range(self):
        """OPUS_GET_FINAL_RANGE per multistream decoder: xor over its substreams (:670-678)"""
        return np.bitwise_xor.reduce(self._ranges, axis=1)

    def _validate(self, pkt, frame_size):
        """opus_multistream_packet_validate (:717-752) + the length / capacity checks of decode_native (:820-835)
        -> ([regular substream packets], code)"""
        nb = self.streams
        if len(pkt) < 2 * nb - 1:
            return None, OPUS_INVALID_PACKET
        subs, pos, samples = [], 0, None
        try:
            for s in range(nb):
                if pos >= len(pkt):
                    return None, OPUS_INVALID_PACKET
                toc, frames, off = parse_packet(pkt[pos:], s + 1 != nb)
                ns = len(frames) * _samples_per_frame(toc, self.fs)
                if ns * 25 > self.fs * 3 or (samples is not None and ns != samples):
                    return None, OPUS_INVALID_PACKET
                samples = ns
                subs.append(regular_packet(toc, frames))
                pos += off
        except PacketParseError as e:
            return None, e.code
        if samples > frame_size:
            return None, OPUS_BUFFER_TOO_SMALL
        return subs, 0

    def _decode(self, packets, frame_size, i16, fec=False):
        assert len(packets) == self.n
        if frame_size <= 0:
            raise ValueError("frame_size")
        frame_size = min(frame_size, self.fs // 25 * 3)
        dt = np.int16 if i16 else np.float32
        pcm = np.zeros((self.n, frame_size * self.channels), dtype=dt)
        samples = np.zeros(self.n, dtype=np.int32)
        subs = [None] * self.n     # per decoder: list of regular substream packets, or "plc"
        for m, pkt in enumerate(packets):
            if pkt is None or len(pkt) == 0:
                subs[m] = "plc"
                continue
            got, code = self._validate(bytes(pkt), frame_size)
            if got is None:
                samples[m] = code   # nothing was decoded: the decoders of this multistream decoder are untouched
            else:
                subs[m] = got
        live = [m for m in range(self.n) if subs[m] is not None]
        for dec, first, count, width in ((self._st, 0, self.coupled, 2), (self._mo, self.coupled, self.mono, 1)):
            if dec is None or not live:
                continue
            pk, ids, fl = [], [], []
            for m in live:
                for k in range(count):
                    lost = subs[m] == "plc"
                    pk.append(b"" if lost else subs[m][first + k])
                    ids.append(m * count + k)
                    fl.append(2 if lost else (1 if fec else 0))
            offsets = np.zeros(len(pk) + 1, dtype=np.uint32)
            offsets[1:] = np.cumsum([len(p) for p in pk])
            arena = np.frombuffer(b"".join(pk) + b"\0", dtype=np.uint8)[:int(offsets[-1])].copy() if offsets[-1] else np.zeros(1, dtype=np.uint8)
            ids = np.array(ids, dtype=np.uint32)
            flags = np.array(fl, dtype=np.uint8)
            fn = dec.decode if i16 else dec.decode_float
            out, smp, rng = fn(arena, offsets, ids, frame_size, flags=flags)
            i = 0
            for m in live:
                for k in range(count):
                    s = first + k
                    self._ranges[m, s] = rng[i]
                    if smp[i] < 0:
                        if samples[m] >= 0:
                            samples[m] = smp[i]
                    else:
                        if samples[m] >= 0:
                            samples[m] = smp[i]
                        row = out[i].reshape(-1, width)[:smp[i]]
                        dst = pcm[m].reshape(-1, self.channels)
                        for c, mp in enumerate(self.mapping):   # copy_channel_out per mapped channel (:880-930)
                            if s < self.coupled and mp in (2 * s, 2 * s + 1):
                                dst[:smp[i], c] = row[:, mp & 1]
                            elif s >= self.coupled and mp == s + self.coupled:
                                dst[:smp[i], c] = row[:, 0]
                    i += 1
        return pcm, samples

    def decode_float(self, packets, frame_size, fec=False):
        """opus_multistream_decode_float (:983-1000) for one packet (bytes, or None = lost) per decoder
        -> (pcm float32 [M, frame_size * channels], samples int32 [M]: samples per channel, or the Opus error code)"""
        return self._decode(packets, frame_size, False, fec)

    def decode(self, packets, frame_size, fec=False):
        """opus_multistream_decode (:953-970): int16, every substream through the soft clipper"""
        return self._decode(packets, frame_size, True, fec)


def pack_multistream(sub_packets):
    """N regular substream packets -> one multistream packet (the first N-1 re-framed self-delimited; test and tooling aid,
    the inverse of what `parse_packet(.., True)` + `regular_packet` undo)."""
    out = b""
    for i, p in enumerate(sub_packets):
        if i + 1 == len(sub_packets):
            out += bytes(p)
            break
        toc, frames, _ = parse_packet(bytes(p), False)
        code = toc & 3
        if code == 0:
            out += bytes([toc]) + _size_bytes(len(frames[0])) + frames[0]
        elif code == 1:
            out += bytes([toc]) + _size_bytes(len(frames[0])) + frames[0] + frames[1]
        elif code == 2:
            out += bytes([toc]) + _size_bytes(len(frames[0])) + _size_bytes(len(frames[1])) + frames[0] + frames[1]
        else:   # re-framed as code 3 VBR without padding
            body = bytes([toc, 0x80 | len(frames)])
            for f in frames:
                body += _size_bytes(len(f))
            out += body + b"".join(frames)
    return out


def matrix_out_float(matrix, x):
    """mapping_matrix_multiply_channel_out_float (src/mapping_matrix.rs:188-217) over every input row in turn: matrix int16
    [rows, cols] (Q15), x float32 [frames, n_in] -> float32 [frames, rows]; output += (cell / 32768) * sample, one input channel
    after the other, every product and sum rounded to float32 like the reference's loop."""
    rows = matrix.shape[0]
    out = np.zeros((x.shape[0], rows), dtype=np.float32)
    for c in range(x.shape[1]):
        coeff = matrix[:, c].astype(np.float32) / np.float32(32768.0)
        out += coeff[None, :] * x[:, c:c + 1].astype(np.float32)
    return out


def matrix_out_short(matrix, x16):
    """mapping_matrix_multiply_channel_out_short (:251-283): x16 int16 [frames, n_in] -> int16 [frames, rows], each term
    (cell * sample + 16384) >> 15, the int16 sum wrapping."""
    rows = matrix.shape[0]
    out = np.zeros((x16.shape[0], rows), dtype=np.int32)
    for c in range(x16.shape[1]):
        delta = (matrix[:, c].astype(np.int32)[None, :] * x16[:, c:c + 1].astype(np.int32) + 16384) >> 15
        out = ((out + delta + 32768) & 0xFFFF) - 32768
    return out.astype(np.int16)


class ProjectionBatchDecoder:
    """M independent `OpusProjectionDecoder`s (`opus_projection_decoder_create`, src/projection.rs:533-570): a multistream decoder
    with the identity mapping, its decoded channels mixed into `channels` outputs by the demixing matrix (little-endian int16,
    column-major, `channels` rows x `streams + coupled_streams` columns)."""

    def __init__(self, n_decoders, sample_rate, channels, streams, coupled_streams, demixing_matrix, device=0, **kw):
        n_in = streams + coupled_streams
        raw = bytes(demixing_matrix)
        if len(raw) != 2 * n_in * channels:
            raise ValueError("demixing matrix size")                      # OPUS_BAD_ARG (:541-547)
        self.matrix = np.frombuffer(raw, dtype="<i2").reshape(n_in, channels).T.copy()     # [rows = channels, cols = n_in]
        self.channels, self.n_in = channels, n_in
        self._ms = MultistreamBatchDecoder(n_decoders, sample_rate, channels, streams, coupled_streams, list(range(channels)), device=device, **kw)

    def close(self):
        self._ms.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def final_range(self):
        return self._ms.final_range()

    def _decode(self, packets, frame_size, i16, fec):
        pcm, samples = self._ms._decode(packets, frame_size, i16, fec)
        n_used = min(self.n_in, self.channels)     # decoded channel c reaches the mixer when an output channel maps to it
        out = np.zeros_like(pcm)
        for m in range(len(packets)):
            if samples[m] <= 0:
                continue
            x = pcm[m].reshape(-1, self.channels)[:samples[m], :n_used]
            y = matrix_out_short(self.matrix, x) if i16 else matrix_out_float(self.matrix, x)
            out[m].reshape(-1, self.channels)[:samples[m]] = y
        return out, samples

    def decode_float(self, packets, frame_size, fec=False):
        """opus_projection_decode_float (:661-706)"""
        return self._decode(packets, frame_size, False, fec)

    def decode(self, packets, frame_size, fec=False):
        """opus_projection_decode (:614-659): int16, the substreams through the soft clipper, Q15 mixing with rounding"""
        return self._decode(packets, frame_size, True, fec)
