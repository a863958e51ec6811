// Host side of the multi-frame front end: Opus packets with more than one frame
// (TOC codes 1-3, padding) are cut into their frames on the host, and every frame travels to the device as its own
// code-0 "virtual packet" (`TOC & ~3` followed by the frame bytes) that occupies the next tick of the stream's lane.
//
// Replaces, for the batch entry points that take HOST buffers:
//   opus_packet_parse_impl (framing, padding, CBR / VBR sizes)       src/packet.rs:316-498
//   the frame loop and the up-front capacity check of decode_native   src/opus_decoder.rs:1296-1367
// Plain C++ (no CUDA) so that the CPU test-suite can exercise it through tests/emul.
#pragma once
#include <stdint.h>
#include <string.h>

#include <vector>

namespace mbh {

enum { H_OK = 0, H_BUFFER_TOO_SMALL = -2, H_INVALID_PACKET = -4 };

// Samples per frame at 48 kHz for a TOC byte (src/packet.rs:180-236).
inline int toc_frame_samples(uint8_t toc) {
  if (toc & 0x80) return 120 << ((toc >> 3) & 3);          // CELT-only: 2.5 / 5 / 10 / 20 ms
  if ((toc & 0x60) == 0x60) return (toc & 0x08) ? 960 : 480;  // hybrid: 10 / 20 ms
  const int sc = (toc >> 3) & 3;                            // SILK-only: 10 / 20 / 40 / 60 ms
  return sc == 3 ? 2880 : 480 << sc;
}

// One- or two-byte frame length (RFC 6716 section 3.2.1; src/packet.rs:279-314). Returns bytes consumed, 0 on error.
inline int read_frame_len(const uint8_t* d, int avail, int* len) {
  if (avail < 1) return 0;
  if (d[0] < 252) { *len = d[0]; return 1; }
  if (avail < 2) return 0;
  *len = d[0] + 4 * (int)d[1];
  return 2;
}

struct Frames {
  int count;
  int off[48];   // byte offset of each frame inside the packet
  int len[48];
};

// Frame table of one packet (not self-delimited). Returns H_OK or H_INVALID_PACKET.
inline int split_packet(const uint8_t* pkt, int n, Frames* out) {
  if (n < 1) return H_INVALID_PACKET;
  const uint8_t toc = pkt[0];
  int pos = 1, left = n - 1, count = 1, last = left;
  bool cbr = false;
  switch (toc & 3) {
    case 0: break;
    case 1:
      count = 2; cbr = true;
      if (left & 1) return H_INVALID_PACKET;
      last = left / 2;
      out->len[0] = last;
      break;
    case 2: {
      count = 2;
      int l0, used = read_frame_len(pkt + pos, left, &l0);
      if (!used) return H_INVALID_PACKET;
      pos += used; left -= used;
      if (l0 > left) return H_INVALID_PACKET;
      out->len[0] = l0;
      last = left - l0;
    } break;
    default: {
      if (left < 1) return H_INVALID_PACKET;
      const uint8_t fc = pkt[pos++];
      left--;
      count = fc & 0x3f;
      if (count == 0 || count * toc_frame_samples(toc) > 5760) return H_INVALID_PACKET;
      if (fc & 0x40) {                       // padding: 255 means "254 and another length byte follows"
        for (;;) {
          if (left < 1) return H_INVALID_PACKET;
          const int b = pkt[pos++];
          left--;
          const int pad = b == 255 ? 254 : b;
          if (left < pad) return H_INVALID_PACKET;
          left -= pad;
          if (b != 255) break;
        }
      }
      cbr = !(fc & 0x80);
      if (cbr) {
        last = left / count;
        if (last * count != left) return H_INVALID_PACKET;
        for (int i = 0; i + 1 < count; i++) out->len[i] = last;
      } else {
        last = left;
        for (int i = 0; i + 1 < count; i++) {
          int li, used = read_frame_len(pkt + pos, left, &li);
          if (!used) return H_INVALID_PACKET;
          pos += used; left -= used;
          if (li > left) return H_INVALID_PACKET;
          out->len[i] = li;
          last -= used + li;
          if (last < 0) return H_INVALID_PACKET;
        }
      }
    }
  }
  if (last < 0 || last > 1275) return H_INVALID_PACKET;
  out->len[count - 1] = last;
  out->count = count;
  for (int i = 0; i < count; i++) {
    if (pos + out->len[i] > n) return H_INVALID_PACKET;
    out->off[i] = pos;
    pos += out->len[i];
  }
  return H_OK;
}

// Flag bits of a virtual packet (the low two are the caller's: bit0 lost, bit1 FEC).
enum { VF_CONTINUATION = 4,   // frame k > 0 of its packet: skipped if an earlier frame of the packet failed
       VF_REJECT_SHIFT = 4 }; // bits 4..7: -(libopus code) decided on the host (invalid framing, buffer too small)

struct Split {
  std::vector<uint8_t> arena;       // virtual packets back to back
  std::vector<uint32_t> offsets;    // n_virtual + 1
  std::vector<uint32_t> stream_ids; // n_virtual
  std::vector<uint8_t> flags;       // n_virtual
  std::vector<uint32_t> out_map;    // 2 * n_virtual: packet row, element offset inside the row
  std::vector<uint32_t> first;      // n_packets + 1: virtual index of each packet's first frame
  void clear() { arena.clear(); offsets.clear(); stream_ids.clear(); flags.clear(); out_map.clear(); first.clear(); }
};

// What the host has to remember of a stream to cut a LOST packet into concealment frames the way decode_native /
// decode_frame(None) do (src/opus_decoder.rs:491-575, 1186-1222): the frame size and the mode of the last packet the decoder
// took in. Both are functions of the packets' TOC bytes alone (opus_decoder.rs:1322-1328).
struct StreamTrack { int32_t frame_size = 0, mode = 0; };   // mode: 0 nothing yet, 1000 SILK, 1001 hybrid, 1002 CELT

inline bool needs_split(const uint8_t* arena, const uint32_t* offsets, const uint8_t* flags, uint32_t n_packets) {
  for (uint32_t p = 0; p < n_packets; p++) {
    if (offsets[p + 1] > offsets[p] && (arena[offsets[p]] & 3u)) return true;
    if (flags && (flags[p] & 3u)) return true;
    if (flags && offsets[p + 1] == offsets[p]) return true;   // an empty packet is a lost packet (src/opus_decoder.rs:1186)
  }
  return false;
}

// cap_per_ch: PCM capacity of a packet row per channel (the reference's `frame_size`); CC: decoder channels.
// track: one StreamTrack per stream id (null: lost packets conceal one frame of the row's capacity).
inline void split_batch(const uint8_t* arena, const uint32_t* offsets, const uint32_t* stream_ids, const uint8_t* flags,
                        uint32_t n_packets, int cap_per_ch, int CC, Split* S, StreamTrack* track = nullptr) {
  S->clear();
  S->offsets.push_back(0);
  S->arena.reserve(offsets[n_packets] + n_packets);
  auto emit = [&](uint32_t p, uint8_t fl, uint32_t elem_off, const uint8_t* toc_and_frame, uint8_t toc, const uint8_t* frame, int len) {
    if (toc_and_frame) S->arena.insert(S->arena.end(), toc_and_frame, toc_and_frame + len);
    else {
      S->arena.push_back(toc);
      if (len > 0) S->arena.insert(S->arena.end(), frame, frame + len);
    }
    S->offsets.push_back((uint32_t)S->arena.size());
    S->stream_ids.push_back(stream_ids[p]);
    S->flags.push_back(fl);
    S->out_map.push_back(p);
    S->out_map.push_back(elem_off);
  };
  for (uint32_t p = 0; p < n_packets; p++) {
    S->first.push_back((uint32_t)S->stream_ids.size());
    const uint8_t* pkt = arena + offsets[p];
    const int n = (int)(offsets[p + 1] - offsets[p]);
    const uint8_t fl = flags ? (uint8_t)(flags[p] & 3u) : 0;
    StreamTrack* tr = track ? track + stream_ids[p] : nullptr;
    // Lost packet: decode_native conceals `want` samples (a multiple of 2.5 ms, else BadArgument) frame by frame;
    // decode_frame(None) takes at most the last packet's frame size at a time, at most 20 ms, and only whole CELT / SILK
    // frame sizes. Each concealment frame becomes a virtual packet whose two bytes are its sample count.
    auto conceal = [&](int want, int k0) {
      const int fs_state = tr && tr->frame_size > 0 ? tr->frame_size : 120;
      const int md = tr ? tr->mode : 1002;
      int done = 0, k = k0;
      while (done < want) {
        int a = want - done;
        if (a > 5760) a = 5760;
        if (a > fs_state) a = fs_state;
        if (md != 0) {
          if (a > 960) a = 960;
          else if (a < 960) {
            if (a > 480) a = 480;
            else if (md != 1002 && a > 240 && a < 480) a = 240;
          }
        }
        const uint8_t sz[2] = {(uint8_t)(a & 255), (uint8_t)(a >> 8)};
        emit(p, (uint8_t)(2u | (k ? VF_CONTINUATION : 0)), (uint32_t)(done * CC), sz, 0, nullptr, 2);
        done += a; k++;
      }
      return k;
    };
    if ((fl & 2u) || n == 0) {
      if (cap_per_ch <= 0 || cap_per_ch % 120) { emit(p, (uint8_t)(1u << VF_REJECT_SHIFT), 0, nullptr, 0, nullptr, 0); continue; }
      conceal(cap_per_ch, 0);
      continue;
    }
    if (fl & 1u) {
      // In-band FEC (decode_native, src/opus_decoder.rs:1240-1316): the row receives the frame BEFORE this packet, rebuilt from
      // the packet's LBRR data. Whatever the row holds beyond one frame of the packet is concealed first; a CELT-only packet
      // (or a decoder whose last packet was CELT-only) has no such data and the whole row is concealed.
      if (cap_per_ch <= 0 || cap_per_ch % 120) { emit(p, (uint8_t)(1u << VF_REJECT_SHIFT), 0, nullptr, 0, nullptr, 0); continue; }
      Frames F;
      F.count = 0;
      const int code = split_packet(pkt, n, &F);
      if (code != H_OK) { emit(p, (uint8_t)((-code) << VF_REJECT_SHIFT), 0, nullptr, (uint8_t)(pkt[0] & ~3u), nullptr, 0); continue; }
      const int per = toc_frame_samples(pkt[0]);
      const int pmode = (pkt[0] & 0x80) ? 1002 : ((pkt[0] & 0x60) == 0x60 ? 1001 : 1000);
      if (cap_per_ch < per || pmode == 1002 || (tr && tr->mode == 1002)) { conceal(cap_per_ch, 0); continue; }
      const int lead = cap_per_ch - per;
      const int k = conceal(lead, 0);
      if (tr) { tr->frame_size = per; tr->mode = pmode; }
      emit(p, (uint8_t)(1u | (k ? VF_CONTINUATION : 0)), (uint32_t)(lead * CC), nullptr, (uint8_t)(pkt[0] & ~3u), pkt + F.off[0], F.len[0]);
      continue;
    }
    if (n == 0 || fl || !(pkt[0] & 3u)) {   // single frame / FEC request: as it is
      if (n > 0 && !fl && tr && toc_frame_samples(pkt[0]) <= cap_per_ch && n - 1 <= 1275) {
        tr->frame_size = toc_frame_samples(pkt[0]);
        tr->mode = (pkt[0] & 0x80) ? 1002 : ((pkt[0] & 0x60) == 0x60 ? 1001 : 1000);
      }
      emit(p, fl, 0, pkt, 0, nullptr, n);
      continue;
    }
    Frames F;
    F.count = 0;
    int code = split_packet(pkt, n, &F);
    const int per = toc_frame_samples(pkt[0]);
    if (code == H_OK && F.count * per > cap_per_ch) code = H_BUFFER_TOO_SMALL;   // src/opus_decoder.rs:1318-1320
    if (code != H_OK) { emit(p, (uint8_t)((-code) << VF_REJECT_SHIFT), 0, nullptr, (uint8_t)(pkt[0] & ~3u), nullptr, 0); continue; }
    if (tr) { tr->frame_size = per; tr->mode = (pkt[0] & 0x80) ? 1002 : ((pkt[0] & 0x60) == 0x60 ? 1001 : 1000); }
    for (int k = 0; k < F.count; k++)
      emit(p, k ? VF_CONTINUATION : 0, (uint32_t)(k * per * CC), nullptr, (uint8_t)(pkt[0] & ~3u), pkt + F.off[k], F.len[k]);
  }
  S->first.push_back((uint32_t)S->stream_ids.size());
}

// Per-packet results from per-frame results: the sample counts add up, the first failing frame decides the code,
// final_range is the last decoded frame's (src/opus_decoder.rs:1339-1366).
inline void merge_results(const Split& S, uint32_t n_packets, const int32_t* v_samples, const uint32_t* v_range, int32_t* samples_out,
                          uint32_t* final_range_out) {
  for (uint32_t p = 0; p < n_packets; p++) {
    int32_t total = 0;
    uint32_t rng = 0;
    for (uint32_t v = S.first[p]; v < S.first[p + 1]; v++) {
      rng = v_range[v];
      if (v_samples[v] < 0) { total = v_samples[v]; break; }
      total += v_samples[v];
    }
    samples_out[p] = total;
    final_range_out[p] = rng;
  }
}

}  // namespace mbh
