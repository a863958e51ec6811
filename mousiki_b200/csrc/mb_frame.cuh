// Phase 0/1 entry for one grid cell: Opus TOC + framing, then the CELT entropy parse.
//
// Replaces the per-packet front end of the reference:
//   opus_packet_get_mode / bandwidth / samples_per_frame / nb_channels   src/packet.rs:30-236
//   opus_packet_parse_impl (code 0 framing)                              src/packet.rs:316-498
//   decode_native_with / decode_frame argument and mode checks           src/opus_decoder.rs:453-600, 1139-1367
// Frame kinds that are not built yet are reported per packet as BD_OPUS_UNIMPLEMENTED and leave the
// stream state untouched: packets that carry a redundancy frame (row F1), lost packets / FEC (row F2). Packets with more than
// one frame (codes 1-3) arrive here already cut into code-0 frames by the host entry points (mb_host_packets.h); on the
// device-resident grid entry point they are reported as BD_OPUS_UNIMPLEMENTED.
#pragma once
#include "mb_celt_parse.cuh"
#include "mb_silk_parse.cuh"

namespace mb {

// cap_per_ch: PCM capacity per channel of the caller's buffer (the reference's `frame_size`).
// prev_pkt: the packet in the lane's previous tick of this pass (0xffffffff: none); prev_mode: the stream's prev_mode word
// when this is the pass's first tick, else null. Only a hint ("the previous packet was hybrid"): a SILK-only packet after
// a hybrid one needs the CELT fade-out frame (src/opus_decoder.rs:868-893), whose record is parsed here; whether it is
// used is decided by the stream pass (k_energy), which knows the real prev_mode.
MB_TBL uint8_t kCeltSilenceFrame[2] = {0xFF, 0xFF};
MB_DEV void parse_cell_body(const uint8_t* arena, const uint32_t* offsets, uint32_t pkt, uint32_t* rec_base, int CC, uint32_t quirks,
                            int disable_inv, int cap_per_ch, const uint8_t* flags, uint32_t prev_pkt, const int32_t* prev_mode,
                            uint32_t* rec_all, const AuxPool& aux) {
  if (pkt == 0xffffffffu) { rec_base[S_STATUS * 32] = 1u << 15; return; }
  // bits 4..7 of the flags: the host front end already rejected the packet (mb_host_packets.h: invalid framing,
  // row too small for all its frames); the cell only exists to report the code with the stream's final_range
  if (flags && (flags[pkt] >> 4)) { rec_base[S_STATUS * 32] = (uint32_t)(uint8_t)(int8_t)(-(int)(flags[pkt] >> 4)); return; }
  const uint32_t off = offsets[pkt];
  const int len = (int)(offsets[pkt + 1] - off);
  if (len == 0 && !flags) { rec_base[S_STATUS * 32] = 1u << 15; return; }
  int code = MB_OK;
  const uint8_t* data = arena + off;
  // The overlap-add form of the inverse MDCT (quirk Q2 cleared) is built for coded CELT-only frames: concealment, in-band FEC and
  // packets with a SILK part are refused per packet there.
  const bool q2_off = !(quirks & Q2_MDCT_ZERO_FILL);
  // (below 48 kHz lost packets are concealed and in-band FEC is decoded like at 48 kHz)
  if (q2_off && ((flags && (flags[pkt] & 3u)) || len == 0)) {
    rec_base[S_STATUS * 32] = (uint32_t)(uint8_t)(int8_t)MB_UNIMPLEMENTED;
    return;
  }
  if ((flags && (flags[pkt] & 2u)) || len == 0) {
    // lost packet (or one concealment frame of it, cut by the host): no payload. The two bytes the host put in its place
    // say how many samples to conceal (mb_host_packets.h); the stream pass decides how (k_energy).
    int n = len >= 2 ? (int)data[0] | ((int)data[1] << 8) : cap_per_ch;
    if (n > cap_per_ch) n = cap_per_ch;
    rec_base[S_RANGE * 32] = (uint32_t)n;
    rec_base[S_STATUS * 32] = 1u << 29;
    return;
  }
  // bit0: decode the packet's in-band FEC (LBRR) data in place of the packet (decode_native, src/opus_decoder.rs:1240-1316).
  // Only SILK / hybrid packets carry any; a CELT-only one is concealed like a lost packet (:1247-1250).
  const int fec = (flags && (flags[pkt] & 1u)) ? 1 : 0;
  if (fec && (data[0] & 0x80u)) {
    rec_base[S_RANGE * 32] = (uint32_t)cap_per_ch;
    rec_base[S_STATUS * 32] = 1u << 29;
    return;
  }
  uint32_t toc = 0;
  if (code == MB_OK) {
    toc = data[0];
    if (toc & 3u) code = MB_UNIMPLEMENTED;              // multi-frame packets (row F1)
    else if (q2_off && (len - 1 <= 1 || !(toc & 0x80u))) code = MB_UNIMPLEMENTED;
    else if (len - 1 > 1275) code = MB_INVALID_PACKET;  // src/packet.rs:465: a frame is at most 1275 bytes
    else if (!(toc & 0x80u)) {
      if (len - 1 <= 1) {                                    // a frame of at most one byte is a concealed frame (src/opus_decoder.rs:491-496)
        const int sc = (int)((toc >> 3) & 3u);
        const int n = (toc & 0x60u) == 0x60u ? ((toc & 0x08u) ? 960 : 480) : (sc == 3 ? 2880 : 480 << sc);
        if (n > cap_per_ch) code = MB_BUFFER_TOO_SMALL;
        else { rec_base[S_RANGE * 32] = (uint32_t)n; rec_base[S_STATUS * 32] = 1u << 29; return; }
      } else if ((toc & 0x60u) == 0x60u) {
        // hybrid: SILK WB payload, optional redundancy header, then CELT bands 17.. from the same range decoder
        // (src/opus_decoder.rs:596-985)
        const int lm = (toc & 0x08u) ? 3 : 2;                // 20 / 10 ms
        const int C = (toc & 4u) ? 2 : 1;
        if ((SHORT_MDCT << lm) > cap_per_ch) code = MB_BUFFER_TOO_SMALL;
        else {
          RangeDec ec;
          ec.open(data + 1, (uint32_t)(len - 1));
          silk_parse_payload(ec, 16, lm == 3 ? 4 : 2, 1, C, quirks, rec_base + (S_SILK_HYB + 1) * 32, fec);
          if (fec) {
            // the CELT part of a packet decoded from its LBRR data is concealed (opus_decoder.rs:944-952): no CELT parse; the
            // cell carries the SILK record, the final range behind it, and the band range of the packet's bandwidth
            rec_base[S_SILK_HYB * 32] = 16u | ((uint32_t)(lm == 3 ? 4 : 2) << 5) | (1u << 8) | ((uint32_t)C << 10);
            rec_base[(S_SILK_HYB + 342) * 32] = ec.rng;
            rec_base[S_RANGE * 32] = ec.rng;
            rec_base[S_HDR0 * 32] = (uint32_t)lm | ((uint32_t)(C - 1) << 2) | (17u << 26);
            rec_base[S_STATUS * 32] = (1u << 14) | (3u << 16) | (1u << 30) | ((uint32_t)((toc & 0x10u) ? 21 : 19) << 8);
            return;
          }
          int redundancy = 0, flen = len - 1;
          const int end = (toc & 0x10u) ? 21 : 19;
          uint32_t kind = 0;
          rec_base[S_AUXREF * 32] = 0u;
          if (ec.tell() + 17 + 20 <= 8 * flen) redundancy = ec.bit_logp(12);
          if (redundancy) {
            // a 5 ms CELT frame behind the hybrid frame, for the switch to / from CELT-only (opus_decoder.rs:746-797): its own
            // range decoder, all bands from 0; it goes to an auxiliary cell (mb_common.cuh)
            const int celt_to_silk = ec.bit_logp(1);
            const int bytes = (int)ec.uint_(256) + 2;
            const int cutoff = flen - bytes;
            if (bytes > flen) code = MB_BAD_ARG;
            else if (cutoff * 8 < ec.tell()) code = MB_UNIMPLEMENTED;   // (the reference then conceals the CELT part)
            else {
              const uint32_t a = rec_all ? aux_alloc(aux) : 0xffffffffu;
              if (a == 0xffffffffu) code = MB_UNIMPLEMENTED;            // pool of auxiliary cells spent in this pass
              else {
                celt_parse_frame<0>(data + 1 + cutoff, bytes, 1, C, end, quirks, disable_inv, rec_all + rec_index(a, 0));
                rec_base[S_AUXREF * 32] = (a << 3) | ((uint32_t)celt_to_silk << 1) | 1u;
                kind = celt_to_silk ? 2u : 3u;
                ec.shrink((uint32_t)bytes);
                flen = cutoff;
              }
            }
          }
          if (code == MB_OK) {
            rec_base[S_SILK_HYB * 32] = 16u | ((uint32_t)(lm == 3 ? 4 : 2) << 5) | (1u << 8) | ((uint32_t)C << 10);
            celt_parse_frame<17>(data + 1, flen, lm, C, end, quirks, disable_inv, rec_base, &ec, (3u << 16) | (kind << 19));
            return;
          }
        }
      } else {
        bool after_hybrid = prev_mode && *prev_mode == 1001;
        if (prev_pkt != 0xffffffffu && flags && (flags[prev_pkt] & 2u)) after_hybrid = true;   // lost: the mode before it is the stream
                                                                                            // pass's to know; make room for the frame
        else if (prev_pkt != 0xffffffffu && offsets[prev_pkt + 1] > offsets[prev_pkt]) {
          const uint32_t ptoc = arena[offsets[prev_pkt]];
          after_hybrid = !(ptoc & 0x80u) && (ptoc & 0x60u) == 0x60u;
        }
        SilkRedundancy red;
        const uint32_t st_silk = silk_parse_frame(data + 1, len - 1, toc, quirks, cap_per_ch, rec_base, &red, fec);
        // status bits 19-20: the CELT frame that shares the cell: 1 fade-out silence, 2 redundancy CELT -> SILK (played
        // before the SILK frame), 3 redundancy SILK -> CELT (after it). Where the cell has no room for the redundancy frame (40 /
        // 60 ms packets; right after a hybrid packet, when the silence frame takes the place) it goes to an auxiliary cell:
        // S_AUXREF = (cell << 3) | silence frame in the own cell << 2 | celt_to_silk << 1 | 1.
        rec_base[S_AUXREF * 32] = 0u;
        if (st_silk & (1u << 14)) {
          const int C = (toc & 4u) ? 2 : 1, end = ((toc >> 5) & 3u) == 0 ? 13 : 17;   // NB 13, MB / WB 17 (opus_decoder.rs:861-866)
          const uint32_t keep = (1u << 16) | (1u << 18) | (st_silk & (1u << 30));   // SILK part, its layout, decoded from LBRR data
          const bool own_room = (st_silk & (1u << 18)) != 0;
          if (red.bytes > 0 && own_room && !after_hybrid)
            celt_parse_frame<0>(data + 1 + red.cutoff, red.bytes, 1, C, end, quirks, disable_inv, rec_base, nullptr,
                                keep | ((red.celt_to_silk ? 2u : 3u) << 19));
          else {
            const bool silence = own_room && after_hybrid;
            if (red.bytes > 0) {
              const uint32_t a = rec_all ? aux_alloc(aux) : 0xffffffffu;
              if (a == 0xffffffffu) {   // pool of auxiliary cells spent in this pass
                rec_base[S_STATUS * 32] = (st_silk & ~(255u | (1u << 14))) | (uint32_t)(uint8_t)(int8_t)MB_UNIMPLEMENTED;
                return;
              }
              celt_parse_frame<0>(data + 1 + red.cutoff, red.bytes, 1, C, end, quirks, disable_inv, rec_all + rec_index(a, 0));
              rec_base[S_AUXREF * 32] = (a << 3) | ((uint32_t)silence << 2) | ((uint32_t)red.celt_to_silk << 1) | 1u;
            }
            if (silence) celt_parse_frame<0>(kCeltSilenceFrame, 2, 0, C, end, quirks, disable_inv, rec_base, nullptr, keep | (1u << 19));
            if (!own_room) {   // 40 / 60 ms packet right after a hybrid one: the fade-out frame goes to an auxiliary cell as well
              rec_base[S_AUXSIL * 32] = 0u;
              const uint32_t a2 = (after_hybrid && rec_all) ? aux_alloc(aux) : 0xffffffffu;
              if (a2 != 0xffffffffu) {
                celt_parse_frame<0>(kCeltSilenceFrame, 2, 0, C, end, quirks, disable_inv, rec_all + rec_index(a2, 0));
                rec_base[S_AUXSIL * 32] = (a2 << 1) | 1u;
              }
            }
          }
        }
        return;
      }
    }
  }
  if (code == MB_OK) {
    const int lm = (int)((toc >> 3) & 3u);
    const int N = SHORT_MDCT << lm;
    if (N > cap_per_ch) code = MB_BUFFER_TOO_SMALL;     // src/opus_decoder.rs:1318-1320
    else if (len - 1 <= 1) { rec_base[S_RANGE * 32] = (uint32_t)N; rec_base[S_STATUS * 32] = 1u << 29; return; }   // concealed frame (:491-496)
    else {
      const int bw = (int)((toc >> 5) & 3u);
      const int end = bw == 0 ? 13 : bw == 1 ? 17 : bw == 2 ? 19 : 21;
      const int C = (toc & 4u) ? 2 : 1;
      celt_parse_frame<0>(data + 1, len - 1, lm, C, end, quirks, disable_inv, rec_base);
      (void)CC;
      return;
    }
  }
  rec_base[S_STATUS * 32] = (uint32_t)(uint8_t)(int8_t)code;
}

// The cell's status word also carries the packet's TOC byte (bits 21-28): the stream pass derives the decoder's bandwidth
// query from it (Decoder::bandwidth, src/codec.rs:941-953).
#define MB_ST_TOC(s) (((s) >> 21) & 255u)
#define MB_ST_LOST_BIT(s) (((s) >> 29) & 1u)
MB_DEV void parse_cell(const uint8_t* arena, const uint32_t* offsets, uint32_t pkt, uint32_t* rec_base, int CC, uint32_t quirks,
                       int disable_inv, int cap_per_ch, const uint8_t* flags, uint32_t prev_pkt = 0xffffffffu,
                       const int32_t* prev_mode = nullptr, uint32_t* rec_all = nullptr, const AuxPool& aux = AuxPool{0, 0, nullptr}) {
  parse_cell_body(arena, offsets, pkt, rec_base, CC, quirks, disable_inv, cap_per_ch, flags, prev_pkt, prev_mode, rec_all, aux);
  if (pkt != 0xffffffffu && !(flags && ((flags[pkt] >> 4) || (flags[pkt] & 2u))) && offsets[pkt + 1] > offsets[pkt] &&
      !MB_ST_LOST_BIT(rec_base[S_STATUS * 32]))
    rec_base[S_STATUS * 32] |= (uint32_t)arena[offsets[pkt]] << 21;
}

}  // namespace mb
