// Phase 2a — per-stream sequential bookkeeping, one lane per stream, ticks in order.
//
// Replaces the state-carrying scalar parts of the reference's CELT/Opus decoder:
//   unquant_coarse_energy / unquant_fine_energy / unquant_energy_finalise (float recurrences)
//                                      src/celt/quant_bands.rs:902-950, 1850-1970
//   energy history rotation            src/celt/celt_decoder.rs:4023-4060
//   rng / final_range / error plumbing src/celt/celt_decoder.rs:4062,4128; src/opus_decoder.rs:978-986,1114-1130
// Everything heavier (PVQ, IMDCT) is independent of the previous frame once the band energies
// and the LCG seed are known, so this tiny serial pass is what lets phase 2b run all ticks at once.
#pragma once
#include "mb_common.cuh"

namespace mb {

MB_TBL float kPredCoef[4] = {29440.f / 32768.f, 26112.f / 32768.f, 21248.f / 32768.f, 16384.f / 32768.f};
MB_TBL float kBetaCoef[4] = {30147.f / 32768.f, 22282.f / 32768.f, 12124.f / 32768.f, 6554.f / 32768.f};
MB_TBL float kEMeans[25] = {6.4375f, 6.25f, 5.75f, 5.3125f, 5.0625f, 4.8125f, 4.5f, 4.375f, 4.875f, 4.6875f, 4.5625f, 4.4375f, 4.875f,
                            4.625f, 4.3125f, 4.5f, 4.375f, 4.625f, 4.75f, 4.4375f, 3.75f, 3.75f, 3.75f, 3.75f, 3.75f};

// Persistent per-stream state (HBM, one record per stream). Reference: OpusDecoder
// (src/opus_decoder.rs:200-241) + OpusCustomDecoder (src/celt/types.rs:793-859), minus the
// signal history which lives in its own array (see mb_celt_tail.cuh).
struct StreamState {
  float old_ebands[2 * NB_EBANDS];
  float old_log_e[2 * NB_EBANDS];
  float old_log_e2[2 * NB_EBANDS];
  float background_log_e[2 * NB_EBANDS];
  float preemph_mem[2];
  // post-filter parameters, one private copy per channel lane of phase 2c (identical values)
  struct Pf { float gain, gain_old; int32_t period, period_old, tapset, tapset_old; } pf[2];
  uint32_t rng;            // CELT decoder rng (seed of the next frame's noise generator)
  uint32_t range_final;    // Opus-level final range of the last successfully decoded packet
  int32_t loss_duration;
  int32_t prev_mode;       // 0 none, else MODE_*
  int32_t last_packet_duration;
  float softclip_mem[2];
  int32_t celt_error;
  int32_t prev_redundancy;   // the last packet ended in a SILK -> CELT redundancy frame (opus_decoder.rs:1117)
  int32_t fold_serial;       // the stream has decoded a SILK-side frame: its CELT-only frames keep the fold buffer exact too
  int32_t bandwidth;         // 0 before the first packet, else 1101 (NB) .. 1105 (FB) of the last packet (opus_decoder.rs:1322)
  // packet-loss concealment (src/celt/types.rs:807-816, celt_decoder.rs:1429-1855)
  int32_t skip_plc;          // no usable history for the pitch-based concealment (set by a reset and by the noise concealment)
  int32_t prefilter_and_fold;// bookkeeping only: under quirk Q2 the next inverse MDCT zero-fills what prefilter_and_fold writes
  int32_t last_pitch_index;  // pitch of the first concealed frame of a loss burst
  int32_t celt_end;          // the CELT decoder's end band (set from the bandwidth of the last coded packet; a lost frame keeps it)
  float plc_lpc[2][24];      // concealment synthesis filter per channel
};

MB_DEV void stream_state_reset(StreamState& s) {
  for (int i = 0; i < 2 * NB_EBANDS; i++) { s.old_ebands[i] = 0.f; s.old_log_e[i] = -28.f; s.old_log_e2[i] = -28.f; s.background_log_e[i] = 0.f; }
  s.preemph_mem[0] = s.preemph_mem[1] = 0.f;
  for (int c = 0; c < 2; c++) { s.pf[c].gain = s.pf[c].gain_old = 0.f; s.pf[c].period = s.pf[c].period_old = 0; s.pf[c].tapset = s.pf[c].tapset_old = 0; }
  s.rng = 0; s.range_final = 0; s.loss_duration = 0; s.prev_mode = 0; s.last_packet_duration = 0;
  s.softclip_mem[0] = s.softclip_mem[1] = 0.f;
  s.celt_error = 0;
  s.prev_redundancy = 0; s.fold_serial = 0;
  s.bandwidth = 0;
  s.skip_plc = 1; s.prefilter_and_fold = 0; s.last_pitch_index = 0; s.celt_end = NB_EBANDS;
  for (int c = 0; c < 2; c++) for (int i = 0; i < 24; i++) s.plc_lpc[c][i] = 0.f;
}

// per-frame side data produced here and consumed by phases 2b / 2c
struct FrameMeta {
  uint32_t seed_in;      // LCG seed for quant_all_bands (= rng left by the previous frame)
  uint32_t sample_off;   // first sample of this frame in the lane's time series (excluding HIST)
  uint32_t lane;         // lane (stream slot) of this call
  uint32_t flags;        // bit0: frame runs through phases 2b/2c; bit1: SILK synthesis (k_silk); bit2: CELT state reset before
                         // this frame; bit3: the CELT output is added to the SILK output already in the PCM row (hybrid, fade-out);
                         // bit4: redundancy frame: the CELT output goes to the cell's spectrum row (dead by then) and k_fade
                         // cross-fades it into the SILK row; bit5: SILK decoder reset before the packet; bit6 / bit7: the
                         // fade at the start (CELT -> SILK) / at the end (SILK -> CELT) of the row applies
};
// FrameMeta.flags, upper bits: 28 the frame's inverse MDCT also keeps its 120-sample overlap tail behind the frame in the
// time series (the frame that follows conceals by pitch repetition and reads it); 30 concealed CELT frame (lost packet);
// 31 concealed by pitch repetition (k_comb_plc synthesises it; seed_in carries loss_duration) — without bit 31 a concealed
// frame is band-shaped noise and runs through k_plan / k_bands / k_imdct like a coded frame (seed_in = noise seed).
// 29: k_fade finishes the packet's row (cross-fade with a side row, then the decode gain): nobody else applies the gain to it.
enum : uint32_t { MB_FL_KEEP_TAIL = 1u << 28, MB_FL_FADE_ROW = 1u << 29, MB_FL_PLC = 1u << 30, MB_FL_PLC_PITCH = 1u << 31 };
// bit 10, only on a CELT-only frame of the serial band pass (bit 9 set: bits 10-27 carry nothing there): the packet opens with a
// cross-fade from 5 ms concealed in the old mode, SILK-only or hybrid (opus_decoder.rs:575-593, 1092-1112). The stream's k_silk lane
// conceals into the side row of the cell side_of[] names (bit 1 makes the lane stop at the cell), k_fade blends the first 2.5 ms.
enum : uint32_t { MB_FL_TRANS_IN = 1u << 10 };
// S_STATUS bit 29: lost frame (no payload); S_RANGE then holds the number of samples to conceal
#define MB_ST_LOST(s) (((s) >> 29) & 1u)
// S_STATUS bit 30: the cell's SILK record comes from the packet's LBRR data (in-band FEC)
#define MB_ST_FEC(s) (((s) >> 30) & 1u)
enum { EN_STRIDE = 4 * 2 * NB_EBANDS, EN_GAINS = 3 * 2 * NB_EBANDS,   // per-frame energy scratch: E, prev1 (old_log_e), prev2 (old_log_e2), band gains
      
       MB_FL_FOLD = 8u | 16u | 512u,
       MB_FL_F32ROW = 256u };  // the packet's PCM goes through the float row (k_deemph / k_fade / k_clip_i16), set on all its frames   // FrameMeta.flags: the frame's band pass runs per stream, in tick order, on the persistent fold buffer
// flags bit 9 (512): CELT-only frame behind a SILK-side frame of the same pass; bits 10-18 / 19-27 (units of 4 words): fold-buffer words
// [lo, hi) a CELT-only frame of the parallel band pass writes back (the part no later frame of the pass overwrites)  // per-frame energy scratch: E, prev1 (old_log_e), prev2 (old_log_e2)

struct Team { int tid, n; };

// The CELT frames of a lane in decode order, written by the stream pass (k_energy) and walked by every kernel that is serial
// in time (k_bands_hyb, k_comb, k_comb_plc, k_deemph): cell of the k-th frame = order[k * n_lanes + lane], k < order_n[lane].
// Almost always these are the lane's cells that hold a CELT frame, in tick order; auxiliary cells (mb_common.cuh) slot in
// where their frame is decoded.
struct EnergyParams {
  uint32_t n_lanes, n_ticks;
  const uint32_t* pkt_map;
  uint32_t* rec;
  StreamState* st_all;
  const uint32_t* lane_stream;
  float* en;
  FrameMeta* meta;
  int32_t* samples_out;
  uint32_t* final_range_out;
  uint32_t quirks;
  const uint8_t* flags;
  int serial_all, force_float_row, CC;
  uint32_t* order; uint32_t* order_n;   // see above; order has room for 2 * n_ticks entries per lane
  uint32_t* side_of;                    // per main cell: the cell whose side row holds the audio k_fade blends in (0xffffffff: none)
  uint32_t* aux_main;                   // per auxiliary cell whose frame belongs in a packet's row: that packet's main cell
  AuxPool aux;
};

// Shared-memory working set of one stream: the four energy histories, this frame's coarse symbols.
struct EnergySmem { float E[2 * NB_EBANDS], le[2 * NB_EBANDS], le2[2 * NB_EBANDS], bg[2 * NB_EBANDS], q[2 * NB_EBANDS]; };

// One team (a warp; one thread in the host emulation) = one stream, its ticks in order. Frame (grid cell) of tick t is
// f = t * n_lanes + lane; its packet index (for the per-packet outputs) is pkt_map ? pkt_map[f] : f. Cells without a packet
// carry S_STATUS bit 15. The front-end bookkeeping is scalar and runs redundantly on every lane (the words it reads are
// the same for the whole team); the 42 band energies are spread over the lanes: the coarse-energy predictor's running
// term prev[c] of band i is the left-to-right sum of the lower bands' symbols, which every lane rebuilds for its own
// band in the reference's order, so the result is bit-identical to the serial loop.
MB_DEV void celt_energy_stream(Team tm, EnergySmem* sm, uint32_t lane, const EnergyParams& Q) {
  const uint32_t n_lanes = Q.n_lanes, n_ticks = Q.n_ticks, quirks = Q.quirks;
  const uint32_t* pkt_map = Q.pkt_map;
  uint32_t* rec = Q.rec;
  float* en = Q.en;
  FrameMeta* meta = Q.meta;
  int32_t* samples_out = Q.samples_out;
  uint32_t* final_range_out = Q.final_range_out;
  const uint8_t* flags = Q.flags;
  const int serial_all = Q.serial_all, force_float_row = Q.force_float_row, CC = Q.CC;
  StreamState& st = Q.st_all[Q.lane_stream ? Q.lane_stream[lane] : lane];
  const bool lead = tm.tid == 0;
  uint32_t n_order = 0;
  auto emit_frame = [&](uint32_t cell) {   // the next CELT frame of the lane in decode order
    if (lead) Q.order[(size_t)n_order * n_lanes + lane] = cell;
    n_order++;
  };
  for (int i = tm.tid; i < 2 * NB_EBANDS; i += tm.n) {
    sm->E[i] = st.old_ebands[i]; sm->le[i] = st.old_log_e[i]; sm->le2[i] = st.old_log_e2[i]; sm->bg[i] = st.background_log_e[i];
  }
  uint32_t st_rng = st.rng, st_range_final = st.range_final;
  int st_loss_duration = st.loss_duration, st_prev_mode = st.prev_mode, st_last_duration = st.last_packet_duration;
  int st_celt_error = st.celt_error, st_prev_redundancy = st.prev_redundancy, st_fold_serial = st.fold_serial, st_bandwidth = st.bandwidth;
  int st_skip_plc = st.skip_plc, st_prefold = st.prefilter_and_fold, st_celt_end = st.celt_end;
  MB_SYNC();
  uint32_t sample_off = 0;
  // Band pass of CELT-only frames. The reference's fold buffer outlives the frame (mb_celt_bands.cuh), so only a pass that
  // walks a stream's frames in order reproduces it exactly: that is what runs whenever the call has enough streams to
  // fill the machine that way (serial_all), and for every stream that has decoded a SILK-side frame (hybrid, fade-out,
  // redundancy: the frames that read the buffer). Otherwise CELT-only frames run all at once and write back the part of
  // their footprint that survives the pass: exact for CELT-only streams, whose frames never read the buffer.
  bool silk_side_seen = serial_all || st_fold_serial;
  bool pkt_failed = false;   // an earlier frame of the same multi-frame packet failed: the reference returns there
                             // (src/opus_decoder.rs:1339-1350), so the remaining frames must not touch the stream
  // Output route of the packet, decided once for all its frames (MB_FL_F32ROW): SILK-only packets may leave k_silk as
  // int16 directly (the i16 API), but only if NO frame of the packet shares its row with a CELT frame and no decode gain
  // applies; otherwise every frame goes through the float row and the packet through the clipper.
  uint32_t pkt_route = 0;
  for (uint32_t t = 0; t < n_ticks; t++) {
    const uint32_t f = t * n_lanes + lane;
    uint32_t* r = rec + rec_index(f, 0);
    uint32_t status = r[S_STATUS * 32];
    FrameMeta m;
    m.seed_in = st_rng; m.sample_off = sample_off; m.lane = lane; m.flags = 0;
    if (lead) Q.side_of[f] = 0xffffffffu;
    if (MB_ST_EMPTY(status)) { if (lead) meta[f] = m; continue; }
    const uint32_t p = pkt_map ? pkt_map[f] : f;
    const bool continuation = flags && (flags[p] & 4u);
    if (!continuation) {
      pkt_failed = false;
      // (a SILK / hybrid packet right after a CELT-only one fades in from a concealed CELT frame: float row)
      bool celt_part = force_float_row != 0 || (MB_ST_DECODED(status) && ((status >> 16) & 1u) && st_prev_mode == 1002);
      for (uint32_t t2 = t; t2 < n_ticks && !celt_part; t2++) {
        const uint32_t f2 = t2 * n_lanes + lane;
        if (t2 > t) {
          const uint32_t p2 = pkt_map ? pkt_map[f2] : f2;
          if (!flags || p2 == 0xffffffffu || !(flags[p2] & 4u)) break;
        }
        const uint32_t s2 = rec[rec_index(f2, S_STATUS)];
        celt_part = MB_ST_DECODED(s2) && (!((s2 >> 16) & 1u) || ((s2 >> 17) & 1u) || ((s2 >> 19) & 3u) || (rec[rec_index(f2, S_AUXREF)] & 1u) ||
                                          (!((s2 >> 18) & 1u) && (rec[rec_index(f2, S_AUXSIL)] & 1u)));
        // (concealment frames come first in their packet: they all see the mode the stream is in now)
        if (MB_ST_LOST(s2) && (st_prev_redundancy || st_prev_mode != 1000)) celt_part = true;
      }
      pkt_route = celt_part ? MB_FL_F32ROW : 0u;
    }
    // ---- CELT concealment of one frame (celt_decode_lost, src/celt/celt_decoder.rs:1429-1855), bands [start, st_celt_end):
    // band-shaped noise through k_plan / k_bands / k_imdct, or (full-band streams, short losses) pitch repetition in
    // k_comb_plc. `extra`: flag bits of the SILK side of the frame (hybrid). Leaves m, the energies and the CELT state updated.
    auto celt_conceal = [&](uint32_t fc, uint32_t* rc, uint32_t statusc, FrameMeta& mc, int audiosize, int lmq, int start, uint32_t extra) {
      const int end = st_celt_end, Mq = 1 << lmq;
      mc.sample_off = sample_off; mc.lane = lane;
      const int noise_based = st_loss_duration >= 40 || start != 0 || st_skip_plc;
      // a synthetic frame header for the kernels downstream: LM, the DECODER's channel count, band range; no leaves, no nodes
      if (lead) {
        rc[S_HDR0 * 32] = (uint32_t)lmq | ((uint32_t)(CC - 1) << 2) | (2u << 6) | ((uint32_t)start << 26);
        rc[S_HDR1 * 32] = 0u;
        rc[S_STATUS * 32] = (statusc & ~(31u << 8)) | ((uint32_t)end << 8);
        rc[S_PLAN * 32] = 0u;
      }
      float* ef = en + (size_t)fc * EN_STRIDE;
      if (noise_based) {
        const float decay = st_loss_duration == 0 ? 1.5f : 0.5f;
        MB_SYNC();
        for (int idx = tm.tid; idx < 2 * NB_EBANDS; idx += tm.n) {
          const int c = idx >= NB_EBANDS, i = idx - c * NB_EBANDS;
          float e = sm->E[idx];
          if (c < CC && i >= start && i < end) { e = fmax32(sm->bg[idx], e - decay); sm->E[idx] = e; }
          ef[idx] = e;
          ef[EN_GAINS + idx] = celt_exp2(fmin32(e + kEMeans[i], 32.0f));
        }
        MB_SYNC();
        // the noise generator runs channel by channel over bands start .. eff_end (celt_decoder.rs:1511-1531)
        int total = 0;
        for (int i = start; i < imin(end, NB_EBANDS); i++) total += (eband(i + 1) - eband(i)) << lmq;
        mc.seed_in = st_rng;
        uint32_t sd = st_rng;
        for (int k = 0; k < CC * total; k++) sd = sd * 1664525u + 1013904223u;
        st_rng = sd;
        st_prefold = 0; st_skip_plc = 1;
        mc.flags = 1u | MB_FL_PLC | MB_FL_F32ROW | (silk_side_seen ? 512u : 0u) | extra;
      } else {
        mc.seed_in = (uint32_t)st_loss_duration;
        st_prefold = 1;
        mc.flags = 1u | MB_FL_PLC | MB_FL_PLC_PITCH | MB_FL_F32ROW | extra;
      }
      st_loss_duration = imin(10000, st_loss_duration + Mq);
      sample_off += (uint32_t)audiosize;
      emit_frame(fc);
    };
    if (MB_ST_LOST(status) && !pkt_failed) {
      // ---- lost packet / frame of at most one byte: decode_frame(None) (src/opus_decoder.rs:491-575). The stream conceals in
      // the mode of its last packet (CELT-only after a SILK -> CELT redundancy frame): SILK in k_silk (mb_silk_plc.cuh), CELT
      // as above, hybrid both (the CELT part from band 17, always noise).
      const int audiosize = (int)r[S_RANGE * 32];            // in 48 kHz samples; the row takes audiosize / ds_out of them
      const int ds_out = quirk_ds(Q.quirks);
      const int md = st_prev_redundancy ? 1002 : st_prev_mode;
      if (md == 0) {   // nothing decoded yet: silence (opus_decoder.rs:527-540)
        m.flags = MB_FL_PLC | MB_FL_F32ROW;   // concealed, but no CELT frame: k_fade zero-fills seed_in samples of the row
        m.seed_in = (uint32_t)(audiosize / ds_out);
        st_prev_mode = 0; st_prev_redundancy = 0; st_range_final = 0; st_last_duration = audiosize / ds_out;
        if (lead) { meta[f] = m; samples_out[p] = audiosize / ds_out; final_range_out[p] = 0; }
        continue;
      }

      int lmq = -1;
      for (int q = 0; q <= MAX_LM; q++) if ((SHORT_MDCT << q) == audiosize) lmq = q;
      if (lmq < 0) {   // (the host front end only cuts whole frame sizes)
        if (lead) { meta[f] = m; samples_out[p] = MB_UNIMPLEMENTED; final_range_out[p] = st_range_final; }
        pkt_failed = true;
        continue;
      }
      if (md == 1000) m.flags = 2u | MB_FL_PLC | pkt_route;
      else celt_conceal(f, r, status, m, audiosize, lmq, md == 1001 ? 17 : 0, md == 1001 ? (2u | 8u) : 0u);
      st_prev_mode = md; st_prev_redundancy = 0; st_range_final = 0; st_last_duration = audiosize / ds_out;
      if (lead) { meta[f] = m; samples_out[p] = audiosize / ds_out; final_range_out[p] = 0; }
      continue;
    }
    if (pkt_failed || !MB_ST_DECODED(status)) {  // rejected before the decoder ran (unsupported mode, PLC, bad TOC)
      if (lead) {
        meta[f] = m;
        samples_out[p] = pkt_failed ? MB_UNIMPLEMENTED : MB_ST_CODE(status);
        final_range_out[p] = st_range_final;
      }
      pkt_failed = true;
      continue;
    }
    {   // opus_packet_get_bandwidth (src/packet.rs:177-204) of the packet the decoder has just taken in
      const uint32_t toc = (status >> 21) & 255u, two = (toc >> 5) & 3u;
      st_bandwidth = (toc & 0x80u) ? (two == 0 ? 1101 : 1102 + (int)two) : ((toc & 0x60u) == 0x60u ? ((toc & 0x10u) ? 1105 : 1104) : 1101 + (int)two);
    }
    uint32_t h0 = r[S_HDR0 * 32];
    const bool silk_bit = ((status >> 16) & 1u) != 0, hybrid = ((status >> 17) & 1u) != 0;
    // bit 18: SILK-only cell whose record sits in the hybrid layout; bits 19-20: the CELT frame sharing the cell (parse_cell):
    // 1 the 2.5 ms silence frame that lets the CELT overlap fade out after a hybrid packet (opus_decoder.rs:868-893, parsed
    // on a hint, used only if the stream really comes from a hybrid packet), 2 / 3 a 5 ms redundancy frame for a
    // CELT -> SILK / SILK -> CELT switch (opus_decoder.rs:746-797, 1012-1080).
    const bool silk_layout = ((status >> 18) & 1u) != 0;
    const int aux = (int)((status >> 19) & 3u);
    // (40 / 60 ms packets keep the fade-out frame in an auxiliary cell)
    const uint32_t asil = (silk_bit && !hybrid && !silk_layout) ? r[S_AUXSIL * 32] : 0u;
    const bool sil_aux = (asil & 1u) != 0;
    const bool fadeout = (aux == 1 || sil_aux) && st_prev_mode == 1001;
    // SILK-only cell whose redundancy frame sits in an auxiliary cell (40 / 60 ms packets, packets right after a hybrid one)
    const uint32_t aref = (silk_bit && !hybrid) ? r[S_AUXREF * 32] : 0u;
    const bool red_aux = (aref & 1u) != 0;
    const bool red = aux >= 2 || red_aux;
    const bool silk_only = silk_bit && !hybrid && !fadeout && !red;
    // Mode switches that need concealment (decode_frame(None), opus_decoder.rs:575-593, 798-809) are row F2: reported per
    // packet without touching the stream. A redundancy frame makes the switch clean: CELT -> SILK when the SILK packet
    // carries one, SILK -> CELT when the previous SILK packet ended in one.
    bool reject;
    const bool trans_to_celt = !silk_bit && (st_prev_mode == 1000 || st_prev_mode == 1001) && !st_prev_redundancy;
    if (!silk_bit) reject = false;
    else if (red) reject = !hybrid && st_prev_mode == 1001 && !(red_aux && fadeout);   // after hybrid the cell must hold the fade-out frame
    else reject = silk_only && st_prev_mode == 1001;
    // below 48 kHz a stream stays in one mode: plain SILK-only packets, or CELT-only ones (row F4; the fades, redundancy frames and
    // concealed transition frames of a mode switch are not built at those rates)
#ifdef MB_F4_ONE_MODE   // (kept for bisecting: below 48 kHz a stream stays in one mode)
    if (quirk_ds(Q.quirks) > 1) {
      const bool plain_hybrid = hybrid && !red && !fadeout && (st_prev_mode == 0 || st_prev_mode == 1001);
      const bool plain_silk = silk_only && (st_prev_mode == 0 || st_prev_mode == 1000);
      if ((silk_bit && !plain_hybrid && !plain_silk) || trans_to_celt) reject = true;
    }
#endif
    if (reject) {
      if (lead) {
        meta[f] = m;
        samples_out[p] = MB_UNIMPLEMENTED;
        final_range_out[p] = st_range_final;
      }
      pkt_failed = true;
      continue;
    }
    // SILK side of the cell: header and final range sit in the hybrid layout when the cell uses it
    const uint32_t silk_h0 = silk_layout ? r[S_SILK_HYB * 32] : h0;
    const uint32_t silk_rng = silk_layout ? r[(S_SILK_HYB + 342) * 32] : r[S_RANGE * 32];
    const int silk_n = (int)((silk_h0 >> 8) & 3u) * (int)((silk_h0 >> 5) & 7u) * 240;   // frames x subframes: 2, 4, 8, 12 -> 10 .. 60 ms
    const uint32_t silk_reset = (silk_bit && st_prev_mode == 1002) ? 32u : 0u;          // opus_decoder.rs:694-696
    // CELT-only -> SILK / hybrid without a redundancy frame (opus_decoder.rs:575-593, 798-809, 1083-1105): the reference conceals
    // 5 ms in the old mode (celt_decode_lost on the CELT state as it stands) and fades from it into the packet's audio. The
    // concealed frame goes to an auxiliary cell, its audio to that cell's side row; k_fade blends it in like a CELT -> SILK
    // redundancy frame.
    uint32_t trans_flags = 0;
    if (silk_bit && !red && st_prev_mode == 1002) {
      uint32_t ac = 0xffffffffu;
      if (lead) ac = aux_alloc(Q.aux);
#if MB_ON_DEVICE
      ac = __shfl_sync(0xffffffffu, ac, 0);
#endif
      if (ac == 0xffffffffu) {   // pool of auxiliary cells spent in this pass
        if (lead) { meta[f] = m; samples_out[p] = MB_UNIMPLEMENTED; final_range_out[p] = st_range_final; }
        pkt_failed = true;
        continue;
      }
      FrameMeta ma;
      ma.seed_in = st_rng; ma.sample_off = sample_off; ma.lane = lane; ma.flags = 0;
      celt_conceal(ac, rec + rec_index(ac, 0), 0u, ma, 2 * SHORT_MDCT, 1, 0, 16u);
      if (lead) { meta[ac] = ma; Q.side_of[f] = ac; }
      trans_flags = 64u | MB_FL_FADE_ROW | MB_FL_F32ROW;
      silk_side_seen = true; st_fold_serial = 1;
      m.seed_in = st_rng; m.sample_off = sample_off;
    }
    if (silk_only) {  // SILK-only packet: synthesis runs in k_silk (mb_silk_out.cuh); here only the front-end bookkeeping
      const int N = silk_n;
      m.flags = 2u | silk_reset | pkt_route | trans_flags;
      st_prev_mode = 1000;
      st_prev_redundancy = 0;
      st_range_final = silk_rng;
      st_last_duration = N / quirk_ds(Q.quirks);
      if (lead) {
        meta[f] = m;
        samples_out[p] = N / quirk_ds(Q.quirks);
        final_range_out[p] = st_range_final;
      }
      continue;
    }
    // A mode change resets the CELT decoder before the frame (opus_decoder.rs:895-898): SILK-only -> hybrid, CELT-only -> hybrid
    // (behind the CELT -> SILK redundancy frame, which still runs on the old state), and before a SILK -> CELT redundancy frame
    // (:1040-1046); not after a packet that ended in one. The energy side is reset here; the tail kernels reset their own state
    // (post-filter, history, de-emphasis memory) on flag bit 2 of the frame.
    auto celt_reset_now = [&]() {
      MB_SYNC();
      for (int i = tm.tid; i < 2 * NB_EBANDS; i += tm.n) { sm->E[i] = 0.f; sm->le[i] = -28.f; sm->le2[i] = -28.f; sm->bg[i] = 0.f; }
      st_rng = 0; st_loss_duration = 0;
      st_skip_plc = 1; st_prefold = 0;   // (last_pitch_index and the concealment filter are only read after a frame has set them)
    };
    // ---- one coded CELT frame: the record of cell fc (the packet's own cell or an auxiliary one), energies, state rotation,
    // its place in the lane's time series. Returns the frame's length in samples.
    auto celt_coded = [&](uint32_t fc, const uint32_t* rc, uint32_t statusc, uint32_t flags_c) -> int {
      FrameMeta mc;
      mc.seed_in = st_rng; mc.sample_off = sample_off; mc.lane = lane; mc.flags = flags_c | MB_FL_F32ROW;
      const uint32_t h0c = rc[S_HDR0 * 32];
      const int lm = (int)MB_H0_LM(h0c), C = (int)MB_H0_C(h0c), start = (int)MB_H0_START(h0c), end = (int)MB_ST_END(statusc);
      const int intra = (int)MB_H0_INTRA(h0c), silence = (int)MB_H0_SILENCE(h0c), transient = (int)MB_H0_TRANSIENT(h0c);
      const int M = 1 << lm;
      float* E = sm->E;
      st_celt_end = end;
      if (st_loss_duration == 0) st_skip_plc = 0;      // prepare_frame (celt_decoder.rs:2875-2877)
      MB_SYNC();
      if (C == 1) {   // celt_decoder.rs:2890-2896
        for (int i = tm.tid; i < NB_EBANDS; i += tm.n) E[i] = fmax32(E[i], E[NB_EBANDS + i]);
        MB_SYNC();
      }
      if (!intra && st_loss_duration != 0) {
        // the first coded frame after a loss starts from safer energies (celt_decoder.rs:2943-2969), both channels
        const int missing = imin(10, st_loss_duration >> lm);
        const float safety = lm == 0 ? 1.5f : (lm == 1 ? 0.5f : 0.0f);
        for (int idx = tm.tid; idx < 2 * NB_EBANDS; idx += tm.n) {
          const int i = idx >= NB_EBANDS ? idx - NB_EBANDS : idx;
          if (i < start || i >= end) continue;
          float e0 = E[idx];
          const float e1 = sm->le[idx], e2 = sm->le2[idx];
          if (e0 < fmax32(e1, e2)) {
            float slope = fmax32(e1 - e0, 0.5f * (e2 - e0));
            slope = fmin32(slope, 2.0f);
            e0 -= fmax32((float)(missing + 1) * slope, 0.0f);
            e0 = fmax32(e0, -20.0f);
          } else e0 = fmin32(e0, fmin32(e1, e2));
          E[idx] = e0 - safety;
        }
        MB_SYNC();
      }
      // this frame's symbols: every lane fetches the words of its own (channel, band) slots
      uint32_t wmine[2] = {0, 0};   // team of 32: slots tid and tid + 32; team of one: read again below
      for (int idx = tm.tid, k = 0; idx < 2 * NB_EBANDS; idx += tm.n, k++) {
        const int c = idx >= NB_EBANDS, i = idx - c * NB_EBANDS;
        uint32_t w = 0;
        if (c < C && i >= start && i < end) w = rc[(S_ENERGY + idx) * 32];
        if (k < 2) wmine[k] = w;
        sm->q[idx] = (float)(int)(int16_t)(w & 0xffffu);
      }
      MB_SYNC();
      // coarse + fine + finalise (three separate float additions per band, in the reference's order)
      {
        const float coef = intra ? 0.f : kPredCoef[lm];
        const float beta = intra ? 4915.f / 32768.f : kBetaCoef[lm];
        for (int idx = tm.tid, k = 0; idx < 2 * NB_EBANDS; idx += tm.n, k++) {
          const int c = idx >= NB_EBANDS, i = idx - c * NB_EBANDS;
          if (c >= C || i < start || i >= end) continue;
          const float* qc = sm->q + c * NB_EBANDS;
          float prev = 0.f;
          if (quirks & Q8_COARSE_CLAMP) { for (int j = start; j < i; j++) prev += qc[j] - beta * qc[j]; }
          else for (int j = start; j < i; j++) prev = prev + qc[j] - beta * qc[j];
          const float q = qc[i];
          const float o = fmax32(E[idx], -9.0f);
          float tmp = coef * o + prev + q;
          if (quirks & Q8_COARSE_CLAMP) tmp = tmp < -28.f ? -28.f : (tmp > 28.f ? 28.f : tmp);
          const uint32_t w = (tm.n >= 32 && k < 2) ? wmine[k] : rc[(S_ENERGY + idx) * 32];
          const int fq = (int)((w >> 24) & 15u);
          if (fq > 0) {
            const float scale = (float)(1 << (14 - fq)) * (1.0f / 16384.0f);
            tmp += ((float)(int)((w >> 16) & 255u) + 0.5f) * scale - 0.5f;
          }
          // finalise: every (band, channel) receives at most one extra bit, so the two priority passes of the reference
          // collapse into one addition
          if (w & (1u << 28)) {
            const float scale = (float)(1 << (14 - fq - 1)) * (1.0f / 16384.0f);
            tmp += ((float)(int)((w >> 29) & 1u) - 0.5f) * scale;
          }
          E[idx] = tmp;
        }
      }
      MB_SYNC();
      float* ef = en + (size_t)fc * EN_STRIDE;
      const float increase = (float)imin(st_loss_duration + M, 160) * 0.001f;
      // (two sweeps with the team's barrier between them: a lane reads E[i] of the other channel before anyone rewrites it.
      //  The host emulation, a team of one, runs the first sweep into a scratch copy.)
#if MB_ON_DEVICE
      float ne[2], nl1[2], nl2[2];
#else
      float ne[2 * NB_EBANDS], nl1[2 * NB_EBANDS], nl2[2 * NB_EBANDS];
#endif
      for (int idx = tm.tid, k = 0; idx < 2 * NB_EBANDS; idx += tm.n, k++) {
        const int c = idx >= NB_EBANDS, i = idx - c * NB_EBANDS;
        if (MB_H0_ACON(h0c)) { ef[2 * NB_EBANDS + idx] = sm->le[idx]; ef[4 * NB_EBANDS + idx] = sm->le2[idx]; }
        float e = silence ? -28.0f : E[idx];
        ef[idx] = e;
        ef[EN_GAINS + idx] = celt_exp2(fmin32(e + kEMeans[i], 32.0f));   // denormalise_bands' band gain (bands.rs:4172-4234)
        // state rotation (celt_decoder.rs:4023-4062)
        if (C == 1 && c == 1) e = silence ? -28.0f : E[i];
        float l1 = sm->le[idx], l2 = sm->le2[idx];
        if (transient) l1 = fmin32(l1, e); else { l2 = l1; l1 = e; }
        sm->bg[idx] = fmin32(sm->bg[idx] + increase, e);
        if (i < start || i >= end) { e = 0.f; l1 = l2 = -28.0f; }
        ne[k] = e; nl1[k] = l1; nl2[k] = l2;
      }
      MB_SYNC();
      for (int idx = tm.tid, k = 0; idx < 2 * NB_EBANDS; idx += tm.n, k++) { E[idx] = ne[k]; sm->le[idx] = nl1[k]; sm->le2[idx] = nl2[k]; }
      st_rng = rc[S_RANGE * 32];
      st_loss_duration = 0; st_prefold = 0;
      if (statusc & (1u << 13)) st_celt_error = 1;
      if (lead) meta[fc] = mc;
      emit_frame(fc);
      sample_off += (uint32_t)(M * SHORT_MDCT);                 // the CELT time line advances by the CELT frame
      return M * SHORT_MDCT;
    };
    const bool celt_reset = (hybrid && st_prev_mode != 1001 && st_prev_mode > 0 && !st_prev_redundancy) || (aux == 3 && !hybrid) || trans_to_celt;
    if (MB_ST_FEC(status) && hybrid) {
      // hybrid packet decoded from its LBRR data: SILK from the record, the CELT part concealed (opus_decoder.rs:944-952)
      if (celt_reset) { celt_reset_now(); m.seed_in = 0; }
      const int lmq = (int)MB_H0_LM(h0);
      st_celt_end = (int)MB_ST_END(status);
      celt_conceal(f, r, status, m, SHORT_MDCT << lmq, lmq, 17, 2u | 8u | (celt_reset ? 4u : 0u) | silk_reset | trans_flags);
      silk_side_seen = true; st_fold_serial = 1;
      const int Nq = (SHORT_MDCT << lmq) / quirk_ds(Q.quirks);
      st_prev_mode = 1001; st_prev_redundancy = 0; st_range_final = silk_rng; st_last_duration = Nq;
      if (lead) { meta[f] = m; samples_out[p] = Nq; final_range_out[p] = st_range_final; }
      continue;
    }
    if (hybrid && red) {
      // ---- hybrid packet with a redundancy frame (opus_decoder.rs:746-797, 1012-1080): two CELT frames, the 5 ms redundancy
      // frame (all bands, its own range decoder; parsed into an auxiliary cell) before or after the packet's own.
      const uint32_t aref = r[S_AUXREF * 32], ac = aref >> 3;
      const bool c2s = ((aref >> 1) & 1u) != 0;
      const uint32_t* ra = rec + rec_index((aref & 1u) ? ac : f, 0);
      const uint32_t status_a = ra[S_STATUS * 32];
      if (!(aref & 1u) || !MB_ST_DECODED(status_a)) {
        if (lead) { meta[f] = m; samples_out[p] = (aref & 1u) ? MB_ST_CODE(status_a) : MB_UNIMPLEMENTED; final_range_out[p] = st_range_final; }
        pkt_failed = true;
        continue;
      }
      silk_side_seen = true; st_fold_serial = 1;
      uint32_t red_rng = 0;
      if (c2s) { celt_coded(ac, ra, status_a, 1u | 16u); red_rng = st_rng; }   // on the old CELT state
      if (celt_reset) celt_reset_now();
      uint32_t fl = 1u | 2u | 8u | (celt_reset ? 4u : 0u) | silk_reset | MB_FL_FADE_ROW;
      if (!c2s) fl |= 128u;
      else if (st_prev_mode != 1000 || st_prev_redundancy) fl |= 64u;                 // opus_decoder.rs:1066
      const int N = celt_coded(f, r, status, fl);
      const uint32_t celt_rng = st_rng;
      if (!c2s) { celt_reset_now(); celt_coded(ac, ra, status_a, 1u | 16u | 4u); red_rng = st_rng; }
      if (lead) Q.side_of[f] = ac;
      const int code = MB_ST_CODE(status) < 0 ? MB_ST_CODE(status) : MB_ST_CODE(status_a);
      if (code < 0) {
        if (lead) { samples_out[p] = code; final_range_out[p] = st_range_final; }
      } else {
        st_prev_mode = 1001; st_prev_redundancy = !c2s;
        st_range_final = celt_rng ^ red_rng;
        st_last_duration = N / quirk_ds(Q.quirks);
        if (lead) { samples_out[p] = N / quirk_ds(Q.quirks); final_range_out[p] = st_range_final; }
      }
      continue;
    }
    // the 2.5 ms fade-out frame after a hybrid packet (opus_decoder.rs:868-893), added onto the start of the packet's row; `own`:
    // the flags of the packet's own cell besides the CELT bit
    auto run_silence = [&](uint32_t own) {
      if (sil_aux) {
        const uint32_t sc = asil >> 1;
        const uint32_t* rs = rec + rec_index(sc, 0);
        celt_coded(sc, rs, rs[S_STATUS * 32], 1u | 8u | (own & MB_FL_FADE_ROW));
        m.flags = own | 8u | MB_FL_F32ROW;
        if (lead) { Q.aux_main[sc - Q.aux.base] = f; meta[f] = m; }
      } else celt_coded(f, r, status, 1u | 8u | own);
    };
    if (fadeout && !red && sil_aux) {   // 40 / 60 ms SILK-only packet after a hybrid one
      silk_side_seen = true; st_fold_serial = 1;
      run_silence(2u | silk_reset);
      st_prev_mode = 1000; st_prev_redundancy = 0; st_range_final = silk_rng; st_last_duration = silk_n / quirk_ds(Q.quirks);
      if (lead) { samples_out[p] = silk_n / quirk_ds(Q.quirks); final_range_out[p] = st_range_final; }
      continue;
    }
    if (red_aux) {
      // ---- SILK-only packet whose redundancy frame sits in an auxiliary cell; its own cell may hold the fade-out frame after a
      // hybrid packet. Order of the reference (opus_decoder.rs:1012-1046): CELT -> SILK frame, fade-out, SILK -> CELT frame.
      const uint32_t ac = aref >> 3;
      const bool c2s = ((aref >> 1) & 1u) != 0;
      const uint32_t* ra = rec + rec_index(ac, 0);
      const uint32_t status_a = ra[S_STATUS * 32];
      if (!MB_ST_DECODED(status_a)) {
        if (lead) { meta[f] = m; samples_out[p] = MB_ST_CODE(status_a); final_range_out[p] = st_range_final; }
        pkt_failed = true;
        continue;
      }
      silk_side_seen = true; st_fold_serial = 1;
      const uint32_t fades = !c2s ? 128u : ((st_prev_mode != 1000 || st_prev_redundancy) ? 64u : 0u);   // opus_decoder.rs:1066
      const bool with_silence = fadeout && !(c2s && st_prev_redundancy);                                // :868-870
      const uint32_t own = 2u | silk_reset | fades | MB_FL_FADE_ROW | MB_FL_F32ROW;
      uint32_t red_rng = 0;
      if (c2s) { celt_coded(ac, ra, status_a, 1u | 16u); red_rng = st_rng; }
      if (with_silence) run_silence(own);
      else { m.flags = own; if (lead) meta[f] = m; }
      if (!c2s) { celt_reset_now(); celt_coded(ac, ra, status_a, 1u | 16u | 4u); red_rng = st_rng; }
      if (lead) Q.side_of[f] = ac;
      const int code = MB_ST_CODE(status_a);
      if (code < 0) {
        if (lead) { samples_out[p] = code; final_range_out[p] = st_range_final; }
      } else {
        st_prev_mode = 1000; st_prev_redundancy = !c2s;
        st_range_final = silk_rng ^ red_rng;
        st_last_duration = silk_n / quirk_ds(Q.quirks);
        if (lead) { samples_out[p] = silk_n / quirk_ds(Q.quirks); final_range_out[p] = st_range_final; }
      }
      continue;
    }
    uint32_t trans_in = 0;
    if (trans_to_celt) {
      // ---- SILK-only / hybrid -> CELT-only without a redundancy frame (opus_decoder.rs:575-593, 895-898, 1092-1112): the reference
      // first conceals min(frame, 5 ms) in the old mode — decode_frame(None): the SILK concealment, for a hybrid stream plus the
      // CELT noise concealment from band 17 added on — resets the CELT decoder, decodes the packet and fades from the concealed
      // audio into it over 2.5 ms. Neither decoder's state after that concealment is ever read again (CELT is reset right away,
      // SILK on its next use, :694-696). The concealed audio goes to the side row of an auxiliary cell: the stream's k_silk lane
      // writes the SILK part there, a hybrid stream's CELT part is an ordinary concealed frame in that cell (accumulating side row).
      uint32_t ac = 0xffffffffu;
      if (lead) ac = aux_alloc(Q.aux);
#if MB_ON_DEVICE
      ac = __shfl_sync(0xffffffffu, ac, 0);
#endif
      if (ac == 0xffffffffu) {   // pool of auxiliary cells spent in this pass
        if (lead) { meta[f] = m; samples_out[p] = MB_UNIMPLEMENTED; final_range_out[p] = st_range_final; }
        pkt_failed = true;
        continue;
      }
      const int lmt = imin((int)MB_H0_LM(h0), 1);   // transition_size = min(audiosize, 5 ms)
      if (st_prev_mode == 1001) {
        FrameMeta ma;
        ma.seed_in = st_rng; ma.sample_off = sample_off; ma.lane = lane; ma.flags = 0;
        celt_conceal(ac, rec + rec_index(ac, 0), 0u, ma, SHORT_MDCT << lmt, lmt, 17, 8u | 16u);
        if (lead) meta[ac] = ma;
      }
      if (lead) Q.side_of[f] = ac;
      trans_in = 2u | MB_FL_TRANS_IN | MB_FL_FADE_ROW | MB_FL_F32ROW;
      silk_side_seen = true; st_fold_serial = 1;
    }
    if (celt_reset) celt_reset_now();
    uint32_t fl;
    if (red) {
      if (lead) Q.side_of[f] = f;
      fl = 1u | 2u | 16u | (celt_reset ? 4u : 0u) | silk_reset | MB_FL_FADE_ROW;
      if (aux == 3) fl |= 128u;
      else if (st_prev_mode != 1000 || st_prev_redundancy) fl |= 64u;   // opus_decoder.rs:1066
    } else fl = (hybrid || fadeout) ? (1u | 2u | 8u | (celt_reset ? 4u : 0u) | silk_reset | trans_flags) : ((silk_side_seen ? (1u | 512u) : 1u) | (trans_to_celt ? (4u | trans_in) : 0u));
    if (fl & (8u | 16u)) { silk_side_seen = true; st_fold_serial = 1; }
    const int Nc = celt_coded(f, r, status, fl);
    const int N = (fadeout || red) ? silk_n : Nc;   // samples the packet delivers
    int code = MB_ST_CODE(status);
    if (code < 0) {  // e.g. bit-budget overrun: state has advanced, the call reports the error (celt_decoder.rs:4128)
      if (lead) { samples_out[p] = code; final_range_out[p] = st_range_final; }
    } else {
      st_prev_mode = (fadeout || red) ? 1000 : (hybrid ? 1001 : 1002);
      st_prev_redundancy = aux == 3;
      st_range_final = fadeout ? silk_rng : (red ? (silk_rng ^ st_rng) : st_rng);
      const int Nout = N / quirk_ds(Q.quirks);   // samples at the decoder's output rate (only coded CELT-only frames get here below 48 kHz)
      st_last_duration = Nout;
      if (lead) { samples_out[p] = Nout; final_range_out[p] = st_range_final; }
    }
  }
  MB_SYNC();
  for (int i = tm.tid; i < 2 * NB_EBANDS; i += tm.n) {
    st.old_ebands[i] = sm->E[i]; st.old_log_e[i] = sm->le[i]; st.old_log_e2[i] = sm->le2[i]; st.background_log_e[i] = sm->bg[i];
  }
  if (lead) {
    st.rng = st_rng; st.range_final = st_range_final; st.loss_duration = st_loss_duration; st.prev_mode = st_prev_mode;
    st.last_packet_duration = st_last_duration; st.celt_error = st_celt_error; st.prev_redundancy = st_prev_redundancy;
    st.fold_serial = st_fold_serial; st.bandwidth = st_bandwidth;
    st.skip_plc = st_skip_plc; st.prefilter_and_fold = st_prefold; st.celt_end = st_celt_end;
    // Fold-buffer write-back of the parallel band pass (mb_celt_bands.cuh): walking the lane's CELT-only frames backwards,
    // each one keeps the words of its footprint that no later frame of the pass covers. A frame's footprint is the first
    // C * norm_len words of the reference's storage (bands.rs:2614-2625), capped at what hybrid frames can read.
    // Overlap tails (MB_FL_KEEP_TAIL): the frame in front of a frame concealed by pitch repetition, and the lane's last frame
    // of the pass (the next call may start with a loss), keep the 120 samples their inverse MDCT leaves behind the frame.
    bool need_tail = true;
    for (uint32_t k = n_order; k-- > 0;) {
      const uint32_t f = Q.order[(size_t)k * n_lanes + lane];
      const uint32_t fl = meta[f].flags;
      if (fl & MB_FL_PLC_PITCH) { need_tail = true; continue; }
      if (need_tail) { meta[f].flags = fl | MB_FL_KEEP_TAIL; need_tail = false; }
    }
    int covered = 0;
    for (uint32_t k = n_order; k-- > 0 && covered < (int)FOLD_WORDS;) {
      const uint32_t f = Q.order[(size_t)k * n_lanes + lane];
      const uint32_t fl = meta[f].flags;
      if ((fl & MB_FL_FOLD) || (fl & MB_FL_PLC)) continue;   // (a concealed frame leaves the fold buffer alone)
      const uint32_t h0 = rec[rec_index(f, S_HDR0)];
      const int top = imin((int)MB_H0_C(h0) * ((int)eband(NB_EBANDS - 1) << (int)MB_H0_LM(h0)), (int)FOLD_WORDS);   // even
      if (top > covered) {
        meta[f].flags = fl | ((uint32_t)(covered >> 2) << 10) | ((uint32_t)(top >> 2) << 19);
        covered = top;
      }
    }
    Q.order_n[lane] = n_order;
  }
}

}  // namespace mb
