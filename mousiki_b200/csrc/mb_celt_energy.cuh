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
  int32_t reserved[3];
};

MB_DEV void stream_state_reset(StreamState& s) {
  for (int i = 0; i < 2 * NB_EBANDS; i++) { s.old_ebands[i] = 0.f; s.old_log_e[i] = -28.f; s.old_log_e2[i] = -28.f; s.background_log_e[i] = 0.f; }
  s.preemph_mem[0] = s.preemph_mem[1] = 0.f;
  for (int c = 0; c < 2; c++) { s.pf[c].gain = s.pf[c].gain_old = 0.f; s.pf[c].period = s.pf[c].period_old = 0; s.pf[c].tapset = s.pf[c].tapset_old = 0; }
  s.rng = 0; s.range_final = 0; s.loss_duration = 0; s.prev_mode = 0; s.last_packet_duration = 0;
  s.softclip_mem[0] = s.softclip_mem[1] = 0.f;
  s.celt_error = 0;
  s.reserved[0] = s.reserved[1] = s.reserved[2] = 0;
}

// per-frame side data produced here and consumed by phases 2b / 2c
struct FrameMeta {
  uint32_t seed_in;      // LCG seed for quant_all_bands (= rng left by the previous frame)
  uint32_t sample_off;   // first sample of this frame in the lane's time series (excluding HIST)
  uint32_t lane;         // lane (stream slot) of this call
  uint32_t flags;        // bit0: frame runs through phases 2b/2c; bit1: SILK synthesis (k_silk); bit2: CELT state reset before
                         // this frame; bit3: the CELT output is added to the SILK output already in the PCM row (hybrid)
};
enum { EN_STRIDE = 3 * 2 * NB_EBANDS };  // per-frame energy scratch: E, prev1 (old_log_e), prev2 (old_log_e2)

// One lane = one stream. Frame (grid cell) of tick t is f = t * n_lanes + lane; its packet index (for the
// per-packet outputs) is pkt_map ? pkt_map[f] : f. Cells without a packet carry S_STATUS bit 15.
MB_DEV void celt_energy_lane(uint32_t lane, uint32_t n_lanes, uint32_t n_ticks, const uint32_t* pkt_map, const uint32_t* rec,
                             StreamState* st_all, const uint32_t* lane_stream, float* en, FrameMeta* meta, int32_t* samples_out,
                             uint32_t* final_range_out, uint32_t quirks, const uint8_t* flags = nullptr) {
  StreamState& st = st_all[lane_stream ? lane_stream[lane] : lane];
  uint32_t sample_off = 0;
  bool pkt_failed = false;   // an earlier frame of the same multi-frame packet failed: the reference returns there
                             // (src/opus_decoder.rs:1339-1350), so the remaining frames must not touch the stream
  for (uint32_t t = 0; t < n_ticks; t++) {
    const uint32_t f = t * n_lanes + lane;
    const uint32_t* r = rec + rec_index(f, 0);
    uint32_t status = r[S_STATUS * 32];
    FrameMeta m;
    m.seed_in = st.rng; m.sample_off = sample_off; m.lane = lane; m.flags = 0;
    if (MB_ST_EMPTY(status)) { meta[f] = m; continue; }
    const uint32_t p = pkt_map ? pkt_map[f] : f;
    const bool continuation = flags && (flags[p] & 4u);
    if (!continuation) pkt_failed = false;
    if (pkt_failed || !MB_ST_DECODED(status)) {  // rejected before the decoder ran (unsupported mode, PLC, bad TOC)
      meta[f] = m;
      samples_out[p] = pkt_failed ? MB_UNIMPLEMENTED : MB_ST_CODE(status);
      final_range_out[p] = st.range_final;
      pkt_failed = true;
      continue;
    }
    uint32_t h0 = r[S_HDR0 * 32];
    // Mode switches that go through decode_frame(None) = concealment (opus_decoder.rs:575-593, 798-809: CELT-only <->
    // SILK-only / hybrid) or through the 2.5 ms CELT silence frame (hybrid -> SILK-only, :868-893) are rows F1/F2: reported
    // per packet without touching the stream.
    const bool silk_bit = ((status >> 16) & 1u) != 0, hybrid = ((status >> 17) & 1u) != 0;
    // bit 18: a SILK-only cell in the hybrid record layout whose CELT slots hold the 2.5 ms silence frame that lets the CELT
    // overlap fade out after a hybrid packet (opus_decoder.rs:868-893; parsed on a hint, see parse_cell)
    const bool has_fadeout = ((status >> 18) & 1u) != 0;
    const bool fadeout = has_fadeout && st.prev_mode == 1001;
    const bool silk_only = silk_bit && !hybrid && !fadeout;
    if ((!silk_bit && (st.prev_mode == 1000 || st.prev_mode == 1001)) || (silk_bit && st.prev_mode == 1002) ||
        (silk_only && st.prev_mode == 1001)) {
      meta[f] = m;
      samples_out[p] = MB_UNIMPLEMENTED;
      final_range_out[p] = st.range_final;
      pkt_failed = true;
      continue;
    }
    // SILK side of the cell: header and final range sit in the hybrid layout when the cell carries the fade-out record
    const uint32_t silk_h0 = has_fadeout ? r[S_SILK_HYB * 32] : h0;
    const int silk_n = (int)((silk_h0 >> 8) & 3u) * (int)((silk_h0 >> 5) & 7u) * 240;   // frames x subframes: 2, 4, 8, 12 -> 10 .. 60 ms
    if (silk_only) {  // SILK-only packet: synthesis runs in k_silk (mb_silk_out.cuh); here only the front-end bookkeeping
      const int N = silk_n;
      m.flags = 2;
      meta[f] = m;
      st.prev_mode = 1000;
      st.range_final = has_fadeout ? r[(S_SILK_HYB + 342) * 32] : r[S_RANGE * 32];
      st.last_packet_duration = N;
      samples_out[p] = N;
      final_range_out[p] = st.range_final;
      continue;
    }
    // SILK-only -> hybrid: the CELT decoder is reset before the frame (opus_decoder.rs:895-898). The energy side is reset
    // here; the tail kernels reset their own state (post-filter, history, de-emphasis memory) on flag bit 2.
    const bool celt_reset = hybrid && st.prev_mode == 1000;
    if (celt_reset) {
      for (int i = 0; i < 2 * NB_EBANDS; i++) { st.old_ebands[i] = 0.f; st.old_log_e[i] = -28.f; st.old_log_e2[i] = -28.f; st.background_log_e[i] = 0.f; }
      st.rng = 0; st.loss_duration = 0;
      m.seed_in = 0;
    }
    const int lm = (int)MB_H0_LM(h0), C = (int)MB_H0_C(h0), start = (int)MB_H0_START(h0), end = (int)MB_ST_END(status);
    const int intra = (int)MB_H0_INTRA(h0), silence = (int)MB_H0_SILENCE(h0), transient = (int)MB_H0_TRANSIENT(h0);
    const int M = 1 << lm;
    float* E = st.old_ebands;
    if (C == 1)
      for (int i = 0; i < NB_EBANDS; i++) E[i] = fmax32(E[i], E[NB_EBANDS + i]);
    // coarse + fine + finalise (three separate float additions per band, in the reference's order)
    {
      float prev[2] = {0.f, 0.f};
      const float coef = intra ? 0.f : kPredCoef[lm];
      const float beta = intra ? 4915.f / 32768.f : kBetaCoef[lm];
      for (int i = start; i < end; i++)
        for (int c = 0; c < C; c++) {
          int idx = c * NB_EBANDS + i;
          float q = (float)(int)(int16_t)(r[(S_ENERGY + idx) * 32] & 0xffffu);
          float o = fmax32(E[idx], -9.0f);
          float tmp = coef * o + prev[c] + q;
          if (quirks & Q8_COARSE_CLAMP) {
            tmp = tmp < -28.f ? -28.f : (tmp > 28.f ? 28.f : tmp);
            prev[c] += q - beta * q;
          } else prev[c] = prev[c] + q - beta * q;
          E[idx] = tmp;
        }
      for (int i = start; i < end; i++)
        for (int c = 0; c < C; c++) {
          int idx = c * NB_EBANDS + i;
          uint32_t w = r[(S_ENERGY + idx) * 32];
          int fq = (int)((w >> 24) & 15u);
          if (fq > 0) {
            float scale = (float)(1 << (14 - fq)) * (1.0f / 16384.0f);
            E[idx] += ((float)(int)((w >> 16) & 255u) + 0.5f) * scale - 0.5f;
          }
        }
      // finalise: every (band, channel) receives at most one extra bit, so the two priority passes
      // of the reference collapse into one sweep
      for (int i = start; i < end; i++)
        for (int c = 0; c < C; c++) {
          int idx = c * NB_EBANDS + i;
          uint32_t w = r[(S_ENERGY + idx) * 32];
          if (w & (1u << 28)) {
            int fq = (int)((w >> 24) & 15u);
            float scale = (float)(1 << (14 - fq - 1)) * (1.0f / 16384.0f);
            E[idx] += ((float)(int)((w >> 29) & 1u) - 0.5f) * scale;
          }
        }
    }
    float* ef = en + (size_t)f * EN_STRIDE;
    if (MB_H0_ACON(h0))
      for (int i = 0; i < 2 * NB_EBANDS; i++) { ef[2 * NB_EBANDS + i] = st.old_log_e[i]; ef[4 * NB_EBANDS + i] = st.old_log_e2[i]; }
    if (silence)
      for (int i = 0; i < 2 * NB_EBANDS; i++) E[i] = -28.0f;
    for (int i = 0; i < 2 * NB_EBANDS; i++) ef[i] = E[i];
    // state rotation (celt_decoder.rs:4023-4062)
    if (C == 1)
      for (int i = 0; i < NB_EBANDS; i++) E[NB_EBANDS + i] = E[i];
    if (transient) {
      for (int i = 0; i < 2 * NB_EBANDS; i++) st.old_log_e[i] = fmin32(st.old_log_e[i], E[i]);
    } else {
      for (int i = 0; i < 2 * NB_EBANDS; i++) { st.old_log_e2[i] = st.old_log_e[i]; st.old_log_e[i] = E[i]; }
    }
    float increase = (float)imin(st.loss_duration + M, 160) * 0.001f;
    for (int i = 0; i < 2 * NB_EBANDS; i++) st.background_log_e[i] = fmin32(st.background_log_e[i] + increase, E[i]);
    for (int c = 0; c < 2; c++) {
      for (int i = 0; i < start; i++) { E[c * NB_EBANDS + i] = 0.f; st.old_log_e[c * NB_EBANDS + i] = st.old_log_e2[c * NB_EBANDS + i] = -28.0f; }
      for (int i = end; i < NB_EBANDS; i++) { E[c * NB_EBANDS + i] = 0.f; st.old_log_e[c * NB_EBANDS + i] = st.old_log_e2[c * NB_EBANDS + i] = -28.0f; }
    }
    st.rng = r[S_RANGE * 32];
    st.loss_duration = 0;
    if (status & (1u << 13)) st.celt_error = 1;
    const int N = fadeout ? silk_n : M * SHORT_MDCT;          // samples the packet delivers
    m.flags = (hybrid || fadeout) ? (1u | 2u | 8u | (celt_reset ? 4u : 0u)) : 1u;   // bit0 CELT, bit1 SILK, bit2 CELT reset, bit3 accumulate
    meta[f] = m;
    sample_off += (uint32_t)(M * SHORT_MDCT);                 // the CELT time line advances by the CELT frame
    int code = MB_ST_CODE(status);
    if (code < 0) {  // e.g. bit-budget overrun: state has advanced, the call reports the error (celt_decoder.rs:4128)
      samples_out[p] = code;
      final_range_out[p] = st.range_final;
    } else {
      st.prev_mode = fadeout ? 1000 : (hybrid ? 1001 : 1002);
      st.range_final = fadeout ? r[(S_SILK_HYB + 342) * 32] : st.rng;
      st.last_packet_duration = N;
      samples_out[p] = N;
      final_range_out[p] = st.range_final;
    }
  }
}

}  // namespace mb
