// Row A2 — mousiki's LIGHTWEIGHT decoder, batched: one lane per stream, a stream's packets in order.
//
// Replaces (per stream):
//   decoder::Decoder::{decode, decode_float32, decode_int24}   src/decoder.rs:142-262
//   silk::decoder::Decoder::decode                               src/silk/decoder.rs:971-1163 (+ :213-969, :1166-1800)
//   resample::up (zero-order hold x3)                            src/resample.rs:3-15
//   convert_float32_le_to_signed16_le / _to_signed24             src/bitdepth.rs:15-99
// This is the float, pion-derived SILK decoder behind the README's "lightweight" API: SILK-only, mono,
// one 20 ms frame per packet, no LBRR; it is NOT bit-compatible with the full decoder's fixed-point SILK path (k_silk).
// Every packet yields 320 samples of the stream's persistent `silk_buffer` held three times each (960 samples): a NB / MB
// frame rewrites only the first 160 / 240 of them, exactly as the reference does.
//
// The entropy walk is inherently serial and the volume of this path is small ("trivial volume"), so a lane
// runs a whole frame: symbols through the shared range decoder (mb_rangedec.cuh) with the reference's distributions in
// inverse-cumulative byte form (lite_tables_data.h), integer NLSF -> LPC, then the float LTP / LPC recursions in the
// reference's operation order (no contraction: -fmad=false), which is what makes the output bit-identical.
#pragma once
#include "mb_silk_parse.cuh"
#include "lite_tables_data.h"

namespace mb {

enum { LITE_BUF = 320, LITE_FINAL = 306, LITE_MAX_RES_LAG = 290, LITE_OUT = 960 };
// per-packet status: 960, or the reference's error (DecoderError / silk DecodeError variants; include/mousiki_b200.h)
enum { LITE_TOO_SHORT = -1, LITE_FRAME_CODE = -2, LITE_CONFIG_MODE = -3, LITE_FRAME_DURATION = -4, LITE_STEREO = -5, LITE_LBRR = -7 };

struct LiteState {   // src/silk/decoder.rs:84-94, src/decoder.rs:137-140
  float silk_buffer[LITE_BUF];
  float final_out[LITE_FINAL];
  float prev_lpc[16];
  int16_t n0_q15[16];
  int32_t prev_lpc_len, previous_log_gain, have_decoded;
};
MB_DEV void lite_state_reset(LiteState& s) {
  for (int i = 0; i < LITE_BUF; i++) s.silk_buffer[i] = 0.f;
  for (int i = 0; i < LITE_FINAL; i++) s.final_out[i] = 0.f;
  for (int i = 0; i < 16; i++) { s.prev_lpc[i] = 0.f; s.n0_q15[i] = 0; }
  s.prev_lpc_len = 0; s.previous_log_gain = 0; s.have_decoded = 0;
}

struct LiteFrame {   // what the entropy walk hands to the synthesis
  float gain_q16[4], aq12[2][16], ltp_scale_q14;
  int32_t exc_q23[LITE_BUF];
  int pitch[4];
  int8_t taps[4][5];
  int n_aq, d_lpc, w_q2, voiced, lag_max, n;
};

MB_DEV int lite_ilog64(uint64_t x) { int n = 0; while (x) { n++; x >>= 1; } return n; }   // src/math.rs:3-8 (64 - leading_zeros)
MB_DEV float lite_clamp1(float v) { return v < -1.0f ? -1.0f : (v > 1.0f ? 1.0f : v); }

// NLSF_Q15 -> Q12 prediction coefficients as floats (src/silk/decoder.rs:1424-1553: cosine interpolation, the P / Q
// polynomial recursion, up to ten rounds of bandwidth expansion; the "prediction gain" step only rounds to Q12)
MB_DEV void lite_nlsf_to_aq12(const int16_t* nq, int d, bool wide, float* aq12) {
  using namespace lite_tab;
  int32_t c[16], p[10], q[10], a[16];
  for (int k = 0; k < 16; k++) c[k] = 0;
  for (int k = 0; k < d; k++) {
    const int i = nq[k] >> 8, f = nq[k] & 255;
    const int32_t cv = Q12_COSINE_TABLE_FOR_LSFCONVERION[i], cn = Q12_COSINE_TABLE_FOR_LSFCONVERION[i + 1];
    const int o = wide ? LSF_ORDERING_FOR_POLYNOMIAL_EVALUATION_WIDEBAND[k] : LSF_ORDERING_FOR_POLYNOMIAL_EVALUATION_NARROWBAND_AND_MEDIUMBAND[k];
    c[o] = (cv * 256 + (cn - cv) * f + 4) >> 3;
  }
  const int d2 = d / 2;
  for (int k = 0; k < 10; k++) p[k] = q[k] = 0;
  p[0] = q[0] = 1 << 16; p[1] = -c[0]; q[1] = -c[1];
  for (int k = 1; k < d2; k++) {
    const int64_t ce = c[2 * k], co = c[2 * k + 1];
    p[k + 1] = p[k - 1] * 2 - (int32_t)((ce * (int64_t)p[k] + 32768) >> 16);
    q[k + 1] = q[k - 1] * 2 - (int32_t)((co * (int64_t)q[k] + 32768) >> 16);
    for (int j = k; j >= 2; j--) {
      p[j] += p[j - 2] - (int32_t)((ce * (int64_t)p[j - 1] + 32768) >> 16);
      q[j] += q[j - 2] - (int32_t)((co * (int64_t)q[j - 1] + 32768) >> 16);
    }
    p[1] -= c[2 * k];
    q[1] -= c[2 * k + 1];
  }
  for (int k = 0; k < d2; k++) {
    const int32_t dq = q[k + 1] - q[k], sp = p[k + 1] + p[k];
    a[k] = -dq - sp;
    a[d - k - 1] = dq - sp;
  }
  int round = 0;
  for (; round < 10; round++) {
    uint32_t maxabs = 0;
    int at = 0;
    for (int i = 0; i < d; i++) {
      const uint32_t v = a[i] < 0 ? 0u - (uint32_t)a[i] : (uint32_t)a[i];
      if (v > maxabs) { maxabs = v; at = i; }
    }
    uint32_t m12 = (maxabs + 16) >> 5;
    if (m12 > 163838u) m12 = 163838u;
    if (m12 <= 32767u) break;
    uint32_t den = (m12 * (uint32_t)(at + 1)) >> 2;
    if (den < 1) den = 1;
    const uint32_t sc0 = 65470u - ((m12 - 32767u) << 14) / den;
    uint32_t sc = sc0;
    for (int k = 0; k < d; k++) {
      a[k] = (int32_t)(((int64_t)a[k] * (int64_t)sc) >> 16);
      sc = (sc0 * sc + 32768u) >> 16;
    }
  }
  if (round == 9)
    for (int i = 0; i < d; i++) {
      int32_t v = (a[i] + 16) >> 5;
      v = v < -32768 ? -32768 : (v > 32767 ? 32767 : v);
      a[i] = v << 5;
    }
  for (int k = 0; k < d; k++) aq12[k] = (float)((a[k] + 16) >> 5);
}

// The entropy walk of one SILK frame (src/silk/decoder.rs:1001-1118) and everything integer that follows from it.
// Returns 0, or LITE_LBRR before any state is touched.
MB_DEV int lite_parse(RangeDec& ec, LiteState& st, int bw, LiteFrame& F) {
  using namespace lite_tab;
  const bool wide = bw == 2;
  const int vad = ec.bit_logp(1), lbrr = ec.bit_logp(1);
  if (lbrr) return LITE_LBRR;
  int signal, qoff;   // 0 inactive, 1 unvoiced, 2 voiced; 0 low, 1 high (:225-266)
  if (!vad) { signal = 0; qoff = ec.icdf(FRAME_TYPE_VAD_INACTIVE, 8) == 0 ? 0 : 1; }
  else { const int t = ec.icdf(FRAME_TYPE_VAD_ACTIVE, 8); signal = t < 2 ? 1 : 2; qoff = t & 1; }
  F.voiced = signal == 2;
  // subframe gains (:269-362)
  for (int s = 0; s < 4; s++) {
    int lg;
    if (s == 0) {
      const uint8_t* msb = signal == 0 ? INDEPENDENT_QUANTIZATION_GAIN_MSB_INACTIVE : signal == 2 ? INDEPENDENT_QUANTIZATION_GAIN_MSB_VOICED
                                                                                                  : INDEPENDENT_QUANTIZATION_GAIN_MSB_UNVOICED;
      int gi = ec.icdf(msb, 8) << 3;
      gi |= ec.icdf(INDEPENDENT_QUANTIZATION_GAIN_LSB, 8);
      lg = st.have_decoded ? imax(gi, st.previous_log_gain - 16) : gi;
    } else {
      const int dl = ec.icdf(DELTA_QUANTIZATION_GAIN, 8);
      lg = imin(63, imax(0, imax(2 * dl - 16, st.previous_log_gain + dl - 4)));
    }
    st.previous_log_gain = lg;
    const int q7 = ((0x1D1C71 * lg) >> 16) + 2090, i = q7 >> 7, f = q7 & 127;
    F.gain_q16[s] = (float)((1 << i) + (((-174 * f * (128 - f)) >> 16) + f) * ((1 << i) >> 7));
  }
  // normalised LSFs: stage 1, stage 2 with backwards prediction, reconstruction, stabilisation (:364-517, :1555-1795)
  const int d = wide ? 16 : 10;
  F.d_lpc = d;
  const int i1 = ec.icdf(wide ? (F.voiced ? NORMALIZED_LSF_STAGE_1_INDEX_WIDEBAND_VOICED : NORMALIZED_LSF_STAGE_1_INDEX_WIDEBAND_UNVOICED)
                              : (F.voiced ? NORMALIZED_LSF_STAGE_1_INDEX_NARROWBAND_OR_MEDIUMBAND_VOICED
                                          : NORMALIZED_LSF_STAGE_1_INDEX_NARROWBAND_OR_MEDIUMBAND_UNVOICED), 8);
  int i2[16];
  for (int k = 0; k < d; k++) {
    const int ctx = wide ? NORMALIZED_LSF_STAGE_TWO_INDEX_WIDEBAND[i1][k] : NORMALIZED_LSF_STAGE_TWO_INDEX_NARROWBAND_OR_MEDIUMBAND[i1][k];
    int v = ec.icdf(NORMALIZED_LSF_STAGE_2_INDEX + NORMALIZED_LSF_STAGE_2_INDEX_OFF[ctx], 8) - 4;
    if (v == -4) v -= ec.icdf(NORMALIZED_LSF_STAGE_2_INDEX_EXTENSION, 8);
    else if (v == 4) v += ec.icdf(NORMALIZED_LSF_STAGE_2_INDEX_EXTENSION, 8);
    i2[k] = v;
  }
  int16_t nlsf[16];
  {
    const int qstep = wide ? 9830 : 11796;
    int res_next = 0;   // res_Q10[k + 1]
    int16_t res[16];
    for (int k = d - 1; k >= 0; k--) {
      int first = 0;
      if (k + 1 < d) {
        const int sel = wide ? PREDICTION_WEIGHT_SELECTION_FOR_WIDEBAND_NORMALIZED_LSF[i1][k] : PREDICTION_WEIGHT_SELECTION_FOR_NARROWBAND_AND_MEDIUMBAND_NORMALIZED_LSF[i1][k];
        const int pred = wide ? PREDICTION_WEIGHT_FOR_WIDEBAND_NORMALIZED_LSF[sel][k] : PREDICTION_WEIGHT_FOR_NARROWBAND_AND_MEDIUMBAND_NORMALIZED_LSF[sel][k];
        first = (res_next * pred) >> 8;
      }
      const int v = i2[k], sg = (v > 0) - (v < 0);
      res[k] = (int16_t)(first + ((((v << 10) - sg * 102) * qstep) >> 16));
      res_next = res[k];
    }
    for (int k = 0; k < d; k++) {
      const int cb = wide ? NORMALIZED_LSF_STAGE_ONE_WIDEBAND[i1][k] : NORMALIZED_LSF_STAGE_ONE_NARROWBAND_OR_MEDIUMBAND[i1][k];
      const int lo = k != 0 ? (wide ? NORMALIZED_LSF_STAGE_ONE_WIDEBAND[i1][k - 1] : NORMALIZED_LSF_STAGE_ONE_NARROWBAND_OR_MEDIUMBAND[i1][k - 1]) : 0;
      const int hi = k + 1 != d ? (wide ? NORMALIZED_LSF_STAGE_ONE_WIDEBAND[i1][k + 1] : NORMALIZED_LSF_STAGE_ONE_NARROWBAND_OR_MEDIUMBAND[i1][k + 1]) : 256;
      const uint64_t w2 = (uint64_t)(1024 / (cb - lo) + 1024 / (hi - cb)) << 16;
      const int il = lite_ilog64(w2);
      const int64_t f = (int64_t)((w2 >> (il - 8)) & 127);
      int64_t y = (il & 1) ? 32768 : 46214;
      y >>= (32 - il) >> 1;
      const int16_t w_q9 = (int16_t)(y + ((213 * f * y) >> 16));
      const int32_t r = (cb << 7) + ((int32_t)res[k] << 14) / (int32_t)w_q9;
      nlsf[k] = (int16_t)(r < 0 ? 0 : (r > 32767 ? 32767 : r));
    }
    const int16_t* nd = wide ? MINIMUM_SPACING_FOR_NORMALIZED_LSCOEFFICIENTS_WIDEBAND : MINIMUM_SPACING_FOR_NORMALIZED_LSCOEFFICIENTS_NARROWBAND_AND_MEDIUMBAND;
    bool settled = false;
    for (int adj = 0; adj <= 19 && !settled; adj++) {
      int at = 0, worst = 0x7fffffff;
      for (int x = 0; x <= d; x++) {
        const int sp = (x == d ? 32768 : nlsf[x]) - (x != 0 ? nlsf[x - 1] : 0) - nd[x];
        if (sp < worst) { at = x; worst = sp; }
      }
      if (worst >= 0) { settled = true; break; }
      if (at == 0) { nlsf[0] = nd[0]; continue; }
      if (at == d) { nlsf[d - 1] = (int16_t)(32768 - nd[d]); continue; }
      int lo = nd[at] >> 1, hi = 32768 - (nd[at] >> 1);
      for (int k = 0; k < at; k++) lo += nd[k];
      for (int k = at + 1; k <= d; k++) hi -= nd[k];
      int c = (nlsf[at - 1] + nlsf[at] + 1) >> 1;
      c = c < lo ? lo : (c > hi ? hi : c);
      nlsf[at - 1] = (int16_t)(c - (nd[at] >> 1));
      nlsf[at] = (int16_t)(nlsf[at - 1] + nd[at]);
    }
    if (!settled) {   // the fallback after 20 adjustments: sort, then enforce the spacing from both ends
      for (int i = 1; i < d; i++) {
        const int16_t cur = nlsf[i];
        int j = i;
        while (j > 0 && nlsf[j - 1] > cur) { nlsf[j] = nlsf[j - 1]; j--; }
        nlsf[j] = cur;
      }
      for (int k = 0; k < d; k++) {
        const int16_t lo = (int16_t)((k != 0 ? nlsf[k - 1] : 0) + nd[k]);
        if (nlsf[k] < lo) nlsf[k] = lo;
      }
      for (int k = d - 1; k >= 0; k--) {
        const int16_t hi = (int16_t)((k == d - 1 ? 32768 : nlsf[k + 1]) - nd[k + 1]);
        if (nlsf[k] > hi) nlsf[k] = hi;
      }
    }
  }
  // interpolation with the previous frame's LSFs (:520-545), LPC sets (:547-561)
  F.w_q2 = ec.icdf(NORMALIZED_LSF_INTERPOLATION_INDEX, 8);
  F.n_aq = 0;
  if (F.w_q2 != 4 && st.have_decoded) {
    int16_t n1[16];
    for (int k = 0; k < d; k++) {
      const int prev = st.n0_q15[k];
      n1[k] = (int16_t)(prev + ((F.w_q2 * ((int)nlsf[k] - prev)) >> 2));
    }
    lite_nlsf_to_aq12(n1, d, wide, F.aq12[F.n_aq++]);
  }
  lite_nlsf_to_aq12(nlsf, d, wide, F.aq12[F.n_aq++]);
  for (int k = 0; k < d; k++) st.n0_q15[k] = nlsf[k];   // (the reference stores them at the end of decode; nothing reads them in between)
  // pitch lags, LTP filters and scaling (:563-675)
  F.lag_max = 0;
  F.ltp_scale_q14 = 15565.0f;
  if (F.voiced) {
    const int khz = bw == 0 ? 8 : bw == 1 ? 12 : 16;
    const int hi = ec.icdf(PRIMARY_PITCH_LAG_HIGH_PART, 8);
    const int lo = ec.icdf(bw == 0 ? PRIMARY_PITCH_LAG_LOW_PART_NARROWBAND : bw == 1 ? PRIMARY_PITCH_LAG_LOW_PART_MEDIUMBAND : PRIMARY_PITCH_LAG_LOW_PART_WIDEBAND, 8);
    const int lag_index = (int16_t)(hi * (khz >> 1) + lo);
    const int contour = (int8_t)ec.icdf(bw == 0 ? SUBFRAME_PITCH_CONTOUR_NARROWBAND20_MS : SUBFRAME_PITCH_CONTOUR_MEDIUMBAND_OR_WIDEBAND20_MS, 8);
    const int min_lag = 2 * khz, max_lag = 18 * khz, lag = min_lag + lag_index;   // silk_decode_pitch, src/silk/decode_pitch.rs:22-89
    for (int k = 0; k < 4; k++) {
      const int off = khz == 8 ? silk_CB_lags_stage2[k * 11 + contour] : silk_CB_lags_stage3[k * 34 + contour];
      F.pitch[k] = imin(max_lag, imax(min_lag, lag + off));
    }
    F.lag_max = max_lag;
    const int per = ec.icdf(PERIODICITY_INDEX, 8);
    for (int s = 0; s < 4; s++) {
      const int fi = ec.icdf(per == 0 ? LTP_FILTER_INDEX0 : per == 1 ? LTP_FILTER_INDEX1 : LTP_FILTER_INDEX2, 8);
      for (int t = 0; t < 5; t++)
        F.taps[s][t] = per == 0 ? CODEBOOK_LTP_FILTER_PERIODICITY_INDEX_0[fi][t] : per == 1 ? CODEBOOK_LTP_FILTER_PERIODICITY_INDEX_1[fi][t]
                                                                                             : CODEBOOK_LTP_FILTER_PERIODICITY_INDEX_2[fi][t];
    }
    const int si = ec.icdf(LTP_SCALING_PARAMETER, 8);
    F.ltp_scale_q14 = si == 0 ? 15565.0f : si == 1 ? 12288.0f : 8192.0f;
  }
  uint32_t seed = (uint32_t)ec.icdf(LINEAR_CONGRUENTIAL_GENERATOR_SEED, 8);
  // excitation (:1166-1422): pulse counts per shell block, positions by recursive halving, LSBs, signs, then the Q23 values
  const int blocks = bw == 0 ? 10 : bw == 1 ? 15 : 20, len = blocks * 16;
  const int rate = ec.icdf(F.voiced ? RATE_LEVEL_VOICED : RATE_LEVEL_UNVOICED, 8);
  uint8_t count[20], lsbs[20];
  for (int b = 0; b < blocks; b++) {
    int pc = ec.icdf(PULSE_COUNT + PULSE_COUNT_OFF[imin(rate, 10)], 8), nl = 0;
    if (pc == 17) {
      while (pc == 17 && nl < 10) { pc = ec.icdf(PULSE_COUNT + PULSE_COUNT_OFF[9], 8); nl++; }
      if (nl == 10) pc = ec.icdf(PULSE_COUNT + PULSE_COUNT_OFF[10], 8);
    }
    count[b] = (uint8_t)pc; lsbs[b] = (uint8_t)nl;
  }
  int32_t* e = F.exc_q23;
  for (int i = 0; i < len; i++) e[i] = 0;
  for (int b = 0; b < blocks; b++) {
    if (count[b] == 0) continue;
    // breadth of 16 -> 8 -> 4 -> 2 -> 1 in the reference's depth-first order: an explicit stack of (count, width)
    int pos = b * 16, sp = 0;
    uint8_t stk_c[8], stk_w[8];
    stk_c[0] = count[b]; stk_w[0] = 16; sp = 1;
    while (sp) {
      const int cnt = stk_c[--sp], w = stk_w[sp];
      if (w == 1) { e[pos++] = cnt; continue; }
      int left = 0;
      if (cnt != 0) {
        const uint8_t* tab = w == 16 ? PULSE_COUNT_SPLIT16_SAMPLE_PARTITIONS : w == 8 ? PULSE_COUNT_SPLIT8_SAMPLE_PARTITIONS
                             : w == 4 ? PULSE_COUNT_SPLIT4_SAMPLE_PARTITIONS : PULSE_COUNT_SPLIT2_SAMPLE_PARTITIONS;
        const uint16_t* off = w == 16 ? PULSE_COUNT_SPLIT16_SAMPLE_PARTITIONS_OFF : w == 8 ? PULSE_COUNT_SPLIT8_SAMPLE_PARTITIONS_OFF
                              : w == 4 ? PULSE_COUNT_SPLIT4_SAMPLE_PARTITIONS_OFF : PULSE_COUNT_SPLIT2_SAMPLE_PARTITIONS_OFF;
        left = ec.icdf(tab + off[imin(cnt - 1, 15)], 8);
      }
      const int right = cnt > left ? cnt - left : 0;
      // the right half waits while the left half is walked to its leaves
      stk_c[sp] = (uint8_t)right; stk_w[sp++] = (uint8_t)(w >> 1);
      stk_c[sp] = (uint8_t)left; stk_w[sp++] = (uint8_t)(w >> 1);
    }
  }
  for (int i = 0; i < len; i++)
    for (int l = 0; l < lsbs[i >> 4]; l++) e[i] = (e[i] << 1) | ec.icdf(EXCITATION_LSB, 8);
  {   // src/silk/code_signs.rs:84-125, sum = count | lsbs << 5 (src/silk/decoder.rs:133-141)
    const uint8_t* base = silk_sign_iCDF + 7 * (qoff + (signal << 1));
    for (int b = 0; b < blocks; b++) {
      const int total = (int)count[b] | ((int)lsbs[b] << 5);
      if (total <= 0) continue;
      uint8_t tab[2];
      tab[0] = base[imin(total & 0x1F, 6)]; tab[1] = 0;
      for (int i = b * 16; i < b * 16 + 16; i++)
        if ((int16_t)e[i] > 0) e[i] = (int16_t)e[i] * (int16_t)((ec.icdf(tab, 8) << 1) - 1);
        else e[i] = (int16_t)e[i];
    }
  }
  const int offset_q23 = signal == 2 ? (qoff ? 25 : 8) : (qoff ? 60 : 25);
  for (int i = 0; i < len; i++) {
    const int32_t raw = e[i];
    int32_t v = (raw << 8) - ((raw > 0) - (raw < 0)) * 20 + offset_q23;
    seed = seed * 196314165u + 907633515u;
    if (seed & 0x80000000u) v = -v;
    seed += (uint32_t)raw;
    e[i] = v;
  }
  F.n = bw == 0 ? 40 : bw == 1 ? 60 : 80;
  return 0;
}

// Frame reconstruction (src/silk/decoder.rs:689-969). `res` holds the excitation / LTP residual of the frame, `lag` the
// rewhitened history in front of it (res_lag), `lpc` the unclamped synthesis output; out = the stream's silk_buffer.
MB_DEV void lite_synthesis(LiteState& st, const LiteFrame& F, float* res, float* lag, float* lpc) {
  float* out = st.silk_buffer;
  const int n = F.n, frame = 4 * n, d = F.d_lpc;
  for (int i = 0; i < frame; i++) res[i] = (float)F.exc_q23[i] * (1.0f / 8388608.0f);
  for (int i = 0; i < frame; i++) lpc[i] = 0.f;
  const int lag_len = imin(F.lag_max + 2, (int)LITE_MAX_RES_LAG);
  for (int i = 0; i < lag_len; i++) lag[i] = 0.f;
  for (int s = 0; s < 4; s++) {
    const float* aq = F.aq12[(s > 1 && F.n_aq > 1) ? 1 : 0];
    const int j = n * s;
    const float gain = F.gain_q16[s];
    if (F.voiced) {   // ltp_synthesis, :689-827
      float scale = F.ltp_scale_q14;
      int out_end;
      if (s < 2 || F.w_q2 == 4) out_end = -s * n;
      else { scale = 16384.0f; out_end = -(s - 2) * n; }
      const int pitch = F.pitch[s];
      for (int i = -pitch - 2; i < out_end; i++) {   // rewhiten the signal the long-term predictor reaches back into
        const int idx = i + j;
        if (idx >= frame) continue;
        float v;
        if (idx >= 0) v = out[idx];
        else {
          if (lag_len + idx < 0) continue;
          v = LITE_FINAL + idx < 0 ? 0.0f : st.final_out[LITE_FINAL + idx];
        }
        for (int k = 0; k < d; k++) {
          const int oi = idx - k - 1;
          const float ov = oi >= 0 ? out[oi] : (LITE_FINAL + oi < 0 ? 0.0f : st.final_out[LITE_FINAL + oi]);
          v -= ov * (aq[k] / 4096.0f);
        }
        v = lite_clamp1(v);
        if (gain != 0.0f) {
          v *= (4.0f * scale) / gain;
          if (idx >= 0) res[idx] = v; else lag[lag_len + idx] = v;
        }
      }
      if (s > 0 && gain != 0.0f) {
        const float g = F.gain_q16[s - 1] / gain;
        for (int i = out_end; i < 0; i++) {
          const int ri = j + i;
          if (ri < 0) { if (lag_len + ri >= 0) lag[lag_len + ri] *= g; }
          else if (ri < frame) res[ri] *= g;
        }
      }
      for (int si = j; si < j + n; si++) {
        float sum = res[si];
        for (int t = 0; t < 5; t++) {
          const int ri = si - pitch + 2 - t;
          const float v = ri < 0 ? (lag_len + ri >= 0 ? lag[lag_len + ri] : 0.0f) : (ri < frame ? res[ri] : 0.0f);
          sum += v * ((float)F.taps[s][t] / 128.0f);
        }
        res[si] = sum;
      }
    }
    const float gf = gain / 65536.0f;   // lpc_synthesis, :829-886
    for (int i = 0; i < n; i++) {
      const int si = j + i;
      float v = gf * res[si];
      for (int k = 0; k < d; k++) {
        const int li = si - k - 1;
        float cur = 0.0f;
        if (li >= 0) cur = lpc[li];
        else if (s == 0 && i < st.prev_lpc_len) {
          const int idx = st.prev_lpc_len - 1 + i - k;
          if (idx >= 0 && idx < st.prev_lpc_len) cur = st.prev_lpc[idx];
        }
        v += cur * (aq[k] / 4096.0f);
      }
      lpc[si] = v;
      out[si] = lite_clamp1(v);
    }
  }
  // the reference keeps the last d_lpc unclamped samples only when sample 319 was written by a frame that is not the
  // stream's first (:873-883): wideband frames, from the second one on
  if (st.have_decoded && frame == LITE_BUF) {
    for (int k = 0; k < d; k++) st.prev_lpc[k] = lpc[frame - d + k];
    st.prev_lpc_len = d;
  }
  for (int i = 0; i < LITE_FINAL; i++) st.final_out[i] = out[LITE_BUF - LITE_FINAL + i];
  st.have_decoded = 1;
}

// One packet of one stream (decode_internal, src/decoder.rs:150-217). Scratch: 3 x 320 floats.
MB_DEV int lite_decode_packet(LiteState& st, const uint8_t* pkt, uint32_t len, LiteFrame& F, float* res, float* lag, float* lpc) {
  if (len == 0) return LITE_TOO_SHORT;
  const int toc = pkt[0], config = toc >> 3;
  if (toc & 3) return LITE_FRAME_CODE;
  if (config > 11) return LITE_CONFIG_MODE;
  if ((config & 3) != 1) return LITE_FRAME_DURATION;
  if (toc & 4) return LITE_STEREO;
  RangeDec ec;
  ec.open(pkt + 1, len - 1);
  const int rc = lite_parse(ec, st, config >> 2, F);
  if (rc) return rc;
  lite_synthesis(st, F, res, lag, lpc);
  return LITE_OUT;
}

MB_DEV int32_t lite_floor_i32(float v) {   // src/bitdepth.rs:92-99
  const int32_t t = (int32_t)v;
  return (v >= 0.0f || (float)t == v) ? t : t - 1;
}

}  // namespace mb
