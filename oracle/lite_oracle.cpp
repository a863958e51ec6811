// ORACLE (test infrastructure only — never linked into the product library).
//
// CPU restatement of mousiki's LIGHTWEIGHT decoder: `decoder::Decoder::{decode,
// decode_float32, decode_int24}` (src/decoder.rs:142-262) over the float, pion-derived SILK decoder
// `silk::decoder::Decoder::decode` (src/silk/decoder.rs:971-1163): SILK-only, mono, one 20 ms frame per packet, no LBRR,
// float LTP / LPC synthesis, zero-order-hold x3 to 48 kHz (src/resample.rs:3-15), floor(x * 32767) to i16
// (src/bitdepth.rs:15-51).  It is NOT the full decoder's fixed-point SILK path (silk_oracle.cpp) and is not
// bit-compatible with it.  Tables: lite_tables_data.h (src/silk/icdf.rs, src/silk/codebook.rs, verbatim).
//
// Pinned on the reference's own vectors (tests/test_lite_oracle.py): src/decoder.rs:275-325 (first 12 samples, 1e-6),
// the Go fixtures src/silk/decoder.rs:2732-2780 (EXPECTED_DECODE_OUT_0 / _1, two consecutive frames), the unit vectors
// :2479-2700, :2925-2970 (gains, stage-2 residual, NLSF, stabilisation, A32_Q17, Aq12, pitch lags, LTP taps).
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "ec.h"
#include "lite_tables_data.h"
#include "silk_oracle.h"
namespace lite_silk {
#define SILK_TABLE static const
#include "silk_tables_data.h"   // silk_sign_iCDF (src/silk/tables_pulses_per_block.rs), the one libopus-layout table this path reads
}

namespace {
using namespace lite_tab;
using namespace orc;
using lite_silk::silk_sign_iCDF;

enum { SUBFR = 4, MAX_SHELL = 20, MAX_EXC = 320, MAX_RES_LAG = 290, FINAL_N = 306, BUF_N = 320 };
enum { SIG_INACTIVE = 0, SIG_UNVOICED = 1, SIG_VOICED = 2 };
enum { BW_NARROW = 0, BW_MEDIUM = 1, BW_WIDE = 2 };
// error codes of the C entry points (DecoderError / DecodeError variants, src/decoder.rs:15-21, src/silk/decoder.rs:182-188)
enum { LITE_OK = 0, LITE_TOO_SHORT = -1, LITE_FRAME_CODE = -2, LITE_CONFIG_MODE = -3, LITE_FRAME_DURATION = -4, LITE_STEREO = -5,
       LITE_OUT_TOO_SMALL = -6, LITE_LBRR = -7, LITE_PITCH = -8 };

// src/silk/range_decoder.rs:51-73 over EcDec (ec_decode / ec_dec_update)
static uint32_t sym(EcDec& ec, const Icdf& t) {
  const uint32_t s = ec.decode((uint32_t)t.total);
  int index = 0;
  while (index < t.n && (uint32_t)t.c[index] <= s) index++;
  if (index == t.n) index = 0;
  const uint32_t high = t.c[index], low = index > 0 ? t.c[index - 1] : 0;
  ec.update(low, high, (uint32_t)t.total);
  return (uint32_t)index;
}
static int ilog_isize(int64_t x) { return x <= 0 ? 0 : 64 - __builtin_clzll((unsigned long long)x); }   // src/math.rs:3-8
static int sign_i32(int v) { return v < 0 ? -1 : (v > 0 ? 1 : 0); }                                      // src/math.rs:10-16
static float clamp1(float v) { return v < -1.0f ? -1.0f : (v > 1.0f ? 1.0f : v); }                       // f32::clamp(-1, 1)

struct Frame {
  int signal, qoff, bw, d_lpc, w_q2, n_aq;
  float gain_q16[SUBFR], aq12[2][16];
  bool has_pitch;
  int lag_max, pitch_lags[SUBFR];
  int8_t b_q7[SUBFR][5];
  bool has_ltp;
  float ltp_scale_q14;
  int exc_len;
  int32_t exc[MAX_EXC];
};

struct LiteDecoder {   // src/silk/decoder.rs:84-94 + src/decoder.rs:137-140
  bool have_decoded = false, prev_voiced = false;
  int previous_log_gain = 0;
  float final_out[FINAL_N];
  int16_t n0_q15[16];
  int n0_len = 0;
  float prev_lpc[16];
  int prev_lpc_len = 0;
  float silk_buffer[BUF_N];
  Frame last;   // (not reference state: what the last frame parsed to, for the unit vectors)
  LiteDecoder() {
    memset(&last, 0, sizeof last);
    memset(final_out, 0, sizeof final_out); memset(n0_q15, 0, sizeof n0_q15);
    memset(prev_lpc, 0, sizeof prev_lpc); memset(silk_buffer, 0, sizeof silk_buffer);
  }
};

// src/silk/decoder.rs:269-362
static void decode_gains(LiteDecoder& d, EcDec& ec, int signal, float gain_q16[SUBFR]) {
  int log_gain = 0;
  for (int s = 0; s < SUBFR; s++) {
    if (s == 0) {
      int gi = (int)sym(ec, signal == SIG_INACTIVE ? INDEPENDENT_QUANTIZATION_GAIN_MSB_INACTIVE
                            : signal == SIG_VOICED ? INDEPENDENT_QUANTIZATION_GAIN_MSB_VOICED : INDEPENDENT_QUANTIZATION_GAIN_MSB_UNVOICED);
      gi = (gi << 3) | (int)sym(ec, INDEPENDENT_QUANTIZATION_GAIN_LSB);
      log_gain = d.have_decoded ? (gi > d.previous_log_gain - 16 ? gi : d.previous_log_gain - 16) : gi;
    } else {
      const int delta = (int)sym(ec, DELTA_QUANTIZATION_GAIN);
      int a = 2 * delta - 16, b = d.previous_log_gain + delta - 4;
      int v = a > b ? a : b;
      log_gain = v < 0 ? 0 : (v > 63 ? 63 : v);
    }
    d.previous_log_gain = log_gain;
    const int in_log_q7 = ((0x1D1C71 * log_gain) >> 16) + 2090;
    const int i = in_log_q7 >> 7, f = in_log_q7 & 127;
    gain_q16[s] = (float)((1 << i) + (((-174 * f * (128 - f)) >> 16) + f) * ((1 << i) >> 7));
  }
}

// src/silk/decoder.rs:404-517: stage-2 indices and the backwards-predicted residual
static int decode_lsf_stage2(EcDec& ec, int bw, uint32_t i1, int16_t res_q10[16]) {
  const bool wide = bw == BW_WIDE;
  const int d_lpc = wide ? 16 : 10;
  int8_t i2[16] = {0};
  for (int i = 0; i < d_lpc; i++) {
    const int cbk = wide ? NORMALIZED_LSF_STAGE_TWO_INDEX_WIDEBAND[i1][i] : NORMALIZED_LSF_STAGE_TWO_INDEX_NARROWBAND_OR_MEDIUMBAND[i1][i];
    i2[i] = (int8_t)((int)sym(ec, NORMALIZED_LSF_STAGE_2_INDEX[cbk]) - 4);
    if (i2[i] == -4) i2[i] = (int8_t)(i2[i] - (int)sym(ec, NORMALIZED_LSF_STAGE_2_INDEX_EXTENSION));
    else if (i2[i] == 4) i2[i] = (int8_t)(i2[i] + (int)sym(ec, NORMALIZED_LSF_STAGE_2_INDEX_EXTENSION));
  }
  const int64_t qstep = wide ? 9830 : 11796;
  memset(res_q10, 0, 16 * sizeof(int16_t));
  for (int k = d_lpc - 1; k >= 0; k--) {
    int64_t first = 0;
    if (k + 1 < d_lpc) {
      const int64_t pred_q8 = wide ? PREDICTION_WEIGHT_FOR_WIDEBAND_NORMALIZED_LSF[PREDICTION_WEIGHT_SELECTION_FOR_WIDEBAND_NORMALIZED_LSF[i1][k]][k]
                                   : PREDICTION_WEIGHT_FOR_NARROWBAND_AND_MEDIUMBAND_NORMALIZED_LSF[PREDICTION_WEIGHT_SELECTION_FOR_NARROWBAND_AND_MEDIUMBAND_NORMALIZED_LSF[i1][k]][k];
      first = ((int64_t)res_q10[k + 1] * pred_q8) >> 8;
    }
    const int64_t i2k = i2[k];
    const int64_t second = (((i2k << 10) - ((i2k > 0) - (i2k < 0)) * 102) * qstep) >> 16;
    res_q10[k] = (int16_t)(first + second);
  }
  return d_lpc;
}

// src/silk/decoder.rs:1717-1795
static void lsf_coefficients(int d_lpc, int16_t* nlsf_q15, int bw, const int16_t* res_q10, uint32_t i1) {
  const bool wide = bw == BW_WIDE;
  for (int k = 0; k < d_lpc; k++) {
    auto cb1 = [&](int kk) -> int64_t { return wide ? NORMALIZED_LSF_STAGE_ONE_WIDEBAND[i1][kk] : NORMALIZED_LSF_STAGE_ONE_NARROWBAND_OR_MEDIUMBAND[i1][kk]; };
    int64_t km1 = 0, kp1 = 256;
    if (k != 0) km1 = cb1(k - 1);
    if (k + 1 != d_lpc) kp1 = cb1(k + 1);
    const uint64_t w2_q18 = (uint64_t)(1024 / (cb1(k) - km1) + 1024 / (kp1 - cb1(k))) << 16;
    const int i = ilog_isize((int64_t)w2_q18);
    const int64_t f = (int64_t)((w2_q18 >> (i - 8)) & 127);
    int64_t y = (i & 1) ? 32768 : 46214;
    y >>= (32 - i) >> 1;
    const int16_t w_q9 = (int16_t)(y + ((213 * f * y) >> 16));
    const int32_t result = ((int32_t)cb1(k) << 7) + ((int32_t)res_q10[k] << 14) / (int32_t)w_q9;
    nlsf_q15[k] = (int16_t)(result < 0 ? 0 : (result > 32767 ? 32767 : result));
  }
}

// src/silk/decoder.rs:1555-1715
static void lsf_stabilize(int16_t* nlsf, int d_lpc, int bw) {
  const int32_t* nd = bw == BW_WIDE ? MINIMUM_SPACING_FOR_NORMALIZED_LSCOEFFICIENTS_WIDEBAND : MINIMUM_SPACING_FOR_NORMALIZED_LSCOEFFICIENTS_NARROWBAND_AND_MEDIUMBAND;
  for (int adj = 0; adj <= 19; adj++) {
    int64_t i = 0, i_value = INT64_MAX;
    for (int x = 0; x <= d_lpc; x++) {
      const int64_t prev = x != 0 ? nlsf[x - 1] : 0, cur = x == d_lpc ? 32768 : nlsf[x];
      const int64_t spacing = cur - prev - nd[x];
      if (spacing < i_value) { i = x; i_value = spacing; }
    }
    if (i_value >= 0) return;
    if (i == 0) { nlsf[0] = (int16_t)nd[0]; continue; }
    if (i == d_lpc) { nlsf[d_lpc - 1] = (int16_t)(32768 - nd[d_lpc]); continue; }
    int32_t min_center = nd[i] >> 1;
    for (int k = 0; k <= i - 1; k++) min_center += nd[k];
    int32_t max_center = 32768 - (nd[i] >> 1);
    for (int k = (int)i + 1; k <= d_lpc; k++) max_center -= nd[k];
    int32_t center = (int32_t)(((int64_t)nlsf[i - 1] + (int64_t)nlsf[i] + 1) >> 1);
    center = center < min_center ? min_center : (center > max_center ? max_center : center);   // i32::clamp
    nlsf[i - 1] = (int16_t)(center - (nd[i] >> 1));
    nlsf[i] = (int16_t)(nlsf[i - 1] + (int16_t)nd[i]);
  }
  for (int i = 1; i < d_lpc; i++) {
    int j = i;
    const int16_t cur = nlsf[i];
    while (j > 0 && nlsf[j - 1] > cur) { nlsf[j] = nlsf[j - 1]; j--; }
    nlsf[j] = cur;
  }
  for (int k = 0; k <= d_lpc - 1; k++) {
    const int16_t prev = k != 0 ? nlsf[k - 1] : 0;
    const int16_t lo = (int16_t)(prev + (int16_t)nd[k]);
    nlsf[k] = nlsf[k] > lo ? nlsf[k] : lo;
  }
  for (int k = d_lpc - 1; k >= 0; k--) {
    const int64_t next = k == d_lpc - 1 ? 32768 : nlsf[k + 1];
    const int16_t hi = (int16_t)(next - nd[k + 1]);
    nlsf[k] = nlsf[k] < hi ? nlsf[k] : hi;
  }
}

// src/silk/decoder.rs:1424-1485
static void lsf_to_a32_q17(const int16_t* n_q15, int d_lpc, int bw, int32_t* a32_q17) {
  int32_t c_q17[16] = {0};
  const int32_t* ordering = bw == BW_WIDE ? LSF_ORDERING_FOR_POLYNOMIAL_EVALUATION_WIDEBAND : LSF_ORDERING_FOR_POLYNOMIAL_EVALUATION_NARROWBAND_AND_MEDIUMBAND;
  for (int k = 0; k < d_lpc; k++) {
    const int i = n_q15[k] >> 8, f = n_q15[k] & 255;
    const int32_t cv = Q12_COSINE_TABLE_FOR_LSFCONVERION[i], cn = Q12_COSINE_TABLE_FOR_LSFCONVERION[i + 1];
    c_q17[ordering[k]] = (cv * 256 + (cn - cv) * f + 4) >> 3;
  }
  const int d2 = d_lpc / 2;
  int32_t p[10] = {0}, q[10] = {0};
  p[0] = 1 << 16; q[0] = 1 << 16; p[1] = -c_q17[0]; q[1] = -c_q17[1];
  for (int k = 1; k < d2; k++) {
    const int64_t ce = c_q17[2 * k], co = c_q17[2 * k + 1];
    p[k + 1] = p[k - 1] * 2 - (int32_t)((ce * (int64_t)p[k] + 32768) >> 16);
    q[k + 1] = q[k - 1] * 2 - (int32_t)((co * (int64_t)q[k] + 32768) >> 16);
    for (int j = k; j >= 2; j--) {
      p[j] += p[j - 2] - (int32_t)((ce * (int64_t)p[j - 1] + 32768) >> 16);
      q[j] += q[j - 2] - (int32_t)((co * (int64_t)q[j - 1] + 32768) >> 16);
    }
    p[1] -= c_q17[2 * k];
    q[1] -= c_q17[2 * k + 1];
  }
  for (int k = 0; k < d2; k++) {
    const int32_t dq = q[k + 1] - q[k], sp = p[k + 1] + p[k];
    a32_q17[k] = -dq - sp;
    a32_q17[d_lpc - k - 1] = dq - sp;
  }
}

// src/silk/decoder.rs:1487-1533
static void limit_a32_range(int32_t* a, int len) {
  int round = 0;
  while (round < 10) {
    uint32_t maxabs = 0;
    int maxidx = 0;
    for (int i = 0; i < len; i++) {
      const uint32_t v = a[i] < 0 ? (uint32_t)(-(int64_t)a[i]) : (uint32_t)a[i];
      if (v > maxabs) { maxabs = v; maxidx = i; }
    }
    uint32_t maxabs_q12 = (maxabs + 16) >> 5;
    if (maxabs_q12 > 163838) maxabs_q12 = 163838;
    if (maxabs_q12 > 32767) {
      uint32_t sc[16] = {0};
      const uint32_t num = (maxabs_q12 - 32767) << 14;
      uint32_t den = (maxabs_q12 * (uint32_t)(maxidx + 1)) >> 2;
      if (den < 1) den = 1;
      sc[0] = 65470 - num / den;
      for (int k = 0; k < len; k++) {
        a[k] = (int32_t)(((int64_t)a[k] * (int64_t)sc[k]) >> 16);
        if (k + 1 < len) sc[k + 1] = (sc[0] * sc[k] + 32768) >> 16;
      }
    } else {
      break;
    }
    round++;
  }
  if (round == 9)
    for (int i = 0; i < len; i++) {
      int32_t q12 = (a[i] + 16) >> 5;
      q12 = q12 < -32768 ? -32768 : (q12 > 32767 ? 32767 : q12);
      a[i] = q12 << 5;
    }
}

// src/silk/decoder.rs:547-561 (generate_a_q12) with :1535-1553 (the "prediction gain" step only converts to Q12 floats)
static void make_aq12(const int16_t* n_q15, int d_lpc, int bw, float* aq12) {
  int32_t a32[16];
  lsf_to_a32_q17(n_q15, d_lpc, bw, a32);
  limit_a32_range(a32, d_lpc);
  for (int k = 0; k < d_lpc; k++) aq12[k] = (float)((a32[k] + 16) >> 5);
}

// src/silk/decoder.rs:1191-1422: shell blocks, pulse positions, LSBs, signs, then the Q23 excitation with the LCG
static void decode_excitation(EcDec& ec, Frame& F, uint32_t seed, int shell_blocks, uint32_t rate_level) {
  uint8_t pulse_counts[MAX_SHELL] = {0}, lsb_counts[MAX_SHELL] = {0};
  uint32_t ri = rate_level > 10 ? 10 : rate_level;
  for (int b = 0; b < shell_blocks; b++) {
    uint8_t pc = (uint8_t)sym(ec, PULSE_COUNT[ri]);
    if (pc == 17) {
      uint8_t lsb = 0;
      while (pc == 17 && lsb < 10) { pc = (uint8_t)sym(ec, PULSE_COUNT[9]); lsb++; }
      lsb_counts[b] = lsb;
      if (lsb == 10) pc = (uint8_t)sym(ec, PULSE_COUNT[10]);
    }
    pulse_counts[b] = pc;
  }
  const int len = shell_blocks * 16;
  static const int offs[3][2] = {{25, 60}, {25, 60}, {8, 25}};
  const int offset_q23 = offs[F.signal][F.qoff];
  int32_t e_raw[MAX_EXC];
  memset(e_raw, 0, sizeof e_raw);
  auto part = [&](const Icdf* ctx, uint8_t block, uint8_t halves[2]) {   // :1405-1422
    if (block == 0) { halves[0] = halves[1] = 0; return; }
    int index = (int)block - 1;
    if (index > 15) index = 15;
    const uint8_t left = (uint8_t)sym(ec, ctx[index]);
    halves[0] = left;
    halves[1] = block > left ? (uint8_t)(block - left) : 0;   // saturating_sub
  };
  for (int b = 0; b < shell_blocks; b++) {
    const uint8_t pulses = pulse_counts[b];
    if (pulses == 0) continue;
    int e = b * 16;
    uint8_t p16[2], p8[2], p4[2], p2[2];
    part(PULSE_COUNT_SPLIT16_SAMPLE_PARTITIONS, pulses, p16);
    for (int a = 0; a < 2; a++) {
      part(PULSE_COUNT_SPLIT8_SAMPLE_PARTITIONS, p16[a], p8);
      for (int c = 0; c < 2; c++) {
        part(PULSE_COUNT_SPLIT4_SAMPLE_PARTITIONS, p8[c], p4);
        for (int g = 0; g < 2; g++) {
          part(PULSE_COUNT_SPLIT2_SAMPLE_PARTITIONS, p4[g], p2);
          e_raw[e++] = p2[0];
          e_raw[e++] = p2[1];
        }
      }
    }
  }
  for (int i = 0; i < len; i++)
    for (int l = 0; l < lsb_counts[i / 16]; l++) e_raw[i] = (e_raw[i] << 1) | (int32_t)sym(ec, EXCITATION_LSB);
  // signs: src/silk/code_signs.rs:84-125 with sum_pulses = pulse_count | lsb_count << 5 (:133-141)
  {
    const uint8_t* icdf_ptr = silk_sign_iCDF + 7 * (F.qoff + (F.signal << 1));
    const int num_blocks = (len + 8) >> 4;
    int pi = 0;
    for (int b = 0; b < num_blocks && b < shell_blocks; b++) {
      const int total = (int)pulse_counts[b] | ((int)lsb_counts[b] << 5);
      if (total > 0) {
        int ti = total & 0x1F;
        if (ti > 6) ti = 6;
        const uint8_t icdf[2] = {icdf_ptr[ti], 0};
        const int end = pi + 16 < len ? pi + 16 : len;
        for (int i = pi; i < end; i++)
          if ((int16_t)e_raw[i] > 0) e_raw[i] = (int16_t)e_raw[i] * (((int16_t)ec.icdf(icdf, 8) << 1) - 1);
      }
      pi += 16;
    }
  }
  F.exc_len = len;
  for (int i = 0; i < len; i++) {
    const int32_t raw = e_raw[i];
    int32_t value = (raw << 8) - sign_i32(raw) * 20 + offset_q23;
    seed = seed * 196314165u + 907633515u;
    if (seed & 0x80000000u) value = -value;
    seed += (uint32_t)raw;
    F.exc[i] = value;
  }
}

// src/silk/decoder.rs:689-827
static void ltp_synthesis(LiteDecoder& d, float* out, const Frame& F, int n, int j, int s, const float* a_q12, int n_a, float* res, int res_len,
                          float* res_lag, int res_lag_len) {
  float ltp_scale = F.ltp_scale_q14;
  int out_end;
  if (s < 2 || F.w_q2 == 4) out_end = -s * n;
  else { ltp_scale = 16384.0f; out_end = -(s - 2) * n; }
  const int pitch = F.pitch_lags[s];
  for (int i = -pitch - 2; i < out_end; i++) {
    const int index = i + j;
    if (index >= res_len) continue;
    bool to_lag = false;
    float res_val;
    int res_index;
    if (index >= 0) {
      if (index >= BUF_N) continue;
      res_val = out[index]; res_index = index;
    } else {
      const int lag_index = res_lag_len + index;
      if (lag_index < 0 || lag_index >= res_lag_len) continue;
      const int fi = FINAL_N + index;
      res_val = (fi < 0 || fi >= FINAL_N) ? 0.0f : d.final_out[fi];
      to_lag = true; res_index = lag_index;
    }
    for (int k = 0; k < n_a && k < F.d_lpc; k++) {
      const int oi = index - k - 1;
      float ov;
      if (oi >= 0) ov = out[oi];
      else { const int fi = FINAL_N + oi; ov = (fi < 0 || fi >= FINAL_N) ? 0.0f : d.final_out[fi]; }
      res_val -= ov * (a_q12[k] / 4096.0f);
    }
    res_val = clamp1(res_val);
    const float gain = F.gain_q16[s];
    if (gain != 0.0f) {
      res_val *= (4.0f * ltp_scale) / gain;
      if (!to_lag) res[res_index] = res_val;
      else if (res_index < res_lag_len) res_lag[res_index] = res_val;
    }
  }
  if (s > 0) {
    const float cur = F.gain_q16[s];
    if (cur != 0.0f) {
      const float scaled = F.gain_q16[s - 1] / cur;
      for (int idx = out_end; idx < 0; idx++) {
        const int ri = j + idx;
        if (ri < 0) { const int li = res_lag_len + ri; if (li >= 0 && li < res_lag_len) res_lag[li] *= scaled; }
        else if (ri < res_len) res[ri] *= scaled;
      }
    }
  }
  for (int si = j; si < j + n; si++) {
    float sum = res[si];
    for (int t = 0; t < 5; t++) {
      const int ri = si - pitch + 2 - t;
      float v;
      if (ri < 0) { const int li = res_lag_len + ri; v = (li >= 0 && li < res_lag_len) ? res_lag[li] : 0.0f; }
      else v = ri < res_len ? res[ri] : 0.0f;
      sum += v * ((float)F.b_q7[s][t] / 128.0f);
    }
    res[si] = sum;
  }
}

// src/silk/decoder.rs:829-886
static void lpc_synthesis(LiteDecoder& d, float* out, const Frame& F, int n, int s, const float* a_q12, int n_a, const float* res, int res_len,
                          float* lpc, int frame_samples) {
  const int j = n * s;
  const float gain_factor = F.gain_q16[s] / 65536.0f;
  for (int i = 0; i < n; i++) {
    const int si = j + i;
    if (si >= res_len || si >= frame_samples) break;
    float v = gain_factor * res[si];
    for (int k = 0; k < n_a && k < F.d_lpc; k++) {
      const int li = si - k - 1;
      float cur;
      if (li >= 0) cur = lpc[li];
      else if (s == 0 && i < d.prev_lpc_len) {
        const int idx = d.prev_lpc_len - 1 + i - k;
        cur = (idx >= 0 && idx < d.prev_lpc_len) ? d.prev_lpc[idx] : 0.0f;
      } else cur = 0.0f;
      v += cur * (a_q12[k] / 4096.0f);
    }
    lpc[si] = v;
    if (si < BUF_N) out[si] = clamp1(v);
    if (s == SUBFR - 1 && d.have_decoded && si == BUF_N - 1 && F.d_lpc <= frame_samples) {
      memcpy(d.prev_lpc, lpc + frame_samples - F.d_lpc, F.d_lpc * sizeof(float));
      d.prev_lpc_len = F.d_lpc;
    }
  }
}

// src/silk/decoder.rs:971-1163 (+ :889-969 silk_frame_reconstruction); out = the decoder's 320-sample silk_buffer
static int silk_lite_decode(LiteDecoder& d, const uint8_t* in, int len, bool stereo, uint32_t ns, int bw) {
  float* out = d.silk_buffer;
  const int n = bw == BW_NARROW ? 40 : bw == BW_MEDIUM ? 60 : 80;
  if (ns != 20000000u) return LITE_FRAME_DURATION;
  if (stereo) return LITE_STEREO;
  EcDec ec;
  ec.init(in, (uint32_t)len);
  const bool vad = ec.bit_logp(1) == 1, lbrr = ec.bit_logp(1) == 1;
  if (lbrr) return LITE_LBRR;
  Frame F;
  memset(&F, 0, sizeof F);
  F.bw = bw;
  {   // :225-266
    const uint32_t t = sym(ec, vad ? FRAME_TYPE_VAD_ACTIVE : FRAME_TYPE_VAD_INACTIVE);
    if (!vad) { F.signal = SIG_INACTIVE; F.qoff = t == 0 ? 0 : 1; }
    else { F.signal = t < 2 ? SIG_UNVOICED : SIG_VOICED; F.qoff = (int)(t & 1); }
  }
  decode_gains(d, ec, F.signal, F.gain_q16);
  const bool voiced = F.signal == SIG_VOICED;
  const uint32_t i1 = sym(ec, bw == BW_WIDE ? (voiced ? NORMALIZED_LSF_STAGE_1_INDEX_WIDEBAND_VOICED : NORMALIZED_LSF_STAGE_1_INDEX_WIDEBAND_UNVOICED)
                                          : (voiced ? NORMALIZED_LSF_STAGE_1_INDEX_NARROWBAND_OR_MEDIUMBAND_VOICED
                                                    : NORMALIZED_LSF_STAGE_1_INDEX_NARROWBAND_OR_MEDIUMBAND_UNVOICED));
  int16_t res_q10[16], nlsf[16] = {0};
  F.d_lpc = decode_lsf_stage2(ec, bw, i1, res_q10);
  lsf_coefficients(F.d_lpc, nlsf, bw, res_q10, i1);
  lsf_stabilize(nlsf, F.d_lpc, bw);
  F.w_q2 = (int)sym(ec, NORMALIZED_LSF_INTERPOLATION_INDEX);   // :520-545
  F.n_aq = 0;
  if (!(F.w_q2 == 4 || !d.have_decoded)) {
    int16_t n1[16];
    for (int k = 0; k < F.d_lpc; k++) {
      const int32_t prev = d.n0_q15[k], diff = (int32_t)nlsf[k] - prev;
      n1[k] = (int16_t)(prev + ((F.w_q2 * diff) >> 2));
    }
    make_aq12(n1, F.d_lpc, bw, F.aq12[F.n_aq++]);
  }
  make_aq12(nlsf, F.d_lpc, bw, F.aq12[F.n_aq++]);
  if (voiced) {   // :563-621
    static const Icdf* low[3] = {&PRIMARY_PITCH_LAG_LOW_PART_NARROWBAND, &PRIMARY_PITCH_LAG_LOW_PART_MEDIUMBAND, &PRIMARY_PITCH_LAG_LOW_PART_WIDEBAND};
    static const int scale[3] = {4, 6, 8}, lmin[3] = {16, 24, 32}, lmax[3] = {144, 216, 288}, khz[3] = {8, 12, 16};
    const uint32_t hi = sym(ec, PRIMARY_PITCH_LAG_HIGH_PART), lo = sym(ec, *low[bw]);
    const uint32_t lag = hi * scale[bw] + lo + lmin[bw];
    const int contour = (int8_t)sym(ec, bw == BW_NARROW ? SUBFRAME_PITCH_CONTOUR_NARROWBAND20_MS : SUBFRAME_PITCH_CONTOUR_MEDIUMBAND_OR_WIDEBAND20_MS);
    int32_t lags[SILK_MAX_SUBFR] = {0};
    silk_decode_pitch((int16_t)(lag - lmin[bw]), contour, lags, khz[bw], SUBFR);
    for (int k = 0; k < SUBFR; k++) F.pitch_lags[k] = lags[k];
    F.has_pitch = true; F.lag_max = lmax[bw];
    // :623-658
    const uint32_t per = sym(ec, PERIODICITY_INDEX);
    const Icdf& ficdf = per == 0 ? LTP_FILTER_INDEX0 : per == 1 ? LTP_FILTER_INDEX1 : LTP_FILTER_INDEX2;
    for (int s = 0; s < SUBFR; s++) {
      const uint32_t fi = sym(ec, ficdf);
      for (int t = 0; t < 5; t++)
        F.b_q7[s][t] = (int8_t)(per == 0 ? CODEBOOK_LTP_FILTER_PERIODICITY_INDEX_0[fi][t] : per == 1 ? CODEBOOK_LTP_FILTER_PERIODICITY_INDEX_1[fi][t]
                                                                                                    : CODEBOOK_LTP_FILTER_PERIODICITY_INDEX_2[fi][t]);
    }
    F.has_ltp = true;
    static const float sc[3] = {15565.0f, 12288.0f, 8192.0f};
    F.ltp_scale_q14 = sc[sym(ec, LTP_SCALING_PARAMETER)];
  } else {
    F.ltp_scale_q14 = 15565.0f;
  }
  const uint32_t seed = sym(ec, LINEAR_CONGRUENTIAL_GENERATOR_SEED);
  const int shell_blocks = bw == BW_NARROW ? 10 : bw == BW_MEDIUM ? 15 : 20;
  const uint32_t rate_level = sym(ec, voiced ? RATE_LEVEL_VOICED : RATE_LEVEL_UNVOICED);
  decode_excitation(ec, F, seed, shell_blocks, rate_level);
  // silk_frame_reconstruction, :889-969
  {
    const int frame_samples = n * SUBFR;
    float lpc[320] = {0}, res[MAX_EXC] = {0}, res_lag[MAX_RES_LAG] = {0};
    const int res_len = F.exc_len < MAX_EXC ? F.exc_len : MAX_EXC;
    for (int i = 0; i < res_len; i++) res[i] = (float)F.exc[i] * (1.0f / 8388608.0f);
    int res_lag_len = F.lag_max + 2;
    if (res_lag_len > MAX_RES_LAG) res_lag_len = MAX_RES_LAG;
    for (int s = 0; s < SUBFR; s++) {
      const int aq_index = (s > 1 && F.n_aq > 1) ? 1 : 0;
      const float* aq = F.aq12[aq_index < F.n_aq - 1 ? aq_index : F.n_aq - 1];
      const int j = n * s;
      if (voiced && F.has_ltp && F.has_pitch) ltp_synthesis(d, out, F, n, j, s, aq, F.d_lpc, res, res_len, res_lag, res_lag_len);
      lpc_synthesis(d, out, F, n, s, aq, F.d_lpc, res, res_len, lpc, frame_samples);
    }
  }
  d.last = F;
  d.prev_voiced = voiced;
  memcpy(d.n0_q15, nlsf, F.d_lpc * sizeof(int16_t));
  d.n0_len = F.d_lpc;
  memcpy(d.final_out, out + (BUF_N - FINAL_N), FINAL_N * sizeof(float));
  d.have_decoded = true;
  return LITE_OK;
}

// src/decoder.rs:150-217 (decode_internal)
static int lite_decode_internal(LiteDecoder& d, const uint8_t* in, int len) {
  if (len <= 0) return LITE_TOO_SHORT;
  const int toc = in[0], config = toc >> 3;
  if ((toc & 3) != 0) return LITE_FRAME_CODE;
  if (config > 11) return LITE_CONFIG_MODE;   // hybrid / CELT
  static const uint32_t ns_tab[4] = {10000000u, 20000000u, 40000000u, 60000000u};
  return silk_lite_decode(d, in + 1, len - 1, (toc & 4) != 0, ns_tab[config & 3], config >> 2);
}
static int32_t floor_to_i32(float v) {   // src/bitdepth.rs:92-99
  const int32_t t = (int32_t)v;
  return (v >= 0.0f || (float)t == v) ? t : t - 1;
}
}  // namespace

extern "C" {
void* orl_create() { return new LiteDecoder(); }
void orl_destroy(void* h) { delete (LiteDecoder*)h; }
// decoder::Decoder::decode_float32 (src/decoder.rs:235-248): 960 floats
int orl_decode_f32(void* h, const uint8_t* pkt, int len, float* out) {
  LiteDecoder& d = *(LiteDecoder*)h;
  const int rc = lite_decode_internal(d, pkt, len);
  if (rc != LITE_OK) return rc;
  for (int i = 0; i < BUF_N; i++) out[3 * i] = out[3 * i + 1] = out[3 * i + 2] = d.silk_buffer[i];
  return 960;
}
// decoder::Decoder::decode (src/decoder.rs:220-233): 960 little-endian i16
int orl_decode_i16(void* h, const uint8_t* pkt, int len, int16_t* out) {
  LiteDecoder& d = *(LiteDecoder*)h;
  const int rc = lite_decode_internal(d, pkt, len);
  if (rc != LITE_OK) return rc;
  for (int i = 0; i < BUF_N; i++) {
    const int16_t v = (int16_t)floor_to_i32(d.silk_buffer[i] * 32767.0f);
    out[3 * i] = out[3 * i + 1] = out[3 * i + 2] = v;
  }
  return 960;
}
// decoder::Decoder::decode_int24 (src/decoder.rs:252-260): rint(x * 2^23)
int orl_decode_i24(void* h, const uint8_t* pkt, int len, int32_t* out) {
  LiteDecoder& d = *(LiteDecoder*)h;
  const int rc = lite_decode_internal(d, pkt, len);
  if (rc != LITE_OK) return rc;
  for (int i = 0; i < BUF_N; i++) {
    const int32_t v = (int32_t)rint((double)d.silk_buffer[i] * 8388608.0);
    out[3 * i] = out[3 * i + 1] = out[3 * i + 2] = v;
  }
  return 960;
}
void orl_silk_buffer(void* h, float* out320) { memcpy(out320, ((LiteDecoder*)h)->silk_buffer, BUF_N * sizeof(float)); }
// the SILK layer alone (silk::decoder::Decoder::decode), for the Go fixtures of src/silk/decoder.rs:2732-2780
int orl_silk_decode(void* h, const uint8_t* frame, int len, int stereo, uint32_t ns, int bw, float* out320) {
  LiteDecoder& d = *(LiteDecoder*)h;
  const int rc = silk_lite_decode(d, frame, len, stereo != 0, ns, bw);
  if (rc == LITE_OK) memcpy(out320, d.silk_buffer, BUF_N * sizeof(float));
  return rc;
}
// unit hooks for the reference's unit vectors
void orl_lsf_coefficients(int d_lpc, int bw, const int16_t* res_q10, uint32_t i1, int16_t* nlsf_q15) { lsf_coefficients(d_lpc, nlsf_q15, bw, res_q10, i1); }
void orl_lsf_stabilize(int16_t* nlsf_q15, int d_lpc, int bw) { lsf_stabilize(nlsf_q15, d_lpc, bw); }
void orl_lsf_to_lpc(const int16_t* nlsf_q15, int d_lpc, int bw, int32_t* a32_q17, float* aq12) {
  lsf_to_a32_q17(nlsf_q15, d_lpc, bw, a32_q17);
  int32_t t[16];
  memcpy(t, a32_q17, d_lpc * sizeof(int32_t));
  limit_a32_range(t, d_lpc);
  for (int k = 0; k < d_lpc; k++) aq12[k] = (float)((t[k] + 16) >> 5);
}
// what the last decoded frame parsed to: signal, qoff, w_q2, d_lpc, lag_max, pitch_lags[4], taps[20], exc_len, exc[320];
// gains[4], ltp_scale, then the Q12 sets (n_aq x 16)
void orl_last_frame(void* h, int32_t* ints /*30 + 320*/, float* floats /*4 + 1 + 1 + 32*/) {
  const Frame& F = ((LiteDecoder*)h)->last;
  ints[0] = F.signal; ints[1] = F.qoff; ints[2] = F.w_q2; ints[3] = F.d_lpc; ints[4] = F.lag_max;
  for (int k = 0; k < 4; k++) ints[5 + k] = F.pitch_lags[k];
  for (int k = 0; k < 20; k++) ints[9 + k] = F.b_q7[k / 5][k % 5];
  ints[29] = F.exc_len;
  for (int k = 0; k < MAX_EXC; k++) ints[30 + k] = F.exc[k];
  for (int k = 0; k < 4; k++) floats[k] = F.gain_q16[k];
  floats[4] = F.ltp_scale_q14; floats[5] = (float)F.n_aq;
  for (int k = 0; k < 32; k++) floats[6 + k] = F.aq12[k / 16][k % 16];
}
}
