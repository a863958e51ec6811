// ORACLE (test infrastructure only — never linked into the product library).
// CPU restatement of mousiki's fixed-point SILK decoder (normal-decode path). Every function cites the reference
// file it follows (paths relative to /root/reference/src/silk). Integer arithmetic is restated with the exact
// widths/wrapping the Rust code uses (release build: `+`/`-`/`<<` on i32 wrap).
#include "silk_oracle.h"

#include <string.h>

#define SILK_TABLE static const
#include "silk_tables_data.h"

namespace orc {

// ---- fixed-point helpers (macros.rs / per-file private helpers) ----
static inline int32_t w_add(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
static inline int32_t w_sub(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }
static inline int32_t w_mul(int32_t a, int32_t b) { return (int32_t)((uint32_t)a * (uint32_t)b); }
static inline int32_t w_shl(int32_t a, int s) { return (int32_t)((uint32_t)a << (s & 31)); }
static inline int32_t smulwb(int32_t a, int32_t b) { return (int32_t)(((int64_t)a * (int16_t)b) >> 16); }
static inline int32_t smulww(int32_t a, int32_t b) { return (int32_t)(((int64_t)a * b) >> 16); }
static inline int32_t smlawb(int32_t acc, int32_t a, int32_t b) { return w_add(acc, smulwb(a, b)); }
static inline int32_t smulbb(int32_t a, int32_t b) { return (int32_t)(int16_t)a * (int32_t)(int16_t)b; }
static inline int32_t smlabb(int32_t acc, int32_t a, int32_t b) { return w_add(acc, smulbb(a, b)); }
static inline int32_t smmul(int32_t a, int32_t b) { return (int32_t)(((int64_t)a * b) >> 32); }
static inline int32_t rshift_round(int32_t v, int s) { return s == 1 ? (v >> 1) + (v & 1) : ((v >> (s - 1)) + 1) >> 1; }
static inline int64_t rshift_round64(int64_t v, int s) { return s == 1 ? (v >> 1) + (v & 1) : ((v >> (s - 1)) + 1) >> 1; }
static inline int16_t sat16(int32_t v) { return (int16_t)(v > 32767 ? 32767 : (v < -32768 ? -32768 : v)); }
static inline int32_t sat32_64(int64_t v) { return (int32_t)(v > INT32_MAX ? INT32_MAX : (v < INT32_MIN ? INT32_MIN : v)); }
static inline int32_t lshift_sat32(int32_t v, int s) { return sat32_64((int64_t)v << s); }
static inline int clz32(int32_t v) { return v == 0 ? 32 : __builtin_clz((uint32_t)v); }
static inline int32_t iabs(int32_t v) { return v < 0 ? (int32_t)(0u - (uint32_t)v) : v; }

// lpc_inv_pred_gain.rs:128-150 (silk_INVERSE32_varQ)
static int32_t inverse32_varq(int32_t b32, int qres) {
  if (b32 == 0 || qres <= 0) return 0;
  int32_t ab = iabs(b32); if (ab < 1) ab = 1;
  int b_headroom = clz32(ab) - 1;
  int32_t b32_nrm = w_shl(b32, b_headroom);
  int32_t b32_inv = (INT32_MAX >> 2) / (b32_nrm >> 16);
  int32_t result = w_shl(b32_inv, 16);
  int32_t err_q32 = w_shl(w_sub(1 << 29, (int32_t)(((int64_t)b32_nrm * b32_inv) >> 16)), 3);
  result = w_add(result, (int32_t)(((int64_t)err_q32 * b32_inv) >> 16));
  int lshift = 61 - b_headroom - qres;
  if (lshift <= 0) return lshift_sat32(result, -lshift);
  return lshift < 32 ? result >> lshift : 0;
}

// stereo_find_predictor.rs:134-166 (silk_DIV32_varQ)
static int32_t div32_varq(int32_t a32, int32_t b32, int qres) {
  int32_t abs_a = a32 == INT32_MIN ? INT32_MAX : iabs(a32), abs_b = b32 == INT32_MIN ? INT32_MAX : iabs(b32);
  int a_headroom = clz32(abs_a) - 1, b_headroom = clz32(abs_b) - 1;
  int32_t a_norm = w_shl(a32, a_headroom), b_norm = w_shl(b32, b_headroom);
  int32_t b_inv = (INT32_MAX >> 2) / (b_norm >> 16);
  int32_t result = smulwb(a_norm, b_inv);
  a_norm = w_sub(a_norm, w_shl(smmul(b_norm, result), 3));
  result = smlawb(result, a_norm, b_inv);
  int lshift = 29 + a_headroom - b_headroom - qres;
  if (lshift < 0) return lshift_sat32(result, -lshift);
  return lshift < 32 ? result >> lshift : 0;
}

// log2lin.rs:14-35
int32_t silk_log2lin(int32_t in_log_q7) {
  if (in_log_q7 < 0) return 0;
  if (in_log_q7 >= 3967) return INT32_MAX;
  int32_t out = 1 << (in_log_q7 >> 7);
  int32_t frac = in_log_q7 & 0x7f;
  int32_t corr = frac + (int32_t)(((int64_t)(frac * (128 - frac)) * -174) >> 16);
  if (in_log_q7 < 2048) out += (int32_t)(((int64_t)out * corr) >> 7);
  else out += (int32_t)((int64_t)(out >> 7) * corr);
  return out;
}

// ---- codebooks ----
static const SilkNlsfCb CB_WB = {32, 16, 9830, silk_NLSF_CB1_WB_Q8, silk_NLSF_CB1_WB_Wght_Q9, silk_NLSF_CB1_iCDF_WB, silk_NLSF_PRED_WB_Q8,
                                 silk_NLSF_CB2_SELECT_WB, silk_NLSF_CB2_iCDF_WB, silk_NLSF_DELTA_MIN_WB_Q15};
static const SilkNlsfCb CB_NB_MB = {32, 10, 11796, silk_NLSF_CB1_NB_MB_Q8, silk_NLSF_CB1_NB_MB_Wght_Q9, silk_NLSF_CB1_iCDF_NB_MB,
                                    silk_NLSF_PRED_NB_MB_Q8, silk_NLSF_CB2_SELECT_NB_MB, silk_NLSF_CB2_iCDF_NB_MB,
                                    silk_NLSF_DELTA_MIN_NB_MB_Q15};
const SilkNlsfCb* silk_cb_wb() { return &CB_WB; }
const SilkNlsfCb* silk_cb_nb_mb() { return &CB_NB_MB; }

// nlsf_unpack.rs:19-52
static void nlsf_unpack(int16_t* ec_ix, uint8_t* pred_q8, const SilkNlsfCb* cb, int cb1_index) {
  const uint8_t* sel = cb->ec_sel + cb1_index * cb->order / 2;
  for (int i = 0; i < cb->order; i += 2) {
    uint8_t e = *sel++;
    ec_ix[i] = (int16_t)(((e >> 1) & 7) * 9);
    pred_q8[i] = cb->pred_q8[i + (e & 1) * (cb->order - 1)];
    ec_ix[i + 1] = (int16_t)(((e >> 5) & 7) * 9);
    pred_q8[i + 1] = cb->pred_q8[i + ((e >> 4) & 1) * (cb->order - 1) + 1];
  }
}

// nlsf_stabilize.rs:20-105
void silk_nlsf_stabilize(int16_t* nlsf, const int16_t* dmin, int L) {
  for (int loops = 0; loops < 20; loops++) {
    int32_t min_diff = nlsf[0] - dmin[0];
    int I = 0;
    for (int i = 1; i < L; i++) {
      int32_t d = nlsf[i] - (nlsf[i - 1] + dmin[i]);
      if (d < min_diff) { min_diff = d; I = i; }
    }
    int32_t d = (1 << 15) - (nlsf[L - 1] + dmin[L]);
    if (d < min_diff) { min_diff = d; I = L; }
    if (min_diff >= 0) return;
    if (I == 0) nlsf[0] = dmin[0];
    else if (I == L) nlsf[L - 1] = sat16((1 << 15) - dmin[L]);
    else {
      int32_t min_c = 0, max_c = 1 << 15;
      for (int k = 0; k < I; k++) min_c += dmin[k];
      min_c += dmin[I] >> 1;
      for (int k = L; k > I; k--) max_c -= dmin[k];
      max_c -= dmin[I] >> 1;
      int32_t c = ((int32_t)nlsf[I - 1] + nlsf[I] + 1) >> 1;
      c = c < min_c ? min_c : (c > max_c ? max_c : c);  // (Rust clamp; min_c <= max_c for the shipped tables)
      nlsf[I - 1] = sat16(c - (dmin[I] >> 1));
      nlsf[I] = sat16(nlsf[I - 1] + dmin[I]);
    }
  }
  for (int i = 1; i < L; i++) {  // insertion sort (sort.rs)
    int16_t v = nlsf[i]; int j = i - 1;
    while (j >= 0 && v < nlsf[j]) { nlsf[j + 1] = nlsf[j]; j--; }
    nlsf[j + 1] = v;
  }
  if (nlsf[0] < dmin[0]) nlsf[0] = dmin[0];
  for (int i = 1; i < L; i++) { int16_t m = sat16((int32_t)nlsf[i - 1] + dmin[i]); if (nlsf[i] < m) nlsf[i] = m; }
  { int16_t m = sat16((1 << 15) - dmin[L]); if (nlsf[L - 1] > m) nlsf[L - 1] = m; }
  for (int i = L - 2; i >= 0; i--) { int16_t m = sat16((int32_t)nlsf[i + 1] - dmin[i + 1]); if (nlsf[i] > m) nlsf[i] = m; }
}

// nlsf_decode.rs:27-107
void silk_nlsf_decode(int16_t* nlsf_q15, const int8_t* indices, const SilkNlsfCb* cb, uint32_t quirks) {
  int order = cb->order;
  int16_t ec_ix[SILK_MAX_ORDER]; uint8_t pred_q8[SILK_MAX_ORDER]; int16_t res_q10[SILK_MAX_ORDER];
  nlsf_unpack(ec_ix, pred_q8, cb, indices[0]);
  int32_t out_q10 = 0;
  for (int i = order - 1; i >= 0; i--) {
    int32_t prod = (quirks & QUIRK_S1_NLSF_SMULBB16) ? (int32_t)(int16_t)((int16_t)out_q10 * (int16_t)pred_q8[i])
                                                     : smulbb(out_q10, pred_q8[i]);
    int32_t pred_q10 = prod >> 8;
    int32_t q = w_shl(indices[i + 1], 10);
    if (q > 0) q -= 102; else if (q < 0) q += 102;
    out_q10 = smlawb(pred_q10, q, cb->quant_step_size_q16);
    res_q10[i] = sat16(out_q10);
  }
  const uint8_t* cb1 = cb->cb1_nlsf_q8 + indices[0] * order;
  const int16_t* w = cb->cb1_wght_q9 + indices[0] * order;
  for (int i = 0; i < order; i++) {
    int32_t v = w_add(w_shl(res_q10[i], 14) / (int32_t)w[i], w_shl((int32_t)cb1[i], 7));
    nlsf_q15[i] = (int16_t)(v < 0 ? 0 : (v > 32767 ? 32767 : v));
  }
  silk_nlsf_stabilize(nlsf_q15, cb->delta_min_q15, order);
}

// bwexpander_32.rs:11-25
static void bwexpander_32(int32_t* ar, int d, int32_t chirp_q16) {
  int32_t cm1 = w_sub(chirp_q16, 1 << 16);
  for (int i = 0; i < d - 1; i++) {
    ar[i] = smulww(chirp_q16, ar[i]);
    chirp_q16 = w_add(chirp_q16, (int32_t)rshift_round64((int64_t)chirp_q16 * cm1, 16));
  }
  ar[d - 1] = smulww(chirp_q16, ar[d - 1]);
}
// bwexpander.rs:11-23
static void bwexpander(int16_t* ar, int d, int32_t chirp_q16) {
  int32_t cm1 = chirp_q16 - (1 << 16);
  for (int i = 0; i < d - 1; i++) {
    ar[i] = (int16_t)rshift_round64((int64_t)chirp_q16 * ar[i], 16);
    chirp_q16 += (int32_t)rshift_round64((int64_t)chirp_q16 * cm1, 16);
  }
  ar[d - 1] = (int16_t)rshift_round64((int64_t)chirp_q16 * ar[d - 1], 16);
}

// lpc_fit.rs:21-68
static void lpc_fit(int16_t* a_qout, int32_t* a_qin, int qout, int qin, int d) {
  int clipped = 1;
  for (int it = 0; it < 10; it++) {
    int32_t maxabs = 0; int idx = 0;
    for (int k = 0; k < d; k++) { int32_t a = iabs(a_qin[k]); if (a > maxabs) { maxabs = a; idx = k; } }
    int32_t m = rshift_round(maxabs, qin - qout);
    if (m <= 32767) { clipped = 0; break; }
    if (m > 163838) m = 163838;
    int32_t den = (m * (idx + 1)) >> 2; if (den < 1) den = 1;
    bwexpander_32(a_qin, d, 65470 - ((m - 32767) << 14) / den);
  }
  if (clipped) {
    for (int k = 0; k < d; k++) { a_qout[k] = sat16(rshift_round(a_qin[k], qin - qout)); a_qin[k] = w_shl(a_qout[k], qin - qout); }
  } else {
    for (int k = 0; k < d; k++) a_qout[k] = (int16_t)rshift_round(a_qin[k], qin - qout);
  }
}

// lpc_inv_pred_gain.rs:22-104 (QA = 24)
int32_t silk_lpc_inverse_pred_gain(const int16_t* a_q12, int order) {
  const int32_t A_LIMIT = (int32_t)((((int64_t)1 << 24) * 3999 + 2000) / 4000), MIN_INV = (int32_t)((((int64_t)1 << 30) + 5000) / 10000);
  int32_t a[24]; int32_t dc = 0;
  for (int k = 0; k < order; k++) { dc += a_q12[k]; a[k] = (int32_t)a_q12[k] << 12; }
  if (dc >= 4096) return 0;
  int32_t inv_gain = 1 << 30;
  for (int k = order - 1; k > 0; k--) {
    if (a[k] > A_LIMIT || a[k] < -A_LIMIT) return 0;
    int32_t rc_q31 = -w_shl(a[k], 31 - 24);
    int32_t rc_mult1 = (1 << 30) - smmul(rc_q31, rc_q31);
    inv_gain = w_shl(smmul(inv_gain, rc_mult1), 2);
    if (inv_gain < MIN_INV) return 0;
    int mult2q = 32 - clz32(iabs(rc_mult1));
    int32_t rc_mult2 = inverse32_varq(rc_mult1, mult2q + 30);
    for (int n = 0; n < (k + 1) >> 1; n++) {
      int32_t t1 = a[n], t2 = a[k - n - 1];
      int64_t u = rshift_round64((int64_t)sat32_64((int64_t)t1 - (int32_t)rshift_round64((int64_t)t2 * rc_q31, 31)) * rc_mult2, mult2q);
      if (u > INT32_MAX || u < INT32_MIN) return 0;
      a[n] = (int32_t)u;
      u = rshift_round64((int64_t)sat32_64((int64_t)t2 - (int32_t)rshift_round64((int64_t)t1 * rc_q31, 31)) * rc_mult2, mult2q);
      if (u > INT32_MAX || u < INT32_MIN) return 0;
      a[k - n - 1] = (int32_t)u;
    }
  }
  if (a[0] > A_LIMIT || a[0] < -A_LIMIT) return 0;
  int32_t rc_q31 = -w_shl(a[0], 31 - 24);
  int32_t rc_mult1 = (1 << 30) - smmul(rc_q31, rc_q31);
  inv_gain = w_shl(smmul(inv_gain, rc_mult1), 2);
  return inv_gain < MIN_INV ? 0 : inv_gain;
}

// nlsf2a.rs:25-117 (QA = 16)
static void nlsf2a_find_poly(int32_t* out, const int32_t* clsf, int dd) {
  out[0] = 1 << 16;
  out[1] = -clsf[0];
  for (int k = 1; k < dd; k++) {
    int32_t f = clsf[2 * k];
    out[k + 1] = w_sub(w_shl(out[k - 1], 1), (int32_t)rshift_round64((int64_t)f * out[k], 16));
    for (int n = k; n > 1; n--) out[n] = w_sub(w_add(out[n], out[n - 2]), (int32_t)rshift_round64((int64_t)f * out[n - 1], 16));
    out[1] = w_sub(out[1], f);
  }
}
void silk_nlsf2a(int16_t* a_q12, const int16_t* nlsf, int d, uint32_t quirks) {
  static const uint8_t ord16[16] = {0, 15, 8, 7, 4, 11, 12, 3, 2, 13, 10, 5, 6, 9, 14, 1}, ord10[10] = {0, 9, 6, 3, 4, 5, 8, 1, 2, 7};
  const uint8_t* ordering = d == 16 ? ord16 : ord10;
  int32_t cos_qa[SILK_MAX_ORDER];
  for (int k = 0; k < d; k++) {
    int src = (quirks & QUIRK_S2_NLSF2A_ORDER) ? ordering[k] : k;
    int32_t v = nlsf[src];
    int f_int = v >> 8; int32_t f_frac = v - (f_int << 8);
    int32_t c = silk_LSFCosTab_FIX_Q12[f_int], delta = (int16_t)(silk_LSFCosTab_FIX_Q12[f_int + 1] - silk_LSFCosTab_FIX_Q12[f_int]);
    cos_qa[ordering[k]] = (int32_t)rshift_round64(((int64_t)c << 8) + (int64_t)delta * f_frac, 20 - 16);
  }
  int dd = d / 2;
  int32_t P[9], Q[9], a32[SILK_MAX_ORDER];
  nlsf2a_find_poly(P, cos_qa, dd);
  nlsf2a_find_poly(Q, cos_qa + 1, dd);
  for (int k = 0; k < dd; k++) {
    int32_t pt = w_add(P[k + 1], P[k]), qt = w_sub(Q[k + 1], Q[k]);
    a32[k] = w_sub(w_sub(0, qt), pt);
    a32[d - k - 1] = w_sub(qt, pt);
  }
  lpc_fit(a_q12, a32, 12, 17, d);
  for (int i = 0; i < 16; i++) {
    if (silk_lpc_inverse_pred_gain(a_q12, d) != 0) return;
    bwexpander_32(a32, d, 65536 - (2 << i));
    for (int k = 0; k < d; k++) a_q12[k] = (int16_t)rshift_round(a32[k], 17 - 12);
  }
}

// gain_quant.rs:78-98
void silk_gains_dequant(int32_t* gain_q16, const int8_t* ind, int8_t* prev_ind, int conditional, int nb_subfr) {
  const int N_LEVELS = 64, MIN_DELTA = -4, MAX_DELTA = 36;
  const int32_t OFFSET = (2 * 128) / 6 + 16 * 128, INV_SCALE_Q16 = (65536 * (((88 - 2) * 128) / 6)) / (N_LEVELS - 1);
  int32_t prev = *prev_ind;
  for (int k = 0; k < nb_subfr; k++) {
    if (k == 0 && !conditional) {
      prev -= 16;
      if (prev < ind[k]) prev = ind[k];
    } else {
      int32_t t = ind[k] + MIN_DELTA, thr = 2 * MAX_DELTA - N_LEVELS + prev;
      prev += t > thr ? (t << 1) - thr : t;
    }
    prev = prev < 0 ? 0 : (prev > N_LEVELS - 1 ? N_LEVELS - 1 : prev);
    int32_t lg = smulwb(INV_SCALE_Q16, prev) + OFFSET;
    gain_q16[k] = silk_log2lin(lg < 3967 ? lg : 3967);
  }
  *prev_ind = (int8_t)prev;
}

// decode_pitch.rs:22-89
void silk_decode_pitch(int lag_index, int contour, int32_t* pitch_lags, int fs_khz, int nb_subfr) {
  int min_lag = 2 * fs_khz, max_lag = 18 * fs_khz, lag = min_lag + lag_index;
  for (int k = 0; k < nb_subfr; k++) {
    int off;
    if (fs_khz == 8) off = nb_subfr == 4 ? silk_CB_lags_stage2[k * 11 + contour] : silk_CB_lags_stage2_10_ms[k * 3 + contour];
    else off = nb_subfr == 4 ? silk_CB_lags_stage3[k * 34 + contour] : silk_CB_lags_stage3_10_ms[k * 12 + contour];
    int v = lag + off;
    pitch_lags[k] = v < min_lag ? min_lag : (v > max_lag ? max_lag : v);
  }
  for (int k = nb_subfr; k < SILK_MAX_SUBFR; k++) pitch_lags[k] = 0;
}

// ---- entropy side ----
// decode_indices.rs:109-269
static void decode_indices(SilkChannel* st, EcDec* ec, int frame_index, int decode_lbrr, int cond, SilkIndices* ix) {
  *ix = SilkIndices();
  int raw = (decode_lbrr || st->vad_flags[frame_index]) ? ec->icdf(silk_type_offset_VAD_iCDF, 8) + 2 : ec->icdf(silk_type_offset_no_VAD_iCDF, 8);
  ix->signal_type = (int8_t)(raw >> 1);
  ix->quant_offset_type = (int8_t)(raw & 1);
  if (cond == SILK_CODE_CONDITIONALLY) ix->gains[0] = (int8_t)ec->icdf(silk_delta_gain_iCDF, 8);
  else {
    int msb = ec->icdf(silk_gain_iCDF + 8 * ix->signal_type, 8);
    ix->gains[0] = (int8_t)((msb << 3) | ec->icdf(silk_uniform8_iCDF, 8));
  }
  for (int k = 1; k < st->nb_subfr; k++) ix->gains[k] = (int8_t)ec->icdf(silk_delta_gain_iCDF, 8);
  const SilkNlsfCb* cb = st->cb;
  ix->nlsf[0] = (int8_t)ec->icdf(cb->cb1_icdf + (ix->signal_type >> 1) * cb->n_vectors, 8);
  int16_t ec_ix[SILK_MAX_ORDER]; uint8_t pred_q8[SILK_MAX_ORDER];
  nlsf_unpack(ec_ix, pred_q8, cb, ix->nlsf[0]);
  for (int i = 0; i < cb->order; i++) {
    int s = ec->icdf(cb->ec_icdf + ec_ix[i], 8);
    if (s == 0) s -= ec->icdf(silk_NLSF_EXT_iCDF, 8);
    else if (s == 8) s += ec->icdf(silk_NLSF_EXT_iCDF, 8);
    ix->nlsf[i + 1] = (int8_t)(s - 4);
  }
  ix->nlsf_interp_q2 = st->nb_subfr == 4 ? (int8_t)ec->icdf(silk_NLSF_interpolation_factor_iCDF, 8) : 4;
  if (ix->signal_type == SILK_TYPE_VOICED) {
    int absolute = 1;
    if (cond == SILK_CODE_CONDITIONALLY && st->ec_prev_signal_type == SILK_TYPE_VOICED) {
      int delta = ec->icdf(silk_pitch_delta_iCDF, 8);
      if (delta > 0) {
        int v = st->ec_prev_lag_index + (delta - 9);
        ix->lag_index = sat16(v);  // i16 saturating_add
        absolute = 0;
      }
    }
    if (absolute) {
      int hi = ec->icdf(silk_pitch_lag_iCDF, 8);
      ix->lag_index = (int16_t)(hi * (st->fs_khz >> 1) + ec->icdf(st->pitch_lag_low_bits_icdf, 8));
    }
    st->ec_prev_lag_index = ix->lag_index;
    ix->contour_index = (int8_t)ec->icdf(st->pitch_contour_icdf, 8);
    ix->per_index = (int8_t)ec->icdf(silk_LTP_per_index_iCDF, 8);
    const uint8_t* g = ix->per_index == 0 ? silk_LTP_gain_iCDF_0 : (ix->per_index == 1 ? silk_LTP_gain_iCDF_1 : silk_LTP_gain_iCDF_2);
    for (int k = 0; k < st->nb_subfr; k++) ix->ltp[k] = (int8_t)ec->icdf(g, 8);
    ix->ltp_scale_index = cond == SILK_CODE_INDEPENDENTLY ? (int8_t)ec->icdf(silk_LTPscale_iCDF, 8) : 0;
  }
  ix->seed = (int8_t)ec->icdf(silk_uniform4_iCDF, 8);
  st->ec_prev_signal_type = ix->signal_type;
}

// shell_coder.rs:36-45, 94-126
static inline void split(EcDec* ec, int total, const uint8_t* table, int* a, int* b) {
  if (total > 0) { *a = ec->icdf(table + silk_shell_code_table_offsets[total], 8); *b = total - *a; }
  else { *a = 0; *b = 0; }
}
static void shell_decoder(int16_t* out, EcDec* ec, int total) {
  int p3[2], p2[4], p1[8], t[16];
  split(ec, total, silk_shell_code_table3, &p3[0], &p3[1]);
  split(ec, p3[0], silk_shell_code_table2, &p2[0], &p2[1]);
  split(ec, p2[0], silk_shell_code_table1, &p1[0], &p1[1]);
  split(ec, p1[0], silk_shell_code_table0, &t[0], &t[1]);
  split(ec, p1[1], silk_shell_code_table0, &t[2], &t[3]);
  split(ec, p2[1], silk_shell_code_table1, &p1[2], &p1[3]);
  split(ec, p1[2], silk_shell_code_table0, &t[4], &t[5]);
  split(ec, p1[3], silk_shell_code_table0, &t[6], &t[7]);
  split(ec, p3[1], silk_shell_code_table2, &p2[2], &p2[3]);
  split(ec, p2[2], silk_shell_code_table1, &p1[4], &p1[5]);
  split(ec, p1[4], silk_shell_code_table0, &t[8], &t[9]);
  split(ec, p1[5], silk_shell_code_table0, &t[10], &t[11]);
  split(ec, p2[3], silk_shell_code_table1, &p1[6], &p1[7]);
  split(ec, p1[6], silk_shell_code_table0, &t[12], &t[13]);
  split(ec, p1[7], silk_shell_code_table0, &t[14], &t[15]);
  for (int i = 0; i < 16; i++) out[i] = (int16_t)t[i];
}

// decode_pulses.rs:50-158 + code_signs.rs:93-125. `pulses` holds the padded length (multiple of 16).
static void decode_pulses(EcDec* ec, int16_t* pulses, int signal_type, int quant_offset_type, int frame_length, uint32_t quirks) {
  int rate = ec->icdf(silk_rate_levels_iCDF + 9 * (signal_type >> 1), 8);
  int nblocks = frame_length >> 4;
  if (nblocks * 16 < frame_length) nblocks++;
  int sum[20], nls[20];
  for (int i = 0; i < nblocks; i++) {
    nls[i] = 0;
    int s = ec->icdf(silk_pulses_per_block_iCDF + 18 * rate, 8);
    while (s == 17) {
      nls[i]++;
      s = ec->icdf(silk_pulses_per_block_iCDF + 18 * 9 + (nls[i] == 10), 8);
    }
    sum[i] = s;
  }
  for (int i = 0; i < nblocks; i++) {
    if (sum[i] > 0) shell_decoder(pulses + 16 * i, ec, sum[i]);
    else memset(pulses + 16 * i, 0, 16 * sizeof(int16_t));
  }
  for (int i = 0; i < nblocks; i++) {
    if (nls[i] == 0) continue;
    for (int k = 0; k < 16; k++) {
      int32_t q = pulses[16 * i + k];
      for (int j = 0; j < nls[i]; j++) q = (q << 1) + ec->icdf(silk_lsb_iCDF, 8);
      pulses[16 * i + k] = (int16_t)q;
    }
    sum[i] |= nls[i] << 5;
  }
  // signs
  uint8_t icdf[2] = {0, 0};
  const uint8_t* base = silk_sign_iCDF + 7 * (quant_offset_type + (signal_type << 1));
  int nsb = (frame_length + 8) >> 4;
  for (int i = 0; i < nsb; i++) {
    if (sum[i] <= 0) continue;
    int ti = sum[i] & 0x1f; if (ti > 6) ti = 6;
    icdf[0] = base[ti];
    int end = 16 * i + 16;
    if ((quirks & QUIRK_S3_SIGNS_FRAMELEN) && end > frame_length) end = frame_length;
    for (int k = 16 * i; k < end; k++)
      if (pulses[k] > 0) pulses[k] = (int16_t)(pulses[k] * ((ec->icdf(icdf, 8) << 1) - 1));
  }
}

// stereo_decode_pred.rs:21-56 (icdf contexts of icdf.rs:128-153 == the tables_other.rs iCDF tables)
static void stereo_decode_pred(EcDec* ec, int32_t pred_q13[2]) {
  int ix[2][3];
  int n = ec->icdf(silk_stereo_pred_joint_iCDF, 8);
  ix[0][2] = n / 5; ix[1][2] = n - 5 * ix[0][2];
  for (int c = 0; c < 2; c++) { ix[c][0] = ec->icdf(silk_uniform3_iCDF, 8); ix[c][1] = ec->icdf(silk_uniform5_iCDF, 8); }
  for (int c = 0; c < 2; c++) {
    ix[c][0] += 3 * ix[c][2];
    int32_t low = silk_stereo_pred_quant_Q13[ix[c][0]];
    int32_t step = smulwb(silk_stereo_pred_quant_Q13[ix[c][0] + 1] - low, ((1 << 15) + 2) / 5);
    pred_q13[c] = smlabb(low, step, 2 * ix[c][1] + 1);
  }
  pred_q13[0] = w_sub(pred_q13[0], pred_q13[1]);
}

// ---- parameters + synthesis ----
// decode_parameters.rs:79-199
static void decode_parameters(SilkChannel* st, SilkControl* c, int cond, uint32_t quirks, int16_t* nlsf_trace) {
  int8_t lgi = (int8_t)st->last_gain_index;
  silk_gains_dequant(c->gains_q16, st->indices.gains, &lgi, cond == SILK_CODE_CONDITIONALLY, st->nb_subfr);
  st->last_gain_index = lgi;
  for (int k = st->nb_subfr; k < SILK_MAX_SUBFR; k++) c->gains_q16[k] = 0;
  int order = st->lpc_order;
  int16_t nlsf[SILK_MAX_ORDER] = {0}, nlsf0[SILK_MAX_ORDER];
  silk_nlsf_decode(nlsf, st->indices.nlsf, st->cb, quirks);
  if (nlsf_trace) memcpy(nlsf_trace, nlsf, sizeof(nlsf));
  memset(c->pred_coef_q12, 0, sizeof(c->pred_coef_q12));
  silk_nlsf2a(c->pred_coef_q12[1], nlsf, order, quirks);
  if (st->first_frame_after_reset) st->indices.nlsf_interp_q2 = 4;
  if (st->indices.nlsf_interp_q2 < 4) {
    for (int i = 0; i < order; i++)
      nlsf0[i] = sat16(st->prev_nlsf_q15[i] + ((st->indices.nlsf_interp_q2 * ((int32_t)nlsf[i] - st->prev_nlsf_q15[i])) >> 2));
    silk_nlsf2a(c->pred_coef_q12[0], nlsf0, order, quirks);
  } else memcpy(c->pred_coef_q12[0], c->pred_coef_q12[1], order * sizeof(int16_t));
  memset(st->prev_nlsf_q15, 0, sizeof(st->prev_nlsf_q15));
  memcpy(st->prev_nlsf_q15, nlsf, order * sizeof(int16_t));
  if (st->loss_count > 0) { bwexpander(c->pred_coef_q12[0], order, 63570); bwexpander(c->pred_coef_q12[1], order, 63570); }
  memset(c->pitch_l, 0, sizeof(c->pitch_l));
  memset(c->ltp_coef_q14, 0, sizeof(c->ltp_coef_q14));
  c->ltp_scale_q14 = 0;
  if (st->indices.signal_type == SILK_TYPE_VOICED) {
    silk_decode_pitch(st->indices.lag_index, st->indices.contour_index, c->pitch_l, st->fs_khz, st->nb_subfr);
    const int8_t* vq = st->indices.per_index == 0 ? silk_LTP_gain_vq_0 : (st->indices.per_index == 1 ? silk_LTP_gain_vq_1 : silk_LTP_gain_vq_2);
    for (int k = 0; k < st->nb_subfr; k++)
      for (int i = 0; i < SILK_LTP_ORDER; i++) c->ltp_coef_q14[k * SILK_LTP_ORDER + i] = (int16_t)(vq[st->indices.ltp[k] * SILK_LTP_ORDER + i] << 7);
    c->ltp_scale_q14 = silk_LTPScales_table_Q14[st->indices.ltp_scale_index];
  } else st->indices.per_index = 0;
}

// lpc_analysis_filter.rs:18-45
static void lpc_analysis_filter(int16_t* out, const int16_t* in, const int16_t* B, int len, int d) {
  for (int ix = d; ix < len; ix++) {
    int32_t acc = (int32_t)in[ix - 1] * B[0];
    for (int k = 1; k < d; k++) acc = w_add(acc, (int32_t)in[ix - 1 - k] * B[k]);
    out[ix] = sat16(rshift_round(w_sub(w_shl(in[ix], 12), acc), 12));
  }
  for (int i = 0; i < d && i < len; i++) out[i] = 0;
}

// decode_core.rs:28-330
static void decode_core(SilkChannel* st, SilkControl* c, int16_t* xq, const int16_t* pulses) {
  int16_t sLTP[320]; int32_t sLTP_Q15[320 + 320], res_Q14[80], sLPC[80 + 16];
  memcpy(sLPC, st->s_lpc_q14, sizeof(st->s_lpc_q14));
  int32_t offset_q10 = silk_Quantization_Offsets_Q10[(st->indices.signal_type >> 1) * 2 + st->indices.quant_offset_type];
  int interp_flag = st->indices.nlsf_interp_q2 < 4;
  int32_t seed = st->indices.seed;
  for (int i = 0; i < st->frame_length; i++) {
    seed = w_add(907633515, w_mul(seed, 196314165));
    int32_t e = (int32_t)pulses[i] << 14;
    if (e > 0) e -= 80 << 4; else if (e < 0) e += 80 << 4;
    e += offset_q10 << 4;
    if (seed < 0) e = -e;
    st->exc_q14[i] = e;
    seed = w_add(seed, pulses[i]);
  }
  int sLTP_buf_idx = st->ltp_mem_length;
  const int32_t* pexc = st->exc_q14;
  int16_t* pxq = xq;
  for (int k = 0; k < st->nb_subfr; k++) {
    const int16_t* A = c->pred_coef_q12[k >> 1];
    int16_t* B = c->ltp_coef_q14 + k * SILK_LTP_ORDER;
    int32_t gain_q16 = c->gains_q16[k], gain_q10 = gain_q16 >> 6;
    int32_t inv_gain_q31 = inverse32_varq(gain_q16, 47);
    int32_t gain_adj = 1 << 16;
    if (gain_q16 != st->prev_gain_q16) {
      gain_adj = div32_varq(st->prev_gain_q16, gain_q16, 16);
      for (int i = 0; i < 16; i++) sLPC[i] = smulww(gain_adj, sLPC[i]);
    }
    st->prev_gain_q16 = gain_q16;
    int signal_type = st->indices.signal_type;
    if (st->loss_count > 0 && st->prev_signal_type == SILK_TYPE_VOICED && signal_type != SILK_TYPE_VOICED && k < 2) {
      memset(B, 0, SILK_LTP_ORDER * sizeof(int16_t));
      B[2] = 4096;
      signal_type = SILK_TYPE_VOICED;
      c->pitch_l[k] = st->lag_prev;
    }
    const int32_t* pres;
    if (signal_type == SILK_TYPE_VOICED) {
      int lag = c->pitch_l[k];
      if (k == 0 || (k == 2 && interp_flag)) {
        int start = st->ltp_mem_length - lag - st->lpc_order - SILK_LTP_ORDER / 2;
        if (k == 2) memcpy(st->out_buf + st->ltp_mem_length, xq, 2 * st->subfr_length * sizeof(int16_t));
        lpc_analysis_filter(sLTP + start, st->out_buf + start + k * st->subfr_length, A, st->ltp_mem_length - start, st->lpc_order);
        if (k == 0) inv_gain_q31 = w_shl(smulwb(inv_gain_q31, c->ltp_scale_q14), 2);
        for (int i = 0; i < lag + SILK_LTP_ORDER / 2; i++) sLTP_Q15[sLTP_buf_idx - i - 1] = smulwb(inv_gain_q31, sLTP[st->ltp_mem_length - i - 1]);
      } else if (gain_adj != (1 << 16)) {
        for (int i = 0; i < lag + SILK_LTP_ORDER / 2; i++) sLTP_Q15[sLTP_buf_idx - i - 1] = smulww(gain_adj, sLTP_Q15[sLTP_buf_idx - i - 1]);
      }
      int pl = sLTP_buf_idx - lag + SILK_LTP_ORDER / 2;
      for (int i = 0; i < st->subfr_length; i++) {
        int32_t p = 2;
        p = smlawb(p, sLTP_Q15[pl], B[0]);
        p = smlawb(p, sLTP_Q15[pl - 1], B[1]);
        p = smlawb(p, sLTP_Q15[pl - 2], B[2]);
        p = smlawb(p, sLTP_Q15[pl - 3], B[3]);
        p = smlawb(p, sLTP_Q15[pl - 4], B[4]);
        pl++;
        res_Q14[i] = w_add(pexc[i], w_shl(p, 1));
        sLTP_Q15[sLTP_buf_idx++] = w_shl(res_Q14[i], 1);
      }
      pres = res_Q14;
    } else pres = pexc;
    for (int i = 0; i < st->subfr_length; i++) {
      int32_t p = st->lpc_order >> 1;
      for (int j = 0; j < st->lpc_order; j++) p = smlawb(p, sLPC[16 + i - 1 - j], A[j]);
      sLPC[16 + i] = sat32_64((int64_t)pres[i] + lshift_sat32(p, 4));
      pxq[i] = sat16(rshift_round(smulww(sLPC[16 + i], gain_q10), 8));
    }
    memmove(sLPC, sLPC + st->subfr_length, 16 * sizeof(int32_t));
    pexc += st->subfr_length;
    pxq += st->subfr_length;
  }
  memcpy(st->s_lpc_q14, sLPC, sizeof(st->s_lpc_q14));
}

// decoder_set_fs.rs:137-198
int SilkChannel::set_fs(int fs_khz_, int fs_api_hz_) {
  if (fs_khz_ != 8 && fs_khz_ != 12 && fs_khz_ != 16) return -1;
  int sl = 5 * fs_khz_, fl = nb_subfr * sl;
  if (fs_khz != fs_khz_ || fs_api_hz != fs_api_hz_) {
    if (resampler.init(fs_khz_ * 1000, fs_api_hz_) < 0) return -1;
    fs_api_hz = fs_api_hz_;
  }
  if (fs_khz != fs_khz_ || frame_length != fl) {
    pitch_contour_icdf = fs_khz_ == 8 ? (nb_subfr == 4 ? silk_pitch_contour_NB_iCDF : silk_pitch_contour_10_ms_NB_iCDF)
                                      : (nb_subfr == 4 ? silk_pitch_contour_iCDF : silk_pitch_contour_10_ms_iCDF);
    if (fs_khz != fs_khz_) {
      ltp_mem_length = 20 * fs_khz_;
      if (fs_khz_ == 16) { lpc_order = 16; cb = &CB_WB; } else { lpc_order = 10; cb = &CB_NB_MB; }
      pitch_lag_low_bits_icdf = fs_khz_ == 8 ? silk_uniform4_iCDF : (fs_khz_ == 12 ? silk_uniform6_iCDF : silk_uniform8_iCDF);
      first_frame_after_reset = 1; lag_prev = 100; last_gain_index = 10; prev_signal_type = 0;
      memset(out_buf, 0, sizeof(out_buf));
      memset(s_lpc_q14, 0, sizeof(s_lpc_q14));
    }
    fs_khz = fs_khz_; subfr_length = sl; frame_length = fl;
  }
  return 0;
}

// ---- resampler (decoder side, up-sampling and copy only) ----
// resampler.rs:87-176
int SilkResampler::init(int fs_in, int fs_out) {
  static const int in_rates[3] = {8000, 12000, 16000}, out_rates[5] = {8000, 12000, 16000, 24000, 48000};
  int ii = -1, oi = -1;
  for (int i = 0; i < 3; i++) if (in_rates[i] == fs_in) ii = i;
  for (int i = 0; i < 5; i++) if (out_rates[i] == fs_out) oi = i;
  if (ii < 0 || oi < 0) return -1;
  *this = SilkResampler();
  input_delay = silk_delay_matrix_dec[ii * 5 + oi];
  fs_in_khz = fs_in / 1000; fs_out_khz = fs_out / 1000; batch_size = fs_in_khz * 10;
  mode = fs_out == fs_in ? 0 : (fs_out == 2 * fs_in ? 1 : (fs_out > fs_in ? 2 : 3));
  if (mode == 3) {   // down_fir_config (resampler.rs:246-264); a decoder meets 3/4 (16 -> 12), 2/3 (12 -> 8) and 1/2 (16 -> 8)
    if ((int64_t)fs_out * 4 == (int64_t)fs_in * 3) { down_fracs = 3; down_order = 18; down_coefs = silk_Resampler_3_4_COEFS; }
    else if ((int64_t)fs_out * 3 == (int64_t)fs_in * 2) { down_fracs = 2; down_order = 18; down_coefs = silk_Resampler_2_3_COEFS; }
    else if (fs_out * 2 == fs_in) { down_fracs = 1; down_order = 24; down_coefs = silk_Resampler_1_2_COEFS; }
    else return -2;
  }
  int up2x = mode == 2;
  int64_t inv = ((int64_t)fs_in << (16 + up2x)) / fs_out;
  while (((inv * fs_out) >> 16) < ((int64_t)fs_in << up2x)) inv++;
  inv_ratio_q16 = (int32_t)inv;
  return 0;
}
// resampler_private_up2_hq.rs:31-117
static void up2_hq(int32_t* S, int16_t* out, const int16_t* in, int len) {
  for (int k = 0; k < len; k++) {
    int32_t in32 = (int32_t)in[k] << 10, Y, X, o1, o2;
    Y = w_sub(in32, S[0]); X = smulwb(Y, silk_resampler_up2_hq_0[0]); o1 = w_add(S[0], X); S[0] = w_add(in32, X);
    Y = w_sub(o1, S[1]); X = smulwb(Y, silk_resampler_up2_hq_0[1]); o2 = w_add(S[1], X); S[1] = w_add(o1, X);
    Y = w_sub(o2, S[2]); X = smlawb(Y, Y, silk_resampler_up2_hq_0[2]); o1 = w_add(S[2], X); S[2] = w_add(o2, X);
    out[2 * k] = sat16(rshift_round(o1, 10));
    Y = w_sub(in32, S[3]); X = smulwb(Y, silk_resampler_up2_hq_1[0]); o1 = w_add(S[3], X); S[3] = w_add(in32, X);
    Y = w_sub(o1, S[4]); X = smulwb(Y, silk_resampler_up2_hq_1[1]); o2 = w_add(S[4], X); S[4] = w_add(o1, X);
    Y = w_sub(o2, S[5]); X = smlawb(Y, Y, silk_resampler_up2_hq_1[2]); o1 = w_add(S[5], X); S[5] = w_add(o2, X);
    out[2 * k + 1] = sat16(rshift_round(o1, 10));
  }
}
// resampler_private_iir_fir.rs:80-197
static int iir_fir(SilkResampler* R, int16_t* out, const int16_t* in, int in_len) {
  if (in_len <= 0) return 0;
  int16_t buf[2 * 160 + 8];
  memcpy(buf, R->s_fir, sizeof(R->s_fir));
  int produced = 0, n = 0;
  while (in_len > 0) {
    n = in_len < R->batch_size ? in_len : R->batch_size;
    up2_hq(R->s_iir, buf + 8, in, n);
    int32_t max_idx = n << 17;
    for (int32_t idx = 0; idx < max_idx; idx = w_add(idx, R->inv_ratio_q16)) {
      int ti = smulwb(idx & 0xFFFF, 12);
      const int16_t* b = buf + (idx >> 16);
      const int16_t* f = silk_resampler_frac_FIR_12 + 4 * ti; const int16_t* r = silk_resampler_frac_FIR_12 + 4 * (11 - ti);
      int32_t acc = smulbb(b[0], f[0]);
      acc = smlabb(acc, b[1], f[1]); acc = smlabb(acc, b[2], f[2]); acc = smlabb(acc, b[3], f[3]);
      acc = smlabb(acc, b[4], r[3]); acc = smlabb(acc, b[5], r[2]); acc = smlabb(acc, b[6], r[1]); acc = smlabb(acc, b[7], r[0]);
      out[produced++] = sat16(w_add(acc, 1 << 14) >> 15);
    }
    in += n; in_len -= n;
    if (in_len > 0) memmove(buf, buf + 2 * n, 8 * sizeof(int16_t));
  }
  memcpy(R->s_fir, buf + 2 * n, sizeof(R->s_fir));
  return produced;
}
// resampler_private_ar2.rs:19-38 + resampler_private_down_fir.rs:139-284
static int down_fir(SilkResampler* R, int16_t* out, const int16_t* in, int in_len) {
  if (in_len <= 0) return 0;
  const int order = R->down_order, half = order / 2, fracs = R->down_fracs;
  const int16_t* fir = R->down_coefs + 2;
  const int32_t a0 = R->down_coefs[0], a1 = R->down_coefs[1];
  int32_t buf[160 + 36];
  memcpy(buf, R->s_down, order * sizeof(int32_t));
  int produced = 0, n = 0;
  while (in_len > 0) {
    n = in_len < R->batch_size ? in_len : R->batch_size;
    for (int k = 0; k < n; k++) {   // second-order AR filter, Q8 out
      int32_t o32 = w_add(R->s_iir[0], w_shl((int32_t)in[k], 8));
      buf[order + k] = o32;
      o32 = w_shl(o32, 2);
      R->s_iir[0] = smlawb(R->s_iir[1], o32, a0);
      R->s_iir[1] = smulwb(o32, a1);
    }
    const int32_t max_idx = n << 16;
    for (int32_t idx = 0; idx < max_idx; idx = w_add(idx, R->inv_ratio_q16)) {
      const int32_t* b = buf + (idx >> 16);
      int32_t acc;
      if (order == 18) {
        const int ind = smulwb(idx & 0xFFFF, fracs);
        const int16_t* f = fir + half * ind;
        const int16_t* r = fir + half * (fracs - 1 - ind);
        acc = smulwb(b[0], f[0]);
        for (int k = 1; k < half; k++) acc = smlawb(acc, b[k], f[k]);
        for (int k = 0; k < half; k++) acc = smlawb(acc, b[order - 1 - k], r[k]);
      } else {
        acc = smulwb(w_add(b[0], b[order - 1]), fir[0]);
        for (int k = 1; k < half; k++) acc = smlawb(acc, w_add(b[k], b[order - 1 - k]), fir[k]);
      }
      out[produced++] = sat16(w_add(acc, 1 << 5) >> 6);
    }
    in += n; in_len -= n;
    if (in_len > 0) memmove(buf, buf + n, order * sizeof(int32_t));
  }
  memcpy(R->s_down, buf + n, order * sizeof(int32_t));
  return produced;
}
// resampler.rs:279-369
int SilkResampler::run(int16_t* out, const int16_t* in, int in_len) {
  int ns = fs_in_khz - input_delay;
  memcpy(delay_buf + input_delay, in, ns * sizeof(int16_t));
  const int16_t* in2 = in + ns; int len2 = in_len - input_delay - ns; if (len2 < 0) len2 = 0;
  int produced = 0;
  if (mode == 0) {
    memcpy(out, delay_buf, fs_in_khz * sizeof(int16_t)); produced = fs_in_khz;
    memcpy(out + produced, in2, len2 * sizeof(int16_t)); produced += len2;
  } else if (mode == 1) {
    up2_hq(s_iir, out, delay_buf, fs_in_khz); produced = fs_out_khz;
    up2_hq(s_iir, out + produced, in2, len2); produced += 2 * len2;
  } else if (mode == 2) {
    produced = iir_fir(this, out, delay_buf, fs_in_khz);
    produced += iir_fir(this, out + produced, in2, len2);
  } else {
    produced = down_fir(this, out, delay_buf, fs_in_khz);
    produced += down_fir(this, out + produced, in2, len2);
  }
  if (input_delay > 0) memcpy(delay_buf, in + in_len - input_delay, input_delay * sizeof(int16_t));
  return produced;
}

// ---- packet-loss concealment (plc.rs) and comfort noise (cng.rs) ----
static inline int32_t silk_rand(int32_t seed) { return w_add(907633515, w_mul(seed, 196314165)); }   // cng.rs:371-373
static inline int32_t rshift_any(int32_t v, int s) { return s <= 0 ? w_shl(v, -s) : v >> s; }           // plc.rs:372-378
static int32_t sqrt_approx(int32_t x) {   // plc.rs:448-457
  if (x <= 0) return 0;
  int lz = clz32(x);
  uint32_t ux = (uint32_t)x;
  int rot = (24 - lz) & 31;
  int32_t frac = (int32_t)(((ux >> rot) | (rot ? ux << (32 - rot) : 0)) & 0x7f);
  int32_t y = (lz & 1) ? 32768 : 46214;
  y >>= lz >> 1;
  return smlawb(y, y, smulbb(213, frac));
}
static void sum_sqr_shift(const int16_t* x, int len, int32_t* energy, int* shift_out) {   // sum_sqr_shift.rs:12-75
  if (len == 0) { *energy = 0; *shift_out = 0; return; }
  auto acc_energy = [&](int32_t acc, int shift) {
    int i = 0;
    for (; i + 1 < len; i += 2) {
      uint32_t pe = (uint32_t)((int32_t)x[i] * x[i]) + (uint32_t)((int32_t)x[i + 1] * x[i + 1]);
      acc = w_add(acc, (int32_t)(shift <= 0 ? pe : (shift >= 32 ? 0u : pe >> shift)));
    }
    if (i < len) {
      uint32_t pe = (uint32_t)((int32_t)x[i] * x[i]);
      acc = w_add(acc, (int32_t)(shift <= 0 ? pe : (shift >= 32 ? 0u : pe >> shift)));
    }
    return acc;
  };
  int shift = 31 - __builtin_clz((uint32_t)len);
  int32_t e = acc_energy(len, shift);
  int lz = e == 0 ? 32 : __builtin_clz((uint32_t)e);
  shift = shift + 3 - lz; if (shift < 0) shift = 0;
  *energy = acc_energy(0, shift);
  *shift_out = shift;
}

void SilkPlc::reset(int frame_length) {   // decoder_state.rs:77-95
  *this = SilkPlc();
  pitch_l_q8 = frame_length << 7;
}
void SilkCng::reset(int lpc_order) {   // cng.rs:62-80
  int32_t step = 32767 / (lpc_order + 1), acc = 0;
  for (int i = 0; i < SILK_MAX_ORDER; i++) {
    if (i < lpc_order) { acc += step; smth_nlsf_q15[i] = (int16_t)acc; } else smth_nlsf_q15[i] = 0;
  }
  smth_gain_q16 = 0; rand_seed = 3176576;
  memset(synth_state, 0, sizeof(synth_state));
}

// plc.rs:117-184 (silk_plc_update)
static void plc_update(SilkChannel* st, const SilkControl* c) {
  SilkPlc* plc = &st->plc;
  st->prev_signal_type = st->indices.signal_type;
  const int nb = st->nb_subfr;
  int32_t ltp_gain_q14 = 0;
  if (st->indices.signal_type == SILK_TYPE_VOICED) {
    const int32_t target = c->pitch_l[nb - 1];
    for (int j = 0; j < nb && j * st->subfr_length < target; j++) {
      const int sf = nb - 1 - j;
      int32_t tg = 0;
      for (int i = 0; i < SILK_LTP_ORDER; i++) tg += c->ltp_coef_q14[sf * SILK_LTP_ORDER + i];
      if (tg > ltp_gain_q14) { ltp_gain_q14 = tg; plc->pitch_l_q8 = w_shl(c->pitch_l[sf], 8); }
    }
    memset(plc->ltp_coef_q14, 0, sizeof(plc->ltp_coef_q14));
    plc->ltp_coef_q14[SILK_LTP_ORDER / 2] = sat16(ltp_gain_q14);
    if (ltp_gain_q14 < 11469) {
      int32_t scale_q10 = (11469 << 10) / (ltp_gain_q14 > 1 ? ltp_gain_q14 : 1);
      for (int i = 0; i < SILK_LTP_ORDER; i++) plc->ltp_coef_q14[i] = sat16(smulbb(plc->ltp_coef_q14[i], scale_q10) >> 10);
    } else if (ltp_gain_q14 > 15565) {
      int32_t scale_q14 = (15565 << 14) / (ltp_gain_q14 > 1 ? ltp_gain_q14 : 1);
      for (int i = 0; i < SILK_LTP_ORDER; i++) plc->ltp_coef_q14[i] = sat16(smulbb(plc->ltp_coef_q14[i], scale_q14) >> 14);
    }
  } else {
    plc->pitch_l_q8 = w_shl(smulbb(st->fs_khz, 18), 8);
    memset(plc->ltp_coef_q14, 0, sizeof(plc->ltp_coef_q14));
  }
  memset(plc->prev_lpc_q12, 0, sizeof(plc->prev_lpc_q12));
  memcpy(plc->prev_lpc_q12, c->pred_coef_q12[1], st->lpc_order * sizeof(int16_t));
  plc->prev_ltp_scale_q14 = sat16(c->ltp_scale_q14);
  plc->prev_gain_q16[0] = c->gains_q16[nb - 2];
  plc->prev_gain_q16[1] = c->gains_q16[nb - 1];
  plc->subfr_length = st->subfr_length;
  plc->nb_subfr = nb;
}

// plc.rs:186-345 (silk_plc_conceal)
static void plc_conceal(SilkChannel* st, SilkControl* c, int16_t* frame) {
  SilkPlc* plc = &st->plc;
  const int frame_length = st->frame_length, nb_subfr = st->nb_subfr, subfr_length = st->subfr_length;
  const int ltp_mem = st->ltp_mem_length, order = st->lpc_order, fs_khz = st->fs_khz;
  const int32_t prev_gain_q10[2] = {plc->prev_gain_q16[0] >> 6, plc->prev_gain_q16[1] >> 6};
  // silk_plc_energy (plc.rs:347-368)
  int32_t e1, e2; int s1, s2;
  {
    int16_t buf[2][80];
    for (int k = 0; k < 2; k++) {
      const int base = (k + nb_subfr - 2) * subfr_length;
      for (int i = 0; i < subfr_length; i++) {
        const int idx = base + i;
        const int32_t v = idx < SILK_MAX_FRAME ? st->exc_q14[idx] : 0;
        buf[k][i] = sat16(smulww(v, prev_gain_q10[k]) >> 8);
      }
    }
    sum_sqr_shift(buf[0], subfr_length, &e1, &s1);
    sum_sqr_shift(buf[1], subfr_length, &e2, &s2);
  }
  // pick_rand_slice (plc.rs:370-388): the PLC state's own subframe geometry, clamped into the excitation buffer
  int32_t rand_slice[128];
  {
    const int second = !(rshift_any(e1, s2) < rshift_any(e2, s1));
    const int sl = plc->subfr_length > 1 ? plc->subfr_length : 1, nb = plc->nb_subfr > 2 ? plc->nb_subfr : 2;
    const int base = second ? nb * sl : (nb - 1) * sl;
    int start = base - 128; if (start < 0) start = 0;
    if (start > SILK_MAX_FRAME - 128) start = SILK_MAX_FRAME - 128;
    memcpy(rand_slice, st->exc_q14 + start, sizeof(rand_slice));
  }
  if (st->first_frame_after_reset) memset(plc->prev_lpc_q12, 0, sizeof(plc->prev_lpc_q12));
  int16_t* B = plc->ltp_coef_q14;
  int32_t rand_scale_q14 = plc->rand_scale_q14;
  static const int16_t HARM_ATT[2] = {32440, 31130}, RAND_V[2] = {31130, 26214}, RAND_UV[2] = {32440, 29491};
  const int att = st->loss_count < 1 ? st->loss_count : 1;
  const int32_t harm_gain_q15 = HARM_ATT[att];
  const int32_t rand_gain_q15 = st->prev_signal_type == SILK_TYPE_VOICED ? RAND_V[att] : RAND_UV[att];
  bwexpander(plc->prev_lpc_q12, order, 64881);
  int16_t a_q12[SILK_MAX_ORDER] = {0};
  memcpy(a_q12, plc->prev_lpc_q12, order * sizeof(int16_t));
  if (st->loss_count == 0) {
    rand_scale_q14 = 1 << 14;
    if (st->prev_signal_type == SILK_TYPE_VOICED) {
      for (int i = 0; i < SILK_LTP_ORDER; i++) rand_scale_q14 -= B[i];
      if (rand_scale_q14 < 3277) rand_scale_q14 = 3277;
      rand_scale_q14 = smulbb(rand_scale_q14, plc->prev_ltp_scale_q14) >> 14;
    } else {
      const int32_t inv_gain_q30 = silk_lpc_inverse_pred_gain(plc->prev_lpc_q12, order);
      int32_t down = (1 << 30) >> 3; if (inv_gain_q30 < down) down = inv_gain_q30;
      if (down < ((1 << 30) >> 8)) down = (1 << 30) >> 8;
      down = w_shl(down, 3);
      rand_scale_q14 = smulwb(down, rand_gain_q15) >> 14;
    }
  }
  int32_t rand_seed = plc->rand_seed;
  int lag = rshift_round(plc->pitch_l_q8, 8);
  int32_t s_ltp_q14[320 + SILK_MAX_FRAME] = {0};
  int16_t s_ltp[320] = {0};
  const int idx = ltp_mem - lag - order - SILK_LTP_ORDER / 2;
  lpc_analysis_filter(s_ltp + idx, st->out_buf + idx, a_q12, ltp_mem - idx, order);
  int32_t inv_gain_q30 = inverse32_varq(plc->prev_gain_q16[1], 46);
  if (inv_gain_q30 > (INT32_MAX >> 1)) inv_gain_q30 = INT32_MAX >> 1;
  for (int i = idx + order; i < ltp_mem; i++) s_ltp_q14[i] = smulwb(inv_gain_q30, s_ltp[i]);
  int buf_idx = ltp_mem;
  for (int k = 0; k < nb_subfr; k++) {
    for (int i = 0; i < subfr_length; i++) {
      int32_t pred = 2;
      for (int t = 0; t < SILK_LTP_ORDER; t++) pred = smlawb(pred, s_ltp_q14[buf_idx - lag + t - SILK_LTP_ORDER / 2], B[t]);
      rand_seed = silk_rand(rand_seed);
      s_ltp_q14[buf_idx] = lshift_sat32(smlawb(pred, rand_slice[(rand_seed >> 25) & 127], rand_scale_q14), 2);
      buf_idx++;
    }
    for (int t = 0; t < SILK_LTP_ORDER; t++) B[t] = sat16(smulbb(harm_gain_q15, B[t]) >> 15);
    rand_scale_q14 = smulbb(rand_scale_q14, rand_gain_q15) >> 15;
    plc->pitch_l_q8 = w_add(plc->pitch_l_q8, smulwb(plc->pitch_l_q8, 655));
    const int32_t max_lag_q8 = w_shl(smulbb(18, fs_khz), 8);
    if (plc->pitch_l_q8 > max_lag_q8) plc->pitch_l_q8 = max_lag_q8;
    lag = rshift_round(plc->pitch_l_q8, 8);
  }
  int32_t* s_lpc = s_ltp_q14 + ltp_mem - SILK_MAX_ORDER;
  memcpy(s_lpc, st->s_lpc_q14, sizeof(st->s_lpc_q14));
  for (int i = 0; i < frame_length; i++) {
    int32_t pred = order >> 1;
    for (int j = 0; j < order; j++) pred = smlawb(pred, s_lpc[SILK_MAX_ORDER + i - j - 1], a_q12[j]);
    const int32_t upd = sat32_64((int64_t)s_lpc[SILK_MAX_ORDER + i] + lshift_sat32(pred, 4));
    s_lpc[SILK_MAX_ORDER + i] = upd;
    frame[i] = sat16(rshift_round(smulww(upd, prev_gain_q10[1]), 8));
  }
  memcpy(st->s_lpc_q14, s_lpc + frame_length, sizeof(st->s_lpc_q14));
  plc->rand_seed = rand_seed;
  plc->rand_scale_q14 = sat16(rand_scale_q14);
  for (int k = 0; k < SILK_MAX_SUBFR; k++) c->pitch_l[k] = lag;
}

// plc.rs:46-72 (silk_plc)
static void silk_plc(SilkChannel* st, SilkControl* c, int16_t* frame, int lost) {
  if (st->fs_khz != st->plc.fs_khz) { st->plc.reset(st->frame_length); st->plc.fs_khz = st->fs_khz; }
  if (lost) {
    plc_conceal(st, c, frame);
    st->loss_count = st->loss_count == INT32_MAX ? INT32_MAX : st->loss_count + 1;
  } else plc_update(st, c);
}

// plc.rs:74-115 (silk_plc_glue_frames)
static void plc_glue_frames(SilkChannel* st, int16_t* frame, int length) {
  SilkPlc* plc = &st->plc;
  if (st->loss_count > 0) {
    sum_sqr_shift(frame, length, &plc->conc_energy, &plc->conc_energy_shift);
    plc->last_frame_lost = 1;
    return;
  }
  if (!plc->last_frame_lost) return;
  int32_t energy; int eshift;
  sum_sqr_shift(frame, length, &energy, &eshift);
  if (eshift > plc->conc_energy_shift) plc->conc_energy >>= eshift - plc->conc_energy_shift;
  else if (eshift < plc->conc_energy_shift) energy >>= plc->conc_energy_shift - eshift;
  if (energy > plc->conc_energy) {
    const int lz = clz32(plc->conc_energy > 1 ? plc->conc_energy : 1) - 1;
    plc->conc_energy = w_shl(plc->conc_energy, lz);
    const int rs = 24 - lz > 0 ? 24 - lz : 0;
    if (rs > 0) energy >>= rs;
    const int32_t frac_q24 = plc->conc_energy / (energy > 1 ? energy : 1);
    int32_t gain_q16 = w_shl(sqrt_approx(frac_q24), 4);
    int32_t slope_q16 = ((1 << 16) - gain_q16) / length;
    slope_q16 = w_shl(slope_q16, 2);
    for (int i = 0; i < length; i++) {
      frame[i] = sat16(smulwb(gain_q16, frame[i]));
      gain_q16 = sat32_64((int64_t)gain_q16 + slope_q16);
      if (gain_q16 > (1 << 16)) break;
    }
  }
  plc->last_frame_lost = 0;
}

// cng.rs:145-370 (apply_cng)
static void apply_cng(SilkChannel* st, const SilkControl* c, int16_t* frame, uint32_t quirks) {
  SilkCng* g = &st->cng;
  const int order = st->lpc_order, length = st->frame_length;
  if (st->fs_khz != g->fs_khz) {
    if (g->fs_khz == -1) memset(g->synth_state, 0, order * sizeof(int32_t));
    else g->reset(order);
    memset(g->smth_nlsf_q15, 0, sizeof(g->smth_nlsf_q15));
    memcpy(g->smth_nlsf_q15, st->prev_nlsf_q15, order * sizeof(int16_t));
    g->fs_khz = st->fs_khz;
  }
  if (st->loss_count == 0 && st->prev_signal_type == SILK_TYPE_INACTIVE) {
    for (int i = 0; i < order; i++) {
      const int32_t cur = g->smth_nlsf_q15[i];
      g->smth_nlsf_q15[i] = (int16_t)(cur + smulwb((int32_t)st->prev_nlsf_q15[i] - cur, 16348));
    }
    int32_t max_gain = 0; int strongest = 0;
    for (int i = 0; i < st->nb_subfr; i++) if (c->gains_q16[i] > max_gain) { max_gain = c->gains_q16[i]; strongest = i; }
    memmove(g->exc_buf_q14 + st->subfr_length, g->exc_buf_q14, (size_t)st->subfr_length * (st->nb_subfr - 1) * sizeof(int32_t));
    memcpy(g->exc_buf_q14, st->exc_q14 + strongest * st->subfr_length, st->subfr_length * sizeof(int32_t));
    for (int i = 0; i < st->nb_subfr; i++) {
      g->smth_gain_q16 = w_add(g->smth_gain_q16, smulwb(w_sub(c->gains_q16[i], g->smth_gain_q16), 4634));
      if (smulww(g->smth_gain_q16, 46396) > c->gains_q16[i]) g->smth_gain_q16 = c->gains_q16[i];
    }
  }
  if (st->loss_count > 0) {
    int32_t sig[SILK_MAX_FRAME + SILK_MAX_ORDER] = {0};
    int32_t gain_q16 = smulww(st->plc.rand_scale_q14, st->plc.prev_gain_q16[1]);
    auto smultt = [](int32_t a, int32_t b) { return w_mul(a >> 16, b >> 16); };
    if (gain_q16 >= (1 << 21) || g->smth_gain_q16 > (1 << 23)) {
      gain_q16 = smultt(gain_q16, gain_q16);
      gain_q16 = w_sub(smultt(g->smth_gain_q16, g->smth_gain_q16), w_shl(gain_q16, 5));
      gain_q16 = lshift_sat32(sqrt_approx(gain_q16), 16);
    } else {
      gain_q16 = smulww(gain_q16, gain_q16);
      gain_q16 = w_sub(smulww(g->smth_gain_q16, g->smth_gain_q16), w_shl(gain_q16, 5));
      gain_q16 = lshift_sat32(sqrt_approx(gain_q16), 8);
    }
    const int32_t gain_q10 = gain_q16 >> 6;
    int mask = 255;
    while (mask > length) mask >>= 1;
    int32_t seed = g->rand_seed;
    for (int i = 0; i < length; i++) {
      seed = silk_rand(seed);
      int ix = (seed >> 24) & mask;
      if (ix < 0) ix = 0; if (ix > SILK_MAX_FRAME - 1) ix = SILK_MAX_FRAME - 1;
      sig[SILK_MAX_ORDER + i] = g->exc_buf_q14[ix];
    }
    g->rand_seed = seed;
    int16_t a_q12[SILK_MAX_ORDER] = {0};
    silk_nlsf2a(a_q12, g->smth_nlsf_q15, order, quirks);
    memcpy(sig, g->synth_state, sizeof(g->synth_state));
    for (int i = 0; i < length; i++) {
      int32_t pred = order >> 1;
      for (int j = 0; j < order; j++) pred = smlawb(pred, sig[SILK_MAX_ORDER + i - 1 - j], a_q12[j]);
      const int32_t upd = sat32_64((int64_t)sig[SILK_MAX_ORDER + i] + lshift_sat32(pred, 4));
      sig[SILK_MAX_ORDER + i] = upd;
      frame[i] = sat16((int32_t)frame[i] + sat16(rshift_round(smulww(upd, gain_q10), 8)));
    }
    memcpy(g->synth_state, sig + length, sizeof(g->synth_state));
  } else memset(g->synth_state, 0, order * sizeof(int32_t));
}

// ---- frame + API level ----
// decode_frame.rs:54-153
static void decode_frame(SilkDecoder* D, int n, EcDec* ec, int16_t* out, int lost_flag, int cond) {
  SilkChannel* st = &D->ch[n];
  SilkControl ctrl;
  int16_t pulses[SILK_MAX_FRAME + 16];
  int16_t nlsf_tr[SILK_MAX_ORDER] = {0};
  const int fi = st->n_frames_decoded;
  const int decode_payload = lost_flag == SILK_FLAG_DECODE_NORMAL || (lost_flag == SILK_FLAG_DECODE_LBRR && st->lbrr_flags[fi] != 0);
  if (decode_payload) {
    decode_indices(st, ec, fi, lost_flag == SILK_FLAG_DECODE_LBRR, cond, &st->indices);
    decode_pulses(ec, pulses, st->indices.signal_type, st->indices.quant_offset_type, st->frame_length, D->quirks);
    decode_parameters(st, &ctrl, cond, D->quirks, nlsf_tr);
    decode_core(st, &ctrl, out, pulses);
    silk_plc(st, &ctrl, out, 0);
    st->loss_count = 0;
    st->first_frame_after_reset = 0;
  } else {
    memset(pulses, 0, sizeof(pulses));
    silk_plc(st, &ctrl, out, 1);
  }
  int mv = st->ltp_mem_length - st->frame_length;  // refresh_output_buffer (decode_frame.rs:207-230)
  memmove(st->out_buf, st->out_buf + st->frame_length, mv * sizeof(int16_t));
  memcpy(st->out_buf + mv, out, st->frame_length * sizeof(int16_t));
  apply_cng(st, &ctrl, out, D->quirks);
  plc_glue_frames(st, out, st->frame_length);
  st->lag_prev = ctrl.pitch_l[st->nb_subfr - 1];
  if (D->trace) {
    D->trace->indices[n] = st->indices;
    memcpy(D->trace->pulses[n], pulses, sizeof(int16_t) * SILK_MAX_FRAME);
    memcpy(D->trace->nlsf_q15[n], nlsf_tr, sizeof(nlsf_tr));
    D->trace->ctrl[n] = ctrl;
    memcpy(D->trace->out[n], out, sizeof(int16_t) * st->frame_length);
  }
}

static void channel_reset(SilkChannel* c) {  // init_decoder.rs:14-26
  *c = SilkChannel();
  c->cng.reset(c->lpc_order);
  c->plc.reset(c->frame_length);
}

// dec_api.rs:102-123: reset_decoder keeps n_channels_api/internal, init_decoder sets both to 1
void SilkDecoder::reset(int init) {
  uint32_t q = quirks; SilkTrace* t = trace; int a = n_channels_api, b = n_channels_internal;
  *this = SilkDecoder();
  channel_reset(&ch[0]); channel_reset(&ch[1]);
  quirks = q; trace = t;
  if (!init) { n_channels_api = a; n_channels_internal = b; }
}

// stereo_ms_to_lr.rs:26-143
static void ms_to_lr(SilkDecoder* D, int16_t* x1, int16_t* x2, const int32_t pred_q13[2], int fs_khz, int frame_length) {
  memcpy(x1, D->s_mid, 4); memcpy(x2, D->s_side, 4);
  memcpy(D->s_mid, x1 + frame_length, 4); memcpy(D->s_side, x2 + frame_length, 4);
  int32_t p0 = D->pred_prev_q13[0], p1 = D->pred_prev_q13[1];
  int interp = 8 * fs_khz;
  int32_t denom = (1 << 16) / interp;
  int32_t d0 = rshift_round(smulbb(w_sub(pred_q13[0], p0), denom), 16), d1 = rshift_round(smulbb(w_sub(pred_q13[1], p1), denom), 16);
  int ni = interp < frame_length ? interp : frame_length;
  for (int n = 0; n < frame_length; n++) {
    if (n < ni) { p0 = w_add(p0, d0); p1 = w_add(p1, d1); }
    else { p0 = pred_q13[0]; p1 = pred_q13[1]; }
    int32_t sum = w_shl(w_add(w_add(x1[n], x1[n + 2]), w_shl(x1[n + 1], 1)), 9);
    sum = smlawb(w_shl(x2[n + 1], 8), sum, p0);
    sum = smlawb(sum, w_shl(x1[n + 1], 11), p1);
    x2[n + 1] = sat16(rshift_round(sum, 8));
  }
  D->pred_prev_q13[0] = sat16(pred_q13[0]); D->pred_prev_q13[1] = sat16(pred_q13[1]);
  for (int n = 0; n < frame_length; n++) {
    int32_t s = x1[n + 1] + x2[n + 1], d = x1[n + 1] - x2[n + 1];
    x1[n + 1] = sat16(s); x2[n + 1] = sat16(d);
  }
}

// dec_api.rs:132-441
int SilkDecoder::decode(SilkDecControl* c, int lost_flag, int new_packet, EcDec* ec, int16_t* samples_out) {
  int nci = c->n_channels_internal;
  if (new_packet) for (int n = 0; n < nci; n++) ch[n].n_frames_decoded = 0;
  if (nci > n_channels_internal) {
    if (quirks & QUIRK_S4_STEREO_REINIT) for (int n = 0; n < nci; n++) channel_reset(&ch[n]);
    else channel_reset(&ch[1]);
  }
  int stereo_to_mono = c->n_channels_internal == 1 && n_channels_internal == 2 && c->internal_sample_rate == 1000 * ch[0].fs_khz;
  if (ch[0].n_frames_decoded == 0) {
    for (int n = 0; n < nci; n++) {
      switch (c->payload_size_ms) {
        case 0: case 10: ch[n].n_frames_per_packet = 1; ch[n].nb_subfr = 2; break;
        case 20: ch[n].n_frames_per_packet = 1; ch[n].nb_subfr = 4; break;
        case 40: ch[n].n_frames_per_packet = 2; ch[n].nb_subfr = 4; break;
        case 60: ch[n].n_frames_per_packet = 3; ch[n].nb_subfr = 4; break;
        default: return -203;
      }
    }
  }
  for (int n = 0; n < nci; n++) {
    int fs_khz = (c->internal_sample_rate >> 10) + 1;
    if (ch[n].set_fs(fs_khz, c->api_sample_rate) < 0) return -200;
  }
  if (c->n_channels_api == 2 && c->n_channels_internal == 2 && (n_channels_api == 1 || n_channels_internal == 1)) {
    pred_prev_q13[0] = pred_prev_q13[1] = 0;
    s_side[0] = s_side[1] = 0;
    ch[1].resampler = ch[0].resampler;
  }
  n_channels_api = c->n_channels_api;
  n_channels_internal = c->n_channels_internal;
  int decode_only_middle = prev_decode_only_middle;
  uint32_t q = quirks;
  if (lost_flag != SILK_FLAG_PACKET_LOST && ch[0].n_frames_decoded == 0) {  // decode_vad_and_lbrr (dec_api.rs:443-530)
    for (int n = 0; n < nci; n++) {
      for (int i = 0; i < ch[n].n_frames_per_packet; i++) ch[n].vad_flags[i] = ec->bit_logp(1);
      ch[n].lbrr_flag = ec->bit_logp(1);
    }
    for (int n = 0; n < nci; n++) {
      memset(ch[n].lbrr_flags, 0, sizeof(ch[n].lbrr_flags));
      if (ch[n].lbrr_flag) {
        if (ch[n].n_frames_per_packet == 1) ch[n].lbrr_flags[0] = 1;
        else {
          int sym = ec->icdf(ch[n].n_frames_per_packet == 2 ? silk_LBRR_flags_2_iCDF : silk_LBRR_flags_3_iCDF, 8) + 1;
          for (int i = 0; i < ch[n].n_frames_per_packet; i++) ch[n].lbrr_flags[i] = (sym >> i) & 1;
        }
      }
    }
    // the LBRR payloads are only skipped over on the way to the regular frames (dec_api.rs:478-480)
    for (int i = 0; lost_flag == SILK_FLAG_DECODE_NORMAL && i < ch[0].n_frames_per_packet; i++) {
      for (int n = 0; n < nci; n++) {
        if (!ch[n].lbrr_flags[i]) continue;
        if (nci == 2 && n == 0) {
          int32_t skip[2];
          stereo_decode_pred(ec, skip);
          if (ch[1].lbrr_flags[i] == 0) decode_only_middle = ec->icdf(silk_stereo_only_code_mid_iCDF, 8);
        }
        int cond = (i > 0 && ch[n].lbrr_flags[i - 1]) ? SILK_CODE_CONDITIONALLY : SILK_CODE_INDEPENDENTLY;
        decode_indices(&ch[n], ec, i, 1, cond, &ch[n].indices);
        int16_t tmp[SILK_MAX_FRAME + 16];
        decode_pulses(ec, tmp, ch[n].indices.signal_type, ch[n].indices.quant_offset_type, ch[n].frame_length, q);
      }
    }
  }
  int32_t ms_pred_q13[2] = {0, 0};
  if (nci == 2) {
    const int fi = ch[0].n_frames_decoded;
    const int decode_ms = lost_flag == SILK_FLAG_DECODE_NORMAL || (lost_flag == SILK_FLAG_DECODE_LBRR && ch[0].lbrr_flags[fi] == 1);
    if (decode_ms) {
      stereo_decode_pred(ec, ms_pred_q13);
      const int need_mid_only = lost_flag == SILK_FLAG_DECODE_NORMAL ? ch[1].vad_flags[fi] == 0 : ch[1].lbrr_flags[fi] == 0;
      decode_only_middle = need_mid_only ? ec->icdf(silk_stereo_only_code_mid_iCDF, 8) : 0;
    } else { ms_pred_q13[0] = pred_prev_q13[0]; ms_pred_q13[1] = pred_prev_q13[1]; }
  } else decode_only_middle = 0;
  if (nci == 2 && !decode_only_middle && prev_decode_only_middle) {
    memset(ch[1].out_buf, 0, sizeof(ch[1].out_buf));
    memset(ch[1].s_lpc_q14, 0, sizeof(ch[1].s_lpc_q14));
    ch[1].lag_prev = 100; ch[1].last_gain_index = 10; ch[1].prev_signal_type = 0; ch[1].first_frame_after_reset = 1;
  }
  int frame_length = ch[0].frame_length;
  int16_t bufs[2][SILK_MAX_FRAME + 2];
  memset(bufs, 0, sizeof(bufs));
  int has_side;
  if (lost_flag == SILK_FLAG_DECODE_NORMAL) has_side = !decode_only_middle;
  else if (lost_flag == SILK_FLAG_PACKET_LOST) has_side = !prev_decode_only_middle;
  else has_side = !prev_decode_only_middle || (nci == 2 && ch[1].lbrr_flags[ch[1].n_frames_decoded] == 1);
  for (int n = 0; n < nci; n++) {
    if (n == 0 || has_side) {
      int fi = ch[0].n_frames_decoded - n, cond;   // conditional_coding_for_channel (dec_api.rs:560-583)
      if (fi <= 0) cond = SILK_CODE_INDEPENDENTLY;
      else if (lost_flag == SILK_FLAG_DECODE_LBRR) cond = ch[n].lbrr_flags[fi - 1] ? SILK_CODE_CONDITIONALLY : SILK_CODE_INDEPENDENTLY;
      else if (n > 0 && prev_decode_only_middle) cond = SILK_CODE_INDEPENDENTLY_NO_LTP_SCALING;
      else cond = SILK_CODE_CONDITIONALLY;
      decode_frame(this, n, ec, bufs[n] + 2, lost_flag, cond);
    } else memset(bufs[n] + 2, 0, frame_length * sizeof(int16_t));
    ch[n].n_frames_decoded++;
  }
  if (c->n_channels_api == 2 && nci == 2) ms_to_lr(this, bufs[0], bufs[1], ms_pred_q13, ch[0].fs_khz, frame_length);
  else {
    memcpy(bufs[0], s_mid, 4);
    memcpy(s_mid, bufs[0] + frame_length, 4);
  }
  int n_api = (int)((int64_t)frame_length * c->api_sample_rate / (ch[0].fs_khz * 1000));
  int16_t res[960];
  int active = c->n_channels_api < nci ? c->n_channels_api : nci;
  for (int n = 0; n < active; n++) {
    ch[n].resampler.run(res, bufs[n] + 1, frame_length);
    if (c->n_channels_api == 2) for (int i = 0; i < n_api; i++) samples_out[n + 2 * i] = res[i];
    else memcpy(samples_out, res, n_api * sizeof(int16_t));
  }
  if (c->n_channels_api == 2 && nci == 1) {
    if (stereo_to_mono) {
      ch[1].resampler.run(res, bufs[0] + 1, frame_length);
      for (int i = 0; i < n_api; i++) samples_out[1 + 2 * i] = res[i];
    } else for (int i = 0; i < n_api; i++) samples_out[1 + 2 * i] = samples_out[2 * i];
  }
  if (ch[0].prev_signal_type == SILK_TYPE_VOICED) { static const int mult[3] = {6, 4, 3}; c->prev_pitch_lag = ch[0].lag_prev * mult[(ch[0].fs_khz - 8) >> 2]; }
  else c->prev_pitch_lag = 0;
  if (lost_flag == SILK_FLAG_PACKET_LOST) { for (int n = 0; n < nci; n++) ch[n].last_gain_index = 10; }
  else prev_decode_only_middle = decode_only_middle;
  return n_api;
}

}  // namespace orc
