// Phase 2 for SILK-only packets: parameter reconstruction + synthesis, one lane per stream, its packets in tick order.
//
// Replaces (reference, src/silk/): decode_parameters.rs:79-199 (gain_quant.rs:78-98, nlsf_decode.rs:27-107,
// nlsf_stabilize.rs:20-105, nlsf2a.rs:25-117, lpc_fit.rs, lpc_inv_pred_gain.rs, bwexpander*.rs, decode_pitch.rs),
// decode_core.rs:28-330 (excitation, LTP re-whitening through lpc_analysis_filter.rs, order-10/16 LPC synthesis),
// decode_frame.rs:54-153 + decoder_set_fs.rs:137-198 (state machine), stereo_ms_to_lr.rs:26-143, the decoder-side
// resampler (resampler.rs:279-369, resampler_private_iir_fir.rs, resampler_private_up2_hq.rs) and the channel plumbing
// of silk_decode (dec_api.rs:132-441), then silk_int16_to_res (opus_decoder.rs:1524-1533).
//
// The work is integer and strictly sequential in time inside one stream (LPC/LTP recursion, IIR resampler), so the
// parallelism is the batch: a lane owns a stream, keeps the LTP history, the scaled LTP state and the resampler window
// in local memory (lane-interleaved, so a warp's accesses at equal indices coalesce), the 16 LPC taps and the
// synthesis history of the running subframe in registers. Concealment and comfort noise: mb_silk_plc.cuh.
#pragma once
#include "mb_silk_parse.cuh"

namespace mb {

// ---- fixed point helpers with the reference's widths (wrapping 32-bit unless said otherwise)
MB_DEV int32_t s_add(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
MB_DEV int32_t s_sub(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }
MB_DEV int32_t s_shl(int32_t a, int s) { return (int32_t)((uint32_t)a << s); }
MB_DEV int32_t s_mulwb(int32_t a, int32_t b) { return (int32_t)(((int64_t)a * (int16_t)b) >> 16); }
MB_DEV int32_t s_mulww(int32_t a, int32_t b) { return (int32_t)(((int64_t)a * b) >> 16); }
MB_DEV int32_t s_mlawb(int32_t acc, int32_t a, int32_t b) { return s_add(acc, s_mulwb(a, b)); }
MB_DEV int32_t s_mulbb(int32_t a, int32_t b) { return (int32_t)(int16_t)a * (int32_t)(int16_t)b; }
MB_DEV int32_t s_mmul(int32_t a, int32_t b) { return (int32_t)(((int64_t)a * b) >> 32); }
MB_DEV int32_t s_rr(int32_t v, int s) { return s == 1 ? (v >> 1) + (v & 1) : ((v >> (s - 1)) + 1) >> 1; }
MB_DEV int64_t s_rr64(int64_t v, int s) { return s == 1 ? (v >> 1) + (v & 1) : ((v >> (s - 1)) + 1) >> 1; }
MB_DEV int32_t s_sat16(int32_t v) { return v > 32767 ? 32767 : (v < -32768 ? -32768 : v); }
MB_DEV int32_t s_sat32(int64_t v) { return (int32_t)(v > 2147483647LL ? 2147483647LL : (v < -2147483647LL - 1 ? -2147483647LL - 1 : v)); }
MB_DEV int s_clz(int32_t v) { return 32 - ilog32((uint32_t)v); }
MB_DEV int32_t s_abs(int32_t v) { return v < 0 ? (int32_t)(0u - (uint32_t)v) : v; }

// silk_INVERSE32_varQ (lpc_inv_pred_gain.rs:128-150)
MB_DEV int32_t silk_inverse32(int32_t b32, int qres) {
  if (b32 == 0 || qres <= 0) return 0;
  const int hr = s_clz(imax(s_abs(b32), 1)) - 1;
  const int32_t nrm = s_shl(b32, hr);
  const int32_t inv = 536870911 / (nrm >> 16);
  int32_t res = s_shl(inv, 16);
  const int32_t err = s_shl(s_sub(1 << 29, (int32_t)(((int64_t)nrm * inv) >> 16)), 3);
  res = s_add(res, (int32_t)(((int64_t)err * inv) >> 16));
  const int ls = 61 - hr - qres;
  if (ls <= 0) return s_sat32((int64_t)res << -ls);
  return ls < 32 ? res >> ls : 0;
}
// silk_DIV32_varQ (stereo_find_predictor.rs:134-166)
MB_DEV int32_t silk_div32(int32_t a32, int32_t b32, int qres) {
  const int32_t aa = a32 == (-2147483647 - 1) ? 2147483647 : s_abs(a32), ab = b32 == (-2147483647 - 1) ? 2147483647 : s_abs(b32);
  const int ah = s_clz(aa) - 1, bh = s_clz(ab) - 1;
  int32_t an = s_shl(a32, ah);
  const int32_t bn = s_shl(b32, bh);
  const int32_t binv = 536870911 / (bn >> 16);
  int32_t res = s_mulwb(an, binv);
  an = s_sub(an, s_shl(s_mmul(bn, res), 3));
  res = s_mlawb(res, an, binv);
  const int ls = 29 + ah - bh - qres;
  if (ls < 0) return s_sat32((int64_t)res << -ls);
  return ls < 32 ? res >> ls : 0;
}
// log2lin.rs:14-35
MB_DEV int32_t silk_log2lin(int32_t x) {
  if (x < 0) return 0;
  if (x >= 3967) return 2147483647;
  int32_t out = 1 << (x >> 7);
  const int32_t fr = x & 0x7f;
  const int32_t corr = fr + (int32_t)(((int64_t)(fr * (128 - fr)) * -174) >> 16);
  if (x < 2048) out += (int32_t)(((int64_t)out * corr) >> 7);
  else out += (int32_t)((int64_t)(out >> 7) * corr);
  return out;
}

// ---- persistent per-stream state (HBM). Reference: silk Decoder (dec_api.rs:35-46), DecoderState
// (decoder_state.rs:99-131), DecoderSampleRateState (decoder_set_fs.rs:63-124), Resampler (resampler.rs:40-48).
// The output history (the last 20 ms per channel, decoder_set_fs.rs out_buf) lives in its own structure-of-arrays
// buffer, see silk_work_load / silk_work_store in mb_silk_out.cuh.
struct SilkChanState {
  int32_t s_lpc_q14[16];
  int16_t prev_nlsf_q15[16];
  int32_t res_iir[6];            // up2_hq all-pass state of the API resampler
  int16_t res_fir[8];            // FIR interpolator tail
  int16_t res_delay[16];         // input delay line (fs_khz samples)
  int32_t res_down[24];          // down-sampling kernel's FIR tail (output rate below the internal rate; AR2 state in res_iir[0..2))
  int32_t prev_gain_q16;
  int16_t fs_khz, nb_subfr, frame_length, lag_prev;
  int8_t first_frame_after_reset, last_gain_index, prev_signal_type, res_ready;
  // packet-loss concealment (PLC state: decoder_state.rs:23-95, plc.rs) and comfort noise (cng.rs:31-80). The two excitation
  // buffers the reference keeps beside them (exc_Q14 of the last decoded frame, the CNG excitation FIFO) live in their own
  // structure-of-arrays buffer in HBM, see SilkExcRef in mb_silk_plc.cuh.
  int32_t loss_count;
  int32_t plc_pitch_l_q8, plc_rand_seed, plc_conc_energy, plc_prev_gain_q16[2];
  int16_t plc_ltp_coef_q14[5], plc_prev_lpc_q12[16], plc_rand_scale_q14, plc_prev_ltp_scale_q14;
  int16_t plc_fs_khz, plc_nb_subfr, plc_subfr_length, plc_conc_energy_shift;
  int32_t cng_synth_state[16], cng_smth_gain_q16, cng_rand_seed;
  int16_t cng_smth_nlsf_q15[16], cng_fs_khz;
  int8_t plc_last_frame_lost, exc_dirty;   // exc_dirty: the channel's excitation buffers in HBM must read as zero (cleared lazily)
};
struct SilkState {
  SilkChanState ch[2];
  int16_t pred_prev_q13[2], s_mid[2], s_side[2];
  int8_t n_channels_api, n_channels_internal, prev_decode_only_middle, pad;
};
MB_DEV void silk_plc_reset(SilkChanState& c, int frame_length) {   // decoder_state.rs:77-95 (fs_khz is the caller's)
  c.plc_pitch_l_q8 = frame_length << 7; c.plc_rand_seed = 0; c.plc_conc_energy = 0; c.plc_prev_gain_q16[0] = c.plc_prev_gain_q16[1] = 65536;
  for (int i = 0; i < 5; i++) c.plc_ltp_coef_q14[i] = 0;
  for (int i = 0; i < 16; i++) c.plc_prev_lpc_q12[i] = 0;
  c.plc_rand_scale_q14 = 0; c.plc_prev_ltp_scale_q14 = 0; c.plc_nb_subfr = 2; c.plc_subfr_length = 20; c.plc_conc_energy_shift = 0;
  c.plc_last_frame_lost = 0;
}
MB_DEV void silk_cng_reset(SilkChanState& c, int order) {        // cng.rs:62-80 (the excitation FIFO is left alone)
  const int32_t step = 32767 / (order + 1);
  int32_t acc = 0;
  for (int i = 0; i < 16; i++) {
    if (i < order) { acc += step; c.cng_smth_nlsf_q15[i] = (int16_t)acc; } else c.cng_smth_nlsf_q15[i] = 0;
    c.cng_synth_state[i] = 0;
  }
  c.cng_smth_gain_q16 = 0; c.cng_rand_seed = 3176576;
}
MB_DEV void silk_chan_reset(SilkChanState& c) {   // init_decoder.rs:14-26 + DecoderState::default
  for (int i = 0; i < 16; i++) { c.s_lpc_q14[i] = 0; c.prev_nlsf_q15[i] = 0; c.res_delay[i] = 0; }
  for (int i = 0; i < 6; i++) c.res_iir[i] = 0;
  for (int i = 0; i < 8; i++) c.res_fir[i] = 0;
  for (int i = 0; i < 24; i++) c.res_down[i] = 0;
  c.prev_gain_q16 = 65536;
  c.fs_khz = 0; c.nb_subfr = 4; c.frame_length = 0; c.lag_prev = 0;
  c.first_frame_after_reset = 1; c.last_gain_index = 0; c.prev_signal_type = 0; c.res_ready = 0;
  c.loss_count = 0;
  silk_plc_reset(c, 0);
  c.plc_fs_khz = 0;
  silk_cng_reset(c, 16);
  c.cng_fs_khz = -1;
  c.exc_dirty = 1;
}
MB_DEV void silk_state_reset(SilkState& s, int init) {   // dec_api.rs:102-123
  silk_chan_reset(s.ch[0]); silk_chan_reset(s.ch[1]);
  s.pred_prev_q13[0] = s.pred_prev_q13[1] = 0; s.s_mid[0] = s.s_mid[1] = 0; s.s_side[0] = s.s_side[1] = 0;
  s.prev_decode_only_middle = 0; s.pad = 0;
  if (init) { s.n_channels_api = 1; s.n_channels_internal = 1; }
}

// ---- NLSF -> LPC
// nlsf_stabilize.rs:20-105
MB_DEV_NOINLINE void silk_nlsf_stabilize(int16_t* x, const int16_t* dmin, int L) {
  for (int loops = 0; loops < 20; loops++) {
    int32_t best = x[0] - dmin[0];
    int I = 0;
    for (int i = 1; i < L; i++) {
      const int32_t d = x[i] - (x[i - 1] + dmin[i]);
      if (d < best) { best = d; I = i; }
    }
    const int32_t dl = (1 << 15) - (x[L - 1] + dmin[L]);
    if (dl < best) { best = dl; I = L; }
    if (best >= 0) return;
    if (I == 0) x[0] = dmin[0];
    else if (I == L) x[L - 1] = (int16_t)s_sat16((1 << 15) - dmin[L]);
    else {
      int32_t lo = 0, hi = 1 << 15;
      for (int k = 0; k < I; k++) lo += dmin[k];
      lo += dmin[I] >> 1;
      for (int k = L; k > I; k--) hi -= dmin[k];
      hi -= dmin[I] >> 1;
      int32_t c = ((int32_t)x[I - 1] + x[I] + 1) >> 1;
      c = c < lo ? lo : (c > hi ? hi : c);
      x[I - 1] = (int16_t)s_sat16(c - (dmin[I] >> 1));
      x[I] = (int16_t)s_sat16(x[I - 1] + dmin[I]);
    }
  }
  for (int i = 1; i < L; i++) {
    const int16_t v = x[i];
    int j = i - 1;
    while (j >= 0 && v < x[j]) { x[j + 1] = x[j]; j--; }
    x[j + 1] = v;
  }
  if (x[0] < dmin[0]) x[0] = dmin[0];
  for (int i = 1; i < L; i++) { const int16_t m = (int16_t)s_sat16((int32_t)x[i - 1] + dmin[i]); if (x[i] < m) x[i] = m; }
  { const int16_t m = (int16_t)s_sat16((1 << 15) - dmin[L]); if (x[L - 1] > m) x[L - 1] = m; }
  for (int i = L - 2; i >= 0; i--) { const int16_t m = (int16_t)s_sat16((int32_t)x[i + 1] - dmin[i + 1]); if (x[i] > m) x[i] = m; }
}

// nlsf_decode.rs:27-107 (+ nlsf_unpack.rs:19-52 for the predictor select)
MB_DEV_NOINLINE void silk_nlsf_decode(int16_t* nlsf, const int8_t* idx, bool wb, uint32_t quirks) {
  const int order = wb ? 16 : 10;
  const uint8_t* sel = (wb ? silk_NLSF_CB2_SELECT_WB : silk_NLSF_CB2_SELECT_NB_MB) + idx[0] * (order >> 1);
  const uint8_t* predtab = wb ? silk_NLSF_PRED_WB_Q8 : silk_NLSF_PRED_NB_MB_Q8;
  const int32_t step = wb ? 9830 : 11796;
  int16_t res[16];
  int32_t out = 0;
  for (int i = order - 1; i >= 0; i--) {
    const uint32_t e = sel[i >> 1];
    const int pick = (i & 1) ? (int)((e >> 4) & 1u) : (int)(e & 1u);
    const int32_t pred = predtab[i + pick * (order - 1)];
    const int32_t prod = (quirks & S1_NLSF_SMULBB16) ? (int32_t)(int16_t)((int16_t)out * (int16_t)pred) : s_mulbb(out, pred);
    int32_t q = s_shl(idx[i + 1], 10);
    if (q > 0) q -= 102; else if (q < 0) q += 102;
    out = s_mlawb(prod >> 8, q, step);
    res[i] = (int16_t)s_sat16(out);
  }
  const uint8_t* cb1 = (wb ? silk_NLSF_CB1_WB_Q8 : silk_NLSF_CB1_NB_MB_Q8) + idx[0] * order;
  const int16_t* wgt = (wb ? silk_NLSF_CB1_WB_Wght_Q9 : silk_NLSF_CB1_NB_MB_Wght_Q9) + idx[0] * order;
  for (int i = 0; i < order; i++) {
    const int32_t v = s_add(s_shl(res[i], 14) / (int32_t)wgt[i], s_shl((int32_t)cb1[i], 7));
    nlsf[i] = (int16_t)(v < 0 ? 0 : (v > 32767 ? 32767 : v));
  }
  silk_nlsf_stabilize(nlsf, wb ? silk_NLSF_DELTA_MIN_WB_Q15 : silk_NLSF_DELTA_MIN_NB_MB_Q15, order);
}

// bwexpander_32.rs:11-25
MB_DEV void silk_bwexpander32(int32_t* ar, int d, int32_t chirp) {
  const int32_t cm1 = s_sub(chirp, 1 << 16);
  for (int i = 0; i < d - 1; i++) {
    ar[i] = s_mulww(chirp, ar[i]);
    chirp = s_add(chirp, (int32_t)s_rr64((int64_t)chirp * cm1, 16));
  }
  ar[d - 1] = s_mulww(chirp, ar[d - 1]);
}

// bwexpander.rs:11-27 (16-bit coefficients)
MB_DEV void silk_bwexpander16(int16_t* ar, int d, int32_t chirp) {
  const int32_t cm1 = s_sub(chirp, 1 << 16);
  for (int i = 0; i < d - 1; i++) {
    ar[i] = (int16_t)s_rr64((int64_t)chirp * ar[i], 16);
    chirp = s_add(chirp, (int32_t)s_rr64((int64_t)chirp * cm1, 16));
  }
  ar[d - 1] = (int16_t)s_rr64((int64_t)chirp * ar[d - 1], 16);
}

// lpc_inv_pred_gain.rs:22-104
MB_DEV_NOINLINE int32_t silk_inv_pred_gain(const int16_t* a_q12, int order) {
  const int32_t A_LIMIT = 16773022, MIN_INV = 107374;   // 0.99975 in Q24, 1e-4 in Q30
  int32_t a[16];
  int32_t dc = 0;
  for (int k = 0; k < order; k++) { dc += a_q12[k]; a[k] = (int32_t)a_q12[k] << 12; }
  if (dc >= 4096) return 0;
  int32_t inv_gain = 1 << 30;
  for (int k = order - 1; k > 0; k--) {
    if (a[k] > A_LIMIT || a[k] < -A_LIMIT) return 0;
    const int32_t rc = -s_shl(a[k], 7);
    const int32_t m1 = (1 << 30) - s_mmul(rc, rc);
    inv_gain = s_shl(s_mmul(inv_gain, m1), 2);
    if (inv_gain < MIN_INV) return 0;
    const int mq = 32 - s_clz(s_abs(m1));
    const int32_t m2 = silk_inverse32(m1, mq + 30);
    for (int n = 0; n < (k + 1) >> 1; n++) {
      const int32_t t1 = a[n], t2 = a[k - n - 1];
      int64_t u = s_rr64((int64_t)s_sat32((int64_t)t1 - (int32_t)s_rr64((int64_t)t2 * rc, 31)) * m2, mq);
      if (u > 2147483647LL || u < -2147483648LL) return 0;
      a[n] = (int32_t)u;
      u = s_rr64((int64_t)s_sat32((int64_t)t2 - (int32_t)s_rr64((int64_t)t1 * rc, 31)) * m2, mq);
      if (u > 2147483647LL || u < -2147483648LL) return 0;
      a[k - n - 1] = (int32_t)u;
    }
  }
  if (a[0] > A_LIMIT || a[0] < -A_LIMIT) return 0;
  const int32_t rc = -s_shl(a[0], 7);
  const int32_t m1 = (1 << 30) - s_mmul(rc, rc);
  inv_gain = s_shl(s_mmul(inv_gain, m1), 2);
  return inv_gain < MIN_INV ? 0 : inv_gain;
}

// nlsf2a.rs:25-139 (+ lpc_fit.rs:21-68)
MB_DEV void silk_find_poly(int32_t* out, const int32_t* c, int dd) {
  out[0] = 1 << 16;
  out[1] = -c[0];
  for (int k = 1; k < dd; k++) {
    const int32_t f = c[2 * k];
    out[k + 1] = s_sub(s_shl(out[k - 1], 1), (int32_t)s_rr64((int64_t)f * out[k], 16));
    for (int n = k; n > 1; n--) out[n] = s_sub(s_add(out[n], out[n - 2]), (int32_t)s_rr64((int64_t)f * out[n - 1], 16));
    out[1] = s_sub(out[1], f);
  }
}
MB_TBL uint8_t kSilkOrd16[16] = {0, 15, 8, 7, 4, 11, 12, 3, 2, 13, 10, 5, 6, 9, 14, 1};
MB_TBL uint8_t kSilkOrd10[10] = {0, 9, 6, 3, 4, 5, 8, 1, 2, 7};
MB_DEV_NOINLINE void silk_nlsf2a(int16_t* a_q12, const int16_t* nlsf, int d, uint32_t quirks) {
  const uint8_t* ordering = d == 16 ? kSilkOrd16 : kSilkOrd10;
  int32_t cosq[16];
  for (int k = 0; k < d; k++) {
    const int dst = ordering[k];
    const int32_t v = nlsf[(quirks & S2_NLSF2A_ORDER) ? dst : k];
    const int fi = v >> 8;
    const int32_t ff = v - (fi << 8);
    const int32_t c0 = silk_LSFCosTab_FIX_Q12[fi], dl = (int16_t)(silk_LSFCosTab_FIX_Q12[fi + 1] - silk_LSFCosTab_FIX_Q12[fi]);
    cosq[dst] = (int32_t)s_rr64(((int64_t)c0 << 8) + (int64_t)dl * ff, 4);
  }
  const int dd = d >> 1;
  int32_t P[9], Q[9], a32[16];
  silk_find_poly(P, cosq, dd);
  silk_find_poly(Q, cosq + 1, dd);
  for (int k = 0; k < dd; k++) {
    const int32_t pt = s_add(P[k + 1], P[k]), qt = s_sub(Q[k + 1], Q[k]);
    a32[k] = s_sub(s_sub(0, qt), pt);
    a32[d - k - 1] = s_sub(qt, pt);
  }
  {  // lpc_fit(a_q12, a32, QOUT = 12, QIN = 17)
    bool clipped = true;
    for (int it = 0; it < 10; it++) {
      int32_t maxabs = 0;
      int idx = 0;
      for (int k = 0; k < d; k++) { const int32_t v = s_abs(a32[k]); if (v > maxabs) { maxabs = v; idx = k; } }
      int32_t m = s_rr(maxabs, 5);
      if (m <= 32767) { clipped = false; break; }
      if (m > 163838) m = 163838;
      const int32_t den = imax((m * (idx + 1)) >> 2, 1);
      silk_bwexpander32(a32, d, 65470 - ((m - 32767) << 14) / den);
    }
    if (clipped) {
      for (int k = 0; k < d; k++) { a_q12[k] = (int16_t)s_sat16(s_rr(a32[k], 5)); a32[k] = s_shl(a_q12[k], 5); }
    } else {
      for (int k = 0; k < d; k++) a_q12[k] = (int16_t)s_rr(a32[k], 5);
    }
  }
  for (int i = 0; i < 16; i++) {
    if (silk_inv_pred_gain(a_q12, d) != 0) return;
    silk_bwexpander32(a32, d, 65536 - (2 << i));
    for (int k = 0; k < d; k++) a_q12[k] = (int16_t)s_rr(a32[k], 5);
  }
}

// ---- one SILK frame of one channel --------------------------------------------------------------------------------
struct SilkFrameParams {     // decoder_control.rs
  int32_t gains_q16[4], pitch_l[4];
  int16_t a_q12[2][16], b_q14[20];
  int32_t ltp_scale_q14;
  int32_t seed_at[4];        // excitation generator state at the start of each subframe (written by the synthesis core)
};

// decode_parameters.rs:79-199
MB_DEV_NOINLINE void silk_frame_params(SilkChanState& st, SilkSide& sd, SilkFrameParams& P, uint32_t quirks) {
  const int nb = st.nb_subfr, fs = st.fs_khz;
  const bool wb = fs == 16;
  const int order = wb ? 16 : 10;
  {  // gain_quant.rs:78-98
    int32_t prev = st.last_gain_index;
    for (int k = 0; k < 4; k++) {
      if (k >= nb) { P.gains_q16[k] = 0; continue; }
      if (k == 0 && sd.cond != SILK_COND_COND) prev = imax(prev - 16, (int)sd.gains[0]);
      else {
        const int32_t t = sd.gains[k] - 4, thr = 2 * 36 - 64 + prev;
        prev += t > thr ? (t << 1) - thr : t;
      }
      prev = prev < 0 ? 0 : (prev > 63 ? 63 : prev);
      const int32_t lg = s_mulwb(1907825, prev) + 2090;
      P.gains_q16[k] = silk_log2lin(lg < 3967 ? lg : 3967);
    }
    st.last_gain_index = (int8_t)prev;
  }
  int16_t nlsf[16];
  for (int i = 0; i < 16; i++) { nlsf[i] = 0; P.a_q12[0][i] = 0; P.a_q12[1][i] = 0; }
  silk_nlsf_decode(nlsf, sd.nlsf, wb, quirks);
  silk_nlsf2a(P.a_q12[1], nlsf, order, quirks);
  if (st.first_frame_after_reset) sd.interp_q2 = 4;
  if (sd.interp_q2 < 4) {
    int16_t n0[16];
    for (int i = 0; i < order; i++)
      n0[i] = (int16_t)s_sat16(st.prev_nlsf_q15[i] + ((sd.interp_q2 * ((int32_t)nlsf[i] - st.prev_nlsf_q15[i])) >> 2));
    silk_nlsf2a(P.a_q12[0], n0, order, quirks);
  } else {
    for (int i = 0; i < order; i++) P.a_q12[0][i] = P.a_q12[1][i];
  }
  for (int i = 0; i < 16; i++) st.prev_nlsf_q15[i] = i < order ? nlsf[i] : 0;
  if (st.loss_count > 0) { silk_bwexpander16(P.a_q12[0], order, 63570); silk_bwexpander16(P.a_q12[1], order, 63570); }   // decode_parameters.rs:156-160
  for (int k = 0; k < 4; k++) P.pitch_l[k] = 0;
  for (int k = 0; k < 20; k++) P.b_q14[k] = 0;
  P.ltp_scale_q14 = 0;
  if (sd.signal_type == SILK_T_VOICED) {
    const int min_lag = 2 * fs, max_lag = 18 * fs, lag = min_lag + sd.lag_index;   // decode_pitch.rs:22-50
    for (int k = 0; k < nb; k++) {
      int off;
      if (fs == 8) off = nb == 4 ? silk_CB_lags_stage2[k * 11 + sd.contour] : silk_CB_lags_stage2_10_ms[k * 3 + sd.contour];
      else off = nb == 4 ? silk_CB_lags_stage3[k * 34 + sd.contour] : silk_CB_lags_stage3_10_ms[k * 12 + sd.contour];
      const int v = lag + off;
      P.pitch_l[k] = v < min_lag ? min_lag : (v > max_lag ? max_lag : v);
    }
    const int8_t* vq = sd.per_index == 0 ? silk_LTP_gain_vq_0 : (sd.per_index == 1 ? silk_LTP_gain_vq_1 : silk_LTP_gain_vq_2);
    for (int k = 0; k < nb; k++)
      for (int i = 0; i < 5; i++) P.b_q14[k * 5 + i] = (int16_t)(vq[sd.ltp[k] * 5 + i] << 7);
    P.ltp_scale_q14 = silk_LTPScales_table_Q14[sd.ltp_scale];
  } else sd.per_index = 0;
}

}  // namespace mb
