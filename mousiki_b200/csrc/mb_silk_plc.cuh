// SILK packet-loss concealment, comfort noise and the glue between a concealed and a decoded frame, for the per-stream
// lane of k_silk. Replaces (reference, src/silk/): plc.rs:46-457 (silk_plc, plc_update, plc_conceal, plc_glue_frames,
// silk_plc_energy, pick_rand_slice, sqrt_approx), cng.rs:62-373 (reset, apply_cng), sum_sqr_shift.rs:12-75, and the
// hooks of decode_frame.rs:117-151.
//
// Where the state lives: the small PLC / CNG fields travel with SilkChanState. The two 320-word excitation buffers of a
// channel (exc_Q14 of the last decoded frame, read by the concealment; the comfort-noise excitation FIFO) stay in HBM,
// structure-of-arrays over streams like the output history, and a decoded frame does not touch them while it runs: the
// excitation is a pure function of the frame's pulses (decode_core.rs:93-111) and the cell records outlive the pass, so the
// lane only notes where the data is (SilkWork::last_exc, ::cng_pend) and replays it into HBM when somebody is about to
// read it: in front of a concealed frame, and when the pass ends (the next call may open with a loss).
#pragma once
#include "mb_silk_core.cuh"

namespace mb {

// word k of buffer `kind` (0: exc_Q14, 1: CNG FIFO) of channel n of stream sid: base[((n * 2 + kind) * 320 + k) * n_streams + sid]
struct SilkExcRef { int32_t* base; size_t n_streams, sid; };
MB_DEV int32_t* silk_exc_buf(const SilkExcRef& X, int n, int kind) { return X.base + (size_t)((n * 2 + kind) * 320) * X.n_streams + X.sid; }
// A channel reset clears both buffers (init_decoder.rs:14-26); done on first use after the reset.
MB_DEV void silk_exc_touch(SilkChanState& c, const SilkExcRef& X, int n) {
  if (!c.exc_dirty) return;
  int32_t* p = silk_exc_buf(X, n, 0);
  for (int k = 0; k < 640; k++) p[(size_t)k * X.n_streams] = 0;
  c.exc_dirty = 0;
}

MB_DEV int32_t silk_rand(int32_t seed) { return s_add(907633515, (int32_t)((uint32_t)seed * 196314165u)); }   // cng.rs:371-373
MB_DEV int32_t s_lshift_sat32(int32_t v, int s) { return s_sat32((int64_t)v << s); }
MB_DEV int32_t s_rshift_any(int32_t v, int s) { return s <= 0 ? s_shl(v, -s) : v >> s; }                        // plc.rs:372-378
MB_DEV int32_t silk_sqrt_approx(int32_t x) {   // plc.rs:448-457
  if (x <= 0) return 0;
  const int lz = s_clz(x);
  const uint32_t ux = (uint32_t)x;
  const int rot = (24 - lz) & 31;
  const int32_t frac = (int32_t)(((ux >> rot) | (rot ? ux << (32 - rot) : 0u)) & 0x7fu);
  int32_t y = (lz & 1) ? 32768 : 46214;
  y >>= lz >> 1;
  return s_mlawb(y, y, s_mulbb(213, frac));
}

// sum_sqr_shift.rs:12-75 over x[0 .. len)
MB_DEV int32_t silk_energy_pass(const int16_t* x, int len, int32_t acc, int shift) {
  int i = 0;
  for (; i + 1 < len; i += 2) {
    const uint32_t pe = (uint32_t)((int32_t)x[i] * x[i]) + (uint32_t)((int32_t)x[i + 1] * x[i + 1]);
    acc = s_add(acc, (int32_t)(shift <= 0 ? pe : (shift >= 32 ? 0u : pe >> shift)));
  }
  if (i < len) {
    const uint32_t pe = (uint32_t)((int32_t)x[i] * x[i]);
    acc = s_add(acc, (int32_t)(shift <= 0 ? pe : (shift >= 32 ? 0u : pe >> shift)));
  }
  return acc;
}
MB_DEV_NOINLINE void silk_sum_sqr_shift(const int16_t* x, int len, int32_t* energy, int* shift_out) {
  if (len == 0) { *energy = 0; *shift_out = 0; return; }
  int shift = ilog32((uint32_t)len) - 1;
  const int32_t e = silk_energy_pass(x, len, len, shift);
  shift = imax(shift + 3 - s_clz(e), 0);
  *energy = silk_energy_pass(x, len, 0, shift);
  *shift_out = shift;
}

// The frame's excitation (decode_core.rs:93-111) for samples [from, to), written to dst[(i - from) * stride].
MB_DEV_NOINLINE void silk_exc_replay(const SilkExcSrc& src, int32_t* dst, size_t stride) {
  const uint32_t* pulse_words = src.pulse_words;   // stride 32
  const int from = src.from, to = src.from + src.len;
  const int32_t offset_q10 = silk_Quantization_Offsets_Q10[(src.signal_type >> 1) * 2 + src.quant_offset];
  int32_t seed = src.seed;
  for (int i = from; i < to; i++) {
    const uint32_t w = pulse_words[(i >> 1) * 32];
    const int32_t pulse = (int16_t)((i & 1) ? (w >> 16) : (w & 0xffffu));
    seed = silk_rand(seed);
    int32_t e = pulse << 14;
    if (e > 0) e -= 1280; else if (e < 0) e += 1280;
    e += offset_q10 << 4;
    if (seed < 0) e = -e;
    dst[(size_t)(i - from) * stride] = e;
    seed = s_add(seed, pulse);
  }
}

// ---- deferred writes of the two excitation buffers
MB_DEV void silk_work_drop(SilkWork& W, int n) { W.last_exc[n].valid = 0; W.cng_n[n] = 0; }   // channel reset: nothing to carry over
MB_DEV_NOINLINE void silk_exc_flush(SilkChanState& c, SilkWork& W, const SilkExcRef& X, int n) {
  if (W.last_exc[n].valid) {
    silk_exc_touch(c, X, n);
    silk_exc_replay(W.last_exc[n], silk_exc_buf(X, n, 0), X.n_streams);
    W.last_exc[n].valid = 0;
  }
}
// cng.rs:196-206: per push the FIFO moves up by one subframe and takes the frame's strongest subframe in front. The pending
// pushes all have one geometry (silk_cng_push), so they are applied at once: push j-th from last fills block j, what was
// there moves up by the number of pushes (nothing moves when they fill the whole FIFO).
MB_DEV_NOINLINE void silk_cng_flush(SilkChanState& c, SilkWork& W, const SilkExcRef& X, int n) {
  const int cnt = W.cng_n[n];
  if (cnt == 0) return;
  silk_exc_touch(c, X, n);
  int32_t* fifo = silk_exc_buf(X, n, 1);
  const size_t xs = X.n_streams;
  const int sub = W.cng_pend[n][0].sub, nb = W.cng_pend[n][0].nb, m = imin(cnt, nb);
  for (int i = sub * (nb - m) - 1; i >= 0; i--) fifo[(size_t)(i + m * sub) * xs] = fifo[(size_t)i * xs];
  for (int j = 0; j < m; j++) silk_exc_replay(W.cng_pend[n][cnt - 1 - j].src, fifo + (size_t)(j * sub) * xs, xs);
  W.cng_n[n] = 0;
}
MB_DEV void silk_cng_push(SilkChanState& c, SilkWork& W, const SilkExcRef& X, int n, const SilkCngPush& u) {
  int cnt = W.cng_n[n];
  if (cnt > 0 && (W.cng_pend[n][0].sub != u.sub || W.cng_pend[n][0].nb != u.nb)) { silk_cng_flush(c, W, X, n); cnt = 0; }
  if (cnt == 4) { for (int j = 0; j < 3; j++) W.cng_pend[n][j] = W.cng_pend[n][j + 1]; cnt = 3; }   // four pushes bury everything older
  W.cng_pend[n][cnt] = u;
  W.cng_n[n] = cnt + 1;
}

// silk_plc entry (plc.rs:46-72): the PLC state follows the channel's sampling rate
MB_DEV void silk_plc_sync_fs(SilkChanState& c) {
  if (c.fs_khz != c.plc_fs_khz) { silk_plc_reset(c, c.frame_length); c.plc_fs_khz = c.fs_khz; }
}

// plc.rs:117-184 (silk_plc_update), after a decoded frame
MB_DEV_NOINLINE void silk_plc_update(SilkChanState& c, const SilkSide& sd, const SilkFrameParams& P) {
  silk_plc_sync_fs(c);
  const int nb = c.nb_subfr, sub = 5 * c.fs_khz, order = c.fs_khz == 16 ? 16 : 10;
  c.prev_signal_type = sd.signal_type;
  int32_t ltp_gain_q14 = 0;
  if (sd.signal_type == SILK_T_VOICED) {
    const int32_t target = P.pitch_l[nb - 1];
    for (int j = 0; j < nb && j * sub < target; j++) {
      const int sf = nb - 1 - j;
      int32_t tg = 0;
      for (int i = 0; i < 5; i++) tg += P.b_q14[sf * 5 + i];
      if (tg > ltp_gain_q14) { ltp_gain_q14 = tg; c.plc_pitch_l_q8 = s_shl(P.pitch_l[sf], 8); }
    }
    for (int i = 0; i < 5; i++) c.plc_ltp_coef_q14[i] = 0;
    c.plc_ltp_coef_q14[2] = (int16_t)s_sat16(ltp_gain_q14);
    if (ltp_gain_q14 < 11469) {
      const int32_t scale_q10 = (11469 << 10) / imax(ltp_gain_q14, 1);
      for (int i = 0; i < 5; i++) c.plc_ltp_coef_q14[i] = (int16_t)s_sat16(s_mulbb(c.plc_ltp_coef_q14[i], scale_q10) >> 10);
    } else if (ltp_gain_q14 > 15565) {
      const int32_t scale_q14 = (15565 << 14) / imax(ltp_gain_q14, 1);
      for (int i = 0; i < 5; i++) c.plc_ltp_coef_q14[i] = (int16_t)s_sat16(s_mulbb(c.plc_ltp_coef_q14[i], scale_q14) >> 14);
    }
  } else {
    c.plc_pitch_l_q8 = s_shl(s_mulbb(c.fs_khz, 18), 8);
    for (int i = 0; i < 5; i++) c.plc_ltp_coef_q14[i] = 0;
  }
  for (int i = 0; i < 16; i++) c.plc_prev_lpc_q12[i] = i < order ? P.a_q12[1][i] : (int16_t)0;
  c.plc_prev_ltp_scale_q14 = (int16_t)s_sat16(P.ltp_scale_q14);
  c.plc_prev_gain_q16[0] = P.gains_q16[nb - 2];
  c.plc_prev_gain_q16[1] = P.gains_q16[nb - 1];
  c.plc_subfr_length = (int16_t)sub;
  c.plc_nb_subfr = (int16_t)nb;
}

// plc.rs:186-345 (silk_plc_conceal): one concealed frame of the channel, written behind the history in the ring like a
// decoded frame. sltp: the lane's 640-word scratch. Returns the lag the frame ends on (decoder_control pitch_l).
MB_DEV_NOINLINE int silk_plc_conceal(SilkChanState& c, SilkWork& W, const SilkExcRef& X, int n, int16_t* __restrict__ h, int pos,
                                     int32_t* __restrict__ sltp) {
  silk_plc_sync_fs(c);
  silk_exc_flush(c, W, X, n);
  silk_exc_touch(c, X, n);
  const int fs = c.fs_khz, nb = c.nb_subfr, sub = 5 * fs, L = sub * nb, ltp_mem = 20 * fs, order = fs == 16 ? 16 : 10;
  const int32_t* exc = silk_exc_buf(X, n, 0);
  const size_t xs = X.n_streams;
  const int32_t pg_q10[2] = {c.plc_prev_gain_q16[0] >> 6, c.plc_prev_gain_q16[1] >> 6};
  // silk_plc_energy (plc.rs:347-368): the last two subframes of the stored excitation at the last two gains
  int32_t e1, e2; int s1, s2;
  {
    int16_t buf[80];
    for (int k = 0; k < 2; k++) {
      const int base = (k + nb - 2) * sub;
      for (int i = 0; i < sub; i++) {
        const int idx = base + i;
        const int32_t v = idx < 320 ? exc[(size_t)idx * xs] : 0;
        buf[i] = (int16_t)s_sat16(s_mulww(v, pg_q10[k]) >> 8);
      }
      if (k == 0) silk_sum_sqr_shift(buf, sub, &e1, &s1); else silk_sum_sqr_shift(buf, sub, &e2, &s2);
    }
  }
  // pick_rand_slice (plc.rs:370-388): 128 excitation samples, positioned by the PLC state's own subframe geometry
  int rand_start;
  {
    const int second = !(s_rshift_any(e1, s2) < s_rshift_any(e2, s1));
    const int sl = imax((int)c.plc_subfr_length, 1), pnb = imax((int)c.plc_nb_subfr, 2);
    rand_start = imin(imax((second ? pnb * sl : (pnb - 1) * sl) - 128, 0), 320 - 128);
  }
  if (c.first_frame_after_reset) for (int i = 0; i < 16; i++) c.plc_prev_lpc_q12[i] = 0;
  int32_t rand_scale_q14 = c.plc_rand_scale_q14;
  const int att = c.loss_count < 1 ? c.loss_count : 1;
  const int32_t harm_gain_q15 = att ? 31130 : 32440;
  const int32_t rand_gain_q15 = c.prev_signal_type == SILK_T_VOICED ? (att ? 26214 : 31130) : (att ? 29491 : 32440);
  silk_bwexpander16(c.plc_prev_lpc_q12, order, 64881);
  int32_t A[16];                                   // pre-shifted like the decoder core's taps
  for (int j = 0; j < 16; j++) A[j] = j < order ? s_shl(c.plc_prev_lpc_q12[j], 16) : 0;
  if (c.loss_count == 0) {
    rand_scale_q14 = 1 << 14;
    if (c.prev_signal_type == SILK_T_VOICED) {
      for (int i = 0; i < 5; i++) rand_scale_q14 -= c.plc_ltp_coef_q14[i];
      rand_scale_q14 = imax(rand_scale_q14, 3277);
      rand_scale_q14 = s_mulbb(rand_scale_q14, c.plc_prev_ltp_scale_q14) >> 14;
    } else {
      const int32_t inv_gain_q30 = silk_inv_pred_gain(c.plc_prev_lpc_q12, order);
      int32_t down = imin((1 << 30) >> 3, inv_gain_q30);
      down = imax(down, (1 << 30) >> 8);
      down = s_shl(down, 3);
      rand_scale_q14 = s_mulwb(down, rand_gain_q15) >> 14;
    }
  }
  int32_t rand_seed = c.plc_rand_seed;
  int lag = s_rr(c.plc_pitch_l_q8, 8);
  for (int i = 0; i < 640; i++) sltp[i] = 0;
  {
    // re-whitened history in the gain-normalised domain: samples [idx + order, ltp_mem)
    const int idx = ltp_mem - lag - order - 2;
    int32_t inv_gain_q30 = silk_inverse32(c.plc_prev_gain_q16[1], 46);
    inv_gain_q30 = imin(inv_gain_q30, 2147483647 >> 1);
    for (int ix = idx + order; ix < ltp_mem; ix++) {
      if (ix < 0) continue;
      int32_t acc = 0;
      for (int k = 0; k < order; k++) {
        const int j = ix - 1 - k;
        const int32_t v = j >= 0 ? (int32_t)h[(pos + j) & SILK_RING_MASK] : 0;
        acc = s_add(acc, v * (A[k] >> 16));
      }
      const int32_t r = s_sat16(s_rr(s_sub(s_shl(h[(pos + ix) & SILK_RING_MASK], 12), acc), 12));
      sltp[ix] = s_mulwb(inv_gain_q30, r);
    }
  }
  int16_t B[5];
  for (int t = 0; t < 5; t++) B[t] = c.plc_ltp_coef_q14[t];
  int buf_idx = ltp_mem;
  const int32_t* rs = exc + (size_t)rand_start * xs;
  for (int k = 0; k < nb; k++) {
    for (int i = 0; i < sub; i++) {
      int32_t pred = 2;
      for (int t = 0; t < 5; t++) {
        const int j = buf_idx - lag + t - 2;
        pred = s_mlawb(pred, (j >= 0 && j < 640) ? sltp[j] : 0, B[t]);
      }
      rand_seed = silk_rand(rand_seed);
      sltp[buf_idx] = s_lshift_sat32(s_mlawb(pred, rs[(size_t)((rand_seed >> 25) & 127) * xs], rand_scale_q14), 2);
      buf_idx++;
    }
    for (int t = 0; t < 5; t++) B[t] = (int16_t)s_sat16(s_mulbb(harm_gain_q15, B[t]) >> 15);
    rand_scale_q14 = s_mulbb(rand_scale_q14, rand_gain_q15) >> 15;
    c.plc_pitch_l_q8 = s_add(c.plc_pitch_l_q8, s_mulwb(c.plc_pitch_l_q8, 655));
    c.plc_pitch_l_q8 = imin(c.plc_pitch_l_q8, s_shl(s_mulbb(18, fs), 8));
    lag = s_rr(c.plc_pitch_l_q8, 8);
  }
  for (int t = 0; t < 5; t++) c.plc_ltp_coef_q14[t] = B[t];
  // short-term synthesis over the generated excitation, in place
  int32_t lpc[16];
  for (int j = 0; j < 16; j++) lpc[j] = c.s_lpc_q14[15 - j];
  for (int i = 0; i < L; i++) {
    int32_t pred = order >> 1;
    for (int j = 0; j < 16; j++) pred = s_add(pred, s_mulhi(lpc[j], A[j]));
    const int32_t upd = s_sat32((int64_t)sltp[ltp_mem + i] + s_lshift_sat32(pred, 4));
    for (int j = 15; j > 0; j--) lpc[j] = lpc[j - 1];
    lpc[0] = upd;
    h[(pos + ltp_mem + i) & SILK_RING_MASK] = (int16_t)s_sat16(s_rr(s_mulww(upd, pg_q10[1]), 8));
  }
  for (int j = 0; j < 16; j++) c.s_lpc_q14[15 - j] = lpc[j];
  c.plc_rand_seed = rand_seed;
  c.plc_rand_scale_q14 = (int16_t)s_sat16(rand_scale_q14);
  c.loss_count = c.loss_count == 2147483647 ? 2147483647 : c.loss_count + 1;
  return lag;
}

// plc.rs:74-115 (silk_plc_glue_frames) on the frame that leaves the channel (after comfort noise)
MB_DEV_NOINLINE void silk_plc_glue(SilkChanState& c, int16_t* frame, int length) {
  if (c.loss_count > 0) {
    int sh;
    silk_sum_sqr_shift(frame, length, &c.plc_conc_energy, &sh);
    c.plc_conc_energy_shift = (int16_t)sh;
    c.plc_last_frame_lost = 1;
    return;
  }
  if (!c.plc_last_frame_lost) return;
  int32_t energy; int eshift;
  silk_sum_sqr_shift(frame, length, &energy, &eshift);
  if (eshift > c.plc_conc_energy_shift) c.plc_conc_energy >>= eshift - c.plc_conc_energy_shift;
  else if (eshift < c.plc_conc_energy_shift) energy >>= c.plc_conc_energy_shift - eshift;
  if (energy > c.plc_conc_energy) {
    const int lz = s_clz(imax(c.plc_conc_energy, 1)) - 1;
    c.plc_conc_energy = s_shl(c.plc_conc_energy, lz);
    const int rs = imax(24 - lz, 0);
    if (rs > 0) energy >>= rs;
    const int32_t frac_q24 = c.plc_conc_energy / imax(energy, 1);
    int32_t gain_q16 = s_shl(silk_sqrt_approx(frac_q24), 4);
    int32_t slope_q16 = ((1 << 16) - gain_q16) / length;
    slope_q16 = s_shl(slope_q16, 2);
    for (int i = 0; i < length; i++) {
      frame[i] = (int16_t)s_sat16(s_mulwb(gain_q16, frame[i]));
      gain_q16 = s_sat32((int64_t)gain_q16 + slope_q16);
      if (gain_q16 > (1 << 16)) break;
    }
  }
  c.plc_last_frame_lost = 0;
}

// cng.rs:145-233, the part that runs on every frame: rate change, and the estimator update on a decoded inactive frame.
// pulse_words / sd: the frame's record (null for a concealed frame); gains: this frame's subframe gains.
MB_DEV_NOINLINE void silk_cng_track(SilkChanState& c, SilkWork& W, const SilkExcRef& X, int n, const uint32_t* pulse_words, const SilkSide* sd,
                                    const int32_t* gains_q16, const int32_t* seed_at) {
  const int order = c.fs_khz == 16 ? 16 : 10, nb = c.nb_subfr, sub = 5 * c.fs_khz;
  if (c.fs_khz != c.cng_fs_khz) {
    if (c.cng_fs_khz == -1) { for (int i = 0; i < order; i++) c.cng_synth_state[i] = 0; }
    else silk_cng_reset(c, order);
    for (int i = 0; i < 16; i++) c.cng_smth_nlsf_q15[i] = i < order ? c.prev_nlsf_q15[i] : (int16_t)0;
    c.cng_fs_khz = c.fs_khz;
  }
  if (c.loss_count == 0 && c.prev_signal_type == SILK_T_INACTIVE && pulse_words) {
    for (int i = 0; i < order; i++) {
      const int32_t cur = c.cng_smth_nlsf_q15[i];
      c.cng_smth_nlsf_q15[i] = (int16_t)(cur + s_mulwb((int32_t)c.prev_nlsf_q15[i] - cur, 16348));
    }
    int32_t max_gain = 0; int strongest = 0;
    for (int i = 0; i < nb; i++) if (gains_q16[i] > max_gain) { max_gain = gains_q16[i]; strongest = i; }
    SilkCngPush u;
    u.src.pulse_words = pulse_words; u.src.from = (int16_t)(strongest * sub); u.src.len = (int16_t)sub;
    u.src.seed = seed_at[strongest]; u.src.signal_type = sd->signal_type; u.src.quant_offset = sd->quant_offset; u.src.valid = 1;
    u.sub = (int16_t)sub; u.nb = (int16_t)nb;
    silk_cng_push(c, W, X, n, u);
    for (int i = 0; i < nb; i++) {
      c.cng_smth_gain_q16 = s_add(c.cng_smth_gain_q16, s_mulwb(s_sub(gains_q16[i], c.cng_smth_gain_q16), 4634));
      if (s_mulww(c.cng_smth_gain_q16, 46396) > gains_q16[i]) c.cng_smth_gain_q16 = gains_q16[i];
    }
  }
  if (c.loss_count == 0) for (int i = 0; i < order; i++) c.cng_synth_state[i] = 0;
}

// cng.rs:235-370: comfort noise added to a concealed frame. sig: the lane's 640-word scratch.
MB_DEV_NOINLINE void silk_cng_add(SilkChanState& c, SilkWork& W, const SilkExcRef& X, int n, int16_t* frame, int32_t* __restrict__ sig,
                                  uint32_t quirks) {
  const int order = c.fs_khz == 16 ? 16 : 10, length = c.frame_length;
  silk_cng_flush(c, W, X, n);
  silk_exc_touch(c, X, n);
  const int32_t* fifo = silk_exc_buf(X, n, 1);
  const size_t xs = X.n_streams;
  int32_t gain_q16 = s_mulww(c.plc_rand_scale_q14, c.plc_prev_gain_q16[1]);
  if (gain_q16 >= (1 << 21) || c.cng_smth_gain_q16 > (1 << 23)) {
    gain_q16 = (int32_t)((uint32_t)(gain_q16 >> 16) * (uint32_t)(gain_q16 >> 16));
    gain_q16 = s_sub((int32_t)((uint32_t)(c.cng_smth_gain_q16 >> 16) * (uint32_t)(c.cng_smth_gain_q16 >> 16)), s_shl(gain_q16, 5));
    gain_q16 = s_lshift_sat32(silk_sqrt_approx(gain_q16), 16);
  } else {
    gain_q16 = s_mulww(gain_q16, gain_q16);
    gain_q16 = s_sub(s_mulww(c.cng_smth_gain_q16, c.cng_smth_gain_q16), s_shl(gain_q16, 5));
    gain_q16 = s_lshift_sat32(silk_sqrt_approx(gain_q16), 8);
  }
  const int32_t gain_q10 = gain_q16 >> 6;
  int mask = 255;
  while (mask > length) mask >>= 1;
  int32_t seed = c.cng_rand_seed;
  for (int i = 0; i < length; i++) {
    seed = silk_rand(seed);
    const int ix = imin(imax((seed >> 24) & mask, 0), 319);
    sig[16 + i] = fifo[(size_t)ix * xs];
  }
  c.cng_rand_seed = seed;
  int16_t a_q12[16];
  for (int i = 0; i < 16; i++) a_q12[i] = 0;
  silk_nlsf2a(a_q12, c.cng_smth_nlsf_q15, order, quirks);
  int32_t A[16], lpc[16];
  for (int j = 0; j < 16; j++) { A[j] = j < order ? s_shl(a_q12[j], 16) : 0; lpc[j] = c.cng_synth_state[15 - j]; }
  for (int i = 0; i < length; i++) {
    int32_t pred = order >> 1;
    for (int j = 0; j < 16; j++) pred = s_add(pred, s_mulhi(lpc[j], A[j]));
    const int32_t upd = s_sat32((int64_t)sig[16 + i] + s_lshift_sat32(pred, 4));
    for (int j = 15; j > 0; j--) lpc[j] = lpc[j - 1];
    lpc[0] = upd;
    frame[i] = (int16_t)s_sat16((int32_t)frame[i] + s_sat16(s_rr(s_mulww(upd, gain_q10), 8)));
  }
  for (int j = 0; j < 16; j++) c.cng_synth_state[15 - j] = lpc[j];
}

}  // namespace mb
