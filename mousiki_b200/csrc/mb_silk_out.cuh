// Phase 2 for SILK-only packets, output side: channel plumbing of silk_decode, mid/side -> left/right, the API
// resampler to 48 kHz and the PCM store; one call per packet from the per-stream lane loop (k_silk).
// Reference: src/silk/dec_api.rs:132-441, stereo_ms_to_lr.rs:26-143, resampler.rs:87-176,279-369,
// resampler_private_iir_fir.rs:80-197, resampler_private_up2_hq.rs:31-117, decoder_set_fs.rs:137-198,
// opus_decoder.rs:596-745 (SILK arm of decode_frame), :1524-1533 (silk_int16_to_res).
//
// The resampler runs as a stream: every input sample goes through the 2x all-pass up-sampler (state in registers) into
// a 10-sample register window, and the FIR interpolator emits the 3 / 4 / 6 output samples whose 8-tap window has just
// become complete. The reference's batch structure (1 ms from the delay line, then 10 ms batches, each restarting the
// Q16 phase accumulator) is kept exactly; what disappears is the 2x up-sampled scratch array. Output samples are
// collected 32 at a time in a lane-private shared-memory row and leave as 128-bit stores (a full 128-byte line per
// lane for mono float output).
#pragma once
#include "mb_celt_tail.cuh"
#include "mb_silk_plc.cuh"
#include "mb_celt_energy.cuh"

namespace mb {

// decoder_set_fs.rs:137-198 for a fixed 48 kHz API rate. Returns true when the output history was cleared.
MB_DEV bool silk_set_fs(SilkChanState& c, int fs_khz, int nb_subfr) {
  bool cleared = false;
  if (c.fs_khz != fs_khz || !c.res_ready) {   // resampler init (always the IIR/FIR up-sampler here: 8/12/16 -> 48 kHz)
    for (int i = 0; i < 6; i++) c.res_iir[i] = 0;
    for (int i = 0; i < 8; i++) c.res_fir[i] = 0;
    for (int i = 0; i < 16; i++) c.res_delay[i] = 0;
    for (int i = 0; i < 24; i++) c.res_down[i] = 0;
    c.res_ready = 1;
  }
  if (c.fs_khz != fs_khz) {
    c.first_frame_after_reset = 1; c.lag_prev = 100; c.last_gain_index = 10; c.prev_signal_type = 0;
    for (int i = 0; i < 16; i++) c.s_lpc_q14[i] = 0;
    cleared = true;
  }
  c.fs_khz = (int16_t)fs_khz; c.nb_subfr = (int16_t)nb_subfr; c.frame_length = (int16_t)(5 * fs_khz * nb_subfr);
  return cleared;
}
MB_DEV void silk_ring_clear(int16_t* h, int pos) {
  for (int i = 0; i < 320; i++) h[(pos + i) & SILK_RING_MASK] = 0;
}

// ---- PCM sink: 32 samples staged per lane, then written out
struct SilkPcmOut {
  float* f32;      // interleaved packet row, or null
  int16_t* i16;    // interleaved packet row, or null
  int CC;          // API channels (interleave stride)
  float gain; int apply_gain;
  int gain_from;   // first sample position (per channel) the gain applies to: 0, or 120 when k_deemph finishes the first
                   // 2.5 ms of the row (CELT fade-out after a hybrid packet) and applies the gain there itself
  int fs_out_khz = 48;   // the decoder's output rate (row F4): 48, or 24 / 16 / 12 / 8 for SILK-only streams
};
struct SilkSink {
  int32_t* stage;  // 32 words, lane-private (shared memory on the device)
  int fill, pos, ch, dup;   // pos: sample position (per channel) of stage[0]; dup: mono stream into a stereo row
  int limit;                // samples per channel the row takes; what lies beyond is dropped
  const SilkPcmOut* o;
};
MB_DEV float silk_to_float(const SilkPcmOut& o, int32_t v, int pos) {
  float x = (float)v * (1.0f / 32768.0f);
  if (o.apply_gain && pos >= o.gain_from) x *= o.gain;      // opus_decoder.rs:1376-1392
  return x;
}
MB_DEV_NOINLINE void silk_sink_flush(SilkSink& s) {
  const SilkPcmOut& o = *s.o;
  const int n = imin(s.fill, imax(s.limit - s.pos, 0));
  if (o.CC == 1 || s.dup) {
    const int rep = s.dup ? 2 : 1;
    if (o.f32) {
      float* dst = o.f32 + (size_t)s.pos * rep;
#if MB_ON_DEVICE
      if (n == 32 && (((uintptr_t)dst) & 15u) == 0) {
        if (rep == 1) {
#pragma unroll
          for (int q = 0; q < 8; q++)
            ((float4*)dst)[q] = make_float4(silk_to_float(o, s.stage[4 * q], s.pos + (4 * q)), silk_to_float(o, s.stage[4 * q + 1], s.pos + (4 * q + 1)),
                                            silk_to_float(o, s.stage[4 * q + 2], s.pos + (4 * q + 2)), silk_to_float(o, s.stage[4 * q + 3], s.pos + (4 * q + 3)));
        } else {
#pragma unroll
          for (int q = 0; q < 16; q++) {
            const float a = silk_to_float(o, s.stage[2 * q], s.pos + (2 * q)), b = silk_to_float(o, s.stage[2 * q + 1], s.pos + (2 * q + 1));
            ((float4*)dst)[q] = make_float4(a, a, b, b);
          }
        }
      } else
#endif
        for (int i = 0; i < n; i++) { const float x = silk_to_float(o, s.stage[i], s.pos + (i)); for (int r = 0; r < rep; r++) dst[i * rep + r] = x; }
    } else {
      int16_t* dst = o.i16 + (size_t)s.pos * rep;
#if MB_ON_DEVICE
      if (n == 32 && (((uintptr_t)dst) & 15u) == 0) {
        if (rep == 1) {
#pragma unroll
          for (int q = 0; q < 4; q++) {
            uint32_t w[4];
#pragma unroll
            for (int j = 0; j < 4; j++) w[j] = (uint32_t)(uint16_t)s.stage[8 * q + 2 * j] | ((uint32_t)(uint16_t)s.stage[8 * q + 2 * j + 1] << 16);
            ((uint4*)dst)[q] = make_uint4(w[0], w[1], w[2], w[3]);
          }
        } else {
#pragma unroll
          for (int q = 0; q < 8; q++) {
            uint32_t w[4];
#pragma unroll
            for (int j = 0; j < 4; j++) { const uint32_t v = (uint16_t)s.stage[4 * q + j]; w[j] = v | (v << 16); }
            ((uint4*)dst)[q] = make_uint4(w[0], w[1], w[2], w[3]);
          }
        }
      } else
#endif
        for (int i = 0; i < n; i++) for (int r = 0; r < rep; r++) dst[i * rep + r] = (int16_t)s.stage[i];
    }
  } else {  // one channel of an interleaved stereo row
    if (o.f32) for (int i = 0; i < n; i++) o.f32[(size_t)(s.pos + i) * 2 + s.ch] = silk_to_float(o, s.stage[i], s.pos + (i));
    else for (int i = 0; i < n; i++) o.i16[(size_t)(s.pos + i) * 2 + s.ch] = (int16_t)s.stage[i];
  }
  s.pos += s.fill; s.fill = 0;
}
MB_DEV void silk_sink_put(SilkSink& s, int32_t v) {
  s.stage[s.fill++] = v;
  if (s.fill == 32) silk_sink_flush(s);
}

// input of the resampler: sample k of the frame, read from a ring (mono path: one sample of delay through s_mid) or
// from the left/right staging rows
struct SilkSrc {
  const int16_t* p; int base, mask, first, has_first;
  MB_MEM int32_t at(int k) const { return (has_first && k == 0) ? first : (int32_t)p[(base + k) & mask]; }
};

// resampler.rs:279-369 + resampler_private_iir_fir.rs:80-197 + resampler_private_up2_hq.rs:31-117, streamed.
// `fir`: the 12 x 4 interpolation table (shared-memory copy on the device). Returns the number of samples produced.
// Output rates other than 48 kHz (row F4; resampler.rs:87-176): the same delayed input stream feeds a plain copy when the
// output rate equals the internal one, the 2x all-pass up-sampler alone when it is twice that, else the up-sampler + FIR
// interpolator with the ratio's own step. Output rates BELOW the internal rate (the down-FIR kernels) are refused at the parser.
MB_DEV_NOINLINE int silk_resample48(SilkChanState& c, const SilkSrc& in, int n_in, const int16_t* fir, SilkSink& sink) {
  const int fs = c.fs_khz, fs_out = sink.o->fs_out_khz;
  int delay = fs == 8 ? 0 : (fs == 12 ? 4 : 7);                 // DELAY_MATRIX_DEC[.][48 kHz] (resampler.rs:228)
  int32_t inc = fs == 8 ? 21846 : (fs == 12 ? 32768 : 43691);   // compute_inv_ratio_q16(fs, 48000, up2x = 1)
  int32_t S0 = c.res_iir[0], S1 = c.res_iir[1], S2 = c.res_iir[2], S3 = c.res_iir[3], S4 = c.res_iir[4], S5 = c.res_iir[5];
  const int32_t c00 = s_shl(silk_resampler_up2_hq_0[0], 16), c01 = s_shl(silk_resampler_up2_hq_0[1], 16), c02 = s_shl(silk_resampler_up2_hq_0[2], 16);
  const int32_t c10 = s_shl(silk_resampler_up2_hq_1[0], 16), c11 = s_shl(silk_resampler_up2_hq_1[1], 16), c12 = s_shl(silk_resampler_up2_hq_1[2], 16);
  if (fs_out != 48) {
    const int ii = fs == 8 ? 0 : (fs == 12 ? 1 : 2), oi = fs_out == 8 ? 0 : (fs_out == 12 ? 1 : (fs_out == 16 ? 2 : 3));
    delay = (int)silk_delay_matrix_dec[ii * 5 + oi];
    if (fs_out == fs || fs_out == 2 * fs) {
      const bool up2 = fs_out != fs;
      int made = 0;
      for (int t = 0; t < n_in; t++) {
        const int32_t x = t < delay ? (int32_t)c.res_delay[t] : in.at(t - delay);
        if (!up2) { silk_sink_put(sink, x); made++; continue; }
        const int32_t in32 = x << 10;
        int32_t Y, X, o1, o2;
        Y = s_sub(in32, S0); X = s_mulhi(Y, c00); o1 = s_add(S0, X); S0 = s_add(in32, X);
        Y = s_sub(o1, S1); X = s_mulhi(Y, c01); o2 = s_add(S1, X); S1 = s_add(o1, X);
        Y = s_sub(o2, S2); X = s_add(Y, s_mulhi(Y, c02)); o1 = s_add(S2, X); S2 = s_add(o2, X);
        silk_sink_put(sink, s_sat16(s_rr(o1, 10)));
        Y = s_sub(in32, S3); X = s_mulhi(Y, c10); o1 = s_add(S3, X); S3 = s_add(in32, X);
        Y = s_sub(o1, S4); X = s_mulhi(Y, c11); o2 = s_add(S4, X); S4 = s_add(o1, X);
        Y = s_sub(o2, S5); X = s_add(Y, s_mulhi(Y, c12)); o1 = s_add(S5, X); S5 = s_add(o2, X);
        silk_sink_put(sink, s_sat16(s_rr(o1, 10)));
        made += 2;
      }
      c.res_iir[0] = S0; c.res_iir[1] = S1; c.res_iir[2] = S2; c.res_iir[3] = S3; c.res_iir[4] = S4; c.res_iir[5] = S5;
      int16_t keep[16];
      for (int i = 0; i < delay; i++) keep[i] = (int16_t)in.at(n_in - delay + i);
      for (int i = 0; i < delay; i++) c.res_delay[i] = keep[i];
      return made;
    }
    if (fs_out < fs) {
      // Output rate below the internal rate (resampler_private_down_fir.rs:139-284, resampler_private_ar2.rs:19-38): a second-order
      // AR filter (Q8), then a polyphase FIR stepped by the ratio; batches of 1 ms (the delay buffer), then 10 ms, each with
      // the previous batch's last `order` filtered samples in front. 16 -> 12 kHz and 12 -> 8 kHz: 18 taps, 3 / 2 phases; 16 -> 8: 24.
      const int16_t* coefs = fs_out * 4 == fs * 3 ? silk_Resampler_3_4_COEFS : (fs_out * 3 == fs * 2 ? silk_Resampler_2_3_COEFS : silk_Resampler_1_2_COEFS);
      const int fracs = fs_out * 4 == fs * 3 ? 3 : (fs_out * 3 == fs * 2 ? 2 : 1);
      const int order = fs_out * 2 == fs ? 24 : 18, half = order / 2;
      const int16_t* firc = coefs + 2;
      const int32_t a0 = coefs[0], a1 = coefs[1];
      int64_t invd = ((int64_t)fs << 16) / fs_out;                       // compute_inv_ratio_q16, no 2x stage
      while (((invd * fs_out) >> 16) < (int64_t)fs) invd++;
      const int32_t incd = (int32_t)invd;
      int32_t buf[160 + 24];
      for (int i = 0; i < order; i++) buf[i] = c.res_down[i];
      int32_t A0 = c.res_iir[0], A1 = c.res_iir[1];
      int made = 0, t = 0, n = fs;
      while (t < n_in) {
        if (n > n_in - t) n = n_in - t;
        for (int k = 0; k < n; k++, t++) {
          const int32_t x = t < delay ? (int32_t)c.res_delay[t] : in.at(t - delay);
          int32_t o32 = s_add(A0, s_shl(x, 8));
          buf[order + k] = o32;
          o32 = s_shl(o32, 2);
          A0 = s_mlawb(A1, o32, a0);
          A1 = s_mulwb(o32, a1);
        }
        const int32_t max_idx = n << 16;
        for (int32_t idx = 0; idx < max_idx; idx = s_add(idx, incd)) {
          const int32_t* b = buf + (idx >> 16);
          int32_t acc;
          if (order == 18) {
            const int ind = s_mulwb(idx & 0xFFFF, fracs);
            const int16_t* f = firc + half * ind;
            const int16_t* r = firc + half * (fracs - 1 - ind);
            acc = s_mulwb(b[0], f[0]);
            for (int k = 1; k < half; k++) acc = s_mlawb(acc, b[k], f[k]);
            for (int k = 0; k < half; k++) acc = s_mlawb(acc, b[order - 1 - k], r[k]);
          } else {
            acc = s_mulwb(s_add(b[0], b[order - 1]), firc[0]);
            for (int k = 1; k < half; k++) acc = s_mlawb(acc, s_add(b[k], b[order - 1 - k]), firc[k]);
          }
          silk_sink_put(sink, s_sat16(s_add(acc, 1 << 5) >> 6));
          made++;
        }
        for (int i = 0; i < order; i++) buf[i] = buf[n + i];
        n = 10 * fs;
      }
      for (int i = 0; i < order; i++) c.res_down[i] = buf[i];
      c.res_iir[0] = A0; c.res_iir[1] = A1;
      int16_t keep[16];
      for (int i = 0; i < delay; i++) keep[i] = (int16_t)in.at(n_in - delay + i);
      for (int i = 0; i < delay; i++) c.res_delay[i] = keep[i];
      return made;
    }
    // compute_inv_ratio_q16 (resampler.rs:178-200) with the 2x stage in front
    int64_t inv = ((int64_t)fs << 17) / fs_out;
    while (((inv * fs_out) >> 16) < ((int64_t)fs << 1)) inv++;
    inc = (int32_t)inv;
  }
  int32_t w[10];                     // the newest 10 up-sampled values, w[9] newest
#pragma unroll
  for (int j = 0; j < 8; j++) w[2 + j] = c.res_fir[j];
  w[0] = w[1] = 0;
  int produced = 0;
  int t = 0;                         // position in the delayed input stream: res_delay[0, delay) ++ in[0, n_in - delay)
  int n = fs;                        // first batch: one millisecond
  int32_t x_next = delay > 0 ? (int32_t)c.res_delay[0] : in.at(0);
  while (t < n_in) {
    if (n > n_in - t) n = n_in - t;
    const int32_t max_idx = n << 17;
    int32_t idx = 0;
    for (int k = -1; k < n; k++) {
      if (k >= 0) {
        const int32_t x = x_next;    // fetched one sample ahead: the load is off the filter's dependency chain
        t++;
        if (t < n_in) x_next = t < delay ? (int32_t)c.res_delay[t] : in.at(t - delay);
        const int32_t in32 = x << 10;
        int32_t Y, X, o1, o2;
        Y = s_sub(in32, S0); X = s_mulhi(Y, c00); o1 = s_add(S0, X); S0 = s_add(in32, X);
        Y = s_sub(o1, S1); X = s_mulhi(Y, c01); o2 = s_add(S1, X); S1 = s_add(o1, X);
        Y = s_sub(o2, S2); X = s_add(Y, s_mulhi(Y, c02)); o1 = s_add(S2, X); S2 = s_add(o2, X);
        const int32_t u0 = s_sat16(s_rr(o1, 10));
        Y = s_sub(in32, S3); X = s_mulhi(Y, c10); o1 = s_add(S3, X); S3 = s_add(in32, X);
        Y = s_sub(o1, S4); X = s_mulhi(Y, c11); o2 = s_add(S4, X); S4 = s_add(o1, X);
        Y = s_sub(o2, S5); X = s_add(Y, s_mulhi(Y, c12)); o1 = s_add(S5, X); S5 = s_add(o2, X);
        const int32_t u1 = s_sat16(s_rr(o1, 10));
#pragma unroll
        for (int j = 0; j < 8; j++) w[j] = w[j + 2];
        w[8] = u0; w[9] = u1;
      }
      // window position of w[j] inside the batch buffer (tail = 0..7, sample k -> 8 + 2k, 9 + 2k): 2k + j. An output with
      // integer position b reads buffer[b .. b + 8): complete once b + 7 <= 2k + 9, i.e. b <= 2k + 2; b = 2k + 1 starts
      // at w[1], b = 2k + 2 at w[2].
      while (idx < max_idx && (idx >> 16) <= 2 * k + 2) {
        const bool hi = (idx >> 16) == 2 * k + 2;
        const int ti = ((idx & 0xffff) * 12) >> 16;
        const int16_t* f = fir + 4 * ti;
        const int16_t* r = fir + 4 * (11 - ti);
        int32_t acc = (hi ? w[2] : w[1]) * f[0];
        acc += (hi ? w[3] : w[2]) * f[1]; acc += (hi ? w[4] : w[3]) * f[2]; acc += (hi ? w[5] : w[4]) * f[3];
        acc += (hi ? w[6] : w[5]) * r[3]; acc += (hi ? w[7] : w[6]) * r[2]; acc += (hi ? w[8] : w[7]) * r[1];
        acc += (hi ? w[9] : w[8]) * r[0];
        silk_sink_put(sink, s_sat16((acc + (1 << 14)) >> 15));
        produced++;
        idx += inc;
      }
    }
    n = 10 * fs;
  }
#pragma unroll
  for (int j = 0; j < 8; j++) c.res_fir[j] = (int16_t)w[2 + j];
  c.res_iir[0] = S0; c.res_iir[1] = S1; c.res_iir[2] = S2; c.res_iir[3] = S3; c.res_iir[4] = S4; c.res_iir[5] = S5;
  for (int i = 0; i < delay; i++) c.res_delay[i] = (int16_t)in.at(n_in - delay + i);
  return produced;
}

// stereo_ms_to_lr.rs:26-143; x1/x2: 2 history samples + frame
MB_DEV_NOINLINE void silk_ms_to_lr(SilkState& S, int16_t* x1, int16_t* x2, const int32_t pred_q13[2], int fs_khz, int n) {
  x1[0] = S.s_mid[0]; x1[1] = S.s_mid[1]; x2[0] = S.s_side[0]; x2[1] = S.s_side[1];
  S.s_mid[0] = x1[n]; S.s_mid[1] = x1[n + 1]; S.s_side[0] = x2[n]; S.s_side[1] = x2[n + 1];
  int32_t p0 = S.pred_prev_q13[0], p1 = S.pred_prev_q13[1];
  const int interp = 8 * fs_khz;
  const int32_t denom = (1 << 16) / interp;
  const int32_t d0 = s_rr(s_mulbb(s_sub(pred_q13[0], p0), denom), 16), d1 = s_rr(s_mulbb(s_sub(pred_q13[1], p1), denom), 16);
  for (int i = 0; i < n; i++) {
    if (i < interp) { p0 = s_add(p0, d0); p1 = s_add(p1, d1); }
    else { p0 = pred_q13[0]; p1 = pred_q13[1]; }
    int32_t sum = s_shl(s_add(s_add(x1[i], x1[i + 2]), s_shl(x1[i + 1], 1)), 9);
    sum = s_mlawb(s_shl(x2[i + 1], 8), sum, p0);
    sum = s_mlawb(sum, s_shl(x1[i + 1], 11), p1);
    x2[i + 1] = (int16_t)s_sat16(s_rr(sum, 8));
  }
  S.pred_prev_q13[0] = (int16_t)s_sat16(pred_q13[0]); S.pred_prev_q13[1] = (int16_t)s_sat16(pred_q13[1]);
  for (int i = 0; i < n; i++) {
    const int32_t a = x1[i + 1], b = x2[i + 1];
    x1[i + 1] = (int16_t)s_sat16(a + b); x2[i + 1] = (int16_t)s_sat16(a - b);
  }
}

// Output history in HBM: 2 channels x 160 words (2 samples each) per stream, structure-of-arrays over streams —
// word k of channel n of stream sid at hist[(n * 160 + k) * n_streams + sid] — so the lanes of a warp move it with
// coalesced 128-byte transactions. While a lane runs, its rings are the live copy; the side channel is only loaded
// (and stored back) when a packet actually uses it.
struct SilkHistRef { uint32_t* base; size_t n_streams, sid; };
MB_DEV void silk_hist_load(SilkWork& W, const SilkHistRef& H, int n) {
  W.pos[n] = 0;
  const uint32_t* g = H.base + (size_t)n * 160 * H.n_streams + H.sid;
  for (int k = 0; k < 160; k++) {
    const uint32_t v = g[(size_t)k * H.n_streams];
    W.hist[n][2 * k] = (int16_t)(v & 0xffffu); W.hist[n][2 * k + 1] = (int16_t)(v >> 16);
  }
}
MB_DEV void silk_hist_store(const SilkWork& W, const SilkHistRef& H, int n) {
  uint32_t* g = H.base + (size_t)n * 160 * H.n_streams + H.sid;
  const int p = W.pos[n];
  for (int k = 0; k < 160; k++)
    g[(size_t)k * H.n_streams] = (uint32_t)(uint16_t)W.hist[n][(p + 2 * k) & SILK_RING_MASK] |
                                 ((uint32_t)(uint16_t)W.hist[n][(p + 2 * k + 1) & SILK_RING_MASK] << 16);
}

// What kind of packet the lane is looking at (the lost_flag of silk_decode, dec_api.rs:132-150)
enum { SILK_PKT_NORMAL = 0, SILK_PKT_LOST = 1, SILK_PKT_FEC = 2 };
struct SilkPacketCtl {
  int mode;        // SILK_PKT_*
  int samples48;   // lost packet: samples per channel to conceal at 48 kHz (120 .. 960); else unused
  int limit;       // samples per channel the row takes (a concealed frame shorter than 10 ms still runs a 10 ms SILK frame)
};
// frame word 2 of a SILK record (SL_BASE + i * SL_FRAME_WORDS + 2): bit0 decode_only_middle, and for a packet decoded from
// its LBRR data (FEC): bit1 mid-only flag not coded (the stream's previous one stands), bit2 stereo predictor not coded
enum { SILK_FR_DOM = 1, SILK_FR_DOM_KEPT = 2, SILK_FR_PRED_KEPT = 4 };

// One SILK packet of one stream: silk_decode (dec_api.rs:132-441) once per 10/20 ms SILK frame in the packet; a lost packet
// (no record) is one concealed frame. stage: 32 lane-private words; fir: interpolation table; side_live: the side
// channel's ring holds its history. Returns samples per channel produced at 48 kHz.
// h0: the SILK header word; frames: the first frame record (SILK-only: slots S_HDR0 / SL_BASE, hybrid: S_SILK_HYB / + 1).
MB_DEV_NOINLINE int silk_synth_packet(SilkState& S, SilkWork& W, const SilkHistRef& H, const SilkExcRef& X, int& side_live, uint32_t h0,
                                      const uint32_t* frames, uint32_t quirks, const SilkPcmOut& o, int32_t* stage, const int16_t* fir,
                                      const SilkPacketCtl& ctl) {
  const bool lost = ctl.mode == SILK_PKT_LOST, fec = ctl.mode == SILK_PKT_FEC;
  if (lost) {   // the packet's geometry is the decoder's own (opus_decoder.rs:651-676): rate and channels of the last packet
    const int fs0 = S.ch[0].fs_khz;
    if (fs0 == 0) return 0;
    h0 = (uint32_t)fs0 | ((ctl.samples48 >= 960 ? 4u : 2u) << 5) | (1u << 8) | ((uint32_t)S.n_channels_internal << 10);
  }
  const int fs = (int)MB_SH_FS(h0), nb = (int)MB_SH_SUBFR(h0), n_frames = (int)MB_SH_FRAMES(h0), nci = (int)MB_SH_NCI(h0);
  const int CC = o.CC;
  const int L = 5 * fs * nb, ltp_mem = 20 * fs;
  int pos = 0;
  if (nci == 2 && !side_live) { silk_hist_load(W, H, 1); side_live = 1; }
  for (int i = 0; i < n_frames; i++) {
    const uint32_t* fr = frames + (i * SL_FRAME_WORDS) * 32;
    if (nci > S.n_channels_internal) {                       // dec_api.rs:150-154
      if (quirks & S4_STEREO_REINIT) { silk_chan_reset(S.ch[0]); silk_ring_clear(W.hist[0], W.pos[0]); silk_work_drop(W, 0); }
      silk_chan_reset(S.ch[1]); silk_ring_clear(W.hist[1], W.pos[1]); silk_work_drop(W, 1);
    }
    const bool stereo_to_mono = nci == 1 && S.n_channels_internal == 2 && fs == S.ch[0].fs_khz;
    for (int n = 0; n < nci; n++)
      if (silk_set_fs(S.ch[n], fs, nb)) silk_ring_clear(W.hist[n], W.pos[n]);
    if (CC == 2 && nci == 2 && (S.n_channels_api == 1 || S.n_channels_internal == 1)) {   // dec_api.rs:181-191
      S.pred_prev_q13[0] = S.pred_prev_q13[1] = 0; S.s_side[0] = S.s_side[1] = 0;
      for (int k = 0; k < 6; k++) S.ch[1].res_iir[k] = S.ch[0].res_iir[k];
      for (int k = 0; k < 8; k++) S.ch[1].res_fir[k] = S.ch[0].res_fir[k];
      for (int k = 0; k < 16; k++) S.ch[1].res_delay[k] = S.ch[0].res_delay[k];
      for (int k = 0; k < 24; k++) S.ch[1].res_down[k] = S.ch[0].res_down[k];
      S.ch[1].res_ready = 1;
    }
    S.n_channels_api = (int8_t)CC; S.n_channels_internal = (int8_t)nci;
    // stereo predictor and mid-only flag: coded, or the stream's previous ones (lost frame; LBRR frame without them)
    const uint32_t w2 = lost ? (uint32_t)(SILK_FR_DOM_KEPT | SILK_FR_PRED_KEPT) : fr[64];
    int32_t pred[2];
    if (w2 & SILK_FR_PRED_KEPT) { pred[0] = S.pred_prev_q13[0]; pred[1] = S.pred_prev_q13[1]; }
    else { pred[0] = (int32_t)fr[0]; pred[1] = (int32_t)fr[32]; }
    const int dom = nci == 2 ? ((w2 & SILK_FR_DOM_KEPT) ? (int)S.prev_decode_only_middle : (int)(w2 & SILK_FR_DOM)) : 0;
    if (nci == 2 && !dom && S.prev_decode_only_middle) {     // dec_api.rs:263-271
      SilkChanState& c = S.ch[1];
      silk_ring_clear(W.hist[1], W.pos[1]);
      for (int k = 0; k < 16; k++) c.s_lpc_q14[k] = 0;
      c.lag_prev = 100; c.last_gain_index = 10; c.prev_signal_type = 0; c.first_frame_after_reset = 1;
    }
    int side_coded = 0;
    if (fec && nci == 2) { SilkSide t; silk_side_load(t, fr + (3 + SL_CH_WORDS) * 32); side_coded = t.coded; }
    // dec_api.rs:296-306
    const bool has_side = lost ? !S.prev_decode_only_middle : (fec ? (!S.prev_decode_only_middle || side_coded) : !dom);
    bool modified[2] = {false, false};   // the frame that leaves the channel differs from the one its history keeps: W.lr[n] + 2
    for (int n = 0; n < nci; n++) {
      int16_t* h = W.hist[n];
      SilkChanState& c = S.ch[n];
      const int hp = W.pos[n];
      if (n > 0 && !has_side) continue;    // side channel absent: the frame is zeros and the channel is left alone (dec_api.rs:318-321)
      bool coded = false;
      SilkFrameParams P;
      if (!lost) {
        const uint32_t* chw = fr + (3 + n * SL_CH_WORDS) * 32;
        SilkSide sd;
        silk_side_load(sd, chw);
        coded = sd.coded != 0;
        if (coded) {
          silk_frame_params(c, sd, P, quirks);
          silk_frame_core(c, sd, P, chw + SL_IDX_WORDS * 32, h, hp, W.sltp_q15);
          silk_plc_update(c, sd, P);                       // plc.rs:117-184 (prev_signal_type among the rest)
          c.loss_count = 0;
          c.first_frame_after_reset = 0;
          c.lag_prev = (int16_t)P.pitch_l[nb - 1];
          W.pos[n] = (hp + L) & SILK_RING_MASK;            // decode_frame.rs:207-230: the history window slides by a frame
          {   // the frame's excitation: in the record until somebody needs it in HBM (silk_exc_flush). A shorter frame leaves
              // the tail of a longer one standing in the reference's buffer (decode_core.rs:93-111 writes frame_length words),
              // and the concealment reads across it (plc.rs:370-388): the longer one goes to HBM before it is forgotten.
            if (W.last_exc[n].valid && W.last_exc[n].len > L) silk_exc_flush(c, W, X, n);
            SilkExcSrc& e = W.last_exc[n];
            e.pulse_words = chw + SL_IDX_WORDS * 32; e.from = 0; e.len = (int16_t)L;
            e.seed = sd.seed; e.signal_type = sd.signal_type; e.quant_offset = sd.quant_offset; e.valid = 1;
          }
          silk_cng_track(c, W, X, n, chw + SL_IDX_WORDS * 32, &sd, P.gains_q16, P.seed_at);
          modified[n] = c.plc_last_frame_lost != 0;
        }
      }
      if (!coded) {   // concealed frame (decode_frame.rs:140-151)
        c.lag_prev = (int16_t)silk_plc_conceal(c, W, X, n, h, hp, W.sltp_q15);
        W.pos[n] = (hp + L) & SILK_RING_MASK;
        silk_cng_track(c, W, X, n, nullptr, nullptr, nullptr, nullptr);
        modified[n] = true;
      }
      if (modified[n]) {
        const int b = W.pos[n] + ltp_mem - L;
        for (int k = 0; k < L; k++) W.lr[n][2 + k] = h[(b + k) & SILK_RING_MASK];
        if (!coded) silk_cng_add(c, W, X, n, W.lr[n] + 2, W.sltp_q15, quirks);
        silk_plc_glue(c, W.lr[n] + 2, L);
      }
    }
    SilkSink sink;
    sink.stage = stage; sink.fill = 0; sink.o = &o; sink.dup = 0; sink.limit = ctl.limit;
    int n_out = 0;
    if (CC == 2 && nci == 2) {
      for (int n = 0; n < 2; n++) {
        if (modified[n]) continue;
        const bool present = n == 0 || has_side;
        const int b = W.pos[n] + ltp_mem - L;              // first sample of the frame just written
        for (int k = 0; k < L; k++) W.lr[n][2 + k] = present ? W.hist[n][(b + k) & SILK_RING_MASK] : (int16_t)0;
      }
      silk_ms_to_lr(S, W.lr[0], W.lr[1], pred, fs, L);
      for (int n = 0; n < 2; n++) {
        SilkSrc src; src.p = W.lr[n] + 1; src.base = 0; src.mask = 0xffff; src.first = 0; src.has_first = 0;
        sink.pos = pos; sink.ch = n;
        n_out = silk_resample48(S.ch[n], src, L, fir, sink);
        if (sink.fill) silk_sink_flush(sink);
      }
    } else {
      // mid only: one sample of delay through s_mid (dec_api.rs:341-347), read straight from the ring
      const int b = W.pos[0] + ltp_mem - L;
      SilkSrc src; src.p = W.hist[0]; src.base = b - 1; src.mask = SILK_RING_MASK; src.first = S.s_mid[1]; src.has_first = 1;
      int16_t m0 = W.hist[0][(b + L - 2) & SILK_RING_MASK], m1 = W.hist[0][(b + L - 1) & SILK_RING_MASK];
      if (modified[0]) { src.p = W.lr[0]; src.base = 1; src.mask = 0xffff; m0 = W.lr[0][L]; m1 = W.lr[0][L + 1]; }
      if (CC == 2 && stereo_to_mono) {
        sink.pos = pos; sink.ch = 0;
        n_out = silk_resample48(S.ch[0], src, L, fir, sink);
        if (sink.fill) silk_sink_flush(sink);
        sink.pos = pos; sink.ch = 1;
        silk_resample48(S.ch[1], src, L, fir, sink);
        if (sink.fill) silk_sink_flush(sink);
      } else {
        sink.pos = pos; sink.ch = 0; sink.dup = CC == 2;
        n_out = silk_resample48(S.ch[0], src, L, fir, sink);
        if (sink.fill) silk_sink_flush(sink);
      }
      S.s_mid[0] = m0; S.s_mid[1] = m1;
    }
    if (lost) { for (int n = 0; n < nci; n++) S.ch[n].last_gain_index = 10; }   // dec_api.rs:430-436
    else S.prev_decode_only_middle = (int8_t)dom;
    pos += n_out;
  }
  return pos;
}

// ---- the per-stream lane of k_silk: the stream's SILK-side cells in tick order -------------------------------------------
struct SilkLaneParams {
  uint32_t n_lanes, n_ticks;
  const uint32_t* pkt_map; const uint32_t* lane_stream; const uint32_t* rec; const FrameMeta* meta;
  SilkState* st_all; uint32_t* silk_hist; int32_t* silk_exc; uint32_t n_streams;
  float* pcm_f32; int16_t* pcm_i16; uint32_t pcm_stride; int CC; uint32_t quirks; float gain; int apply_gain;
  const uint32_t* out_map;
  float* xbuf; const uint32_t* side_of;   // side rows (mb_celt_tail.cuh): where a SILK -> CELT switch without redundancy conceals into
};
// W: the lane's working set (local memory on the device); stage: 32 lane-private words; fir: interpolation table.
MB_DEV void silk_lane(const SilkLaneParams& q, uint32_t lane, SilkWork& W, int32_t* stage, const int16_t* fir) {
  bool any = false;
  for (uint32_t t = 0; t < q.n_ticks && !any; t++) any = (q.meta[t * q.n_lanes + lane].flags & 2u) != 0;
  if (!any) return;
  const uint32_t sid = q.lane_stream ? q.lane_stream[lane] : lane;
  SilkState* gs = q.st_all + sid;
  SilkState S = *gs;
  SilkHistRef H = {q.silk_hist, q.n_streams, sid};
  SilkExcRef X = {q.silk_exc, q.n_streams, sid};
  int side_live = 0;
  silk_hist_load(W, H, 0);
  silk_work_drop(W, 0); silk_work_drop(W, 1);
  for (uint32_t t = 0; t < q.n_ticks; t++) {
    const uint32_t f = t * q.n_lanes + lane;
    const uint32_t fl = q.meta[f].flags;
    if (!(fl & 2u)) continue;
    const uint32_t p = q.pkt_map ? q.pkt_map[f] : f;
    const uint32_t* rb = q.rec + rec_index(f, 0);
    const uint32_t status = rb[S_STATUS * 32];
    if ((fl & 1u) && (fl & 512u) && (fl & MB_FL_TRANS_IN) && !(fl & (8u | 16u))) {
      // a CELT-only packet right after a SILK-only / hybrid one, no redundancy frame (opus_decoder.rs:575-593): min(frame, 5 ms)
      // concealed by the SILK decoder as it stands (decode_frame(None): a 10 ms SILK frame, of which the row takes that much),
      // into the side row k_fade blends from. The SILK state is not read again before its next reset (:694-696).
      SilkPcmOut o;
      o.f32 = side_row(q.xbuf, q.side_of[f]); o.i16 = nullptr;
      o.CC = q.CC; o.gain = q.gain; o.apply_gain = 0; o.gain_from = 0;
      SilkPacketCtl ctl;
      ctl.mode = SILK_PKT_LOST;
      ctl.samples48 = imin(SHORT_MDCT << (int)MB_H0_LM(rb[S_HDR0 * 32]), 2 * SHORT_MDCT);
      ctl.limit = ctl.samples48 / quirk_ds(q.quirks);
      o.fs_out_khz = 48 / quirk_ds(q.quirks);
      silk_synth_packet(S, W, H, X, side_live, 0u, rb, q.quirks, o, stage, fir, ctl);
      continue;
    }
    const bool to_float = (fl & MB_FL_F32ROW) != 0;   // the packet goes through the float row (a CELT frame shares one of its rows, or a decode gain applies)
    const bool shared = (fl & (8u | 16u)) != 0;       // a CELT frame shares THIS row (hybrid, fade-out, redundancy): k_deemph / k_fade finish it
    const bool hyb_layout = ((status >> 17) & 3u) != 0;       // hybrid, or a one-frame SILK-only cell
    if (fl & 32u) {   // CELT -> SILK: silk decoder reset (opus_decoder.rs:694-696)
      silk_state_reset(S, 0);
      silk_ring_clear(W.hist[0], W.pos[0]); silk_ring_clear(W.hist[1], W.pos[1]);
      silk_work_drop(W, 0); silk_work_drop(W, 1);
      side_live = 1;
    }
    SilkPcmOut o;
    // SILK-only cells are final: int16 directly on the i16 path
    o.f32 = (to_float || !q.pcm_i16) ? q.pcm_f32 + pcm_elem_off(q.out_map, p, q.pcm_stride) : nullptr;
    o.i16 = (to_float || !q.pcm_i16) ? nullptr : q.pcm_i16 + pcm_elem_off(q.out_map, p, q.pcm_stride);
    // decode gain: once on the finished sum. Hybrid: k_deemph does it for the whole row; fade-out: k_deemph for the first
    // 120 samples, here for the rest; redundancy: k_fade for the whole row
    const bool fadeout = (fl & 8u) && !((status >> 17) & 1u) && !MB_ST_LOST(status);
    o.CC = q.CC; o.gain = q.gain; o.apply_gain = ((shared && !fadeout) || (fl & MB_FL_FADE_ROW)) ? 0 : q.apply_gain;
    o.gain_from = fadeout ? SHORT_MDCT / quirk_ds(q.quirks) : 0;
    o.fs_out_khz = 48 / quirk_ds(q.quirks);
    SilkPacketCtl ctl;
    ctl.mode = MB_ST_LOST(status) ? SILK_PKT_LOST : (MB_ST_FEC(status) ? SILK_PKT_FEC : SILK_PKT_NORMAL);
    ctl.samples48 = ctl.mode == SILK_PKT_LOST ? (int)rb[S_RANGE * 32] : 0;
    ctl.limit = ctl.mode == SILK_PKT_LOST ? ctl.samples48 / quirk_ds(q.quirks) : 0x7fffffff;   // (the row counts output samples)
    silk_synth_packet(S, W, H, X, side_live, rb[(hyb_layout ? S_SILK_HYB : S_HDR0) * 32], rb + (hyb_layout ? S_SILK_HYB + 1 : SL_BASE) * 32, q.quirks,
                      o, stage, fir, ctl);
  }
  // the pass ends: what the next call may need of this one's records goes to HBM
  for (int n = 0; n < 2; n++) { silk_exc_flush(S.ch[n], W, X, n); silk_cng_flush(S.ch[n], W, X, n); }
  silk_hist_store(W, H, 0);
  if (side_live) silk_hist_store(W, H, 1);
  *gs = S;
}

}  // namespace mb
