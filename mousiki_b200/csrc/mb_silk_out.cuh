// Phase 2 for SILK-only packets, output side: channel plumbing of silk_decode, mid/side -> left/right, the API
// resampler to 48 kHz and the PCM store; and the per-stream lane loop that walks a stream's packets in tick order.
// Reference: src/silk/dec_api.rs:132-441, stereo_ms_to_lr.rs:26-143, resampler.rs:87-176,279-369,
// resampler_private_iir_fir.rs:80-197, resampler_private_up2_hq.rs:31-117, decoder_set_fs.rs:137-198,
// opus_decoder.rs:596-745 (SILK arm of decode_frame), :1524-1533 (silk_int16_to_res).
#pragma once
#include "mb_silk_synth.cuh"
#include "mb_celt_energy.cuh"

namespace mb {

// decoder_set_fs.rs:137-198 for a fixed 48 kHz API rate. The resampler (always the IIR/FIR up-sampler here:
// 8/12/16 -> 48 kHz) is re-initialised when the internal rate changes or the channel was reset.
MB_DEV void silk_set_fs(SilkChanState& c, int fs_khz, int nb_subfr) {
  const int frame_length = 5 * fs_khz * nb_subfr;
  if (c.fs_khz != fs_khz || !c.res_ready) {
    for (int i = 0; i < 6; i++) c.res_iir[i] = 0;
    for (int i = 0; i < 8; i++) c.res_fir[i] = 0;
    for (int i = 0; i < 16; i++) c.res_delay[i] = 0;
    c.res_ready = 1;
  }
  if (c.fs_khz != fs_khz) {
    c.first_frame_after_reset = 1; c.lag_prev = 100; c.last_gain_index = 10; c.prev_signal_type = 0;
    for (int i = 0; i < 320; i++) c.out_buf[i] = 0;
    for (int i = 0; i < 16; i++) c.s_lpc_q14[i] = 0;
  }
  c.fs_khz = (int16_t)fs_khz; c.nb_subfr = (int16_t)nb_subfr; c.frame_length = (int16_t)frame_length;
}

// where the samples of one packet go
struct SilkPcmOut {
  float* f32;      // interleaved, or null
  int16_t* i16;    // interleaved, or null
  int CC;          // API channels (interleave stride)
  float gain; int apply_gain;
};
MB_DEV void silk_store_sample(const SilkPcmOut& o, int pos, int c, int32_t v) {
  const size_t at = (size_t)pos * o.CC + c;
  if (o.f32) {
    float x = (float)v * (1.0f / 32768.0f);
    if (o.apply_gain) x *= o.gain;      // opus_decoder.rs:1376-1392
    o.f32[at] = x;
  } else o.i16[at] = (int16_t)v;
}

// resampler_private_up2_hq.rs:31-117, one input sample -> two output samples
MB_DEV void silk_up2(int32_t* S, int32_t in, int16_t* out2) {
  const int32_t in32 = in << 10;
  int32_t Y, X, o1, o2;
  Y = s_sub(in32, S[0]); X = s_mulwb(Y, silk_resampler_up2_hq_0[0]); o1 = s_add(S[0], X); S[0] = s_add(in32, X);
  Y = s_sub(o1, S[1]); X = s_mulwb(Y, silk_resampler_up2_hq_0[1]); o2 = s_add(S[1], X); S[1] = s_add(o1, X);
  Y = s_sub(o2, S[2]); X = s_mlawb(Y, Y, silk_resampler_up2_hq_0[2]); o1 = s_add(S[2], X); S[2] = s_add(o2, X);
  out2[0] = (int16_t)s_sat16(s_rr(o1, 10));
  Y = s_sub(in32, S[3]); X = s_mulwb(Y, silk_resampler_up2_hq_1[0]); o1 = s_add(S[3], X); S[3] = s_add(in32, X);
  Y = s_sub(o1, S[4]); X = s_mulwb(Y, silk_resampler_up2_hq_1[1]); o2 = s_add(S[4], X); S[4] = s_add(o1, X);
  Y = s_sub(o2, S[5]); X = s_mlawb(Y, Y, silk_resampler_up2_hq_1[2]); o1 = s_add(S[5], X); S[5] = s_add(o2, X);
  out2[1] = (int16_t)s_sat16(s_rr(o1, 10));
}

// resampler_private_iir_fir.rs:80-197: `n_in` samples in, batches of 10 ms; returns the number of samples written.
// up: lane-private window of 8 + 2*batch samples; iir/fir: the persistent filter state (registers while running).
MB_DEV_NOINLINE int silk_iir_fir(int32_t* iir, int16_t* fir, int16_t* up, const int16_t* in, int n_in, int batch, int32_t inc_q16,
                                 const SilkPcmOut& o, int pos, int c) {
  if (n_in <= 0) return 0;
  for (int i = 0; i < 8; i++) up[i] = fir[i];
  int produced = 0, n = 0;
  while (n_in > 0) {
    n = n_in < batch ? n_in : batch;
    for (int k = 0; k < n; k++) silk_up2(iir, in[k], up + 8 + 2 * k);
    const int32_t max_idx = n << 17;
    for (int32_t idx = 0; idx < max_idx; idx += inc_q16) {
      const int ti = ((idx & 0xffff) * 12) >> 16;
      const int16_t* b = up + (idx >> 16);
      const int16_t* f = silk_resampler_frac_FIR_12 + 4 * ti;
      const int16_t* r = silk_resampler_frac_FIR_12 + 4 * (11 - ti);
      int32_t acc = (int32_t)b[0] * f[0];
      acc += (int32_t)b[1] * f[1]; acc += (int32_t)b[2] * f[2]; acc += (int32_t)b[3] * f[3];
      acc += (int32_t)b[4] * r[3]; acc += (int32_t)b[5] * r[2]; acc += (int32_t)b[6] * r[1]; acc += (int32_t)b[7] * r[0];
      silk_store_sample(o, pos + produced, c, s_sat16((acc + (1 << 14)) >> 15));
      produced++;
    }
    in += n; n_in -= n;
    if (n_in > 0) for (int i = 0; i < 8; i++) up[i] = up[2 * n + i];
  }
  for (int i = 0; i < 8; i++) fir[i] = up[2 * n + i];
  return produced;
}

// resampler.rs:279-369 (silk_resampler) for the up-sampling kernels: 1 ms from the delay line, then the rest.
MB_DEV int silk_resample48(SilkChanState& c, int16_t* up, const int16_t* in, int n_in, const SilkPcmOut& o, int pos, int ch) {
  const int fs = c.fs_khz;
  const int delay = fs == 8 ? 0 : (fs == 12 ? 4 : 7);                 // DELAY_MATRIX_DEC[.][48 kHz] (resampler.rs:228)
  const int32_t inc = fs == 8 ? 21846 : (fs == 12 ? 32768 : 43691);   // compute_inv_ratio_q16(fs, 48000, up2x = 1)
  const int ns = fs - delay;
  for (int i = 0; i < ns; i++) c.res_delay[delay + i] = in[i];
  int produced = silk_iir_fir(c.res_iir, c.res_fir, up, c.res_delay, fs, 10 * fs, inc, o, pos, ch);
  produced += silk_iir_fir(c.res_iir, c.res_fir, up, in + ns, n_in - fs, 10 * fs, inc, o, pos + produced, ch);
  for (int i = 0; i < delay; i++) c.res_delay[i] = in[n_in - delay + i];
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

// One SILK-only packet of one stream: silk_decode (dec_api.rs:132-441) once per 10/20 ms SILK frame in the packet.
// Returns samples per channel written at 48 kHz.
MB_DEV_NOINLINE int silk_synth_packet(SilkState& S, SilkWork& W, const uint32_t* rec_base, uint32_t quirks, const SilkPcmOut& o) {
  const uint32_t h0 = rec_base[S_HDR0 * 32];
  const int fs = (int)MB_SH_FS(h0), nb = (int)MB_SH_SUBFR(h0), n_frames = (int)MB_SH_FRAMES(h0), nci = (int)MB_SH_NCI(h0);
  const int CC = o.CC;
  const int L = 5 * fs * nb, ltp_mem = 20 * fs;
  int pos = 0;
  for (int i = 0; i < n_frames; i++) {
    const uint32_t* fr = rec_base + (SL_BASE + i * SL_FRAME_WORDS) * 32;
    if (nci > S.n_channels_internal) {                       // dec_api.rs:150-154
      if (quirks & S4_STEREO_REINIT) silk_chan_reset(S.ch[0]);
      silk_chan_reset(S.ch[1]);
    }
    const bool stereo_to_mono = nci == 1 && S.n_channels_internal == 2 && fs == S.ch[0].fs_khz;
    for (int n = 0; n < nci; n++) silk_set_fs(S.ch[n], fs, nb);
    if (CC == 2 && nci == 2 && (S.n_channels_api == 1 || S.n_channels_internal == 1)) {   // dec_api.rs:181-191
      S.pred_prev_q13[0] = S.pred_prev_q13[1] = 0; S.s_side[0] = S.s_side[1] = 0;
      for (int k = 0; k < 6; k++) S.ch[1].res_iir[k] = S.ch[0].res_iir[k];
      for (int k = 0; k < 8; k++) S.ch[1].res_fir[k] = S.ch[0].res_fir[k];
      for (int k = 0; k < 16; k++) S.ch[1].res_delay[k] = S.ch[0].res_delay[k];
      S.ch[1].res_ready = 1;
    }
    S.n_channels_api = (int8_t)CC; S.n_channels_internal = (int8_t)nci;
    const int32_t pred[2] = {(int32_t)fr[0], (int32_t)fr[32]};
    const int dom = (int)fr[64];
    if (nci == 2 && !dom && S.prev_decode_only_middle) {     // dec_api.rs:263-271
      SilkChanState& c = S.ch[1];
      for (int k = 0; k < 320; k++) c.out_buf[k] = 0;
      for (int k = 0; k < 16; k++) c.s_lpc_q14[k] = 0;
      c.lag_prev = 100; c.last_gain_index = 10; c.prev_signal_type = 0; c.first_frame_after_reset = 1;
    }
    for (int n = 0; n < nci; n++) {
      const uint32_t* chw = fr + (3 + n * SL_CH_WORDS) * 32;
      int16_t* h = W.hist[n];
      SilkChanState& c = S.ch[n];
      if (n == 0 || !dom) {
        SilkSide sd;
        silk_side_load(sd, chw);
        SilkFrameParams P;
        for (int k = 0; k < ltp_mem; k++) h[k] = c.out_buf[k];
        silk_frame_params(c, sd, P, quirks);
        silk_frame_core(c, sd, P, chw + SL_IDX_WORDS * 32, h, W.sltp_q15);
        c.prev_signal_type = sd.signal_type;               // plc.rs:123
        c.first_frame_after_reset = 0;
        for (int k = 0; k < ltp_mem; k++) c.out_buf[k] = h[k + L];   // decode_frame.rs:207-230
        c.lag_prev = (int16_t)P.pitch_l[nb - 1];
      } else {
        for (int k = 0; k < L; k++) h[ltp_mem + k] = 0;
      }
      for (int k = 0; k < L; k++) W.lr[n][2 + k] = h[ltp_mem + k];
    }
    if (CC == 2 && nci == 2) silk_ms_to_lr(S, W.lr[0], W.lr[1], pred, fs, L);
    else {
      W.lr[0][0] = S.s_mid[0]; W.lr[0][1] = S.s_mid[1];
      S.s_mid[0] = W.lr[0][L]; S.s_mid[1] = W.lr[0][L + 1];
    }
    const int active = CC < nci ? CC : nci;
    int n_out = 0;
    for (int n = 0; n < active; n++) n_out = silk_resample48(S.ch[n], W.up, W.lr[n] + 1, L, o, pos, n);
    if (CC == 2 && nci == 1) {
      if (stereo_to_mono) silk_resample48(S.ch[1], W.up, W.lr[0] + 1, L, o, pos, 1);
      else {
        for (int k = 0; k < n_out; k++) {
          if (o.f32) o.f32[(size_t)(pos + k) * 2 + 1] = o.f32[(size_t)(pos + k) * 2];
          else o.i16[(size_t)(pos + k) * 2 + 1] = o.i16[(size_t)(pos + k) * 2];
        }
      }
    }
    S.prev_decode_only_middle = (int8_t)dom;
    pos += n_out;
  }
  return pos;
}

}  // namespace mb
