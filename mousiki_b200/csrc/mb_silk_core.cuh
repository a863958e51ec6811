// SILK synthesis core for one lane: excitation -> (LTP) -> LPC synthesis of one frame of one channel.
// Replaces decode_core.rs:28-330 and the re-whitening filter lpc_analysis_filter.rs:18-45.
//
// Layout choices (lane-private, see mb_silk_synth.cuh header):
//  * the output history is a 1024-entry ring per channel: [pos, pos + ltp_mem) are the last 20 ms, the frame under
//    construction is appended behind it, and "shifting the history" (decode_frame.rs:207-230) is pos += frame_length;
//  * the 5 LTP taps and the order + 1 re-whitening taps slide over register windows: one local-memory load per sample
//    instead of 5 / 17;
//  * every Q16 x 16-bit product (silk_SMLAWB) is one IMAD.HI: (a * b) >> 16 == mulhi(a, b << 16) for 16-bit b, with the
//    coefficients pre-shifted once per subframe;
//  * the 10/16 short-term products are summed as four independent partial sums (wrapping adds commute).
#pragma once
#include "mb_silk_synth.cuh"

namespace mb {

enum { SILK_RING = 1024, SILK_RING_MASK = SILK_RING - 1 };

// Where a decoded frame's excitation can be rebuilt from while the pass lasts (mb_silk_plc.cuh): its pulses in the cell record
struct SilkExcSrc { const uint32_t* pulse_words; int32_t seed; int16_t from, len; int8_t signal_type, quant_offset, valid; };  // seed: at sample `from`
struct SilkCngPush { SilkExcSrc src; int16_t sub, nb; };

struct SilkWork {
  SilkExcSrc last_exc[2];      // per channel: the last decoded frame of the pass, its excitation not yet in HBM
  SilkCngPush cng_pend[2][4];  // per channel: comfort-noise FIFO pushes not yet applied in HBM, oldest first
  int cng_n[2];
  int16_t hist[2][SILK_RING];  // per channel output ring (history + running frame)
  int32_t sltp_q15[640];       // gain-normalised LTP state of the running frame (decode_core.rs s_ltp_q15)
  int16_t lr[2][322];          // stereo only: mid/side -> left/right staging (2 history samples + frame)
  int pos[2];                  // ring index of the oldest history sample of each channel
};

MB_DEV int32_t s_mulhi(int32_t a, int32_t b) {
#if MB_ON_DEVICE
  return __mulhi(a, b);
#else
  return (int32_t)(((int64_t)a * b) >> 32);
#endif
}

// `pulse_words`: the frame's pulses in the cell record (2 x i16 per word, stride 32 words).
// `h`/`pos`: this channel's ring; the frame is written to h[(pos + ltp_mem + n) & mask].
MB_DEV_NOINLINE void silk_frame_core(SilkChanState& st, const SilkSide& sd, SilkFrameParams& P, const uint32_t* __restrict__ pulse_words,
                                     int16_t* __restrict__ h, int pos, int32_t* __restrict__ sltp_q15) {
  const int fs = st.fs_khz, nb = st.nb_subfr, sub = 5 * fs, ltp_mem = 20 * fs, order = fs == 16 ? 16 : 10;
  const int32_t offset_q10 = silk_Quantization_Offsets_Q10[(sd.signal_type >> 1) * 2 + sd.quant_offset];
  const bool interp = sd.interp_q2 < 4;
  const bool voiced = sd.signal_type == SILK_T_VOICED;
  // first frame after a loss that ended a voiced stretch: its first two subframes keep a quarter-gain pitch predictor at
  // the previous lag (decode_core.rs:140-152)
  const bool trans = st.loss_count > 0 && st.prev_signal_type == SILK_T_VOICED && !voiced;
  int32_t seed = sd.seed;
  int32_t lpc[16];                   // synthesis history, lpc[j] = sLPC_Q14 of sample (i - 1 - j)
#pragma unroll
  for (int j = 0; j < 16; j++) lpc[j] = st.s_lpc_q14[15 - j];
  int sltp_idx = ltp_mem;
  int32_t prev_gain = st.prev_gain_q16;
  for (int k = 0; k < nb; k++) {
    P.seed_at[k] = seed;
    int32_t A[16];                   // short-term taps, pre-shifted for IMAD.HI; zero above the order
#pragma unroll
    for (int j = 0; j < 16; j++) A[j] = j < order ? s_shl(P.a_q12[k >> 1][j], 16) : 0;
    const int32_t gain_q16 = P.gains_q16[k], gain_q10 = gain_q16 >> 6;
    int32_t inv_gain_q31 = silk_inverse32(gain_q16, 47);
    int32_t gain_adj = 1 << 16;
    if (gain_q16 != prev_gain) {
      gain_adj = silk_div32(prev_gain, gain_q16, 16);
#pragma unroll
      for (int j = 0; j < 16; j++) lpc[j] = s_mulww(gain_adj, lpc[j]);
    }
    prev_gain = gain_q16;
    int32_t B0 = 0, B1 = 0, B2 = 0, B3 = 0, B4 = 0;          // LTP taps, pre-shifted
    int32_t l0 = 0, l1 = 0, l2 = 0, l3 = 0, l4 = 0;          // sltp_q15[pl], [pl-1], .. [pl-4]
    int32_t lnext = 0;
    int pl = 0;
    const bool vk = voiced || (trans && k < 2);
    if (trans && k < 2) P.pitch_l[k] = st.lag_prev;
    if (vk) {
      const int lag = P.pitch_l[k];
      if (k == 0 || (k == 2 && interp)) {
        // re-whitening: analysis filter over the last lag + 2 samples before subframe k (at k == 2 that includes the two
        // subframes just synthesised, which sit right behind the history in the ring), scaled into the normalised domain
        if (k == 0) inv_gain_q31 = s_shl(s_mulwb(inv_gain_q31, P.ltp_scale_q14), 2);
        const int base = pos + k * sub;                      // in[x] = h[(base + x) & mask]
        int32_t win[17];                                     // win[j] = in[ix - j], j = 0..16
#pragma unroll
        for (int j = 0; j < 17; j++) win[j] = h[(base + ltp_mem - 1 - j) & SILK_RING_MASK];
        for (int i = 0; i < lag + 2; i++) {
          int32_t a0 = 0, a1 = 0;
#pragma unroll
          for (int j = 0; j < 16; j += 2) {
            a0 = s_add(a0, win[j + 1] * (A[j] >> 16));
            a1 = s_add(a1, win[j + 2] * (A[j + 1] >> 16));
          }
          const int32_t r = s_sat16(s_rr(s_sub(s_shl(win[0], 12), s_add(a0, a1)), 12));
          sltp_q15[sltp_idx - 1 - i] = s_mulwb(inv_gain_q31, r);
#pragma unroll
          for (int j = 0; j < 16; j++) win[j] = win[j + 1];
          win[16] = h[(base + ltp_mem - 1 - i - 17) & SILK_RING_MASK];
        }
      } else if (gain_adj != (1 << 16)) {
        for (int i = 0; i < lag + 2; i++) sltp_q15[sltp_idx - 1 - i] = s_mulww(gain_adj, sltp_q15[sltp_idx - 1 - i]);
      }
      pl = sltp_idx - lag + 2;
      const int16_t* B = P.b_q14 + 5 * k;
      if (voiced) { B0 = s_shl(B[0], 16); B1 = s_shl(B[1], 16); B2 = s_shl(B[2], 16); B3 = s_shl(B[3], 16); B4 = s_shl(B[4], 16); }
      else B2 = s_shl(4096, 16);
      l0 = sltp_q15[pl]; l1 = sltp_q15[pl - 1]; l2 = sltp_q15[pl - 2]; l3 = sltp_q15[pl - 3]; l4 = sltp_q15[pl - 4];
      lnext = sltp_q15[pl + 1];   // one sample ahead of the window (written at least lag - 3 >= 13 samples ago)
    }
    const int out0 = pos + ltp_mem + k * sub;
    uint32_t w_next = pulse_words[((k * sub) >> 1) * 32];            // two pulses per record word, fetched one step ahead
    for (int i = 0; i < sub; i += 2) {
      const uint32_t w = w_next;
      if (i + 2 < sub) w_next = pulse_words[((k * sub + i + 2) >> 1) * 32];
#pragma unroll
      for (int u = 0; u < 2; u++) {
        // excitation (decode_core.rs:93-111)
        const int32_t pulse = (int16_t)(u ? (w >> 16) : (w & 0xffffu));
        seed = s_add(907633515, (int32_t)((uint32_t)seed * 196314165u));
        int32_t e = pulse << 14;
        if (e > 0) e -= 1280; else if (e < 0) e += 1280;
        e += offset_q10 << 4;
        if (seed < 0) e = -e;
        seed = s_add(seed, pulse);
        int32_t res = e;
        if (vk) {
          int32_t p = 2;
          p = s_add(p, s_mulhi(l0, B0)); p = s_add(p, s_mulhi(l1, B1)); p = s_add(p, s_mulhi(l2, B2));
          p = s_add(p, s_mulhi(l3, B3)); p = s_add(p, s_mulhi(l4, B4));
          res = s_add(e, s_shl(p, 1));
          sltp_q15[sltp_idx++] = s_shl(res, 1);
          pl++;
          l4 = l3; l3 = l2; l2 = l1; l1 = l0; l0 = lnext;
          lnext = sltp_q15[pl + 1];
        }
        int32_t p0 = order >> 1, p1 = 0, p2 = 0, p3 = 0;
#pragma unroll
        for (int j = 0; j < 16; j += 4) {
          p0 = s_add(p0, s_mulhi(lpc[j], A[j])); p1 = s_add(p1, s_mulhi(lpc[j + 1], A[j + 1]));
          p2 = s_add(p2, s_mulhi(lpc[j + 2], A[j + 2])); p3 = s_add(p3, s_mulhi(lpc[j + 3], A[j + 3]));
        }
        const int32_t pred = s_add(s_add(p0, p1), s_add(p2, p3));
        const int32_t s = s_sat32((int64_t)res + s_sat32((int64_t)pred << 4));
#pragma unroll
        for (int j = 15; j > 0; j--) lpc[j] = lpc[j - 1];
        lpc[0] = s;
        h[(out0 + i + u) & SILK_RING_MASK] = (int16_t)s_sat16(s_rr(s_mulww(s, gain_q10), 8));
      }
    }
  }
  st.prev_gain_q16 = prev_gain;
#pragma unroll
  for (int j = 0; j < 16; j++) st.s_lpc_q14[15 - j] = lpc[j];
}

}  // namespace mb
