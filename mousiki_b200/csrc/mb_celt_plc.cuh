// CELT packet-loss concealment by pitch repetition, and the time-domain tail of the streams that need it.
//
// Replaces:
//   celt_decode_lost, pitch branch                 src/celt/celt_decoder.rs:1576-1846
//   celt_plc_pitch_search                          src/celt/celt_decoder.rs:960-1029
//   pitch_downsample, celt_fir5, pitch_search,
//   find_best_pitch, celt_pitch_xcorr              src/celt/pitch.rs:118-187, 392-424, 557-696, 1358-1384, 1418-1545
//   celt_autocorr, celt_lpc, celt_iir              src/celt/lpc.rs:146-190, 343-376, 465-530
// A lost CELT frame is rare, so this is the slow path: one warp per stream that has such a frame in the pass walks the
// stream's frames in order — post-filter for the coded frames exactly like k_comb, pitch concealment for the lost ones —
// directly on the stream's linear time series in global memory, where the reference's history buffer (decode_mem: 2048
// samples and the 120-sample overlap region) is simply "the 2168 samples in front of the lost frame". Every sum keeps the
// reference's left-to-right order: the lanes share out the LAGS of a correlation, never the terms of one sum.
#pragma once
#include "mb_celt_tail.cuh"

namespace mb {

enum { PLC_LPC_ORDER = 24, PLC_LAG_MAX = 720, PLC_LAG_MIN = 100, PLC_HIST = 2048 };

struct PlcSmem {
  float lpraw[PLC_HIST / 2];     // down-sampled sum of the channels, before and ...
  float lp[PLC_HIST / 2];        // ... after the whitening FIR; later the windowed excitation history
  float exc[MAX_PERIOD + PLC_LPC_ORDER];
  float work[MAX_N + OVERLAP + 8];   // FIR output / products on their way into a serial sum
  float xcorr[(PLC_LAG_MAX - PLC_LAG_MIN) / 2];
  float ac[PLC_LPC_ORDER + 1];
  float lpc[PLC_LPC_ORDER];
  float red[4];
  float save[OVERLAP];           // the next frame's first samples while the concealed frame's overlap region sits on them
  int best[2];
};

MB_DEV float plc_serial_sum(const float* p, int n) { float e = 0.f; MB_ROLLED for (int i = 0; i < n; i++) e += p[i]; return e; }

// find_best_pitch (pitch.rs:118-187); one lane
MB_DEV void plc_find_best_pitch(const float* xcorr, const float* y, int ystride, int len, int max_pitch, int* best) {
  float syy = 1.0f;
  MB_ROLLED
  for (int j = 0; j < len; j++) syy += y[j * ystride] * y[j * ystride];
  float bn0 = -1.0f, bd0 = 0.0f, bn1 = -1.0f, bd1 = 0.0f;
  best[0] = 0; best[1] = max_pitch > 1 ? 1 : 0;
  MB_ROLLED
  for (int i = 0; i < max_pitch; i++) {
    const float corr = xcorr[i];
    if (corr > 0.0f) {
      float c16 = corr;
      c16 *= 1e-12f;
      const float num = c16 * c16;
      if (num * bd1 > bn1 * syy) {
        if (num * bd0 > bn0 * syy) { bn1 = bn0; bd1 = bd0; best[1] = best[0]; bn0 = num; bd0 = syy; best[0] = i; }
        else { bn1 = num; bd1 = syy; best[1] = i; }
      }
    }
    const float a = y[(i + len) * ystride], b = y[i * ystride];
    syy += a * a - b * b;
    if (syy < 1.0f) syy = 1.0f;
  }
}

// celt_lpc (lpc.rs:146-190); one lane
MB_DEV void plc_celt_lpc(float* lpc, const float* ac, int order) {
  for (int i = 0; i < order; i++) lpc[i] = 0.f;
  const float ac0 = ac[0];
  if (ac0 <= 1e-10f) return;
  float error = ac0;
  MB_ROLLED
  for (int i = 0; i < order; i++) {
    float rr = 0.f;
    MB_ROLLED
    for (int j = 0; j < i; j++) rr = fma32(lpc[j], ac[i - j], rr);
    rr += ac[i + 1];
    const float r = -(rr / error);
    lpc[i] = r;
    MB_ROLLED
    for (int j = 0; j < (i + 1) >> 1; j++) {
      const float t1 = lpc[j], t2 = lpc[i - 1 - j];
      lpc[j] = fma32(r, t2, t1);
      lpc[i - 1 - j] = fma32(r, t1, t2);
    }
    const float r2 = r * r;
    error = fma32(-r2, error, error);
    if (error <= 0.001f * ac0) break;
  }
}

// celt_autocorr without window (pitch_downsample) or on an already windowed copy: ac[k] = xcorr over n - lag terms, plus
// the tail terms, each a left-to-right sum; lane k takes lag k.
MB_DEV void plc_autocorr(Team tm, const float* x, float* ac, int lag, int n) {
  const int fast_n = n - lag;
  MB_ROLLED
  for (int k = tm.tid; k <= lag; k += tm.n) {
    float s = 0.f;
    MB_ROLLED
    for (int j = 0; j < fast_n; j++) s += x[j] * x[j + k];
    float d = 0.f;
    MB_ROLLED
    for (int i = k + fast_n; i < n; i++) d += x[i] * x[i - k];
    ac[k] = s + d;
  }
  MB_SYNC();
}

// celt_plc_pitch_search (celt_decoder.rs:960-1029). d0 / d1: the channels' 2048 history samples (d1 null for mono).
MB_DEV_NOINLINE int plc_pitch_search(Team tm, PlcSmem* sm, const float* d0, const float* d1) {
  const int half = PLC_HIST / 2;
  // pitch_downsample (pitch.rs:1418-1545)
  MB_ROLLED
  for (int i = tm.tid; i < half; i += tm.n) {
    float v = i ? 0.25f * d0[2 * i - 1] + 0.25f * d0[2 * i + 1] + 0.5f * d0[2 * i] : 0.25f * d0[1] + 0.5f * d0[0];
    if (d1) v += i ? 0.25f * d1[2 * i - 1] + 0.25f * d1[2 * i + 1] + 0.5f * d1[2 * i] : 0.25f * d1[1] + 0.5f * d1[0];
    sm->lpraw[i] = v;
  }
  MB_SYNC();
  plc_autocorr(tm, sm->lpraw, sm->ac, 4, half);
  if (tm.tid == 0) {
    float* ac = sm->ac;
    ac[0] *= 1.0001f;
    for (int i = 1; i <= 4; i++) { const float coeff = 0.008f * (float)i; ac[i] -= ac[i] * coeff * coeff; }
    float lpc[4];
    plc_celt_lpc(lpc, ac, 4);
    float tmp = 1.0f;
    for (int i = 0; i < 4; i++) { tmp *= 0.9f; lpc[i] *= tmp; }
    const float c1 = 0.8f;
    sm->lpc[0] = lpc[0] + 0.8f; sm->lpc[1] = fma32(c1, lpc[0], lpc[1]); sm->lpc[2] = fma32(c1, lpc[1], lpc[2]);
    sm->lpc[3] = fma32(c1, lpc[2], lpc[3]); sm->lpc[4] = c1 * lpc[3];
  }
  MB_SYNC();
  // celt_fir5 (pitch.rs:1358-1384): the memories are the five previous INPUT samples
  MB_ROLLED
  for (int i = tm.tid; i < half; i += tm.n) {
    float sum = sm->lpraw[i];
    MB_ROLLED
    for (int k = 0; k < 5; k++) sum = fma32(sm->lpc[k], i - 1 - k >= 0 ? sm->lpraw[i - 1 - k] : 0.f, sum);
    sm->lp[i] = sum;
  }
  MB_SYNC();
  // pitch_search (pitch.rs:557-696): x = lp + 360, y = lp, len = 1328, max_pitch = 620
  const float* x = sm->lp + (PLC_LAG_MAX >> 1);
  const float* y = sm->lp;
  const int len = PLC_HIST - PLC_LAG_MAX, max_pitch = PLC_LAG_MAX - PLC_LAG_MIN;
  MB_ROLLED
  for (int i = tm.tid; i < (max_pitch >> 2); i += tm.n) {   // quarter rate: every second sample of the half-rate signals
    float s = 0.f;
    MB_ROLLED
    for (int j = 0; j < (len >> 2); j++) s += x[2 * j] * y[2 * (i + j)];
    sm->xcorr[i] = s;
  }
  MB_SYNC();
  if (tm.tid == 0) plc_find_best_pitch(sm->xcorr, y, 2, len >> 2, max_pitch >> 2, sm->best);
  MB_SYNC();
  const int b0 = sm->best[0], b1 = sm->best[1];
  MB_ROLLED
  for (int i = tm.tid; i < (max_pitch >> 1); i += tm.n) {
    float v = 0.f;
    if (!(abs(i - 2 * b0) > 2 && abs(i - 2 * b1) > 2)) {
      float s = 0.f;
      MB_ROLLED
      for (int j = 0; j < (len >> 1); j++) s += x[j] * y[i + j];
      v = s < -1.0f ? -1.0f : s;
    }
    sm->xcorr[i] = v;
  }
  MB_SYNC();
  if (tm.tid == 0) {
    plc_find_best_pitch(sm->xcorr, y, 1, len >> 1, max_pitch >> 1, sm->best);
    const int b = sm->best[0];
    int res = 2 * b;
    if (b > 0 && b < (max_pitch >> 1) - 1) {
      const float a = sm->xcorr[b - 1], bb = sm->xcorr[b], c = sm->xcorr[b + 1];
      int offset = 0;
      if ((c - a) > 0.7f * (bb - a)) offset = 1;
      else if ((a - c) > 0.7f * (bb - c)) offset = -1;
      res = 2 * b - offset;
    }
    sm->best[0] = PLC_LAG_MAX - res;
  }
  MB_SYNC();
  return sm->best[0];
}

// One channel of the pitch branch of celt_decode_lost (celt_decoder.rs:1625-1750). `fr` points at the lost frame's first
// sample in the channel's linear series: fr[-2048 .. 0) is the history, fr[0 .. 120) the overlap region the previous
// frame left; on return fr[0 .. N + 120) holds the concealed frame and its own overlap region.
MB_DEV_NOINLINE void plc_pitch_channel(Team tm, PlcSmem* sm, float* fr, int N, int pitch_index, int loss_duration, float* lpc_state) {
  const float fade = loss_duration == 0 ? 1.0f : 0.8f;
  const int pi = imin(imin(imax(pitch_index, PLC_LAG_MIN), PLC_LAG_MAX), MAX_PERIOD);
  const int exc_length = imin(2 * pi, MAX_PERIOD);
  float* exc = sm->exc;
  MB_ROLLED
  for (int i = tm.tid; i < MAX_PERIOD + PLC_LPC_ORDER; i += tm.n) exc[i] = fr[OVERLAP - MAX_PERIOD - PLC_LPC_ORDER + i];
  MB_SYNC();
  if (loss_duration == 0) {
    // windowed copy of the last 1024 samples, autocorrelation, lag window, Levinson-Durbin (celt_decoder.rs:1637-1654)
    float* w = sm->lp;
    MB_ROLLED
    for (int i = tm.tid; i < MAX_PERIOD; i += tm.n) {
      float v = exc[PLC_LPC_ORDER + i];
      if (i < OVERLAP) v *= bits2f(kWindow120Bits[i]);
      if (MAX_PERIOD - 1 - i < OVERLAP) v *= bits2f(kWindow120Bits[MAX_PERIOD - 1 - i]);
      w[i] = v;
    }
    MB_SYNC();
    plc_autocorr(tm, w, sm->ac, PLC_LPC_ORDER, MAX_PERIOD);
    if (tm.tid == 0) {
      sm->ac[0] *= 1.0001f;
      for (int i = 1; i <= PLC_LPC_ORDER; i++) { const float f = 0.008f * 0.008f * (float)(i * i); sm->ac[i] -= sm->ac[i] * f; }
      plc_celt_lpc(sm->lpc, sm->ac, PLC_LPC_ORDER);
      for (int i = 0; i < PLC_LPC_ORDER; i++) lpc_state[i] = sm->lpc[i];
    }
  } else {
    for (int i = tm.tid; i < PLC_LPC_ORDER; i += tm.n) sm->lpc[i] = lpc_state[i];
  }
  MB_SYNC();
  const float* lpc = sm->lpc;
  // excitation = history through the analysis filter (celt_decoder.rs:1658-1669)
  const int s0 = MAX_PERIOD - exc_length;
  MB_ROLLED
  for (int i = tm.tid; i < exc_length; i += tm.n) {
    float acc = exc[PLC_LPC_ORDER + s0 + i];
    MB_ROLLED
    for (int t = 0; t < PLC_LPC_ORDER; t++) acc += lpc[t] * exc[PLC_LPC_ORDER + s0 + i - 1 - t];
    sm->work[i] = acc;
  }
  MB_SYNC();
  MB_ROLLED
  for (int i = tm.tid; i < exc_length; i += tm.n) exc[PLC_LPC_ORDER + s0 + i] = sm->work[i];
  MB_SYNC();
  // decay from the energies of the last two half-periods (celt_decoder.rs:1671-1684)
  for (int q = tm.tid; q < 2; q += tm.n) {
    const int dl = exc_length / 2;
    const float* p = exc + PLC_LPC_ORDER + MAX_PERIOD - (q + 1) * dl;
    float e = 1.0f;
    MB_ROLLED
    for (int i = 0; i < dl; i++) e += p[i] * p[i];
    sm->red[q] = e;
  }
  MB_SYNC();
  const float e2 = sm->red[1];
  const float e1 = fmin32(sm->red[0], e2);
  const float decay = celt_sqrt((0.5f * e1) / e2);
  // extrapolation with a decaying gain; s1 = energy of the period it repeats (celt_decoder.rs:1692-1709)
  const int xo = MAX_PERIOD - pi, xl = N + OVERLAP;
  MB_ROLLED
  for (int i = tm.tid; i < xl; i += tm.n) {
    const int wraps = i / pi, j = i - wraps * pi;
    float att = fade * decay;
    MB_ROLLED
    for (int w = 0; w < wraps; w++) att *= decay;
    const float ref = fr[-MAX_PERIOD + xo + j];
    sm->work[i] = ref * ref;
    fr[i] = att * exc[PLC_LPC_ORDER + xo + j];
  }
  MB_SYNC();
  if (tm.tid == 0) {
    sm->red[0] = plc_serial_sum(sm->work, xl);
    // celt_iir (lpc.rs:343-376): the synthesis filter, memory = the last 24 samples of the history, newest first
    float mem[PLC_LPC_ORDER];
    for (int i = 0; i < PLC_LPC_ORDER; i++) mem[i] = fr[-1 - i];
    MB_ROLLED
    for (int i = 0; i < xl; i++) {
      float acc = fr[i];
#pragma unroll
      for (int t = 0; t < PLC_LPC_ORDER; t++) acc -= lpc[t] * mem[t];
#pragma unroll
      for (int t = PLC_LPC_ORDER - 1; t >= 1; t--) mem[t] = mem[t - 1];
      mem[0] = acc;
      fr[i] = acc;
    }
  }
  MB_SYNC();
  MB_ROLLED
  for (int i = tm.tid; i < xl; i += tm.n) { const float v = fr[i]; sm->work[i] = v * v; }
  MB_SYNC();
  if (tm.tid == 0) sm->red[1] = plc_serial_sum(sm->work, xl);
  MB_SYNC();
  const float s1 = sm->red[0], s2 = sm->red[1];
  if (s1 > 0.2f * s2) {
    if (s1 < s2) {
      const float ratio = celt_sqrt((0.5f * s1 + 1.0f) / (s2 + 1.0f));
      MB_ROLLED
      for (int i = tm.tid; i < xl; i += tm.n) fr[i] *= i < OVERLAP ? 1.0f - bits2f(kWindow120Bits[i]) * (1.0f - ratio) : ratio;
    }
  } else {
    MB_ROLLED
    for (int i = tm.tid; i < xl; i += tm.n) fr[i] = 0.f;
  }
  MB_SYNC();
}

// The whole tail of one stream that has a pitch-concealed frame in this pass: history in, frames in tick order (post-filter
// or concealment), history out. Both channels; the series rows carry SER_PREFIX samples of room in front of the first frame.
MB_DEV void celt_comb_plc_lane(Team tm, PlcSmem* sm, uint32_t lane, const TailParams& P) {
  const int CC = P.CC;
  bool any = false;
  const uint32_t n_frames = P.order_n[lane];
  for (uint32_t k = tm.tid; k < n_frames; k += tm.n) any |= (P.meta[P.order[(size_t)k * P.n_lanes + lane]].flags & MB_FL_PLC_PITCH) != 0;
#if MB_ON_DEVICE
  any = __any_sync(0xffffffffu, any);
#endif
  if (!any) return;
  const uint32_t sid = P.lane_stream ? P.lane_stream[lane] : lane;
  StreamState& st = P.st_all[sid];
  float* ser[2];
  for (int c = 0; c < CC; c++) {
    ser[c] = P.series + ((size_t)lane * CC + c) * P.series_stride + SER_PREFIX;
    const float* hist = P.hist + ((size_t)sid * CC + c) * HIST_G;
    MB_ROLLED
    for (int i = tm.tid; i < PLC_HIST; i += tm.n) ser[c][i - PLC_HIST] = hist[i];
  }
  bool first_celt = true;
  uint32_t total = 0;
  // A CELT decoder reset clears the decoder's memory: everything in front of the frame must read as zero from then on. The lane
  // works in place in the series, where the frames in front still wait for the de-emphasis: they are parked in the stream's
  // history buffer (free while the lane runs) and put back before the next reset / at the end.
  int reset_at = -(1 << 30), stash_at = 0;
  bool stashed = false;
  auto unstash = [&]() {
    if (!stashed) return;
    for (int c = 0; c < CC; c++) {
      const float* hs = P.hist + ((size_t)sid * CC + c) * HIST_G;
      MB_ROLLED
      for (int i = tm.tid; i < PLC_HIST; i += tm.n) ser[c][stash_at - PLC_HIST + i] = hs[i];
    }
    stashed = false;
    MB_SYNC();
  };
  MB_SYNC();
  for (uint32_t k = 0; k < n_frames; k++) {
    const uint32_t f = P.order[(size_t)k * P.n_lanes + lane];
    const FrameMeta m = P.meta[f];
    const uint32_t* r = P.rec + rec_index(f, 0);
    const uint32_t h0 = r[S_HDR0 * 32], h1 = r[S_HDR1 * 32];
    const int lm = (int)MB_H0_LM(h0), N = SHORT_MDCT << lm;
    if (m.flags & 4u) {  // CELT decoder reset before this frame
      unstash();
      for (int c = 0; c < CC; c++) {
        float* hs = P.hist + ((size_t)sid * CC + c) * HIST_G;
        float* x = ser[c] + m.sample_off;
        MB_ROLLED
        for (int i = tm.tid; i < PLC_HIST; i += tm.n) { hs[i] = x[i - PLC_HIST]; x[i - PLC_HIST] = 0.f; }
        StreamState::Pf& pf = st.pf[c];
        if (tm.tid == 0) { pf.period = pf.period_old = 0; pf.tapset = pf.tapset_old = 0; pf.gain = pf.gain_old = 0.f; }
      }
      stashed = true; stash_at = (int)m.sample_off; reset_at = (int)m.sample_off;
      MB_SYNC();
    }
    if (m.flags & MB_FL_PLC_PITCH) {
      if (first_celt)   // the overlap region the previous call's last frame left
        for (int c = 0; c < CC; c++)
          for (int i = tm.tid; i < OVERLAP; i += tm.n) ser[c][m.sample_off + i] = P.hist[((size_t)sid * CC + c) * HIST_G + PLC_HIST + i];
      MB_SYNC();
      const int loss_duration = (int)m.seed_in;
      int pitch_index = st.last_pitch_index;
      if (loss_duration == 0) {
        pitch_index = plc_pitch_search(tm, sm, ser[0] + m.sample_off - PLC_HIST, CC == 2 ? ser[1] + m.sample_off - PLC_HIST : nullptr);
        if (tm.tid == 0) st.last_pitch_index = pitch_index;
      }
      // The concealed frame comes with its own overlap region, N .. N + 120 — where the NEXT frame's inverse MDCT output
      // already sits if that frame has one (in the reference the next frame's transform simply overwrites the region later).
      bool next_has_output = false;
      if (k + 1 < n_frames) next_has_output = !(P.meta[P.order[(size_t)(k + 1) * P.n_lanes + lane]].flags & MB_FL_PLC_PITCH);
      for (int c = 0; c < CC; c++) {
        float* fr = ser[c] + m.sample_off;
        if (next_has_output) {
          for (int i = tm.tid; i < OVERLAP; i += tm.n) sm->save[i] = fr[N + i];
          MB_SYNC();
        }
        plc_pitch_channel(tm, sm, fr, N, pitch_index, loss_duration, st.plc_lpc[c]);
        if (next_has_output) {
          for (int i = tm.tid; i < OVERLAP; i += tm.n) fr[N + i] = sm->save[i];
          MB_SYNC();
        }
      }
    } else if (!(m.flags & MB_FL_PLC)) {
      for (int c = 0; c < CC; c++) {
        StreamState::Pf& pf = st.pf[c];
        float* x = ser[c] + m.sample_off;
        const int new_pitch = (int)MB_H1_PITCH(h1), new_tapset = (int)MB_H0_TAPSET(h0);
        const float new_gain = 0.09375f * (float)MB_H0_QG1(h0);
        const int p1 = imax(pf.period, COMB_MINPERIOD), p0 = imax(pf.period_old, COMB_MINPERIOD);
        const float g1 = pf.gain, g0 = pf.gain_old;
        const int t1 = pf.tapset, t0 = pf.tapset_old;
        MB_SYNC();
        const int first = imin(SHORT_MDCT, N);
        const int second = (lm != 0 && first < N) ? N - first : 0;
        team_comb_filter(tm, x, first, p0, p1, g0, g1, imax(t0, 0), imax(t1, 0));
        if (second) team_comb_filter(tm, x + first, second, p1, new_pitch, g1, new_gain, imax(t1, 0), imax(new_tapset, 0));
        MB_SYNC();
        if (tm.tid == 0) {
          pf.period_old = p1; pf.gain_old = g1; pf.tapset_old = t1;
          pf.period = new_pitch; pf.gain = new_gain; pf.tapset = new_tapset;
          if (lm != 0) { pf.period_old = new_pitch; pf.gain_old = new_gain; pf.tapset_old = new_tapset; }
        }
        MB_SYNC();
      }
    }
    first_celt = false;
    total = m.sample_off + (uint32_t)N;
  }
  MB_SYNC();
  unstash();
  for (int c = 0; c < CC; c++) {   // history out: the 2048 samples in front of `total` (zeros in front of a reset) and the overlap region behind it
    float* hist = P.hist + ((size_t)sid * CC + c) * HIST_G;
    MB_ROLLED
    for (int i = tm.tid; i < HIST_G; i += tm.n) {
      const int sidx = (int)total - PLC_HIST + i;
      hist[i] = sidx < reset_at ? 0.f : ser[c][sidx];
    }
  }
}

}  // namespace mb
