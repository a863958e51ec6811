// Phase 2c — time-domain tail, one lane per stream-channel, ticks in order.
//
// Replaces:
//   comb_filter_in_place / comb_filter_const_in_place   src/celt/celt.rs:133-174, 332-443
//   post-filter parameter rotation                       src/celt/celt_decoder.rs:3902-3937, 4011-4021
//   deemphasis (+ stereo fast path)                      src/celt/celt_decoder.rs:2157-2299
//   apply_decode_gain, float2int16                       src/opus_decoder.rs:1376-1418, src/celt/float_cast.rs:38-41
// Both filters are recursive along time (the comb filter reads its own output T samples back, the
// de-emphasis is a one-pole IIR), so the time axis stays serial inside a lane and the batch supplies
// the parallelism. The lane's signal lives in one contiguous time series per channel
// ([HIST history | frames of this call]) so the post-filter history is just a negative offset.
#pragma once
#include "mb_celt_energy.cuh"

namespace mb {

MB_TBL float kTapsetGains[9] = {0.3066406250f, 0.2170410156f, 0.1296386719f, 0.4638671875f, 0.2680664062f, 0.f,
                                0.7998046875f, 0.1000976562f, 0.f};

MB_DEV void comb_const(float* x, int n, int T, float g10, float g11, float g12) {
  float x4 = x[-T - 2], x3 = x[-T - 1], x2 = x[-T], x1 = x[-T + 1];
  for (int i = 0; i < n; i++) {
    float x0 = x[i - T + 2];
    float acc = fma32(g10, x2, x[i]);
    acc = fma32(g11, x1 + x3, acc);
    acc = fma32(g12, x0 + x4, acc);
    x[i] = acc;
    x4 = x3; x3 = x2; x2 = x1; x1 = x0;
  }
}
// x points at the first sample to filter; history is at negative offsets.
MB_DEV void comb_filter(float* x, int n, int T0, int T1, float g0, float g1, int tapset0, int tapset1) {
  if (n == 0) return;
  if (g0 == 0.f && g1 == 0.f) return;
  T0 = imax(T0, COMB_MINPERIOD);
  T1 = imax(T1, COMB_MINPERIOD);
  float g00 = g0 * kTapsetGains[tapset0 * 3], g01 = g0 * kTapsetGains[tapset0 * 3 + 1], g02 = g0 * kTapsetGains[tapset0 * 3 + 2];
  float g10 = g1 * kTapsetGains[tapset1 * 3], g11 = g1 * kTapsetGains[tapset1 * 3 + 1], g12 = g1 * kTapsetGains[tapset1 * 3 + 2];
  float x1 = x[-T1 + 1], x2 = x[-T1], x3 = x[-T1 - 1], x4 = x[-T1 - 2];
  int overlap = imin(OVERLAP, n);
  if (g0 == g1 && T0 == T1 && tapset0 == tapset1) overlap = 0;
  for (int i = 0; i < overlap; i++) {
    float x0 = x[i - T1 + 2];
    float w = bits2f(kWindow120Bits[i]);
    float f = w * w;
    float omf = 1.0f - f;
    float acc = fma32(omf * g00, x[i - T0], x[i]);
    acc = fma32(omf * g01, x[i - T0 + 1] + x[i - T0 - 1], acc);
    acc = fma32(omf * g02, x[i - T0 + 2] + x[i - T0 - 2], acc);
    acc = fma32(f * g10, x2, acc);
    acc = fma32(f * g11, x1 + x3, acc);
    acc = fma32(f * g12, x0 + x4, acc);
    x[i] = acc;
    x4 = x3; x3 = x2; x2 = x1; x1 = x0;
  }
  if (g1 == 0.f) return;
  if (overlap < n) comb_const(x + overlap, n - overlap, T1, g10, g11, g12);
}

struct TailParams {
  const uint32_t* rec;
  const FrameMeta* meta;
  const uint32_t* pkt_map;     // optional grid cell -> packet index (identity when null)
  const uint32_t* lane_stream; // optional lane -> stream
  StreamState* st_all;
  float* hist;                 // [stream][CC][HIST] persistent post-filtered history
  float* series;               // [lane][CC][series_stride]
  uint32_t series_stride;
  uint32_t n_lanes, n_ticks;
  int CC;
  float gain;                  // linear decode gain (1.0 when gain_q8 == 0)
  int apply_gain;
  float* pcm_f32;              // [packet][pcm_stride] interleaved, or null
  uint32_t pcm_stride;         // elements between packets
};

// lane = stream slot, c = channel. The two channels of a stream run in neighbouring threads; each owns
// a private copy of the (identical) post-filter parameters so that no thread reads what another writes.
MB_DEV void celt_tail_lane(uint32_t lane, int c, const TailParams& P) {
  const int CC = P.CC;
  const uint32_t sid = P.lane_stream ? P.lane_stream[lane] : lane;
  StreamState& st = P.st_all[sid];
  float* ser = P.series + ((size_t)lane * CC + c) * P.series_stride;
  float* hist = P.hist + ((size_t)sid * CC + c) * HIST;
  for (int i = 0; i < HIST; i++) ser[i] = hist[i];
  StreamState::Pf& pf = st.pf[c];
  int pf_period = pf.period, pf_period_old = pf.period_old, pf_tapset = pf.tapset, pf_tapset_old = pf.tapset_old;
  float pf_gain = pf.gain, pf_gain_old = pf.gain_old;
  float mem = st.preemph_mem[c];
  uint32_t total = 0;
  for (uint32_t t = 0; t < P.n_ticks; t++) {
    const uint32_t f = t * P.n_lanes + lane;
    const FrameMeta m = P.meta[f];
    if (!(m.flags & 1u)) continue;
    const uint32_t* r = P.rec + rec_index(f, 0);
    const uint32_t h0 = r[S_HDR0 * 32], h1 = r[S_HDR1 * 32];
    const int lm = (int)MB_H0_LM(h0), N = SHORT_MDCT << lm;
    const int new_pitch = (int)MB_H1_PITCH(h1), new_tapset = (int)MB_H0_TAPSET(h0);
    const float new_gain = 0.09375f * (float)MB_H0_QG1(h0);
    float* x = ser + HIST + m.sample_off;
    pf_period = imax(pf_period, COMB_MINPERIOD);
    pf_period_old = imax(pf_period_old, COMB_MINPERIOD);
    int first = imin(SHORT_MDCT, N);
    comb_filter(x, first, pf_period_old, pf_period, pf_gain_old, pf_gain, imax(pf_tapset_old, 0), imax(pf_tapset, 0));
    if (lm != 0 && first < N)
      comb_filter(x + first, N - first, pf_period, new_pitch, pf_gain, new_gain, imax(pf_tapset, 0), imax(new_tapset, 0));
    pf_period_old = pf_period; pf_gain_old = pf_gain; pf_tapset_old = pf_tapset;
    pf_period = new_pitch; pf_gain = new_gain; pf_tapset = new_tapset;
    if (lm != 0) { pf_period_old = pf_period; pf_gain_old = pf_gain; pf_tapset_old = pf_tapset; }
    // de-emphasis + scaling + interleave (celt_decoder.rs:2157-2299)
    const float coef0 = 0.8500061035f, inv = 1.0f / 32768.0f;
    if (P.pcm_f32) {
      const uint32_t p = P.pkt_map ? P.pkt_map[f] : f;
      float* out = P.pcm_f32 + (size_t)p * P.pcm_stride + c;
      for (int j = 0; j < N; j++) {
        float tv = x[j] + mem + 1.0e-30f;
        mem = coef0 * tv;
        float v = tv * inv;
        if (P.apply_gain) v *= P.gain;
        out[(size_t)j * CC] = v;
      }
    } else {
      for (int j = 0; j < N; j++) { float tv = x[j] + mem + 1.0e-30f; mem = coef0 * tv; }
    }
    total = m.sample_off + (uint32_t)N;
  }
  // persist: last HIST post-filtered samples + filter memories
  for (int i = 0; i < HIST; i++) hist[i] = ser[total + i];
  st.preemph_mem[c] = mem;
  pf.period = pf_period; pf.period_old = pf_period_old; pf.tapset = pf_tapset; pf.tapset_old = pf_tapset_old;
  pf.gain = pf_gain; pf.gain_old = pf_gain_old;
}

// float -> int16 (src/celt/float_cast.rs:38-41)
MB_DEV int16_t float2int16(float v) {
  float s = v * 32768.0f;
  s = s < -32768.0f ? -32768.0f : (s > 32767.0f ? 32767.0f : s);
#if MB_ON_DEVICE
  return (int16_t)__float2int_rn(s);
#else
  return (int16_t)lrintf(s);
#endif
}

// opus_pcm_soft_clip_impl (src/opus.rs:16-140), one lane per stream, all its packets of this call in
// order, operating in place on the interleaved float PCM of each packet.
MB_DEV void soft_clip_packet(float* x_, int N, int C, float* declip_mem) {
  if (C < 1 || N < 1) return;
  for (int i = 0; i < N * C; i++) x_[i] = x_[i] < -2.f ? -2.f : (x_[i] > 2.f ? 2.f : x_[i]);
  for (int c = 0; c < C; c++) {
    float* x = x_ + c;
    float a = declip_mem[c];
    for (int i = 0; i < N; i++) {
      if (x[i * C] * a >= 0) break;
      x[i * C] = x[i * C] + a * x[i * C] * x[i * C];
    }
    int curr = 0;
    float x0 = x[0];
    for (;;) {
      int i;
      for (i = curr; i < N; i++)
        if (!(x[i * C] >= -1.f && x[i * C] <= 1.f)) break;
      if (i == N) { a = 0; break; }
      int peak_pos = i, start = i, end = i;
      float maxval = fabsf(x[i * C]);
      while (start > 0 && x[i * C] * x[(start - 1) * C] >= 0) start--;
      while (end < N && x[i * C] * x[end * C] >= 0) {
        if (fabsf(x[end * C]) > maxval) { maxval = fabsf(x[end * C]); peak_pos = end; }
        end++;
      }
      int special = (start == 0 && x[i * C] * x[0] >= 0);
      a = (maxval - 1) / (maxval * maxval);
      a += a * 2.4e-7f;
      if (x[i * C] > 0) a = -a;
      for (i = start; i < end; i++) x[i * C] = x[i * C] + a * x[i * C] * x[i * C];
      if (special && peak_pos >= 2) {
        float offset = x0 - x[0];
        float delta = offset / (float)peak_pos;
        for (i = curr; i < peak_pos; i++) {
          offset -= delta;
          x[i * C] += offset;
          x[i * C] = x[i * C] < -1.f ? -1.f : (x[i * C] > 1.f ? 1.f : x[i * C]);
        }
      }
      curr = end;
      if (curr == N) break;
    }
    declip_mem[c] = a;
  }
}

}  // namespace mb
