// Phase 2c — time-domain tail, ticks in order: (i) post-filter, one warp per stream-channel; (ii) de-emphasis, one
// lane per stream-channel over shared-memory tiles.
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
  const uint32_t* out_map = nullptr;  // see pcm_elem_off
  float* xbuf = nullptr;              // [frame][XBUF_STRIDE]: spectrum rows; their upper half is the cell's side row (below)
  const uint32_t* order = nullptr;    // the lane's CELT frames in decode order (mb_celt_energy.cuh)
  const uint32_t* order_n = nullptr;
  const uint32_t* side_of = nullptr;  // per main cell: the cell whose side row k_fade blends into the packet's row
  const uint32_t* aux_main = nullptr; // per auxiliary cell whose frame belongs in a packet's row: that packet's main cell
  uint32_t aux_base = 0;
  // int16 output path: k_deemph also stores the int16 form of a plain coded CELT-only frame and notes per main cell whether the
  // clipper still has work there (0: not judged, 1: int16 stored and no sample beyond +-1, 2: go through k_clip_i16)
  int16_t* pcm_i16 = nullptr;
  uint32_t* clip_need = nullptr;
  // overlap-add form of the inverse MDCT (quirk Q2 cleared, mb_celt_imdct.cuh): the tails the units stored, [cell][CC][64]
  const float* tails = nullptr;
  int ola = 0;
  int ds = 1;   // output rate 48000 / ds: the de-emphasis filter runs over every sample, every ds-th one is kept (celt_decoder.rs:2232-2297)
};
// the cell whose packet row a frame of cell f is written to
MB_DEV uint32_t row_cell(const TailParams& P, uint32_t f) { return f >= P.n_lanes * P.n_ticks ? P.aux_main[f - P.aux_base] : f; }
// Side row of a cell: 5 ms of audio that is not the packet's own (a redundancy frame), kept for k_fade. It lives in the upper
// half of the cell's spectrum row: such a cell's own CELT frame is at most 5 ms long and only uses the first 2 x 240 floats.
enum { SIDE_OFF = MAX_N };
MB_DEV float* side_row(float* xbuf, uint32_t cell) { return xbuf + (size_t)cell * XBUF_STRIDE + SIDE_OFF; }

// ---------------------------------------------------------------- phase 2c(i): post-filter, one warp per stream-channel
// The comb filter reads its own output T-2 .. T+2 samples back with T >= 15, so a run of min(T0, T1) - 2 >= 13
// consecutive samples has no dependency inside the run: the warp computes such a run in parallel, then the next.
// The channel's signal lives in a shared-memory buffer [HIST history | current frame].
struct CombSmem { float buf[HIST + MAX_N]; };

// x = first sample to filter inside the shared buffer; the history sits at negative offsets.
MB_DEV void team_comb_filter(Team tm, float* x, int n, int T0, int T1, float g0, float g1, int tapset0, int tapset1) {
  if (n == 0) return;
  if (g0 == 0.f && g1 == 0.f) return;
  T0 = imax(T0, COMB_MINPERIOD);
  T1 = imax(T1, COMB_MINPERIOD);
  const float g00 = g0 * kTapsetGains[tapset0 * 3], g01 = g0 * kTapsetGains[tapset0 * 3 + 1], g02 = g0 * kTapsetGains[tapset0 * 3 + 2];
  const float g10 = g1 * kTapsetGains[tapset1 * 3], g11 = g1 * kTapsetGains[tapset1 * 3 + 1], g12 = g1 * kTapsetGains[tapset1 * 3 + 2];
  int overlap = imin(OVERLAP, n);
  if (g0 == g1 && T0 == T1 && tapset0 == tapset1) overlap = 0;
  const int run0 = imin(T0, T1) - 2;
  for (int s = 0; s < overlap; s += run0) {
    const int e = imin(s + run0, overlap);
    for (int i = s + tm.tid; i < e; i += tm.n) {
      const float w = bits2f(kWindow120Bits[i]);
      const float f = w * w;
      const float omf = 1.0f - f;
      float acc = fma32(omf * g00, x[i - T0], x[i]);
      acc = fma32(omf * g01, x[i - T0 + 1] + x[i - T0 - 1], acc);
      acc = fma32(omf * g02, x[i - T0 + 2] + x[i - T0 - 2], acc);
      acc = fma32(f * g10, x[i - T1], acc);
      acc = fma32(f * g11, x[i - T1 + 1] + x[i - T1 - 1], acc);
      acc = fma32(f * g12, x[i - T1 + 2] + x[i - T1 - 2], acc);
      x[i] = acc;
    }
    MB_SYNC();
  }
  if (g1 == 0.f) return;
  const int run1 = T1 - 2;
  for (int s = overlap; s < n; s += run1) {
    const int e = imin(s + run1, n);
    for (int i = s + tm.tid; i < e; i += tm.n) {
      float acc = fma32(g10, x[i - T1], x[i]);
      acc = fma32(g11, x[i - T1 + 1] + x[i - T1 - 1], acc);
      acc = fma32(g12, x[i - T1 + 2] + x[i - T1 - 2], acc);
      x[i] = acc;
    }
    MB_SYNC();
  }
}

MB_DEV void celt_comb_channel(Team tm, CombSmem* sm, uint32_t lane, int c, const TailParams& P) {
  const int CC = P.CC;
  {  // a lane with a frame concealed by pitch repetition belongs to k_comb_plc (mb_celt_plc.cuh)
    bool any = false;
    for (uint32_t k = tm.tid; k < P.order_n[lane]; k += tm.n) any |= (P.meta[P.order[(size_t)k * P.n_lanes + lane]].flags & MB_FL_PLC_PITCH) != 0;
#if MB_ON_DEVICE
    any = __any_sync(0xffffffffu, any);
#endif
    if (any) return;
  }
  const uint32_t sid = P.lane_stream ? P.lane_stream[lane] : lane;
  StreamState& st = P.st_all[sid];
  float* ser = P.series + ((size_t)lane * CC + c) * P.series_stride + SER_PREFIX;
  float* hist = P.hist + ((size_t)sid * CC + c) * HIST_G;
  float* buf = sm->buf;
  for (int i = tm.tid; i < HIST; i += tm.n) buf[i] = hist[HIST_G - OVERLAP - HIST + i];
  StreamState::Pf& pf = st.pf[c];
  int pf_period = pf.period, pf_period_old = pf.period_old, pf_tapset = pf.tapset, pf_tapset_old = pf.tapset_old;
  float pf_gain = pf.gain, pf_gain_old = pf.gain_old;
  int total = 0, reset_at = -1;   // samples of CELT signal this pass; where the last decoder reset fell
  // overlap-add form: the 60 unwindowed samples the previous frame's inverse MDCT left behind it — the persisted overlap region
  // for the first frame of the pass, nothing after a decoder reset, else the previous unit's entry in the tail store
  const float* prev_tail = hist + HIST_G - OVERLAP;
  bool tail_zero = false;
  MB_SYNC();
  const uint32_t n_frames = P.order_n[lane];
  for (uint32_t k = 0; k < n_frames; k++) {
    const uint32_t f = P.order[(size_t)k * P.n_lanes + lane];
    const FrameMeta m = P.meta[f];
    const uint32_t* r = P.rec + rec_index(f, 0);
    const uint32_t h0 = r[S_HDR0 * 32], h1 = r[S_HDR1 * 32];
    const int lm = (int)MB_H0_LM(h0), N = SHORT_MDCT << lm;
    if (m.flags & 4u) {  // CELT decoder reset before this frame (SILK-only -> hybrid)
      for (int i = tm.tid; i < HIST; i += tm.n) buf[i] = 0.f;
      pf_period = pf_period_old = 0; pf_tapset = pf_tapset_old = 0; pf_gain = pf_gain_old = 0.f;
      reset_at = (int)m.sample_off;
      tail_zero = true;
      MB_SYNC();
    }
    float* xg = ser + m.sample_off;
    float* x = buf + HIST;
    for (int i = tm.tid; i < N; i += tm.n) x[i] = xg[i];
    MB_SYNC();
    if (P.ola) {
      // out[i] = x2*w2 - x1*w1, out[119-i] = x2*w1 + x1*w2 (clt_mdct_backward, mdct.rs:376-388): the unit left -x1*w1 and x1*w2
      // (x2 = 0); adding the x2 products gives the same two roundings. Then the clamp the unit held back for these samples.
      const float SIG_SAT = 536870911.0f;
      for (int i = tm.tid; i < OVERLAP / 2; i += tm.n) {
        const float t = tail_zero ? 0.f : prev_tail[i];
        const float w1 = bits2f(kWindow120Bits[i]), w2 = bits2f(kWindow120Bits[OVERLAP - 1 - i]);
        const float a = t * w2 + x[i], b = t * w1 + x[OVERLAP - 1 - i];
        x[i] = a < -SIG_SAT ? -SIG_SAT : (a > SIG_SAT ? SIG_SAT : a);
        x[OVERLAP - 1 - i] = b < -SIG_SAT ? -SIG_SAT : (b > SIG_SAT ? SIG_SAT : b);
      }
      prev_tail = P.tails + ((size_t)f * CC + c) * 64;
      tail_zero = false;
      MB_SYNC();
    }
    if (!(m.flags & MB_FL_PLC)) {   // (a concealed frame is neither post-filtered nor does it move the filter's parameters)
      const int new_pitch = (int)MB_H1_PITCH(h1), new_tapset = (int)MB_H0_TAPSET(h0);
      const float new_gain = 0.09375f * (float)MB_H0_QG1(h0);
      pf_period = imax(pf_period, COMB_MINPERIOD);
      pf_period_old = imax(pf_period_old, COMB_MINPERIOD);
      const int first = imin(SHORT_MDCT, N);
      const int second = (lm != 0 && first < N) ? N - first : 0;
      const int touched = P.ola || (pf_gain_old != 0.f || pf_gain != 0.f) || (second && (pf_gain != 0.f || new_gain != 0.f));
      team_comb_filter(tm, x, first, pf_period_old, pf_period, pf_gain_old, pf_gain, imax(pf_tapset_old, 0), imax(pf_tapset, 0));
      if (second)
        team_comb_filter(tm, x + first, second, pf_period, new_pitch, pf_gain, new_gain, imax(pf_tapset, 0), imax(new_tapset, 0));
      pf_period_old = pf_period; pf_gain_old = pf_gain; pf_tapset_old = pf_tapset;
      pf_period = new_pitch; pf_gain = new_gain; pf_tapset = new_tapset;
      if (lm != 0) { pf_period_old = pf_period; pf_gain_old = pf_gain; pf_tapset_old = pf_tapset; }
      if (touched)
        for (int i = tm.tid; i < N; i += tm.n) xg[i] = x[i];
    }
    total = (int)m.sample_off + N;
    // slide the window: the last HIST samples become the history of the next frame
    for (int i0 = 0; i0 < HIST; i0 += tm.n) {
      const int i = i0 + tm.tid;
      float v = 0.f;
      if (i < HIST) v = buf[i + N];
      MB_SYNC();
      if (i < HIST) buf[i] = v;
    }
    MB_SYNC();
  }
  // Persistent history, laid out like the reference's decode_mem: the 2048 samples in front of `total` (this pass's series,
  // older ones from the previous history, zeros in front of a decoder reset), then the overlap region the last frame's
  // inverse MDCT left behind it (MB_FL_KEEP_TAIL). In place: ascending chunks, read before write.
  if (total > 0) {
    for (int i0 = 0; i0 < HIST_G; i0 += tm.n) {
      const int i = i0 + tm.tid;
      float v = 0.f;
      if (i < HIST_G) {
        const int s = total - (HIST_G - OVERLAP) + i;
        if (P.ola && s >= total) v = s - total < OVERLAP / 2 ? prev_tail[s - total] : 0.f;   // (the last unit's tail; zeros behind it)
        else if (s >= 0) v = s < reset_at ? 0.f : ser[s];
        else v = reset_at >= 0 ? 0.f : hist[i + total];
      }
      MB_SYNC();
      if (i < HIST_G) hist[i] = v;
    }
  }
  if (tm.tid == 0) {
    pf.period = pf_period; pf.period_old = pf_period_old; pf.tapset = pf_tapset; pf.tapset_old = pf_tapset_old;
    pf.gain = pf_gain; pf.gain_old = pf_gain_old;
  }
}

// ---------------------------------------------------------------- phase 2c(ii): de-emphasis, one lane per stream-channel
// deemphasis (celt_decoder.rs:2157-2299): y = x + mem + 1e-30; mem = 0.8500061 * y; out = y / 32768 (* gain).
struct Deemph {
  float mem, gain;
  int apply_gain;
  // prev/accum: hybrid frames add the CELT signal to the SILK samples already in the row (celt_decoder.rs:2261-2266);
  // the decode gain then applies to the sum (opus_decoder.rs:1376-1392)
  MB_MEM float step(float xv, float prev = 0.f, bool accum = false) {
    const float tv = xv + mem + 1.0e-30f;
    mem = 0.8500061035f * tv;
    float v = tv * (1.0f / 32768.0f);
    if (accum) v = prev + v;
    if (apply_gain) v *= gain;
    return v;
  }
};

// Plain per-lane form (host emulation, and the reference shape of the device tile code in mb_api.cu).
MB_DEV void celt_deemph_lane(uint32_t lane, int c, const TailParams& P) {
  const int CC = P.CC;
  const uint32_t sid = P.lane_stream ? P.lane_stream[lane] : lane;
  StreamState& st = P.st_all[sid];
  const float* ser = P.series + ((size_t)lane * CC + c) * P.series_stride + SER_PREFIX;
  Deemph d = {st.preemph_mem[c], P.gain, P.apply_gain};
  const uint32_t n_frames = P.order_n[lane];
  for (uint32_t k = 0; k < n_frames; k++) {
    const uint32_t f = P.order[(size_t)k * P.n_lanes + lane];
    const FrameMeta m = P.meta[f];
    const int N = SHORT_MDCT << (int)MB_H0_LM(P.rec[rec_index(f, S_HDR0)]);
    const float* x = ser + m.sample_off;
    const bool side = (m.flags & 16u) != 0;       // not the packet's audio (redundancy frame): kept aside for celt_fade_cell
    float* out;
    if (side) out = side_row(P.xbuf, f) + c;
    else {
      const uint32_t fo = row_cell(P, f);
      const uint32_t p = P.pkt_map ? P.pkt_map[fo] : fo;
      out = P.pcm_f32 ? P.pcm_f32 + pcm_elem_off(P.out_map, p, P.pcm_stride) + c : nullptr;
    }
    if (m.flags & 4u) d.mem = 0.f;
    const bool accum = (m.flags & 8u) != 0;
    // the decode gain applies once to the finished row: here, unless the row still waits for a cross-fade (k_fade applies it)
    d.apply_gain = (side || (m.flags & MB_FL_FADE_ROW)) ? 0 : P.apply_gain;
    if (P.ds > 1) {   // (hybrid: the kept samples add onto the SILK samples already in the row, celt_decoder.rs:2281-2290)
      for (int j = 0; j < N; j++) {
        if (out && j % P.ds == 0) {
          float* o = out + (size_t)(j / P.ds) * CC;
          *o = d.step(x[j], accum ? *o : 0.f, accum);
        } else d.step(x[j]);
      }
      continue;
    }
    for (int j = 0; j < N; j++) {
      const float v = d.step(x[j], accum && out ? out[(size_t)j * CC] : 0.f, accum);
      if (out) out[(size_t)j * CC] = v;
    }
  }
  st.preemph_mem[c] = d.mem;
}

// Redundancy cross-fades (opus_decoder.rs:1048-1080, smooth_fade :100-135) for channel c of cell f, after the de-emphasis:
// the 5 ms CELT frame kept in the cell's spectrum row replaces / blends into the SILK audio of the row, then the decode
// gain is applied to the finished row. w = window[i]^2 over 120 samples (48 kHz).
MB_DEV void celt_fade_cell(uint32_t f, int c, const TailParams& P) {
  const FrameMeta m = P.meta[f];
  // a lost packet before anything was decoded: silence (opus_decoder.rs:527-540). (Neither CELT nor SILK part: a lost packet of a
  // SILK-only stream carries bit 1 and whatever seed the stream's CELT history left — it is the k_silk lane's, not a zero fill.)
  if ((m.flags & MB_FL_PLC) && !(m.flags & 3u) && P.pcm_f32) {
    const uint32_t p0 = P.pkt_map ? P.pkt_map[f] : f;
    float* row0 = P.pcm_f32 + pcm_elem_off(P.out_map, p0, P.pcm_stride) + c;
    for (int i = 0; i < (int)m.seed_in; i++) row0[(size_t)i * P.CC] = 0.f;
    return;
  }
  if (f >= P.n_lanes * P.n_ticks) return;      // (auxiliary cells have no row of their own)
  const uint32_t src = P.side_of[f];
  if (src == 0xffffffffu || !P.pcm_f32) return;
  const int CC = P.CC;
  // Below 48 kHz (row F4) the rows hold every ds-th sample: 2.5 ms are F = 120 / ds samples and smooth_fade steps through the
  // window by ds (opus_decoder.rs:100-135: inc = 48000 / Fs).
  const int ds = P.ds, F = SHORT_MDCT / ds;
  const uint32_t p = P.pkt_map ? P.pkt_map[f] : f;
  float* row = P.pcm_f32 + pcm_elem_off(P.out_map, p, P.pcm_stride) + c;
  const float* red = side_row(P.xbuf, src) + c;
  if ((m.flags & 1u) && (m.flags & 512u) && (m.flags & MB_FL_TRANS_IN) && !(m.flags & (8u | 16u))) {
    // SILK-only / hybrid -> CELT-only without a redundancy frame (opus_decoder.rs:1083-1112): the audio concealed in the old mode
    // opens the frame and fades into it (frames of 5 ms and more), or fades into its first 2.5 ms (2.5 ms frames)
    const int N = (SHORT_MDCT << (int)MB_H0_LM(P.rec[rec_index(f, S_HDR0)])) / ds;
    const int at = N >= 2 * F ? F : 0;
    for (int i = 0; i < at; i++) row[(size_t)i * CC] = red[(size_t)i * CC];
    for (int i = 0; i < F; i++) {
      const float wv = bits2f(kWindow120Bits[i * ds]), w = wv * wv;
      float* o = row + (size_t)(at + i) * CC;
      *o = w * *o + (1.0f - w) * red[(size_t)(at + i) * CC];
    }
    if (P.apply_gain)
      for (int i = 0; i < N; i++) row[(size_t)i * CC] *= P.gain;
    return;
  }
  const uint32_t status = P.rec[rec_index(f, S_STATUS)];   // (SILK header: in the hybrid layout, or where 40 / 60 ms packets keep it)
  const uint32_t silk_h0 = P.rec[rec_index(f, ((status >> 17) & 3u) ? S_SILK_HYB : S_HDR0)];
  const int N = (int)((silk_h0 >> 8) & 3u) * (int)((silk_h0 >> 5) & 7u) * 240 / ds;   // samples of the finished row (the SILK side's)
  if (m.flags & 128u) {          // SILK -> CELT: the end of the row fades into the second half of the redundancy frame
    for (int i = 0; i < F; i++) {
      const float wv = bits2f(kWindow120Bits[i * ds]), w = wv * wv;
      float* o = row + (size_t)(N - F + i) * CC;
      *o = w * red[(size_t)(F + i) * CC] + (1.0f - w) * *o;
    }
  } else if (m.flags & 64u) {    // CELT -> SILK: the redundancy frame opens the row and fades into the SILK audio
    for (int i = 0; i < F; i++) row[(size_t)i * CC] = red[(size_t)i * CC];
    for (int i = 0; i < F; i++) {
      const float wv = bits2f(kWindow120Bits[i * ds]), w = wv * wv;
      float* o = row + (size_t)(F + i) * CC;
      *o = w * *o + (1.0f - w) * red[(size_t)(F + i) * CC];
    }
  }
  if (P.apply_gain)
    for (int i = 0; i < N; i++) row[(size_t)i * CC] *= P.gain;
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

// opus_pcm_soft_clip_impl (src/opus.rs:16-140) for one channel of one packet, in place on interleaved float PCM
// (x = first sample of the channel, C = interleave stride); `a` is the channel's declip_mem. The +-2 clamp of the
// whole packet has to be applied before (soft_clip_packet / the device kernel do it in parallel).
MB_DEV void soft_clip_channel(float* x, int N, int C, float& a_io) {
  float a = a_io;
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
  a_io = a;
}

MB_DEV void soft_clip_packet(float* x_, int N, int C, float* declip_mem) {
  if (C < 1 || N < 1) return;
  for (int i = 0; i < N * C; i++) x_[i] = x_[i] < -2.f ? -2.f : (x_[i] > 2.f ? 2.f : x_[i]);
  for (int c = 0; c < C; c++) soft_clip_channel(x_ + c, N, C, declip_mem[c]);
}

}  // namespace mb
