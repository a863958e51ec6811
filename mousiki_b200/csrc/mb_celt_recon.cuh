// Phase 2b — spectrum reconstruction + inverse MDCT, one thread team (CTA) per frame,
// every frame of every stream concurrently.
//
// Replaces the float half of the reference's CELT frame decoder:
//   decode_pulses/cwrsi + alg_unquant + normalise_residual + exp_rotation + extract_collapse_mask
//                                      src/celt/cwrs.rs:293-452, src/celt/vq.rs:116-221,332-355,1013-1037,1164-1192
//   quant_all_bands resynthesis (folding, noise fill, haar1, hadamard (de)interleave, stereo_merge)
//                                      src/celt/bands.rs:1917-3082, 3477-3505, 3708-3822
//   anti_collapse, denormalise_bands   src/celt/bands.rs:3220-3349, 4172-4234
//   celt_synthesis / clt_mdct_backward / kiss_fft butterflies
//                                      src/celt/celt_decoder.rs:473-800, src/celt/mdct.rs:197-388,
//                                      src/celt/kiss_fft.rs:748-795, 900-1340
// Stage A decodes every PVQ leaf of the frame in parallel (one thread per leaf: the codeword walk
// and the spreading rotation are serial inside a leaf but leaves are independent). Stage B replays
// the band/split tree in order (folding needs earlier bands and the LCG seed) with all threads
// executing the same control flow and sharing the vector work. Stage C is a shared-memory
// mixed-radix FFT (radix order 4,2,4,3,5 for 480 points, as the static mode prescribes) with
// pre/post rotation and the window fold.
#pragma once
#include "mb_celt_energy.cuh"
#include "mb_celt_parse.cuh"

namespace mb {

MB_TBL float kEMeans[25] = {6.4375f, 6.25f, 5.75f, 5.3125f, 5.0625f, 4.8125f, 4.5f, 4.375f, 4.875f, 4.6875f, 4.5625f, 4.4375f, 4.875f,
                            4.625f, 4.3125f, 4.5f, 4.375f, 4.625f, 4.75f, 4.4375f, 3.75f, 3.75f, 3.75f, 3.75f, 3.75f};
MB_TBL uint8_t kOrdery[30] = {1, 0, 3, 0, 2, 1, 7, 0, 4, 3, 6, 1, 5, 2, 15, 0, 8, 7, 12, 3, 11, 4, 14, 1, 9, 6, 13, 2, 10, 5};
MB_TBL uint8_t kBitInterleave[16] = {0, 1, 1, 1, 2, 3, 3, 3, 2, 3, 3, 3, 2, 3, 3, 3};
MB_TBL uint8_t kBitDeinterleave[16] = {0x00, 0x03, 0x0C, 0x0F, 0x30, 0x33, 0x3C, 0x3F, 0xC0, 0xC3, 0xCC, 0xCF, 0xF0, 0xF3, 0xFC, 0xFF};
MB_TBL uint8_t kFftFactors[4][10] = {{5, 96, 3, 32, 4, 8, 2, 4, 4, 1}, {5, 48, 3, 16, 4, 4, 4, 1, 0, 0}, {5, 24, 3, 8, 2, 4, 4, 1, 0, 0}, {5, 12, 3, 4, 4, 1, 0, 0, 0, 0}};
MB_TBL uint16_t kMdctTwOff[4] = {0, 960, 1440, 1680};
// band of a 2.5 ms-resolution bin (bins 0..99; eBands5ms boundaries)
MB_TBL uint8_t kBandOfBin[100] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 12, 12, 13, 13, 13, 13, 14, 14, 14, 14, 15, 15, 15, 15, 15, 15,
                                  16, 16, 16, 16, 16, 16, 17, 17, 17, 17, 17, 17, 17, 17, 18, 18, 18, 18, 18, 18, 18, 18, 18, 18, 18, 18,
                                  19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19, 19,
                                  20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20, 20};

// Shared-memory working set of one frame (floats unless noted). ~24 KB.
struct ReconSmem {
  float X[2 * MAX_N];          // normalised spectrum, then freq
  float W[2 * MAX_N];          // norm / norm2 fold buffers (2 x 800) during stage B, FFT work (2 x 480 complex) in stage C
  float tmp[208];              // hadamard / lowband scratch
  float lowband_scratch[176];
  uint16_t leaf_cm[CAP_LEAVES];
  float red[4];                // scalar broadcast slots
  uint8_t collapse[2 * NB_EBANDS + 6];
};

struct Team { int tid, n; };

// ---------------------------------------------------------------- stage A: one PVQ leaf (one thread)
// cwrsi (cwrs.rs:293-406) writing pulses straight into the spectrum as floats (|v| <= 128, exact).
MB_DEV float pvq_decode_leaf(int n, int k, uint32_t i, float* y) {
  float yy = 0.f;
  int o = 0;
  while (n > 2) {
    int val;
    if (k >= n) {
      uint32_t p = pvq_u(n, k + 1);
      int neg = i >= p;
      if (neg) i -= p;
      int k0 = k;
      uint32_t q = pvq_u(n, n);
      if (q > i) {
        k = n;
        do p = pvq_u(--k, n); while (p > i);
      } else {
        for (p = pvq_u(n, k); p > i; p = pvq_u(n, k)) k--;
      }
      i -= p;
      val = neg ? -(k0 - k) : (k0 - k);
    } else {
      uint32_t p = pvq_u(k, n), q = pvq_u(k + 1, n);
      if (p <= i && i < q) { i -= p; val = 0; }
      else {
        int neg = i >= q;
        if (neg) i -= q;
        int k0 = k;
        do p = pvq_u(--k, n); while (p > i);
        i -= p;
        val = neg ? -(k0 - k) : (k0 - k);
      }
    }
    y[o++] = (float)val;
    yy += (float)val * (float)val;
    n--;
  }
  uint32_t p = 2u * (uint32_t)k + 1u;
  int neg = i >= p;
  if (neg) i -= p;
  int k0 = k;
  k = (int)((i + 1u) >> 1);
  if (k) i -= 2u * (uint32_t)k - 1u;
  int val = neg ? -(k0 - k) : (k0 - k);
  y[o++] = (float)val;
  yy += (float)val * (float)val;
  val = i == 0 ? k : -k;
  y[o] = (float)val;
  yy += (float)val * (float)val;
  return yy;
}

MB_DEV void exp_rotation1(float* x, int len, int stride, float c, float s) {
  if (stride == 0 || len <= stride) return;
  float ms = -s;
  for (int i = 0; i < len - stride; i++) {
    float x1 = x[i], x2 = x[i + stride];
    x[i + stride] = c * x2 + s * x1;
    x[i] = c * x1 + ms * x2;
  }
  for (int i = len - 2 * stride - 1; i >= 0; i--) {
    float x1 = x[i], x2 = x[i + stride];
    x[i + stride] = c * x2 + s * x1;
    x[i] = c * x1 + ms * x2;
  }
}

// alg_unquant for one leaf: pulses -> unit-norm * gain -> inverse spreading rotation; returns collapse mask.
MB_DEV uint32_t pvq_leaf(float* x, int n, int k, int B, float gain, int spread, uint32_t index, int32_t* iy_dbg) {
  float ryy = pvq_decode_leaf(n, k, index, x);
  if (iy_dbg) for (int i = 0; i < n; i++) iy_dbg[i] = (int32_t)x[i];
  uint32_t mask = 1;
  if (B > 1) {  // extract_collapse_mask on the integer pulses
    int n0 = n / B;
    mask = 0;
    for (int i = 0; i < B; i++) {
      int any = 0;
      for (int j = 0; j < n0; j++) any |= x[i * n0 + j] != 0.f;
      mask |= (uint32_t)any << i;
    }
  }
  float scale = celt_rsqrt(ryy) * gain;
  for (int i = 0; i < n; i++) x[i] = scale * x[i];
  // exp_rotation(x, n, -1, B, k, spread)  (vq.rs:157-221)
  if (!(2 * k >= n || spread == 0)) {
    int factor = spread == 1 ? 15 : (spread == 2 ? 10 : 5);
    float g = (1.0f * (float)n) / (float)(n + factor * k);
    float theta = 0.5f * g * g;
    float c = celt_cos_norm(theta), s = celt_cos_norm(1.0f - theta);
    int stride2 = 0;
    if (n >= 8 * B) {
      stride2 = 1;
      while ((stride2 * stride2 + stride2) * B + (B >> 2) < n) stride2++;
    }
    int ld = n / B;
    for (int i = 0; i < B; i++) {
      float* b = x + i * ld;
      if (stride2) exp_rotation1(b, ld, stride2, s, c);
      exp_rotation1(b, ld, 1, c, s);
    }
  }
  return mask;
}

// ---------------------------------------------------------------- stage B helpers (team-wide, uniform control flow)
struct Walk {
  Team tm;
  ReconSmem* sm;
  uint32_t seed;
  int node_pos, leaf_pos;
  int tf_change, spread;
  const uint32_t* nodes;     // node stream of this frame in the record tile (word i at nodes[i * 32])
  int n_nodes;
  uint32_t cache0, cache1;   // device: lane l holds node (base + l) and (base + 32 + l) of the current window
};

// Next word of the node stream (consumed in execution order by every lane at once). On the device the warp keeps a
// 64-node window in two registers per lane and broadcasts by shuffle; the next window is fetched one window ahead.
MB_DEV uint32_t next_node(Walk& w) {
  const int p = w.node_pos++;
#if MB_ON_DEVICE
  const int slot = p & 31;
  if (slot == 0 && p != 0) {
    w.cache0 = w.cache1;
    const int q = p + 32 + w.tm.tid;
    w.cache1 = q < w.n_nodes ? __ldg(w.nodes + (size_t)q * 32) : 0u;
  }
  return __shfl_sync(0xffffffffu, w.cache0, slot);
#else
  return w.nodes[(size_t)p * 32];
#endif
}
MB_DEV void open_nodes(Walk& w, const uint32_t* nodes, int n_nodes) {
  w.nodes = nodes; w.n_nodes = n_nodes; w.node_pos = 0; w.cache0 = w.cache1 = 0;
#if MB_ON_DEVICE
  if (w.tm.tid < n_nodes) w.cache0 = __ldg(nodes + (size_t)w.tm.tid * 32);
  if (w.tm.tid + 32 < n_nodes) w.cache1 = __ldg(nodes + (size_t)(w.tm.tid + 32) * 32);
#endif
}

MB_DEV uint32_t lcg_jump(uint32_t seed, int steps) { return kLcgA[steps] * seed + kLcgC[steps]; }

MB_DEV void team_haar1(Walk& w, float* x, int n0, int stride) {
  if (stride == 0 || n0 < 2) return;
  int half = n0 / 2;
  const float scale = 0.70710678f;
  int total = stride * half;
  for (int idx = w.tm.tid; idx < total; idx += w.tm.n) {
    int i = idx / half, j = idx - i * half;
    int i0 = stride * 2 * j + i, i1 = i0 + stride;
    float t1 = scale * x[i0], t2 = scale * x[i1];
    x[i0] = t1 + t2;
    x[i1] = t1 - t2;
  }
  MB_SYNC();
}
MB_DEV const uint8_t* ordery_for(int stride) { return kOrdery + (stride == 2 ? 0 : stride == 4 ? 2 : stride == 8 ? 6 : 14); }
// x (n0*stride elements) is permuted through sm->tmp
MB_DEV void team_deinterleave(Walk& w, float* x, int n0, int stride, int hadamard) {
  int n = n0 * stride;
  if (stride == 0 || n == 0) return;
  float* tmp = w.sm->tmp;
  const uint8_t* ord = ordery_for(stride);
  for (int idx = w.tm.tid; idx < n; idx += w.tm.n) {
    int i = idx / n0, j = idx - i * n0;
    int dst = (hadamard ? (int)ord[i] : i) * n0 + j;
    tmp[dst] = x[j * stride + i];
  }
  MB_SYNC();
  for (int idx = w.tm.tid; idx < n; idx += w.tm.n) x[idx] = tmp[idx];
  MB_SYNC();
}
MB_DEV void team_interleave(Walk& w, float* x, int n0, int stride, int hadamard) {
  int n = n0 * stride;
  if (stride == 0 || n == 0) return;
  float* tmp = w.sm->tmp;
  const uint8_t* ord = ordery_for(stride);
  for (int idx = w.tm.tid; idx < n; idx += w.tm.n) {
    int i = idx / n0, j = idx - i * n0;
    int src = (hadamard ? (int)ord[i] : i) * n0 + j;
    tmp[j * stride + i] = x[src];
  }
  MB_SYNC();
  for (int idx = w.tm.tid; idx < n; idx += w.tm.n) x[idx] = tmp[idx];
  MB_SYNC();
}
// renormalise_vector (vq.rs:1073-1088): the energy sum keeps the reference's left-to-right order
MB_DEV void team_renormalise(Walk& w, float* x, int n, float gain) {
  MB_SYNC();
  if (w.tm.tid == 0) {
    float e = 0.f;
    for (int i = 0; i < n; i++) e += x[i] * x[i];
    e = 1e-15f + e;
    w.sm->red[0] = celt_rsqrt(e) * gain;
  }
  MB_SYNC();
  float scale = w.sm->red[0];
  for (int i = w.tm.tid; i < n; i += w.tm.n) x[i] *= scale;
  MB_SYNC();
}

// quant_partition, resynthesis side. Consumes the node stream in execution order.
MB_DEV_NOINLINE uint32_t recon_partition(Walk& w, float* x, int n, int B, float* lowband, int lm, float gain, uint32_t fill) {
  struct Frame { float* x; float* lb; int n, B, B0, lm; float gmid, gside; uint32_t fill, cm; uint8_t order, st; };
  Frame st[6];
  int sp = 0;
  st[0].x = x; st[0].lb = lowband; st[0].n = n; st[0].B = B; st[0].lm = lm; st[0].gmid = gain; st[0].fill = fill; st[0].st = 0;
  uint32_t ret = 0;
  while (sp >= 0) {
    Frame& F = st[sp];
    if (F.st == 0) {
      uint32_t node = next_node(w);
      uint32_t type = MB_ND_TYPE(node);
      if (type == ND_SPLIT) {
        int half = F.n >> 1, itheta = (int)MB_ND_ITHETA(node);
        F.B0 = F.B;
        if (F.B == 1) F.fill = (F.fill & 1u) | (F.fill << 1);
        F.B = (F.B + 1) >> 1;
        uint32_t bmask = mask_from_bits(F.B);
        int imid, iside;
        if (itheta == 0) { imid = 32767; iside = 0; F.fill &= bmask; }
        else if (itheta == 16384) { imid = 0; iside = 32767; F.fill &= F.B >= 32 ? 0xffffffffu : bmask << F.B; }
        else { imid = bitexact_cos(itheta); iside = bitexact_cos(16384 - itheta); }
        float g = F.gmid;
        F.gmid = g * ((float)imid / 32768.0f);
        F.gside = g * ((float)iside / 32768.0f);
        F.order = (uint8_t)MB_ND_ORDER(node);
        F.st = 1;
        Frame& Cn = st[sp + 1];
        Cn.n = half; Cn.B = F.B; Cn.lm = F.lm - 1; Cn.st = 0;
        if (F.order) { Cn.x = F.x; Cn.lb = F.lb; Cn.gmid = F.gmid; Cn.fill = F.fill; }
        else { Cn.x = F.x + half; Cn.lb = F.lb ? F.lb + half : nullptr; Cn.gmid = F.gside; Cn.fill = F.fill >> F.B; }
        sp++;
      } else {
        uint32_t cm = 0;
        if (type == ND_LEAF_PULSE) {
          cm = w.sm->leaf_cm[w.leaf_pos++];
        } else {  // no pulses: fold / noise fill (bands.rs:2104-2127)
          uint32_t cm_mask = mask_from_bits(F.B);
          uint32_t fl = F.fill & cm_mask;
          float* xx = F.x;
          int nn = F.n;
          if (!fl) {
            for (int j = w.tm.tid; j < nn; j += w.tm.n) xx[j] = 0.f;
            MB_SYNC();
          } else if (F.lb) {
            const float* lb = F.lb;
            for (int j = w.tm.tid; j < nn; j += w.tm.n) {
              uint32_t s = lcg_jump(w.seed, j + 1);
              xx[j] = lb[j] + ((s & 0x8000u) ? 1.0f / 256.0f : -(1.0f / 256.0f));
            }
            w.seed = lcg_jump(w.seed, nn);
            cm = fl;
            team_renormalise(w, xx, nn, F.gmid);
          } else {
            for (int j = w.tm.tid; j < nn; j += w.tm.n) {
              uint32_t s = lcg_jump(w.seed, j + 1);
              xx[j] = (float)((int32_t)s >> 20);
            }
            w.seed = lcg_jump(w.seed, nn);
            cm = cm_mask;
            team_renormalise(w, xx, nn, F.gmid);
          }
        }
        ret = cm;
        sp--;
      }
    } else if (F.st == 1) {
      int half = F.n >> 1;
      F.cm = F.order ? ret : ret << (F.B0 >> 1);
      Frame& Cn = st[sp + 1];
      Cn.n = half; Cn.B = F.B; Cn.lm = F.lm - 1; Cn.st = 0;
      if (F.order) { Cn.x = F.x + half; Cn.lb = F.lb ? F.lb + half : nullptr; Cn.gmid = F.gside; Cn.fill = F.fill >> F.B; }
      else { Cn.x = F.x; Cn.lb = F.lb; Cn.gmid = F.gmid; Cn.fill = F.fill; }
      F.st = 2;
      sp++;
    } else {
      ret = F.cm | (F.order ? ret << (F.B0 >> 1) : ret);
      sp--;
    }
  }
  return ret;
}

// quant_band, resynthesis side (bands.rs:2133-2291)
MB_DEV_NOINLINE uint32_t recon_band(Walk& w, float* x, int n, int B, float* lowband, int lm, float* lowband_out, float gain,
                           float* lowband_scratch, uint32_t fill) {
  const int n0 = n, long_blocks = B == 1;
  if (n == 1) {
    uint32_t node = next_node(w);
    if (w.tm.tid == 0) {
      x[0] = (MB_ND_SIGN(node) & 1u) ? -1.0f : 1.0f;
      if (lowband_out) lowband_out[0] = x[0];
    }
    MB_SYNC();
    return 1;
  }
  int tf_change = w.tf_change;
  int n_b = n / B;
  int recombine = tf_change > 0 ? tf_change : 0, time_divide = 0;
  if (lowband && lowband_scratch && (recombine || ((n_b & 1) == 0 && tf_change < 0) || B > 1)) {
    for (int j = w.tm.tid; j < n; j += w.tm.n) lowband_scratch[j] = lowband[j];
    MB_SYNC();
    lowband = lowband_scratch;
  }
  for (int k = 0; k < recombine; k++) {
    if (lowband) team_haar1(w, lowband, n >> k, 1 << k);
    fill = kBitInterleave[fill & 0xF] | (uint32_t)kBitInterleave[(fill >> 4) & 0xF] << 2;
  }
  B >>= recombine;
  n_b <<= recombine;
  while ((n_b & 1) == 0 && tf_change < 0) {
    if (lowband) team_haar1(w, lowband, n_b, B);
    fill |= fill << B;
    B <<= 1;
    n_b >>= 1;
    time_divide++;
    tf_change++;
  }
  const int B0 = B, n_b0 = n_b;
  if (B0 > 1 && lowband) team_deinterleave(w, lowband, n_b >> recombine, B0 << recombine, long_blocks);
  uint32_t cm = recon_partition(w, x, n, B, lowband, lm, gain, fill);
  MB_SYNC();
  if (B0 > 1) team_interleave(w, x, n_b >> recombine, B0 << recombine, long_blocks);
  n_b = n_b0;
  B = B0;
  for (int k = 0; k < time_divide; k++) {
    B >>= 1;
    n_b <<= 1;
    if (B > 0) cm |= cm >> B;
    team_haar1(w, x, n_b, imax(B, 1));
  }
  for (int k = 0; k < recombine; k++) {
    cm = kBitDeinterleave[cm & 0xF];
    team_haar1(w, x, n0 >> k, 1 << k);
  }
  B <<= recombine;
  if (lowband_out) {
    float scale = celt_sqrt((float)n0);
    for (int j = w.tm.tid; j < n0; j += w.tm.n) lowband_out[j] = scale * x[j];
    MB_SYNC();
  }
  return cm & mask_from_bits(B);
}

// quant_band_stereo, resynthesis side (bands.rs:2294-2514, stereo_merge :3477-3505)
MB_DEV uint32_t recon_band_stereo(Walk& w, float* x, float* y, int n, int B, float* lowband, int lm, float* lowband_out,
                                  float* lowband_scratch, uint32_t fill) {
  if (n == 1) {
    uint32_t node = next_node(w);
    if (w.tm.tid == 0) {
      x[0] = (MB_ND_SIGN(node) & 1u) ? -1.0f : 1.0f;
      y[0] = (MB_ND_SIGN(node) & 2u) ? -1.0f : 1.0f;
      if (lowband_out) lowband_out[0] = x[0];
    }
    MB_SYNC();
    return 1;
  }
  const uint32_t orig_fill = fill;
  uint32_t node = next_node(w);
  const int itheta = (int)MB_ND_ITHETA(node), inv = (int)MB_ND_INV(node);
  int imid, iside;
  {
    uint32_t bmask = mask_from_bits(B);
    if (itheta == 0) { imid = 32767; iside = 0; fill &= bmask; }
    else if (itheta == 16384) { imid = 0; iside = 32767; fill &= B >= 32 ? 0xffffffffu : bmask << B; }
    else { imid = bitexact_cos(itheta); iside = bitexact_cos(16384 - itheta); }
  }
  const float mid = (float)imid / 32768.0f, side = (float)iside / 32768.0f;
  uint32_t cm;
  if (n == 2) {
    int c = itheta > 8192;
    float* x2 = c ? y : x;
    float* y2 = c ? x : y;
    int sv = 1 - 2 * (int)(MB_ND_SIGN(node) & 1u);
    cm = recon_band(w, x2, n, B, lowband, lm, lowband_out, 1.0f, lowband_scratch, orig_fill);
    if (w.tm.tid == 0) {
      y2[0] = -(float)sv * x2[1];
      y2[1] = (float)sv * x2[0];
      x[0] *= mid; x[1] *= mid; y[0] *= side; y[1] *= side;
      float t = x[0]; x[0] = t - y[0]; y[0] += t;
      t = x[1]; x[1] = t - y[1]; y[1] += t;
    }
    MB_SYNC();
  } else {
    if (MB_ND_ORDER(node)) {
      cm = recon_band(w, x, n, B, lowband, lm, lowband_out, 1.0f, lowband_scratch, fill);
      cm |= recon_band(w, y, n, B, nullptr, lm, nullptr, side, nullptr, fill >> B);
    } else {
      cm = recon_band(w, y, n, B, nullptr, lm, nullptr, side, nullptr, fill >> B);
      cm |= recon_band(w, x, n, B, lowband, lm, lowband_out, 1.0f, lowband_scratch, fill);
    }
    // stereo_merge: the two inner products keep the reference's summation order (one lane each)
    MB_SYNC();
    for (int q = w.tm.tid; q < 2; q += w.tm.n) {
      const float* o = q ? y : x;
      float acc = 0.f;
      for (int j = 0; j < n; j++) acc += y[j] * o[j];
      w.sm->red[q] = acc;
    }
    MB_SYNC();
    float lgain_, rgain_;
    int degenerate_;
    {
      float xp = w.sm->red[0] * mid, sd = w.sm->red[1];
      float mid2 = mid * mid;
      float el = mid2 + sd - 2.0f * xp, er = mid2 + sd + 2.0f * xp;
      degenerate_ = er < 6e-4f || el < 6e-4f;
      lgain_ = degenerate_ ? 0.f : celt_rsqrt(el);
      rgain_ = degenerate_ ? 0.f : celt_rsqrt(er);
    }
    MB_SYNC();
    const float lgain = lgain_, rgain = rgain_;
    const int degenerate = degenerate_;
    for (int j = w.tm.tid; j < n; j += w.tm.n) {
      if (degenerate) y[j] = x[j];
      else {
        float l = mid * x[j], r = y[j];
        x[j] = lgain * (l - r);
        y[j] = rgain * (l + r);
      }
    }
    MB_SYNC();
  }
  if (inv) {
    for (int j = w.tm.tid; j < n; j += w.tm.n) y[j] = -y[j];
    MB_SYNC();
  }
  return cm;
}

// ---------------------------------------------------------------- stage C: FFT + IMDCT
struct cpx { float r, i; };
MB_DEV cpx c_mul(cpx a, cpx b) { cpx o; o.r = fma32(a.r, b.r, -(a.i * b.i)); o.i = fma32(a.r, b.i, a.i * b.r); return o; }
MB_DEV cpx c_add(cpx a, cpx b) { cpx o; o.r = a.r + b.r; o.i = a.i + b.i; return o; }
MB_DEV cpx c_sub(cpx a, cpx b) { cpx o; o.r = a.r - b.r; o.i = a.i - b.i; return o; }
MB_DEV cpx fft_tw(int idx) { cpx o; o.r = bits2f(kFftTwiddles480Bits[2 * idx]); o.i = bits2f(kFftTwiddles480Bits[2 * idx + 1]); return o; }

// One butterfly of radix p at (group g, position k); p points at the group's base.
MB_DEV void bfly_one(cpx* p, int radix, int k, int m, int fstride) {
  if (radix == 4) {
    if (m == 1) {
      cpx s0 = c_sub(p[0], p[2]), s1 = c_add(p[1], p[3]), s1b = c_sub(p[1], p[3]);
      cpx f0 = c_add(p[0], p[2]);
      p[2] = c_sub(f0, s1);
      p[0] = c_add(f0, s1);
      cpx a = {s0.r + s1b.i, s0.i - s1b.r}, b = {s0.r - s1b.i, s0.i + s1b.r};
      p[1] = a; p[3] = b;
    } else {
      cpx s0 = c_mul(p[k + m], fft_tw(k * fstride)), s1 = c_mul(p[k + 2 * m], fft_tw(2 * k * fstride)),
          s2 = c_mul(p[k + 3 * m], fft_tw(3 * k * fstride));
      cpx s5 = c_sub(p[k], s1);
      cpx f0 = c_add(p[k], s1);
      cpx s3 = c_add(s0, s2), s4 = c_sub(s0, s2);
      p[k + 2 * m] = c_sub(f0, s3);
      p[k] = c_add(f0, s3);
      cpx a = {s5.r + s4.i, s5.i - s4.r}, b = {s5.r - s4.i, s5.i + s4.r};
      p[k + m] = a; p[k + 3 * m] = b;
    }
  } else if (radix == 2) {  // always the m == 4 form in the static modes (kiss_fft.rs:900-966)
    const float t8 = 0.70710678118654752440f;
    cpx v = p[4 + k], t;
    if (k == 0) t = v;
    else if (k == 1) { t.r = (v.r + v.i) * t8; t.i = (v.i - v.r) * t8; }
    else if (k == 2) { t.r = v.i; t.i = -v.r; }
    else { t.r = (v.i - v.r) * t8; t.i = -(v.i + v.r) * t8; }
    p[4 + k] = c_sub(p[k], t);
    p[k] = c_add(p[k], t);
  } else if (radix == 3) {
    cpx epi3 = fft_tw(fstride * m);
    cpx s1 = c_mul(p[m + k], fft_tw(k * fstride)), s2 = c_mul(p[2 * m + k], fft_tw(2 * k * fstride));
    cpx s3 = c_add(s1, s2), s0 = c_sub(s1, s2);
    cpx fm = {p[k].r - 0.5f * s3.r, p[k].i - 0.5f * s3.i};
    s0.r *= epi3.i; s0.i *= epi3.i;
    cpx f0 = c_add(p[k], s3);
    cpx a = {fm.r + s0.i, fm.i - s0.r}, b = {fm.r - s0.i, fm.i + s0.r};
    p[2 * m + k] = a; p[m + k] = b; p[k] = f0;
  } else {  // radix 5
    cpx ya = fft_tw(fstride * m), yb = fft_tw(fstride * 2 * m);
    cpx s0 = p[k];
    cpx s1 = c_mul(p[m + k], fft_tw(k * fstride)), s2 = c_mul(p[2 * m + k], fft_tw(2 * k * fstride));
    cpx s3 = c_mul(p[3 * m + k], fft_tw(3 * k * fstride)), s4 = c_mul(p[4 * m + k], fft_tw(4 * k * fstride));
    cpx s7 = c_add(s1, s4), s10 = c_sub(s1, s4), s8 = c_add(s2, s3), s9 = c_sub(s2, s3);
    cpx f0 = c_add(s0, c_add(s7, s8));
    cpx s5 = {s0.r + fma32(s7.r, ya.r, s8.r * yb.r), s0.i + fma32(s7.i, ya.r, s8.i * yb.r)};
    cpx s6 = {fma32(s10.i, ya.i, s9.i * yb.i), -fma32(s10.r, ya.i, s9.r * yb.i)};
    p[m + k] = c_sub(s5, s6);
    p[4 * m + k] = c_add(s5, s6);
    cpx s11 = {s0.r + fma32(s7.r, yb.r, s8.r * ya.r), s0.i + fma32(s7.i, yb.r, s8.i * ya.r)};
    cpx s12 = {fma32(s9.i, ya.i, (-s10.i) * yb.i), fma32(s10.r, yb.i, (-s9.r) * ya.i)};
    p[2 * m + k] = c_add(s11, s12);
    p[3 * m + k] = c_sub(s11, s12);
    p[k] = f0;
  }
}

// `nt` independent transforms of size nfft = 480 >> shift laid out back to back in `f`.
MB_DEV void team_fft(Team tm, cpx* f, int shift, int nt) {
  const uint8_t* fac = kFftFactors[shift];
  const int nfft = 480 >> shift;
  int fstride[6];
  fstride[0] = 1;
  int stages = 0, m;
  do {
    int p = fac[2 * stages];
    m = fac[2 * stages + 1];
    fstride[stages + 1] = fstride[stages] * p;
    stages++;
  } while (m != 1);
  m = fac[2 * stages - 1];
  for (int i = stages - 1; i >= 0; i--) {
    int mm = i != 0 ? fac[2 * i - 1] : 1;
    int radix = fac[2 * i], groups = fstride[i], fs = fstride[i] << shift;
    int per = groups * m;  // butterflies per transform in this stage
    for (int idx = tm.tid; idx < per * nt; idx += tm.n) {
      int t = idx / per, r = idx - t * per;
      int g = r / m, k = r - g * m;
      bfly_one(f + t * nfft + g * mm, radix, k, m, fs);
    }
    MB_SYNC();
    m = mm;
  }
}

// ---------------------------------------------------------------- whole frame
struct ReconParams {
  const uint32_t* rec;
  const float* en;
  const FrameMeta* meta;
  float* series;            // [lane][CC][series_stride]
  uint32_t series_stride;
  int CC;
  uint32_t quirks;
  int32_t* iy_dbg;          // optional [frame][2*960] decoded pulses (parity tests), else null
  float* x_dbg;             // optional [frame][2*960] normalised spectrum after anti-collapse
};

MB_DEV void celt_recon_frame(Team tm, ReconSmem* sm, uint32_t f, const ReconParams& P) {
  const FrameMeta meta = P.meta[f];
  if (!(meta.flags & 1u)) return;
  const uint32_t* r = P.rec + rec_index(f, 0);
  const uint32_t h0 = r[S_HDR0 * 32], h1 = r[S_HDR1 * 32], status = r[S_STATUS * 32];
  const int lm = (int)MB_H0_LM(h0), C = (int)MB_H0_C(h0), start = (int)MB_H0_START(h0), end = (int)MB_ST_END(status);
  const int transient = (int)MB_H0_TRANSIENT(h0), silence = (int)MB_H0_SILENCE(h0), spread = (int)MB_H0_SPREAD(h0);
  const int M = 1 << lm, N = M * SHORT_MDCT, CC = P.CC;
  const int n_nodes = (int)MB_H1_NNODES(h1), n_leaves = (int)MB_H1_NLEAVES(h1);
  float* X = sm->X;
  float* Y = C == 2 ? X + N : nullptr;

  for (int i = tm.tid; i < C * N; i += tm.n) X[i] = 0.f;
  MB_SYNC();
  // ---- stage A: all PVQ leaves in parallel
  for (int l = tm.tid; l < n_leaves; l += tm.n) {
    uint32_t index = r[(S_LEAVES + 3 * l) * 32];
    float gain = bits2f(r[(S_LEAVES + 3 * l + 1) * 32]);
    uint32_t g = r[(S_LEAVES + 3 * l + 2) * 32];
    int xo = (int)MB_LF_XOFF(g);     // offset in the packed spectrum (channel pitch N)
    int ch = xo >= N ? 1 : 0;
    uint32_t cm = pvq_leaf(X + xo, (int)MB_LF_N(g), (int)MB_LF_K(g), (int)MB_LF_B(g), gain, spread, index,
                           P.iy_dbg ? P.iy_dbg + (size_t)f * 2 * MAX_N + ch * MAX_N + (xo - ch * N) : nullptr);
    sm->leaf_cm[l] = (uint16_t)cm;
  }
  MB_SYNC();
  // ---- stage B: band walk
  uint32_t seed;
  uint8_t* collapse = sm->collapse;   // every lane stores the same value: no ordering needed between lanes
  {
    Walk w;
    w.tm = tm; w.sm = sm; w.seed = meta.seed_in; w.leaf_pos = 0; w.spread = spread; w.tf_change = 0;
    open_nodes(w, r + S_NODES * 32, n_nodes);
    const int B = transient ? M : 1;
    const int norm_offset = M * eband(start);
    const int norm_len = imax(M * eband(NB_EBANDS - 1) - norm_offset, 0);
    float* norm = sm->W;
    float* norm2 = norm + norm_len;
    // (CELT-only frames start at band 0: every fold source is written earlier in the same frame, so
    //  the fold buffer needs no cross-frame persistence; hybrid frames will carry it in stream state)
    for (int i = tm.tid; i < 2 * norm_len; i += tm.n) norm[i] = 0.f;
    for (int i = tm.tid; i < 2 * NB_EBANDS; i += tm.n) collapse[i] = 0;
    MB_SYNC();
    for (int i = start; i < end; i++) {
      const uint32_t ba = r[(S_BANDA + i) * 32];
      const int band_start = M * eband(i), n = M * eband(i + 1) - band_start;
      const int last = i + 1 == end;
      w.tf_change = MB_BA_TF(ba);
      if (i == start + 1) {  // special_hybrid_folding (bands.rs:3517-3555)
        int n1 = M * (eband(start + 1) - eband(start)), n2 = M * (eband(start + 2) - eband(start + 1));
        if (n2 > n1) {
          int cnt = n2 - n1;
          for (int j = tm.tid; j < cnt; j += tm.n) {
            norm[n1 + j] = norm[2 * n1 - n2 + j];
            if (MB_H0_DUAL(h0)) norm2[n1 + j] = norm2[2 * n1 - n2 + j];
          }
          MB_SYNC();
        }
      }
      uint32_t x_cm = 0, y_cm = 0;
      if (MB_BA_HASFOLD(ba)) {
        for (int fb = (int)MB_BA_FS(ba); fb < (int)MB_BA_FE(ba); fb++) { x_cm |= collapse[fb * C]; y_cm |= collapse[fb * C + C - 1]; }
      } else x_cm = y_cm = mask_from_bits(B);
      if (MB_BA_AVG(ba)) {
        for (int j = tm.tid; j < norm_len; j += tm.n) norm[j] = (norm[j] + norm2[j]) * 0.5f;
        MB_SYNC();
      }
      const int in1 = (int)MB_BA_IN1(ba);
      float* lb_in = in1 ? norm + (in1 - 1) : nullptr;
      float* lb_out = MB_BA_OUT(ba) ? norm + (band_start - norm_offset) : nullptr;
      float* scratch = last ? nullptr : sm->lowband_scratch;
      float* xb = X + band_start;
      float* yb = Y ? Y + band_start : nullptr;
      if (MB_BA_DUAL(ba)) {
        x_cm = recon_band(w, xb, n, B, lb_in, lm, lb_out, 1.0f, scratch, x_cm);
        y_cm = recon_band(w, yb, n, B, in1 ? norm2 + (in1 - 1) : nullptr, lm, lb_out ? norm2 + (band_start - norm_offset) : nullptr, 1.0f,
                          scratch, y_cm);
      } else if (Y) {
        x_cm = recon_band_stereo(w, xb, yb, n, B, lb_in, lm, lb_out, scratch, x_cm | y_cm);
        y_cm = x_cm;
      } else {
        x_cm = recon_band(w, xb, n, B, lb_in, lm, lb_out, 1.0f, scratch, x_cm | y_cm);
        y_cm = x_cm;
      }
      collapse[i * C] = (uint8_t)x_cm;
      collapse[i * C + C - 1] = (uint8_t)y_cm;
    }
    seed = w.seed;
  }
  MB_SYNC();
  const float* E = P.en + (size_t)f * EN_STRIDE;
  // ---- anti_collapse (bands.rs:3220-3349)
  if (MB_H0_ACON(h0)) {
    const float* prev1 = E + 2 * NB_EBANDS;
    const float* prev2 = E + 4 * NB_EBANDS;
    for (int i = start; i < end; i++) {
      int N0 = eband(i + 1) - eband(i);
      int pulses_i = (int)r[(S_BANDB + i) * 32];
      int depth = (int)((uint32_t)(1 + pulses_i) / (uint32_t)N0) >> lm;
      float thresh = 0.5f * celt_exp2(-0.125f * (float)depth);
      float sqrt_1 = celt_rsqrt((float)(N0 << lm));
      for (int c = 0; c < C; c++) {
        float p1 = prev1[c * NB_EBANDS + i], p2 = prev2[c * NB_EBANDS + i];
        if (C == 1) { p1 = fmax32(p1, prev1[NB_EBANDS + i]); p2 = fmax32(p2, prev2[NB_EBANDS + i]); }
        float Ediff = E[c * NB_EBANDS + i] - fmin32(p1, p2);
        if (Ediff < 0.f) Ediff = 0.f;
        float rr = 2.0f * celt_exp2(-Ediff);
        if (lm == 3) rr *= 1.41421356f;
        rr = fmin32(rr, thresh);
        rr *= sqrt_1;
        float* Xb = X + c * N + (eband(i) << lm);
        int renorm = 0;
        for (int k = 0; k < M; k++) {
          if (!(collapse[i * C + c] & (1u << k))) {
            for (int j = tm.tid; j < N0; j += tm.n) {
              uint32_t s = lcg_jump(seed, j + 1);
              Xb[(j << lm) + k] = (s & 0x8000u) ? rr : -rr;
            }
            seed = lcg_jump(seed, N0);
            renorm = 1;
          }
        }
        if (renorm) {
          Walk w; w.tm = tm; w.sm = sm;
          team_renormalise(w, Xb, N0 << lm, 1.0f);
        }
      }
    }
    MB_SYNC();
  }
  if (P.x_dbg) {
    for (int i = tm.tid; i < C * N; i += tm.n) P.x_dbg[(size_t)f * 2 * MAX_N + (i / N) * MAX_N + (i % N)] = X[i];
  }
  // ---- denormalise_bands (bands.rs:4172-4234), in place: X becomes freq
  {
    const int eff_end = imin(end, NB_EBANDS);
    float* gains = sm->tmp;  // per (channel, band) linear gain, computed once
    for (int idx = tm.tid; idx < C * NB_EBANDS; idx += tm.n) {
      int b = idx % NB_EBANDS;
      gains[idx] = celt_exp2(fmin32(E[idx] + kEMeans[b], 32.0f));
    }
    MB_SYNC();
    const int lo = silence ? 0 : M * eband(start), hi = silence ? 0 : M * eband(eff_end);
    for (int idx = tm.tid; idx < C * N; idx += tm.n) {
      int c = idx >= N, j = idx - c * N;
      float v = 0.f;
      if (j >= lo && j < hi) v = X[idx] * gains[c * NB_EBANDS + kBandOfBin[j >> lm]];
      X[idx] = v;
    }
    MB_SYNC();
    if (CC == 1 && C == 2) {
      for (int j = tm.tid; j < N; j += tm.n) X[j] = 0.5f * (X[j] + X[N + j]);
      MB_SYNC();
    }
  }
  // ---- inverse MDCT (mdct.rs:197-388 with the reference's zero-filled overlap)
  {
    const int nch = (CC == 2 && C == 2) ? 2 : 1;       // distinct spectra to transform
    const int nbk = transient ? M : 1;                  // short blocks per channel
    const int shift = transient ? MAX_LM : MAX_LM - lm;
    const int n2 = 960 >> shift, n4 = n2 >> 1;          // per-block MDCT size / FFT size
    const int nt = nch * nbk;                           // transforms, each n4 complex points
    cpx* F = (cpx*)sm->W;
    const uint32_t* twb = kMdctTwiddles960Bits + kMdctTwOff[shift];
    const int16_t* br = shift == 0 ? kFftBitrev480 : shift == 1 ? kFftBitrev240 : shift == 2 ? kFftBitrev120 : kFftBitrev60;
    // pre-rotation + bit-reversal + conjugate (ifft = conj . fft . conj)
    for (int idx = tm.tid; idx < nt * n4; idx += tm.n) {
      int t = idx / n4, i = idx - t * n4;
      int c = t / nbk, b = t - c * nbk;
      const float* in = X + c * N + b;          // element k of block b is in[k * nbk]
      float x1 = in[(2 * i) * nbk], x2 = in[(n2 - 1 - 2 * i) * nbk];
      float t0 = bits2f(twb[i]), t1 = bits2f(twb[n4 + i]);
      cpx v;
      v.r = x2 * t0 + x1 * t1;
      v.i = -(x1 * t0 - x2 * t1);
      F[t * n4 + br[i]] = v;
    }
    MB_SYNC();
    team_fft(tm, F, shift, nt);
    // post-rotation and window fold against the zero-filled overlap into shared memory (X is free once the
    // pre-rotation has consumed it), then one coalesced pass to the lane's time series
    float* O = X;
    for (int idx = tm.tid; idx < nt * n4; idx += tm.n) {
      int t = idx / n4, i = idx - t * n4;
      int c = t / nbk, b = t - c * nbk;
      cpx v = F[t * n4 + i];
      v.i = -v.i;
      float t0 = bits2f(twb[i]), t1 = bits2f(twb[n4 + i]);
      float yr = v.r * t0 + v.i * t1;     // temp[2i]
      float yi = v.r * t1 - v.i * t0;     // temp[n2-1-2i]
      float* ob = O + c * N + b * n2;
      const int lim = nbk > 1 ? 120 : N;
#pragma unroll
      for (int q = 0; q < 2; q++) {
        int k = q ? n2 - 1 - 2 * i : 2 * i;
        float val = q ? yi : yr;
        // temp[k] lands at out[60 + k]; the fold rewrites out[0..120) from temp[0..60)
        if (k < 60) {
          float w1 = bits2f(kWindow120Bits[59 - k]), w2 = bits2f(kWindow120Bits[60 + k]);
          ob[59 - k] = 0.f * w2 - val * w1;     // out[i] = x2*w2 - x1*w1 with x2 = 0
          ob[60 + k] = 0.f * w1 + val * w2;     // out[119-i] = x2*w1 + x1*w2
        } else if (60 + k < lim) {
          ob[60 + k] = val;
        }
      }
    }
    MB_SYNC();
    const float SIG_SAT = 536870911.0f;
    for (int cc = 0; cc < CC; cc++) {
      const float* src = O + (nch == 2 ? cc : 0) * N;
      float* out = P.series + ((size_t)meta.lane * CC + cc) * P.series_stride + HIST + meta.sample_off;
      for (int i = tm.tid; i < N; i += tm.n) {
        float o = src[i];
        out[i] = o < -SIG_SAT ? -SIG_SAT : (o > SIG_SAT ? SIG_SAT : o);
      }
    }
  }
}

}  // namespace mb
