// Phase 2b(i) — PVQ leaves: one thread per leaf, every leaf of every frame concurrently.
//
// Replaces:
//   decode_pulses / cwrsi_pvq                 src/celt/cwrs.rs:293-452
//   alg_unquant, normalise_residual           src/celt/vq.rs:1013-1037, 332-355
//   exp_rotation / exp_rotation1 (inverse)    src/celt/vq.rs:116-221
//   extract_collapse_mask                     src/celt/vq.rs:1164-1192
// A leaf's index -> pulse vector walk and its spreading rotation are serial inside the leaf, but leaves do not
// depend on each other, on the band order, or on decoder state: phase 1 recorded (index, N, K, B, gain, offset) for
// each. A CTA stages one frame's spectrum in shared memory, its threads decode the leaves in place, and the frame
// leaves the SM as one coalesced row of the spectrum buffer; the collapse mask of each leaf replaces the leaf's gain
// word in the record (phase 2b(ii) needs only the mask).
#pragma once
#include "mb_celt_energy.cuh"
#include "mb_celt_parse.cuh"

namespace mb {

enum { XBUF_STRIDE = 2 * MAX_N };   // floats per frame in the spectrum buffer (channel pitch N, packed)


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


struct PvqParams {
  uint32_t* rec;
  const FrameMeta* meta;    // null: decide from the record alone (see celt_pvq_frame)
  float* xbuf;              // [frame][XBUF_STRIDE]
  int32_t* iy_dbg;          // optional [frame][2*960] decoded pulses (parity tests), else null
};

// One frame: `tm` threads share its leaves; X is the CTA's staging row (C*N floats).
MB_DEV void celt_pvq_frame(Team tm, float* X, uint32_t f, const PvqParams& P) {
  uint32_t* r = P.rec + rec_index(f, 0);
  if (P.meta) { if (!(P.meta[f].flags & 1u)) return; }
  else {
    // no frame meta yet (the kernel runs beside k_energy): every cell whose record holds a parsed CELT part is
    // unpacked; a cell the stream pass rejects afterwards (mode switch, failed packet) is simply never consumed
    const uint32_t status = r[S_STATUS * 32];
    if (!MB_ST_DECODED(status) || (((status >> 16) & 1u) && !((status >> 17) & 1u) && !((status >> 19) & 3u))) return;   // SILK only, no CELT record
  }
  const uint32_t h0 = r[S_HDR0 * 32], h1 = r[S_HDR1 * 32];
  const int lm = (int)MB_H0_LM(h0), C = (int)MB_H0_C(h0), spread = (int)MB_H0_SPREAD(h0);
  const int N = SHORT_MDCT << lm;
  const int n_leaves = (int)MB_H1_NLEAVES(h1);
  for (int l = tm.tid; l < n_leaves; l += tm.n) {
    uint32_t index = r[(S_LEAVES + 3 * l) * 32];
    float gain = bits2f(r[(S_LEAVES + 3 * l + 1) * 32]);
    uint32_t g = r[(S_LEAVES + 3 * l + 2) * 32];
    int xo = (int)MB_LF_XOFF(g);     // offset in the packed spectrum (channel pitch N)
    int ch = xo >= N ? 1 : 0;
    uint32_t cm = pvq_leaf(X + xo, (int)MB_LF_N(g), (int)MB_LF_K(g), (int)MB_LF_B(g), gain, spread, index,
                           P.iy_dbg ? P.iy_dbg + (size_t)f * 2 * MAX_N + ch * MAX_N + (xo - ch * N) : nullptr);
    r[(S_LEAVES + 3 * l + 1) * 32] = cm;
  }
  MB_CTA_SYNC();
  float* out = P.xbuf + (size_t)f * XBUF_STRIDE;
  for (int i = tm.tid; i < C * N; i += tm.n) out[i] = X[i];
}

}  // namespace mb
