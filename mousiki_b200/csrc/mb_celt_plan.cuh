// Phase 2b(ii) — band plan: one lane per frame resolves every integer decision of the band resynthesis and
// writes a flat list of vector operations for phase 2b(iii) to execute.
//
// Replaces the control half (no float vector is touched here) of
//   quant_all_bands (decode/resynthesis)      src/celt/bands.rs:2575-3082
//   quant_band / quant_band_stereo            src/celt/bands.rs:2133-2291, 2294-2514
//   quant_partition                           src/celt/bands.rs:1917-2131
//   special_hybrid_folding                    src/celt/bands.rs:3517-3555
//   anti_collapse (which blocks, which seeds) src/celt/bands.rs:3220-3349
// Why a separate pass: fill masks, collapse masks, fold sources and the noise generator's seed at each use are
// integer functions of the split tree (phase 1) and of the leaves' collapse masks (phase 2b(i)). Walking the tree is
// branchy scalar work — ideal for one lane per frame, 32 frames per warp — while the vector work it steers is
// regular. Resolving it here also removes dead work: the reference transforms the fold source (`lowband`) of every
// band; the plan keeps those operations only for bands that really fold.
#pragma once
#include "mb_celt_pvq.cuh"

namespace mb {

// ---- operation stream (32-bit words). word0: type[0..3] a[4..12] n[13..21] p[22..31]; extra words per type.
// `a` is an offset into the band window (x at 0, y at WIN_Y), `n` a length in samples.
enum {
  OP_END_BAND = 0,
  OP_SKIP = 1,      // n = words to skip after this one
  OP_COPY_LB = 2,   // scratch[0..n) = norm[w1 & 2047 ..] (w1 bit 11: second fold buffer); p & 1: the other way  +1 word
  OP_HAAR = 3,      // haar1 over n samples, stride 1 << (p >> 1), buffer p & 1 (0 window at a, 1 scratch)
  OP_DEINT = 4,     // deinterleave_hadamard: stride 1 << ((p >> 1) & 7), hadamard p >> 4, buffer p & 1
  OP_INTER = 5,     // interleave_hadamard, same fields
  OP_ZERO = 6,      // x[a..a+n) = 0
  OP_FOLD = 7,      // x = scratch[p..] + noise sign, renormalise to gain w1; seed w2 (p = offset in scratch)  +2 words
  OP_NOISE = 8,     // x = LCG noise, renormalise to gain w1; seed w2                                             +2 words
  OP_N1 = 9,        // x[a] = p & 1 ? -1 : +1
  OP_LBOUT = 10,    // norm[w1 & 2047 ..] = sqrt(n) * x[a..a+n) (w1 bit 11: second fold buffer)                   +1 word
  OP_MERGE = 11,    // stereo_merge(x, y, mid = (w1 & 0xffff) / 32768, n); p & 1: negate y afterwards            +1 word
  OP_STEREO2 = 12,  // the N = 2 stereo case; p: c[0] sign[1] inv[2]; w1 = imid | iside << 16                      +1 word
  OP_AVG = 13,      // norm = (norm + norm2) / 2 over n | p << 9 samples
  OP_HYB = 14,      // special_hybrid_folding: w1 = n1 | cnt << 11 | dual << 22                                    +1 word
  OP_ANTI = 15,     // anti_collapse of one (band, channel): p = lm | block mask << 2; w1 seed, w2 r bits          +2 words
  WIN_Y = 176,
};
#define MB_OP(type, a, n, p) ((uint32_t)(type) | ((uint32_t)(a) << 4) | ((uint32_t)(n) << 13) | ((uint32_t)(p) << 22))
#define MB_OP_TYPE(w) ((w) & 15u)
#define MB_OP_A(w) (((w) >> 4) & 511u)
#define MB_OP_N(w) (((w) >> 13) & 511u)
#define MB_OP_P(w) (((w) >> 22) & 1023u)

MB_TBL uint8_t kBitInterleave[16] = {0, 1, 1, 1, 2, 3, 3, 3, 2, 3, 3, 3, 2, 3, 3, 3};
MB_TBL uint8_t kBitDeinterleave[16] = {0x00, 0x03, 0x0C, 0x0F, 0x30, 0x33, 0x3C, 0x3F, 0xC0, 0xC3, 0xCC, 0xCF, 0xF0, 0xF3, 0xFC, 0xFF};

MB_DEV uint32_t lcg_jump(uint32_t seed, int steps) { return kLcgA[steps] * seed + kLcgC[steps]; }
MB_DEV int ilog2_exact(int v) { return ilog32((uint32_t)v) - 1; }

struct Plan {
  const uint32_t* nodes;     // word i at nodes[i * 32]
  const uint32_t* leaves;    // leaf l: collapse mask at leaves[(3 * l + 1) * 32]
  uint32_t* ops;             // word i at ops[i * 32]
  int node_pos, leaf_pos, n_ops, n_fold, overflow;
  uint32_t seed;
  int tf_change;
  int lb_inplace;            // last band of a hybrid frame: the fold source is transformed in place (see plan_band)

  MB_MEM void put(uint32_t w) {
    if (n_ops < CAP_OPS) ops[(size_t)n_ops * 32] = w; else overflow = 1;
    n_ops++;
  }
  MB_MEM void set(int pos, uint32_t w) { if (pos < CAP_OPS) ops[(size_t)pos * 32] = w; }
  MB_MEM uint32_t node() { return nodes[(size_t)(node_pos++) * 32]; }
  MB_MEM uint32_t leaf_cm() { return leaves[(size_t)(3 * (leaf_pos++) + 1) * 32]; }
};

// quant_partition: a = window offset of the partition, lb = offset of its fold source in the scratch (-1: none)
MB_DEV uint32_t plan_partition(Plan& P, int a, int n, int B, int lb, float gain, uint32_t fill) {
  struct Frame { int a, lb, n, B, B0; float gmid, gside; uint32_t fill, cm; uint8_t order, st; };
  Frame st[6];
  int sp = 0;
  st[0].a = a; st[0].lb = lb; st[0].n = n; st[0].B = B; st[0].gmid = gain; st[0].fill = fill; st[0].st = 0;
  uint32_t ret = 0;
  while (sp >= 0) {
    Frame& F = st[sp];
    if (F.st == 0) {
      const uint32_t node = P.node();
      const uint32_t type = MB_ND_TYPE(node);
      if (type == ND_SPLIT) {
        const int half = F.n >> 1, itheta = (int)MB_ND_ITHETA(node);
        F.B0 = F.B;
        if (F.B == 1) F.fill = (F.fill & 1u) | (F.fill << 1);
        F.B = (F.B + 1) >> 1;
        const uint32_t bmask = mask_from_bits(F.B);
        int imid, iside;
        if (itheta == 0) { imid = 32767; iside = 0; F.fill &= bmask; }
        else if (itheta == 16384) { imid = 0; iside = 32767; F.fill &= F.B >= 32 ? 0xffffffffu : bmask << F.B; }
        else { imid = bitexact_cos(itheta); iside = bitexact_cos(16384 - itheta); }
        const float g = F.gmid;
        F.gmid = g * ((float)imid / 32768.0f);
        F.gside = g * ((float)iside / 32768.0f);
        F.order = (uint8_t)MB_ND_ORDER(node);
        F.st = 1;
        Frame& Cn = st[sp + 1];
        Cn.n = half; Cn.B = F.B; Cn.st = 0;
        if (F.order) { Cn.a = F.a; Cn.lb = F.lb; Cn.gmid = F.gmid; Cn.fill = F.fill; }
        else { Cn.a = F.a + half; Cn.lb = F.lb >= 0 ? F.lb + half : -1; Cn.gmid = F.gside; Cn.fill = F.fill >> F.B; }
        sp++;
      } else {
        uint32_t cm = 0;
        if (type == ND_LEAF_PULSE) {
          cm = P.leaf_cm();
        } else {  // no pulses: fold / noise fill (bands.rs:2104-2127)
          const uint32_t cm_mask = mask_from_bits(F.B);
          const uint32_t fl = F.fill & cm_mask;
          if (!fl) {
            P.put(MB_OP(OP_ZERO, F.a, F.n, 0));
          } else {
            if (F.lb >= 0) { P.put(MB_OP(OP_FOLD, F.a, F.n, F.lb)); P.n_fold++; cm = fl; }
            else { P.put(MB_OP(OP_NOISE, F.a, F.n, 0)); cm = cm_mask; }
            P.put(f2bits(F.gmid));
            P.put(P.seed);
            P.seed = lcg_jump(P.seed, F.n);
          }
        }
        ret = cm;
        sp--;
      }
    } else if (F.st == 1) {
      const int half = F.n >> 1;
      F.cm = F.order ? ret : ret << (F.B0 >> 1);
      Frame& Cn = st[sp + 1];
      Cn.n = half; Cn.B = F.B; Cn.st = 0;
      if (F.order) { Cn.a = F.a + half; Cn.lb = F.lb >= 0 ? F.lb + half : -1; Cn.gmid = F.gside; Cn.fill = F.fill >> F.B; }
      else { Cn.a = F.a; Cn.lb = F.lb; Cn.gmid = F.gmid; Cn.fill = F.fill; }
      F.st = 2;
      sp++;
    } else {
      ret = F.cm | (F.order ? ret << (F.B0 >> 1) : ret);
      sp--;
    }
  }
  return ret;
}

// quant_band. lb_src: offset of the fold source in the fold buffer (-1 none), lb_which: 0 norm / 1 norm2;
// lb_out: destination offset of this band's contribution to the fold buffer (-1 none).
MB_DEV_NOINLINE uint32_t plan_band(Plan& P, int a, int n, int B, int lb_src, int lb_which, int lb_out, int out_which, float gain, uint32_t fill) {
  const int n0 = n, long_blocks = B == 1;
  if (n == 1) {
    const uint32_t node = P.node();
    P.put(MB_OP(OP_N1, a, 1, MB_ND_SIGN(node) & 1u));
    if (lb_out >= 0) { P.put(MB_OP(OP_LBOUT, a, 1, 0)); P.put((uint32_t)lb_out | ((uint32_t)out_which << 11)); }
    return 1;
  }
  int tf_change = P.tf_change;
  int n_b = n / B;
  const int recombine = tf_change > 0 ? tf_change : 0;
  int time_divide = 0;
  const int have_lb = lb_src >= 0;
  const int lb_begin = P.n_ops;
  if (have_lb) { P.put(MB_OP(OP_COPY_LB, 0, n, 0)); P.put((uint32_t)lb_src | ((uint32_t)lb_which << 11)); }
  for (int k = 0; k < recombine; k++) {
    if (have_lb) P.put(MB_OP(OP_HAAR, 0, n, 1 | (k << 1)));
    fill = kBitInterleave[fill & 0xF] | (uint32_t)kBitInterleave[(fill >> 4) & 0xF] << 2;
  }
  B >>= recombine;
  n_b <<= recombine;
  while ((n_b & 1) == 0 && tf_change < 0) {
    if (have_lb) P.put(MB_OP(OP_HAAR, 0, n, 1 | (ilog2_exact(B) << 1)));
    fill |= fill << B;
    B <<= 1;
    n_b >>= 1;
    time_divide++;
    tf_change++;
  }
  const int B0 = B, n_b0 = n_b;
  const int il_stride_log = ilog2_exact(B0 << recombine);
  if (B0 > 1 && have_lb) P.put(MB_OP(OP_DEINT, 0, n, 1 | (il_stride_log << 1) | (long_blocks << 4)));
  // The reference hands the last band no scratch, so the transforms above run on the fold buffer itself
  // (bands.rs:2176-2200); its fold buffer outlives the frame (types.rs:400), and the next hybrid frame of the stream reads it.
  const int write_back = have_lb && P.lb_inplace && P.n_ops > lb_begin + 2;
  if (write_back) { P.put(MB_OP(OP_COPY_LB, 0, n, 1)); P.put((uint32_t)lb_src | ((uint32_t)lb_which << 11)); }
  const int lb_end = P.n_ops;
  const int folds_before = P.n_fold;
  uint32_t cm = plan_partition(P, a, n, B, have_lb ? 0 : -1, gain, fill);
  if (have_lb && !write_back && P.n_fold == folds_before) P.set(lb_begin, MB_OP(OP_SKIP, 0, lb_end - lb_begin - 1, 0));  // nobody folded: dead work
  if (B0 > 1) P.put(MB_OP(OP_INTER, a, n, (il_stride_log << 1) | (long_blocks << 4)));
  n_b = n_b0;
  B = B0;
  for (int k = 0; k < time_divide; k++) {
    B >>= 1;
    n_b <<= 1;
    if (B > 0) cm |= cm >> B;
    P.put(MB_OP(OP_HAAR, a, n, ilog2_exact(imax(B, 1)) << 1));
  }
  for (int k = 0; k < recombine; k++) {
    cm = kBitDeinterleave[cm & 0xF];
    P.put(MB_OP(OP_HAAR, a, n, k << 1));
  }
  B <<= recombine;
  if (lb_out >= 0) { P.put(MB_OP(OP_LBOUT, a, n0, 0)); P.put((uint32_t)lb_out | ((uint32_t)out_which << 11)); }
  return cm & mask_from_bits(B);
}

// quant_band_stereo (x at window offset 0, y at WIN_Y)
MB_DEV_NOINLINE uint32_t plan_band_stereo(Plan& P, int n, int B, int lb_src, int lb_out, uint32_t fill) {
  if (n == 1) {
    const uint32_t node = P.node();
    P.put(MB_OP(OP_N1, 0, 1, MB_ND_SIGN(node) & 1u));
    P.put(MB_OP(OP_N1, WIN_Y, 1, (MB_ND_SIGN(node) >> 1) & 1u));
    if (lb_out >= 0) { P.put(MB_OP(OP_LBOUT, 0, 1, 0)); P.put((uint32_t)lb_out); }
    return 1;
  }
  const uint32_t orig_fill = fill;
  const uint32_t node = P.node();
  const int itheta = (int)MB_ND_ITHETA(node), inv = (int)MB_ND_INV(node);
  int imid, iside;
  {
    const uint32_t bmask = mask_from_bits(B);
    if (itheta == 0) { imid = 32767; iside = 0; fill &= bmask; }
    else if (itheta == 16384) { imid = 0; iside = 32767; fill &= B >= 32 ? 0xffffffffu : bmask << B; }
    else { imid = bitexact_cos(itheta); iside = bitexact_cos(16384 - itheta); }
  }
  const float side = (float)iside / 32768.0f;
  uint32_t cm;
  if (n == 2) {
    const int c = itheta > 8192;
    cm = plan_band(P, c ? WIN_Y : 0, n, B, lb_src, 0, lb_out, 0, 1.0f, orig_fill);
    P.put(MB_OP(OP_STEREO2, 0, 2, (uint32_t)c | ((MB_ND_SIGN(node) & 1u) << 1) | ((uint32_t)inv << 2)));
    P.put((uint32_t)imid | ((uint32_t)iside << 16));
  } else {
    if (MB_ND_ORDER(node)) {
      cm = plan_band(P, 0, n, B, lb_src, 0, lb_out, 0, 1.0f, fill);
      cm |= plan_band(P, WIN_Y, n, B, -1, 0, -1, 0, side, fill >> B);
    } else {
      cm = plan_band(P, WIN_Y, n, B, -1, 0, -1, 0, side, fill >> B);
      cm |= plan_band(P, 0, n, B, lb_src, 0, lb_out, 0, 1.0f, fill);
    }
    P.put(MB_OP(OP_MERGE, 0, n, inv));
    P.put((uint32_t)imid);
  }
  return cm;
}

struct PlanParams {
  uint32_t* rec;
  const float* en;
  const FrameMeta* meta;
};

// The whole frame (band loop of quant_all_bands, then the anti-collapse plan).
MB_DEV void celt_plan_frame(uint32_t f, const PlanParams& Q) {
  const FrameMeta meta = Q.meta[f];
  if (!(meta.flags & 1u)) return;
  uint32_t* r = Q.rec + rec_index(f, 0);
  const uint32_t h0 = r[S_HDR0 * 32], status = r[S_STATUS * 32];
  const int lm = (int)MB_H0_LM(h0), C = (int)MB_H0_C(h0), start = (int)MB_H0_START(h0), end = (int)MB_ST_END(status);
  const int transient = (int)MB_H0_TRANSIENT(h0), acon = (int)MB_H0_ACON(h0);
  const int M = 1 << lm;
  Plan P;
  P.nodes = r + S_NODES * 32; P.leaves = r + S_LEAVES * 32; P.ops = r + S_OPS * 32;
  P.node_pos = 0; P.leaf_pos = 0; P.n_ops = 0; P.n_fold = 0; P.overflow = 0; P.seed = meta.seed_in; P.tf_change = 0; P.lb_inplace = 0;
  if (meta.flags & MB_FL_PLC) {
    // Concealed frame. Pitch repetition is synthesised in the time domain (k_comb_plc): nothing to plan. The noise
    // concealment (celt_decoder.rs:1507-1533) fills every band of every channel with unit-norm noise; the generator runs
    // channel by channel, so the second channel's seeds continue where the first channel's bands end.
    if (meta.flags & MB_FL_PLC_PITCH) return;
    const int eff_end = imin(end, NB_EBANDS);
    uint32_t seed_c[2];
    seed_c[0] = meta.seed_in;
    uint32_t sd = meta.seed_in;
    for (int i = start; i < eff_end; i++) sd = lcg_jump(sd, M * (eband(i + 1) - eband(i)));
    seed_c[1] = sd;
    for (int i = start; i < eff_end; i++) {
      const int n = M * (eband(i + 1) - eband(i));
      for (int c = 0; c < C; c++) {
        P.put(MB_OP(OP_NOISE, c * WIN_Y, n, 0));
        P.put(f2bits(1.0f));
        P.put(seed_c[c]);
        seed_c[c] = lcg_jump(seed_c[c], n);
      }
      P.put(MB_OP(OP_END_BAND, 0, 0, 0));
    }
    r[S_PLAN * 32] = (uint32_t)P.n_ops;
    return;
  }
  uint8_t collapse[2 * NB_EBANDS];
  uint16_t anti_pos[NB_EBANDS];
  for (int i = 0; i < 2 * NB_EBANDS; i++) collapse[i] = 0;
  const int B = transient ? M : 1;
  const int norm_offset = M * eband(start);
  const int norm_len = imax(M * eband(NB_EBANDS - 1) - norm_offset, 0);
  for (int i = start; i < end; i++) {
    const uint32_t ba = r[(S_BANDA + i) * 32];
    const int band_start = M * eband(i), n = M * eband(i + 1) - band_start;
    P.tf_change = MB_BA_TF(ba);
    P.lb_inplace = i + 1 == end;   // (CELT-only frames too: their fold buffer can be read by a later hybrid frame of the stream)
    if (i == start + 1) {  // special_hybrid_folding (bands.rs:3517-3555)
      const int n1 = M * (eband(start + 1) - eband(start)), n2 = M * (eband(start + 2) - eband(start + 1));
      if (n2 > n1) { P.put(MB_OP(OP_HYB, 0, 0, 0)); P.put((uint32_t)n1 | ((uint32_t)(n2 - n1) << 11) | (MB_H0_DUAL(h0) << 22)); }
    }
    uint32_t x_cm = 0, y_cm = 0;
    if (MB_BA_HASFOLD(ba)) {
      for (int fb = (int)MB_BA_FS(ba); fb < (int)MB_BA_FE(ba); fb++) { x_cm |= collapse[fb * C]; y_cm |= collapse[fb * C + C - 1]; }
    } else x_cm = y_cm = mask_from_bits(B);
    if (MB_BA_AVG(ba)) P.put(MB_OP(OP_AVG, 0, norm_len & 511, norm_len >> 9));
    const int in1 = (int)MB_BA_IN1(ba);
    const int lb_src = in1 ? in1 - 1 : -1;
    const int lb_out = MB_BA_OUT(ba) ? band_start - norm_offset : -1;
    if (MB_BA_DUAL(ba)) {
      x_cm = plan_band(P, 0, n, B, lb_src, 0, lb_out, 0, 1.0f, x_cm);
      y_cm = plan_band(P, WIN_Y, n, B, lb_src, 1, lb_out, 1, 1.0f, y_cm);
    } else if (C == 2) {
      x_cm = plan_band_stereo(P, n, B, lb_src, lb_out, x_cm | y_cm);
      y_cm = x_cm;
    } else {
      x_cm = plan_band(P, 0, n, B, lb_src, 0, lb_out, 0, 1.0f, x_cm | y_cm);
      y_cm = x_cm;
    }
    collapse[i * C] = (uint8_t)x_cm;
    collapse[i * C + C - 1] = (uint8_t)y_cm;
    anti_pos[i] = (uint16_t)imin(P.n_ops, 0xffff);
    if (acon)
      for (int c = 0; c < C; c++) { P.put(MB_OP(OP_SKIP, 0, 2, 0)); P.put(0); P.put(0); }
    P.put(MB_OP(OP_END_BAND, 0, 0, 0));
  }
  // ---- anti_collapse plan (bands.rs:3220-3349): which blocks of which band get noise, at which seed, at which level
  if (acon) {
    const float* E = Q.en + (size_t)f * EN_STRIDE;
    const float* prev1 = E + 2 * NB_EBANDS;
    const float* prev2 = E + 4 * NB_EBANDS;
    uint32_t seed = P.seed;
    for (int i = start; i < end; i++) {
      const int N0 = eband(i + 1) - eband(i);
      const int pulses_i = (int)r[(S_BANDB + i) * 32];
      const int depth = (int)((uint32_t)(1 + pulses_i) / (uint32_t)N0) >> lm;
      const float thresh = 0.5f * celt_exp2(-0.125f * (float)depth);
      const float sqrt_1 = celt_rsqrt((float)(N0 << lm));
      for (int c = 0; c < C; c++) {
        const uint32_t missing = ~(uint32_t)collapse[i * C + c] & mask_from_bits(M);
        if (!missing) continue;
        float p1 = prev1[c * NB_EBANDS + i], p2 = prev2[c * NB_EBANDS + i];
        if (C == 1) { p1 = fmax32(p1, prev1[NB_EBANDS + i]); p2 = fmax32(p2, prev2[NB_EBANDS + i]); }
        float Ediff = E[c * NB_EBANDS + i] - fmin32(p1, p2);
        if (Ediff < 0.f) Ediff = 0.f;
        float rr = 2.0f * celt_exp2(-Ediff);
        if (lm == 3) rr *= 1.41421356f;
        rr = fmin32(rr, thresh);
        rr *= sqrt_1;
        const int pos = (int)anti_pos[i] + 3 * c;
        P.set(pos, MB_OP(OP_ANTI, c * WIN_Y, N0 << lm, (uint32_t)lm | (missing << 2)));
        P.set(pos + 1, seed);
        P.set(pos + 2, f2bits(rr));
        for (int k = 0; k < M; k++)
          if (missing & (1u << k)) seed = lcg_jump(seed, N0);
      }
    }
  }
  r[S_PLAN * 32] = P.overflow ? 0xffffffffu : (uint32_t)P.n_ops;
}

}  // namespace mb
