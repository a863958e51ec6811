// Phase 2b(iii) — band resynthesis: one warp per frame executes the frame's operation list (mb_celt_plan.cuh) band by
// band on a small shared-memory window, then scales the band to its energy.
//
// Replaces the float half of
//   quant_band / quant_band_stereo / quant_partition resynthesis (fold, noise fill, haar1, hadamard
//   (de)interleave, stereo_merge)            src/celt/bands.rs:1917-3082, 3477-3505, 3708-3836
//   renormalise_vector                       src/celt/vq.rs:1073-1088
//   anti_collapse (noise injection)          src/celt/bands.rs:3220-3349
//   denormalise_bands                        src/celt/bands.rs:4172-4234
// Only the band being rebuilt (<= 176 bins per channel) and the fold buffer live on chip; the spectrum row itself
// stays in the spectrum buffer, where the PVQ pass left the pulse leaves and this pass leaves `freq`, in place.
#pragma once
#include "mb_celt_plan.cuh"

namespace mb {

MB_TBL uint8_t kOrdery[30] = {1, 0, 3, 0, 2, 1, 7, 0, 4, 3, 6, 1, 5, 2, 15, 0, 8, 7, 12, 3, 11, 4, 14, 1, 9, 6, 13, 2, 10, 5};

// Shared-memory working set of one frame: 9.4 KB.
enum { NORM_MAX = 624 };       // 8 * eband(20): the longest fold buffer (bands.rs:2614-2617)
struct alignas(16) BandsSmem {
  float win2[2][2 * WIN_Y];    // x | y of the current band, and of the next one on its way in (cp.async)
  float scratch[176];          // transformed copy of the fold source
  float tmp[176];              // products on their way into a serial sum (and the permutation scratch of the host emulation)
  float norm[2 * NORM_MAX];    // fold buffers (norm | norm2)
  float gains[2 * NB_EBANDS];
};

// Window of operation words: lane l holds words (wb + l) and (wb + 32 + l); broadcast by shuffle.
struct OpReader {
  const uint32_t* ops;
  int n_ops, pos, wb, tid;
  uint32_t c0, c1;
  MB_MEM uint32_t fetch(int q) const {
#if MB_ON_DEVICE
    return q < n_ops ? __ldg(ops + (size_t)q * 32) : 0u;
#else
    return 0u;
#endif
  }
  MB_MEM void open(const uint32_t* o, int n, int lane) {
    ops = o; n_ops = n; pos = 0; wb = 0; tid = lane;
    c0 = fetch(lane); c1 = fetch(lane + 32);
  }
  MB_MEM uint32_t next() {
    const int p = pos++;
#if MB_ON_DEVICE
    if (p - wb >= 32) { wb += 32; c0 = c1; c1 = fetch(wb + 32 + tid); }
    return __shfl_sync(0xffffffffu, c0, p - wb);
#else
    return ops[(size_t)p * 32];
#endif
  }
  MB_MEM void skip(int n) {
    pos += n;
#if MB_ON_DEVICE
    if (pos - wb >= 64) { wb = pos & ~31; c0 = fetch(wb + tid); c1 = fetch(wb + 32 + tid); }
#endif
  }
};

// Left-to-right sum of p[0..n) — the reference's summation order (renormalise_vector, stereo_merge), which fixes the
// rounding. The products were formed by the whole team; what is left for one lane is the chain of dependent additions,
// fed by 128-bit shared-memory loads. p must be 16-byte aligned.
MB_DEV float serial_sum(const float* p, int n) {
  float e = 0.f;
  int i = 0;
#if MB_ON_DEVICE
  MB_ROLLED
  for (; i + 4 <= n; i += 4) {
    const float4 v = *(const float4*)(p + i);
    e += v.x; e += v.y; e += v.z; e += v.w;
  }
#endif
  MB_ROLLED
  for (; i < n; i++) e += p[i];
  return e;
}
MB_DEV float team_bcast(Team tm, float v, int src) {
#if MB_ON_DEVICE
  return __shfl_sync(0xffffffffu, v, src);
#else
  return v;
#endif
}

MB_DEV_NOINLINE void team_haar1(Team tm, float* x, int n0, int stride) {
  if (n0 < 2) return;
  const int half = n0 / 2;
  const float scale = 0.70710678f;
  const int total = stride * half;
  const int sl = ilog2_exact(stride);   // stride is a power of two; the butterflies are independent, any order will do
  MB_ROLLED
  for (int idx = tm.tid; idx < total; idx += tm.n) {
    int i = idx & (stride - 1), j = idx >> sl;
    int i0 = stride * 2 * j + i, i1 = i0 + stride;
    float t1 = scale * x[i0], t2 = scale * x[i1];
    x[i0] = t1 + t2;
    x[i1] = t1 - t2;
  }
  MB_SYNC();
}
MB_DEV const uint8_t* ordery_for(int stride) { return kOrdery + (stride == 2 ? 0 : stride == 4 ? 2 : stride == 8 ? 6 : 14); }
// (de)interleave_hadamard: a permutation of at most 176 samples, through the scratch row
MB_DEV_NOINLINE void team_permute(Team tm, float* x, float* tmp, int n0, int stride, int hadamard, int inverse) {
  const int n = n0 * stride;
  const uint8_t* ord = ordery_for(stride);
  const int sl = ilog2_exact(stride);
  MB_ROLLED
  MB_ROLLED
  for (int idx = tm.tid; idx < n; idx += tm.n) {
    const int i = idx & (stride - 1), j = idx >> sl;
    const int o = (hadamard ? (int)ord[i] : i) * n0 + j;
    if (inverse) tmp[idx] = x[o]; else tmp[o] = x[idx];      // interleave: x'[idx] = x[o];  deinterleave: x'[o] = x[idx]
  }
  MB_SYNC();
  MB_ROLLED
  MB_ROLLED
  for (int idx = tm.tid; idx < n; idx += tm.n) x[idx] = tmp[idx];
  MB_SYNC();
}
MB_DEV void team_deinterleave(Team tm, float* x, float* tmp, int n0, int stride, int hadamard) { team_permute(tm, x, tmp, n0, stride, hadamard, 0); }
MB_DEV void team_interleave(Team tm, float* x, float* tmp, int n0, int stride, int hadamard) { team_permute(tm, x, tmp, n0, stride, hadamard, 1); }
// renormalise_vector (vq.rs:1073-1088): the energy sum keeps the reference's left-to-right order
#ifndef MB_BANDS_SUM_MODE
#define MB_BANDS_SUM_MODE 0   // 0: one lane multiplies and adds (measured faster: 13.7 vs 14.1 ms per step); 1: the team multiplies, one lane adds (128-bit loads)
#endif
MB_DEV_NOINLINE void team_renormalise(Team tm, BandsSmem* sm, float* x, int n, float gain) {
  MB_SYNC();
  float scale = 0.f;
#if MB_BANDS_SUM_MODE == 1
  MB_ROLLED
  for (int i = tm.tid; i < n; i += tm.n) { const float v = x[i]; sm->tmp[i] = v * v; }
  MB_SYNC();
  if (tm.tid == 0) scale = celt_rsqrt(1e-15f + serial_sum(sm->tmp, n)) * gain;
#else
  if (tm.tid == 0) {
    float e = 0.f;
    for (int i = 0; i < n; i++) e += x[i] * x[i];
    scale = celt_rsqrt(1e-15f + e) * gain;
  }
#endif
  scale = team_bcast(tm, scale, 0);
  MB_ROLLED
  for (int i = tm.tid; i < n; i += tm.n) x[i] *= scale;
  MB_SYNC();
}

struct BandsParams {
  const uint32_t* rec;
  const float* en;
  const FrameMeta* meta;
  float* xbuf;              // [frame][XBUF_STRIDE]: pulse leaves in, freq out
  float* x_dbg;             // optional [frame][2*960] normalised spectrum after anti-collapse
  float* fold_all = nullptr;           // [stream][FOLD_WORDS]: write-back target of the parallel pass (see celt_bands_frame)
  const uint32_t* lane_stream = nullptr;
};

// fold: the stream's persistent fold buffer (FOLD_WORDS floats) for hybrid frames, else null.  The reference keeps
// `scratch.norm_storage` in the decoder and never clears it (src/celt/bands.rs:2614-2625, src/celt/types.rs:400,416):
// a hybrid frame starts at band 17, so its first folds read what EARLIER frames of the stream left there.  Frames that
// carry a MB_FL_FOLD flag (hybrid; the CELT frames that share a SILK packet's cell; CELT-only frames behind one of those in
// the same pass) therefore run only when `fold` is given (tick order per stream); the parallel pass skips them and, for
// the CELT-only frames it runs, stores the words of the buffer that survive the pass (a later hybrid frame may read them).
MB_DEV void celt_bands_frame(Team tm, BandsSmem* sm, uint32_t f, const BandsParams& P, bool serial = false) {
  const FrameMeta meta = P.meta[f];
  if (!(meta.flags & 1u) || (meta.flags & MB_FL_PLC_PITCH)) return;   // (a frame concealed by pitch repetition has no spectrum)
  if (((meta.flags & MB_FL_FOLD) != 0) != serial) return;
  const uint32_t* r = P.rec + rec_index(f, 0);
  const uint32_t h0 = r[S_HDR0 * 32], status = r[S_STATUS * 32], plan = r[S_PLAN * 32];
  const int lm = (int)MB_H0_LM(h0), C = (int)MB_H0_C(h0), start = (int)MB_H0_START(h0), end = (int)MB_ST_END(status);
  const int M = 1 << lm, N = M * SHORT_MDCT;
  if (plan == 0xffffffffu) return;   // operation list overflow (cannot happen for legal trees; the frame stays silent)
  float* xrow = P.xbuf + (size_t)f * XBUF_STRIDE;
  const float* E = P.en + (size_t)f * EN_STRIDE;
  float* norm = sm->norm;

  // (CELT-only frames start at band 0: every fold source is written earlier in the same frame, so the fold buffer
  //  needs no cross-frame persistence)
  const int norm_len = M * (eband(NB_EBANDS - 1) - eband(start));   // the reference's norm2 = norm + norm_len
  // sm->norm is the reference's flat storage: norm at 0, norm2 at norm_len. On the serial path it is the stream's buffer
  // (loaded by the caller, left as the previous frame left it); the parallel path starts from zeros.
  float* norm2 = norm + norm_len;
  if (!serial) {
    MB_ROLLED
    for (int i = tm.tid; i < 2 * NORM_MAX; i += tm.n) norm[i] = 0.f;
  }
  // band gains exp2(min(E + eMeans, 32)) (bands.rs:4172-4234): computed by the stream pass, which holds E anyway
  MB_ROLLED
  for (int idx = tm.tid; idx < C * NB_EBANDS; idx += tm.n) sm->gains[idx] = E[EN_GAINS + idx];
  OpReader ops;
  ops.open(r + S_OPS * 32, (int)plan, tm.tid);
  MB_SYNC();
  // The band window is double-buffered: while band b is rebuilt, band b+1 streams in from the spectrum row (cp.async),
  // so the global-memory latency of the row is paid once per frame instead of once per band.
  auto fetch = [&](int band, float* w) {
    const int b0 = M * eband(band), nb = M * eband(band + 1) - b0;
    MB_ROLLED
    for (int idx = tm.tid; idx < C * nb; idx += tm.n) {
      int c = idx >= nb, j = idx - c * nb;
#if MB_ON_DEVICE
      const uint32_t sa = (uint32_t)__cvta_generic_to_shared(w + c * WIN_Y + j);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(sa), "l"(xrow + c * N + b0 + j));
#else
      w[c * WIN_Y + j] = xrow[c * N + b0 + j];
#endif
    }
#if MB_ON_DEVICE
    asm volatile("cp.async.commit_group;\n" ::);
#endif
  };
  if (start < end) fetch(start, sm->win2[0]);
  for (int band = start; band < end; band++) {
    const int band_start = M * eband(band), n = M * eband(band + 1) - band_start;
    float* const win = sm->win2[(band - start) & 1];
    if (band + 1 < end) fetch(band + 1, sm->win2[(band + 1 - start) & 1]);
#if MB_ON_DEVICE
    if (band + 1 < end) asm volatile("cp.async.wait_group 1;\n" ::); else asm volatile("cp.async.wait_group 0;\n" ::);
#endif
    MB_SYNC();
    for (;;) {
      const uint32_t op = ops.next();
      const uint32_t type = MB_OP_TYPE(op);
      if (type == OP_END_BAND) break;
      const int a = (int)MB_OP_A(op), len = (int)MB_OP_N(op);
      const uint32_t p = MB_OP_P(op);
      float* x = win + a;
      switch (type) {
        case OP_SKIP: ops.skip(len); break;
        case OP_COPY_LB: {
          const uint32_t w1 = ops.next();
          float* src = ((w1 >> 11) & 1u ? norm2 : norm) + (w1 & 2047u);
          float* from = (p & 1u) ? sm->scratch : src;
          float* to = (p & 1u) ? src : sm->scratch;
          MB_ROLLED
          for (int j = tm.tid; j < len; j += tm.n) to[j] = from[j];
          MB_SYNC();
        } break;
        case OP_HAAR: {
          const int stride = 1 << (p >> 1);
          team_haar1(tm, (p & 1u) ? sm->scratch : x, len / stride, stride);
        } break;
        case OP_DEINT: {
          const int stride = 1 << ((p >> 1) & 7u);
          team_deinterleave(tm, (p & 1u) ? sm->scratch : x, sm->tmp, len / stride, stride, (int)(p >> 4) & 1);
        } break;
        case OP_INTER: {
          const int stride = 1 << ((p >> 1) & 7u);
          team_interleave(tm, (p & 1u) ? sm->scratch : x, sm->tmp, len / stride, stride, (int)(p >> 4) & 1);
        } break;
        case OP_ZERO:
          MB_ROLLED
          for (int j = tm.tid; j < len; j += tm.n) x[j] = 0.f;
          MB_SYNC();
          break;
        case OP_FOLD: {
          const float gain = bits2f(ops.next());
          const uint32_t seed = ops.next();
          const float* lb = sm->scratch + p;
          MB_ROLLED
          for (int j = tm.tid; j < len; j += tm.n) {
            uint32_t s = lcg_jump(seed, j + 1);
            x[j] = lb[j] + ((s & 0x8000u) ? 1.0f / 256.0f : -(1.0f / 256.0f));
          }
          team_renormalise(tm, sm, x, len, gain);
        } break;
        case OP_NOISE: {
          const float gain = bits2f(ops.next());
          const uint32_t seed = ops.next();
          MB_ROLLED
          for (int j = tm.tid; j < len; j += tm.n) {
            uint32_t s = lcg_jump(seed, j + 1);
            x[j] = (float)((int32_t)s >> 20);
          }
          team_renormalise(tm, sm, x, len, gain);
        } break;
        case OP_N1:
          if (tm.tid == 0) x[0] = (p & 1u) ? -1.0f : 1.0f;
          MB_SYNC();
          break;
        case OP_LBOUT: {
          const uint32_t w1 = ops.next();
          float* dst = ((w1 >> 11) & 1u ? norm2 : norm) + (w1 & 2047u);
          const float scale = celt_sqrt((float)len);
          MB_ROLLED
          for (int j = tm.tid; j < len; j += tm.n) dst[j] = scale * x[j];
          MB_SYNC();
        } break;
        case OP_MERGE: {  // stereo_merge (bands.rs:3477-3505)
          const float mid = (float)(int)(ops.next() & 0xffffu) / 32768.0f;
          float* xx = win;
          float* yy = win + WIN_Y;
          // the two inner products keep the reference's summation order, one lane per sum
          float acc = 0.f, xp, sd;
#if MB_BANDS_SUM_MODE == 1
          // (products by the team; the fold source in `scratch` is dead by the time a band's two channels are merged)
          MB_ROLLED
          for (int j = tm.tid; j < len; j += tm.n) { const float yv = yy[j]; sm->tmp[j] = yv * xx[j]; sm->scratch[j] = yv * yv; }
          MB_SYNC();
          if (tm.tid < 2) acc = serial_sum(tm.tid ? sm->scratch : sm->tmp, len);
          if (tm.n == 1) { xp = acc * mid; sd = serial_sum(sm->scratch, len); }
          else { xp = team_bcast(tm, acc, 0) * mid; sd = team_bcast(tm, acc, 1); }
#else
          if (tm.tid < 2) { const float* o = tm.tid ? yy : xx; for (int j = 0; j < len; j++) acc += yy[j] * o[j]; }
          if (tm.n == 1) { xp = acc * mid; sd = 0.f; for (int j = 0; j < len; j++) sd += yy[j] * yy[j]; }
          else { xp = team_bcast(tm, acc, 0) * mid; sd = team_bcast(tm, acc, 1); }
#endif
          const float mid2 = mid * mid;
          const float el = mid2 + sd - 2.0f * xp, er = mid2 + sd + 2.0f * xp;
          const int degenerate = er < 6e-4f || el < 6e-4f;
          const float lgain = degenerate ? 0.f : celt_rsqrt(el), rgain = degenerate ? 0.f : celt_rsqrt(er);
          const int inv = (int)(p & 1u);
          MB_ROLLED
          for (int j = tm.tid; j < len; j += tm.n) {
            float yv;
            if (degenerate) yv = xx[j];
            else {
              float l = mid * xx[j], rv = yy[j];
              xx[j] = lgain * (l - rv);
              yv = rgain * (l + rv);
            }
            yy[j] = inv ? -yv : yv;
          }
          MB_SYNC();
        } break;
        case OP_STEREO2: {
          const uint32_t w1 = ops.next();
          if (tm.tid == 0) {
            const float mid = (float)(int)(w1 & 0xffffu) / 32768.0f, side = (float)(int)(w1 >> 16) / 32768.0f;
            float* xx = win;
            float* yy = win + WIN_Y;
            const int c = (int)(p & 1u);
            float* x2 = c ? yy : xx;
            float* y2 = c ? xx : yy;
            const int sv = 1 - 2 * (int)((p >> 1) & 1u);
            y2[0] = -(float)sv * x2[1];
            y2[1] = (float)sv * x2[0];
            xx[0] *= mid; xx[1] *= mid; yy[0] *= side; yy[1] *= side;
            float t = xx[0]; xx[0] = t - yy[0]; yy[0] += t;
            t = xx[1]; xx[1] = t - yy[1]; yy[1] += t;
            if ((p >> 2) & 1u) { yy[0] = -yy[0]; yy[1] = -yy[1]; }
          }
          MB_SYNC();
        } break;
        case OP_AVG: {
          const int cnt = len | ((int)p << 9);
          MB_ROLLED
          for (int j = tm.tid; j < cnt; j += tm.n) norm[j] = (norm[j] + norm2[j]) * 0.5f;
          MB_SYNC();
        } break;
        case OP_HYB: {
          const uint32_t w1 = ops.next();
          const int n1 = (int)(w1 & 2047u), cnt = (int)((w1 >> 11) & 2047u), dual = (int)((w1 >> 22) & 1u);
          const int n2 = n1 + cnt;
          MB_ROLLED
          for (int j = tm.tid; j < cnt; j += tm.n) {
            norm[n1 + j] = norm[2 * n1 - n2 + j];
            if (dual) norm2[n1 + j] = norm2[2 * n1 - n2 + j];
          }
          MB_SYNC();
        } break;
        case OP_ANTI: {
          uint32_t seed = ops.next();
          const float rr = bits2f(ops.next());
          const int alm = (int)(p & 3u), N0 = len >> alm;
          const uint32_t missing = p >> 2;
          for (int k = 0; k < (1 << alm); k++) {
            if (!(missing & (1u << k))) continue;
            MB_ROLLED
            for (int j = tm.tid; j < N0; j += tm.n) {
              uint32_t s = lcg_jump(seed, j + 1);
              x[(j << alm) + k] = (s & 0x8000u) ? rr : -rr;
            }
            seed = lcg_jump(seed, N0);
          }
          team_renormalise(tm, sm, x, len, 1.0f);
        } break;
        default: break;
      }
    }
    // ---- denormalise_bands for this band, in place in the spectrum row
    if (P.x_dbg)
      MB_ROLLED
      for (int idx = tm.tid; idx < C * n; idx += tm.n) {
        int c = idx >= n, j = idx - c * n;
        P.x_dbg[(size_t)f * 2 * MAX_N + c * MAX_N + band_start + j] = win[c * WIN_Y + j];
      }
    MB_ROLLED
    for (int idx = tm.tid; idx < C * n; idx += tm.n) {
      int c = idx >= n, j = idx - c * n;
      xrow[c * N + band_start + j] = win[c * WIN_Y + j] * sm->gains[c * NB_EBANDS + band];
    }
    MB_SYNC();
  }
  // write back on the parallel path: the words of this frame's footprint that no later frame of the pass overwrites
  // (k_energy worked the interval out: flags bits 10-18 / 19-27, in units of 4 words)
  if (!serial && P.fold_all) {
    const int lo = 4 * (int)((meta.flags >> 10) & 511u), hi = (meta.flags & MB_FL_PLC) ? 0 : 4 * (int)((meta.flags >> 19) & 511u);
    if (hi > lo) {
      float* dst = P.fold_all + (size_t)(P.lane_stream ? P.lane_stream[meta.lane] : meta.lane) * FOLD_WORDS;
      for (int w = lo + tm.tid; w < hi; w += tm.n) dst[w] = norm[w];
      MB_SYNC();
    }
  }
}

}  // namespace mb
