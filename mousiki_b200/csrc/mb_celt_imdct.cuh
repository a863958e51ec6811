// Phase 2b(iv) — inverse MDCT: one warp per (frame, channel), every frame of every stream concurrently.
//
// Replaces:
//   celt_synthesis / apply_inverse_mdct       src/celt/celt_decoder.rs:473-507, 573-800
//   clt_mdct_backward, pre/post rotation      src/celt/mdct.rs:197-388 (with the reference's zero-filled overlap)
//   KissFftState::ifft, kf_bfly2/3/4/5        src/celt/kiss_fft.rs:729-795, 900-1340
// The spectrum row (freq, written in place by the band pass) is staged in shared memory, pre-rotated into bit-reversed
// order, transformed by a shared-memory mixed-radix FFT that keeps the static mode's radix order (4,2,4,3,5 for 480
// points — the rounding order is part of parity), post-rotated, folded against the window, and leaves as one coalesced
// row of the lane's time series.
#pragma once
#include "mb_celt_bands.cuh"

namespace mb {

MB_TBL uint8_t kFftFactors[4][10] = {{5, 96, 3, 32, 4, 8, 2, 4, 4, 1}, {5, 48, 3, 16, 4, 4, 4, 1, 0, 0}, {5, 24, 3, 8, 2, 4, 4, 1, 0, 0}, {5, 12, 3, 4, 4, 1, 0, 0, 0, 0}};
MB_TBL uint16_t kMdctTwOff[4] = {0, 960, 1440, 1680};

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
    // (idx -> transform, group, position without integer division: every m below the radix-5 stage is a power of two, the
    //  radix-5 stage has a single group, and idx / per over this small range is exact in float)
    const int lm2 = ilog2_exact(m);
    const float inv_per = 1.0f / (float)per;
    for (int idx = tm.tid; idx < per * nt; idx += tm.n) {
      int t = 0, r = idx;
      if (nt > 1) { t = (int)(((float)idx + 0.5f) * inv_per); r = idx - t * per; }
      int g = 0, k = r;
      if (groups > 1) { g = r >> lm2; k = r & (m - 1); }
      bfly_one(f + t * nfft + g * mm, radix, k, m, fs);
    }
    MB_SYNC();
    m = mm;
  }
}
MB_DEV int div_small(int idx, int d, float inv_d) { (void)d; return (int)(((float)idx + 0.5f) * inv_d); }   // exact for idx < 4096, d >= 12

struct ImdctSmem {
  float io[MAX_N];     // freq in, time-domain samples out
  float F[MAX_N];      // FFT work: 480 complex points
  float tail[64];      // the 60 samples the last block writes beyond the frame (the reference's overlap region, clt_mdct_backward
                       // writes n2 + overlap / 2 samples): only a frame concealed by pitch repetition ever reads them
};

struct ImdctParams {
  const uint32_t* rec;
  const FrameMeta* meta;
  const float* xbuf;        // [frame][XBUF_STRIDE] freq
  float* series;            // [lane][CC][series_stride]
  uint32_t series_stride;
  int CC;
  // Overlap-add form (quirk Q2 cleared: clt_mdct_backward as libopus has it, without the reference's zero fill, mdct.rs:373):
  // a block's first 60 samples take in the 60 unwindowed samples the block before it left behind. Between the short blocks of
  // a frame that happens here; across frames the stream's post-filter pass (serial in time, celt_comb_channel) adds the term
  // from the tail every unit stores in `tails` ([cell][CC][TAIL_STRIDE]) and clamps the frame's first 120 samples.
  float* tails = nullptr;
  int ola = 0;
  int ds = 1;               // output rate 48000 / ds: bins from N / ds on are cleared (denormalise_bands, bands.rs:4198-4204)
};
enum { TAIL_STRIDE = 64 };

// The transform itself, in place on sm->io: N = 120 << lm frequency bins in (short blocks interleaved as the band pass
// leaves them), N time-domain samples out (clt_mdct_backward against the reference's zero-filled overlap).
MB_DEV void celt_imdct_core(Team tm, ImdctSmem* sm, int lm, int transient) {
  const int M = 1 << lm, N = M * SHORT_MDCT;
  float* X = sm->io;
  const int nbk = transient ? M : 1;                  // short blocks
  const int shift = transient ? MAX_LM : MAX_LM - lm;
  const int n2 = 960 >> shift, n4 = n2 >> 1;          // per-block MDCT size / FFT size
  cpx* F = (cpx*)sm->F;
  const uint32_t* twb = kMdctTwiddles960Bits + kMdctTwOff[shift];
  const int16_t* br = shift == 0 ? kFftBitrev480 : shift == 1 ? kFftBitrev240 : shift == 2 ? kFftBitrev120 : kFftBitrev60;
  // pre-rotation + bit-reversal + conjugate (ifft = conj . fft . conj)
  const float inv_n4 = 1.0f / (float)n4;
  for (int idx = tm.tid; idx < nbk * n4; idx += tm.n) {
    int b = nbk > 1 ? div_small(idx, n4, inv_n4) : 0, i = idx - b * n4;
    const float* in = X + b;          // element k of block b is in[k * nbk]
    float x1 = in[(2 * i) * nbk], x2 = in[(n2 - 1 - 2 * i) * nbk];
    float t0 = bits2f(twb[i]), t1 = bits2f(twb[n4 + i]);
    cpx v;
    v.r = x2 * t0 + x1 * t1;
    v.i = -(x1 * t0 - x2 * t1);
    F[b * n4 + br[i]] = v;
  }
  MB_SYNC();
  team_fft(tm, F, shift, nbk);
  // post-rotation and window fold against the zero-filled overlap, back into the staging row
  float* O = X;
  for (int idx = tm.tid; idx < nbk * n4; idx += tm.n) {
    int b = nbk > 1 ? div_small(idx, n4, inv_n4) : 0, i = idx - b * n4;
    cpx v = F[b * n4 + i];
    v.i = -v.i;
    float t0 = bits2f(twb[i]), t1 = bits2f(twb[n4 + i]);
    float yr = v.r * t0 + v.i * t1;     // temp[2i]
    float yi = v.r * t1 - v.i * t0;     // temp[n2-1-2i]
    float* ob = O + b * n2;
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
      } else if (b == nbk - 1) {
        sm->tail[b * n2 + 60 + k - N] = val;
      }
    }
  }
  MB_SYNC();
}

// ---------------------------------------------------------------- warp-resident form: what k_imdct runs
// The same butterflies in the same order (each one's arithmetic is what parity needs; which thread computes it is free),
// laid out so that a warp keeps the whole transform in registers.  kiss_fft's input permutation puts input i = 15 r + c
// (c = d1 + 5 d2) at position d1 * 3L + d2 * L + w of a transform of 15 L points (L = 32 >> shift): lane r of a block
// therefore pre-rotates 15 consecutive inputs; the radix-4 / radix-2 stages below the radix-3 stage (a transform of L
// points across the L lanes, one per c) run on warp shuffles — a radix-4 butterfly as two exchange steps (partner two
// digits away: a0 +- a2, a1 +- a3; then the neighbour: the four outputs), which leaves outputs 1 and 2 in each other's
// lanes, so that after all stages lane r holds position w = bit-reversed r — and the radix-3 and radix-5 stages find
// all their operands in the lane's own 15 registers.  Short blocks of a transient frame take 4 lanes each: 8 of them
// fill the warp.  Twiddles and the window come from one 11.5 KB block in shared memory that a CTA fetches once with a
// bulk async copy; the spectrum rows arrive the same way, one row ahead of the transform.
struct ImdctTables {
  float mdct_tw[1800];   // kMdctTwiddles960Bits: N = 1920 / 960 / 480 / 240 back to back
  float fft_tw[960];     // kFftTwiddles480Bits: (re, im) pairs
  float window[120];     // kWindow120Bits
};
static_assert(sizeof(ImdctTables) % 16 == 0, "bulk copies move multiples of 16 bytes");

enum { IMDCT_ROW = MAX_N + 64 };   // freq in, time-domain samples out, then the 60 samples the last block writes beyond the
                                   // frame (the reference's overlap region; only a pitch concealment ever reads them)

#if MB_ON_DEVICE
enum { MB_NL = 1 };
#define MB_EACH_LANE(tm, ln) for (int ln = (tm).tid, mb_once_ = 1; mb_once_; mb_once_ = 0)
#define MB_LIDX(ln) 0
MB_DEV cpx lane_xchg(const cpx* snd, int ln, int mask) {
  cpx o;
  o.r = __shfl_xor_sync(0xffffffffu, snd[0].r, mask);
  o.i = __shfl_xor_sync(0xffffffffu, snd[0].i, mask);
  return o;
}
#else
enum { MB_NL = 32 };   // the host emulation walks the 32 lanes of the warp phase by phase
#define MB_EACH_LANE(tm, ln) for (int ln = 0; ln < 32; ln++)
#define MB_LIDX(ln) (ln)
MB_DEV cpx lane_xchg(const cpx* snd, int ln, int mask) { return snd[ln ^ mask]; }
#endif

MB_DEV float neg_if(float x, bool c) { return bits2f(f2bits(x) ^ (c ? 0x80000000u : 0u)); }
// A radix-4 butterfly (kf_bfly4) as two exchange steps; x - y is x + (-y) exactly and x + y is y + x.
// Step 1, partner two digits away: the lane whose digit has the high bit clear forms a0 + a2 (or a1 + a3), its partner
// a0 - a2 (a1 - a3): either way what it received plus or minus what it holds.
MB_DEV cpx r4_step1(cpx own, cpx got, bool hi) {
  cpx o;
  o.r = got.r + neg_if(own.r, hi);
  o.i = got.i + neg_if(own.i, hi);
  return o;
}
// Step 2, the neighbour: (sum02, sum13) -> outputs 0 and 2; (dif02, dif13) -> outputs 1 and 3 with dif13 turned by -i,
// which its holder does before sending (r4_turn).
MB_DEV cpx r4_turn(cpx x, bool c) {
  cpx o;
  o.r = c ? x.i : x.r;
  o.i = c ? -x.r : x.i;
  return o;
}
MB_DEV cpx r4_step2(cpx sent, cpx got, bool lo) {
  cpx o;
  o.r = got.r + neg_if(sent.r, lo);
  o.i = got.i + neg_if(sent.i, lo);
  return o;
}
MB_DEV cpx tw_at(const ImdctTables* T, int idx) { cpx o; o.r = T->fft_tw[2 * idx]; o.i = T->fft_tw[2 * idx + 1]; return o; }
MB_DEV int bitrev_low(int r, int bits) {
#if MB_ON_DEVICE
  return (int)(__brev((unsigned)r) >> (32 - bits));
#else
  int w = 0;
  for (int b = 0; b < bits; b++) w |= ((r >> b) & 1) << (bits - 1 - b);
  return w;
#endif
}

// In place on io[0 .. IMDCT_ROW): N = 120 << lm frequency bins in (short blocks interleaved as the band pass leaves
// them), N time-domain samples out (clt_mdct_backward against the reference's zero-filled overlap), then the 60
// unwindowed samples behind the frame.
// SH / NB >= 0 fix the transform size (480 >> SH points) and the number of blocks at compile time (addresses become
// immediates); -1 reads them from (lm, transient).
template <int SH, int NB>
MB_DEV void celt_imdct_core_t(Team tm, float* io, const ImdctTables* T, int lm, int transient, int ola) {
  const int M = 1 << lm;
  const int nbk = NB > 0 ? NB : (transient ? M : 1);
  const int shift = SH >= 0 ? SH : (transient ? MAX_LM : MAX_LM - lm);
  const int N = M * SHORT_MDCT;
  const int n2 = 960 >> shift, n4 = n2 >> 1;
  const int lgL = 5 - shift, L = 1 << lgL;
  const float* twb = T->mdct_tw + (1920 - (1920 >> shift));
  const int pA = 3 - shift;                                   // radix 4, m = 1: the top two bits of r
  const bool hasB = shift == 0 || shift == 2, hasC = shift <= 1;
  const int pB = shift == 0 ? 2 : 0;                          // radix 2, m = 4
  const float t8 = 0.70710678118654752440f;
  cpx v[MB_NL][15], snd[MB_NL], twC[MB_NL];
  int kB[MB_NL];
  MB_EACH_LANE(tm, ln) {
    const int r = ln & (L - 1), w = bitrev_low(r, lgL);
    kB[MB_LIDX(ln)] = ((ln >> pB) & 1) ? (w & 3) : 0;         // the upper operand of the radix-2 stage is turned by W8^k
    twC[MB_LIDX(ln)] = tw_at(T, hasC ? (r & 3) * (w & (shift == 0 ? 7 : 3)) * (15 << shift) : 0);
  }
#pragma unroll
  for (int c = 0; c < 15; c++) {
    // pre-rotation (+ conjugate: ifft = conj . fft . conj) of input 15 r + c
    MB_EACH_LANE(tm, ln) {
      const int blk = ln >> lgL, r = ln & (L - 1);
      cpx x; x.r = 0.f; x.i = 0.f;
      if (blk < nbk) {
        const int i = 15 * r + c;
        const float* in = io + blk;          // element k of block b is in[k * nbk]
        const float x1 = in[(2 * i) * nbk], x2 = in[(n2 - 1 - 2 * i) * nbk];
        const float t0 = twb[i], t1 = twb[n4 + i];
        x.r = x2 * t0 + x1 * t1;
        x.i = -(x1 * t0 - x2 * t1);
      }
      snd[MB_LIDX(ln)] = x;
    }
    MB_EACH_LANE(tm, ln) v[MB_LIDX(ln)][c] = r4_step1(snd[MB_LIDX(ln)], lane_xchg(snd, ln, 2 << pA), ((ln >> (pA + 1)) & 1) != 0);
    MB_EACH_LANE(tm, ln) snd[MB_LIDX(ln)] = r4_turn(v[MB_LIDX(ln)][c], ((ln >> pA) & 3) == 3);
    MB_EACH_LANE(tm, ln) v[MB_LIDX(ln)][c] = r4_step2(snd[MB_LIDX(ln)], lane_xchg(snd, ln, 1 << pA), ((ln >> pA) & 1) != 0);
    if (hasB) {                             // radix 2, m = 4 (kiss_fft.rs:900-966)
      MB_EACH_LANE(tm, ln) {
        const cpx x = v[MB_LIDX(ln)][c];
        const int k = kB[MB_LIDX(ln)];
        const float a = (x.r + x.i) * t8, b = (x.i - x.r) * t8;
        cpx t;
        t.r = (k & 1) ? ((k & 2) ? b : a) : ((k & 2) ? x.i : x.r);
        t.i = (k & 1) ? ((k & 2) ? -a : b) : ((k & 2) ? -x.r : x.i);
        snd[MB_LIDX(ln)] = t;
      }
      MB_EACH_LANE(tm, ln) {
        const cpx mine = snd[MB_LIDX(ln)], got = lane_xchg(snd, ln, 1 << pB);
        v[MB_LIDX(ln)][c].r = got.r + neg_if(mine.r, ((ln >> pB) & 1) != 0);   // p[k] + t below, p[k] - t above
        v[MB_LIDX(ln)][c].i = got.i + neg_if(mine.i, ((ln >> pB) & 1) != 0);
      }
    }
    if (hasC) {                             // radix 4, m = 8 (480 points) or m = 4 (240 points): the two low bits of r
      MB_EACH_LANE(tm, ln) {
        const cpx x = v[MB_LIDX(ln)][c];
        snd[MB_LIDX(ln)] = (ln & 3) ? c_mul(x, twC[MB_LIDX(ln)]) : x;
      }
      MB_EACH_LANE(tm, ln) v[MB_LIDX(ln)][c] = r4_step1(snd[MB_LIDX(ln)], lane_xchg(snd, ln, 2), (ln & 2) != 0);
      MB_EACH_LANE(tm, ln) snd[MB_LIDX(ln)] = r4_turn(v[MB_LIDX(ln)][c], (ln & 3) == 3);
      MB_EACH_LANE(tm, ln) v[MB_LIDX(ln)][c] = r4_step2(snd[MB_LIDX(ln)], lane_xchg(snd, ln, 1), (ln & 1) != 0);
    }
  }
  MB_SYNC();                                // every lane is done reading the input row: the output goes into the same row
  MB_EACH_LANE(tm, ln) {
    const int blk = ln >> lgL, r = ln & (L - 1), w = bitrev_low(r, lgL);
    if (blk < nbk) {
      cpx* p = v[MB_LIDX(ln)];
      {                                     // radix 3, m = L, k = w: p[k] = v[d1], p[m + k] = v[d1 + 5], p[2m + k] = v[d1 + 10]
        const int fs = 5 << shift;
        const cpx tw1 = tw_at(T, w * fs), tw2 = tw_at(T, 2 * w * fs);
        const float epi3i = T->fft_tw[2 * 160 + 1];
#pragma unroll
        for (int d1 = 0; d1 < 5; d1++) {
          const cpx p0 = p[d1];
          cpx s1 = c_mul(p[d1 + 5], tw1), s2 = c_mul(p[d1 + 10], tw2);
          cpx s3 = c_add(s1, s2), s0 = c_sub(s1, s2);
          cpx fm = {p0.r - 0.5f * s3.r, p0.i - 0.5f * s3.i};
          s0.r *= epi3i; s0.i *= epi3i;
          cpx f0 = c_add(p0, s3);
          cpx a = {fm.r + s0.i, fm.i - s0.r}, b = {fm.r - s0.i, fm.i + s0.r};
          p[d1 + 10] = a; p[d1 + 5] = b; p[d1] = f0;
        }
      }
      const cpx ya = tw_at(T, 96), yb = tw_at(T, 192);
      float* ob = io + blk * n2 + 60;       // temp[k] of this block lands at ob[k]
#pragma unroll
      for (int d2 = 0; d2 < 3; d2++) {      // radix 5, m = 3L, k = d2 L + w: p[q m + k] = v[q + 5 d2]
        const int k = d2 * L + w, fs = 1 << shift;
        cpx* q = p + 5 * d2;
        cpx s0 = q[0];
        cpx s1 = c_mul(q[1], tw_at(T, k * fs)), s2 = c_mul(q[2], tw_at(T, 2 * k * fs));
        cpx s3 = c_mul(q[3], tw_at(T, 3 * k * fs)), s4 = c_mul(q[4], tw_at(T, 4 * k * fs));
        cpx s7 = c_add(s1, s4), s10 = c_sub(s1, s4), s8 = c_add(s2, s3), s9 = c_sub(s2, s3);
        cpx f0 = c_add(s0, c_add(s7, s8));
        cpx s5 = {s0.r + fma32(s7.r, ya.r, s8.r * yb.r), s0.i + fma32(s7.i, ya.r, s8.i * yb.r)};
        cpx s6 = {fma32(s10.i, ya.i, s9.i * yb.i), -fma32(s10.r, ya.i, s9.r * yb.i)};
        q[1] = c_sub(s5, s6);
        q[4] = c_add(s5, s6);
        cpx s11 = {s0.r + fma32(s7.r, yb.r, s8.r * ya.r), s0.i + fma32(s7.i, yb.r, s8.i * ya.r)};
        cpx s12 = {fma32(s9.i, ya.i, (-s10.i) * yb.i), fma32(s10.r, yb.i, (-s9.r) * ya.i)};
        q[2] = c_add(s11, s12);
        q[3] = c_sub(s11, s12);
        q[0] = f0;
        // post-rotation: temp[2j] and temp[n2-1-2j].  A short block's upper half lands on the next block's first 60
        // samples, which that block's fold below rewrites (the reference's zero-filled overlap, Q2); the last block's
        // upper half is the overlap region behind the frame.
#pragma unroll
        for (int d1 = 0; d1 < 5; d1++) {
          const int j = d1 * 3 * L + k;
          cpx y = q[d1];
          y.i = -y.i;
          const float t0 = twb[j], t1 = twb[n4 + j];
          ob[2 * j] = y.r * t0 + y.i * t1;
          ob[n2 - 1 - 2 * j] = y.r * t1 - y.i * t0;
        }
      }
    }
  }
  MB_SYNC();
  // window fold of each block's first 60 samples against the zero-filled overlap — or, in the overlap-add form, against the
  // 60 unwindowed samples the block before it wrote there (the frame's first block meets the previous frame's later)
  for (int b = 0; b < nbk; b++) {
    float* ob = io + b * n2;
    for (int k = tm.tid; k < 60; k += tm.n) {
      const float val = ob[60 + k];
      const float x2 = (ola && b > 0) ? ob[59 - k] : 0.f;
      const float w1 = T->window[59 - k], w2 = T->window[60 + k];
      ob[59 - k] = x2 * w2 - val * w1;     // out[i] = x2*w2 - x1*w1
      ob[60 + k] = x2 * w1 + val * w2;     // out[119-i] = x2*w1 + x1*w2
    }
  }
  MB_SYNC();
}

// The two shapes nearly every 20 ms frame takes get their own instantiation; the rest share the generic one.
MB_DEV void celt_imdct_core_w(Team tm, float* io, const ImdctTables* T, int lm, int transient, int ola = 0) {
  if (lm == 3 && !transient) celt_imdct_core_t<0, 1>(tm, io, T, lm, transient, ola);
  else if (lm == 3) celt_imdct_core_t<3, 8>(tm, io, T, lm, transient, ola);
  else celt_imdct_core_t<-1, 0>(tm, io, T, lm, transient, ola);
}

#if !MB_ON_DEVICE
static const ImdctTables* imdct_tables_host() {
  static ImdctTables* t = nullptr;
  if (!t) {
    t = new ImdctTables;
    for (int i = 0; i < 1800; i++) t->mdct_tw[i] = bits2f(kMdctTwiddles960Bits[i]);
    for (int i = 0; i < 960; i++) t->fft_tw[i] = bits2f(kFftTwiddles480Bits[i]);
    for (int i = 0; i < 120; i++) t->window[i] = bits2f(kWindow120Bits[i]);
  }
  return t;
}
#endif

// Unit u of frame f: coded channel u when both the packet and the decoder are stereo, else the single (down-mixed or
// duplicated) spectrum.
MB_DEV void celt_imdct_unit(Team tm, ImdctSmem* sm, uint32_t f, int u, const ImdctParams& P) {
  const FrameMeta meta = P.meta[f];
  if (!(meta.flags & 1u) || (meta.flags & MB_FL_PLC_PITCH)) return;
  const uint32_t* r = P.rec + rec_index(f, 0);
  const uint32_t h0 = r[S_HDR0 * 32], status = r[S_STATUS * 32];
  const int lm = (int)MB_H0_LM(h0), C = (int)MB_H0_C(h0), start = (int)MB_H0_START(h0), end = (int)MB_ST_END(status);
  const int transient = (int)MB_H0_TRANSIENT(h0), silence = (int)MB_H0_SILENCE(h0);
  const int M = 1 << lm, N = M * SHORT_MDCT, CC = P.CC;
  const int nch = (CC == 2 && C == 2) ? 2 : 1;
  if (u >= nch) return;
  const float* xrow = P.xbuf + (size_t)f * XBUF_STRIDE;
  const int planned = r[S_PLAN * 32] != 0xffffffffu;
  const int lo = (silence || !planned) ? 0 : M * eband(start), hi = (silence || !planned) ? 0 : M * eband(imin(end, NB_EBANDS));
  const int downmix = CC == 1 && C == 2;
  float* X = sm->io;
  for (int i = tm.tid; i < N; i += tm.n) {
    float v = 0.f;
    if (i >= lo && i < hi) {
      v = xrow[u * N + i];
      if (downmix) v = 0.5f * (v + xrow[N + i]);
    }
    X[i] = v;
  }
  MB_SYNC();
  celt_imdct_core(tm, sm, lm, transient);
  float* O = X;
  const float SIG_SAT = 536870911.0f;
  for (int cc = 0; cc < CC; cc++) {
    if (nch == 2 && cc != u) continue;
    float* out = P.series + ((size_t)meta.lane * CC + cc) * P.series_stride + SER_PREFIX + meta.sample_off;
    for (int i = tm.tid; i < N; i += tm.n) {
      float o = O[i];
      out[i] = o < -SIG_SAT ? -SIG_SAT : (o > SIG_SAT ? SIG_SAT : o);
    }
    // The frame after this one is concealed by pitch repetition: it reads the decoder's overlap region too, which sits
    // behind the frame in the reference's linear history (decode_mem[2048 .. 2168)) and here at the head of the lost
    // frame's own stretch of the time series: 60 unwindowed samples, then zeros (the reference's zero fill, Q2).
    if (meta.flags & MB_FL_KEEP_TAIL)
      for (int i = tm.tid; i < OVERLAP; i += tm.n) out[N + i] = i < OVERLAP / 2 ? sm->tail[i] : 0.f;
  }
}

// The same unit on the warp-resident core, in two halves so that k_imdct can fetch the next row while this one is
// transformed: imdct_unit_plan reads the headers (what to fetch), imdct_unit_run transforms the staged row and writes it out.
struct ImdctUnit {
  const float* src;      // first spectrum row of the unit in xbuf (nullptr: nothing to do)
  float* out;            // (meta.lane, channel 0) row position in the time series
  uint32_t flags, cell;
  int N, lm, transient, lo, hi, downmix, nch, u;
};
MB_DEV ImdctUnit imdct_unit_plan(uint32_t f, int u, const ImdctParams& P) {
  ImdctUnit U;
  U.src = nullptr;
  const FrameMeta meta = P.meta[f];
  if (!(meta.flags & 1u) || (meta.flags & MB_FL_PLC_PITCH)) return U;
  const uint32_t* r = P.rec + rec_index(f, 0);
  const uint32_t h0 = r[S_HDR0 * 32], status = r[S_STATUS * 32];
  const int lm = (int)MB_H0_LM(h0), C = (int)MB_H0_C(h0), start = (int)MB_H0_START(h0), end = (int)MB_ST_END(status);
  const int silence = (int)MB_H0_SILENCE(h0);
  const int M = 1 << lm, CC = P.CC;
  U.lm = lm; U.transient = (int)MB_H0_TRANSIENT(h0); U.N = M * SHORT_MDCT; U.u = u; U.flags = meta.flags; U.cell = f;
  U.nch = (CC == 2 && C == 2) ? 2 : 1;
  if (u >= U.nch) return U;
  const int planned = r[S_PLAN * 32] != 0xffffffffu;
  U.lo = (silence || !planned) ? 0 : M * eband(start);
  U.hi = (silence || !planned) ? 0 : M * eband(imin(end, NB_EBANDS));
  if (P.ds > 1) U.hi = imin(U.hi, U.N / P.ds);
  U.downmix = CC == 1 && C == 2;
  U.out = P.series + ((size_t)meta.lane * CC) * P.series_stride + SER_PREFIX + meta.sample_off;
  U.src = P.xbuf + (size_t)f * XBUF_STRIDE + u * U.N;
  return U;
}
// io holds the unit's spectrum row (U.src[0 .. N), as fetched); bins outside [lo, hi) are cleared here.
MB_DEV void imdct_unit_run(Team tm, float* io, const ImdctTables* T, const ImdctUnit& U, const ImdctParams& P) {
  const int N = U.N, CC = P.CC;
  if (U.downmix)
    for (int i = tm.tid; i < N; i += tm.n) io[i] = 0.5f * (io[i] + U.src[N + i]);
  for (int i = tm.tid; i < U.lo; i += tm.n) io[i] = 0.f;
  for (int i = (U.hi < U.lo ? U.lo : U.hi) + tm.tid; i < N; i += tm.n) io[i] = 0.f;
  MB_SYNC();
  celt_imdct_core_w(tm, io, T, U.lm, U.transient, P.ola);
  const float SIG_SAT = 536870911.0f;
  for (int cc = 0; cc < CC; cc++) {
    if (U.nch == 2 && cc != U.u) continue;
    float* out = U.out + (size_t)cc * P.series_stride;
    if (P.ola) {
      // overlap-add form: the frame's first 120 samples still miss the previous frame's term — they leave unclamped, the
      // post-filter pass finishes them; the 60 samples behind the frame go to the tail store for the frame that follows
      float* tl = P.tails + ((size_t)U.cell * CC + cc) * TAIL_STRIDE;
      for (int i = tm.tid; i < OVERLAP / 2; i += tm.n) tl[i] = io[N + i];
      for (int i = tm.tid; i < N; i += tm.n) {
        float o = io[i];
        out[i] = i < OVERLAP ? o : (o < -SIG_SAT ? -SIG_SAT : (o > SIG_SAT ? SIG_SAT : o));
      }
      continue;
    }
    if ((((uintptr_t)out) & 15u) == 0) {
      for (int i4 = tm.tid; i4 < N / 4; i4 += tm.n) {
        float4v a = ld_f4(io + 4 * i4);
        a.x = a.x < -SIG_SAT ? -SIG_SAT : (a.x > SIG_SAT ? SIG_SAT : a.x);
        a.y = a.y < -SIG_SAT ? -SIG_SAT : (a.y > SIG_SAT ? SIG_SAT : a.y);
        a.z = a.z < -SIG_SAT ? -SIG_SAT : (a.z > SIG_SAT ? SIG_SAT : a.z);
        a.w = a.w < -SIG_SAT ? -SIG_SAT : (a.w > SIG_SAT ? SIG_SAT : a.w);
        st_f4(out + 4 * i4, a);
      }
    } else {
      for (int i = tm.tid; i < N; i += tm.n) {
        float o = io[i];
        out[i] = o < -SIG_SAT ? -SIG_SAT : (o > SIG_SAT ? SIG_SAT : o);
      }
    }
    if (U.flags & MB_FL_KEEP_TAIL)          // see celt_imdct_unit
      for (int i = tm.tid; i < OVERLAP; i += tm.n) out[N + i] = i < OVERLAP / 2 ? io[N + i] : 0.f;
  }
}
// one unit start to finish (host emulation; the device kernel pipelines the two halves)
MB_DEV void celt_imdct_unit_w(Team tm, float* io, const ImdctTables* T, uint32_t f, int u, const ImdctParams& P) {
  const ImdctUnit U = imdct_unit_plan(f, u, P);
  if (!U.src) return;
  for (int i = tm.tid; i < U.N; i += tm.n) io[i] = U.src[i];
  MB_SYNC();
  imdct_unit_run(tm, io, T, U, P);
}

}  // namespace mb
