// ORACLE (test infrastructure only — never linked into the product library).
// CPU restatement of mousiki's float CELT frame decoder. Each function cites the
// reference lines it follows (paths relative to /root/reference/src/celt unless noted).
// Build: -O2 -ffp-contract=off (Rust never contracts a*b+c; FMAs only where the
// reference calls fmaf explicitly).
#include "celt_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "celt_tables_data.h"

namespace orc {

// ---------------------------------------------------------------- small tables
static const float kEMeans[25] = {  // quant_bands.rs:129-132
    6.4375f, 6.25f, 5.75f, 5.3125f, 5.0625f, 4.8125f, 4.5f, 4.375f, 4.875f, 4.6875f, 4.5625f, 4.4375f, 4.875f,
    4.625f, 4.3125f, 4.5f, 4.375f, 4.625f, 4.75f, 4.4375f, 3.75f, 3.75f, 3.75f, 3.75f, 3.75f};
static const float kPredCoef[4] = {29440.f / 32768.f, 26112.f / 32768.f, 21248.f / 32768.f, 16384.f / 32768.f};
static const float kBetaCoef[4] = {30147.f / 32768.f, 22282.f / 32768.f, 12124.f / 32768.f, 6554.f / 32768.f};
static const float kBetaIntra = 4915.f / 32768.f;
static const uint8_t kSmallEnergyIcdf[3] = {2, 1, 0};
static const uint8_t kTrimIcdf[11] = {126, 124, 119, 109, 87, 41, 19, 9, 4, 2, 0};  // celt_decoder.rs:1865
static const uint8_t kSpreadIcdf[4] = {25, 23, 2, 0};
static const uint8_t kTapsetIcdf[3] = {2, 1, 0};
static const int8_t kTfSelect[4][8] = {  // celt.rs:39-44
    {0, -1, 0, -1, 0, -1, 0, -1}, {0, -1, 0, -2, 1, 0, 1, -1}, {0, -2, 0, -3, 2, 0, 1, -1}, {0, -2, 0, -3, 3, 0, 1, -1}};
static const float kTapsetGains[3][3] = {  // celt.rs:23-27
    {0.3066406250f, 0.2170410156f, 0.1296386719f}, {0.4638671875f, 0.2680664062f, 0.f}, {0.7998046875f, 0.1000976562f, 0.f}};
static const uint8_t kLog2Frac[24] = {0, 8, 13, 16, 19, 21, 23, 24, 26, 27, 28, 29, 30, 31, 32, 32, 33, 34, 34, 35, 36, 36, 37, 37};
static const float kPreemph = 0.8500061035f;  // modes.rs:89 (0.850_006_1)
static const float SIG_SAT = 536870911.0f;
static const float VERY_SMALL = 1.0e-30f;
static const float EPSILON = 1e-15f;

static inline float bits2f(uint32_t b) { float f; memcpy(&f, &b, 4); return f; }
static float g_window[OVERLAP];
static float g_mdct_tw[1800];
static float g_fft_tw[960];
static uint32_t g_pvq_u[1488];
static const int kPvqRowOff[15] = {0, 208, 415, 621, 826, 1030, 1233, 1336, 1389, 1421, 1441, 1455, 1464, 1470, 1473};
static const int kPvqRowLen[15] = {209, 208, 207, 206, 205, 204, 104, 54, 33, 21, 15, 10, 7, 4, 1};

static void init_tables_once() {
  static bool done = false;
  if (done) return;
  for (int i = 0; i < OVERLAP; i++) g_window[i] = bits2f(kWindow120Bits[i]);
  for (int i = 0; i < 1800; i++) g_mdct_tw[i] = bits2f(kMdctTwiddles960Bits[i]);
  for (int i = 0; i < 960; i++) g_fft_tw[i] = bits2f(kFftTwiddles480Bits[i]);
  // cwrs_pvq.rs:19-52: U(n,k) = U(n-1,k) + U(n,k-1) + U(n-1,k-1), saturating u64, truncated to u32.
  static uint64_t t[15][209];
  memset(t, 0, sizeof(t));
  t[0][0] = 1;
  for (int n = 1; n <= 14; n++)
    for (int k = 1; k <= 208; k++) {
      unsigned __int128 s = (unsigned __int128)t[n - 1][k] + t[n][k - 1] + t[n - 1][k - 1];
      t[n][k] = s > UINT64_MAX ? UINT64_MAX : (uint64_t)s;
    }
  int o = 0;
  for (int row = 0; row < 15; row++)
    for (int col = row; col < row + kPvqRowLen[row]; col++) g_pvq_u[o++] = (uint32_t)t[row][col];
  done = true;
}

// ---------------------------------------------------------------- math (math.rs:259-334)
static inline float celt_exp2(float x) { return (float)exp(0.6931471805599453094 * (double)x); }
static inline float celt_sqrt(float x) { return (float)sqrt((double)x); }
static inline float celt_rsqrt(float x) { return 1.0f / celt_sqrt(x); }
// musl cosf evaluates in double and rounds once; (float)cos((double)x) is the closest portable match.
static inline float celt_cos_norm(float x) { return (float)cos((double)(0.5f * 3.14159265358979323846f * x)); }
static inline uint32_t lcg_rand(uint32_t s) { return s * 1664525u + 1013904223u; }
static inline int celt_sudiv(int a, int b) { return a / b; }
static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline float fmin32(float a, float b) { return a < b ? a : b; }  // f32::min for non-NaN
static inline float fmax32(float a, float b) { return a > b ? a : b; }
static inline int eb(int i) { return kEBands5ms[i]; }

static unsigned isqrt32(uint32_t v) {  // math.rs:139-160
  unsigned g = 0;
  int bshift = (ec_ilog(v) - 1) >> 1;
  unsigned b = 1u << bshift;
  do {
    uint32_t t = ((g << 1) + b) << bshift;
    if (t <= v) { g += b; v -= t; }
    b >>= 1; bshift--;
  } while (bshift >= 0);
  return g;
}

// ---------------------------------------------------------------- rate.rs:207-295
static inline int get_pulses(int i) { return i < 8 ? i : (8 + (i & 7)) << ((i >> 3) - 1); }
static inline const uint8_t* cache_row(int band, int lm) {
  int ci = kCacheIndex50[(lm + 1) * NB_EBANDS + band];
  return ci < 0 ? nullptr : kCacheBits50 + ci;
}
static int bits2pulses(int band, int lm, int bits) {
  if (bits <= 0) return 0;
  const uint8_t* t = cache_row(band, lm);
  if (!t) return 0;
  int lo = 0, hi = t[0];
  int target = bits - 1;
  for (int i = 0; i < 6; i++) {
    int mid = (lo + hi + 1) >> 1;
    if ((int)t[mid] >= target) hi = mid; else lo = mid;
  }
  int lo_v = lo == 0 ? -1 : t[lo];
  return target - lo_v <= (int)t[hi] - target ? lo : hi;
}
static int pulses2bits(int band, int lm, int pulses) {
  if (pulses == 0) return 0;
  const uint8_t* t = cache_row(band, lm);
  if (!t) return 0;
  return t[pulses] + 1;
}

static void init_caps(int* cap, int lm, int C) {  // celt.rs:799-822
  for (int i = 0; i < NB_EBANDS; i++) {
    int n = (eb(i + 1) - eb(i)) << lm;
    cap[i] = ((kCacheCaps50[NB_EBANDS * (2 * lm + C - 1) + i] + 64) * C * n) >> 2;
  }
}

// rate.rs:505-838
static int interp_bits2pulses(int start, int end, int skip_start, const int* bits1, const int* bits2,
                              const int* thresh, const int* cap, int total, int* balance, int skip_rsv,
                              int* intensity, int intensity_rsv, int* dual_stereo, int dual_stereo_rsv,
                              int* bits, int* ebits, int* fine_priority, int C, int lm, EcDec* dec, bool q7, bool q9) {
  const int ALLOC_STEPS = 6, MAX_FINE_BITS = 8, FINE_OFFSET = 21;
  int alloc_floor = C << BITRES;
  int stereo = C > 1;
  int logM = lm << BITRES;
  int lo = 0, hi = 1 << ALLOC_STEPS;
  for (int i = 0; i < ALLOC_STEPS; i++) {
    int mid = (lo + hi) >> 1, psum = 0, done = 0;
    for (int j = end; j-- > start;) {
      int tmp = bits1[j] + ((mid * bits2[j]) >> ALLOC_STEPS);
      if (tmp >= thresh[j] || done) { done = 1; psum += imin(tmp, cap[j]); }
      else if (tmp >= alloc_floor) psum += alloc_floor;
    }
    if (psum > total) hi = mid; else lo = mid;
  }
  int psum = 0, done = 0;
  for (int j = end; j-- > start;) {
    int tmp = bits1[j] + ((lo * bits2[j]) >> ALLOC_STEPS);
    if (tmp < thresh[j] && !done) tmp = tmp >= alloc_floor ? alloc_floor : 0;
    else done = 1;
    tmp = imin(tmp, cap[j]);
    bits[j] = tmp;
    psum += tmp;
  }
  int coded = end;
  while (coded > start) {
    int j = coded - 1;
    int band_start = eb(start), band_end = eb(coded), band_prev = eb(j);
    int band_width = band_end - band_prev;
    if (j <= skip_start) { total += skip_rsv; break; }
    int left = total - psum;
    int denom = imax(band_end - band_start, 1);
    int per = (int)((uint32_t)imax(left, 0) / (uint32_t)denom);
    left -= denom * per;
    int rem = imax(left - (band_prev - band_start), 0);
    int band_bits = bits[j] + per * band_width + rem;
    if (band_bits >= imax(thresh[j], alloc_floor + (1 << BITRES))) {
      if (dec->bit_logp(1)) break;
      psum += 1 << BITRES;
      band_bits -= 1 << BITRES;
    }
    psum -= bits[j] + intensity_rsv;
    if (intensity_rsv > 0) intensity_rsv = kLog2Frac[j - start];
    psum += intensity_rsv;
    if (band_bits >= alloc_floor) { psum += alloc_floor; bits[j] = alloc_floor; }
    else bits[j] = 0;
    coded--;
  }
  if (intensity_rsv > 0) *intensity = start + (int)dec->uint((uint32_t)(coded + 1 - start));
  else *intensity = 0;
  if (*intensity <= start) {
    total += dual_stereo_rsv;
    // rate.rs:703-715 (Q9): mousiki keeps the reservation, so the dual-stereo flag is still
    // read when intensity<=start; libopus zeroes it here and skips the read.
    if (!q9) dual_stereo_rsv = 0;
  }
  if (dual_stereo_rsv > 0) *dual_stereo = dec->bit_logp(1);
  else *dual_stereo = 0;
  int denom = imax(eb(coded) - eb(start), 1);
  int left = total - psum;
  int per = (int)((uint32_t)imax(left, 0) / (uint32_t)denom);
  left -= denom * per;
  for (int j = start; j < coded; j++) bits[j] += per * (eb(j + 1) - eb(j));
  for (int j = start; j < coded; j++) {
    int add = imin(eb(j + 1) - eb(j), left);
    bits[j] += add;
    left -= add;
  }
  int bal = 0;
  for (int j = start; j < coded; j++) {
    int N0 = eb(j + 1) - eb(j), N = N0 << lm;
    int bit = bits[j] + bal;
    if (N > 1) {
      int excess = imax(bit - cap[j], 0);
      bits[j] = bit - excess;
      int den = C * N + ((C == 2 && N > 2 && !*dual_stereo && j < *intensity) ? 1 : 0);
      int NClogN = den * (kLogN400[j] + logM);
      int offset = (NClogN >> 1) - den * FINE_OFFSET;
      if (N == 2) offset += den << BITRES >> 2;
      if (bits[j] + offset < den * 2 << BITRES) offset += NClogN >> 2;
      else if (bits[j] + offset < den * 3 << BITRES) offset += NClogN >> 3;
      int e = imax(0, bits[j] + offset + (den << (BITRES - 1)));
      e = (int)((uint32_t)e / (uint32_t)den) >> BITRES;
      // rate.rs:777-779 (Q7): mousiki shifts by `stereo` in the test as well; libopus does not.
      int lim = q7 ? (bits[j] >> stereo) >> BITRES : bits[j] >> BITRES;
      if (C * e > lim) e = bits[j] >> stereo >> BITRES;
      e = imin(e, MAX_FINE_BITS);
      fine_priority[j] = e * (den << BITRES) >= bits[j] + offset;
      bits[j] -= C * e << BITRES;
      ebits[j] = e;
      if (excess > 0) {
        int extra_fine = imin(excess >> (stereo + BITRES), MAX_FINE_BITS - ebits[j]);
        ebits[j] += extra_fine;
        int extra_bits = extra_fine * C << BITRES;
        if (extra_bits >= excess - bal) fine_priority[j] = 1;
        bal = excess - extra_bits;
      } else bal = excess;
    } else {
      int excess = imax(0, bit - (C << BITRES));
      bits[j] = bit - excess;
      ebits[j] = 0;
      fine_priority[j] = 1;
      bal = excess;
    }
  }
  *balance = bal;
  for (int j = coded; j < end; j++) {
    ebits[j] = bits[j] >> stereo >> BITRES;
    bits[j] = 0;
    fine_priority[j] = ebits[j] < 1;
  }
  return coded;
}

// rate.rs:843-1011
static int compute_allocation(int start, int end, const int* offsets, const int* cap, int alloc_trim, int* intensity,
                              int* dual_stereo, int total, int* balance, int* pulses, int* ebits, int* fine_priority,
                              int C, int lm, EcDec* dec, bool q7, bool q9) {
  int bits1[NB_EBANDS] = {0}, bits2[NB_EBANDS] = {0}, thresh[NB_EBANDS] = {0}, trim_offset[NB_EBANDS] = {0};
  total = imax(total, 0);
  int skip_start = start;
  int skip_rsv = 0;
  if (total >= 1 << BITRES) { skip_rsv = 1 << BITRES; total -= skip_rsv; }
  int intensity_rsv = 0, dual_stereo_rsv = 0;
  if (C == 2) {
    int cand = kLog2Frac[end - start];
    if (cand <= total) {
      intensity_rsv = cand;
      total -= intensity_rsv;
      if (total >= 1 << BITRES) { dual_stereo_rsv = 1 << BITRES; total -= dual_stereo_rsv; }
    }
  }
  for (int j = start; j < end; j++) {
    int n = eb(j + 1) - eb(j);
    thresh[j] = imax(C << BITRES, (3 * n) << (lm + BITRES) >> 4);
    trim_offset[j] = (C * n * (alloc_trim - 5 - lm) * (end - j - 1) * (1 << (lm + BITRES))) >> 6;
    if ((n << lm) == 1) trim_offset[j] -= C << BITRES;
  }
  const int nvec = 11;
  int lo = 1, hi = nvec - 1;
  while (lo <= hi) {
    int mid = (lo + hi) >> 1, done = 0, psum = 0;
    for (int j = end; j-- > start;) {
      int n = eb(j + 1) - eb(j);
      int bitsj = (C * n * kBandAllocation[mid * NB_EBANDS + j]) << lm >> 2;
      if (bitsj > 0) bitsj = imax(0, bitsj + trim_offset[j]);
      bitsj += offsets[j];
      if (bitsj >= thresh[j] || done) { done = 1; psum += imin(bitsj, cap[j]); }
      else if (bitsj >= C << BITRES) psum += C << BITRES;
    }
    if (psum > total) hi = mid - 1; else lo = mid + 1;
  }
  hi = lo;
  lo -= 1;
  for (int j = start; j < end; j++) {
    int n = eb(j + 1) - eb(j);
    int b1 = (C * n * kBandAllocation[lo * NB_EBANDS + j]) << lm >> 2;
    int b2 = hi >= nvec ? cap[j] : (C * n * kBandAllocation[hi * NB_EBANDS + j]) << lm >> 2;
    if (b1 > 0) b1 = imax(0, b1 + trim_offset[j]);
    if (b2 > 0) b2 = imax(0, b2 + trim_offset[j]);
    if (lo > 0) b1 += offsets[j];
    b2 += offsets[j];
    if (offsets[j] > 0) skip_start = j;
    b2 = imax(0, b2 - b1);
    bits1[j] = b1;
    bits2[j] = b2;
  }
  return interp_bits2pulses(start, end, skip_start, bits1, bits2, thresh, cap, total, balance, skip_rsv, intensity,
                            intensity_rsv, dual_stereo, dual_stereo_rsv, pulses, ebits, fine_priority, C, lm, dec, q7, q9);
}

// ---------------------------------------------------------------- quant_bands.rs
// the symbol and its interval [fl, high) for a decoded position fm (the arithmetic of ec_laplace_decode, quant_bands.rs:90-125)
static int laplace_map(uint32_t fm, uint32_t fs, uint32_t decay, uint32_t* fl_out, uint32_t* high_out) {
  int val = 0;
  uint32_t fl = 0;
  if (fm >= fs) {
    val++;
    fl = fs;
    // laplace_get_freq1 (quant_bands.rs:40-48)
    uint32_t rem = 32768u - 32u - fs;
    fs = (decay >= 16384 ? 0 : (uint32_t)(((uint64_t)rem * (16384 - decay)) >> 15)) + 1;
    while (fs > 1 && fm >= fl + 2 * fs) {
      fs *= 2;
      fl += fs;
      fs = (uint32_t)(((uint64_t)(fs - 2) * decay) >> 15);
      fs += 1;
      val++;
    }
    if (fs <= 1) {
      int di = (int)((fm - fl) >> 1);
      val += di;
      fl += 2 * (uint32_t)di;
    }
    if (fm < fl + fs) val = -val; else fl += fs;
  }
  *fl_out = fl;
  *high_out = fl + fs < 32768u ? fl + fs : 32768u;
  return val;
}
static int laplace_decode(EcDec* dec, uint32_t fs, uint32_t decay) {  // quant_bands.rs:90-125
  uint32_t fl, high;
  const int val = laplace_map(dec->decode_bin(15), fs, decay, &fl, &high);
  dec->update(fl, high, 32768u);
  return val;
}
#ifdef ORC_GEN
// corpus writer: make the next laplace_decode return `want` (or the nearest value the model can express)
static void gen_force_laplace(EcDec* dec, uint32_t fs, uint32_t decay, int want) {
  uint32_t fm = 0, best_fm = 0;
  int best_d = 1 << 30;
  for (int guard = 0; guard < 4096 && fm < 32768u; guard++) {
    uint32_t fl, high;
    const int v = laplace_map(fm, fs, decay, &fl, &high);
    const int d = v > want ? v - want : want - v;
    if (d < best_d) { best_d = d; best_fm = fl; }
    if (d == 0 || high <= fm) break;
    fm = high;
  }
  dec->gen->forced = best_fm;
}
#endif

// quant_bands.rs:902-950. Returns nothing; `qi_out` (optional) receives the raw symbols.
static void unquant_coarse_energy(int start, int end, float* oldE, int intra, EcDec* dec, int C, int lm, bool q8) {
  const uint8_t* prob = kEProbModel + (lm * 2 + intra) * 42;
  float prev[2] = {0, 0};
  float coef = intra ? 0.f : kPredCoef[lm];
  float beta = intra ? kBetaIntra : kBetaCoef[lm];
  int budget = (int)dec->storage * 8;
  for (int i = start; i < end; i++) {
    for (int c = 0; c < C; c++) {
      int idx = c * NB_EBANDS + i;
      int tell = dec->tell();
      int qi;
      if (budget - tell >= 15) {
        int pi = 2 * imin(i, 20);
#ifdef ORC_GEN
        if (dec->gen) {   // corpus writer: the band energy follows the stream's envelope (level, tilt, slow loudness, per-band wander)
          GenCtl* g = dec->gen;
          const float target = g->level + g->loud + g->tilt * (float)(10 - i) * 0.1f + g->band_dev[idx];
          const float pred = coef * fmax32(oldE[idx], -9.0f) + prev[c];
          int want = (int)floorf(target - pred + 0.5f);
          want = want < -12 ? -12 : (want > 12 ? 12 : want);
          gen_force_laplace(dec, (uint32_t)prob[pi] << 7, (uint32_t)prob[pi + 1] << 6, want);
        }
#endif
        qi = laplace_decode(dec, (uint32_t)prob[pi] << 7, (uint32_t)prob[pi + 1] << 6);
      } else if (budget - tell >= 2) {
        qi = dec->icdf(kSmallEnergyIcdf, 2);
        qi = (qi >> 1) ^ -(qi & 1);
      } else if (budget - tell >= 1) {
        qi = -dec->bit_logp(1);
      } else qi = -1;
      oldE[idx] = fmax32(oldE[idx], -9.0f);
      float q = (float)qi;
      float tmp = coef * oldE[idx] + prev[c] + q;
      if (q8) {
        tmp = tmp < -28.f ? -28.f : (tmp > 28.f ? 28.f : tmp);  // f32::clamp
        oldE[idx] = tmp;
        prev[c] += q - beta * q;
      } else {  // libopus float build: no clamp, prev = prev + q - beta*q
        oldE[idx] = tmp;
        prev[c] = prev[c] + q - beta * q;
      }
    }
  }
}

static void unquant_fine_energy(int start, int end, float* oldE, const int* fine_quant, EcDec* dec, int C) {  // :1850-1880
  for (int i = start; i < end; i++) {
    if (fine_quant[i] <= 0) continue;
    float scale = (float)(1 << (14 - fine_quant[i])) * (1.0f / 16384.0f);
    for (int c = 0; c < C; c++) {
      int q2 = (int)dec->bits(fine_quant[i]);
      float offset = ((float)q2 + 0.5f) * scale - 0.5f;
      oldE[c * NB_EBANDS + i] += offset;
    }
  }
}

static void unquant_energy_finalise(int start, int end, float* oldE, const int* fine_quant, const int* fine_priority,
                                    int bits_left, EcDec* dec, int C) {  // :1922-1970
  for (int prio = 0; prio < 2; prio++) {
    for (int i = start; i < end; i++) {
      if (bits_left < C) break;
      if (fine_quant[i] >= 8 || fine_priority[i] != prio) continue;
      int fb = imax(fine_quant[i], 0);
      float scale = (float)(1 << (14 - fb - 1)) * (1.0f / 16384.0f);
      for (int c = 0; c < C; c++) {
        int q2 = (int)dec->bits(1);
        float offset = ((float)q2 - 0.5f) * scale;
        oldE[c * NB_EBANDS + i] += offset;
        bits_left--;
      }
    }
  }
}

// celt_decoder.rs:2704-2748
static void tf_decode(int start, int end, int is_transient, int* tf_res, int lm, EcDec* dec) {
  uint32_t budget = dec->storage * 8;
  uint32_t tell = (uint32_t)dec->tell();
  uint32_t logp = is_transient ? 2 : 4;
  int tf_select_rsv = lm > 0 && tell + logp < budget;
  if (tf_select_rsv) budget -= 1;
  int curr = 0, tf_changed = 0;
  for (int i = start; i < end; i++) {
    if (tell + logp <= budget) {
      curr ^= dec->bit_logp(logp);
      tell = (uint32_t)dec->tell();
      tf_changed |= curr;
    }
    tf_res[i] = curr;
    logp = is_transient ? 4 : 5;
  }
  int tf_select = 0;
  if (tf_select_rsv && kTfSelect[lm][4 * is_transient + 0 + tf_changed] != kTfSelect[lm][4 * is_transient + 2 + tf_changed])
    tf_select = dec->bit_logp(1);
  for (int i = start; i < end; i++) tf_res[i] = kTfSelect[lm][4 * is_transient + 2 * tf_select + tf_res[i]];
}

// ---------------------------------------------------------------- cwrs.rs:262-406
static inline uint32_t pvq_u(int n, int k) {
  int row = imin(n, k), col = imax(n, k);
  return g_pvq_u[kPvqRowOff[row] + col];
}
static float cwrsi(int n, int k, uint32_t i, int* y) {
  float yy = 0;
  int o = 0;
  while (n > 2) {
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
      int val = neg ? -(k0 - k) : (k0 - k);
      y[o] = val;
      yy += (float)val * (float)val;
    } else {
      uint32_t p = pvq_u(k, n), q = pvq_u(k + 1, n);
      if (p <= i && i < q) {
        i -= p;
        y[o] = 0;
      } else {
        int neg = i >= q;
        if (neg) i -= q;
        int k0 = k;
        do p = pvq_u(--k, n); while (p > i);
        i -= p;
        int val = neg ? -(k0 - k) : (k0 - k);
        y[o] = val;
        yy += (float)val * (float)val;
      }
    }
    o++;
    n--;
  }
  uint32_t p = 2 * (uint32_t)k + 1;
  int neg = i >= p;
  if (neg) i -= p;
  int k0 = k;
  k = (int)((i + 1) >> 1);
  if (k) i -= 2 * (uint32_t)k - 1;
  int val = neg ? -(k0 - k) : (k0 - k);
  y[o++] = val;
  yy += (float)val * (float)val;
  val = i == 0 ? k : -k;   // cwrs.rs:395-399: last = if index == 0 { k } else { -k }
  y[o] = val;
  yy += (float)val * (float)val;
  return yy;
}
static float decode_pulses(int* y, int n, int k, EcDec* dec) {  // cwrs.rs:427-452
  uint32_t total = pvq_u(n, k) + pvq_u(n, k + 1);
  return cwrsi(n, k, dec->uint(total), y);
}

// ---------------------------------------------------------------- vq.rs
static void exp_rotation1(float* x, int len, int stride, float c, float s) {  // vq.rs:116-150
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
static void exp_rotation(float* x, int len, int dir, int stride, int k, int spread) {  // vq.rs:157-221
  static const int factor_tab[3] = {15, 10, 5};
  if (len == 0 || stride == 0) return;
  if (2 * k >= len || spread == SPREAD_NONE) return;
  int factor = factor_tab[spread - 1];
  float gain = (1.0f * (float)len) / (float)(len + factor * k);
  float theta = 0.5f * gain * gain;
  float c = celt_cos_norm(theta);
  float s = celt_cos_norm(1.0f - theta);
  int stride2 = 0;
  if (len >= 8 * stride) {
    stride2 = 1;
    while ((stride2 * stride2 + stride2) * stride + (stride >> 2) < len) stride2++;
  }
  int ld = len / stride;
  if (ld == 0) return;
  for (int i = 0; i < stride; i++) {
    float* b = x + i * ld;
    if (dir < 0) {
      if (stride2) exp_rotation1(b, ld, stride2, s, c);
      exp_rotation1(b, ld, 1, c, s);
    } else {
      exp_rotation1(b, ld, 1, c, -s);
      if (stride2) exp_rotation1(b, ld, stride2, s, -c);
    }
  }
}
static uint32_t extract_collapse_mask(const int* iy, int n, int b) {  // vq.rs:1164-1192
  if (b <= 1) return 1;
  int n0 = n / b;
  uint32_t mask = 0;
  for (int i = 0; i < b; i++) {
    int acc = 0;
    for (int j = 0; j < n0; j++) acc |= iy[i * n0 + j];
    if (acc) mask |= 1u << i;
  }
  return mask;
}
static void renormalise_vector(float* x, int n, float gain) {  // vq.rs:1073-1088
  float e = 0.f;
  for (int i = 0; i < n; i++) e += x[i] * x[i];
  e = EPSILON + e;
  float scale = celt_rsqrt(e) * gain;
  for (int i = 0; i < n; i++) x[i] *= scale;
}

// ---------------------------------------------------------------- bands.rs
struct BandCtx {
  int band, intensity, spread, tf_change, remaining_bits, disable_inv;
  uint32_t seed;
  EcDec* dec;
  uint32_t quirks;
  const float* xbase;  // trace: start of the frame's X array (C*N)
  int32_t* iy_out;     // trace: pulses at X positions, may be null
};
struct SplitCtx { int inv, imid, iside, delta, itheta, qalloc; };

static inline int frac_mul16(int a, int b) { return (16384 + (int)(int16_t)a * (int)(int16_t)b) >> 15; }
static int bitexact_cos(int x) {  // bands.rs:536-541
  int x2 = (4096 + x * x) >> 13;
  x2 = (32767 - x2) + frac_mul16(x2, -7651 + frac_mul16(x2, 8277 + frac_mul16(-626, x2)));
  return (int16_t)(1 + x2);
}
static int bitexact_log2tan(int isin, int icos) {  // bands.rs:549-560
  int lc = ec_ilog((uint32_t)icos), ls = ec_ilog((uint32_t)isin);
  icos <<= 15 - lc;
  isin <<= 15 - ls;
  return ((ls - lc) << 11) + frac_mul16(isin, frac_mul16(isin, -2597) + 7932) -
         frac_mul16(icos, frac_mul16(icos, -2597) + 7932);
}
static int compute_qn(int n, int b, int offset, int pulse_cap, int stereo) {  // bands.rs:618-640
  static const int exp2_table8[8] = {16384, 17866, 19483, 21247, 23170, 25267, 27554, 30048};
  int n2 = 2 * n - 1;
  if (stereo && n == 2) n2--;
  int qb = celt_sudiv(b + n2 * offset, n2);
  qb = imin(qb, b - pulse_cap - (4 << BITRES));
  qb = imin(qb, 8 << BITRES);
  if (qb < (1 << BITRES >> 1)) return 1;
  int qn = exp2_table8[qb & 7] >> (14 - (qb >> BITRES));
  return (qn + 1) >> 1 << 1;
}
static inline uint32_t mask_from_bits(int b) { return b <= 0 ? 0 : (b >= 32 ? 0xffffffffu : (1u << b) - 1u); }

static void haar1(float* x, int n0, int stride) {  // bands.rs:3797-3822
  if (stride == 0 || n0 < 2) return;
  int half = n0 / 2;
  const float scale = 0.70710678f;
  for (int i = 0; i < stride; i++)
    for (int j = 0; j < half; j++) {
      int i0 = stride * 2 * j + i, i1 = i0 + stride;
      float t1 = scale * x[i0], t2 = scale * x[i1];
      x[i0] = t1 + t2;
      x[i1] = t1 - t2;
    }
}
static const int kOrdery[4][16] = {{1, 0}, {3, 0, 2, 1}, {7, 0, 4, 3, 6, 1, 5, 2},
                                   {15, 0, 8, 7, 12, 3, 11, 4, 14, 1, 9, 6, 13, 2, 10, 5}};
static const int* ordery_for(int stride) { return stride == 2 ? kOrdery[0] : stride == 4 ? kOrdery[1] : stride == 8 ? kOrdery[2] : kOrdery[3]; }
static void deinterleave_hadamard(float* x, int n0, int stride, int hadamard) {  // bands.rs:3708-3750
  int n = n0 * stride;
  if (stride == 0 || n == 0) return;
  float tmp[208];
  if (hadamard) {
    const int* ord = ordery_for(stride);
    for (int i = 0; i < stride; i++)
      for (int j = 0; j < n0; j++) tmp[ord[i] * n0 + j] = x[j * stride + i];
  } else {
    for (int i = 0; i < stride; i++)
      for (int j = 0; j < n0; j++) tmp[i * n0 + j] = x[j * stride + i];
  }
  memcpy(x, tmp, n * sizeof(float));
}
static void interleave_hadamard(float* x, int n0, int stride, int hadamard) {  // bands.rs:3753-3795
  int n = n0 * stride;
  if (stride == 0 || n == 0) return;
  float tmp[208];
  if (hadamard) {
    const int* ord = ordery_for(stride);
    for (int i = 0; i < stride; i++)
      for (int j = 0; j < n0; j++) tmp[j * stride + i] = x[ord[i] * n0 + j];
  } else {
    for (int i = 0; i < stride; i++)
      for (int j = 0; j < n0; j++) tmp[j * stride + i] = x[i * n0 + j];
  }
  memcpy(x, tmp, n * sizeof(float));
}

// bands.rs:274-495 (decode arm only)
static void compute_theta(BandCtx* ctx, SplitCtx* s, int n, int* b, int B, int B0, int lm, int stereo, uint32_t* fill) {
  EcDec* dec = ctx->dec;
  int band = ctx->band;
  int pulse_cap = kLogN400[band] + lm * (1 << BITRES);
  int offset = (pulse_cap >> 1) - (stereo && n == 2 ? 16 : 4);
  int qn = compute_qn(n, *b, offset, pulse_cap, stereo);
  if (stereo && band >= ctx->intensity) qn = 1;
  int itheta = 0, inv = 0, imid, iside, delta;
  int tell0 = (int)dec->tell_frac();
  if (qn != 1) {
    if (stereo && n > 2) {
      int p0 = 3, x0 = qn / 2, ft = p0 * (x0 + 1) + x0;
      int fs = (int)dec->decode((uint32_t)ft);
      int x = fs < (x0 + 1) * p0 ? fs / p0 : x0 + 1 + (fs - (x0 + 1) * p0);
      int fl, fh;
      if (x <= x0) { fl = p0 * x; fh = p0 * (x + 1); }
      else { int base = (x0 + 1) * p0; fl = base + (x - 1 - x0); fh = base + (x - x0); }
      dec->update((uint32_t)fl, (uint32_t)fh, (uint32_t)ft);
      itheta = x;
    } else if (B0 > 1 || stereo) {
      itheta = (int)dec->uint((uint32_t)(qn + 1));
    } else {
      int hq = qn >> 1, ft = (hq + 1) * (hq + 1);
      int fm = (int)dec->decode((uint32_t)ft);
      int fl, fs;
      if (fm < ((hq * (hq + 1)) >> 1)) {
        itheta = ((int)isqrt32((uint32_t)(8 * fm + 1)) - 1) >> 1;
        fl = (itheta * (itheta + 1)) >> 1;
        fs = itheta + 1;
      } else {
        itheta = (2 * (qn + 1) - (int)isqrt32((uint32_t)(8 * (ft - fm - 1) + 1))) >> 1;
        fl = ft - (((qn + 1 - itheta) * (qn + 2 - itheta)) >> 1);
        fs = qn + 1 - itheta;
      }
      dec->update((uint32_t)fl, (uint32_t)(fl + fs), (uint32_t)ft);
    }
    if (qn > 0) itheta = (int)((uint32_t)(itheta * 16384) / (uint32_t)qn);
  } else if (stereo) {
    if (*b > (2 << BITRES) && ctx->remaining_bits > (2 << BITRES)) inv = dec->bit_logp(2);
    else inv = 0;
    if (ctx->disable_inv) inv = 0;
    itheta = 0;
  }
  int qalloc = (int)dec->tell_frac() - tell0;
  *b -= qalloc;
  uint32_t bmask = mask_from_bits(B);
  if (itheta == 0) {
    imid = 32767; iside = 0; *fill &= bmask; delta = -16384;
  } else if (itheta == 16384) {
    imid = 0; iside = 32767;
    *fill &= B <= 0 ? 0 : (B >= 32 ? 0xffffffffu : bmask << B);
    delta = 16384;
  } else {
    imid = bitexact_cos(itheta);
    iside = bitexact_cos(16384 - itheta);
    delta = frac_mul16(imax((n - 1) << 7, 0), bitexact_log2tan(iside, imid));
  }
  s->inv = inv; s->imid = imid; s->iside = iside; s->delta = delta; s->itheta = itheta; s->qalloc = qalloc;
}

static uint32_t alg_unquant(BandCtx* ctx, float* x, int n, int k, int spread, int B, float gain) {  // vq.rs:1013-1037
  int iy[208];
  float ryy = decode_pulses(iy, n, k, ctx->dec);
  if (ctx->iy_out) for (int i = 0; i < n; i++) ctx->iy_out[(x - ctx->xbase) + i] = iy[i];
  float scale = celt_rsqrt(ryy) * gain;  // normalise_residual vq.rs:332-355
  for (int i = 0; i < n; i++) x[i] = scale * (float)iy[i];
  exp_rotation(x, n, -1, B, k, spread);
  return extract_collapse_mask(iy, n, B);
}

// bands.rs:1917-2131
static uint32_t quant_partition(BandCtx* ctx, float* x, int n, int b, int B, float* lowband, int lm, float gain,
                                uint32_t fill) {
  int band = ctx->band;
  const uint8_t* cache = cache_row(band, lm);
  uint32_t cm = 0;
  int B0 = B;
  if (lm != -1 && n > 2 && cache && b > (int)cache[cache[0]] + 12) {
    int half = n >> 1;
    float* y = x + half;
    float* lb2 = lowband ? lowband + half : nullptr;
    lm -= 1;
    if (B == 1) fill = (fill & 1) | (fill << 1);
    B = (B + 1) >> 1;
    SplitCtx s;
    compute_theta(ctx, &s, half, &b, B, B0, lm, 0, &fill);
    float mid = (float)s.imid / 32768.0f, side = (float)s.iside / 32768.0f;
    int delta = s.delta, itheta = s.itheta;
    if (B0 > 1 && (itheta & 0x3fff)) {
      if (itheta > 8192) delta -= delta >> (4 - lm);
      else delta = imin(0, delta + ((half << BITRES) >> (5 - lm)));
    }
    ctx->remaining_bits -= s.qalloc;
    int mbits = (b - delta) / 2;
    mbits = imax(0, imin(b, mbits));
    int sbits = b - mbits;
    int rebalance = ctx->remaining_bits;
    if (mbits >= sbits) {
      cm = quant_partition(ctx, x, half, mbits, B, lowband, lm, gain * mid, fill);
      rebalance = mbits - (rebalance - ctx->remaining_bits);
      if (rebalance > 3 << BITRES && itheta != 0) sbits += rebalance - (3 << BITRES);
      cm |= quant_partition(ctx, y, half, sbits, B, lb2, lm, gain * side, fill >> B) << (B0 >> 1);
    } else {
      cm = quant_partition(ctx, y, half, sbits, B, lb2, lm, gain * side, fill >> B) << (B0 >> 1);
      rebalance = sbits - (rebalance - ctx->remaining_bits);
      if (rebalance > 3 << BITRES && itheta != 16384) mbits += rebalance - (3 << BITRES);
      cm |= quant_partition(ctx, x, half, mbits, B, lowband, lm, gain * mid, fill);
    }
    return cm;
  }
  int q = bits2pulses(band, lm, b);
  int curr_bits = pulses2bits(band, lm, q);
  ctx->remaining_bits -= curr_bits;
  while (ctx->remaining_bits < 0 && q > 0) {
    ctx->remaining_bits += curr_bits;
    q--;
    curr_bits = pulses2bits(band, lm, q);
    ctx->remaining_bits -= curr_bits;
  }
  if (q != 0) {
    cm = alg_unquant(ctx, x, n, get_pulses(q), ctx->spread, imax(B, 1), gain);
  } else {
    uint32_t cm_mask = mask_from_bits(B);
    fill &= cm_mask;
    if (!fill) {
      for (int j = 0; j < n; j++) x[j] = 0.f;
    } else if (lowband) {
      const float tmp = 1.0f / 256.0f;
      for (int j = 0; j < n; j++) {
        ctx->seed = lcg_rand(ctx->seed);
        x[j] = lowband[j] + ((ctx->seed & 0x8000) ? tmp : -tmp);
      }
      cm = fill;
      renormalise_vector(x, n, gain);
    } else {
      for (int j = 0; j < n; j++) {
        ctx->seed = lcg_rand(ctx->seed);
        x[j] = (float)((int32_t)ctx->seed >> 20);
      }
      cm = cm_mask;
      renormalise_vector(x, n, gain);
    }
  }
  return cm;
}

static uint32_t quant_band_n1(BandCtx* ctx, float* x, float* y, float* lowband_out) {  // bands.rs:642-704
  float* ch[2] = {x, y};
  for (int c = 0; c < (y ? 2 : 1); c++) {
    int sign = 0;
    if (ctx->remaining_bits >= 1 << BITRES) {
      sign = (int)ctx->dec->bits(1);
      ctx->remaining_bits -= 1 << BITRES;
    }
    ch[c][0] = sign ? -1.0f : 1.0f;
  }
  if (lowband_out) lowband_out[0] = x[0];
  return 1;
}

// bands.rs:2133-2291
static uint32_t quant_band(BandCtx* ctx, float* x, int n, int b, int B, float* lowband, int lm, float* lowband_out,
                           float gain, float* lowband_scratch, uint32_t fill) {
  int n0 = n, n_b = n, B0 = B, time_divide = 0, recombine = 0;
  int long_blocks = B0 == 1;
  int tf_change = ctx->tf_change;
  n_b = n_b / B;
  if (n == 1) return quant_band_n1(ctx, x, nullptr, lowband_out);
  if (tf_change > 0) recombine = tf_change;
  if (lowband && lowband_scratch && (recombine || ((n_b & 1) == 0 && tf_change < 0) || B0 > 1)) {
    memcpy(lowband_scratch, lowband, n * sizeof(float));
    lowband = lowband_scratch;
  }
  static const uint8_t bit_interleave[16] = {0, 1, 1, 1, 2, 3, 3, 3, 2, 3, 3, 3, 2, 3, 3, 3};
  for (int k = 0; k < recombine; k++) {
    if (lowband) haar1(lowband, n >> k, 1 << k);
    fill = bit_interleave[fill & 0xF] | (uint32_t)bit_interleave[(fill >> 4) & 0xF] << 2;
  }
  B >>= recombine;
  n_b <<= recombine;
  while ((n_b & 1) == 0 && tf_change < 0) {
    if (lowband) haar1(lowband, n_b, B);
    fill |= fill << B;
    B <<= 1;
    n_b >>= 1;
    time_divide++;
    tf_change++;
  }
  B0 = B;
  int n_b0 = n_b;
  if (B0 > 1 && lowband) deinterleave_hadamard(lowband, n_b >> recombine, B0 << recombine, long_blocks);
  uint32_t cm = quant_partition(ctx, x, n, b, B, lowband, lm, gain, fill);
  // resynth
  if (B0 > 1) interleave_hadamard(x, n_b >> recombine, B0 << recombine, long_blocks);
  n_b = n_b0;
  B = B0;
  for (int k = 0; k < time_divide; k++) {
    B >>= 1;
    n_b <<= 1;
    if (B > 0) cm |= cm >> B;
    haar1(x, n_b, imax(B, 1));
  }
  static const uint8_t bit_deinterleave[16] = {0x00, 0x03, 0x0C, 0x0F, 0x30, 0x33, 0x3C, 0x3F,
                                               0xC0, 0xC3, 0xCC, 0xCF, 0xF0, 0xF3, 0xFC, 0xFF};
  for (int k = 0; k < recombine; k++) {
    cm = bit_deinterleave[cm & 0xF];
    haar1(x, n0 >> k, 1 << k);
  }
  B <<= recombine;
  if (lowband_out) {
    float scale = celt_sqrt((float)n0);
    for (int j = 0; j < n0; j++) lowband_out[j] = scale * x[j];
  }
  cm &= mask_from_bits(B);
  return cm;
}

static void stereo_merge(float* x, float* y, float mid, int n) {  // bands.rs:3477-3505
  float xp = 0, side = 0;
  for (int j = 0; j < n; j++) { float a = y[j]; xp += a * x[j]; side += a * y[j]; }
  xp *= mid;
  float mid2 = mid * mid;
  float el = mid2 + side - 2.0f * xp;
  float er = mid2 + side + 2.0f * xp;
  if (er < 6e-4f || el < 6e-4f) { memcpy(y, x, n * sizeof(float)); return; }
  float lgain = celt_rsqrt(el), rgain = celt_rsqrt(er);
  for (int j = 0; j < n; j++) {
    float l = mid * x[j], r = y[j];
    x[j] = lgain * (l - r);
    y[j] = rgain * (l + r);
  }
}

// bands.rs:2294-2514
static uint32_t quant_band_stereo(BandCtx* ctx, float* x, float* y, int n, int b, int B, float* lowband, int lm,
                                  float* lowband_out, float* lowband_scratch, uint32_t fill) {
  if (n == 1) return quant_band_n1(ctx, x, y, lowband_out);
  uint32_t orig_fill = fill;
  SplitCtx s;
  compute_theta(ctx, &s, n, &b, B, B, lm, 1, &fill);
  float mid = (float)s.imid / 32768.0f, side = (float)s.iside / 32768.0f;
  int itheta = s.itheta;
  uint32_t cm;
  if (n == 2) {
    int mbits = b, sbits = 0;
    if (itheta != 0 && itheta != 16384) sbits = 1 << BITRES;
    mbits -= sbits;
    int c = itheta > 8192;
    ctx->remaining_bits -= s.qalloc + sbits;
    float* x2 = c ? y : x;
    float* y2 = c ? x : y;
    int sign = 0;
    if (sbits) sign = (int)ctx->dec->bits(1);
    int sv = 1 - 2 * sign;
    cm = quant_band(ctx, x2, n, mbits, B, lowband, lm, lowband_out, 1.0f, lowband_scratch, orig_fill);
    y2[0] = -(float)sv * x2[1];
    y2[1] = (float)sv * x2[0];
    x[0] *= mid; x[1] *= mid; y[0] *= side; y[1] *= side;
    float t = x[0]; x[0] = t - y[0]; y[0] += t;
    t = x[1]; x[1] = t - y[1]; y[1] += t;
  } else {
    int mbits = (b - s.delta) / 2;
    mbits = imax(0, imin(b, mbits));
    int sbits = b - mbits;
    ctx->remaining_bits -= s.qalloc;
    int rebalance = ctx->remaining_bits;
    if (mbits >= sbits) {
      cm = quant_band(ctx, x, n, mbits, B, lowband, lm, lowband_out, 1.0f, lowband_scratch, fill);
      rebalance = mbits - (rebalance - ctx->remaining_bits);
      if (rebalance > 3 << BITRES && itheta != 0) sbits += rebalance - (3 << BITRES);
      cm |= quant_band(ctx, y, n, sbits, B, nullptr, lm, nullptr, side, nullptr, fill >> B);
    } else {
      cm = quant_band(ctx, y, n, sbits, B, nullptr, lm, nullptr, side, nullptr, fill >> B);
      rebalance = sbits - (rebalance - ctx->remaining_bits);
      if (rebalance > 3 << BITRES && itheta != 16384) mbits += rebalance - (3 << BITRES);
      cm |= quant_band(ctx, x, n, mbits, B, lowband, lm, lowband_out, 1.0f, lowband_scratch, fill);
    }
  }
  if (n != 2) stereo_merge(x, y, mid, n);
  if (s.inv) for (int j = 0; j < n; j++) y[j] = -y[j];
  return cm;
}

// lowband_in_out_mut (bands.rs:2529-2573). With Q6 the release-build semantics are kept:
// windows that overlap (hybrid band start+1) or overrun the norm buffer yield (None, None).
static void lowband_in_out(float* buf, int buf_len, int in_off, int out_off, int len, bool q6, float** in, float** out) {
  *in = *out = nullptr;
  if (!q6) {
    if (in_off >= 0) *in = buf + in_off;
    if (out_off >= 0) *out = buf + out_off;
    return;
  }
  if (in_off < 0 && out_off < 0) return;
  if (out_off < 0) { if (in_off + len <= buf_len) *in = buf + in_off; return; }
  if (in_off < 0) { if (out_off + len <= buf_len) *out = buf + out_off; return; }
  if (in_off + len > buf_len || out_off + len > buf_len) return;
  if (in_off + len <= out_off || out_off + len <= in_off) { *in = buf + in_off; *out = buf + out_off; }
}

// bands.rs:2575-3082 (decode arm)
static void quant_all_bands(CeltDecoder* st, int start, int end, float* X, float* Y, uint8_t* collapse_masks,
                            const int* pulses, int short_blocks, int spread, int dual_stereo, int intensity,
                            const int* tf_res, int total_bits, int balance, EcDec* dec, int lm, int coded_bands,
                            uint32_t* seed, int32_t* iy_out) {
  if (start >= end) return;
  const uint32_t quirks = st->quirks;
  int C = Y ? 2 : 1;
  int M = 1 << lm;
  int B = short_blocks ? M : 1;
  int norm_offset = M * eb(start);
  int last_band_start = M * eb(NB_EBANDS - 1);
  int norm_len = imax(last_band_start - norm_offset, 0);
  if ((int)st->norm_storage.size() < C * norm_len) st->norm_storage.resize(C * norm_len, 0.0f);
  float* norm = st->norm_storage.data();
  float* norm2 = C == 2 ? norm + norm_len : nullptr;
  float lowband_scratch_buf[22 * 8];
  float* lowband_scratch = lowband_scratch_buf;
  BandCtx ctx;
  ctx.intensity = intensity; ctx.spread = spread; ctx.seed = *seed; ctx.dec = dec; ctx.disable_inv = st->disable_inv;
  ctx.quirks = quirks; ctx.xbase = X; ctx.iy_out = iy_out; ctx.tf_change = 0; ctx.remaining_bits = total_bits; ctx.band = start;
  int lowband_offset = -1;
  int update_lowband = 1;
  for (int i = start; i < end; i++) {
    ctx.band = i;
    int last = i + 1 == end;
    int band_start = M * eb(i), band_end = M * eb(i + 1);
    int n = band_end - band_start;
    int tell = (int)dec->tell_frac();
    if (i != start) balance -= tell;
    int remaining_bits = total_bits - tell - 1;
    ctx.remaining_bits = remaining_bits;
    int b = 0;
    if (i < coded_bands) {
      int curr_balance = celt_sudiv(balance, imin(3, coded_bands - i));
      b = imax(0, imin(16383, imin(remaining_bits + 1, pulses[i] + curr_balance)));
    }
    if ((band_start >= norm_offset + n || i == start + 1) && (update_lowband || lowband_offset < 0)) lowband_offset = i;
    if (i == start + 1) {  // special_hybrid_folding bands.rs:3517-3555
      int n1 = M * (eb(start + 1) - eb(start)), n2 = M * (eb(start + 2) - eb(start + 1));
      if (n2 > n1) {
        memmove(norm + n1, norm + 2 * n1 - n2, (n2 - n1) * sizeof(float));
        if (dual_stereo && norm2) memmove(norm2 + n1, norm2 + 2 * n1 - n2, (n2 - n1) * sizeof(float));
      }
    }
    int tf_change = tf_res[i];
    ctx.tf_change = tf_change;
    if (last) lowband_scratch = nullptr;
    float* x_band = X + band_start;
    float* y_band = Y ? Y + band_start : nullptr;
    int effective_lowband = -1;
    uint32_t x_cm = 0, y_cm = 0;
    if (lowband_offset >= 0 && (spread != SPREAD_AGGRESSIVE || B > 1 || tf_change < 0)) {
      effective_lowband = imax(0, M * eb(lowband_offset) - norm_offset - n);
      int fold_start = lowband_offset;
      if (quirks & QUIRK_Q5_FOLD_START) {
        int threshold = effective_lowband + norm_offset + n;
        while (fold_start > 0) { fold_start--; if (M * eb(fold_start) <= threshold) break; }
      } else {
        while (M * eb(--fold_start) > effective_lowband + norm_offset) {}
      }
      int fold_end = lowband_offset - 1;
      while (++fold_end < i && M * eb(fold_end) < effective_lowband + norm_offset + n) {}
      for (int f = fold_start; f < fold_end; f++) { x_cm |= collapse_masks[f * C]; y_cm |= collapse_masks[f * C + C - 1]; }
    }
    if (effective_lowband < 0) x_cm = y_cm = mask_from_bits(B);
    if (dual_stereo && i == intensity) {
      dual_stereo = 0;
      // bands.rs:2846-2853: mousiki averages the whole norm buffer (libopus stops at the current band).
      if (norm2) for (int j = 0; j < norm_len; j++) norm[j] = (norm[j] + norm2[j]) * 0.5f;
    }
    int out_off = last ? -1 : band_start - norm_offset;
    bool q6 = quirks & QUIRK_Q6_HYBRID_LOWBAND;
    float *lb_in, *lb_out;
    if (dual_stereo) {
      lowband_in_out(norm, norm_len, effective_lowband, out_off, n, q6, &lb_in, &lb_out);
      x_cm = quant_band(&ctx, x_band, n, b / 2, B, lb_in, lm, lb_out, 1.0f, lowband_scratch, x_cm);
      lowband_in_out(norm2, norm_len, effective_lowband, out_off, n, q6, &lb_in, &lb_out);
      y_cm = quant_band(&ctx, y_band, n, b / 2, B, lb_in, lm, lb_out, 1.0f, lowband_scratch, y_cm);
    } else if (Y) {
      lowband_in_out(norm, norm_len, effective_lowband, out_off, n, q6, &lb_in, &lb_out);
      x_cm = quant_band_stereo(&ctx, x_band, y_band, n, b, B, lb_in, lm, lb_out, lowband_scratch, x_cm | y_cm);
      y_cm = x_cm;
    } else {
      lowband_in_out(norm, norm_len, effective_lowband, out_off, n, q6, &lb_in, &lb_out);
      x_cm = quant_band(&ctx, x_band, n, b, B, lb_in, lm, lb_out, 1.0f, lowband_scratch, x_cm | y_cm);
      y_cm = x_cm;
    }
    collapse_masks[i * C] = (uint8_t)x_cm;
    collapse_masks[i * C + C - 1] = (uint8_t)y_cm;
    balance += pulses[i] + tell;
    update_lowband = b > (n << BITRES);
  }
  *seed = ctx.seed;
}

// bands.rs:3220-3349
static void anti_collapse(float* X_, const uint8_t* collapse_masks, int lm, int C, int size, int start, int end,
                          const float* logE, const float* prev1logE, const float* prev2logE, const int* pulses, uint32_t seed) {
  for (int i = start; i < end; i++) {
    int N0 = eb(i + 1) - eb(i);
    int depth = (int)((uint32_t)(1 + pulses[i]) / (uint32_t)N0) >> lm;
    float thresh = 0.5f * celt_exp2(-0.125f * (float)depth);
    float sqrt_1 = celt_rsqrt((float)(N0 << lm));
    for (int c = 0; c < C; c++) {
      float prev1 = prev1logE[c * NB_EBANDS + i], prev2 = prev2logE[c * NB_EBANDS + i];
      if (C == 1) {
        prev1 = fmax32(prev1, prev1logE[NB_EBANDS + i]);
        prev2 = fmax32(prev2, prev2logE[NB_EBANDS + i]);
      }
      float Ediff = logE[c * NB_EBANDS + i] - fmin32(prev1, prev2);
      if (Ediff < 0.f) Ediff = 0.f;
      float r = 2.0f * celt_exp2(-Ediff);
      if (lm == 3) r *= 1.41421356f;
      r = fmin32(r, thresh);
      r *= sqrt_1;
      float* X = X_ + c * size + (eb(i) << lm);
      int renorm = 0;
      for (int k = 0; k < 1 << lm; k++) {
        if (!(collapse_masks[i * C + c] & (1u << k))) {
          for (int j = 0; j < N0; j++) {
            seed = lcg_rand(seed);
            X[(j << lm) + k] = (seed & 0x8000) ? r : -r;
          }
          renorm = 1;
        }
      }
      if (renorm) renormalise_vector(X, N0 << lm, 1.0f);
    }
  }
}

// bands.rs:4172-4234
static void denormalise_bands(const float* X, float* freq, const float* bandLogE, int start, int end, int M,
                              int downsample, int silence) {
  int N = M * SHORT_MDCT;
  int bound = M * eb(end);
  if (downsample != 1) bound = imin(bound, N / downsample);
  if (silence) { bound = 0; start = end = 0; }
  int j = 0;
  for (; j < M * eb(start); j++) freq[j] = 0.f;
  for (int i = start; i < end; i++) {
    int band_end = M * eb(i + 1);
    float g = celt_exp2(fmin32(bandLogE[i] + kEMeans[i], 32.0f));
    for (; j < band_end; j++) freq[j] = X[j] * g;
  }
  for (int k = bound; k < N; k++) freq[k] = 0.f;
}

// ---------------------------------------------------------------- kiss_fft.rs
struct cpx { float r, i; };
static inline cpx c_mul(cpx a, cpx b) {  // kiss_fft.rs:537-542 (explicit fmaf)
  cpx o; o.r = fmaf(a.r, b.r, -(a.i * b.i)); o.i = fmaf(a.r, b.i, a.i * b.r); return o;
}
static inline cpx c_add(cpx a, cpx b) { return {a.r + b.r, a.i + b.i}; }
static inline cpx c_sub(cpx a, cpx b) { return {a.r - b.r, a.i - b.i}; }
static inline cpx tw(int idx) { return {g_fft_tw[2 * idx], g_fft_tw[2 * idx + 1]}; }

static void kf_bfly2(cpx* f, int m, int n) {  // kiss_fft.rs:900-966 (m==4 path in static modes)
  if (m == 1) {
    for (int i = 0; i < n; i++) { cpx t = f[2 * i + 1]; f[2 * i + 1] = c_sub(f[2 * i], t); f[2 * i] = c_add(f[2 * i], t); }
    return;
  }
  const float t8 = 0.70710678118654752440f;
  for (int i = 0; i < n; i++) {
    cpx* p = f + i * 8;
    cpx t = p[4];
    p[4] = c_sub(p[0], t); p[0] = c_add(p[0], t);
    t.r = (p[5].r + p[5].i) * t8; t.i = (p[5].i - p[5].r) * t8;
    p[5] = c_sub(p[1], t); p[1] = c_add(p[1], t);
    t.r = p[6].i; t.i = -p[6].r;
    p[6] = c_sub(p[2], t); p[2] = c_add(p[2], t);
    t.r = (p[7].i - p[7].r) * t8; t.i = -(p[7].i + p[7].r) * t8;
    p[7] = c_sub(p[3], t); p[3] = c_add(p[3], t);
  }
}
static void kf_bfly3(cpx* f, int fstride, int m, int n, int mm) {  // kiss_fft.rs:968-1020
  int m2 = 2 * m;
  cpx epi3 = tw(fstride * m);
  for (int i = 0; i < n; i++) {
    cpx* p = f + i * mm;
    int t1 = 0, t2 = 0;
    for (int k = 0; k < m; k++) {
      cpx s1 = c_mul(p[m + k], tw(t1)), s2 = c_mul(p[m2 + k], tw(t2));
      cpx s3 = c_add(s1, s2), s0 = c_sub(s1, s2);
      t1 += fstride; t2 += fstride * 2;
      cpx fm = {p[k].r - 0.5f * s3.r, p[k].i - 0.5f * s3.i};
      s0.r *= epi3.i; s0.i *= epi3.i;
      cpx f0 = c_add(p[k], s3);
      p[m2 + k] = {fm.r + s0.i, fm.i - s0.r};
      p[m + k] = {fm.r - s0.i, fm.i + s0.r};
      p[k] = f0;
    }
  }
}
static void kf_bfly4(cpx* f, int fstride, int m, int n, int mm) {  // kiss_fft.rs:1022-1165
  if (m == 1) {
    for (int i = 0; i < n; i++) {
      cpx* p = f + i * mm;
      cpx s0 = c_sub(p[0], p[2]), s1 = c_add(p[1], p[3]), s1b = c_sub(p[1], p[3]);
      cpx f0 = c_add(p[0], p[2]);
      p[2] = c_sub(f0, s1);
      f0 = c_add(f0, s1);
      p[1] = {s0.r + s1b.i, s0.i - s1b.r};
      p[3] = {s0.r - s1b.i, s0.i + s1b.r};
      p[0] = f0;
    }
    return;
  }
  int m2 = 2 * m, m3 = 3 * m;
  for (int i = 0; i < n; i++) {
    cpx* p = f + i * mm;
    int t1 = 0, t2 = 0, t3 = 0;
    for (int j = 0; j < m; j++) {
      cpx s0 = c_mul(p[j + m], tw(t1)), s1 = c_mul(p[j + m2], tw(t2)), s2 = c_mul(p[j + m3], tw(t3));
      t1 += fstride; t2 += fstride * 2; t3 += fstride * 3;
      cpx s5 = c_sub(p[j], s1);
      cpx f0 = c_add(p[j], s1);
      cpx s3 = c_add(s0, s2), s4 = c_sub(s0, s2);
      p[j + m2] = c_sub(f0, s3);
      f0 = c_add(f0, s3);
      p[j + m] = {s5.r + s4.i, s5.i - s4.r};
      p[j + m3] = {s5.r - s4.i, s5.i + s4.r};
      p[j] = f0;
    }
  }
}
static void kf_bfly5(cpx* f, int fstride, int m, int n, int mm) {  // kiss_fft.rs:1167-1340
  cpx ya = tw(fstride * m), yb = tw(fstride * 2 * m);
  for (int i = 0; i < n; i++) {
    cpx* p = f + i * mm;
    for (int u = 0; u < m; u++) {
      cpx s0 = p[u];
      cpx s1 = c_mul(p[m + u], tw(u * fstride)), s2 = c_mul(p[2 * m + u], tw(2 * u * fstride));
      cpx s3 = c_mul(p[3 * m + u], tw(3 * u * fstride)), s4 = c_mul(p[4 * m + u], tw(4 * u * fstride));
      cpx s7 = c_add(s1, s4), s10 = c_sub(s1, s4), s8 = c_add(s2, s3), s9 = c_sub(s2, s3);
      cpx f0 = c_add(s0, c_add(s7, s8));
      cpx s5 = {s0.r + fmaf(s7.r, ya.r, s8.r * yb.r), s0.i + fmaf(s7.i, ya.r, s8.i * yb.r)};
      cpx s6 = {fmaf(s10.i, ya.i, s9.i * yb.i), -fmaf(s10.r, ya.i, s9.r * yb.i)};
      p[m + u] = c_sub(s5, s6);
      p[4 * m + u] = c_add(s5, s6);
      cpx s11 = {s0.r + fmaf(s7.r, yb.r, s8.r * ya.r), s0.i + fmaf(s7.i, yb.r, s8.i * ya.r)};
      cpx s12 = {fmaf(s9.i, ya.i, (-s10.i) * yb.i), fmaf(s10.r, yb.i, (-s9.r) * ya.i)};
      p[2 * m + u] = c_add(s11, s12);
      p[3 * m + u] = c_sub(s11, s12);
      p[u] = f0;
    }
  }
}
static const int kFftFactors[4][10] = {{5, 96, 3, 32, 4, 8, 2, 4, 4, 1}, {5, 48, 3, 16, 4, 4, 4, 1}, {5, 24, 3, 8, 2, 4, 4, 1}, {5, 12, 3, 4, 4, 1}};
static const int16_t* const kBitrev[4] = {kFftBitrev480, kFftBitrev240, kFftBitrev120, kFftBitrev60};

static void fft_impl(cpx* fout, int shift) {  // kiss_fft.rs:748-795
  const int* fac = kFftFactors[shift];
  int fstride[8];
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
    int m2 = i != 0 ? fac[2 * i - 1] : 1;
    switch (fac[2 * i]) {
      case 2: kf_bfly2(fout, m, fstride[i]); break;
      case 3: kf_bfly3(fout, fstride[i] << shift, m, fstride[i], m2); break;
      case 4: kf_bfly4(fout, fstride[i] << shift, m, fstride[i], m2); break;
      case 5: kf_bfly5(fout, fstride[i] << shift, m, fstride[i], m2); break;
    }
    m = m2;
  }
}

// mdct.rs:197-290, 362-388. `out_len` = length of the slice the reference passes (Q2 zero-fills all of it).
static void clt_mdct_backward(const float* in, float* out, int out_len, int shift, bool q2) {
  static const int tw_off[5] = {0, 960, 1440, 1680, 1800};
  int n = 1920 >> shift, n2 = n >> 1, n4 = n >> 2;
  if (q2) for (int i = 0; i < out_len; i++) out[i] = 0.f;
  const float* cosp = g_mdct_tw + tw_off[shift];
  const float* sinp = cosp + n4;
  cpx pre[480], f[480];
  for (int i = 0; i < n4; i++) {
    float x1 = in[2 * i], x2 = in[n2 - 1 - 2 * i];
    float t0 = cosp[i], t1 = sinp[i];
    pre[i].r = x2 * t0 + x1 * t1;
    pre[i].i = x1 * t0 - x2 * t1;
  }
  const int16_t* br = kBitrev[shift];
  for (int i = 0; i < n4; i++) { f[br[i]].r = pre[i].r; f[br[i]].i = -pre[i].i; }
  fft_impl(f, shift);
  for (int i = 0; i < n4; i++) f[i].i = -f[i].i;
  float temp[960];
  int pairs = (n4 + 1) >> 1;
  for (int i = 0; i < pairs; i++) {
    float yr_f = f[i].r * cosp[i] + f[i].i * sinp[i];
    float yi_f = f[i].r * sinp[i] - f[i].i * cosp[i];
    int bi = n4 - i - 1;
    float yr_b, yi_b;
    if (bi == i) { yr_b = yr_f; yi_b = yi_f; }
    else {
      yr_b = f[bi].r * cosp[bi] + f[bi].i * sinp[bi];
      yi_b = f[bi].r * sinp[bi] - f[bi].i * cosp[bi];
    }
    temp[2 * i] = yr_f;
    temp[2 * i + 1] = yi_b;
    temp[n2 - 2 * (i + 1)] = yr_b;
    temp[n2 - 2 * (i + 1) + 1] = yi_f;
  }
  int half = OVERLAP >> 1;
  for (int i = 0; i < n2; i++) out[half + i] = temp[i];
  for (int i = 0; i < half; i++) {
    float w1 = g_window[i], w2 = g_window[OVERLAP - 1 - i];
    int yp1 = i, xp1 = OVERLAP - 1 - i;
    float x1 = out[xp1], x2 = out[yp1];
    out[yp1] = x2 * w2 - x1 * w1;
    out[xp1] = x2 * w1 + x1 * w2;
  }
}

// ---------------------------------------------------------------- celt.rs:133-174, 332-443
static void comb_filter_const(float* ch, int os, int n, int T, float g10, float g11, float g12) {
  float x4 = ch[os - T - 2], x3 = ch[os - T - 1], x2 = ch[os - T], x1 = ch[os - T + 1];
  for (int i = 0; i < n; i++) {
    float cur = ch[os + i];
    float x0 = ch[os + i - T + 2];
    float acc = fmaf(g10, x2, cur);
    acc = fmaf(g11, x1 + x3, acc);
    acc = fmaf(g12, x0 + x4, acc);
    ch[os + i] = acc;
    x4 = x3; x3 = x2; x2 = x1; x1 = x0;
  }
}
static void comb_filter_in_place(float* ch, int os, int n, int T0, int T1, float g0, float g1, int tapset0, int tapset1) {
  if (n == 0) return;
  if (g0 == 0.f && g1 == 0.f) return;
  T0 = imax(T0, COMBFILTER_MINPERIOD);
  T1 = imax(T1, COMBFILTER_MINPERIOD);
  float g00 = g0 * kTapsetGains[tapset0][0], g01 = g0 * kTapsetGains[tapset0][1], g02 = g0 * kTapsetGains[tapset0][2];
  float g10 = g1 * kTapsetGains[tapset1][0], g11 = g1 * kTapsetGains[tapset1][1], g12 = g1 * kTapsetGains[tapset1][2];
  float x1 = ch[os - T1 + 1], x2 = ch[os - T1], x3 = ch[os - T1 - 1], x4 = ch[os - T1 - 2];
  int overlap = imin(OVERLAP, n);
  if (g0 == g1 && T0 == T1 && tapset0 == tapset1) overlap = 0;
  for (int i = 0; i < overlap; i++) {
    float x0 = ch[os + i - T1 + 2];
    float f = g_window[i] * g_window[i];
    float omf = 1.0f - f;
    float cur = ch[os + i];
    float p0 = ch[os + i - T0], p1 = ch[os + i - T0 + 1], pm1 = ch[os + i - T0 - 1];
    float p2 = ch[os + i - T0 + 2], pm2 = ch[os + i - T0 - 2];
    float acc = fmaf(omf * g00, p0, cur);
    acc = fmaf(omf * g01, p1 + pm1, acc);
    acc = fmaf(omf * g02, p2 + pm2, acc);
    acc = fmaf(f * g10, x2, acc);
    acc = fmaf(f * g11, x1 + x3, acc);
    acc = fmaf(f * g12, x0 + x4, acc);
    ch[os + i] = acc;
    x4 = x3; x3 = x2; x2 = x1; x1 = x0;
  }
  if (g1 == 0.f) return;
  if (overlap < n) comb_filter_const(ch, os + overlap, n - overlap, T1, g10, g11, g12);
}

// celt_decoder.rs:2157-2299
static void deemphasis(float* const in[2], float* pcm, int N, int CC, int downsample, float* mem, int accum) {
  const float coef0 = kPreemph;
  const float inv = 1.0f / 32768.0f;
  if (downsample == 1 && CC == 2 && !accum) {
    float m0 = mem[0], m1 = mem[1];
    for (int j = 0; j < N; j++) {
      float t0 = in[0][j] + m0 + VERY_SMALL;
      float t1 = in[1][j] + m1 + VERY_SMALL;
      m0 = coef0 * t0;
      m1 = coef0 * t1;
      pcm[2 * j] = t0 * inv;
      pcm[2 * j + 1] = t1 * inv;
    }
    mem[0] = m0; mem[1] = m1;
    return;
  }
  std::vector<float> scratch(N);
  int Nd = N / downsample;
  for (int c = 0; c < CC; c++) {
    const float* x = in[c];
    float m = mem[c];
    int apply_ds = 0;
    if (downsample > 1) {
      for (int j = 0; j < N; j++) { float t = x[j] + m + VERY_SMALL; m = coef0 * t; scratch[j] = t; }
      apply_ds = 1;
    } else if (accum) {
      for (int j = 0; j < N; j++) { float t = x[j] + m + VERY_SMALL; m = coef0 * t; pcm[j * CC + c] = pcm[j * CC + c] + t * inv; }
    } else {
      for (int j = 0; j < N; j++) { float t = x[j] + m + VERY_SMALL; m = coef0 * t; pcm[j * CC + c] = t * inv; }
    }
    mem[c] = m;
    if (apply_ds) {
      for (int j = 0; j < Nd; j++) {
        float v = scratch[j * downsample] * inv;
        if (accum) pcm[j * CC + c] = pcm[j * CC + c] + v; else pcm[j * CC + c] = v;
      }
    }
  }
}

// celt_synthesis (celt_decoder.rs:573-800): denormalise + inverse MDCT of every channel into out_syn[c][0 .. N + overlap)
static void celt_synthesis(const float* X, float* const out_syn[2], const float* oldE, int start, int eff_end, int C, int CC, int is_transient,
                           int lm, int downsample, int silence, uint32_t quirks) {
  const int M = 1 << lm, N = M * SHORT_MDCT;
  int nbk, nb, shift;
  if (is_transient) { nbk = M; nb = SHORT_MDCT; shift = MAX_LM; }
  else { nbk = 1; nb = SHORT_MDCT << lm; shift = MAX_LM - lm; }
  const bool q2 = quirks & QUIRK_Q2_MDCT_ZERO_FILL;
  std::vector<float> freq(N), freq2(N), temp(N);
  auto imdct = [&](const float* fr, float* out) {
    for (int b = 0; b < nbk; b++) {
      for (int k = 0; k < nb; k++) temp[k] = fr[b + k * nbk];
      clt_mdct_backward(temp.data(), out + b * nb, N + OVERLAP - b * nb, shift, q2);
    }
  };
  if (CC == 2 && C == 1) {
    denormalise_bands(X, freq.data(), oldE, start, eff_end, M, downsample, silence);
    imdct(freq.data(), out_syn[0]);
    imdct(freq.data(), out_syn[1]);
  } else if (CC == 1 && C == 2) {
    denormalise_bands(X, freq.data(), oldE, start, eff_end, M, downsample, silence);
    denormalise_bands(X + N, freq2.data(), oldE + NB_EBANDS, start, eff_end, M, downsample, silence);
    for (int i = 0; i < N; i++) freq[i] = 0.5f * (freq[i] + freq2[i]);
    imdct(freq.data(), out_syn[0]);
  } else {
    for (int c = 0; c < CC; c++) {
      denormalise_bands(X + c * N, freq.data(), oldE + c * NB_EBANDS, start, eff_end, M, downsample, silence);
      imdct(freq.data(), out_syn[c]);
    }
  }
  for (int c = 0; c < CC; c++)
    for (int i = 0; i < N; i++) out_syn[c][i] = out_syn[c][i] < -SIG_SAT ? -SIG_SAT : (out_syn[c][i] > SIG_SAT ? SIG_SAT : out_syn[c][i]);
}

// ---------------------------------------------------------------- packet-loss concealment
// pitch.rs:118-187 (find_best_pitch)
static void find_best_pitch(const float* xcorr, const float* y, int len, int max_pitch, int* best_pitch) {
  float syy = 1.0f;
  for (int j = 0; j < len; j++) syy += y[j] * y[j];
  float best_num0 = -1.0f, best_den0 = 0.0f, best_num1 = -1.0f, best_den1 = 0.0f;
  best_pitch[0] = 0; best_pitch[1] = max_pitch > 1 ? 1 : 0;
  for (int i = 0; i < max_pitch; i++) {
    const float corr = xcorr[i];
    if (corr > 0.0f) {
      float corr16 = corr;
      corr16 *= 1e-12f;
      const float num = corr16 * corr16;
      if (num * best_den1 > best_num1 * syy) {
        if (num * best_den0 > best_num0 * syy) {
          best_num1 = best_num0; best_den1 = best_den0; best_pitch[1] = best_pitch[0];
          best_num0 = num; best_den0 = syy; best_pitch[0] = i;
        } else { best_num1 = num; best_den1 = syy; best_pitch[1] = i; }
      }
    }
    syy += y[i + len] * y[i + len] - y[i] * y[i];
    if (syy < 1.0f) syy = 1.0f;
  }
}
// pitch.rs:392-424 (celt_pitch_xcorr; every lag is a plain left-to-right inner product)
static void celt_pitch_xcorr(const float* x, const float* y, int len, int max_pitch, float* xcorr) {
  for (int i = 0; i < max_pitch; i++) {
    float sum = 0.f;
    for (int j = 0; j < len; j++) sum += x[j] * y[i + j];
    xcorr[i] = sum;
  }
}
// lpc.rs:465-530 (celt_autocorr, float)
static void celt_autocorr(const float* x, float* ac, const float* window, int overlap, int lag, int n) {
  const int fast_n = n - lag;
  std::vector<float> buf(x, x + n);
  if (overlap > 0)
    for (int i = 0; i < overlap; i++) { const float w = window[i]; buf[i] *= w; buf[n - i - 1] *= w; }
  celt_pitch_xcorr(buf.data(), buf.data(), fast_n, lag + 1, ac);
  for (int k = 0; k <= lag; k++) {
    float d = 0.f;
    for (int i = k + fast_n; i < n; i++) d += buf[i] * buf[i - k];
    ac[k] += d;
  }
}
// lpc.rs:146-190 (celt_lpc, float; fmaf where the reference calls mul_add_f32)
static void celt_lpc(float* lpc, const float* ac, int order) {
  for (int i = 0; i < order; i++) lpc[i] = 0.f;
  const float ac0 = ac[0];
  if (ac0 <= 1e-10f) return;
  float error = ac0;
  for (int i = 0; i < order; i++) {
    float rr = 0.f;
    for (int j = 0; j < i; j++) rr = fmaf(lpc[j], ac[i - j], rr);
    rr += ac[i + 1];
    const float r = -(rr / error);
    lpc[i] = r;
    for (int j = 0; j < (i + 1) >> 1; j++) {
      const float t1 = lpc[j], t2 = lpc[i - 1 - j];
      lpc[j] = fmaf(r, t2, t1);
      lpc[i - 1 - j] = fmaf(r, t1, t2);
    }
    const float r2 = r * r;
    error = fmaf(-r2, error, error);
    if (error <= 0.001f * ac0) break;
  }
}
// pitch.rs:1418-1545 (pitch_downsample, float) + :1358-1384 (celt_fir5)
static void pitch_downsample(const float* const x[2], int nch, float* x_lp, int len) {
  const int half = len / 2;
  for (int i = 1; i < half; i++) x_lp[i] = 0.25f * x[0][2 * i - 1] + 0.25f * x[0][2 * i + 1] + 0.5f * x[0][2 * i];
  x_lp[0] = 0.25f * x[0][1] + 0.5f * x[0][0];
  if (nch == 2) {
    for (int i = 1; i < half; i++) x_lp[i] += 0.25f * x[1][2 * i - 1] + 0.25f * x[1][2 * i + 1] + 0.5f * x[1][2 * i];
    x_lp[0] += 0.25f * x[1][1] + 0.5f * x[1][0];
  }
  float ac[5] = {0, 0, 0, 0, 0};
  celt_autocorr(x_lp, ac, nullptr, 0, 4, half);
  ac[0] *= 1.0001f;
  for (int i = 1; i <= 4; i++) { const float coeff = 0.008f * (float)i; ac[i] -= ac[i] * coeff * coeff; }
  float lpc[4];
  celt_lpc(lpc, ac, 4);
  float tmp = 1.0f;
  for (int i = 0; i < 4; i++) { tmp *= 0.9f; lpc[i] *= tmp; }
  const float c1 = 0.8f;
  const float num[5] = {lpc[0] + 0.8f, fmaf(c1, lpc[0], lpc[1]), fmaf(c1, lpc[1], lpc[2]), fmaf(c1, lpc[2], lpc[3]), c1 * lpc[3]};
  float m0 = 0, m1 = 0, m2 = 0, m3 = 0, m4 = 0;
  for (int i = 0; i < half; i++) {
    const float cur = x_lp[i];
    float sum = cur;
    sum = fmaf(num[0], m0, sum); sum = fmaf(num[1], m1, sum); sum = fmaf(num[2], m2, sum); sum = fmaf(num[3], m3, sum); sum = fmaf(num[4], m4, sum);
    m4 = m3; m3 = m2; m2 = m1; m1 = m0; m0 = cur;
    x_lp[i] = sum;
  }
}
// pitch.rs:557-696 (pitch_search, float)
static int pitch_search(const float* x_lp, const float* y, int len, int max_pitch) {
  const int lag = len + max_pitch;
  std::vector<float> x4(len >> 2), y4(lag >> 2), xcorr(max_pitch >> 1);
  int best[2] = {0, 0};
  for (int j = 0; j < len >> 2; j++) x4[j] = x_lp[2 * j];
  for (int j = 0; j < lag >> 2; j++) y4[j] = y[2 * j];
  celt_pitch_xcorr(x4.data(), y4.data(), len >> 2, max_pitch >> 2, xcorr.data());
  find_best_pitch(xcorr.data(), y4.data(), len >> 2, max_pitch >> 2, best);
  const int len_half = len >> 1, mp_half = max_pitch >> 1;
  for (int i = 0; i < mp_half; i++) {
    if (abs(i - 2 * best[0]) > 2 && abs(i - 2 * best[1]) > 2) { xcorr[i] = 0.f; continue; }
    float sum = 0.f;
    for (int j = 0; j < len_half; j++) sum += x_lp[j] * y[i + j];
    xcorr[i] = sum < -1.0f ? -1.0f : sum;
  }
  find_best_pitch(xcorr.data(), y, len_half, mp_half, best);
  if (best[0] > 0 && best[0] < mp_half - 1) {
    const float a = xcorr[best[0] - 1], b = xcorr[best[0]], c = xcorr[best[0] + 1];
    int offset = 0;
    if ((c - a) > 0.7f * (b - a)) offset = 1;
    else if ((a - c) > 0.7f * (b - c)) offset = -1;
    return 2 * best[0] - offset;
  }
  return 2 * best[0];
}
// comb_filter with an output buffer and no cross-fade window (celt.rs:176-330 as prefilter_and_fold calls it: overlap 0)
static void comb_filter_out(float* y, const float* x, int xs, int n, int T0, int T1, float g0, float g1, int tapset0, int tapset1) {
  (void)T0; (void)tapset0;
  if (g0 == 0.f && g1 == 0.f) { for (int i = 0; i < n; i++) y[i] = x[xs + i]; return; }
  T1 = imax(T1, COMBFILTER_MINPERIOD);
  if (g1 == 0.f) { for (int i = 0; i < n; i++) y[i] = x[xs + i]; return; }
  const float g10 = g1 * kTapsetGains[tapset1][0], g11 = g1 * kTapsetGains[tapset1][1], g12 = g1 * kTapsetGains[tapset1][2];
  float x4 = x[xs - T1 - 2], x3 = x[xs - T1 - 1], x2 = x[xs - T1], x1 = x[xs - T1 + 1];
  for (int i = 0; i < n; i++) {
    const float x0 = x[xs + i - T1 + 2];
    float acc = fmaf(g10, x2, x[xs + i]);
    acc = fmaf(g11, x1 + x3, acc);
    acc = fmaf(g12, x0 + x4, acc);
    y[i] = acc;
    x4 = x3; x3 = x2; x2 = x1; x1 = x0;
  }
}
// celt_decoder.rs:1327-1380 (prefilter_and_fold, float). Under Q2 the next inverse MDCT zero-fills what this writes.
static void prefilter_and_fold(CeltDecoder* st, int n) {
  const int start = DECODE_BUFFER_SIZE - n;
  float etmp[OVERLAP];
  for (int c = 0; c < st->channels; c++) {
    float* mem = st->decode_mem[c];
    comb_filter_out(etmp, mem, start, OVERLAP, st->postfilter_period_old, st->postfilter_period, -st->postfilter_gain_old, -st->postfilter_gain,
                    imax(st->postfilter_tapset_old, 0), imax(st->postfilter_tapset, 0));
    for (int i = 0; i < OVERLAP / 2; i++) mem[start + i] = g_window[i] * etmp[OVERLAP - 1 - i] + g_window[OVERLAP - 1 - i] * etmp[i];
  }
}
// celt_decoder.rs:1429-1855 (celt_decode_lost, float, no deep PLC)
static void celt_decode_lost(CeltDecoder* st, int N, int lm) {
  const int CC = st->channels;
  const int start_band = imin(imax(st->start_band, 0), NB_EBANDS), end_band = imin(imax(st->end_band, st->start_band), NB_EBANDS);
  const int eff_end = imax(start_band, imin(end_band, NB_EBANDS));
  const int loss_duration = st->loss_duration;
  const int noise_based = loss_duration >= 40 || start_band != 0 || st->skip_plc;
  if (noise_based) {
    for (int c = 0; c < CC; c++) memmove(st->decode_mem[c], st->decode_mem[c] + N, (DECODE_BUFFER_SIZE - N + OVERLAP) * sizeof(float));
    if (st->prefilter_and_fold) prefilter_and_fold(st, N);
    const float decay = loss_duration == 0 ? 1.5f : 0.5f;
    for (int c = 0; c < CC; c++)
      for (int i = start_band; i < end_band; i++) {
        const int idx = c * NB_EBANDS + i;
        st->old_ebands[idx] = fmax32(st->background_log_e[idx], st->old_ebands[idx] - decay);
      }
    uint32_t seed = st->rng;
    std::vector<float> X((size_t)CC * N, 0.f);
    for (int c = 0; c < CC; c++)
      for (int i = start_band; i < eff_end; i++) {
        const int bs = eb(i) << lm;
        int bw = (eb(i + 1) - eb(i)) << lm;
        if (bw > N - bs) bw = N - bs > 0 ? N - bs : 0;
        if (bw == 0) continue;
        float* x = X.data() + c * N + bs;
        for (int j = 0; j < bw; j++) { seed = lcg_rand(seed); x[j] = (float)((int32_t)seed >> 20); }
        renormalise_vector(x, bw, 1.0f);
      }
    st->rng = seed;
    float* out_syn[2];
    for (int c = 0; c < CC; c++) out_syn[c] = st->decode_mem[c] + DECODE_BUFFER_SIZE - N;
    celt_synthesis(X.data(), out_syn, st->old_ebands, start_band, eff_end, CC, CC, 0, lm, st->downsample > 1 ? st->downsample : 1, 0, st->quirks);
    st->prefilter_and_fold = 0;
    st->skip_plc = 1;
  } else {
    enum { LPC_ORDER = 24, PLC_LAG_MAX = 720, PLC_LAG_MIN = 100 };
    int pitch_index;
    if (loss_duration == 0) {
      const float* views[2] = {st->decode_mem[0], st->decode_mem[1]};
      std::vector<float> lp(DECODE_BUFFER_SIZE >> 1);
      pitch_downsample(views, CC, lp.data(), DECODE_BUFFER_SIZE);
      pitch_index = PLC_LAG_MAX - pitch_search(lp.data() + (PLC_LAG_MAX >> 1), lp.data(), DECODE_BUFFER_SIZE - PLC_LAG_MAX, PLC_LAG_MAX - PLC_LAG_MIN);
      st->last_pitch_index = pitch_index;
    } else pitch_index = st->last_pitch_index;
    const float fade = loss_duration == 0 ? 1.0f : 0.8f;
    pitch_index = imin(imax(pitch_index, PLC_LAG_MIN), PLC_LAG_MAX);
    const int pi = imin(pitch_index, MAX_PERIOD);
    const int exc_length = imin(2 * pi, MAX_PERIOD);
    std::vector<float> exc(MAX_PERIOD + LPC_ORDER), fir(exc_length);
    for (int c = 0; c < CC; c++) {
      float* mem = st->decode_mem[c];
      for (int i = 0; i < MAX_PERIOD + LPC_ORDER; i++) exc[i] = mem[DECODE_BUFFER_SIZE + OVERLAP - MAX_PERIOD - LPC_ORDER + i];
      float* lpc = st->lpc[c];
      if (loss_duration == 0) {
        float ac[LPC_ORDER + 1];
        celt_autocorr(exc.data() + LPC_ORDER, ac, g_window, OVERLAP, LPC_ORDER, MAX_PERIOD);
        ac[0] *= 1.0001f;
        for (int i = 1; i <= LPC_ORDER; i++) { const float f = 0.008f * 0.008f * (float)(i * i); ac[i] -= ac[i] * f; }
        celt_lpc(lpc, ac, LPC_ORDER);
      }
      const int s0 = MAX_PERIOD - exc_length;
      for (int i = 0; i < exc_length; i++) {
        float acc = exc[LPC_ORDER + s0 + i];
        for (int t = 0; t < LPC_ORDER; t++) acc += lpc[t] * exc[LPC_ORDER + s0 + i - 1 - t];
        fir[i] = acc;
      }
      for (int i = 0; i < exc_length; i++) exc[LPC_ORDER + s0 + i] = fir[i];
      float e1 = 1.0f, e2 = 1.0f;
      const int dl = exc_length / 2;
      for (int i = 0; i < dl; i++) {
        const float a = exc[LPC_ORDER + MAX_PERIOD - dl + i]; e1 += a * a;
        const float b = exc[LPC_ORDER + MAX_PERIOD - 2 * dl + i]; e2 += b * b;
      }
      e1 = fmin32(e1, e2);
      const float decay = celt_sqrt((0.5f * e1) / e2);
      memmove(mem, mem + N, (DECODE_BUFFER_SIZE - N) * sizeof(float));
      const int xo = MAX_PERIOD - pi, xl = N + OVERLAP, si = DECODE_BUFFER_SIZE - N, rb = DECODE_BUFFER_SIZE - MAX_PERIOD - N + xo;
      float att = fade * decay, s1 = 0.f;
      int j = 0;
      for (int i = 0; i < xl; i++) {
        if (j >= pi) { j -= pi; att *= decay; }
        mem[si + i] = att * exc[LPC_ORDER + xo + j];
        const float r = mem[rb + j];
        s1 += r * r;
        j++;
      }
      float lm_[LPC_ORDER];
      for (int i = 0; i < LPC_ORDER; i++) lm_[i] = mem[si - 1 - i];
      for (int i = 0; i < xl; i++) {   // lpc.rs:343-376 (celt_iir)
        float acc = mem[si + i];
        for (int t = 0; t < LPC_ORDER; t++) acc -= lpc[t] * lm_[t];
        mem[si + i] = acc;
        for (int t = LPC_ORDER - 1; t >= 1; t--) lm_[t] = lm_[t - 1];
        lm_[0] = acc;
      }
      float s2 = 0.f;
      for (int i = 0; i < xl; i++) s2 += mem[si + i] * mem[si + i];
      if (s1 > 0.2f * s2) {
        if (s1 < s2) {
          const float ratio = celt_sqrt((0.5f * s1 + 1.0f) / (s2 + 1.0f));
          for (int i = 0; i < OVERLAP; i++) mem[si + i] *= 1.0f - g_window[i] * (1.0f - ratio);
          for (int i = OVERLAP; i < xl; i++) mem[si + i] *= ratio;
        }
      } else for (int i = 0; i < xl; i++) mem[si + i] = 0.f;
    }
    st->prefilter_and_fold = 1;
  }
  st->loss_duration = imin(10000, loss_duration + (1 << lm));
}

void CeltDecoder::init(int channels_, int downsample_) {
  init_tables_once();
  channels = channels_;
  stream_channels = channels_;
  downsample = downsample_;
  start_band = 0;
  end_band = NB_EBANDS;
  disable_inv = channels_ == 1;  // opus_custom_decoder_init: disable_inv = channels == 1
  reset();
}
void CeltDecoder::reset() {  // types.rs:989-1043
  rng = 0; error = 0; loss_duration = 0;
  skip_plc = 1; prefilter_and_fold = 0; last_pitch_index = 0;
  memset(lpc, 0, sizeof(lpc));
  postfilter_period = postfilter_period_old = 0;
  postfilter_gain = postfilter_gain_old = 0.f;
  postfilter_tapset = postfilter_tapset_old = 0;
  preemph_mem[0] = preemph_mem[1] = 0.f;
  memset(decode_mem, 0, sizeof(decode_mem));
  for (int i = 0; i < 2 * NB_EBANDS; i++) { old_ebands[i] = 0.f; old_log_e[i] = old_log_e2[i] = -28.0f; background_log_e[i] = 0.f; }
}

// window_48000_960.rs (for the front end's smooth_fade, opus_decoder.rs:100-135)
const float* celt_window120() { init_tables_once(); return g_window; }

// celt_decoder.rs:2751-3163 (prepare_frame) + :3209-4137 (celt_decode_with_ec_dred_impl)
int celt_decode_with_ec(CeltDecoder* st, const uint8_t* data, int len, float* pcm, int frame_size, EcDec* ext_dec, int accum) {
  init_tables_once();
  const uint32_t quirks = st->quirks;
  const int CC = st->channels, C = st->stream_channels;
  const int start = st->start_band, end = st->end_band;
  const int downsample = st->downsample;
  int scaled = frame_size * downsample;
  if (scaled > (SHORT_MDCT << MAX_LM)) return -1;
  int lm = -1;
  for (int c = 0; c <= MAX_LM; c++) if ((SHORT_MDCT << c) == scaled) { lm = c; break; }
  if (lm < 0) return -1;
  const int M = 1 << lm, N = M * SHORT_MDCT;
  if (len > 1275) return -1;
  const int eff_end = imin(end, NB_EBANDS);
  if (data == nullptr || len <= 1) {   // packet loss (celt_decoder.rs:2869-2873, 3245-3309)
    celt_decode_lost(st, N, lm);
    float* lost_syn[2];
    for (int c = 0; c < CC; c++) lost_syn[c] = st->decode_mem[c] + DECODE_BUFFER_SIZE - N;
    deemphasis(lost_syn, pcm, N, CC, downsample, st->preemph_mem, accum);
    return N / downsample;
  }
  if (st->loss_duration == 0) st->skip_plc = 0;

  EcDec own;
  EcDec* dec = ext_dec;
  if (!dec) { own.init(data, (uint32_t)len); dec = &own; }
  if (C == 1) for (int i = 0; i < NB_EBANDS; i++) st->old_ebands[i] = fmax32(st->old_ebands[i], st->old_ebands[NB_EBANDS + i]);

  const int len_bits = len * 8;
  int tell = dec->tell();
  int silence = 0;
  if (tell >= len_bits) silence = 1;
  else if (tell == 1) {
#ifdef ORC_GEN
    if (dec->gen) dec->gen->forced = 0;
#endif
    silence = dec->bit_logp(15);
  }
  if (silence) { dec->nbits_total += len_bits - dec->tell(); tell = len_bits; }
  else tell = dec->tell();

  float postfilter_gain = 0.f;
  int postfilter_pitch = 0, postfilter_tapset = 0;
  if (start == 0 && tell + 16 <= len_bits) {
#ifdef ORC_GEN
    if (dec->gen) dec->gen->forced = dec->gen->unit() < dec->gen->p_postfilter;
#endif
    if (dec->bit_logp(1)) {
      int octave = (int)dec->uint(6);
      postfilter_pitch = (16 << octave) + (int)dec->bits(4 + octave) - 1;
#ifdef ORC_GEN
      if (dec->gen) dec->gen->forced = dec->gen->next() % 3u;   // a mild comb (gain 0.09 - 0.28), as on voiced material
#endif
      int qg = (int)dec->bits(3);
      if (dec->tell() + 2 <= len_bits) postfilter_tapset = dec->icdf(kTapsetIcdf, 2);
      postfilter_gain = 0.09375f * (float)(qg + 1);
    }
    tell = dec->tell();
  }
  int is_transient = 0;
#ifdef ORC_GEN
  if (dec->gen) dec->gen->forced = dec->gen->unit() < dec->gen->p_transient;
#endif
  if (lm > 0 && tell + 3 <= len_bits) { is_transient = dec->bit_logp(3); tell = dec->tell(); }
#ifdef ORC_GEN
  if (dec->gen) dec->gen->forced = dec->gen->frame == 0;   // intra energy on the first frame of a stream only
#endif
  const int short_blocks = is_transient ? M : 0;
  int intra = 0;
  if (tell + 3 <= len_bits) intra = dec->bit_logp(3);
#ifdef ORC_GEN
  if (dec->gen) dec->gen->forced = -1;
#endif
  if (!intra && st->loss_duration != 0) {   // celt_decoder.rs:2943-2969: the first good frame after a loss starts from safer energies
    const int missing = imin(10, st->loss_duration >> lm);
    const float safety = lm == 0 ? 1.5f : (lm == 1 ? 0.5f : 0.0f);
    for (int c = 0; c < 2; c++)
      for (int i = start; i < end; i++) {
        const int idx = c * NB_EBANDS + i;
        float e0 = st->old_ebands[idx];
        const float e1 = st->old_log_e[idx], e2 = st->old_log_e2[idx];
        if (e0 < fmax32(e1, e2)) {
          float slope = fmax32(e1 - e0, 0.5f * (e2 - e0));
          slope = fmin32(slope, 2.0f);
          e0 -= fmax32((float)(missing + 1) * slope, 0.0f);
          st->old_ebands[idx] = fmax32(e0, -20.0f);
        } else st->old_ebands[idx] = fmin32(st->old_ebands[idx], fmin32(e1, e2));
        st->old_ebands[idx] -= safety;
      }
  }

  unquant_coarse_energy(start, end, st->old_ebands, intra, dec, C, lm, quirks & QUIRK_Q8_COARSE_CLAMP);
  int tf_res[NB_EBANDS] = {0};
  tf_decode(start, end, is_transient, tf_res, lm, dec);
  tell = dec->tell();
  int spread = SPREAD_NORMAL;
  if (tell + 4 <= len_bits) spread = dec->icdf(kSpreadIcdf, 5);
  int cap[NB_EBANDS], offsets[NB_EBANDS] = {0};
  init_caps(cap, lm, C);
  int dynalloc_logp = 6;
  const int total_bits_full = len_bits << BITRES;
  int dyn_total = total_bits_full;
  int tell_frac = (int)dec->tell_frac();
  for (int i = start; i < end; i++) {
    int width = (C * (eb(i + 1) - eb(i))) << lm;
    int quanta = imin(width << BITRES, imax(6 << BITRES, width));
    int loop_logp = dynalloc_logp, boost = 0;
    while (tell_frac + (loop_logp << BITRES) < dyn_total && boost < cap[i]) {
      int flag = dec->bit_logp(loop_logp);
      tell_frac = (int)dec->tell_frac();
      if (!flag) break;
      boost += quanta;
      dyn_total -= quanta;
      loop_logp = 1;
    }
    offsets[i] = boost;
    if (boost > 0) dynalloc_logp = imax(2, dynalloc_logp - 1);
  }
  int fine_quant[NB_EBANDS] = {0}, pulses[NB_EBANDS] = {0}, fine_priority[NB_EBANDS] = {0};
  int alloc_trim = tell_frac + (6 << BITRES) <= dyn_total ? dec->icdf(kTrimIcdf, 7) : 5;
  // Q1 (celt_decoder.rs:3054-3057): the float build hands the post-dynalloc budget to quant_all_bands.
  const int frame_total_bits = (quirks & QUIRK_Q1_DYNALLOC_BUDGET) ? dyn_total : total_bits_full;
  int bits = total_bits_full - (int)dec->tell_frac() - 1;
  int anti_collapse_rsv = is_transient && lm >= 2 && bits >= ((lm + 2) << BITRES) ? (1 << BITRES) : 0;
  bits -= anti_collapse_rsv;
  int intensity = 0, dual_stereo = 0, balance = 0;
  int coded_bands = compute_allocation(start, end, offsets, cap, alloc_trim, &intensity, &dual_stereo, bits, &balance,
                                       pulses, fine_quant, fine_priority, C, lm, dec, quirks & QUIRK_Q7_EBITS_BUST,
                                       quirks & QUIRK_Q9_DUAL_STEREO_RSV);
  unquant_fine_energy(start, end, st->old_ebands, fine_quant, dec, C);

  const int stride = DECODE_BUFFER_SIZE + OVERLAP;
  for (int c = 0; c < CC; c++) memmove(st->decode_mem[c], st->decode_mem[c] + N, (DECODE_BUFFER_SIZE - N + OVERLAP) * sizeof(float));
  (void)stride;

  uint8_t collapse_masks[2 * NB_EBANDS] = {0};
  std::vector<float> X(C * N, 0.0f);
  std::vector<int32_t> iy;
  if (st->trace) iy.assign(C * N, 0);
  quant_all_bands(st, start, end, X.data(), C == 2 ? X.data() + N : nullptr, collapse_masks, pulses, short_blocks, spread,
                  dual_stereo, imax(intensity, 0), tf_res, frame_total_bits - anti_collapse_rsv, balance, dec, lm,
                  imax(coded_bands, 0), &st->rng, st->trace ? iy.data() : nullptr);
  int anti_collapse_on = 0;
  if (anti_collapse_rsv > 0) anti_collapse_on = (int)dec->bits(1);
  unquant_energy_finalise(start, end, st->old_ebands, fine_quant, fine_priority, len_bits - dec->tell(), dec, C);
  if (anti_collapse_on)
    anti_collapse(X.data(), collapse_masks, lm, C, N, start, end, st->old_ebands, st->old_log_e, st->old_log_e2, pulses, st->rng);
  if (silence) for (int i = 0; i < 2 * NB_EBANDS; i++) st->old_ebands[i] = -28.0f;

  if (st->prefilter_and_fold) prefilter_and_fold(st, N);   // celt_decoder.rs:3641 (after the history has moved, :3354)
  float* out_syn[2];
  for (int c = 0; c < CC; c++) out_syn[c] = st->decode_mem[c] + DECODE_BUFFER_SIZE - N;
  celt_synthesis(X.data(), out_syn, st->old_ebands, start, eff_end, C, CC, is_transient, lm, downsample, silence, quirks);
  if (st->trace) {
    CeltTrace* t = st->trace;
    t->lm = lm; t->C = C; t->N = N; t->silence = silence; t->transient = is_transient; t->intra = intra; t->spread = spread;
    t->anti_collapse_on = anti_collapse_on; t->intensity = intensity; t->dual_stereo = dual_stereo; t->coded_bands = coded_bands;
    t->pf_pitch = postfilter_pitch; t->pf_tapset = postfilter_tapset; t->pf_gain = postfilter_gain;
    memcpy(t->tf_res, tf_res, sizeof(tf_res)); memcpy(t->pulses, pulses, sizeof(pulses)); memcpy(t->fine_quant, fine_quant, sizeof(fine_quant));
    t->iy = iy; t->X = X; t->bandE.assign(st->old_ebands, st->old_ebands + 2 * NB_EBANDS);
    t->syn.resize(CC * (N + OVERLAP));
    for (int c = 0; c < CC; c++) memcpy(t->syn.data() + c * (N + OVERLAP), out_syn[c], (N + OVERLAP) * sizeof(float));
    t->seed_out = st->rng;
  }

  st->postfilter_period = imax(st->postfilter_period, COMBFILTER_MINPERIOD);
  st->postfilter_period_old = imax(st->postfilter_period_old, COMBFILTER_MINPERIOD);
  const int os = DECODE_BUFFER_SIZE - N;
  for (int c = 0; c < CC; c++) {
    int first = imin(SHORT_MDCT, N);
    comb_filter_in_place(st->decode_mem[c], os, first, st->postfilter_period_old, st->postfilter_period, st->postfilter_gain_old,
                         st->postfilter_gain, imax(st->postfilter_tapset_old, 0), imax(st->postfilter_tapset, 0));
    if (lm != 0 && first < N)
      comb_filter_in_place(st->decode_mem[c], os + first, N - first, st->postfilter_period, postfilter_pitch, st->postfilter_gain,
                           postfilter_gain, imax(st->postfilter_tapset, 0), imax(postfilter_tapset, 0));
  }
  st->postfilter_period_old = st->postfilter_period;
  st->postfilter_gain_old = st->postfilter_gain;
  st->postfilter_tapset_old = st->postfilter_tapset;
  st->postfilter_period = postfilter_pitch;
  st->postfilter_gain = postfilter_gain;
  st->postfilter_tapset = postfilter_tapset;
  if (lm != 0) {
    st->postfilter_period_old = st->postfilter_period;
    st->postfilter_gain_old = st->postfilter_gain;
    st->postfilter_tapset_old = st->postfilter_tapset;
  }
  if (C == 1) memcpy(st->old_ebands + NB_EBANDS, st->old_ebands, NB_EBANDS * sizeof(float));
  if (is_transient) {
    for (int i = 0; i < 2 * NB_EBANDS; i++) st->old_log_e[i] = fmin32(st->old_log_e[i], st->old_ebands[i]);
  } else {
    memcpy(st->old_log_e2, st->old_log_e, sizeof(st->old_log_e));
    memcpy(st->old_log_e, st->old_ebands, sizeof(st->old_ebands));
  }
  float increase = (float)imin(st->loss_duration + M, 160) * 0.001f;
  for (int i = 0; i < 2 * NB_EBANDS; i++) st->background_log_e[i] = fmin32(st->background_log_e[i] + increase, st->old_ebands[i]);
  for (int c = 0; c < 2; c++) {
    for (int i = 0; i < start; i++) { st->old_ebands[c * NB_EBANDS + i] = 0.f; st->old_log_e[c * NB_EBANDS + i] = st->old_log_e2[c * NB_EBANDS + i] = -28.0f; }
    for (int i = end; i < NB_EBANDS; i++) { st->old_ebands[c * NB_EBANDS + i] = 0.f; st->old_log_e[c * NB_EBANDS + i] = st->old_log_e2[c * NB_EBANDS + i] = -28.0f; }
  }
  st->rng = dec->rng;
  if (st->trace) {
    st->trace->post.resize(CC * N);
    for (int c = 0; c < CC; c++) memcpy(st->trace->post.data() + c * N, out_syn[c], N * sizeof(float));
  }
  deemphasis(out_syn, pcm, N, CC, downsample, st->preemph_mem, accum);
  st->loss_duration = 0;
  st->prefilter_and_fold = 0;
  if (dec->tell() > 8 * len) return -4;  // InvalidPacket (celt_decoder.rs:4128)
  if (dec->error) st->error = 1;
  return N / downsample;
}

}  // namespace orc

// ---------------------------------------------------------------- unit known-answer hooks (tests/test_unit_kat.py)
// One word-array entry point over the building blocks above, so that the reference's unit tests for this path can be
// replayed on the oracle; the product's kernel bodies answer the same protocol through tests/emul (CPU) and
// bd_debug_kat (GPU). Floats travel as their bit patterns.
//   1 bitexact_cos            in x[n]                              out cos[n]                 bands.rs:4805-4831
//   2 bitexact_log2tan        in (isin, icos)[n]                   out v[n]                   bands.rs:4836-4884
//   3 cwrsi                   in n, k, count, inc, i0              out y[count][n]            cwrs.rs:681-786
//   4 comb_filter_in_place    in n, T0, T1, tap0, tap1, g0, g1, hist, x[hist + n]   out x'[n] celt.rs:868-1190
//   5 ifft (480 >> shift)     in shift, cpx[nfft]                  out cpx[nfft]              kiss_fft.rs:1364-1440
//   6 frame IMDCT             in lm, transient, freq[N]            out time[N]                mdct.rs:1746-1843
//   7 range decoder script    in len, bytes[len] (one per word), ops (op, a, b)[..]           out one word per op
//        op 0 bit_logp(a)  1 icdf(table at word offset a of the input, ftb b)  2 uint(a)  3 bits(a)  4 tell  5 tell_frac
//        6 state: rng  7 val  8 nbits_total  9 offs  10 rem  11 error  12 nend_bits     range.rs:659-695, entdec.rs:327-377
//   8 pulse cache             in (band, lm, pulses)[n]             out (bits, bits2pulses(bits), max pulses of the row)[n]   rate.rs:1424-1465
extern "C" int orc_kat(int what, const uint32_t* in, uint32_t n_in, uint32_t* out, uint32_t n_out) {
  using namespace orc;
  init_tables_once();
  auto F = [](uint32_t w) { return bits2f(w); };
  auto U = [](float f) { uint32_t b; memcpy(&b, &f, 4); return b; };
  switch (what) {
    case 1:
      for (uint32_t i = 0; i < n_in && i < n_out; i++) out[i] = (uint32_t)bitexact_cos((int)in[i]);
      return 0;
    case 2:
      for (uint32_t i = 0; 2 * i + 1 < n_in && i < n_out; i++) out[i] = (uint32_t)bitexact_log2tan((int)in[2 * i], (int)in[2 * i + 1]);
      return 0;
    case 3: {
      const int n = (int)in[0], k = (int)in[1];
      const uint32_t count = in[2], inc = in[3], i0 = in[4];
      std::vector<int> y(n);
      for (uint32_t j = 0; j < count && (j + 1) * n <= n_out; j++) {
        cwrsi(n, k, i0 + j * inc, y.data());
        for (int t = 0; t < n; t++) out[j * n + t] = (uint32_t)y[t];
      }
      return 0;
    }
    case 4: {
      const int n = (int)in[0], T0 = (int)in[1], T1 = (int)in[2], tap0 = (int)in[3], tap1 = (int)in[4], hist = (int)in[7];
      std::vector<float> ch(hist + n);
      for (int i = 0; i < hist + n; i++) ch[i] = F(in[8 + i]);
      comb_filter_in_place(ch.data(), hist, n, T0, T1, F(in[5]), F(in[6]), tap0, tap1);
      for (int i = 0; i < n && i < (int)n_out; i++) out[i] = U(ch[hist + i]);
      return 0;
    }
    case 5: {
      const int shift = (int)in[0], nfft = 480 >> shift;
      cpx f[480];
      const int16_t* br = kBitrev[shift];
      for (int i = 0; i < nfft; i++) { f[br[i]].r = F(in[1 + 2 * i]); f[br[i]].i = -F(in[2 + 2 * i]); }   // kiss_fft.rs:739-745
      fft_impl(f, shift);
      for (int i = 0; i < nfft; i++) { out[2 * i] = U(f[i].r); out[2 * i + 1] = U(-f[i].i); }
      return 0;
    }
    case 6: {
      const int lm = (int)in[0], transient = (int)in[1], M = 1 << lm, N = M * SHORT_MDCT;
      int nbk, nb, shift;
      if (transient) { nbk = M; nb = SHORT_MDCT; shift = MAX_LM; } else { nbk = 1; nb = SHORT_MDCT << lm; shift = MAX_LM - lm; }
      std::vector<float> temp(N), o(N + OVERLAP, 0.f);
      for (int b = 0; b < nbk; b++) {
        for (int k = 0; k < nb; k++) temp[k] = F(in[2 + b + k * nbk]);
        clt_mdct_backward(temp.data(), o.data() + b * nb, N + OVERLAP - b * nb, shift, true);
      }
      for (int i = 0; i < N && i < (int)n_out; i++) out[i] = U(o[i]);
      return 0;
    }
    case 7: {
      const uint32_t len = in[0];
      std::vector<uint8_t> bytes(len);
      for (uint32_t i = 0; i < len; i++) bytes[i] = (uint8_t)in[1 + i];
      EcDec d;
      d.init(bytes.data(), len);
      uint32_t o = 0;
      for (uint32_t p = 1 + len; p + 2 < n_in && o < n_out; p += 3, o++) {
        const uint32_t op = in[p], a = in[p + 1], b = in[p + 2];
        switch (op) {
          case 0: out[o] = (uint32_t)d.bit_logp(a); break;
          case 1: {
            uint8_t tab[64];
            for (int i = 0; i < 64 && a + i < n_in; i++) tab[i] = (uint8_t)in[a + i];
            out[o] = (uint32_t)d.icdf(tab, b);
          } break;
          case 2: out[o] = d.uint(a); break;
          case 3: out[o] = d.bits(a); break;
          case 4: out[o] = (uint32_t)d.tell(); break;
          case 5: out[o] = d.tell_frac(); break;
          case 6: out[o] = d.rng; break;
          case 7: out[o] = d.val; break;
          case 8: out[o] = (uint32_t)d.nbits_total; break;
          case 9: out[o] = d.offs; break;
          case 10: out[o] = (uint32_t)d.rem; break;
          case 11: out[o] = (uint32_t)d.error; break;
          case 12: out[o] = (uint32_t)d.nend_bits; break;
          default: return -1;
        }
      }
      return 0;
    }
    case 8:
      for (uint32_t i = 0; 3 * i + 2 < n_in && 3 * i + 2 < n_out; i++) {
        const int band = (int)in[3 * i], lm = (int)in[3 * i + 1], pulses = (int)in[3 * i + 2];
        const uint8_t* t = cache_row(band, lm);
        const int bits = (t && pulses > t[0]) ? 0 : pulses2bits(band, lm, pulses);
        out[3 * i] = (uint32_t)bits;
        out[3 * i + 1] = (uint32_t)bits2pulses(band, lm, bits);
        out[3 * i + 2] = t ? t[0] : 0u;
      }
      return 0;
  }
  return -1;
}
