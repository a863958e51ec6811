// Phase 1 — CELT entropy parse, one lane per frame.
//
// Replaces the bit-stream reading half of the reference's CELT frame decoder:
//   prepare_frame                     src/celt/celt_decoder.rs:2751-3163
//   tf_decode                         src/celt/celt_decoder.rs:2704-2748
//   unquant_coarse/fine/finalise      src/celt/quant_bands.rs:902-950, 1850-1970 (symbols only)
//   clt_compute_allocation            src/celt/rate.rs:505-1011
//   quant_all_bands / quant_band* / quant_partition / compute_theta (entropy reads and
//   integer bit-budget logic only)    src/celt/bands.rs:274-495, 1917-3082
//   decode_pulses (the index read)    src/celt/cwrs.rs:427-452
// Nothing here touches decoder state or float spectra: every decision below is an integer
// function of the packet bytes, so all frames of all streams parse concurrently. The lane
// writes a compact "program" (header, energy symbols, per-band words, a node stream for the
// split tree and a leaf list with PVQ indices) that phase 2 replays data-parallel.
#pragma once
#include "mb_rangedec.cuh"

namespace mb {

MB_TBL uint8_t kTrimIcdf[11] = {126, 124, 119, 109, 87, 41, 19, 9, 4, 2, 0};
MB_TBL uint8_t kSpreadIcdf[4] = {25, 23, 2, 0};
MB_TBL uint8_t kTapsetIcdf[3] = {2, 1, 0};
MB_TBL uint8_t kSmallEnergyIcdf[3] = {2, 1, 0};
MB_TBL int8_t kTfSelect[32] = {0, -1, 0, -1, 0, -1, 0, -1, 0, -1, 0, -2, 1, 0, 1, -1,
                               0, -2, 0, -3, 2, 0, 1, -1, 0, -2, 0, -3, 3, 0, 1, -1};
MB_TBL uint8_t kLog2Frac[24] = {0, 8, 13, 16, 19, 21, 23, 24, 26, 27, 28, 29, 30, 31, 32, 32, 33, 34, 34, 35, 36, 36, 37, 37};
MB_TBL int16_t kPvqRowOff[15] = {0, 208, 415, 621, 826, 1030, 1233, 1336, 1389, 1421, 1441, 1455, 1464, 1470, 1473};
MB_TBL uint16_t kExp2Table8[8] = {16384, 17866, 19483, 21247, 23170, 25267, 27554, 30048};

MB_DEV uint32_t pvq_u(int n, int k) {   // one load from the padded copy of CELT_PVQ_U_DATA (celt_tables_data.h)
  int row = n < k ? n : k, col = n < k ? k : n;
  return kPvqU2[row * PVQ_U_PITCH + col];
}
MB_DEV int get_pulses(int i) { return i < 8 ? i : (8 + (i & 7)) << ((i >> 3) - 1); }
// pulse cache row of (band, lm); lm may be -1. Returns index into kCacheBits50 or -1.
MB_DEV int cache_row(int band, int lm) { return kCacheIndex50[(lm + 1) * NB_EBANDS + band]; }

MB_DEV int bits2pulses(int row, int bits) {
  if (bits <= 0 || row < 0) return 0;
  const uint8_t* t = kCacheBits50 + row;
  int lo = 0, hi = t[0];
  int target = bits - 1;
  for (int i = 0; i < 6; i++) {
    int mid = (lo + hi + 1) >> 1;
    if ((int)t[mid] >= target) hi = mid; else lo = mid;
  }
  int lo_v = lo == 0 ? -1 : (int)t[lo];
  return target - lo_v <= (int)t[hi] - target ? lo : hi;
}
MB_DEV int pulses2bits(int row, int pulses) { return (pulses == 0 || row < 0) ? 0 : kCacheBits50[row + pulses] + 1; }

MB_DEV int frac_mul16(int a, int b) { return (16384 + (int)(int16_t)a * (int)(int16_t)b) >> 15; }
MB_DEV int bitexact_cos(int x) {
  int x2 = (4096 + x * x) >> 13;
  x2 = (32767 - x2) + frac_mul16(x2, -7651 + frac_mul16(x2, 8277 + frac_mul16(-626, x2)));
  return (int16_t)(1 + x2);
}
MB_DEV int bitexact_log2tan(int isin, int icos) {
  int lc = ilog32((uint32_t)icos), ls = ilog32((uint32_t)isin);
  icos <<= 15 - lc;
  isin <<= 15 - ls;
  return ((ls - lc) << 11) + frac_mul16(isin, frac_mul16(isin, -2597) + 7932) - frac_mul16(icos, frac_mul16(icos, -2597) + 7932);
}
MB_DEV unsigned isqrt32(uint32_t v) {
  unsigned g = 0;
  int bshift = (ilog32(v) - 1) >> 1;
  unsigned b = 1u << bshift;
  do {
    uint32_t t = ((g << 1) + b) << bshift;
    if (t <= v) { g += b; v -= t; }
    b >>= 1; bshift--;
  } while (bshift >= 0);
  return g;
}

// Writer for one frame's record words (tile-of-32 SoA, see mb_common.cuh).
struct RecWriter {
  uint32_t* base;  // &rec[rec_index(f, 0)]
  int n_nodes, n_leaves, overflow;
  MB_MEM void put(int slot, uint32_t v) { base[(size_t)slot * 32u] = v; }
  MB_MEM void node(uint32_t w) {
    if (n_nodes < CAP_NODES) put(S_NODES + n_nodes, w); else overflow = 1;
    n_nodes++;
  }
  MB_MEM void leaf(uint32_t index, float gain, int x_off, int n, int k, int B) {
    if (n_leaves < CAP_LEAVES) {
      int s = S_LEAVES + 3 * n_leaves;
      put(s, index); put(s + 1, f2bits(gain));
      put(s + 2, (uint32_t)x_off | ((uint32_t)n << 11) | ((uint32_t)k << 19) | ((uint32_t)B << 27));
    } else overflow = 1;
    n_leaves++;
  }
};

struct ParseCtx {
  RangeDec dec;
  RecWriter w;
  int band, intensity, remaining_bits, disable_inv, lm_frame;
};
struct Theta { int inv, imid, iside, delta, itheta, qalloc; };

// compute_theta, decode arm, integer part (src/celt/bands.rs:274-495)
MB_DEV void parse_theta(ParseCtx& cx, Theta& s, int n, int& b, int B, int B0, int lm, int stereo) {
  RangeDec& dec = cx.dec;
  int pulse_cap = kLogN400[cx.band] + lm * (1 << BITRES);
  int offset = (pulse_cap >> 1) - (stereo && n == 2 ? 16 : 4);
  // compute_qn (bands.rs:618-640)
  int qn;
  {
    int n2 = 2 * n - 1;
    if (stereo && n == 2) n2--;
    int qb = (b + n2 * offset) / n2;
    qb = imin(qb, b - pulse_cap - (4 << BITRES));
    qb = imin(qb, 8 << BITRES);
    if (qb < (1 << BITRES >> 1)) qn = 1;
    else { qn = kExp2Table8[qb & 7] >> (14 - (qb >> BITRES)); qn = (qn + 1) >> 1 << 1; }
  }
  if (stereo && cx.band >= cx.intensity) qn = 1;
  int itheta = 0, inv = 0;
  int tell0 = (int)dec.tell_frac();
  if (qn != 1) {
    if (stereo && n > 2) {  // step pdf
      int x0 = qn / 2, ft = 3 * (x0 + 1) + x0;
      int fs = (int)dec.decode((uint32_t)ft);
      int x = fs < (x0 + 1) * 3 ? fs / 3 : x0 + 1 + (fs - (x0 + 1) * 3);
      int fl, fh;
      if (x <= x0) { fl = 3 * x; fh = 3 * (x + 1); }
      else { fl = (x0 + 1) * 3 + (x - 1 - x0); fh = fl + 1; }
      dec.update((uint32_t)fl, (uint32_t)fh, (uint32_t)ft);
      itheta = x;
    } else if (B0 > 1 || stereo) {  // uniform pdf
      itheta = (int)dec.uint_((uint32_t)(qn + 1));
    } else {  // triangular pdf
      int hq = qn >> 1, ft = (hq + 1) * (hq + 1);
      int fm = (int)dec.decode((uint32_t)ft);
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
      dec.update((uint32_t)fl, (uint32_t)(fl + fs), (uint32_t)ft);
    }
    itheta = (int)((uint32_t)(itheta * 16384) / (uint32_t)qn);
  } else if (stereo) {
    if (b > (2 << BITRES) && cx.remaining_bits > (2 << BITRES)) inv = dec.bit_logp(2);
    if (cx.disable_inv) inv = 0;
  }
  int qalloc = (int)dec.tell_frac() - tell0;
  b -= qalloc;
  int imid, iside, delta;
  if (itheta == 0) { imid = 32767; iside = 0; delta = -16384; }
  else if (itheta == 16384) { imid = 0; iside = 32767; delta = 16384; }
  else {
    imid = bitexact_cos(itheta);
    iside = bitexact_cos(16384 - itheta);
    delta = frac_mul16(imax((n - 1) << 7, 0), bitexact_log2tan(iside, imid));
  }
  s.inv = inv; s.imid = imid; s.iside = iside; s.delta = delta; s.itheta = itheta; s.qalloc = qalloc;
}

// One quant_band call (mono path; src/celt/bands.rs:2133-2291 + quant_partition :1917-2131),
// integer side only. The split tree is walked with an explicit stack (depth <= LM+2) so that the
// lanes of a warp keep iterating one common loop instead of diverging through recursion.
MB_DEV_NOINLINE void parse_band(ParseCtx& cx, int x_off, int n, int b, int B, int lm, float gain, int tf_change) {
  if (n == 1) {  // quant_band_n1 (bands.rs:642-704), mono call
    uint32_t sign = 0;
    if (cx.remaining_bits >= 1 << BITRES) { sign = cx.dec.raw_bits(1); cx.remaining_bits -= 1 << BITRES; }
    cx.w.node(ND_N1 | (sign << 5));
    return;
  }
  int n_b = n / B;
  if (tf_change > 0) { B >>= tf_change; n_b <<= tf_change; }
  while ((n_b & 1) == 0 && tf_change < 0) { B <<= 1; n_b >>= 1; tf_change++; }

  struct Frame { int n, b, B, lm, x_off, mbits, sbits, rebal, itheta; float gmid, gside; uint8_t order, st; };
  Frame st[6];
  int sp = 0;
  st[0].n = n; st[0].b = b; st[0].B = B; st[0].lm = lm; st[0].x_off = x_off; st[0].gmid = gain; st[0].st = 0;
  // (for st==0 frames `gmid` carries the node's own gain)
  while (sp >= 0) {
    Frame& F = st[sp];
    if (F.st == 0) {
      int row = cache_row(cx.band, F.lm);
      if (F.lm != -1 && F.n > 2 && row >= 0 && F.b > (int)kCacheBits50[row + kCacheBits50[row]] + 12) {
        int half = F.n >> 1, clm = F.lm - 1, B0 = F.B, cB = (B0 + 1) >> 1;
        int bb = F.b;
        Theta t;
        parse_theta(cx, t, half, bb, cB, B0, clm, 0);
        int delta = t.delta;
        if (B0 > 1 && (t.itheta & 0x3fff)) {
          if (t.itheta > 8192) delta -= delta >> (4 - clm);
          else delta = imin(0, delta + ((half << BITRES) >> (5 - clm)));
        }
        cx.remaining_bits -= t.qalloc;
        int mbits = imax(0, imin(bb, (bb - delta) / 2));
        int sbits = bb - mbits;
        float g = F.gmid;
        F.mbits = mbits; F.sbits = sbits; F.rebal = cx.remaining_bits; F.itheta = t.itheta;
        F.order = mbits >= sbits;
        F.gmid = g * ((float)t.imid / 32768.0f);
        F.gside = g * ((float)t.iside / 32768.0f);
        F.st = 1;
        cx.w.node(ND_SPLIT | ((uint32_t)F.order << 3) | ((uint32_t)t.itheta << 7));
        Frame& Cn = st[sp + 1];
        Cn.n = half; Cn.B = cB; Cn.lm = clm; Cn.st = 0;
        if (F.order) { Cn.b = mbits; Cn.x_off = F.x_off; Cn.gmid = F.gmid; }
        else { Cn.b = sbits; Cn.x_off = F.x_off + half; Cn.gmid = F.gside; }
        sp++;
      } else {
        int q = bits2pulses(row, F.b);
        int curr_bits = pulses2bits(row, q);
        cx.remaining_bits -= curr_bits;
        while (cx.remaining_bits < 0 && q > 0) {
          cx.remaining_bits += curr_bits;
          q--;
          curr_bits = pulses2bits(row, q);
          cx.remaining_bits -= curr_bits;
        }
        if (q != 0) {
          int k = get_pulses(q);
          uint32_t idx = cx.dec.uint_(pvq_u(F.n, k) + pvq_u(F.n, k + 1));
          cx.w.leaf(idx, F.gmid, F.x_off, F.n, k, imax(F.B, 1));
          cx.w.node(ND_LEAF_PULSE);
        } else cx.w.node(ND_LEAF_EMPTY);
        sp--;
      }
    } else if (F.st == 1) {
      int half = F.n >> 1;
      Frame& Cn = st[sp + 1];
      Cn.n = half; Cn.B = (F.B + 1) >> 1; Cn.lm = F.lm - 1; Cn.st = 0;
      if (F.order) {
        int rebalance = F.mbits - (F.rebal - cx.remaining_bits);
        int sbits = F.sbits;
        if (rebalance > 3 << BITRES && F.itheta != 0) sbits += rebalance - (3 << BITRES);
        Cn.b = sbits; Cn.x_off = F.x_off + half; Cn.gmid = F.gside;
      } else {
        int rebalance = F.sbits - (F.rebal - cx.remaining_bits);
        int mbits = F.mbits;
        if (rebalance > 3 << BITRES && F.itheta != 16384) mbits += rebalance - (3 << BITRES);
        Cn.b = mbits; Cn.x_off = F.x_off; Cn.gmid = F.gmid;
      }
      F.st = 2;
      sp++;
    } else sp--;
  }
}

// quant_band_stereo, integer side (src/celt/bands.rs:2294-2514)
MB_DEV_NOINLINE void parse_band_stereo(ParseCtx& cx, int x_off, int y_off, int n, int b, int B, int lm, int tf_change) {
  if (n == 1) {
    uint32_t signs = 0;
    for (int c = 0; c < 2; c++)
      if (cx.remaining_bits >= 1 << BITRES) { signs |= cx.dec.raw_bits(1) << c; cx.remaining_bits -= 1 << BITRES; }
    cx.w.node(ND_N1 | (signs << 5));
    return;
  }
  Theta t;
  parse_theta(cx, t, n, b, B, B, lm, 1);
  float side = (float)t.iside / 32768.0f;
  if (n == 2) {
    int sbits = (t.itheta != 0 && t.itheta != 16384) ? 1 << BITRES : 0;
    int mbits = b - sbits;
    int c = t.itheta > 8192;
    cx.remaining_bits -= t.qalloc + sbits;
    uint32_t sign = sbits ? cx.dec.raw_bits(1) : 0u;
    cx.w.node(ND_STEREO | ((uint32_t)t.inv << 4) | (sign << 5) | ((uint32_t)t.itheta << 7));
    parse_band(cx, c ? y_off : x_off, n, mbits, B, lm, 1.0f, tf_change);
    return;
  }
  int mbits = imax(0, imin(b, (b - t.delta) / 2));
  int sbits = b - mbits;
  cx.remaining_bits -= t.qalloc;
  int rebalance = cx.remaining_bits;
  int order = mbits >= sbits;
  cx.w.node(ND_STEREO | ((uint32_t)order << 3) | ((uint32_t)t.inv << 4) | ((uint32_t)t.itheta << 7));
  if (order) {
    parse_band(cx, x_off, n, mbits, B, lm, 1.0f, tf_change);
    rebalance = mbits - (rebalance - cx.remaining_bits);
    if (rebalance > 3 << BITRES && t.itheta != 0) sbits += rebalance - (3 << BITRES);
    parse_band(cx, y_off, n, sbits, B, lm, side, tf_change);
  } else {
    parse_band(cx, y_off, n, sbits, B, lm, side, tf_change);
    rebalance = sbits - (rebalance - cx.remaining_bits);
    if (rebalance > 3 << BITRES && t.itheta != 16384) mbits += rebalance - (3 << BITRES);
    parse_band(cx, x_off, n, mbits, B, lm, 1.0f, tf_change);
  }
}

// Laplace-distributed coarse energy symbol (src/celt/quant_bands.rs:40-48, 90-125)
MB_DEV int laplace_symbol(RangeDec& dec, uint32_t fs, uint32_t decay) {
  int val = 0;
  uint32_t fl = 0;
  uint32_t fm = dec.decode_bin(15);
  if (fm >= fs) {
    val++;
    fl = fs;
    uint32_t rem = 32768u - 32u - fs;
    fs = (decay >= 16384u ? 0u : (uint32_t)(((uint64_t)rem * (16384u - decay)) >> 15)) + 1u;
    while (fs > 1 && fm >= fl + 2 * fs) {
      fs *= 2;
      fl += fs;
      fs = (uint32_t)(((uint64_t)(fs - 2u) * decay) >> 15) + 1u;
      val++;
    }
    if (fs <= 1) {
      int di = (int)((fm - fl) >> 1);
      val += di;
      fl += 2u * (uint32_t)di;
    }
    if (fm < fl + fs) val = -val; else fl += fs;
  }
  uint32_t high = fl + fs < 32768u ? fl + fs : 32768u;
  dec.update(fl, high, 32768u);
  return val;
}

// Whole-frame parse. `rec_base` = &rec[rec_index(f, 0)].
// Returns the status word (also stored in S_STATUS).
// `pre`: a range decoder already positioned behind the SILK payload of a hybrid frame (null: open a fresh one);
// `status_extra`: bits OR-ed into the status word (cell kind).
// START is a template parameter so that the CELT-only instance keeps `start == 0` folded away (the hybrid instance starts at 17).
template <int START>
MB_DEV_NOINLINE uint32_t celt_parse_frame(const uint8_t* data, int len, int lm, int C, int end, uint32_t quirks,
                                          int disable_inv, uint32_t* rec_base, const RangeDec* pre = nullptr, uint32_t status_extra = 0) {
  constexpr int start = START;
  ParseCtx cx;
  cx.w.base = rec_base; cx.w.n_nodes = 0; cx.w.n_leaves = 0; cx.w.overflow = 0;
  cx.disable_inv = disable_inv; cx.lm_frame = lm;
  RangeDec& dec = cx.dec;
  if (START != 0 && pre) dec = *pre; else dec.open(data, (uint32_t)len);
  const int M = 1 << lm;
  const int len_bits = len * 8;

  int tell = dec.tell();
  int silence = 0;
  if (tell >= len_bits) silence = 1;
  else if (tell == 1) silence = dec.bit_logp(15);
  if (silence) { dec.nbits_total += len_bits - dec.tell(); tell = len_bits; }
  else tell = dec.tell();

  int pf_pitch = 0, pf_tapset = 0, pf_qg1 = 0;
  if (start == 0 && tell + 16 <= len_bits) {
    if (dec.bit_logp(1)) {
      int octave = (int)dec.uint_(6);
      pf_pitch = (16 << octave) + (int)dec.raw_bits(4 + octave) - 1;
      pf_qg1 = (int)dec.raw_bits(3) + 1;
      if (dec.tell() + 2 <= len_bits) pf_tapset = dec.icdf(kTapsetIcdf, 2);
    }
    tell = dec.tell();
  }
  int is_transient = 0;
  if (lm > 0 && tell + 3 <= len_bits) { is_transient = dec.bit_logp(3); tell = dec.tell(); }
  int intra = 0;
  if (tell + 3 <= len_bits) intra = dec.bit_logp(3);

  // ---- coarse energy symbols (quant_bands.rs:902-950; the float recurrence runs in phase 2a)
  uint32_t ew[2 * NB_EBANDS];
  MB_ROLLED
  for (int i = 0; i < 2 * NB_EBANDS; i++) ew[i] = 0;
  {
    const uint8_t* prob = kEProbModel + (lm * 2 + intra) * 42;
    MB_ROLLED
    for (int i = start; i < end; i++)
      MB_ROLLED
      for (int c = 0; c < C; c++) {
        int t = dec.tell();
        int qi;
        if (len_bits - t >= 15) {
          int pi = 2 * imin(i, 20);
          qi = laplace_symbol(dec, (uint32_t)prob[pi] << 7, (uint32_t)prob[pi + 1] << 6);
        } else if (len_bits - t >= 2) {
          qi = dec.icdf(kSmallEnergyIcdf, 2);
          qi = (qi >> 1) ^ -(qi & 1);
        } else if (len_bits - t >= 1) qi = -dec.bit_logp(1);
        else qi = -1;
        ew[c * NB_EBANDS + i] = (uint32_t)(uint16_t)(int16_t)qi;
      }
  }
  // ---- tf_decode (celt_decoder.rs:2704-2748)
  int8_t tf_res[NB_EBANDS];
  MB_ROLLED
  for (int i = 0; i < NB_EBANDS; i++) tf_res[i] = 0;
  {
    uint32_t budget = (uint32_t)len_bits;
    uint32_t t = (uint32_t)dec.tell();
    uint32_t logp = is_transient ? 2 : 4;
    int tf_select_rsv = lm > 0 && t + logp < budget;
    if (tf_select_rsv) budget -= 1;
    int curr = 0, tf_changed = 0;
    MB_ROLLED
    for (int i = start; i < end; i++) {
      if (t + logp <= budget) {
        curr ^= dec.bit_logp(logp);
        t = (uint32_t)dec.tell();
        tf_changed |= curr;
      }
      tf_res[i] = (int8_t)curr;
      logp = is_transient ? 4 : 5;
    }
    int tf_select = 0;
    const int8_t* sel = kTfSelect + lm * 8 + 4 * is_transient;
    if (tf_select_rsv && sel[tf_changed] != sel[2 + tf_changed]) tf_select = dec.bit_logp(1);
    MB_ROLLED
    for (int i = start; i < end; i++) tf_res[i] = sel[2 * tf_select + tf_res[i]];
  }
  tell = dec.tell();
  int spread = 2;  // SPREAD_NORMAL
  if (tell + 4 <= len_bits) spread = dec.icdf(kSpreadIcdf, 5);

  // ---- init_caps (celt.rs:799-822) + dynalloc boosts (celt_decoder.rs:3021-3044)
  int cap[NB_EBANDS], offsets[NB_EBANDS];
  MB_ROLLED
  for (int i = 0; i < NB_EBANDS; i++) {
    int n = (eband(i + 1) - eband(i)) << lm;
    cap[i] = ((kCacheCaps50[NB_EBANDS * (2 * lm + C - 1) + i] + 64) * C * n) >> 2;
    offsets[i] = 0;
  }
  const int total_bits_full = len_bits << BITRES;
  int dyn_total = total_bits_full;
  int tell_frac = (int)dec.tell_frac();
  {
    int dynalloc_logp = 6;
    MB_ROLLED
    for (int i = start; i < end; i++) {
      int width = (C * (eband(i + 1) - eband(i))) << lm;
      int quanta = imin(width << BITRES, imax(6 << BITRES, width));
      int loop_logp = dynalloc_logp, boost = 0;
      while (tell_frac + (loop_logp << BITRES) < dyn_total && boost < cap[i]) {
        int flag = dec.bit_logp(loop_logp);
        tell_frac = (int)dec.tell_frac();
        if (!flag) break;
        boost += quanta;
        dyn_total -= quanta;
        loop_logp = 1;
      }
      offsets[i] = boost;
      if (boost > 0) dynalloc_logp = imax(2, dynalloc_logp - 1);
    }
  }
  int alloc_trim = tell_frac + (6 << BITRES) <= dyn_total ? dec.icdf(kTrimIcdf, 7) : 5;
  const int frame_total_bits = (quirks & Q1_DYNALLOC_BUDGET) ? dyn_total : total_bits_full;
  int bits = total_bits_full - (int)dec.tell_frac() - 1;
  int anti_collapse_rsv = is_transient && lm >= 2 && bits >= ((lm + 2) << BITRES) ? (1 << BITRES) : 0;
  bits -= anti_collapse_rsv;

  // ---- bit allocation (rate.rs:843-1011 + interp_bits2pulses :505-838)
  int pulses[NB_EBANDS], fine_quant[NB_EBANDS];
  uint32_t fine_prio_mask = 0;
  int intensity = 0, dual_stereo = 0, balance = 0, coded_bands;
  MB_ROLLED
  for (int i = 0; i < NB_EBANDS; i++) { pulses[i] = 0; fine_quant[i] = 0; }
  {
    int bits1[NB_EBANDS], bits2[NB_EBANDS], thresh[NB_EBANDS], trim_offset[NB_EBANDS];
    int total = imax(bits, 0);
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
    MB_ROLLED
    for (int j = start; j < end; j++) {
      int n = eband(j + 1) - eband(j);
      thresh[j] = imax(C << BITRES, (3 * n) << (lm + BITRES) >> 4);
      trim_offset[j] = (C * n * (alloc_trim - 5 - lm) * (end - j - 1) * (1 << (lm + BITRES))) >> 6;
      if ((n << lm) == 1) trim_offset[j] -= C << BITRES;
    }
    int lo = 1, hi = 10;
    while (lo <= hi) {
      int mid = (lo + hi) >> 1, done = 0, psum = 0;
      MB_ROLLED
      for (int j = end; j-- > start;) {
        int n = eband(j + 1) - eband(j);
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
    MB_ROLLED
    for (int j = start; j < end; j++) {
      int n = eband(j + 1) - eband(j);
      int b1 = (C * n * kBandAllocation[lo * NB_EBANDS + j]) << lm >> 2;
      int b2 = hi >= 11 ? cap[j] : (C * n * kBandAllocation[hi * NB_EBANDS + j]) << lm >> 2;
      if (b1 > 0) b1 = imax(0, b1 + trim_offset[j]);
      if (b2 > 0) b2 = imax(0, b2 + trim_offset[j]);
      if (lo > 0) b1 += offsets[j];
      b2 += offsets[j];
      if (offsets[j] > 0) skip_start = j;
      bits1[j] = b1;
      bits2[j] = imax(0, b2 - b1);
    }
    // interp_bits2pulses
    const int alloc_floor = C << BITRES, stereo = C > 1, logM = lm << BITRES;
    lo = 0; hi = 64;
    MB_ROLLED
    for (int it = 0; it < 6; it++) {
      int mid = (lo + hi) >> 1, psum = 0, done = 0;
      MB_ROLLED
      for (int j = end; j-- > start;) {
        int tmp = bits1[j] + ((mid * bits2[j]) >> 6);
        if (tmp >= thresh[j] || done) { done = 1; psum += imin(tmp, cap[j]); }
        else if (tmp >= alloc_floor) psum += alloc_floor;
      }
      if (psum > total) hi = mid; else lo = mid;
    }
    int psum = 0, done = 0;
    MB_ROLLED
    for (int j = end; j-- > start;) {
      int tmp = bits1[j] + ((lo * bits2[j]) >> 6);
      if (tmp < thresh[j] && !done) tmp = tmp >= alloc_floor ? alloc_floor : 0;
      else done = 1;
      tmp = imin(tmp, cap[j]);
      pulses[j] = tmp;
      psum += tmp;
    }
    int coded = end;
    while (coded > start) {
      int j = coded - 1;
      int band_start = eband(start), band_end = eband(coded), band_prev = eband(j);
      if (j <= skip_start) { total += skip_rsv; break; }
      int left = total - psum;
      int denom = imax(band_end - band_start, 1);
      int per = (int)((uint32_t)imax(left, 0) / (uint32_t)denom);
      left -= denom * per;
      int rem = imax(left - (band_prev - band_start), 0);
      int band_bits = pulses[j] + per * (band_end - band_prev) + rem;
      if (band_bits >= imax(thresh[j], alloc_floor + (1 << BITRES))) {
        if (dec.bit_logp(1)) break;
        psum += 1 << BITRES;
        band_bits -= 1 << BITRES;
      }
      psum -= pulses[j] + intensity_rsv;
      if (intensity_rsv > 0) intensity_rsv = kLog2Frac[j - start];
      psum += intensity_rsv;
      if (band_bits >= alloc_floor) { psum += alloc_floor; pulses[j] = alloc_floor; }
      else pulses[j] = 0;
      coded--;
    }
    if (intensity_rsv > 0) intensity = start + (int)dec.uint_((uint32_t)(coded + 1 - start));
    if (intensity <= start) {
      total += dual_stereo_rsv;
      if (!(quirks & Q9_DUAL_STEREO_RSV)) dual_stereo_rsv = 0;
    }
    if (dual_stereo_rsv > 0) dual_stereo = dec.bit_logp(1);
    {
      int denom = imax(eband(coded) - eband(start), 1);
      int left = total - psum;
      int per = (int)((uint32_t)imax(left, 0) / (uint32_t)denom);
      left -= denom * per;
      MB_ROLLED
      for (int j = start; j < coded; j++) pulses[j] += per * (eband(j + 1) - eband(j));
      MB_ROLLED
      for (int j = start; j < coded; j++) {
        int add = imin(eband(j + 1) - eband(j), left);
        pulses[j] += add;
        left -= add;
      }
    }
    int bal = 0;
    MB_ROLLED
    for (int j = start; j < coded; j++) {
      int N0 = eband(j + 1) - eband(j), N = N0 << lm;
      int bit = pulses[j] + bal;
      int prio;
      if (N > 1) {
        int excess = imax(bit - cap[j], 0);
        int bj = bit - excess;
        int den = C * N + ((C == 2 && N > 2 && !dual_stereo && j < intensity) ? 1 : 0);
        int NClogN = den * (kLogN400[j] + logM);
        int offset = (NClogN >> 1) - den * 21;
        if (N == 2) offset += den << BITRES >> 2;
        if (bj + offset < den * 2 << BITRES) offset += NClogN >> 2;
        else if (bj + offset < den * 3 << BITRES) offset += NClogN >> 3;
        int e = imax(0, bj + offset + (den << (BITRES - 1)));
        e = (int)((uint32_t)e / (uint32_t)den) >> BITRES;
        int lim = (quirks & Q7_EBITS_BUST) ? (bj >> stereo) >> BITRES : bj >> BITRES;
        if (C * e > lim) e = bj >> stereo >> BITRES;
        e = imin(e, 8);
        prio = e * (den << BITRES) >= bj + offset;
        bj -= C * e << BITRES;
        if (excess > 0) {
          int extra_fine = imin(excess >> (stereo + BITRES), 8 - e);
          e += extra_fine;
          int extra_bits = extra_fine * C << BITRES;
          if (extra_bits >= excess - bal) prio = 1;
          bal = excess - extra_bits;
        } else bal = excess;
        pulses[j] = bj;
        fine_quant[j] = e;
      } else {
        int excess = imax(0, bit - (C << BITRES));
        pulses[j] = bit - excess;
        fine_quant[j] = 0;
        prio = 1;
        bal = excess;
      }
      fine_prio_mask |= (uint32_t)prio << j;
    }
    balance = bal;
    MB_ROLLED
    for (int j = coded; j < end; j++) {
      fine_quant[j] = pulses[j] >> stereo >> BITRES;
      pulses[j] = 0;
      fine_prio_mask |= (uint32_t)(fine_quant[j] < 1) << j;
    }
    coded_bands = coded;
  }
  // ---- fine energy bits (quant_bands.rs:1850-1880)
  MB_ROLLED
  for (int i = start; i < end; i++) {
    if (fine_quant[i] <= 0) continue;
    MB_ROLLED
    for (int c = 0; c < C; c++) {
      uint32_t q2 = dec.raw_bits((unsigned)fine_quant[i]);
      ew[c * NB_EBANDS + i] |= (q2 & 255u) << 16;
    }
  }
  MB_ROLLED
  for (int i = start; i < end; i++)
    MB_ROLLED
    for (int c = 0; c < C; c++) ew[c * NB_EBANDS + i] |= (uint32_t)imin(imax(fine_quant[i], 0), 15) << 24;

  // ---- band loop (bands.rs:2575-3082), integer side
  {
    const int B = is_transient ? M : 1;
    const int norm_offset = M * eband(start);
    const int norm_len = imax(M * eband(NB_EBANDS - 1) - norm_offset, 0);
    const int total_bits = frame_total_bits - anti_collapse_rsv;
    const int N = M * SHORT_MDCT;
    cx.intensity = intensity;
    int lowband_offset = -1, update_lowband = 1, dual = dual_stereo;
    MB_ROLLED
    for (int i = start; i < end; i++) {
      cx.band = i;
      int band_start = M * eband(i), n = M * eband(i + 1) - band_start;
      int last = i + 1 == end;
      int tf = (int)dec.tell_frac();
      if (i != start) balance -= tf;
      int remaining = total_bits - tf - 1;
      cx.remaining_bits = remaining;
      int b = 0;
      if (i < coded_bands) {
        int curr_balance = balance / imin(3, coded_bands - i);
        b = imax(0, imin(16383, imin(remaining + 1, pulses[i] + curr_balance)));
      }
      if ((band_start >= norm_offset + n || i == start + 1) && (update_lowband || lowband_offset < 0)) lowband_offset = i;
      int tf_change = tf_res[i];
      int eff = -1, fold_start = 0, fold_end = 0;
      if (lowband_offset >= 0 && (spread != 3 || B > 1 || tf_change < 0)) {
        eff = imax(0, M * eband(lowband_offset) - norm_offset - n);
        fold_start = lowband_offset;
        if (quirks & Q5_FOLD_START) {
          int threshold = eff + norm_offset + n;
          while (fold_start > 0) { fold_start--; if (M * eband(fold_start) <= threshold) break; }
        } else {
          while (M * eband(--fold_start) > eff + norm_offset) {}
        }
        fold_end = lowband_offset - 1;
        while (++fold_end < i && M * eband(fold_end) < eff + norm_offset + n) {}
      }
      int average = 0;
      if (dual && i == intensity) { dual = 0; average = 1; }
      // lowband window validity (bands.rs:2529-2573; Q6 keeps the release-build "overlap -> none" rule)
      int out_off = last ? -1 : band_start - norm_offset;
      int in_ok = eff >= 0, out_ok = out_off >= 0;
      if (quirks & Q6_HYBRID_LOWBAND) {
        if (in_ok && out_ok) {
          if (eff + n > norm_len || out_off + n > norm_len) in_ok = out_ok = 0;
          else if (!(eff + n <= out_off || out_off + n <= eff)) in_ok = out_ok = 0;
        } else if (in_ok) { if (eff + n > norm_len) in_ok = 0; }
        else if (out_ok) { if (out_off + n > norm_len) out_ok = 0; }
      }
      cx.w.put(S_BANDA + i, (uint32_t)(in_ok ? eff + 1 : 0) | ((uint32_t)fold_start << 11) | ((uint32_t)fold_end << 16) |
                                ((uint32_t)(tf_change + 4) << 21) | ((uint32_t)dual << 24) | ((uint32_t)out_ok << 25) |
                                ((uint32_t)average << 26) | ((uint32_t)(eff >= 0) << 27));
      cx.w.put(S_BANDB + i, (uint32_t)pulses[i]);
      if (dual) {
        parse_band(cx, band_start, n, b / 2, B, lm, 1.0f, tf_change);
        parse_band(cx, N + band_start, n, b / 2, B, lm, 1.0f, tf_change);
      } else if (C == 2) {
        parse_band_stereo(cx, band_start, N + band_start, n, b, B, lm, tf_change);
      } else {
        parse_band(cx, band_start, n, b, B, lm, 1.0f, tf_change);
      }
      balance += pulses[i] + tf;
      update_lowband = b > (n << BITRES);
    }
  }
  int anti_collapse_on = 0;
  if (anti_collapse_rsv > 0) anti_collapse_on = (int)dec.raw_bits(1);
  // ---- unquant_energy_finalise bits (quant_bands.rs:1922-1970)
  {
    int bits_left = len_bits - dec.tell();
    MB_ROLLED
    for (int prio = 0; prio < 2; prio++)
      MB_ROLLED
      for (int i = start; i < end; i++) {
        if (bits_left < C) break;
        if (fine_quant[i] >= 8 || (int)((fine_prio_mask >> i) & 1u) != prio) continue;
        MB_ROLLED
        for (int c = 0; c < C; c++) {
          uint32_t q2 = dec.raw_bits(1);
          ew[c * NB_EBANDS + i] |= (1u << 28) | (q2 << 29);
          bits_left--;
        }
      }
  }
  int code = MB_OK;
  if (dec.tell() > 8 * len) code = MB_INVALID_PACKET;
  if (cx.w.overflow) code = MB_INTERNAL_ERROR;
  MB_ROLLED
  for (int i = 0; i < 2 * NB_EBANDS; i++) cx.w.put(S_ENERGY + i, ew[i]);
  cx.w.put(S_HDR0, (uint32_t)lm | ((uint32_t)(C - 1) << 2) | ((uint32_t)silence << 3) | ((uint32_t)is_transient << 4) |
                       ((uint32_t)intra << 5) | ((uint32_t)spread << 6) | ((uint32_t)anti_collapse_on << 8) |
                       ((uint32_t)dual_stereo << 9) | ((uint32_t)intensity << 10) | ((uint32_t)coded_bands << 15) |
                       ((uint32_t)pf_tapset << 20) | ((uint32_t)pf_qg1 << 22) | ((uint32_t)start << 26));
  cx.w.put(S_HDR1, (uint32_t)pf_pitch | ((uint32_t)imin(cx.w.n_nodes, CAP_NODES) << 11) | ((uint32_t)imin(cx.w.n_leaves, CAP_LEAVES) << 22));
  cx.w.put(S_RANGE, dec.rng);
  uint32_t status = (uint32_t)(uint8_t)(int8_t)code | ((uint32_t)end << 8) | ((uint32_t)(dec.error != 0) << 13) | (1u << 14) | status_extra;
  cx.w.put(S_STATUS, status);
  return status;
}

}  // namespace mb
