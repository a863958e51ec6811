// ORACLE (test infrastructure only — never linked into the product library).
// CPU restatement of the reference's range coder:
//   decoder  src/celt/entdec.rs:29-230   (EcDec::new/normalize/decode/decode_bin/update/
//                                         dec_bit_logp/dec_icdf/dec_uint/dec_bits)
//   tell     src/celt/entcode.rs:152-198 (ec_ilog, ec_tell, ec_tell_frac)
//   encoder  src/celt/entenc.rs          (only used by the synthetic corpus writer, tools/)
// Parity pins: tests/test_opus_decode.rs:311-342 (cres), src/range.rs:659-695.
#pragma once
#include <stdint.h>
#include <string.h>

namespace orc {

enum { EC_SYM_BITS = 8, EC_CODE_BITS = 32, EC_SYM_MAX = 255, EC_CODE_SHIFT = 23,
       EC_CODE_EXTRA = 7, EC_UINT_BITS = 8, BITRES = 3 };
static const uint32_t EC_CODE_TOP = 1u << 31;
static const uint32_t EC_CODE_BOT = EC_CODE_TOP >> EC_SYM_BITS;

static inline int ec_ilog(uint32_t v) { return v ? 32 - __builtin_clz(v) : 0; }

#ifdef ORC_GEN
struct EcEnc;
struct EcDec;
// Corpus writer (tools/, bench.py --corpus struct; never part of the checker build): with ORC_GEN defined a decoder may be
// "driven": every symbol it asks for is CHOSEN (drawn from the symbol's own model, or forced by the site that asks) and written
// to a range encoder, whose rng / bit count the decoder mirrors — ec_tell is the same function of both, so the decoder's budget
// logic walks exactly the path a decoder reading the finished bytes will walk.
struct GenCtl {
  EcEnc* enc = nullptr;
  uint64_t s = 0x9e3779b97f4a7c15ull;
  int64_t forced = -1;            // the next draw, when a site wants a particular symbol
  uint32_t next() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (uint32_t)(s >> 16); }
  float unit() { return (float)(next() & 0xffffff) * (1.0f / 16777216.0f); }
  uint32_t draw(uint32_t ft) {
    uint32_t v = forced >= 0 ? (uint32_t)forced : next() % ft;
    forced = -1;
    return v < ft ? v : ft - 1;
  }
  // policy state of the CELT sites (celt_oracle.cpp)
  int frame = 0;
  float level = 0.f, tilt = 0.f, loud = 0.f, band_dev[2 * 21] = {0};
  float p_transient = 0.1f, p_postfilter = 0.3f;
};
#endif

struct EcDec {
#ifdef ORC_GEN
  GenCtl* gen = nullptr;
  inline void gen_sync();
  inline void gen_open(GenCtl* g, uint32_t len);
#endif
  const uint8_t* buf = nullptr;
  uint32_t storage = 0, end_offs = 0, end_window = 0;
  int nend_bits = 0, nbits_total = 0;
  uint32_t offs = 0, rng = 0, val = 0, ext = 0;
  int rem = 0, error = 0;

  int read_byte() { return offs < storage ? buf[offs++] : 0; }
  int read_byte_from_end() { return end_offs < storage ? buf[storage - ++end_offs] : 0; }

  void normalize() {
    while (rng <= EC_CODE_BOT) {
      nbits_total += EC_SYM_BITS;
      rng <<= EC_SYM_BITS;
      uint32_t sym = (uint32_t)rem;
      rem = read_byte();
      sym = ((sym << EC_SYM_BITS) | (uint32_t)rem) >> (EC_SYM_BITS - EC_CODE_EXTRA);
      val = ((val << EC_SYM_BITS) + (EC_SYM_MAX & ~sym)) & (EC_CODE_TOP - 1);
    }
  }
  void init(const uint8_t* b, uint32_t len) {
    buf = b; storage = len; end_offs = 0; end_window = 0; nend_bits = 0;
    nbits_total = EC_CODE_BITS + 1 - ((EC_CODE_BITS - EC_CODE_EXTRA) / EC_SYM_BITS) * EC_SYM_BITS;
    offs = 0; rng = 1u << EC_CODE_EXTRA; ext = 0; error = 0;
    rem = read_byte();
    val = rng - 1 - ((uint32_t)rem >> (EC_SYM_BITS - EC_CODE_EXTRA));
    normalize();
  }
  uint32_t decode(uint32_t ft) {
#ifdef ORC_GEN
    if (gen) return gen->draw(ft);
#endif
    ext = rng / ft;
    uint32_t s = val / ext;
    return ft - (s + 1 < ft ? s + 1 : ft);
  }
  uint32_t decode_bin(unsigned bits) {
#ifdef ORC_GEN
    if (gen) return gen->draw(1u << bits);
#endif
    ext = rng >> bits;
    uint32_t s = val / ext;
    uint32_t top = 1u << bits;
    return top - (s + 1 < top ? s + 1 : top);
  }
  void update(uint32_t fl, uint32_t fh, uint32_t ft) {
#ifdef ORC_GEN
    if (gen) { gen_update(fl, fh, ft); return; }
#endif
    uint32_t s = ext * (ft - fh);
    val -= s;
    rng = fl > 0 ? ext * (fh - fl) : rng - s;
    normalize();
  }
  int bit_logp(unsigned logp) {
#ifdef ORC_GEN
    if (gen) return gen_bit_logp(logp);
#endif
    uint32_t r = rng, d = val, s = r >> logp;
    int ret = d < s;
    if (!ret) val = d - s;
    rng = ret ? s : r - s;
    normalize();
    return ret;
  }
  int icdf(const uint8_t* tab, unsigned ftb) {
#ifdef ORC_GEN
    if (gen) return gen_icdf(tab, ftb);
#endif
    uint32_t s = rng, d = val, r = s >> ftb, t;
    int ret = -1;
    do { t = s; s = r * tab[++ret]; } while (d < s);
    val = d - s; rng = t - s;
    normalize();
    return ret;
  }
  uint32_t bits(unsigned n) {
#ifdef ORC_GEN
    if (gen) return gen_bits(n);
#endif
    uint32_t window = end_window; int available = nend_bits;
    if ((unsigned)available < n) {
      do { window |= (uint32_t)read_byte_from_end() << available; available += EC_SYM_BITS; }
      while (available <= 32 - EC_SYM_BITS);
    }
    uint32_t ret = window & ((1u << n) - 1u);
    window >>= n; available -= n;
    end_window = window; nend_bits = available; nbits_total += n;
    return ret;
  }
  uint32_t uint(uint32_t ft) {
#ifdef ORC_GEN
    if (gen) return gen_uint(ft);
#endif
    ft--;
    int ftb = ec_ilog(ft);
    if (ftb > EC_UINT_BITS) {
      ftb -= EC_UINT_BITS;
      uint32_t ft1 = (ft >> ftb) + 1;
      uint32_t s = decode(ft1);
      update(s, s + 1, ft1);
      uint32_t t = (s << ftb) | bits(ftb);
      if (t <= ft) return t;
      error = 1;
      return ft;
    }
    ft++;
    uint32_t s = decode(ft);
    update(s, s + 1, ft);
    return s;
  }
  int tell() const { return nbits_total - ec_ilog(rng); }
  uint32_t tell_frac() const {
    static const uint32_t corr[8] = {35733, 38967, 42495, 46340, 50535, 55109, 60097, 65535};
    uint32_t nbits = (uint32_t)nbits_total << BITRES;
    int l = ec_ilog(rng);
    uint32_t r = rng >> (l - 16);
    unsigned b = (r >> 12) - 8;
    b += r > corr[b];
    return nbits - (uint32_t)((l << 3) + (int)b);
  }
#ifdef ORC_GEN
  inline void gen_update(uint32_t fl, uint32_t fh, uint32_t ft);
  inline int gen_bit_logp(unsigned logp);
  inline int gen_icdf(const uint8_t* tab, unsigned ftb);
  inline uint32_t gen_bits(unsigned n);
  inline uint32_t gen_uint(uint32_t ft);
#endif
};

// Range ENCODER (libopus-format bitstream writer). Used only to synthesise
// structured test/bench corpora; mirrors the decoder arithmetic above.
struct EcEnc {
  uint8_t* buf = nullptr;
  uint32_t storage = 0, end_offs = 0, end_window = 0;
  int nend_bits = 0, nbits_total = 0;
  uint32_t offs = 0, rng = 0, val = 0, ext = 0;
  int rem = -1, error = 0;

  void init(uint8_t* b, uint32_t size) {
    buf = b; storage = size; end_offs = 0; end_window = 0; nend_bits = 0;
    nbits_total = EC_CODE_BITS + 1; offs = 0; rng = EC_CODE_TOP; rem = -1; val = 0; ext = 0; error = 0;
  }
  int write_byte(unsigned v) { if (offs + end_offs >= storage) return -1; buf[offs++] = (uint8_t)v; return 0; }
  int write_byte_at_end(unsigned v) { if (offs + end_offs >= storage) return -1; buf[storage - ++end_offs] = (uint8_t)v; return 0; }
  void carry_out(int c) {
    if (c != EC_SYM_MAX) {
      int carry = c >> EC_SYM_BITS;
      if (rem >= 0) error |= write_byte(rem + carry);
      if (ext > 0) {
        unsigned sym = (EC_SYM_MAX + carry) & EC_SYM_MAX;
        do error |= write_byte(sym); while (--ext > 0);
      }
      rem = c & EC_SYM_MAX;
    } else ext++;
  }
  void normalize() {
    while (rng <= EC_CODE_BOT) {
      carry_out((int)(val >> EC_CODE_SHIFT));
      val = (val << EC_SYM_BITS) & (EC_CODE_TOP - 1);
      rng <<= EC_SYM_BITS;
      nbits_total += EC_SYM_BITS;
    }
  }
  void encode(uint32_t fl, uint32_t fh, uint32_t ft) {
    uint32_t r = rng / ft;
    if (fl > 0) { val += rng - r * (ft - fl); rng = r * (fh - fl); } else rng -= r * (ft - fh);
    normalize();
  }
  void encode_bin(uint32_t fl, uint32_t fh, unsigned bits) {
    uint32_t r = rng >> bits;
    if (fl > 0) { val += rng - r * ((1u << bits) - fl); rng = r * (fh - fl); } else rng -= r * ((1u << bits) - fh);
    normalize();
  }
  void bit_logp(int v, unsigned logp) {
    uint32_t r = rng, l = val, s = r >> logp;
    r -= s;
    if (v) val = l + r;
    rng = v ? s : r;
    normalize();
  }
  void icdf(int s, const uint8_t* tab, unsigned ftb) {
    uint32_t r = rng >> ftb;
    if (s > 0) { val += rng - r * tab[s - 1]; rng = r * (tab[s - 1] - tab[s]); } else rng -= r * tab[s];
    normalize();
  }
  void bits(uint32_t fl, unsigned n) {
    uint32_t window = end_window; int used = nend_bits;
    if (used + (int)n > 32) {
      do { error |= write_byte_at_end(window & EC_SYM_MAX); window >>= EC_SYM_BITS; used -= EC_SYM_BITS; }
      while (used >= EC_SYM_BITS);
    }
    window |= fl << used; used += n;
    end_window = window; nend_bits = used; nbits_total += n;
  }
  void uint(uint32_t fl, uint32_t ft) {
    ft--;
    int ftb = ec_ilog(ft);
    if (ftb > EC_UINT_BITS) {
      ftb -= EC_UINT_BITS;
      uint32_t ft1 = (ft >> ftb) + 1, fl1 = fl >> ftb;
      encode(fl1, fl1 + 1, ft1);
      bits(fl & ((1u << ftb) - 1u), ftb);
    } else encode(fl, fl + 1, ft + 1);
  }
  int tell() const { return nbits_total - ec_ilog(rng); }
  uint32_t tell_frac() const {
    static const uint32_t corr[8] = {35733, 38967, 42495, 46340, 50535, 55109, 60097, 65535};
    uint32_t nbits = (uint32_t)nbits_total << BITRES;
    int l = ec_ilog(rng);
    uint32_t r = rng >> (l - 16);
    unsigned b = (r >> 12) - 8;
    b += r > corr[b];
    return nbits - (uint32_t)((l << 3) + (int)b);
  }
  void done() {
    int l = EC_CODE_BITS - ec_ilog(rng);
    uint32_t msk = (EC_CODE_TOP - 1) >> l;
    uint32_t end = (val + msk) & ~msk;
    if ((end | msk) >= val + rng) { l++; msk >>= 1; end = (val + msk) & ~msk; }
    while (l > 0) { carry_out((int)(end >> EC_CODE_SHIFT)); end = (end << EC_SYM_BITS) & (EC_CODE_TOP - 1); l -= EC_SYM_BITS; }
    if (rem >= 0 || ext > 0) carry_out(0);
    uint32_t window = end_window; int used = nend_bits;
    while (used >= EC_SYM_BITS) { error |= write_byte_at_end(window & EC_SYM_MAX); window >>= EC_SYM_BITS; used -= EC_SYM_BITS; }
    if (!error) {
      memset(buf + offs, 0, storage - offs - end_offs);
      if (used > 0) {
        if (end_offs >= storage) error = -1;
        else {
          l = -l;
          if (offs + end_offs >= storage && l < used) { window &= (1u << l) - 1; error = -1; }
          buf[storage - end_offs - 1] |= (uint8_t)window;
        }
      }
    }
  }
};

#ifdef ORC_GEN
inline void EcDec::gen_sync() { rng = gen->enc->rng; nbits_total = gen->enc->nbits_total; }
inline void EcDec::gen_open(GenCtl* g, uint32_t len) {
  gen = g; buf = nullptr; storage = len; error = 0; end_offs = 0; end_window = 0; nend_bits = 0; offs = 0; val = 0; ext = 0; rem = 0;
  gen_sync();
}
inline void EcDec::gen_update(uint32_t fl, uint32_t fh, uint32_t ft) { gen->enc->encode(fl, fh, ft); gen_sync(); }
inline int EcDec::gen_bit_logp(unsigned logp) {
  int v = gen->forced >= 0 ? (int)(gen->forced != 0) : (int)((gen->next() & ((1u << logp) - 1u)) == 0);
  gen->forced = -1;
  gen->enc->bit_logp(v, logp);
  gen_sync();
  return v;
}
inline int EcDec::gen_icdf(const uint8_t* tab, unsigned ftb) {
  int n = 0;
  while (tab[n] != 0) n++;
  n++;                                     // symbols 0 .. n-1, tab[n-1] == 0
  int sym;
  if (gen->forced >= 0) { sym = (int)gen->forced; gen->forced = -1; if (sym >= n) sym = n - 1; }
  else {                                   // drawn from the table's own distribution
    const uint32_t ft = 1u << ftb, u = gen->next() % ft;
    sym = 0;
    while (sym < n - 1 && ft - tab[sym] <= u) sym++;
  }
  // (a symbol of probability zero cannot be written: move to a neighbour that has some)
  while (sym > 0 && tab[sym - 1] == tab[sym]) sym--;
  if (sym == 0 && (1u << ftb) == tab[0]) { sym = 1; while (sym < n - 1 && tab[sym - 1] == tab[sym]) sym++; }
  gen->enc->icdf(sym, tab, ftb);
  gen_sync();
  return sym;
}
inline uint32_t EcDec::gen_bits(unsigned n) {
  if (n == 0) return 0;
  uint32_t v = gen->forced >= 0 ? (uint32_t)gen->forced : gen->next();
  gen->forced = -1;
  v &= n >= 32 ? 0xffffffffu : ((1u << n) - 1u);
  gen->enc->bits(v, n);
  nbits_total = gen->enc->nbits_total;
  return v;
}
inline uint32_t EcDec::gen_uint(uint32_t ft) {
  uint32_t t;
  if (gen->forced >= 0) { t = (uint32_t)gen->forced; gen->forced = -1; if (t >= ft) t = ft - 1; }
  else t = (uint32_t)((((uint64_t)gen->next() << 16) ^ gen->next()) % ft);
  gen->enc->uint(t, ft);
  gen_sync();
  return t;
}
#endif

}  // namespace orc
