// Range decoder for one lane (one Opus frame), registers only.
// Replaces: EcDec in reference src/celt/entdec.rs:22-237 and ec_tell/ec_tell_frac in
// src/celt/entcode.rs:152-198. The arithmetic is fixed by the Opus bitstream (RFC 6716 §4.1);
// the byte source is a global-memory pointer into the packet arena (read through the
// read-only path), front bytes for the range coder and tail bytes for the raw bits.
#pragma once
#include "mb_common.cuh"

namespace mb {

struct RangeDec {
  const uint8_t* buf;
  uint32_t storage;      // bytes in the frame
  uint32_t offs;         // next front byte
  uint32_t end_offs;     // bytes consumed from the tail
  uint32_t end_window;   // raw-bit window
  int nend_bits;
  int nbits_total;
  uint32_t rng, val, ext;
  uint32_t pre;          // front byte at `offs`, fetched one step ahead (0 past the end)
  int rem;
  int error;

  MB_MEM uint32_t byte_at(uint32_t i) const {
#if MB_ON_DEVICE
    return __ldg(buf + i);
#else
    return buf[i];
#endif
  }
  MB_MEM void renorm() {
    while (rng <= 0x800000u) {
      nbits_total += 8;
      rng <<= 8;
      uint32_t prev = (uint32_t)rem;
      rem = (int)pre;                      // the byte was fetched when the previous one was consumed
      if (offs < storage) offs++;
      pre = offs < storage ? byte_at(offs) : 0u;
      uint32_t sym = ((prev << 8) | (uint32_t)rem) >> 1;
      val = ((val << 8) + (255u & ~sym)) & 0x7fffffffu;
    }
  }
  MB_MEM void open(const uint8_t* b, uint32_t len) {
    buf = b; storage = len; offs = 0; end_offs = 0; end_window = 0; nend_bits = 0;
    nbits_total = 9;  // EC_CODE_BITS + 1 - ((EC_CODE_BITS - EC_CODE_EXTRA) / EC_SYM_BITS) * EC_SYM_BITS
    rng = 128u; ext = 0; error = 0;
    rem = offs < storage ? (int)byte_at(offs++) : 0;
    pre = offs < storage ? byte_at(offs) : 0u;
    val = rng - 1u - ((uint32_t)rem >> 1);
    renorm();
  }
  // the frame ends n bytes earlier than announced (a redundancy frame sits behind it, src/opus_decoder.rs:776-781)
  MB_MEM void shrink(uint32_t n) {
    storage -= n;
    if (offs >= storage) pre = 0u;
  }
  MB_MEM uint32_t decode(uint32_t ft) {
    ext = rng / ft;
    uint32_t s = val / ext;
    return ft - (s + 1u < ft ? s + 1u : ft);
  }
  MB_MEM uint32_t decode_bin(unsigned bits) {
    ext = rng >> bits;
    uint32_t s = val / ext;
    uint32_t top = 1u << bits;
    return top - (s + 1u < top ? s + 1u : top);
  }
  MB_MEM void update(uint32_t fl, uint32_t fh, uint32_t ft) {
    uint32_t s = ext * (ft - fh);
    val -= s;
    rng = fl > 0 ? ext * (fh - fl) : rng - s;
    renorm();
  }
  MB_MEM int bit_logp(unsigned logp) {
    uint32_t s = rng >> logp;
    int hit = val < s;
    if (!hit) { val -= s; rng -= s; } else rng = s;
    renorm();
    return hit;
  }
  // inverse-CDF symbol; `tab` lives in device global memory (or host memory in the emulation)
  MB_MEM int icdf(const uint8_t* tab, unsigned ftb) {
    uint32_t s = rng, d = val, r = s >> ftb, t;
    int ret = -1;
    do { t = s; s = r * tab[++ret]; } while (d < s);
    val = d - s; rng = t - s;
    renorm();
    return ret;
  }
  MB_MEM uint32_t raw_bits(unsigned n) {
    uint32_t window = end_window;
    int available = nend_bits;
    if ((unsigned)available < n) {
      do {
        uint32_t b = end_offs < storage ? byte_at(storage - ++end_offs) : 0u;
        window |= b << available;
        available += 8;
      } while (available <= 24);
    }
    uint32_t ret = window & ((1u << n) - 1u);
    end_window = window >> n;
    nend_bits = available - (int)n;
    nbits_total += (int)n;
    return ret;
  }
  MB_MEM uint32_t uint_(uint32_t ft) {
    ft--;
    int ftb = ilog32(ft);
    if (ftb > 8) {
      ftb -= 8;
      uint32_t ft1 = (ft >> ftb) + 1u;
      uint32_t s = decode(ft1);
      update(s, s + 1u, ft1);
      uint32_t t = (s << ftb) | raw_bits((unsigned)ftb);
      if (t <= ft) return t;
      error = 1;
      return ft;
    }
    ft++;
    uint32_t s = decode(ft);
    update(s, s + 1u, ft);
    return s;
  }
  MB_MEM int tell() const { return nbits_total - ilog32(rng); }
  MB_MEM uint32_t tell_frac() const {
    // correction thresholds of ec_tell_frac (src/celt/entcode.rs:183-198)
    uint32_t nbits = (uint32_t)nbits_total << BITRES;
    int l = ilog32(rng);
    uint32_t r = rng >> (l - 16);
    uint32_t b = (r >> 12) - 8u;
    uint32_t thr = b == 0 ? 35733u : b == 1 ? 38967u : b == 2 ? 42495u : b == 3 ? 46340u
                 : b == 4 ? 50535u : b == 5 ? 55109u : b == 6 ? 60097u : 65535u;
    b += r > thr;
    return nbits - (uint32_t)((l << 3) + (int)b);
  }
};

}  // namespace mb
