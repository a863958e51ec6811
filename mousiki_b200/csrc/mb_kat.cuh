// Unit known-answer hooks over the kernel bodies (debug entry bd_debug_kat; tests/test_unit_kat.py). The reference's unit
// tests for this path (src/celt/bands.rs:4805-4915, cwrs.rs:681-786, celt.rs:868-1190, kiss_fft.rs:1364-1440,
// mdct.rs:1746-1843, entdec.rs:327-377, range.rs:659-695, rate.rs:1424-1465) are replayed on the very functions the
// kernels call: one warp (one thread in the host emulation) runs one case. Same word protocol as the oracle's orc_kat
// (oracle/celt_oracle.cpp); floats travel as bit patterns.
#pragma once
#include "mb_celt_imdct.cuh"
#include "mb_celt_tail.cuh"

namespace mb {

union KatSmem {
  CombSmem comb;
  ImdctSmem imdct;
  float imdct_row[IMDCT_ROW];
};

// `gbytes`: 64 bytes the range decoder can read through its byte path (global memory on the device).
// `T`: the packed IMDCT table block (shared or global memory on the device).
MB_DEV int kat_run(Team tm, KatSmem* sm, int what, const uint32_t* in, uint32_t n_in, uint32_t* out, uint32_t n_out, uint8_t* gbytes,
                   const ImdctTables* T) {
  switch (what) {
    case 1:
      for (uint32_t i = tm.tid; i < n_in && i < n_out; i += tm.n) out[i] = (uint32_t)bitexact_cos((int)in[i]);
      return 0;
    case 2:
      for (uint32_t i = tm.tid; 2 * i + 1 < n_in && i < n_out; i += tm.n) out[i] = (uint32_t)bitexact_log2tan((int)in[2 * i], (int)in[2 * i + 1]);
      return 0;
    case 3: {   // pvq_decode_leaf, a lane per index
      const int n = (int)in[0], k = (int)in[1];
      const uint32_t count = in[2], inc = in[3], i0 = in[4];
      for (uint32_t j = tm.tid; j < count && (j + 1) * n <= n_out; j += tm.n) {
        float y[176];
        pvq_decode_leaf(n, k, i0 + j * inc, y);
        for (int t = 0; t < n; t++) out[j * n + t] = (uint32_t)(int32_t)y[t];
      }
      return 0;
    }
    case 4: {   // team_comb_filter on [hist | n] staged like k_comb stages a channel
      const int n = (int)in[0], T0 = (int)in[1], T1 = (int)in[2], tap0 = (int)in[3], tap1 = (int)in[4], hist = (int)in[7];
      if (hist > HIST || n > MAX_N) return -1;
      float* x = sm->comb.buf + HIST;
      for (int i = tm.tid; i < hist + n; i += tm.n) x[i - hist] = bits2f(in[8 + i]);
      MB_SYNC();
      team_comb_filter(tm, x, n, T0, T1, bits2f(in[5]), bits2f(in[6]), tap0, tap1);
      MB_SYNC();
      for (int i = tm.tid; i < n && i < (int)n_out; i += tm.n) out[i] = f2bits(x[i]);
      return 0;
    }
    case 5: {   // team_fft as KissFftState::ifft uses it: bit-reversed scatter of the conjugate, transform, conjugate
      const int shift = (int)in[0], nfft = 480 >> shift;
      cpx* F = (cpx*)sm->imdct.F;
      const int16_t* br = shift == 0 ? kFftBitrev480 : shift == 1 ? kFftBitrev240 : shift == 2 ? kFftBitrev120 : kFftBitrev60;
      for (int i = tm.tid; i < nfft; i += tm.n) { cpx v; v.r = bits2f(in[1 + 2 * i]); v.i = -bits2f(in[2 + 2 * i]); F[br[i]] = v; }
      MB_SYNC();
      team_fft(tm, F, shift, 1);
      for (int i = tm.tid; i < nfft; i += tm.n) { out[2 * i] = f2bits(F[i].r); out[2 * i + 1] = f2bits(-F[i].i); }
      return 0;
    }
    case 6: {   // celt_imdct_core
      const int lm = (int)in[0], transient = (int)in[1], N = SHORT_MDCT << lm;
      for (int i = tm.tid; i < N; i += tm.n) sm->imdct.io[i] = bits2f(in[2 + i]);
      MB_SYNC();
      celt_imdct_core(tm, &sm->imdct, lm, transient);
      MB_SYNC();
      for (int i = tm.tid; i < N && i < (int)n_out; i += tm.n) out[i] = f2bits(sm->imdct.io[i]);
      for (int i = tm.tid; i < 60 && N + i < (int)n_out; i += tm.n) out[N + i] = f2bits(sm->imdct.tail[i]);
      return 0;
    }
    case 7: {   // RangeDec script, one lane
      if (tm.tid != 0) return 0;
      const uint32_t len = in[0];
      uint8_t* bytes = gbytes;
      if (len > 64) return -1;
      for (uint32_t i = 0; i < len; i++) bytes[i] = (uint8_t)in[1 + i];
      RangeDec d;
      d.open(bytes, len);
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
          case 2: out[o] = d.uint_(a); break;
          case 3: out[o] = d.raw_bits(a); break;
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
    case 9: {   // celt_imdct_core_w, the warp-resident form k_imdct runs (same words as case 6)
      const int lm = (int)in[0], transient = (int)in[1], N = SHORT_MDCT << lm;
      for (int i = tm.tid; i < N; i += tm.n) sm->imdct_row[i] = bits2f(in[2 + i]);
      MB_SYNC();
      celt_imdct_core_w(tm, sm->imdct_row, T, lm, transient);
      MB_SYNC();
      for (int i = tm.tid; i < N + 60 && i < (int)n_out; i += tm.n) out[i] = f2bits(sm->imdct_row[i]);
      return 0;
    }
    case 8:
      for (uint32_t i = tm.tid; 3 * i + 2 < n_in && 3 * i + 2 < n_out; i += tm.n) {
        const int band = (int)in[3 * i], lm = (int)in[3 * i + 1], pulses = (int)in[3 * i + 2];
        const int row = cache_row(band, lm);
        const int maxp = row >= 0 ? (int)kCacheBits50[row] : 0;
        const int bits = pulses > maxp ? 0 : pulses2bits(row, pulses);
        out[3 * i] = (uint32_t)bits;
        out[3 * i + 1] = (uint32_t)bits2pulses(row, bits);
        out[3 * i + 2] = (uint32_t)maxp;
      }
      return 0;
  }
  return -1;
}

}  // namespace mb
