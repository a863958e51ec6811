// mousiki_b200 — shared definitions for the batched CELT decode kernels (sm_100a).
//
// The kernel bodies in mb_*.cuh are written as plain functions over an explicit
// (thread id, team size) pair so that the very same source compiles
//   * with nvcc as __device__ code called from the __global__ wrappers in mb_kernels.cu, and
//   * with g++ (-DMB_EMUL) as a single-threaded host emulation used by the CPU test-suite
//     (tests/emul/) to check the algorithm against the oracle where no GPU exists.
// The emulation is test infrastructure only; the product library never runs it.
#pragma once
#include <stdint.h>
#include <string.h>
#include <math.h>

#if defined(__CUDACC__) && !defined(MB_EMUL)
#define MB_DEV __device__ __forceinline__
#define MB_MEM __device__ __forceinline__
#define MB_DEV_NOINLINE __device__ __noinline__
#define MB_TBL static __device__ const
#define MB_SYNC() __syncwarp()   /* warp-sized thread teams (mb_celt_bands.cuh, mb_celt_imdct.cuh) */
#define MB_CTA_SYNC() __syncthreads()
#define MB_ON_DEVICE 1
#define MB_ROLLED _Pragma("unroll 1")   /* keep a loop rolled: the hot kernels are bound by instruction fetch, not by loop overhead */
#else
#define MB_DEV static inline
#define MB_MEM inline
#define MB_DEV_NOINLINE static
#define MB_TBL static const
#define MB_SYNC() ((void)0)
#define MB_CTA_SYNC() ((void)0)
#define MB_ON_DEVICE 0
#define MB_ROLLED
#endif

#include "celt_tables_data.h"

namespace mb {

// ---- fixed CELT mode parameters (48 kHz / 960 static mode; reference src/celt/modes.rs:84-105)
enum {
  NB_EBANDS = 21, OVERLAP = 120, SHORT_MDCT = 120, MAX_LM = 3, MAX_N = 960,
  MAX_PERIOD = 1024, COMB_MINPERIOD = 15, BITRES = 3,
  HIST = 1024,  // post-filtered history the comb filter needs (it reaches back T+2 <= 1024): the shared-memory window of k_comb
  HIST_G = 2048 + 120,  // persistent history per stream-channel in HBM, laid out like the reference's decode_mem: the last 2048
                        // output samples, then the 120-sample overlap region (what the pitch concealment reads)
  SER_PREFIX = 2176,    // room in front of the first frame of a lane's time series (k_comb_plc lays the history there)
  SER_SLACK = 128,      // room behind the last frame (the overlap region of the lane's last frame)
};

// ---- quirk switches: bit set = mousiki's float decoder behaviour, clear = libopus
enum : uint32_t {
  Q1_DYNALLOC_BUDGET = 1u << 0, Q2_MDCT_ZERO_FILL = 1u << 1, Q5_FOLD_START = 1u << 2, Q6_HYBRID_LOWBAND = 1u << 3,
  Q7_EBITS_BUST = 1u << 4, Q8_COARSE_CLAMP = 1u << 5, Q9_DUAL_STEREO_RSV = 1u << 6, QUIRKS_MOUSIKI = 0x7f,
  // not a quirk: bits 24-27 of the word the kernels receive carry the decoder's down-sampling factor 48000 / Fs (0 or 1: none),
  // set by bd_create from bd_config.sample_rate_hz (CELT `downsample`, src/celt/celt_decoder.rs:584-776, 2232-2297)
  QX_DS_SHIFT = 24,
};

// ---- Opus error codes (reference src/opus_decoder.rs:275-286)
enum { MB_OK = 0, MB_BAD_ARG = -1, MB_BUFFER_TOO_SMALL = -2, MB_INTERNAL_ERROR = -3, MB_INVALID_PACKET = -4, MB_UNIMPLEMENTED = -5 };

// ---- phase-1 -> phase-2 scratch ("records"): 32-bit words, structure-of-arrays in tiles of 32 frames:
//      word(slot, f) = rec[((f >> 5) * slots + slot) * 32 + (f & 31)]
// so the 32 lanes of a parse warp store 128 contiguous bytes per slot, and a reconstruction block
// fetches one 32-byte sector per slot for 8 neighbouring frames.
enum {
  S_HDR0 = 0, S_HDR1 = 1, S_RANGE = 2, S_STATUS = 3,
  S_ENERGY = 4,                 // 42 words, index c*21+band
  S_BANDA = S_ENERGY + 42,      // 21 words
  S_BANDB = S_BANDA + 21,       // 21 words
  S_NODES = S_BANDB + 21,       // CAP_NODES words
  CAP_NODES = 1344,             // 21 bands x (1 stereo split + 2 x (15 splits + 16 leaves)) rounded up
  CAP_LEAVES = 704,             // 21 bands x 2 x 16 rounded up
  S_LEAVES = S_NODES + CAP_NODES,  // 3 words per leaf: index, gain bits (collapse mask after the PVQ pass), packed geometry
  S_PLAN = S_LEAVES + 3 * CAP_LEAVES,  // 1 word: operation words written by the band plan (0xffffffff: overflow)
  S_OPS = S_PLAN + 1,              // CAP_OPS words: the band plan's operation list (mb_celt_plan.cuh)
  CAP_OPS = 2048,
  S_SILK_HYB = S_OPS + CAP_OPS,    // hybrid cells: SILK header word, then one SILK frame record (mb_silk_parse.cuh)
  REC_SLOTS = S_SILK_HYB + 344,
  FOLD_WORDS = 1248                // fold buffer per stream: the reference's norm_storage, 2 channels x 8 * eband(20) floats
};

// Auxiliary cells: a few packets decode TWO CELT frames (a hybrid packet with a redundancy frame, a mode switch that first
// conceals one frame in the old mode); the second one lives in an auxiliary cell, index aux.base + i, handed out from a per-pass
// pool behind the main cells by whoever finds the need (k_parse, k_energy).
enum { S_AUXREF = S_SILK_HYB + 343 };   // SILK-side cell whose redundancy frame sits in an auxiliary cell: (cell << 3) | own silence frame << 2 | celt_to_silk << 1 | 1; else 0
enum { S_AUXSIL = S_SILK_HYB + 341 };   // 40 / 60 ms SILK cells only (free in their layout): (cell << 1) | 1 = auxiliary cell with the fade-out silence frame
struct AuxPool { uint32_t base, cap; uint32_t* count; };
MB_DEV uint32_t aux_alloc(const AuxPool& A) {   // cell index, or 0xffffffff when the pool is spent
  if (!A.count || A.cap == 0) return 0xffffffffu;
#if MB_ON_DEVICE
  const uint32_t i = atomicAdd(A.count, 1u);
#else
  const uint32_t i = (*A.count)++;
#endif
  return i < A.cap ? A.base + i : 0xffffffffu;
}

// First PCM element of packet p: row p of the caller's buffer, or (calls cut into frames by mb_host_packets.h) row
// out_map[2p] at element out_map[2p+1].
MB_DEV int quirk_ds(uint32_t quirks) { const int d = (int)((quirks >> 24) & 15u); return d > 1 ? d : 1; }
MB_DEV size_t pcm_elem_off(const uint32_t* out_map, uint32_t p, uint32_t stride) {
  return out_map ? (size_t)out_map[2 * p] * stride + out_map[2 * p + 1] : (size_t)p * stride;
}

// S_HDR0 fields
#define MB_H0_LM(h) ((h) & 3u)
#define MB_H0_C(h) ((((h) >> 2) & 1u) + 1u)
#define MB_H0_SILENCE(h) (((h) >> 3) & 1u)
#define MB_H0_TRANSIENT(h) (((h) >> 4) & 1u)
#define MB_H0_INTRA(h) (((h) >> 5) & 1u)
#define MB_H0_SPREAD(h) (((h) >> 6) & 3u)
#define MB_H0_ACON(h) (((h) >> 8) & 1u)
#define MB_H0_DUAL(h) (((h) >> 9) & 1u)
#define MB_H0_INTENSITY(h) (((h) >> 10) & 31u)
#define MB_H0_CODED(h) (((h) >> 15) & 31u)
#define MB_H0_TAPSET(h) (((h) >> 20) & 3u)
#define MB_H0_QG1(h) (((h) >> 22) & 15u)   // 0 = post-filter off, else qg+1
#define MB_H0_START(h) (((h) >> 26) & 31u)
// S_HDR1 fields
#define MB_H1_PITCH(h) ((h) & 2047u)
#define MB_H1_NNODES(h) (((h) >> 11) & 2047u)
#define MB_H1_NLEAVES(h) (((h) >> 22) & 1023u)
// S_STATUS: bits 0..7 signed return code (0 ok), 8..12 end band, 13 range-decoder error flag, 14 frame decoded, 15 empty cell
#define MB_ST_CODE(s) ((int)(int8_t)((s) & 255u))
#define MB_ST_END(s) (((s) >> 8) & 31u)
#define MB_ST_DECODED(s) (((s) >> 14) & 1u)
#define MB_ST_EMPTY(s) (((s) >> 15) & 1u)   // grid cell without a packet

// node stream word: type[0..2] order[3] inv[4] sign[5..6] itheta[7..21]
enum { ND_LEAF_EMPTY = 0, ND_LEAF_PULSE = 1, ND_SPLIT = 2, ND_STEREO = 3, ND_N1 = 4 };
#define MB_ND_TYPE(w) ((w) & 7u)
#define MB_ND_ORDER(w) (((w) >> 3) & 1u)
#define MB_ND_INV(w) (((w) >> 4) & 1u)
#define MB_ND_SIGN(w) (((w) >> 5) & 3u)
#define MB_ND_ITHETA(w) (((w) >> 7) & 32767u)
// leaf geometry word: x_off[0..10] n[11..18] k[19..26] B[27..31]
#define MB_LF_XOFF(w) ((w) & 2047u)
#define MB_LF_N(w) (((w) >> 11) & 255u)
#define MB_LF_K(w) (((w) >> 19) & 255u)
#define MB_LF_B(w) (((w) >> 27) & 31u)
// band word A: in+1[0..10] fold_start[11..15] fold_end[16..20] tf+4[21..23] dual[24] out_valid[25] average[26] has_fold[27]
#define MB_BA_IN1(w) ((w) & 2047u)
#define MB_BA_FS(w) (((w) >> 11) & 31u)
#define MB_BA_FE(w) (((w) >> 16) & 31u)
#define MB_BA_TF(w) ((int)(((w) >> 21) & 7u) - 4)
#define MB_BA_DUAL(w) (((w) >> 24) & 1u)
#define MB_BA_OUT(w) (((w) >> 25) & 1u)
#define MB_BA_AVG(w) (((w) >> 26) & 1u)
#define MB_BA_HASFOLD(w) (((w) >> 27) & 1u)

MB_DEV size_t rec_index(uint32_t f, uint32_t slot) { return ((size_t)(f >> 5) * REC_SLOTS + slot) * 32u + (f & 31u); }

// ---- per-frame descriptor produced by the framing kernel (TOC parse; reference src/packet.rs:30-236)
struct FrameDesc {
  uint32_t byte_off;   // first payload byte of the frame inside the packet arena
  uint16_t len;        // payload bytes
  uint8_t cfg;         // lm[0..1] stream_channels-1[2] mode: 0 celt, 1 silk, 2 hybrid [3..4]
  uint8_t end_band;    // 13/17/19/21 by bandwidth
};

// ---- small helpers
MB_DEV int imin(int a, int b) { return a < b ? a : b; }
MB_DEV int imax(int a, int b) { return a > b ? a : b; }
MB_DEV float fmin32(float a, float b) { return a < b ? a : b; }
MB_DEV float fmax32(float a, float b) { return a > b ? a : b; }
MB_DEV int ilog32(uint32_t v) {
#if MB_ON_DEVICE
  return 32 - __clz((int)v);
#else
  return v ? 32 - __builtin_clz(v) : 0;
#endif
}
MB_DEV float bits2f(uint32_t b) {
#if MB_ON_DEVICE
  return __uint_as_float(b);
#else
  float f; memcpy(&f, &b, 4); return f;
#endif
}
MB_DEV uint32_t f2bits(float f) {
#if MB_ON_DEVICE
  return __float_as_uint(f);
#else
  uint32_t b; memcpy(&b, &f, 4); return b;
#endif
}
// The reference (Rust) never contracts a*b+c; fused multiply-adds appear only where it calls fmaf
// (src/celt/kiss_fft.rs:537-542, src/celt/celt.rs:164-166,414-419). Device code is built with
// -fmad=false and uses this helper at exactly those sites.
MB_DEV float fma32(float a, float b, float c) {
#if MB_ON_DEVICE
  return __fmaf_rn(a, b, c);
#else
  return fmaf(a, b, c);
#endif
}
// f64-evaluate-then-round helpers (reference src/celt/math.rs:259-334 via crate libm)
MB_DEV float celt_exp2(float x) { return (float)exp(0.6931471805599453094 * (double)x); }
// (float)sqrt((double)x) == correctly rounded f32 sqrt (53 >= 2*24+2 bits: double rounding is innocuous for sqrt)
MB_DEV float celt_sqrt(float x) {
#if MB_ON_DEVICE
  return __fsqrt_rn(x);
#else
  return (float)sqrt((double)x);
#endif
}
MB_DEV float celt_rsqrt(float x) {
#if MB_ON_DEVICE
  return __fdiv_rn(1.0f, celt_sqrt(x));
#else
  return 1.0f / celt_sqrt(x);
#endif
}
MB_DEV float celt_cos_norm(float x) { return (float)cos((double)(0.5f * 3.14159265358979323846f * x)); }
// 128-bit row moves (16-byte aligned addresses); the host emulation copies four floats
struct float4v { float x, y, z, w; };
MB_DEV float4v ld_f4(const float* p) {
#if MB_ON_DEVICE
  const float4 t = *(const float4*)p;
  float4v o = {t.x, t.y, t.z, t.w};
#else
  float4v o = {p[0], p[1], p[2], p[3]};
#endif
  return o;
}
MB_DEV void st_f4(float* p, float4v v) {
#if MB_ON_DEVICE
  *(float4*)p = make_float4(v.x, v.y, v.z, v.w);
#else
  p[0] = v.x; p[1] = v.y; p[2] = v.z; p[3] = v.w;
#endif
}
MB_DEV int eband(int i) { return kEBands5ms[i]; }
MB_DEV uint32_t mask_from_bits(int b) { return b <= 0 ? 0u : (b >= 32 ? 0xffffffffu : (1u << b) - 1u); }

}  // namespace mb
