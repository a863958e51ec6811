// mousiki_b200 — CUDA kernels (sm_100a) and the C ABI declared in include/mousiki_b200.h.
//
// Launch sequence of one decode pass over a [ticks x lanes] grid of packets:
//   k_parse    lane per packet          phase 1  (mb_frame.cuh / mb_celt_parse.cuh)
//   k_energy   lane per stream          phase 2a (mb_celt_energy.cuh)
//   k_pvq      CTA per frame, thread per PVQ leaf   phase 2b(i)   (mb_celt_pvq.cuh)
//   k_plan     lane per frame                       phase 2b(ii)  (mb_celt_plan.cuh)
//   k_bands    warp per frame                       phase 2b(iii) (mb_celt_bands.cuh)
//   k_imdct    warp per frame-channel               phase 2b(iv)  (mb_celt_imdct.cuh)
//   k_comb     warp per stream-channel              phase 2c(i)   (mb_celt_tail.cuh)
//   k_deemph   lane per stream-channel, smem tiles  phase 2c(ii)
//   k_clip_i16 warp per stream                          soft clip + float -> int16 (i16 output path only)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -lineinfo (see __graft_entry__.build()).
#include <cuda_runtime.h>
#include <chrono>
#include <functional>
#include <stdio.h>
#include <stdlib.h>

#include <new>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/mousiki_b200.h"
#include "mb_host_packets.h"
#include "mb_celt_imdct.cuh"
#include "mb_celt_tail.cuh"
#include "mb_celt_plc.cuh"
#include "mb_frame.cuh"
#include "mb_silk_out.cuh"
#include "mb_kat.cuh"
#include "mb_lite.cuh"

using namespace mb;

// ------------------------------------------------------------------------------------------ kernels
#ifndef MB_PLAN_MIN_CTAS
#define MB_PLAN_MIN_CTAS 8   // (10: 4.58 ms, 12: 5.51 ms against 4.16 ms per step)
#endif
// Two register budgets of the same parser: 10 CTAs per SM (51 registers) when the launch spans several waves — 8.11 against 8.76 ms
// per step on the headline call, C3 843 k -> 884 k audio-s/s — and 7 CTAs per SM (73 registers) when all of it fits in one wave,
// where the extra residents only slow each other down (the 13-tick passes of the host-buffer call: 9.9 against 10.3 ms).
template <int MIN_CTAS>
__global__ void __launch_bounds__(128, MIN_CTAS) k_parse(const uint8_t* __restrict__ arena, const uint32_t* __restrict__ offsets,
                                               const uint32_t* __restrict__ pkt_map, uint32_t n_cells, uint32_t* __restrict__ rec, int CC,
                                               uint32_t quirks, int disable_inv, int cap_per_ch, const uint8_t* __restrict__ flags,
                                               uint32_t n_lanes, const uint32_t* __restrict__ lane_stream, const StreamState* st,
                                               AuxPool aux) {
  uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_cells) return;
  uint32_t pkt = pkt_map ? pkt_map[f] : f;
  uint32_t prev_pkt = 0xffffffffu;
  const int32_t* prev_mode = nullptr;
  if (f >= n_lanes) prev_pkt = pkt_map ? pkt_map[f - n_lanes] : f - n_lanes;
  else prev_mode = &st[lane_stream ? lane_stream[f] : f].prev_mode;
  parse_cell(arena, offsets, pkt, rec + rec_index(f, 0), CC, quirks, disable_inv, cap_per_ch, flags, prev_pkt, prev_mode, rec, aux);
}

// one warp per stream, 4 streams per CTA
constexpr int ENERGY_WARPS = 4;
__global__ void __launch_bounds__(32 * ENERGY_WARPS) k_energy(EnergyParams Q) {
  __shared__ EnergySmem sm[ENERGY_WARPS];
  const uint32_t lane = blockIdx.x * ENERGY_WARPS + (threadIdx.x >> 5);
  if (lane >= Q.n_lanes) return;
  Team tm = {(int)(threadIdx.x & 31), 32};
  celt_energy_stream(tm, &sm[threadIdx.x >> 5], lane, Q);
}

// The auxiliary cells of the pass (mb_common.cuh): unused ones must look empty to every per-cell kernel.
__global__ void k_aux_clear(uint32_t* rec, FrameMeta* meta, AuxPool A) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) *A.count = 0;
  if (i >= A.cap) return;
  rec[rec_index(A.base + i, S_STATUS)] = 1u << 15;
  meta[A.base + i].flags = 0;
}
// Cell of the u-th thread / warp of a per-cell launch: the pass's main cells, then the auxiliary pool.
__device__ __forceinline__ uint32_t launch_cell(uint32_t u, uint32_t n_main, uint32_t aux_base) { return u < n_main ? u : aux_base + (u - n_main); }

#ifndef MB_PVQ_THREADS
#define MB_PVQ_THREADS 64
#endif
constexpr int PVQ_THREADS = MB_PVQ_THREADS;
// One CTA per PVQ_CELLS consecutive cells: the first lanes read their status words (one 32-byte sector) and the CTA then
// walks the cells that hold a CELT part. A group of SILK-only or empty cells costs one load, not PVQ_CELLS CTA launches
// (the kernel runs beside k_silk, whose long-lived CTAs should not have to compete with a million empty ones); a longer
// walk per CTA costs CELT throughput (per 5e5 frames, beside k_energy: 32 cells 9.7 ms, 8 cells 8.3 ms, 1 cell 7.9 ms).
#ifndef MB_PVQ_CELLS
#define MB_PVQ_CELLS 8
#endif
constexpr uint32_t PVQ_CELLS = MB_PVQ_CELLS;
__global__ void __launch_bounds__(PVQ_THREADS) k_pvq(PvqParams P, uint32_t n_cells, uint32_t n_main, uint32_t aux_base) {
  __shared__ float X[2 * MAX_N];
  __shared__ uint32_t live_s;
  Team tm = {(int)threadIdx.x, (int)blockDim.x};
  const uint32_t base = blockIdx.x * PVQ_CELLS;
  if (threadIdx.x < 32) {
    const uint32_t f = launch_cell(base + threadIdx.x, n_main, aux_base);
    bool live = false;
    if (threadIdx.x < PVQ_CELLS && base + threadIdx.x < n_cells) {
      if (P.meta) live = (P.meta[f].flags & 1u) != 0;
      else {
        const uint32_t status = P.rec[rec_index(f, S_STATUS)];
        live = MB_ST_DECODED(status) && (!((status >> 16) & 1u) || ((status >> 17) & 1u) || ((status >> 19) & 3u));   // has a CELT record
      }
    }
    const uint32_t m = __ballot_sync(0xffffffffu, live);
    if (threadIdx.x == 0) live_s = m;
  }
  __syncthreads();
  uint32_t m = live_s;
  while (m) {
    const int i = __ffs(m) - 1;
    m &= m - 1;
    celt_pvq_frame(tm, X, launch_cell(base + (uint32_t)i, n_main, aux_base), P);
    __syncthreads();
  }
}

__global__ void __launch_bounds__(128, MB_PLAN_MIN_CTAS) k_plan(PlanParams P, uint32_t n_cells, uint32_t n_main, uint32_t aux_base) {
  uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_cells) return;
  celt_plan_frame(launch_cell(f, n_main, aux_base), P);
}

// one warp per frame; 4 frames per CTA
#ifndef MB_BANDS_WARPS
#define MB_BANDS_WARPS 4
#endif
constexpr int BANDS_WARPS = MB_BANDS_WARPS;
__global__ void __launch_bounds__(32 * BANDS_WARPS, 24 / BANDS_WARPS) k_bands(BandsParams P, uint32_t n_cells, uint32_t n_main, uint32_t aux_base) {
  __shared__ BandsSmem sm[BANDS_WARPS];
  const uint32_t f = blockIdx.x * BANDS_WARPS + (threadIdx.x >> 5);
  if (f >= n_cells) return;
  Team tm = {(int)(threadIdx.x & 31), 32};
  celt_bands_frame(tm, &sm[threadIdx.x >> 5], launch_cell(f, n_main, aux_base), P);
}

// Hybrid frames: one warp per lane, the lane's ticks in order, because the reference's fold buffer outlives the frame
// (mb_celt_bands.cuh); `fold` is [stream][FOLD_WORDS].  Lanes without a hybrid frame leave after one pass over the flags.
__global__ void __launch_bounds__(32 * BANDS_WARPS, 24 / BANDS_WARPS) k_bands_hyb(BandsParams P, uint32_t n_lanes, const uint32_t* __restrict__ order,
                                                                 const uint32_t* __restrict__ order_n,
                                                                 const uint32_t* __restrict__ lane_stream, float* fold) {
  __shared__ BandsSmem sm[BANDS_WARPS];
  const uint32_t lane = blockIdx.x * BANDS_WARPS + (threadIdx.x >> 5);
  if (lane >= n_lanes) return;
  Team tm = {(int)(threadIdx.x & 31), 32};
  uint32_t any = 0;
  const uint32_t n_frames = order_n[lane];
  for (uint32_t k = tm.tid; k < n_frames; k += 32) any |= P.meta[order[(size_t)k * n_lanes + lane]].flags & MB_FL_FOLD;
  if (!__any_sync(0xffffffffu, any != 0)) return;
  float* fb = fold + (size_t)(lane_stream ? lane_stream[lane] : lane) * FOLD_WORDS;
  BandsSmem* my = &sm[threadIdx.x >> 5];
  for (int w = tm.tid; w < (int)FOLD_WORDS; w += 32) my->norm[w] = fb[w];
  __syncwarp();
  for (uint32_t k = 0; k < n_frames; k++) {
    const uint32_t f = order[(size_t)k * n_lanes + lane];
    if (P.meta[f].flags & MB_FL_FOLD) celt_bands_frame(tm, my, f, P, true);
  }
  __syncwarp();
  for (int w = tm.tid; w < (int)FOLD_WORDS; w += 32) fb[w] = my->norm[w];
}

// one warp per (frame, channel), 12 warps per CTA, a grid of resident CTAs that walk the units (mb_celt_imdct.cuh).
// The twiddle / window block is packed once per device (k_imdct_tables at bd_create) and every CTA brings it into
// shared memory with one bulk async copy (cp.async.bulk + mbarrier) before its first unit; a warp fetches the spectrum
// row of its next unit the same way (two row buffers and two mbarriers per warp) while it transforms the current one.
constexpr int IMDCT_WARPS = 12;
constexpr int IMDCT_CTAS_PER_SM = 2;
static __device__ __align__(128) ImdctTables g_imdct_tables;
__global__ void k_imdct_tables() {
  for (int i = threadIdx.x; i < 1800; i += blockDim.x) g_imdct_tables.mdct_tw[i] = bits2f(kMdctTwiddles960Bits[i]);
  for (int i = threadIdx.x; i < 960; i += blockDim.x) g_imdct_tables.fft_tw[i] = bits2f(kFftTwiddles480Bits[i]);
  for (int i = threadIdx.x; i < 120; i += blockDim.x) g_imdct_tables.window[i] = bits2f(kWindow120Bits[i]);
}
struct ImdctCtaSmem {
  ImdctTables T;
  float row[IMDCT_WARPS][2][IMDCT_ROW];
  unsigned long long bar_tables;
  unsigned long long bar_row[IMDCT_WARPS][2];
};
static_assert((IMDCT_ROW * sizeof(float)) % 16 == 0, "row buffers are bulk-copy destinations");
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done)
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done)
                 : "r"(bar), "r"(parity)
                 : "memory");
}
__global__ void __launch_bounds__(32 * IMDCT_WARPS, IMDCT_CTAS_PER_SM) k_imdct(ImdctParams P, uint32_t n_units, uint32_t n_main, uint32_t aux_base) {
  extern __shared__ __align__(128) unsigned char imdct_smem_raw[];
  ImdctCtaSmem* S = (ImdctCtaSmem*)imdct_smem_raw;
  const uint32_t wp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_t = (uint32_t)__cvta_generic_to_shared(&S->bar_tables);
  const uint32_t bar_r[2] = {(uint32_t)__cvta_generic_to_shared(&S->bar_row[wp][0]), (uint32_t)__cvta_generic_to_shared(&S->bar_row[wp][1])};
  if (threadIdx.x == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_t));
  if (lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_r[0]));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_r[1]));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) bulk_g2s((uint32_t)__cvta_generic_to_shared(&S->T), &g_imdct_tables, (uint32_t)sizeof(ImdctTables), bar_t);
  Team tm = {(int)lane, 32};
  const uint32_t nu = (uint32_t)P.CC, stride = gridDim.x * IMDCT_WARPS;
  uint32_t unit = blockIdx.x * IMDCT_WARPS + wp;
  auto fetch = [&](const ImdctUnit& U, int b) {
    if (lane == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the row's last readers (generic proxy) come first
      bulk_g2s((uint32_t)__cvta_generic_to_shared(&S->row[wp][b][0]), U.src, (uint32_t)U.N * 4u, bar_r[b]);
    }
  };
  ImdctUnit cur;
  cur.src = nullptr;
  if (unit < n_units) {
    cur = imdct_unit_plan(launch_cell(unit / nu, n_main, aux_base), (int)(unit % nu), P);
    if (cur.src) fetch(cur, 0);
  }
  mbar_wait(bar_t, 0);
  uint32_t phase[2] = {0, 0};
  int b = 0;
  while (unit < n_units) {
    const uint32_t nxt = unit + stride;
    ImdctUnit nx;
    nx.src = nullptr;
    if (nxt < n_units) {
      nx = imdct_unit_plan(launch_cell(nxt / nu, n_main, aux_base), (int)(nxt % nu), P);
      if (nx.src) fetch(nx, b ^ 1);
    }
    if (cur.src) {
      mbar_wait(bar_r[b], phase[b]);
      phase[b] ^= 1;
      imdct_unit_run(tm, S->row[wp][b], &S->T, cur, P);
    }
    __syncwarp();
    cur = nx;
    unit = nxt;
    b ^= 1;
  }
}

// phase 2c(i): one warp per stream-channel, 4 per CTA
#ifndef MB_COMB_WARPS
#define MB_COMB_WARPS 4
#endif
constexpr int COMB_WARPS = MB_COMB_WARPS;
__global__ void __launch_bounds__(32 * COMB_WARPS) k_comb(TailParams P) {
  __shared__ CombSmem sm[COMB_WARPS];
  const uint32_t idx = blockIdx.x * COMB_WARPS + (threadIdx.x >> 5);
  if (idx >= P.n_lanes * (uint32_t)P.CC) return;
  Team tm = {(int)(threadIdx.x & 31), 32};
  celt_comb_channel(tm, &sm[threadIdx.x >> 5], idx / (uint32_t)P.CC, (int)(idx % (uint32_t)P.CC), P);
}

// the slow tail of the streams with a frame concealed by pitch repetition (mb_celt_plc.cuh): one warp per stream
__global__ void __launch_bounds__(32) k_comb_plc(TailParams P) {
  __shared__ PlcSmem sm;
  if (blockIdx.x >= P.n_lanes) return;
  Team tm = {(int)threadIdx.x, 32};
  celt_comb_plc_lane(tm, &sm, blockIdx.x, P);
}

// phase 2c(ii): one lane per stream-channel. A warp owns 32 neighbouring stream-channels (16 stereo streams) and moves
// their signals through a [32 rows x CH samples] shared-memory tile: rows arrive by cp.async as coalesced 16-byte
// pieces (double-buffered: the next tile is in flight while the current one is filtered), the serial one-pole
// recursion runs down each lane's own row with 128-bit shared-memory accesses (row pitch CH + 4 words: conflict-free),
// and the interleaved PCM leaves as coalesced 128-bit stores. Same arithmetic as celt_deemph_lane.
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

constexpr int DE_WARPS = 2;
template <int CH>
__global__ void __launch_bounds__(32 * DE_WARPS) k_deemph(TailParams P) {
  constexpr int PITCH = CH + 4;                 // floats per tile row
  constexpr int PIECES = CH / 4;                // 16-byte pieces per row
  extern __shared__ __align__(16) unsigned char de_smem[];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  float* tiles = (float*)de_smem + (size_t)w * 2 * 32 * PITCH;       // two tiles per warp
  __shared__ const float* row_x_[DE_WARPS][32];
  __shared__ float* row_out_[DE_WARPS][32];
  __shared__ int row_n_[DE_WARPS][32];
  __shared__ int16_t* row_o16_[DE_WARPS][32];
  const float** row_x = row_x_[w];
  float** row_out = row_out_[w];
  int* row_n = row_n_[w];
  int16_t** row_o16 = row_o16_[w];
  const uint32_t n_main = P.n_lanes * P.n_ticks;
  const int CC = P.CC;
  const uint32_t total = P.n_lanes * (uint32_t)CC;
  const uint32_t idx0 = (blockIdx.x * DE_WARPS + w) * 32;
  if (idx0 >= total) return;
  const uint32_t idx = idx0 + l;
  const bool live = idx < total;
  const uint32_t lane = live ? idx / (uint32_t)CC : 0;
  const int c = live ? (int)(idx % (uint32_t)CC) : 0;
  const uint32_t sid = P.lane_stream ? P.lane_stream[lane] : lane;
  StreamState& st = P.st_all[sid];
  const float* ser = P.series + ((size_t)lane * CC + c) * P.series_stride + SER_PREFIX;
  Deemph d = {live ? st.preemph_mem[c] : 0.f, P.gain, P.apply_gain};

  auto issue = [&](int k0, float* tile) {   // rows' [k0, k0 + CH) -> tile, 16 bytes per lane per piece
#pragma unroll 4
    for (int r = 0; r < 32; r++) {
      const float* xr = row_x[r];
      const int nr = row_n[r];
#pragma unroll
      for (int q = l; q < PIECES; q += 32)
        if (k0 + 4 * q < nr) cp_async16(tile + r * PITCH + 4 * q, xr + k0 + 4 * q);
    }
    cp_async_commit();
  };

  const uint32_t my_frames = live ? P.order_n[lane] : 0u;
  const uint32_t max_frames = __reduce_max_sync(0xffffffffu, my_frames);
  for (uint32_t k = 0; k < max_frames; k++) {
    int N = 0, hyb = 0, side = 0, late_gain = 0;
    const float* x = nullptr;
    float* out = nullptr;
    int16_t* out16 = nullptr;            // int16 twin of the row: plain coded CELT-only frames of the int16 path only
    uint32_t need_cell = 0xffffffffu;    // the main cell whose clipper note this frame contributes to
    uint32_t over_mask = 0;              // bit r: row r of the warp holds a sample beyond +-1 (or could not be stored as int16)
    if (k < my_frames) {
      const uint32_t f = P.order[(size_t)k * P.n_lanes + lane];
      const FrameMeta m = P.meta[f];
      N = SHORT_MDCT << (int)MB_H0_LM(P.rec[rec_index(f, S_HDR0)]);
      x = ser + m.sample_off;
      if (m.flags & 16u) { out = side_row(P.xbuf, f); side = 1; }   // not the packet's audio: kept aside for k_fade
      else {
        const uint32_t fo = row_cell(P, f);
        const uint32_t p = P.pkt_map ? P.pkt_map[fo] : fo;
        const size_t eo = pcm_elem_off(P.out_map, p, P.pcm_stride);
        out = P.pcm_f32 ? P.pcm_f32 + eo : nullptr;
        if (P.clip_need && out) {
          need_cell = fo;
          // anything but a coded CELT-only frame alone in its row (SILK part, side / fade rows, concealment) keeps the clipper's route
          const uint32_t other = 2u | 8u | 16u | 32u | 64u | 128u | MB_FL_FADE_ROW | MB_FL_PLC | MB_FL_PLC_PITCH;
          if (P.ds == 1 && f < n_main && !(m.flags & other) && ((((uintptr_t)(P.pcm_i16 + eo)) & 7u) == 0) && ((((uintptr_t)out) & 15u) == 0))
            out16 = P.pcm_i16 + eo;
        }
      }
      if (m.flags & 4u) d.mem = 0.f;                 // CELT reset before this frame
      if ((m.flags & 8u) && out) hyb = 1;            // hybrid: add to the SILK samples already in the row
      late_gain = side || (m.flags & MB_FL_FADE_ROW);   // a row that waits for k_fade gets its gain there
    }
    const unsigned hyb_mask = __ballot_sync(0xffffffffu, hyb);
    const unsigned late_mask = __ballot_sync(0xffffffffu, late_gain);
    d.apply_gain = (hyb || late_gain) ? 0 : P.apply_gain;   // hybrid: the gain applies to SILK + CELT, at the store below
    __syncwarp();
    row_x[l] = x; row_out[l] = out; row_n[l] = N; row_o16[l] = out16;
    const int max_n = __reduce_max_sync(0xffffffffu, N);
    __syncwarp();
    if (max_n == 0) continue;
    issue(0, tiles);
    int buf = 0;
    for (int k0 = 0; k0 < max_n; k0 += CH, buf ^= 1) {
      float* tile = tiles + buf * 32 * PITCH;
      const bool more = k0 + CH < max_n;
      if (more) { issue(k0 + CH, tiles + (buf ^ 1) * 32 * PITCH); cp_async_wait<1>(); } else cp_async_wait<0>();
      __syncwarp();
      {  // the serial recursion, this lane's row
        const int cnt = min(CH, N - k0);
        float4* row = (float4*)(tile + l * PITCH);
        for (int q = 0; 4 * q < cnt; q++) {
          float4 v = row[q];
          v.x = d.step(v.x); v.y = d.step(v.y); v.z = d.step(v.z); v.w = d.step(v.w);
          row[q] = v;
        }
      }
      __syncwarp();
      if (P.ds > 1) {
        // output rates below 48 kHz: every ds-th filtered sample is kept (deemphasis with downsample, celt_decoder.rs:2232-2297)
        const int ds = P.ds;
        for (int r = 0; r < 32; r++) {
          float* o = row_out[r];
          const int cnt = min(CH, row_n[r] - k0);
          if (!o || cnt <= 0) continue;
          const float* tx = tile + r * PITCH;
          const int cr = CC == 2 ? (r & 1) : 0;     // (idx0 is a multiple of 32: row r is channel r % CC of its stream)
          const bool hy = (hyb_mask >> r) & 1u, late = (late_mask >> r) & 1u;
          for (int j = l; j < cnt; j += 32)
            if ((k0 + j) % ds == 0) {
              float* q = o + (size_t)((k0 + j) / ds) * CC + cr;
              float v = tx[j];
              if (hy) { v = *q + v; if (P.apply_gain && !late) v *= P.gain; }   // hybrid: onto the SILK samples, then the gain on the sum
              *q = v;
            }
        }
      } else if (CC == 2) {
#pragma unroll 2
        for (int r = 0; r < 32; r += 2) {
          float* o = row_out[r];
          const int cnt = min(CH, row_n[r] - k0);
          if (!o || cnt <= 0) continue;
          o += (size_t)k0 * 2;
          const float* tx = tile + r * PITCH;
          const float* ty = tx + PITCH;
          if ((hyb_mask >> r) & 1u) {
            for (int e = l; e < 2 * cnt; e += 32) {
              float v = o[e] + ((e & 1) ? ty[e >> 1] : tx[e >> 1]);
              if (P.apply_gain && !((late_mask >> r) & 1u)) v *= P.gain;
              o[e] = v;
            }
          } else if (int16_t* o16 = row_o16[r]) {   // (aligned by construction) float and int16 forms of the same samples
            o16 += (size_t)k0 * 2;
            int ov = 0;
            for (int j = 2 * l; j < cnt; j += 64) {
              const float2 a = *(const float2*)(tx + j), b = *(const float2*)(ty + j);
              *(float4*)(o + 2 * j) = make_float4(a.x, b.x, a.y, b.y);
              ov |= !(a.x >= -1.f && a.x <= 1.f) | !(b.x >= -1.f && b.x <= 1.f) | !(a.y >= -1.f && a.y <= 1.f) | !(b.y >= -1.f && b.y <= 1.f);
              uint2 q;
              q.x = (uint32_t)(uint16_t)float2int16(a.x) | ((uint32_t)(uint16_t)float2int16(b.x) << 16);
              q.y = (uint32_t)(uint16_t)float2int16(a.y) | ((uint32_t)(uint16_t)float2int16(b.y) << 16);
              *(uint2*)(o16 + 2 * j) = q;
            }
            if (__any_sync(0xffffffffu, ov)) over_mask |= 3u << r;
          } else if (((uintptr_t)o & 15u) == 0) {
            for (int j = 2 * l; j < cnt; j += 64) {   // float4 = (x[j], y[j], x[j+1], y[j+1])
              const float2 a = *(const float2*)(tx + j), b = *(const float2*)(ty + j);
              *(float4*)(o + 2 * j) = make_float4(a.x, b.x, a.y, b.y);
            }
          } else {
            for (int e = l; e < 2 * cnt; e += 32) o[e] = (e & 1) ? ty[e >> 1] : tx[e >> 1];
          }
        }
      } else {
#pragma unroll 4
        for (int r = 0; r < 32; r++) {
          float* o = row_out[r];
          const int cnt = min(CH, row_n[r] - k0);
          if (!o || cnt <= 0) continue;
          const float* tx = tile + r * PITCH;
          if ((hyb_mask >> r) & 1u) {
            for (int j = l; j < cnt; j += 32) {
              float v = o[k0 + j] + tx[j];
              if (P.apply_gain && !((late_mask >> r) & 1u)) v *= P.gain;
              o[k0 + j] = v;
            }
          } else if (int16_t* o16 = row_o16[r]) {
            int ov = 0;
            for (int j = l; j < cnt; j += 32) {
              const float v = tx[j];
              o[k0 + j] = v;
              ov |= !(v >= -1.f && v <= 1.f);
              o16[k0 + j] = float2int16(v);
            }
            if (__any_sync(0xffffffffu, ov)) over_mask |= 1u << r;
          } else
            for (int j = l; j < cnt; j += 32) o[k0 + j] = tx[j];
        }
      }
      __syncwarp();
    }
    // the clipper's note for this frame's packet (both channels of a stream agree: over_mask carries the pair's bits)
    if (need_cell != 0xffffffffu) atomicMax(P.clip_need + need_cell, (out16 && !((over_mask >> l) & 1u)) ? 1u : 2u);
  }
  if (live) st.preemph_mem[c] = d.mem;
}

// Redundancy cross-fades: one thread per (cell, channel); cells without a redundancy frame leave at the flag test.
__global__ void __launch_bounds__(128) k_fade(TailParams P, uint32_t n_cells) {
  const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= n_cells * (uint32_t)P.CC) return;
  celt_fade_cell(u / (uint32_t)P.CC, (int)(u % (uint32_t)P.CC), P);
}

// Warp-cooperative form of soft_clip_channel (mb_celt_tail.cuh; opus_pcm_soft_clip_impl, src/opus.rs:16-140): the scans
// for the first sample beyond +-1, for the zero crossings on either side and for the peak are ballots / a max-reduction
// over 32 samples at a time, the correction x + a*x*x is applied in parallel. Comparisons and max are exact, every sample
// sees the same arithmetic as in the serial form, so the result is bit-identical to it. Only the rare linear ramp of the
// "clip before the first zero crossing" case keeps its serial float recurrence (lane 0).
template <class Pred> __device__ __forceinline__ int warp_first(int lo, int hi, int l, Pred pred) {
  for (int b = lo; b < hi; b += 32) {
    const int i = b + l;
    const unsigned m = __ballot_sync(0xffffffffu, i < hi && pred(i));
    if (m) return b + __ffs(m) - 1;
  }
  return hi;
}
template <class Pred> __device__ __forceinline__ int warp_last_below(int hi, int l, Pred pred) {   // largest i < hi with pred(i), else -1
  for (int b = hi; b > 0; b -= 32) {
    const int i = b - 1 - l;
    const unsigned m = __ballot_sync(0xffffffffu, i >= 0 && pred(i));
    if (m) return b - 1 - (__ffs(m) - 1);
  }
  return -1;
}
// Everything behind the first region, all regions at once. A region is a maximal run of samples none of which has the sign
// opposite to its first sample beyond +-1 (zeros join either side and stay zero), its correction depends on nothing but the run's
// largest magnitude — so the runs can be found and corrected independently: lane l walks samples [lo + l*K, lo + (l+1)*K), closes
// the runs that begin and end inside its stretch on its own, and the runs that cross stretches are joined by one walk over the 32
// lanes' summaries (shuffles). Same comparisons, same (maxval - 1) / maxval^2 arithmetic per run as the serial form: bit-identical.
// Returns the coefficient of the run that reaches N (0 if it needs none) — what the serial loop leaves in `a`.
__device__ __noinline__ float soft_clip_runs_warp(float* x, int lo, int N, int C, int l) {
  const int K = (N - lo + 31) >> 5;
  const int b = lo + l * K, e = min(N, b + K);
  auto coef = [](float maxval, int sign) {
    float a = (maxval - 1) / (maxval * maxval);
    a += a * 2.4e-7f;
    return sign > 0 ? -a : a;
  };
  int fs = 0, cs = 0, nch = 0, head_end = e, seg_begin = b;
  float head_max = 0.f, seg_max = 0.f;
  for (int i = b; i < e; i++) {
    const float v = x[i * C];
    const int sg = v > 0.f ? 1 : (v < 0.f ? -1 : 0);
    if (sg != 0 && cs != 0 && sg != cs) {   // a run ends in front of sample i
      if (nch == 0) { head_max = seg_max; head_end = i; }
      else if (seg_max > 1.f) {
        const float a = coef(seg_max, cs);
        for (int k = seg_begin; k < i; k++) { const float u = x[k * C]; x[k * C] = u + a * u * u; }
      }
      nch++; seg_begin = i; seg_max = 0.f;
    }
    if (sg != 0) { cs = sg; if (fs == 0) fs = sg; }
    seg_max = fmaxf(seg_max, fabsf(v));
  }
  if (nch == 0) head_max = seg_max;
  const float tail_max = seg_max;
  const int tail_begin = seg_begin, ls = cs;
  // join the runs across lanes
  int id = 0, csign = 0, my_h = -1, my_t = -1, hs = 0, ts = 0;
  float cmax = 0.f, hfin = 0.f, tfin = 0.f;
  for (int j = 0; j < 32; j++) {
    if (lo + j * K >= N) break;
    const int fsj = __shfl_sync(0xffffffffu, fs, j), lsj = __shfl_sync(0xffffffffu, ls, j), nchj = __shfl_sync(0xffffffffu, nch, j);
    const float hmj = __shfl_sync(0xffffffffu, head_max, j), tmj = __shfl_sync(0xffffffffu, tail_max, j);
    if (fsj != 0 && csign != 0 && fsj != csign) {
      if (id == my_h) { hfin = cmax; hs = csign; }
      if (id == my_t) { tfin = cmax; ts = csign; }
      id++; cmax = 0.f; csign = 0;
    }
    if (j == l) my_h = id;
    cmax = fmaxf(cmax, hmj);
    if (fsj != 0 && csign == 0) csign = fsj;
    if (nchj > 0) {
      if (id == my_h) { hfin = cmax; hs = csign; }
      if (id == my_t) { tfin = cmax; ts = csign; }
      id++; cmax = tmj; csign = lsj;
    }
    if (j == l) my_t = id;
  }
  if (id == my_h) { hfin = cmax; hs = csign; }
  if (id == my_t) { tfin = cmax; ts = csign; }
  if (b < e) {
    if (hfin > 1.f) {
      const float a = coef(hfin, hs);
      for (int k = b; k < head_end; k++) { const float u = x[k * C]; x[k * C] = u + a * u * u; }
    }
    if (nch > 0 && tfin > 1.f) {
      const float a = coef(tfin, ts);
      for (int k = tail_begin; k < e; k++) { const float u = x[k * C]; x[k * C] = u + a * u * u; }
    }
  }
  __syncwarp();
  return cmax > 1.f ? coef(cmax, csign) : 0.f;
}

// all_serial: every region in turn (the form the parallel one is checked against; also taken for rows that hold a NaN)
__device__ __forceinline__ void soft_clip_channel_warp(float* x, int N, int C, float& a_io, int l, bool all_serial) {
  float a = a_io;
  {
    const int first = warp_first(0, N, l, [&](int i) { return x[i * C] * a >= 0; });
    for (int i = l; i < first; i += 32) { const float v = x[i * C]; x[i * C] = v + a * v * v; }
    __syncwarp();
  }
  int curr = 0;
  const float x0 = x[0];
  for (;;) {
    const int i0 = warp_first(curr, N, l, [&](int i) { const float v = x[i * C]; return !(v >= -1.f && v <= 1.f); });
    if (i0 == N) { a = 0; break; }
    const float xi = x[i0 * C];
    const int start = warp_last_below(i0, l, [&](int i) { return !(xi * x[i * C] >= 0); }) + 1;
    if (!all_serial && start != 0) {   // only a region that opens the row needs the serial form (its ramp, its peak position)
      __syncwarp();
      a = soft_clip_runs_warp(x, curr, N, C, l);
      break;
    }
    const int end = warp_first(i0, N, l, [&](int i) { return !(xi * x[i * C] >= 0); });
    // peak: the first position of the largest magnitude in [i0, end)
    float mx = -1.f;
    int mp = 0x7fffffff;
    for (int i = i0 + l; i < end; i += 32) { const float v = fabsf(x[i * C]); if (v > mx) { mx = v; mp = i; } }
    for (int d = 16; d; d >>= 1) {
      const float omx = __shfl_xor_sync(0xffffffffu, mx, d);
      const int omp = __shfl_xor_sync(0xffffffffu, mp, d);
      if (omx > mx || (omx == mx && omp < mp)) { mx = omx; mp = omp; }
    }
    float maxval = fabsf(xi);
    int peak_pos = i0;
    if (mx > maxval) { maxval = mx; peak_pos = mp; }
    const int special = (start == 0 && xi * x[0] >= 0);
    a = (maxval - 1) / (maxval * maxval);
    a += a * 2.4e-7f;
    if (xi > 0) a = -a;
    __syncwarp();
    for (int i = start + l; i < end; i += 32) { const float v = x[i * C]; x[i * C] = v + a * v * v; }
    __syncwarp();
    if (special && peak_pos >= 2) {
      if (l == 0) {
        float offset = x0 - x[0];
        const float delta = offset / (float)peak_pos;
        for (int i = curr; i < peak_pos; i++) {
          offset -= delta;
          float v = x[i * C] + offset;
          x[i * C] = v < -1.f ? -1.f : (v > 1.f ? 1.f : v);
        }
      }
      __syncwarp();
    }
    curr = end;
    if (curr == N) break;
  }
  a_io = a;
}

// opus_decode's output stage (src/opus_decoder.rs:1679-1712): soft clip (src/opus.rs:16-140) then float -> int16
// (src/celt/float_cast.rs:38-41). One warp per stream, its packets in tick order: the packet's interleaved float PCM is
// staged in shared memory; a packet with no sample outside [-1, 1] and no carried-over nonlinearity (the common case)
// skips the serial clipper entirely; otherwise one lane per channel runs it in shared memory. The int16 PCM leaves as
// coalesced 32-bit stores.
constexpr int CLIP_WARPS = 4;
static __device__ unsigned long long g_clip_stats[2];   // packets the clipper skipped (int16 already stored) / worked on; read under MOUSIKI_B200_TIMING
__global__ void __launch_bounds__(32 * CLIP_WARPS) k_clip_i16(uint32_t n_lanes, uint32_t n_ticks, const uint32_t* __restrict__ pkt_map,
                                                              const uint32_t* __restrict__ lane_stream, const FrameMeta* __restrict__ meta,
                                                              const int32_t* __restrict__ samples_out, StreamState* st,
                                                              float* __restrict__ pcm, int16_t* __restrict__ pcm16, uint32_t pcm_stride,
                                                              int CC, const uint32_t* __restrict__ out_map,
                                                              const uint8_t* __restrict__ flags, const uint32_t* __restrict__ clip_need, bool clip_serial) {
  __shared__ __align__(16) float buf_[CLIP_WARPS][2 * MAX_N];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const uint32_t lane = blockIdx.x * CLIP_WARPS + w;
  if (lane >= n_lanes) return;
  float* buf = buf_[w];
  StreamState& s = st[lane_stream ? lane_stream[lane] : lane];
  float mem = l < 2 ? s.softclip_mem[l] : 0.f;
  unsigned n_skipped = 0, n_worked = 0;
  for (uint32_t t = 0; t < n_ticks; t++) {
    const uint32_t f = t * n_lanes + lane;
    const uint32_t fl = meta[f].flags;
    if (!(fl & (3u | MB_FL_PLC))) continue;
    const uint32_t p = pkt_map ? pkt_map[f] : f;
    int n = samples_out[p];
    if (n <= 0) continue;
    // k_deemph's note (TailParams::clip_need): 1 on every frame of the packet = its int16 samples are stored and none lies beyond +-1
    bool stored = clip_need && clip_need[f] == 1u;
    if (out_map) {
      // a packet cut into frames is clipped as a whole (src/opus_decoder.rs:1352-1366): its frames sit in the following
      // ticks and in one contiguous stretch of the row; if one of them failed, the reference never reaches the clipper
      if (flags[p] & 4u) continue;
      bool failed = false;
      uint32_t t2 = t + 1;
      for (; t2 < n_ticks; t2++) {
        const uint32_t p2 = pkt_map[t2 * n_lanes + lane];
        if (p2 == 0xffffffffu || !(flags[p2] & 4u)) break;
        const int n2 = samples_out[p2];
        if (n2 <= 0) failed = true; else n += n2;
        stored = stored && clip_need[t2 * n_lanes + lane] == 1u;
      }
      if (failed) continue;
    }
    // The packet's route (k_energy, the same on all its frames): float row -> clip -> int16, or int16 written by k_silk
    // already. The latter is final unless the previous packet left a non-linearity to carry over (opus.rs:44-60).
    const bool from_float = (fl & MB_FL_F32ROW) != 0;
    // a LOST packet is not soft-clipped, and leaves the clipper's memory cleared (opus_decoder.rs:1186-1222: soft_clip = false)
    const bool lost_pkt = flags && (flags[p] & 3u);   // (nor is one rebuilt from in-band FEC data, :1312)
    if (lost_pkt) mem = 0.f;
    if (!from_float && !__any_sync(0xffffffffu, mem != 0.f)) continue;
    // the common case of the int16 path: nothing to clip, nothing carried over, and the samples are already in place
    if (stored && from_float && !lost_pkt && !__any_sync(0xffffffffu, mem != 0.f)) { n_skipped++; continue; }
    n_worked++;
    const int total = n * CC;
    const size_t eo = pcm_elem_off(out_map, p, pcm_stride);
    float* src = pcm + eo;
    if (!from_float) {   // SILK samples are exact multiples of 2^-15: back to float without loss, into the float scratch row
      const int16_t* s16 = pcm16 + eo;
      for (int i = l; i < total; i += 32) src[i] = (float)s16[i] * (1.0f / 32768.0f);
      __syncwarp();
    }
    if (total > 2 * MAX_N) buf = src;   // long multi-frame packet: clip in place in the float scratch row
    int over = 0, nan = 0;
    const bool vec = ((((uintptr_t)src) & 15u) | (uint32_t)(total & 7)) == 0 && ((((uintptr_t)(pcm16 + eo)) & 15u) == 0);
    if (vec) {  // 128-bit loads, several in flight per lane: the row is read once and nothing else hides its latency
      const float4* s4 = (const float4*)src;
      float4* b4 = (float4*)buf;
      const int total4 = total >> 2;
#pragma unroll 4
      for (int i = l; i < total4; i += 32) {
        float4 v = s4[i];
        over |= !(v.x >= -1.f && v.x <= 1.f) | !(v.y >= -1.f && v.y <= 1.f) | !(v.z >= -1.f && v.z <= 1.f) | !(v.w >= -1.f && v.w <= 1.f);
        nan |= (v.x != v.x) | (v.y != v.y) | (v.z != v.z) | (v.w != v.w);
        v.x = v.x < -2.f ? -2.f : (v.x > 2.f ? 2.f : v.x); v.y = v.y < -2.f ? -2.f : (v.y > 2.f ? 2.f : v.y);
        v.z = v.z < -2.f ? -2.f : (v.z > 2.f ? 2.f : v.z); v.w = v.w < -2.f ? -2.f : (v.w > 2.f ? 2.f : v.w);
        b4[i] = v;
      }
    } else {
      for (int i = l; i < total; i += 32) {
        float v = src[i];
        over |= !(v >= -1.f && v <= 1.f);
        nan |= v != v;
        buf[i] = v < -2.f ? -2.f : (v > 2.f ? 2.f : v);
      }
    }
    over |= mem != 0.f;
    over = __any_sync(0xffffffffu, over) && !lost_pkt;
    __syncwarp();
    if (over) {
      const bool all_serial = clip_serial || __any_sync(0xffffffffu, nan);
      for (int c = 0; c < CC; c++) {
        float a = __shfl_sync(0xffffffffu, mem, c);
        soft_clip_channel_warp(buf + c, n, CC, a, l, all_serial);
        if (l == c) mem = a;
      }
      __syncwarp();
    }
    int16_t* dst = pcm16 + eo;
    if (vec) {
      const float4* b4 = (const float4*)buf;
      for (int i = l; i < (total >> 3); i += 32) {
        const float4 a = b4[2 * i], b = b4[2 * i + 1];
        uint4 o;
        o.x = (uint32_t)(uint16_t)float2int16(a.x) | ((uint32_t)(uint16_t)float2int16(a.y) << 16);
        o.y = (uint32_t)(uint16_t)float2int16(a.z) | ((uint32_t)(uint16_t)float2int16(a.w) << 16);
        o.z = (uint32_t)(uint16_t)float2int16(b.x) | ((uint32_t)(uint16_t)float2int16(b.y) << 16);
        o.w = (uint32_t)(uint16_t)float2int16(b.z) | ((uint32_t)(uint16_t)float2int16(b.w) << 16);
        ((uint4*)dst)[i] = o;
      }
    } else if ((((uintptr_t)dst) & 3u) == 0) {
      for (int i = 2 * l; i + 1 < total; i += 64) {
        const uint32_t lo = (uint16_t)float2int16(buf[i]), hi = (uint16_t)float2int16(buf[i + 1]);
        *(uint32_t*)(dst + i) = lo | (hi << 16);
      }
      if ((total & 1) && l == 0) dst[total - 1] = float2int16(buf[total - 1]);
    } else {
      for (int i = l; i < total; i += 32) dst[i] = float2int16(buf[i]);
    }
    __syncwarp();
    buf = buf_[w];
  }
  if (l < 2) s.softclip_mem[l] = mem;
  if (l == 0) { atomicAdd(&g_clip_stats[0], (unsigned long long)n_skipped); atomicAdd(&g_clip_stats[1], (unsigned long long)n_worked); }
}
// res_to_int24 (src/celt/celt_decoder.rs:3165-3176), in place on the float scratch: same 32-bit containers
__global__ void k_f32_to_i24(float* __restrict__ pcm, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float v = pcm[i] * (32768.0f * 256.0f);
  v = v < -8388608.0f ? -8388608.0f : (v > 8388607.0f ? 8388607.0f : v);
  ((int32_t*)pcm)[i] = __float2int_rn(v);
}
// decode_float leaves the clipper memory cleared (src/opus_decoder.rs:1411-1417)
__global__ void k_clear_softclip(uint32_t n_lanes, const uint32_t* __restrict__ lane_stream, StreamState* st) {
  uint32_t lane = blockIdx.x * blockDim.x + threadIdx.x;
  if (lane >= n_lanes) return;
  StreamState& s = st[lane_stream ? lane_stream[lane] : lane];
  s.softclip_mem[0] = s.softclip_mem[1] = 0.f;
}
__global__ void k_reset(StreamState* st, float* hist, const uint32_t* __restrict__ ids, uint32_t n, int CC) {
  uint32_t i = blockIdx.x;
  if (i >= n) return;
  uint32_t sid = ids ? ids[i] : i;
  if (threadIdx.x == 0) stream_state_reset(st[sid]);
  float* h = hist + (size_t)sid * CC * HIST_G;
  for (int k = threadIdx.x; k < CC * HIST_G; k += blockDim.x) h[k] = 0.f;
}


// SILK-only packets, phase 2: one lane per stream, its packets in tick order (mb_silk_synth.cuh / mb_silk_out.cuh).
// The stream's SILK state is staged in the lane's local memory for the pass and written back at the end.
__global__ void __launch_bounds__(64, 12) k_silk(SilkLaneParams q) {
  __shared__ int32_t stage[64][33];   // 32 staged output samples per lane, pitch 33: conflict-free for lockstep lanes
  __shared__ int16_t fir[48];
  if (threadIdx.x < 48) fir[threadIdx.x] = silk_resampler_frac_FIR_12[threadIdx.x];
  __syncthreads();
  const uint32_t lane = blockIdx.x * blockDim.x + threadIdx.x;
  if (lane >= q.n_lanes) return;
  SilkWork W;
  silk_lane(q, lane, W, stage[threadIdx.x], fir);
}
__global__ void k_silk_reset(SilkState* st, uint32_t* silk_hist, uint32_t n_streams, const uint32_t* __restrict__ ids, uint32_t n) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t sid = ids ? ids[i] : i;
  silk_state_reset(st[sid], 1);   // (marks the excitation buffers for clearing on first use, mb_silk_plc.cuh)
  for (int k = 0; k < 320; k++) silk_hist[(size_t)k * n_streams + sid] = 0;
}

// unit known-answer hook: one warp runs one case on the kernel bodies (mb_kat.cuh)
__global__ void __launch_bounds__(32) k_kat(int what, const uint32_t* __restrict__ in, uint32_t n_in, uint32_t* out, uint32_t n_out,
                                             uint8_t* gbytes, int* rc) {
  __shared__ KatSmem sm;
  Team tm = {(int)threadIdx.x, 32};
  const int r = kat_run(tm, &sm, what, in, n_in, out, n_out, gbytes, &g_imdct_tables);
  if (threadIdx.x == 0) *rc = r;
}

// Row A2, the lightweight decoder (mb_lite.cuh): one lane per stream walks its packets in tick order; after each tick the
// warp writes its 32 rows together (each stream's 320-sample buffer held 3x, converted on the way) so the stores coalesce.
struct LiteParams {
  const uint8_t* arena;
  const uint32_t* offsets;
  LiteState* st;
  void* pcm;          // [packet][960] f32 / i16 / i32
  int32_t* status;    // [packet]
  uint32_t n_lanes, n_ticks;
  int mode;           // 0: decode_float32, 1: decode (i16), 2: decode_int24
};
__global__ void __launch_bounds__(64) k_lite(LiteParams P) {
  const uint32_t lane = blockIdx.x * blockDim.x + threadIdx.x, l32 = threadIdx.x & 31, lane0 = lane - l32;
  const bool live = lane < P.n_lanes;
  LiteFrame F;
  float res[LITE_BUF], lag[LITE_MAX_RES_LAG], lpc[LITE_BUF];
  for (uint32_t t = 0; t < P.n_ticks; t++) {
    int rc = 0;
    if (live) {
      const uint32_t p = t * P.n_lanes + lane, a = P.offsets[p], b = P.offsets[p + 1];
      rc = lite_decode_packet(P.st[lane], P.arena + a, b - a, F, res, lag, lpc);
      P.status[p] = rc;
    }
    __syncwarp();
    for (uint32_t k = 0; k < 32 && lane0 + k < P.n_lanes; k++) {
      const int ok = __shfl_sync(0xffffffffu, rc, (int)k) == LITE_OUT;
      const float* src = P.st[lane0 + k].silk_buffer;
      const size_t row = ((size_t)t * P.n_lanes + lane0 + k) * LITE_OUT;
      for (uint32_t i = l32; i < LITE_OUT; i += 32) {
        const float v = ok ? src[i / 3] : 0.f;   // (rows of refused packets are cleared; the reference leaves `out` alone)
        if (P.mode == 0) ((float*)P.pcm)[row + i] = v;
        else if (P.mode == 1) ((int16_t*)P.pcm)[row + i] = (int16_t)lite_floor_i32(v * 32767.0f);
        else ((int32_t*)P.pcm)[row + i] = (int32_t)rint((double)v * 8388608.0);
      }
    }
    __syncwarp();
  }
}
__global__ void k_lite_reset(LiteState* st, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) lite_state_reset(st[i]);
}

// ------------------------------------------------------------------------------------------ host side
namespace {

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() { return (T*)p; }
};

}  // namespace

struct bd_handle {
  bd_config cfg;
  int device = 0;
  int CC = 2;
  int n_sm = 148;
  int disable_inv = 0;
  float gain = 1.f;
  uint32_t max_frames_per_pass = 0;
  bool want_i24 = false;       // the running host call delivers 24-bit integers (bd_decode_i24)
  std::vector<uint32_t> sched_ids;          // stream ids of the call the kept schedule (h_lane_stream, h_pkt_map) belongs to
  uint32_t sched_n = 0, sched_lanes = 0, sched_ticks = 0;
  bool sched_identity = false;
  const uint8_t* arena_staged = nullptr;    // the host arena whose bytes are already on their way to d_arena (decode_host)
  std::vector<uint8_t> h_toc;               // TOC byte of every packet of the call in flight
  std::function<void()> after_enqueue;      // host work of the call that runs beside the device work
  int ds = 1;                  // 48000 / sample_rate_hz
  bool clip_serial = false;    // MOUSIKI_B200_CLIP_SERIAL=1: the clipper walks the regions of a row one after the other (A/B and test switch)
  bool fused_i16 = true;       // MOUSIKI_B200_FUSED_I16=0: k_deemph leaves every int16 sample to k_clip_i16 (A/B switch)
  bool parallel_bands = false; // MOUSIKI_B200_PARALLEL_BANDS=1: small calls run CELT-only frames through the all-at-once band pass (see k_energy)
  std::string err;
  uint32_t launches = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t copy_stream = nullptr;   // device -> host PCM of pass k overlaps the kernels of pass k + 1
  cudaStream_t side_stream = nullptr;   // the latency-bound per-stream kernels (k_energy, k_silk) run here beside k_pvq
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_silk = nullptr;
  cudaEvent_t ev_clip_done[2] = {nullptr, nullptr};   // the clipper of the pass that last used each half of the frame meta
  cudaStream_t clip_stream = nullptr;   // k_clip_i16 of pass k runs beside the kernels of pass k + 1
  std::vector<cudaEvent_t> pass_events;
  size_t pass_event_next = 0;
  // MOUSIKI_B200_TIMING: timestamps of the pass / clip / copy milestones of the last host call
  bool timeline_on = false;
  std::vector<std::pair<std::string, cudaEvent_t>> timeline;
  void mark(const char* what, size_t k, cudaStream_t st) {
    if (!timeline_on) return;
    cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st);
    timeline.push_back({std::string(what) + " " + std::to_string(k), e});
  }
  struct PassOut { bool on = false; char* host_pcm = nullptr; const char* dev_pcm = nullptr; size_t elem = 0; } pass_out;
  DevBuf states, silk_states, silk_hist, silk_exc, fold, hist, rec, en, meta, series, xbuf, order, side_of, aux_count, aux_main, clip_need, tails;
  // staging for the host-buffer entry points
  DevBuf d_arena, d_offsets, d_pkt_map, d_out_map, d_lane_stream, d_flags, d_pcm, d_pcm16, d_samples, d_range, d_ids, d_dbg_iy, d_dbg_x;
  std::vector<uint32_t> h_pkt_map, h_lane_stream, h_lane_of_stream, h_epoch, h_count, h_tick, v_range;
  std::vector<int32_t> v_samples;
  mbh::Split split;
  std::vector<mbh::StreamTrack> track;   // per stream: frame size / mode of the last packet (cuts lost packets into concealment frames)
  uint32_t epoch = 0;
  int32_t* dbg_iy_host = nullptr;
  float* dbg_x_host = nullptr;
  uint32_t dbg_max = 0;
  // optional per-kernel timing (CUDA events on the launching stream), see bd_profile_*
  bool prof_on = false;
  struct Span { int id; cudaEvent_t a, b; };
  std::vector<Span> prof_spans;
  std::vector<cudaEvent_t> prof_pool;
  cudaEvent_t prof_event() {
    if (!prof_pool.empty()) { cudaEvent_t e = prof_pool.back(); prof_pool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
  }
  void prof_begin(int id, cudaStream_t s) { if (prof_on) { Span sp{id, prof_event(), prof_event()}; cudaEventRecord(sp.a, s); prof_spans.push_back(sp); } }
  void prof_end(cudaStream_t s) { if (prof_on) cudaEventRecord(prof_spans.back().b, s); }
};

#define CK(call)                                                                          \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                       \
      return e_ == cudaErrorMemoryAllocation ? BD_ERR_ALLOC : BD_ERR_CUDA;                \
    }                                                                                     \
  } while (0)

extern "C" {

const char* bd_last_error(const bd_handle* h) { return h ? h->err.c_str() : "null handle"; }
uint32_t bd_last_launch_count(const bd_handle* h) { return h ? h->launches : 0; }

void* bd_alloc_pinned(uint64_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, (size_t)bytes, cudaHostAllocDefault) != cudaSuccess) return nullptr;
  return p;
}
void bd_free_pinned(void* p) { if (p) cudaFreeHost(p); }

int bd_create(const bd_config* cfg, bd_handle** out) {
  if (!cfg || !out) return BD_ERR_BAD_ARG;
  *out = nullptr;
  if (cfg->struct_size != sizeof(bd_config)) return BD_ERR_BAD_ARG;
  if (cfg->channels != 1 && cfg->channels != 2) return BD_ERR_BAD_ARG;
  const int fs = cfg->sample_rate_hz;
  if (!(fs == 48000 || fs == 24000 || fs == 16000 || fs == 12000 || fs == 8000)) return BD_ERR_BAD_ARG;
  // (row F4: below 48 kHz coded CELT-only frames decode — CELT's own down-sampling; the rest is refused per packet, mb_frame.cuh)
  if (cfg->n_streams == 0) return BD_ERR_BAD_ARG;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || cfg->device < 0 || cfg->device >= ndev) return BD_ERR_CUDA;
  bd_handle* h = new (std::nothrow) bd_handle();
  if (!h) return BD_ERR_ALLOC;
  h->cfg = *cfg;
  h->ds = 48000 / fs;
  h->cfg.quirks = (cfg->quirks & 0x00ffffffu) | (h->ds > 1 ? (uint32_t)h->ds << QX_DS_SHIFT : 0u);   // (mb_common.cuh: the kernels' copy carries the factor)
  h->device = cfg->device;
  h->CC = cfg->channels;
  h->disable_inv = cfg->disable_phase_inversion < 0 ? (cfg->channels == 1) : (cfg->disable_phase_inversion != 0);
  h->gain = cfg->gain_q8 == 0 ? 1.f : (float)exp(0.6931471805599453094 * (double)((3.32192809488736234787f / 5120.0f) * (float)cfg->gain_q8));
  h->max_frames_per_pass = cfg->max_frames_per_pass ? cfg->max_frames_per_pass : (1u << 19);   // ~16 GB of per-frame scratch at most: one pass for the headline call (40.9 vs 41.4 ms in two)
  { const char* sb = getenv("MOUSIKI_B200_PARALLEL_BANDS"); h->parallel_bands = sb && atoi(sb) > 0; }
  { const char* sb = getenv("MOUSIKI_B200_CLIP_SERIAL"); h->clip_serial = sb && atoi(sb) > 0; }
  { const char* sb = getenv("MOUSIKI_B200_FUSED_I16"); h->fused_i16 = !(sb && atoi(sb) == 0); }
  const char* env = getenv("MOUSIKI_B200_MAX_FRAMES_PER_PASS");
  if (env && atoi(env) > 0) h->max_frames_per_pass = (uint32_t)atoi(env);
  auto fail = [&](int code) { bd_destroy(h); return code; };
  if (cudaSetDevice(h->device) != cudaSuccess) return fail(BD_ERR_CUDA);
  {
    // The decode kernels outrank the clipper / copy side: a one-wave kernel (k_parse) that has to share its wave with clipper CTAs
    // takes twice as long, while the clipper of pass k has the whole of pass k + 1 to finish in (MOUSIKI_B200_MAIN_PRIO: A/B switch).
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    const char* mp = getenv("MOUSIKI_B200_MAIN_PRIO");
    int prio = mp ? atoi(mp) : (hi < lo ? lo - 1 : lo);
    if (prio < hi) prio = hi;
    if (prio > lo) prio = lo;
    if (cudaStreamCreateWithPriority(&h->own_stream, cudaStreamNonBlocking, prio) != cudaSuccess) return fail(BD_ERR_CUDA);
  }
  if (cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking) != cudaSuccess) return fail(BD_ERR_CUDA);
  {
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);   // hi = numerically smallest = most urgent
    if (cudaStreamCreateWithPriority(&h->side_stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess) return fail(BD_ERR_CUDA);
  }
  if (cudaStreamCreateWithFlags(&h->clip_stream, cudaStreamNonBlocking) != cudaSuccess) return fail(BD_ERR_CUDA);
  if (cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming) != cudaSuccess) return fail(BD_ERR_CUDA);
  if (cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming) != cudaSuccess) return fail(BD_ERR_CUDA);
  if (cudaEventCreateWithFlags(&h->ev_silk, cudaEventDisableTiming) != cudaSuccess) return fail(BD_ERR_CUDA);
  for (int i = 0; i < 2; i++)
    if (cudaEventCreateWithFlags(&h->ev_clip_done[i], cudaEventDisableTiming) != cudaSuccess) return fail(BD_ERR_CUDA);
  if (cudaFuncSetAttribute(k_deemph<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(DE_WARPS * 2 * 32 * (128 + 4) * sizeof(float))) != cudaSuccess)
    return fail(BD_ERR_CUDA);
  if (cudaFuncSetAttribute(k_imdct, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ImdctCtaSmem)) != cudaSuccess) return fail(BD_ERR_CUDA);
  {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, h->device) != cudaSuccess) return fail(BD_ERR_CUDA);
    h->n_sm = prop.multiProcessorCount;
  }
  {   // the packed IMDCT table block of this device, written once (the stream sync below covers it)
    static std::mutex mu;
    static bool ready[64] = {false};
    std::lock_guard<std::mutex> lock(mu);
    if (h->device >= 64 || !ready[h->device]) {
      k_imdct_tables<<<1, 256, 0, h->own_stream>>>();
      if (cudaStreamSynchronize(h->own_stream) != cudaSuccess || cudaGetLastError() != cudaSuccess) return fail(BD_ERR_CUDA);
      if (h->device < 64) ready[h->device] = true;
    }
  }
  if (h->states.ensure((size_t)cfg->n_streams * sizeof(StreamState)) != cudaSuccess) return fail(BD_ERR_ALLOC);
  if (h->hist.ensure((size_t)cfg->n_streams * h->CC * HIST_G * sizeof(float)) != cudaSuccess) return fail(BD_ERR_ALLOC);
  if (h->silk_states.ensure((size_t)cfg->n_streams * sizeof(SilkState)) != cudaSuccess) return fail(BD_ERR_ALLOC);
  if (h->silk_hist.ensure((size_t)cfg->n_streams * 320 * sizeof(uint32_t)) != cudaSuccess) return fail(BD_ERR_ALLOC);
  // excitation of the last decoded SILK frame + comfort-noise FIFO, 2 channels x 2 x 320 words per stream (mb_silk_plc.cuh);
  // never read before the lane that owns a stream has cleared them (SilkChanState::exc_dirty)
  if (h->silk_exc.ensure((size_t)cfg->n_streams * 1280 * sizeof(int32_t)) != cudaSuccess) return fail(BD_ERR_ALLOC);
  if (h->fold.ensure((size_t)cfg->n_streams * FOLD_WORDS * sizeof(float)) != cudaSuccess) return fail(BD_ERR_ALLOC);
  if (cudaMemsetAsync(h->fold.p, 0, (size_t)cfg->n_streams * FOLD_WORDS * sizeof(float), h->own_stream) != cudaSuccess) return fail(BD_ERR_CUDA);
  k_silk_reset<<<(cfg->n_streams + 127) / 128, 128, 0, h->own_stream>>>(h->silk_states.as<SilkState>(), h->silk_hist.as<uint32_t>(),
                                                                        cfg->n_streams, nullptr, cfg->n_streams);
  k_reset<<<cfg->n_streams, 128, 0, h->own_stream>>>(h->states.as<StreamState>(), h->hist.as<float>(), nullptr, cfg->n_streams, h->CC);
  if (cudaStreamSynchronize(h->own_stream) != cudaSuccess || cudaGetLastError() != cudaSuccess) return fail(BD_ERR_CUDA);
  *out = h;
  return BD_OK;
}

void bd_destroy(bd_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  DevBuf* all[] = {&h->order, &h->side_of, &h->aux_count, &h->aux_main, &h->clip_need, &h->tails, &h->states, &h->silk_states, &h->silk_hist, &h->silk_exc, &h->fold, &h->hist, &h->rec, &h->en, &h->meta, &h->series, &h->xbuf, &h->d_arena, &h->d_offsets, &h->d_pkt_map, &h->d_out_map,
                   &h->d_lane_stream, &h->d_flags, &h->d_pcm, &h->d_pcm16, &h->d_samples, &h->d_range, &h->d_ids, &h->d_dbg_iy, &h->d_dbg_x};
  for (DevBuf* b : all) b->release();
  for (auto& sp : h->prof_spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
  for (auto e : h->prof_pool) cudaEventDestroy(e);
  for (auto e : h->pass_events) cudaEventDestroy(e);
  if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
  if (h->side_stream) cudaStreamDestroy(h->side_stream);
  if (h->clip_stream) cudaStreamDestroy(h->clip_stream);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  if (h->ev_silk) cudaEventDestroy(h->ev_silk);
  for (int i = 0; i < 2; i++) if (h->ev_clip_done[i]) cudaEventDestroy(h->ev_clip_done[i]);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  delete h;
}

int bd_reset_streams(bd_handle* h, const uint32_t* ids, uint32_t n) {
  if (!h) return BD_ERR_BAD_ARG;
  CK(cudaSetDevice(h->device));
  const uint32_t* d_ids = nullptr;
  if (ids) {
    for (uint32_t i = 0; i < n; i++)
      if (ids[i] >= h->cfg.n_streams) return BD_ERR_BAD_ARG;
    CK(h->d_ids.ensure((size_t)n * 4));
    CK(cudaMemcpyAsync(h->d_ids.p, ids, (size_t)n * 4, cudaMemcpyHostToDevice, h->own_stream));
    d_ids = h->d_ids.as<uint32_t>();
  } else n = h->cfg.n_streams;
  if (h->track.size() == h->cfg.n_streams)
    for (uint32_t i = 0; i < n; i++) h->track[ids ? ids[i] : i] = mbh::StreamTrack();
  if (n == 0) return BD_OK;
  k_reset<<<n, 128, 0, h->own_stream>>>(h->states.as<StreamState>(), h->hist.as<float>(), d_ids, n, h->CC);
  k_silk_reset<<<(n + 127) / 128, 128, 0, h->own_stream>>>(h->silk_states.as<SilkState>(), h->silk_hist.as<uint32_t>(), h->cfg.n_streams, d_ids, n);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->own_stream));
  return BD_OK;
}

int bd_get_stream_info(bd_handle* h, uint32_t sid, bd_stream_info* out) {
  if (!h || !out || sid >= h->cfg.n_streams) return BD_ERR_BAD_ARG;
  CK(cudaSetDevice(h->device));
  StreamState s;
  CK(cudaMemcpy(&s, h->states.as<StreamState>() + sid, sizeof(s), cudaMemcpyDeviceToHost));
  out->final_range = s.range_final; out->last_packet_duration = s.last_packet_duration; out->prev_mode = s.prev_mode;
  out->celt_error = s.celt_error;
  out->bandwidth = s.bandwidth;
  if (s.prev_mode == 1002) out->pitch = s.pf[0].period;                      // CELT: the post-filter period (OPUS_GET_PITCH)
  else {                                                                     // SILK / hybrid: dec_control.prev_pitch_lag (dec_api.rs:420-427)
    SilkState ss;
    CK(cudaMemcpy(&ss, h->silk_states.as<SilkState>() + sid, sizeof(ss), cudaMemcpyDeviceToHost));
    const int fs = ss.ch[0].fs_khz;
    out->pitch = (ss.ch[0].prev_signal_type == 2 && (fs == 8 || fs == 12 || fs == 16)) ? ss.ch[0].lag_prev * (fs == 8 ? 6 : fs == 12 ? 4 : 3) : 0;
  }
  return BD_OK;
}

// ---- packet queries (src/packet.rs:30-56, 177-280): host-only, no decoder and no device involved
int bd_packet_get_mode(const uint8_t* pkt, uint32_t avail) {
  if (!pkt || !avail) return BD_OPUS_BAD_ARG;
  const uint8_t toc = pkt[0];
  return toc & 0x80 ? 1002 : (toc & 0x60) == 0x60 ? 1001 : 1000;
}
int bd_packet_get_bandwidth(const uint8_t* pkt, uint32_t avail) {
  if (!pkt || !avail) return BD_OPUS_BAD_ARG;
  const uint8_t toc = pkt[0];
  const int two = (toc >> 5) & 3;
  if (toc & 0x80) return two == 0 ? 1101 : 1102 + two;          // CELT has no medium band
  if ((toc & 0x60) == 0x60) return toc & 0x10 ? 1105 : 1104;    // hybrid: SWB or FB
  return 1101 + two;
}
int bd_packet_get_nb_channels(const uint8_t* pkt, uint32_t avail) {
  if (!pkt || !avail) return BD_OPUS_BAD_ARG;
  return pkt[0] & 0x04 ? 2 : 1;
}
int bd_packet_get_samples_per_frame(const uint8_t* pkt, uint32_t avail, uint32_t fs_hz) {
  if (!pkt || !avail) return BD_OPUS_BAD_ARG;
  const uint8_t toc = pkt[0];
  const uint32_t size = (toc >> 3) & 3;
  uint64_t n;
  if (toc & 0x80) n = ((uint64_t)fs_hz << size) / 400;
  else if ((toc & 0x60) == 0x60) n = toc & 0x08 ? fs_hz / 50 : fs_hz / 100;
  else n = size == 3 ? (uint64_t)fs_hz * 60 / 1000 : ((uint64_t)fs_hz << size) / 100;
  return n > 0x7fffffffu ? BD_OPUS_BAD_ARG : (int)n;
}
int bd_packet_get_nb_frames(const uint8_t* pkt, uint32_t avail, uint32_t len) {
  if (!pkt || len == 0 || len > avail) return BD_OPUS_BAD_ARG;
  const int code = pkt[0] & 3;
  if (code == 0) return 1;
  if (code != 3) return 2;
  if (len < 2) return BD_OPUS_INVALID_PACKET;
  return pkt[1] & 0x3f;
}
int bd_packet_get_nb_samples(const uint8_t* pkt, uint32_t avail, uint32_t len, uint32_t fs_hz) {
  const int count = bd_packet_get_nb_frames(pkt, avail, len);
  if (count < 0) return count;
  const int per = bd_packet_get_samples_per_frame(pkt, avail, fs_hz);
  if (per < 0) return per;
  const uint64_t samples = (uint64_t)count * (uint64_t)per;
  if (samples * 25 > (uint64_t)fs_hz * 3) return BD_OPUS_INVALID_PACKET;   // more than 120 ms
  return (int)samples;
}

int bd_profile_enable(bd_handle* h, int on) {
  if (!h) return BD_ERR_BAD_ARG;
  h->prof_on = on != 0;
  return BD_OK;
}
int bd_profile_read(bd_handle* h, double* ms_out, uint32_t* launches_out) {
  if (!h || !ms_out || !launches_out) return BD_ERR_BAD_ARG;
  for (int i = 0; i < BD_K_COUNT; i++) { ms_out[i] = 0; launches_out[i] = 0; }
  for (auto& sp : h->prof_spans) {
    float ms = 0.f;
    cudaEventSynchronize(sp.b);
    if (cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess && sp.id >= 0 && sp.id < BD_K_COUNT) { ms_out[sp.id] += ms; launches_out[sp.id]++; }
    h->prof_pool.push_back(sp.a); h->prof_pool.push_back(sp.b);
  }
  h->prof_spans.clear();
  return BD_OK;
}

int bd_debug_capture(bd_handle* h, int32_t* pulses_host, float* spectrum_host, uint32_t max_packets) {
  if (!h) return BD_ERR_BAD_ARG;
  h->dbg_iy_host = pulses_host; h->dbg_x_host = spectrum_host; h->dbg_max = (pulses_host || spectrum_host) ? max_packets : 0;
  return BD_OK;
}

int bd_debug_kat(int device, int what, const uint32_t* in, uint32_t n_in, uint32_t* out, uint32_t n_out) {
  if (!in || !out) return BD_ERR_BAD_ARG;
  if (cudaSetDevice(device) != cudaSuccess) return BD_ERR_CUDA;
  uint32_t *d_in = nullptr, *d_out = nullptr;
  uint8_t* d_bytes = nullptr;
  int rc = BD_ERR_CUDA, krc = -1;
  if (cudaMalloc(&d_in, (size_t)n_in * 4 + 4) == cudaSuccess && cudaMalloc(&d_out, (size_t)n_out * 4 + 4 + sizeof(int)) == cudaSuccess &&
      cudaMalloc(&d_bytes, 64) == cudaSuccess && cudaMemcpy(d_in, in, (size_t)n_in * 4, cudaMemcpyHostToDevice) == cudaSuccess &&
      cudaMemset(d_out, 0, (size_t)n_out * 4 + 4 + sizeof(int)) == cudaSuccess) {
    int* d_rc = (int*)(d_out + n_out + 1);
    k_imdct_tables<<<1, 256>>>();
    k_kat<<<1, 32>>>(what, d_in, n_in, d_out, n_out, d_bytes, d_rc);
    if (cudaDeviceSynchronize() == cudaSuccess && cudaMemcpy(out, d_out, (size_t)n_out * 4, cudaMemcpyDeviceToHost) == cudaSuccess &&
        cudaMemcpy(&krc, d_rc, sizeof(int), cudaMemcpyDeviceToHost) == cudaSuccess)
      rc = krc == 0 ? BD_OK : BD_ERR_BAD_ARG;
  }
  cudaFree(d_in); cudaFree(d_out); cudaFree(d_bytes);
  return rc;
}

// One or more passes over the grid (ticks are split so that a pass holds at most max_frames_per_pass cells).
// All pointers are device pointers. pkt_map == nullptr: cell f is packet f.
static int decode_grid_device(bd_handle* h, const uint8_t* d_arena, const uint32_t* d_offsets, const uint32_t* d_pkt_map,
                              const uint8_t* d_flags, uint32_t n_lanes, uint32_t n_ticks, const uint32_t* d_lane_stream, float* d_pcm_f32,
                              int16_t* d_pcm_i16, uint32_t pcm_stride, int32_t* d_samples, uint32_t* d_range, uint32_t n_packets,
                              cudaStream_t s, int32_t* d_dbg_iy, float* d_dbg_x, const uint32_t* d_out_map = nullptr) {
  if (n_lanes == 0 || n_ticks == 0) return BD_OK;
  const int CC = h->CC;
  if (pcm_stride < (uint32_t)(SHORT_MDCT / h->ds * CC)) return BD_ERR_BAD_ARG;
  if ((uint64_t)n_lanes > h->max_frames_per_pass) return BD_ERR_BAD_ARG;  // caller splits lanes (lanes are independent)
  uint32_t ticks_per_pass = h->max_frames_per_pass / n_lanes;
  if (ticks_per_pass > n_ticks) ticks_per_pass = n_ticks;
  // PCM going back to the host pass by pass: at least four passes, so that the first copy starts early (the copy stream,
  // not the kernels, ends the call; measured on the headline workload: 71.4 -> 69.1 ms)
  {
    const char* mp = getenv("MOUSIKI_B200_MIN_PASSES");   // (tuning switch)
    const uint32_t min_passes = mp && atoi(mp) > 0 ? (uint32_t)atoi(mp) : 4u;
    if (h->pass_out.on && !d_out_map && n_ticks >= 16 && ticks_per_pass > (n_ticks + min_passes - 1) / min_passes)
      ticks_per_pass = (n_ticks + min_passes - 1) / min_passes;
  }
  const uint32_t cells_max = ticks_per_pass * n_lanes;
  // auxiliary cells (second CELT frame of a packet, mb_common.cuh): a pool per pass behind the main cells; a packet that
  // finds it empty is reported as unimplemented
  const uint32_t aux_cap = cells_max / 8u < 1024u ? 1024u : cells_max / 8u;
  const uint32_t cells_all = cells_max + aux_cap;
  // (a lane's time series also holds its auxiliary frames: 5 ms each, at most one per tick)
  const uint32_t series_stride = SER_PREFIX + ticks_per_pass * (MAX_N + 2 * SHORT_MDCT) + SER_SLACK;
  CK(h->rec.ensure((size_t)((cells_all + 31) / 32) * REC_SLOTS * 32 * sizeof(uint32_t)));
  CK(h->en.ensure((size_t)cells_all * EN_STRIDE * sizeof(float)));
  CK(h->meta.ensure((size_t)2 * cells_all * sizeof(FrameMeta)));   // two halves: the clipper of pass k runs beside pass k+1
  CK(h->series.ensure((size_t)n_lanes * CC * series_stride * sizeof(float)));
  CK(h->xbuf.ensure((size_t)cells_all * XBUF_STRIDE * sizeof(float)));
  CK(h->order.ensure(((size_t)2 * cells_max + n_lanes) * sizeof(uint32_t)));   // [2 * ticks][lanes] cells, then the per-lane counts
  CK(h->side_of.ensure((size_t)cells_max * sizeof(uint32_t)));
  CK(h->aux_count.ensure(sizeof(uint32_t)));
  CK(h->aux_main.ensure((size_t)aux_cap * sizeof(uint32_t)));
  if (d_pcm_i16) CK(h->clip_need.ensure((size_t)2 * cells_max * sizeof(uint32_t)));
  const bool ola = !(h->cfg.quirks & Q2_MDCT_ZERO_FILL);   // overlap-add form of the inverse MDCT (mb_celt_imdct.cuh)
  if (ola) CK(h->tails.ensure((size_t)cells_all * CC * TAIL_STRIDE * sizeof(float)));
  float* pcm_f32 = d_pcm_f32;
  if (d_pcm_i16) {  // the i16 path decodes to a float scratch, soft-clips, then rounds (src/opus_decoder.rs:1679-1712)
    CK(h->d_pcm.ensure((size_t)n_packets * pcm_stride * sizeof(float)));
    pcm_f32 = h->d_pcm.as<float>();
  }
  const int cap_per_ch = (int)(pcm_stride / (uint32_t)CC) * h->ds;   // the parser compares 48 kHz frame sizes with it
  // The clipper runs in line with the decode kernels: beside them (its own stream, MOUSIKI_B200_CLIP_STREAM=1) its CTAs share the
  // one wave k_parse of the next pass fits in and double that kernel's time (measured: 58.6 vs 54.4 ms of device time per call).
  cudaStream_t clip_s = getenv("MOUSIKI_B200_CLIP_STREAM") ? h->clip_stream : s;
  // Pass sizes. With the PCM going back to the host pass by pass, only the last pass's copy is exposed: the tail of the
  // call is cut into a short final pass (a quarter of what is left) so that most of the copy hides behind kernels.
  // Multi-frame calls keep the regular grid of passes (the host scheduler aligned frame groups to it).
  h->pass_event_next = 0;
  std::vector<uint32_t> pass_ticks;
  for (uint32_t left = n_ticks; left > 0;) {
    uint32_t nt = left < ticks_per_pass ? left : ticks_per_pass;
    if (h->pass_out.on && !d_out_map && nt == left && nt >= 8) nt -= nt / 4;
    pass_ticks.push_back(nt);
    left -= nt;
  }
  uint32_t t0 = 0;
  for (size_t pi = 0; pi < pass_ticks.size(); t0 += pass_ticks[pi], pi++) {
    const uint32_t nt = pass_ticks[pi];
    const uint32_t cells = nt * n_lanes;
    FrameMeta* const meta_p = h->meta.as<FrameMeta>() + (pi & 1) * (size_t)cells_all;
    const uint32_t cells_run = cells + aux_cap;   // per-cell launches: the pass's main cells, then the auxiliary pool
    AuxPool aux = {cells_max, aux_cap, h->aux_count.as<uint32_t>()};
    uint32_t* const order_p = h->order.as<uint32_t>();
    uint32_t* const order_n_p = order_p + (size_t)2 * cells_max;
    // (only the auxiliary cells' meta, which the clipper of pass pi - 2 never reads)
    k_aux_clear<<<(aux_cap + 255) / 256, 256, 0, s>>>(h->rec.as<uint32_t>(), meta_p, aux);
    const size_t cell0 = (size_t)t0 * n_lanes;
    // identity mapping: shift the per-packet arrays so that cell index == local packet index
    const uint32_t* map = d_pkt_map ? d_pkt_map + cell0 : nullptr;
    const size_t shift = d_pkt_map ? 0 : cell0;
    const uint32_t* offs = d_offsets + shift;
    const uint8_t* flg = d_flags ? d_flags + shift : nullptr;
    int32_t* samples = d_samples + shift;
    uint32_t* range = d_range + shift;
    float* pcm = pcm_f32 ? pcm_f32 + shift * pcm_stride : nullptr;
    h->prof_begin(BD_K_PARSE, s);
    if ((cells + 127) / 128 > 7u * (uint32_t)h->n_sm)
      k_parse<10><<<(cells + 127) / 128, 128, 0, s>>>(d_arena, offs, map, cells, h->rec.as<uint32_t>(), CC, h->cfg.quirks, h->disable_inv,
                                                      cap_per_ch, flg, n_lanes, d_lane_stream, h->states.as<StreamState>(), aux);
    else
      k_parse<7><<<(cells + 127) / 128, 128, 0, s>>>(d_arena, offs, map, cells, h->rec.as<uint32_t>(), CC, h->cfg.quirks, h->disable_inv,
                                                     cap_per_ch, flg, n_lanes, d_lane_stream, h->states.as<StreamState>(), aux);
    h->prof_end(s);
    // fork: k_pvq needs nothing but the records, so it runs beside k_energy / k_silk, which keep one lane per stream
    // busy for a long time and leave most of the machine idle. The two per-stream kernels go to a second stream of the
    // highest priority, so that their few long-lived CTAs become resident at once instead of queueing behind k_pvq's.
    CK(cudaEventRecord(h->ev_fork, s));
    CK(cudaStreamWaitEvent(h->side_stream, h->ev_fork, 0));
    cudaStream_t ss = h->side_stream;
    // the clipper of pass pi - 2 read this half of the frame meta on its own stream: k_energy may only rewrite it afterwards
    if (d_pcm_i16 && pi >= 2) CK(cudaStreamWaitEvent(ss, h->ev_clip_done[pi & 1], 0));
    h->prof_begin(BD_K_ENERGY, ss);
    EnergyParams eq;
    eq.n_lanes = n_lanes; eq.n_ticks = nt; eq.pkt_map = map; eq.rec = h->rec.as<uint32_t>(); eq.st_all = h->states.as<StreamState>();
    eq.lane_stream = d_lane_stream; eq.en = h->en.as<float>(); eq.meta = meta_p; eq.samples_out = samples; eq.final_range_out = range;
    eq.quirks = h->cfg.quirks; eq.flags = flg;
    // The per-stream band pass (k_bands_hyb) is the exact one: it carries the reference's never-cleared fold buffer through every
    // frame of the stream. The all-at-once pass (k_bands) writes back whole footprints, including words a frame never wrote, so a
    // hybrid / redundancy / transition frame that folds from them later may see other data than the reference does: it is an
    // opt-in for small calls of streams known to stay CELT-only (one warp per lane fills the band-pass slots from 3552 lanes on).
    eq.serial_all = !h->parallel_bands || n_lanes >= 148u * 6u * (uint32_t)BANDS_WARPS;
    eq.force_float_row = h->cfg.gain_q8 != 0 || !d_pcm_i16;   // direct int16 from k_silk: only the i16 API without a decode gain
    eq.CC = CC; eq.order = order_p; eq.order_n = order_n_p; eq.side_of = h->side_of.as<uint32_t>(); eq.aux = aux;
    eq.aux_main = h->aux_main.as<uint32_t>();
    k_energy<<<(n_lanes + ENERGY_WARPS - 1) / ENERGY_WARPS, 32 * ENERGY_WARPS, 0, ss>>>(eq);
    h->prof_end(ss);
    CK(cudaEventRecord(h->ev_join, ss));   // the frame meta and band energies are ready: k_plan may go
    // SILK synthesis (SILK-only and hybrid cells), before the CELT tail so that k_deemph can add the CELT bands of hybrid
    // frames on top. SILK-only samples are exact int16 values scaled by 2^-15: on the i16 path soft clip + float2int16
    // is the identity and the lane stores int16 directly; hybrid cells go through the float row.
    h->prof_begin(BD_K_SILK, ss);
    SilkLaneParams sq;
    sq.n_lanes = n_lanes; sq.n_ticks = nt; sq.pkt_map = map; sq.lane_stream = d_lane_stream; sq.rec = h->rec.as<uint32_t>(); sq.meta = meta_p;
    sq.st_all = h->silk_states.as<SilkState>(); sq.silk_hist = h->silk_hist.as<uint32_t>(); sq.silk_exc = h->silk_exc.as<int32_t>();
    sq.n_streams = h->cfg.n_streams; sq.pcm_f32 = pcm; sq.pcm_i16 = d_pcm_i16 ? d_pcm_i16 + shift * pcm_stride : nullptr;
    sq.pcm_stride = pcm_stride; sq.CC = CC; sq.quirks = h->cfg.quirks; sq.gain = h->gain; sq.apply_gain = h->cfg.gain_q8 != 0; sq.out_map = d_out_map;
    sq.xbuf = h->xbuf.as<float>(); sq.side_of = h->side_of.as<uint32_t>();
    k_silk<<<(n_lanes + 63) / 64, 64, 0, ss>>>(sq);
    h->prof_end(ss);
    CK(cudaEventRecord(h->ev_silk, ss));   // the SILK rows are written: k_deemph (hybrid accumulate) and the output stage may go
    h->launches += 1;
    PvqParams pq;
    pq.rec = h->rec.as<uint32_t>(); pq.meta = nullptr; pq.xbuf = h->xbuf.as<float>();
    pq.iy_dbg = d_dbg_iy ? d_dbg_iy + cell0 * 2 * MAX_N : nullptr;
    h->prof_begin(BD_K_PVQ, s);
    k_pvq<<<(cells_run + PVQ_CELLS - 1) / PVQ_CELLS, PVQ_THREADS, 0, s>>>(pq, cells_run, cells, cells_max);
    h->prof_end(s);
    CK(cudaStreamWaitEvent(s, h->ev_join, 0));   // join 1: everything below needs the frame meta
    PlanParams pl;
    pl.rec = h->rec.as<uint32_t>(); pl.en = h->en.as<float>(); pl.meta = meta_p;
    h->prof_begin(BD_K_PLAN, s);
    k_plan<<<(cells_run + 127) / 128, 128, 0, s>>>(pl, cells_run, cells, cells_max);
    h->prof_end(s);
    BandsParams bp;
    bp.rec = h->rec.as<uint32_t>(); bp.en = h->en.as<float>(); bp.meta = meta_p; bp.xbuf = h->xbuf.as<float>();
    bp.x_dbg = d_dbg_x ? d_dbg_x + cell0 * 2 * MAX_N : nullptr;
    bp.fold_all = h->fold.as<float>(); bp.lane_stream = d_lane_stream;
    h->prof_begin(BD_K_BANDS, s);
    if (!eq.serial_all)   // (with the per-stream pass taking every frame, the all-at-once pass would only look at flags and leave)
      k_bands<<<(cells_run + BANDS_WARPS - 1) / BANDS_WARPS, 32 * BANDS_WARPS, 0, s>>>(bp, cells_run, cells, cells_max);
    k_bands_hyb<<<(n_lanes + BANDS_WARPS - 1) / BANDS_WARPS, 32 * BANDS_WARPS, 0, s>>>(bp, n_lanes, order_p, order_n_p, d_lane_stream, h->fold.as<float>());
    h->prof_end(s);
    ImdctParams ip;
    ip.rec = h->rec.as<uint32_t>(); ip.meta = meta_p; ip.xbuf = h->xbuf.as<float>(); ip.series = h->series.as<float>();
    ip.series_stride = series_stride; ip.CC = CC;
    if (ola) { ip.ola = 1; ip.tails = h->tails.as<float>(); }
    ip.ds = h->ds;
    h->prof_begin(BD_K_IMDCT, s);
    {
      const uint32_t want = (cells_run * CC + IMDCT_WARPS - 1) / IMDCT_WARPS, resident = (uint32_t)(h->n_sm * IMDCT_CTAS_PER_SM);
      k_imdct<<<want < resident ? want : resident, 32 * IMDCT_WARPS, sizeof(ImdctCtaSmem), s>>>(ip, cells_run * CC, cells, cells_max);
    }
    h->prof_end(s);
    TailParams tp;
    tp.rec = h->rec.as<uint32_t>(); tp.meta = meta_p; tp.pkt_map = map; tp.lane_stream = d_lane_stream;
    tp.st_all = h->states.as<StreamState>(); tp.hist = h->hist.as<float>(); tp.series = h->series.as<float>();
    tp.series_stride = series_stride; tp.n_lanes = n_lanes; tp.n_ticks = nt; tp.CC = CC; tp.gain = h->gain;
    tp.apply_gain = h->cfg.gain_q8 != 0; tp.pcm_f32 = pcm; tp.pcm_stride = pcm_stride; tp.out_map = d_out_map; tp.xbuf = h->xbuf.as<float>();
    tp.order = order_p; tp.order_n = order_n_p; tp.side_of = h->side_of.as<uint32_t>();
    tp.aux_main = h->aux_main.as<uint32_t>(); tp.aux_base = cells_max;
    if (ola) { tp.ola = 1; tp.tails = h->tails.as<float>(); }
    tp.ds = h->ds;
    uint32_t* clip_need_p = nullptr;
    if (d_pcm_i16 && h->fused_i16) {   // (double-buffered like the frame meta: the clipper of pass pi reads it beside pass pi + 1; s is behind ev_clip_done via join 1)
      clip_need_p = h->clip_need.as<uint32_t>() + (pi & 1) * (size_t)cells_max;
      CK(cudaMemsetAsync(clip_need_p, 0, (size_t)cells * sizeof(uint32_t), s));
      tp.pcm_i16 = d_pcm_i16 + shift * pcm_stride; tp.clip_need = clip_need_p;
    }
    h->prof_begin(BD_K_COMB, s);
    k_comb<<<(n_lanes * CC + COMB_WARPS - 1) / COMB_WARPS, 32 * COMB_WARPS, 0, s>>>(tp);
    k_comb_plc<<<n_lanes, 32, 0, s>>>(tp);   // lanes with a frame concealed by pitch repetition (the others leave after one look at their flags)
    h->launches += 1;
    h->prof_end(s);
    // join 2: k_silk ran beside k_pvq .. k_comb. (Kernels with a large shared-memory footprint, k_bands / k_imdct, only join
    // it on an SM once that SM is idle enough to change its L1 / shared split; giving k_silk the large split up front makes
    // them overlap but costs k_silk its L1: 15.0 -> 18.1 ms, a net loss.)
    CK(cudaStreamWaitEvent(s, h->ev_silk, 0));
    h->prof_begin(BD_K_DEEMPH, s);
    {
      // tile width by the number of warps: wide tiles (more bytes in flight per warp) while warps are scarce
      const uint32_t de_ctas = (n_lanes * CC + 32 * DE_WARPS - 1) / (32 * DE_WARPS);
      if (de_ctas <= 4u * 148u) k_deemph<128><<<de_ctas, 32 * DE_WARPS, DE_WARPS * 2 * 32 * (128 + 4) * sizeof(float), s>>>(tp);
      else k_deemph<32><<<de_ctas, 32 * DE_WARPS, DE_WARPS * 2 * 32 * (32 + 4) * sizeof(float), s>>>(tp);
    }
    h->prof_end(s);
    k_fade<<<(cells * CC + 127) / 128, 128, 0, s>>>(tp, cells);
    h->launches += eq.serial_all ? 9 : 10;   // parse, energy, pvq, plan, (bands,) bands_hyb, imdct, comb, deemph, fade
    h->mark("kernels", pi, s);
    // ship this pass's PCM (cells [cell0 + c0, cell0 + c0 + nc)) on the copy stream while later kernels compute
    auto next_event = [&](cudaEvent_t* out) -> int {
      const size_t k = h->pass_event_next++;
      while (h->pass_events.size() <= k) { cudaEvent_t e; CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->pass_events.push_back(e); }
      *out = h->pass_events[k];
      return BD_OK;
    };
    // the stream the pass's PCM is final on: the clipper's for int16 output, the main stream for float output
    cudaStream_t fin = d_pcm_i16 ? clip_s : s;
    auto ship = [&](size_t c0, size_t nc) -> int {
      cudaEvent_t ev;
      { int rc = next_event(&ev); if (rc != BD_OK) return rc; }
      CK(cudaEventRecord(ev, fin));
      CK(cudaStreamWaitEvent(h->copy_stream, ev, 0));
      const size_t off = (cell0 + c0) * pcm_stride * h->pass_out.elem, bytes = nc * pcm_stride * h->pass_out.elem;
      h->mark("final", pi, fin);
      CK(cudaMemcpyAsync(h->pass_out.host_pcm + off, h->pass_out.dev_pcm + off, bytes, cudaMemcpyDeviceToHost, h->copy_stream));
      h->mark("copied", pi, h->copy_stream);
      return BD_OK;
    };
    if (d_pcm_i16) {
      // The clipper walks a stream's ticks in order, so it can run tick range by tick range; with the PCM going back to the
      // host, each range is shipped as soon as it is clipped and only the last range's copy stays exposed.
      // The clipper has its own stream: it only touches the pass's float rows, the int16 rows and the streams' clipper
      // memory, so pass k is clipped while the main stream already decodes pass k+1 (the frame meta is double-buffered).
      {
        cudaEvent_t ev;
        { int rc = next_event(&ev); if (rc != BD_OK) return rc; }
        CK(cudaEventRecord(ev, s));
        CK(cudaStreamWaitEvent(clip_s, ev, 0));
      }
      const uint32_t n_chunks = h->pass_out.on ? (nt < 4u ? nt : 4u) : 1u;
      for (uint32_t c = 0; c < n_chunks; c++) {
        const uint32_t ta = (uint32_t)((uint64_t)nt * c / n_chunks), tb = (uint32_t)((uint64_t)nt * (c + 1) / n_chunks);
        const size_t c0 = (size_t)ta * n_lanes;
        h->prof_begin(BD_K_CLIP_I16, fin);
        k_clip_i16<<<(n_lanes + CLIP_WARPS - 1) / CLIP_WARPS, 32 * CLIP_WARPS, 0, fin>>>(
            n_lanes, tb - ta, map ? map + c0 : nullptr, d_lane_stream, meta_p + c0, map ? samples : samples + c0,
            h->states.as<StreamState>(), map ? pcm : pcm + c0 * pcm_stride, d_pcm_i16 + (map ? 0 : (shift + c0) * pcm_stride), pcm_stride, CC,
            d_out_map, flg, clip_need_p ? clip_need_p + c0 : nullptr, h->clip_serial);
        h->prof_end(fin);
        h->launches += 1;
        if (h->pass_out.on) { int rc = ship(c0, (size_t)(tb - ta) * n_lanes); if (rc != BD_OK) return rc; }
      }
      CK(cudaEventRecord(h->ev_clip_done[pi & 1], fin));
    } else {
      k_clear_softclip<<<(n_lanes + 63) / 64, 64, 0, s>>>(n_lanes, d_lane_stream, h->states.as<StreamState>());
      h->launches += 1;
      if (h->pass_out.on) { int rc = ship(0, cells); if (rc != BD_OK) return rc; }
    }
  }
  if (d_pcm_i16) {  // join the clipper back into the caller's stream
    const size_t k = h->pass_event_next++;
    while (h->pass_events.size() <= k) { cudaEvent_t e; CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); h->pass_events.push_back(e); }
    CK(cudaEventRecord(h->pass_events[k], clip_s));
    CK(cudaStreamWaitEvent(s, h->pass_events[k], 0));
  }
  CK(cudaGetLastError());
  return BD_OK;
}

int bd_decode_grid_dev(bd_handle* h, const uint8_t* d_arena, const uint32_t* d_offsets, uint32_t n_lanes, uint32_t n_ticks,
                       const uint32_t* d_lane_stream, float* d_pcm_f32, int16_t* d_pcm_i16, uint32_t pcm_stride, int32_t* d_samples_out,
                       uint32_t* d_final_range_out, void* cuda_stream) {
  if (!h || !d_arena || !d_offsets || !d_samples_out || !d_final_range_out) return BD_ERR_BAD_ARG;
  if (d_pcm_f32 && d_pcm_i16) return BD_ERR_BAD_ARG;
  if (!d_lane_stream && n_lanes > h->cfg.n_streams) return BD_ERR_BAD_ARG;
  CK(cudaSetDevice(h->device));
  h->launches = 0;
  return decode_grid_device(h, d_arena, d_offsets, nullptr, nullptr, n_lanes, n_ticks, d_lane_stream, d_pcm_f32, d_pcm_i16, pcm_stride,
                            d_samples_out, d_final_range_out, n_lanes * n_ticks, (cudaStream_t)cuda_stream, nullptr, nullptr);
}

// out_map (host, 2 words per packet: PCM row, element offset in the row) and n_rows come from the multi-frame split
// (mb_host_packets.h): the "packets" are then frames, several of which share one PCM row. Null: packet p owns row p.
static int decode_host_impl(bd_handle* h, const uint8_t* arena, const uint32_t* offsets, const uint32_t* stream_ids, uint32_t n_packets,
                            const uint8_t* flags, float* pcm_f32, int16_t* pcm_i16, uint32_t pcm_stride, int32_t* samples_out,
                            uint32_t* final_range_out, const uint32_t* out_map, uint32_t n_rows) {
  if (!h || !arena || !offsets || !stream_ids || !samples_out || !final_range_out || (!pcm_f32 && !pcm_i16)) return BD_ERR_BAD_ARG;
  if (n_packets == 0) return BD_OK;
  const auto t_entry = std::chrono::steady_clock::now();
  h->timeline_on = getenv("MOUSIKI_B200_TIMING") != nullptr;
  for (auto& m : h->timeline) cudaEventDestroy(m.second);
  h->timeline.clear();
  CK(cudaSetDevice(h->device));
  h->launches = 0;
  cudaStream_t s = h->own_stream;
  const uint32_t S = h->cfg.n_streams;
  // ---- schedule: lane = order of first appearance of a stream in this call, tick = occurrence number
  // (a serving loop sends the same stream-id pattern call after call: the schedule of the previous call is kept and reused when
  //  the ids compare equal — one pass over them instead of three with scattered accesses)
  const bool sched_hit = !out_map && h->sched_n == n_packets && h->sched_ids.size() == n_packets &&
                         memcmp(h->sched_ids.data(), stream_ids, (size_t)n_packets * 4) == 0;
  uint32_t n_lanes = 0, n_ticks = 0;
  bool identity = false;
  if (sched_hit) { n_lanes = h->sched_lanes; n_ticks = h->sched_ticks; identity = h->sched_identity; }
  else {
  h->sched_n = 0;
  if (h->h_epoch.size() != S) { h->h_epoch.assign(S, 0); h->h_lane_of_stream.assign(S, 0); h->h_count.assign(S, 0); h->epoch = 0; }
  h->epoch++;
  if (h->epoch == 0) { std::fill(h->h_epoch.begin(), h->h_epoch.end(), 0u); h->epoch = 1; }
  h->h_lane_stream.clear();
  for (uint32_t p = 0; p < n_packets; p++) {
    uint32_t sid = stream_ids[p];
    if (sid >= S) return BD_ERR_BAD_ARG;
    if (h->h_epoch[sid] != h->epoch) {
      h->h_epoch[sid] = h->epoch; h->h_count[sid] = 0;
      h->h_lane_of_stream[sid] = (uint32_t)h->h_lane_stream.size();
      h->h_lane_stream.push_back(sid);
    }
  }
  n_lanes = (uint32_t)h->h_lane_stream.size();
  // tick = occurrence number of the stream in this call. The frames of one multi-frame packet must not straddle a
  // pass boundary (k_clip_i16 soft-clips the packet as a whole): the group moves to the next pass, leaving empty cells.
  uint32_t tpp = h->max_frames_per_pass / n_lanes;
  if (tpp == 0) tpp = 1;
  h->h_tick.resize(n_packets);
  for (uint32_t p = 0; p < n_packets; p++) {
    const uint32_t sid = stream_ids[p];
    uint32_t cur = h->h_count[sid];
    if (out_map && !(flags[p] & mbh::VF_CONTINUATION)) {
      uint32_t g = 1;
      while (p + g < n_packets && (flags[p + g] & mbh::VF_CONTINUATION)) g++;
      if (g > tpp) return BD_ERR_BAD_ARG;   // max_frames_per_pass too small for this many lanes
      if (cur % tpp + g > tpp) cur = (cur / tpp + 1) * tpp;
    }
    h->h_tick[p] = cur;
    h->h_count[sid] = cur + 1;
    if (cur + 1 > n_ticks) n_ticks = cur + 1;
  }
  identity = n_ticks == 1 && !out_map;  // every stream at most once: cell == packet
  if (!identity) {
    h->h_pkt_map.assign((size_t)n_lanes * n_ticks, 0xffffffffu);
    bool grid_order = (size_t)n_packets == (size_t)n_lanes * n_ticks && !out_map;
    for (uint32_t p = 0; p < n_packets; p++) {
      const size_t cell = (size_t)h->h_tick[p] * n_lanes + h->h_lane_of_stream[stream_ids[p]];
      h->h_pkt_map[cell] = p;
      grid_order = grid_order && cell == p;
    }
    identity = grid_order;  // packets already arrive tick-major in lane order: no indirection needed
  }
  if (!out_map) {
    h->sched_ids.assign(stream_ids, stream_ids + n_packets);
    h->sched_n = n_packets; h->sched_lanes = n_lanes; h->sched_ticks = n_ticks; h->sched_identity = identity;
  }
  }
  const uint32_t* d_map = nullptr;
  if (!identity) {
    CK(h->d_pkt_map.ensure(h->h_pkt_map.size() * 4));
    CK(cudaMemcpyAsync(h->d_pkt_map.p, h->h_pkt_map.data(), h->h_pkt_map.size() * 4, cudaMemcpyHostToDevice, s));
    d_map = h->d_pkt_map.as<uint32_t>();
  }
  const size_t arena_bytes = offsets[n_packets];
  CK(h->d_arena.ensure(arena_bytes + 16));
  CK(h->d_offsets.ensure((size_t)(n_packets + 1) * 4));
  CK(h->d_lane_stream.ensure((size_t)n_lanes * 4));
  CK(h->d_samples.ensure((size_t)n_packets * 4));
  CK(h->d_range.ensure((size_t)n_packets * 4));
  h->mark("start", 0, s);
  if (h->arena_staged != arena) CK(cudaMemcpyAsync(h->d_arena.p, arena, arena_bytes, cudaMemcpyHostToDevice, s));
  h->arena_staged = nullptr;
  CK(cudaMemcpyAsync(h->d_offsets.p, offsets, (size_t)(n_packets + 1) * 4, cudaMemcpyHostToDevice, s));
  CK(cudaMemcpyAsync(h->d_lane_stream.p, h->h_lane_stream.data(), (size_t)n_lanes * 4, cudaMemcpyHostToDevice, s));
  const uint8_t* d_flags = nullptr;
  // an all-ones flag array is still needed in identity mode to tell "lost packet" from "empty cell"
  std::vector<uint8_t> zero_flags;
  if (!flags) { zero_flags.assign(n_packets, 0); flags = zero_flags.data(); }
  CK(h->d_flags.ensure(n_packets));
  CK(cudaMemcpyAsync(h->d_flags.p, flags, n_packets, cudaMemcpyHostToDevice, s));
  d_flags = h->d_flags.as<uint8_t>();
  const uint32_t* d_out_map = nullptr;
  if (out_map) {
    CK(h->d_out_map.ensure((size_t)n_packets * 8));
    CK(cudaMemcpyAsync(h->d_out_map.p, out_map, (size_t)n_packets * 8, cudaMemcpyHostToDevice, s));
    d_out_map = h->d_out_map.as<uint32_t>();
  }
  const size_t pcm_elems = (size_t)n_rows * pcm_stride;
  float* d_f32 = nullptr;
  int16_t* d_i16 = nullptr;
  if (pcm_f32) {
    // the float output reuses d_pcm; the i16 path uses d_pcm as its float scratch and d_pcm16 as the result
    CK(h->d_pcm.ensure(pcm_elems * sizeof(float)));
    d_f32 = h->d_pcm.as<float>();
  } else {
    CK(h->d_pcm16.ensure(pcm_elems * sizeof(int16_t)));
    d_i16 = h->d_pcm16.as<int16_t>();
  }
  int32_t* d_dbg_iy = nullptr;
  float* d_dbg_x = nullptr;
  const size_t n_cells = (size_t)n_lanes * n_ticks;
  if (h->dbg_max) {
    if (h->dbg_iy_host) { CK(h->d_dbg_iy.ensure(n_cells * 2 * MAX_N * 4)); CK(cudaMemsetAsync(h->d_dbg_iy.p, 0, n_cells * 2 * MAX_N * 4, s)); d_dbg_iy = h->d_dbg_iy.as<int32_t>(); }
    if (h->dbg_x_host) { CK(h->d_dbg_x.ensure(n_cells * 2 * MAX_N * 4)); CK(cudaMemsetAsync(h->d_dbg_x.p, 0, n_cells * 2 * MAX_N * 4, s)); d_dbg_x = h->d_dbg_x.as<float>(); }
  }
  CK(cudaMemsetAsync(h->d_samples.p, 0, (size_t)n_packets * 4, s));
  // cell order == packet order: each pass's PCM is one contiguous block, copied out on a second stream as soon as
  // the pass is done
  const bool overlap_out = identity && !h->want_i24;
  h->pass_out.on = overlap_out;
  h->pass_event_next = 0;
  h->pass_out.host_pcm = pcm_f32 ? (char*)pcm_f32 : (char*)pcm_i16;
  h->pass_out.dev_pcm = pcm_f32 ? (const char*)d_f32 : (const char*)d_i16;
  h->pass_out.elem = pcm_f32 ? sizeof(float) : sizeof(int16_t);
  const auto t_sched = std::chrono::steady_clock::now();
  int rc = decode_grid_device(h, h->d_arena.as<uint8_t>(), h->d_offsets.as<uint32_t>(), d_map, d_flags, n_lanes, n_ticks,
                              h->d_lane_stream.as<uint32_t>(), d_f32, d_i16, pcm_stride, h->d_samples.as<int32_t>(),
                              h->d_range.as<uint32_t>(), n_rows, s, d_dbg_iy, d_dbg_x, d_out_map);
  h->pass_out.on = false;
  if (rc != BD_OK) { cudaStreamSynchronize(h->copy_stream); h->after_enqueue = nullptr; return rc; }
  if (h->want_i24 && pcm_elems) {
    k_f32_to_i24<<<(unsigned)((pcm_elems + 255) / 256), 256, 0, s>>>(d_f32, pcm_elems);
    h->launches += 1;
  }
  if (!overlap_out) {
    if (pcm_f32) CK(cudaMemcpyAsync(pcm_f32, d_f32, pcm_elems * sizeof(float), cudaMemcpyDeviceToHost, s));
    else CK(cudaMemcpyAsync(pcm_i16, d_i16, pcm_elems * sizeof(int16_t), cudaMemcpyDeviceToHost, s));
  }
  CK(cudaMemcpyAsync(samples_out, h->d_samples.p, (size_t)n_packets * 4, cudaMemcpyDeviceToHost, s));
  CK(cudaMemcpyAsync(final_range_out, h->d_range.p, (size_t)n_packets * 4, cudaMemcpyDeviceToHost, s));
  const auto t_enq = std::chrono::steady_clock::now();
  if (h->after_enqueue) { h->after_enqueue(); h->after_enqueue = nullptr; }   // host bookkeeping that nothing on the device waits for
  CK(cudaStreamSynchronize(s));
  const auto t_dev = std::chrono::steady_clock::now();
  CK(cudaStreamSynchronize(h->copy_stream));
  if (getenv("MOUSIKI_B200_TIMING")) {
    auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    fprintf(stderr, "[mousiki_b200] host schedule + staging %.2f ms, enqueue %.2f ms, wait for kernels %.2f ms, wait for the last copy %.2f ms\n",
            ms(t_entry, t_sched), ms(t_sched, t_enq), ms(t_enq, t_dev), ms(t_dev, std::chrono::steady_clock::now()));
    unsigned long long cs[2] = {0, 0};
    if (cudaMemcpyFromSymbol(cs, g_clip_stats, sizeof(cs)) == cudaSuccess)
      fprintf(stderr, "[mousiki_b200] clipper since load: %llu packets skipped (int16 stored by k_deemph), %llu worked on\n", cs[0], cs[1]);
    for (size_t i = 1; i < h->timeline.size(); i++) {
      float t = 0;
      cudaEventElapsedTime(&t, h->timeline[0].second, h->timeline[i].second);
      fprintf(stderr, "[mousiki_b200]   %-12s %8.2f ms\n", h->timeline[i].first.c_str(), t);
    }
  }
  if (h->dbg_max) {
    // debug captures are indexed by grid cell; scatter them back to packet order
    std::vector<int32_t> iy;
    std::vector<float> xs;
    if (d_dbg_iy) { iy.resize(n_cells * 2 * MAX_N); CK(cudaMemcpy(iy.data(), d_dbg_iy, iy.size() * 4, cudaMemcpyDeviceToHost)); }
    if (d_dbg_x) { xs.resize(n_cells * 2 * MAX_N); CK(cudaMemcpy(xs.data(), d_dbg_x, xs.size() * 4, cudaMemcpyDeviceToHost)); }
    for (size_t cidx = 0; cidx < n_cells; cidx++) {
      uint32_t p = identity ? (uint32_t)cidx : h->h_pkt_map[cidx];
      if (p == 0xffffffffu || p >= h->dbg_max) continue;
      if (d_dbg_iy) memcpy(h->dbg_iy_host + (size_t)p * 2 * MAX_N, iy.data() + cidx * 2 * MAX_N, 2 * MAX_N * 4);
      if (d_dbg_x) memcpy(h->dbg_x_host + (size_t)p * 2 * MAX_N, xs.data() + cidx * 2 * MAX_N, 2 * MAX_N * 4);
    }
  }
  return BD_OK;
}

// Packets with more than one frame are cut into frames here (mb_host_packets.h); a call without any takes the
// zero-copy path above.
static int decode_host(bd_handle* h, const uint8_t* arena, const uint32_t* offsets, const uint32_t* stream_ids, uint32_t n_packets,
                       const uint8_t* flags, float* pcm_f32, int16_t* pcm_i16, uint32_t pcm_stride, int32_t* samples_out,
                       uint32_t* final_range_out) {
  if (!h || !arena || !offsets || !stream_ids || !samples_out || !final_range_out || (!pcm_f32 && !pcm_i16)) return BD_ERR_BAD_ARG;
  const auto t_in = std::chrono::steady_clock::now();
  h->after_enqueue = nullptr;
  // The packet bytes start towards the device before the host looks at them: the copy runs beside the walk and the schedule below
  // (a call that turns out to need cutting into frames sends its re-cut arena instead; the early copy is then simply overwritten).
  h->arena_staged = nullptr;
  if (n_packets && offsets[n_packets] >= (1u << 20)) {
    CK(cudaSetDevice(h->device));
    CK(h->d_arena.ensure((size_t)offsets[n_packets] + 16));
    CK(cudaMemcpyAsync(h->d_arena.p, arena, offsets[n_packets], cudaMemcpyHostToDevice, h->own_stream));
    h->arena_staged = arena;
  }
  if (h->track.size() != h->cfg.n_streams) h->track.assign(h->cfg.n_streams, mbh::StreamTrack());
  // One walk over the packets (each TOC byte is a cache miss of its own in a large arena): stream ids in range, does any packet
  // need cutting into frames, and the TOC bytes kept aside for the note below.
  bool split = false;
  h->h_toc.resize(n_packets);
  {
    // (large calls: the walk is spread over a few host threads — it is latency-bound on the TOC misses and sits in front of all device work)
    const uint32_t n_thr = n_packets >= (1u << 16) ? 8u : 1u;
    uint8_t res[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // bit 0: a packet needs cutting, bit 1: a stream id out of range
    auto walk = [&](uint32_t k) {
      const uint32_t p0 = (uint32_t)((uint64_t)n_packets * k / n_thr), p1 = (uint32_t)((uint64_t)n_packets * (k + 1) / n_thr);
      uint8_t r = 0;
      for (uint32_t p = p0; p < p1; p++) {
        if (stream_ids[p] >= h->cfg.n_streams) r |= 2;
        const uint32_t n = offsets[p + 1] - offsets[p];
        const uint8_t toc = n ? arena[offsets[p]] : 0;
        h->h_toc[p] = toc;
        if ((n && (toc & 3u)) || (flags && ((flags[p] & 3u) || n == 0))) r |= 1;   // (mbh::needs_split)
      }
      res[k] = r;
    };
    std::vector<std::thread> thr;
    for (uint32_t k = 1; k < n_thr; k++) thr.emplace_back(walk, k);
    walk(0);
    for (auto& t : thr) t.join();
    uint8_t all = 0;
    for (uint32_t k = 0; k < n_thr; k++) all |= res[k];
    if (all & 2) return BD_ERR_BAD_ARG;
    split = (all & 1) != 0;
  }
  if (n_packets == 0 || !split) {
    // (no split: still remember what the decoders took in, for a later lost packet — after the device work is queued, beside it)
    h->after_enqueue = [=]() {
      for (uint32_t p = 0; p < n_packets; p++) {
        const uint32_t n = offsets[p + 1] - offsets[p];
        if (n == 0 || (flags && (flags[p] & 3u))) continue;
        const uint8_t toc = h->h_toc[p];
        if (mbh::toc_frame_samples(toc) > (int)(pcm_stride / (uint32_t)h->CC) * h->ds || n - 1 > 1275) continue;
        h->track[stream_ids[p]].frame_size = mbh::toc_frame_samples(toc);
        h->track[stream_ids[p]].mode = (toc & 0x80) ? 1002 : ((toc & 0x60) == 0x60 ? 1001 : 1000);
      }
    };
    if (getenv("MOUSIKI_B200_TIMING"))
      fprintf(stderr, "[mousiki_b200] packet walk %.2f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_in).count());
    const int rc0 = decode_host_impl(h, arena, offsets, stream_ids, n_packets, flags, pcm_f32, pcm_i16, pcm_stride, samples_out, final_range_out,
                                     nullptr, n_packets);
    if (h->after_enqueue) { h->after_enqueue(); h->after_enqueue = nullptr; }   // (a call that failed before it got there)
    return rc0;
  }
  mbh::Split& S = h->split;
  // (the splitter counts in 48 kHz samples: capacity scaled up, element offsets of the frames scaled back down)
  mbh::split_batch(arena, offsets, stream_ids, flags, n_packets, (int)(pcm_stride / (uint32_t)h->CC) * h->ds, h->CC, &S, h->track.data());
  if (h->ds > 1)
    for (size_t i = 1; i < S.out_map.size(); i += 2) S.out_map[i] /= (uint32_t)h->ds;
  const uint32_t nv = (uint32_t)S.stream_ids.size();
  h->v_samples.resize(nv);
  h->v_range.resize(nv);
  int rc = decode_host_impl(h, S.arena.data(), S.offsets.data(), S.stream_ids.data(), nv, S.flags.data(), pcm_f32, pcm_i16, pcm_stride,
                            h->v_samples.data(), h->v_range.data(), S.out_map.data(), n_packets);
  if (rc != BD_OK) return rc;
  mbh::merge_results(S, n_packets, h->v_samples.data(), h->v_range.data(), samples_out, final_range_out);
  return BD_OK;
}

int bd_decode_f32(bd_handle* h, const uint8_t* arena, const uint32_t* offsets, const uint32_t* stream_ids, uint32_t n_packets,
                  const uint8_t* flags, float* pcm, uint32_t pcm_stride, int32_t* samples_out, uint32_t* final_range_out) {
  return decode_host(h, arena, offsets, stream_ids, n_packets, flags, pcm, nullptr, pcm_stride, samples_out, final_range_out);
}
int bd_decode_i24(bd_handle* h, const uint8_t* arena, const uint32_t* offsets, const uint32_t* stream_ids, uint32_t n_packets,
                  const uint8_t* flags, int32_t* pcm, uint32_t pcm_stride, int32_t* samples_out, uint32_t* final_range_out) {
  if (!h) return BD_ERR_BAD_ARG;
  h->want_i24 = true;
  const int rc = decode_host(h, arena, offsets, stream_ids, n_packets, flags, (float*)pcm, nullptr, pcm_stride, samples_out, final_range_out);
  h->want_i24 = false;
  return rc;
}
int bd_decode_i16(bd_handle* h, const uint8_t* arena, const uint32_t* offsets, const uint32_t* stream_ids, uint32_t n_packets,
                  const uint8_t* flags, int16_t* pcm, uint32_t pcm_stride, int32_t* samples_out, uint32_t* final_range_out) {
  return decode_host(h, arena, offsets, stream_ids, n_packets, flags, nullptr, pcm, pcm_stride, samples_out, final_range_out);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------ allocation-free entry point
// bd_parse_noalloc: the entropy-parse kernel (phase 1) over a small batch with its record scratch in a file-scope device array
// instead of the handle's cudaMalloc'ed buffers, for callers that cannot allocate (CUDA graph capture, tight serving loops).
// No counterpart in the reference. The array is shared by every call on the GPU
// and keeps the last call's contents: nothing in it has to start at zero (a lane only reads back the words its own parse has just
// written), and calls that overlap on one GPU race on it.
static __device__ __align__(256) uint32_t g_noalloc_rec[(size_t)(BD_NOALLOC_MAX_PACKETS / 32) * REC_SLOTS * 32];   // 12.2 MB

// k_parse for stand-alone packets (every packet its own fresh stream), then the per-packet results straight from the record
__global__ void __launch_bounds__(128, 7) k_parse_noalloc(const uint8_t* __restrict__ arena, const uint32_t* __restrict__ offsets,
                                                        uint32_t n_packets, uint32_t* rec, int CC, uint32_t quirks, int disable_inv,
                                                        int cap_per_ch, int32_t* __restrict__ samples_out,
                                                        uint32_t* __restrict__ final_range_out) {
  const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n_packets) return;
  uint32_t* r = rec + rec_index(f, 0);
  int32_t samples = MB_UNIMPLEMENTED;
  uint32_t range = 0;
  const uint32_t off = offsets[f], len = offsets[f + 1] - off;
  if (len >= 1 && (arena[off] & 0x80u) && !(arena[off] & 3u)) {   // CELT-only, one frame: final_range is the parse's alone
    parse_cell(arena, offsets, f, r, CC, quirks, disable_inv, cap_per_ch, nullptr);
    const uint32_t status = r[S_STATUS * 32];
    const int code = MB_ST_CODE(status);
    if (MB_ST_LOST(status)) samples = (int32_t)r[S_RANGE * 32];   // frame of at most one byte: concealed, no range
    else if (!MB_ST_DECODED(status) || code < 0) samples = code;
    else { samples = SHORT_MDCT << (int)MB_H0_LM(r[S_HDR0 * 32]); range = r[S_RANGE * 32]; }
  }
  samples_out[f] = samples;
  final_range_out[f] = range;
}

namespace {
std::mutex g_noalloc_mu;
uint32_t* g_noalloc_ptr[64] = {nullptr};   // per device: address of g_noalloc_rec on that device
}  // namespace

extern "C" int bd_parse_noalloc(int device, const uint8_t* d_arena, const uint32_t* d_offsets, uint32_t n_packets, int channels,
                                uint32_t quirks, int frame_size, int32_t* d_samples_out, uint32_t* d_final_range_out, void* cuda_stream) {
  if (device < 0 || device >= 64 || !d_arena || !d_offsets || !d_samples_out || !d_final_range_out) return BD_ERR_BAD_ARG;
  if (channels != 1 && channels != 2) return BD_ERR_BAD_ARG;
  if (n_packets > BD_NOALLOC_MAX_PACKETS) return BD_ERR_TOO_LARGE;   // beyond the static scratch: split the batch
  if (n_packets == 0) return BD_OK;
  if (cudaSetDevice(device) != cudaSuccess) return BD_ERR_CUDA;
  uint32_t* rec = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
      // in host code `g_noalloc_rec` is only a handle of the symbol: the device address comes from the runtime, per device
      void* p = nullptr;
      if (cudaGetSymbolAddress(&p, g_noalloc_rec) != cudaSuccess || !p) return BD_ERR_CUDA;
      g_noalloc_ptr[device] = (uint32_t*)p;
    }
    rec = g_noalloc_ptr[device];
  }
  k_parse_noalloc<<<(n_packets + 127) / 128, 128, 0, (cudaStream_t)cuda_stream>>>(d_arena, d_offsets, n_packets, rec, channels, quirks,
                                                                                channels == 1, frame_size, d_samples_out,
                                                                                d_final_range_out);
  return cudaGetLastError() == cudaSuccess ? BD_OK : BD_ERR_CUDA;
}

// ------------------------------------------------------------------------------------------ row A2: lightweight decoder
struct bd_lite_handle {
  int device = 0;
  uint32_t n_streams = 0;
  cudaStream_t stream = nullptr;
  DevBuf st, arena, offsets, pcm, status;
};

extern "C" int bd_lite_create(int device, uint32_t n_streams, bd_lite_handle** out) {
  if (!out || n_streams == 0) return BD_ERR_BAD_ARG;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) return BD_ERR_CUDA;
  if (cudaSetDevice(device) != cudaSuccess) return BD_ERR_CUDA;
  bd_lite_handle* h = new (std::nothrow) bd_lite_handle();
  if (!h) return BD_ERR_ALLOC;
  h->device = device; h->n_streams = n_streams;
  if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return BD_ERR_CUDA; }
  if (h->st.ensure((size_t)n_streams * sizeof(LiteState)) != cudaSuccess) { bd_lite_destroy(h); return BD_ERR_ALLOC; }
  k_lite_reset<<<(n_streams + 127) / 128, 128, 0, h->stream>>>(h->st.as<LiteState>(), n_streams);
  if (cudaStreamSynchronize(h->stream) != cudaSuccess || cudaGetLastError() != cudaSuccess) { bd_lite_destroy(h); return BD_ERR_CUDA; }
  *out = h;
  return BD_OK;
}

extern "C" void bd_lite_destroy(bd_lite_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  if (h->stream) { cudaStreamSynchronize(h->stream); cudaStreamDestroy(h->stream); }
  h->st.release(); h->arena.release(); h->offsets.release(); h->pcm.release(); h->status.release();
  delete h;
}

extern "C" int bd_lite_reset(bd_lite_handle* h) {
  if (!h) return BD_ERR_BAD_ARG;
  if (cudaSetDevice(h->device) != cudaSuccess) return BD_ERR_CUDA;
  k_lite_reset<<<(h->n_streams + 127) / 128, 128, 0, h->stream>>>(h->st.as<LiteState>(), h->n_streams);
  return cudaStreamSynchronize(h->stream) == cudaSuccess && cudaGetLastError() == cudaSuccess ? BD_OK : BD_ERR_CUDA;
}

namespace {
int lite_decode(bd_lite_handle* h, const uint8_t* arena, const uint32_t* offsets, uint32_t n_lanes, uint32_t n_ticks, void* pcm, int32_t* status,
                int mode) {
  if (!h || !offsets || !pcm || !status || n_lanes == 0 || n_lanes > h->n_streams) return BD_ERR_BAD_ARG;
  if (n_ticks == 0) return BD_OK;
  const size_t n = (size_t)n_lanes * n_ticks, bytes = offsets[n], elem = mode == 1 ? 2 : 4;
  if (bytes && !arena) return BD_ERR_BAD_ARG;
  if (cudaSetDevice(h->device) != cudaSuccess) return BD_ERR_CUDA;
  if (h->arena.ensure(bytes + 16) != cudaSuccess || h->offsets.ensure((n + 1) * 4) != cudaSuccess || h->pcm.ensure(n * LITE_OUT * elem) != cudaSuccess ||
      h->status.ensure(n * 4) != cudaSuccess)
    return BD_ERR_ALLOC;
  cudaStream_t s = h->stream;
  if (bytes && cudaMemcpyAsync(h->arena.p, arena, bytes, cudaMemcpyHostToDevice, s) != cudaSuccess) return BD_ERR_CUDA;
  if (cudaMemcpyAsync(h->offsets.p, offsets, (n + 1) * 4, cudaMemcpyHostToDevice, s) != cudaSuccess) return BD_ERR_CUDA;
  LiteParams P;
  P.arena = h->arena.as<uint8_t>(); P.offsets = h->offsets.as<uint32_t>(); P.st = h->st.as<LiteState>(); P.pcm = h->pcm.p;
  P.status = h->status.as<int32_t>(); P.n_lanes = n_lanes; P.n_ticks = n_ticks; P.mode = mode;
  k_lite<<<(n_lanes + 63) / 64, 64, 0, s>>>(P);
  if (cudaGetLastError() != cudaSuccess) return BD_ERR_CUDA;
  if (cudaMemcpyAsync(pcm, h->pcm.p, n * LITE_OUT * elem, cudaMemcpyDeviceToHost, s) != cudaSuccess) return BD_ERR_CUDA;
  if (cudaMemcpyAsync(status, h->status.p, n * 4, cudaMemcpyDeviceToHost, s) != cudaSuccess) return BD_ERR_CUDA;
  return cudaStreamSynchronize(s) == cudaSuccess ? BD_OK : BD_ERR_CUDA;
}
}  // namespace

extern "C" int bd_lite_decode_f32(bd_lite_handle* h, const uint8_t* arena, const uint32_t* offsets, uint32_t n_lanes, uint32_t n_ticks, float* pcm,
                                  int32_t* status) {
  return lite_decode(h, arena, offsets, n_lanes, n_ticks, pcm, status, 0);
}
extern "C" int bd_lite_decode_i16(bd_lite_handle* h, const uint8_t* arena, const uint32_t* offsets, uint32_t n_lanes, uint32_t n_ticks, int16_t* pcm,
                                  int32_t* status) {
  return lite_decode(h, arena, offsets, n_lanes, n_ticks, pcm, status, 1);
}
extern "C" int bd_lite_decode_i24(bd_lite_handle* h, const uint8_t* arena, const uint32_t* offsets, uint32_t n_lanes, uint32_t n_ticks, int32_t* pcm,
                                  int32_t* status) {
  return lite_decode(h, arena, offsets, n_lanes, n_ticks, pcm, status, 2);
}
