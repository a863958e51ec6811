// TEST INFRASTRUCTURE ONLY. Single-threaded host build of the kernel bodies in
// mousiki_b200/csrc/mb_*.cuh (compiled with -DMB_EMUL, team size 1) so that the CPU test-suite can
// check the product's algorithmic logic against the oracle where no GPU exists. Nothing in the
// product library links or calls this file; GPU parity proper is tests/test_gpu_parity.py.
#define MB_EMUL 1
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../mousiki_b200/csrc/mb_host_packets.h"
#include "../../mousiki_b200/csrc/mb_celt_imdct.cuh"
#include "../../mousiki_b200/csrc/mb_celt_tail.cuh"
#include "../../mousiki_b200/csrc/mb_celt_plc.cuh"
#include "../../mousiki_b200/csrc/mb_frame.cuh"
#include "../../mousiki_b200/csrc/mb_silk_out.cuh"
#include "../../mousiki_b200/csrc/mb_kat.cuh"
#include "../../mousiki_b200/csrc/mb_lite.cuh"

using namespace mb;

extern "C" {

// unit known-answer hooks over the kernel bodies (mb_kat.cuh), team of one
int emu_kat(int what, const uint32_t* in, uint32_t n_in, uint32_t* out, uint32_t n_out) {
  static KatSmem sm;
  uint8_t bytes[64];
  Team tm = {0, 1};
  return kat_run(tm, &sm, what, in, n_in, out, n_out, bytes, imdct_tables_host());
}

// Row A2: the body of k_lite (mb_lite.cuh), one stream after the other; mode 0 f32, 1 i16, 2 i24 as in the kernel
int emu_lite_decode(const uint8_t* arena, const uint32_t* offsets, uint32_t n_lanes, uint32_t n_ticks, void* pcm, int32_t* status, int mode) {
  std::vector<LiteState> st(n_lanes);
  for (auto& x : st) lite_state_reset(x);
  LiteFrame* F = new LiteFrame;
  float res[LITE_BUF], lag[LITE_MAX_RES_LAG], lpc[LITE_BUF];
  for (uint32_t t = 0; t < n_ticks; t++)
    for (uint32_t j = 0; j < n_lanes; j++) {
      const uint32_t p = t * n_lanes + j;
      const int rc = lite_decode_packet(st[j], arena + offsets[p], offsets[p + 1] - offsets[p], *F, res, lag, lpc);
      status[p] = rc;
      for (int i = 0; i < LITE_OUT; i++) {
        const float v = rc == LITE_OUT ? st[j].silk_buffer[i / 3] : 0.f;
        if (mode == 0) ((float*)pcm)[(size_t)p * LITE_OUT + i] = v;
        else if (mode == 1) ((int16_t*)pcm)[(size_t)p * LITE_OUT + i] = (int16_t)lite_floor_i32(v * 32767.0f);
        else ((int32_t*)pcm)[(size_t)p * LITE_OUT + i] = (int32_t)rint((double)v * 8388608.0);
      }
    }
  delete F;
  return 0;
}

// decode gain of the following emu_decode_grid calls (Q8 dB, 0 = none; src/opus_decoder.rs:1376-1392)
static int g_gain_q8 = 0;
void emu_set_gain(int gain_q8) { g_gain_q8 = gain_q8; }
// per-packet flags of the following emu_decode_grid calls (bit 1: lost packet, its two payload bytes = samples to conceal;
// null: none), as the host front end hands them to the device
static const uint8_t* g_flags = nullptr;
void emu_set_flags(const uint8_t* flags) { g_flags = flags; }
// decoder output rate 48000 / ds (row F4); pcm_stride and the sample counts are then in output samples, as on the device
static int g_ds = 1;
void emu_set_downsample(int ds) { g_ds = ds > 1 ? ds : 1; }

// debug aid for the PVQ work-balance studies (tools/pvq_balance.py): the (N, K) of every leaf of the first `max_frames` cells
static uint32_t* g_leaf_dump = nullptr;
static uint32_t g_leaf_cap = 0, g_leaf_n = 0;
void emu_leaf_dump(uint32_t* buf, uint32_t cap) { g_leaf_dump = buf; g_leaf_cap = cap; g_leaf_n = 0; }
uint32_t emu_leaf_count() { return g_leaf_n; }

// Decodes a regular [n_ticks][n_lanes] packet grid with one fresh decoder state per lane, phases in
// the same order as the CUDA launcher. pcm: [packet][pcm_stride]. iy_dbg/x_dbg: optional [packet][2*960].
int emu_decode_grid(int CC, uint32_t quirks, const uint8_t* arena, const uint32_t* offsets, uint32_t n_lanes, uint32_t n_ticks,
                    uint32_t pcm_stride, float* pcm, int16_t* pcm16, int32_t* samples_out, uint32_t* final_range_out, int32_t* iy_dbg,
                    float* x_dbg) {
  quirks = (quirks & 0x00ffffffu) | (g_ds > 1 ? (uint32_t)g_ds << QX_DS_SHIFT : 0u);
  const uint32_t F = n_lanes * n_ticks;
  const uint32_t FA = 2 * F;   // main cells, then as many auxiliary cells (mb_common.cuh)
  const size_t tiles = (FA + 31) / 32;
  std::vector<uint32_t> rec(tiles * REC_SLOTS * 32, getenv("MB_EMUL_CLEAN") ? 0u : 0x5a5a5a5au);   // (scratch the device reuses between calls: stale there, garbage here)
  std::vector<StreamState> st(n_lanes);
  for (auto& s : st) stream_state_reset(s);
  std::vector<float> hist((size_t)n_lanes * CC * HIST_G, 0.f);
  std::vector<float> en((size_t)FA * EN_STRIDE, 0.f);
  std::vector<FrameMeta> meta(FA);
  std::vector<uint32_t> order((size_t)2 * F, 0), order_n(n_lanes, 0), side_of(F, 0xffffffffu);
  uint32_t aux_count = 0;
  AuxPool aux = {F, F, &aux_count};
  std::vector<uint32_t> aux_main(F, 0);
  for (uint32_t a = F; a < FA; a++) { rec[rec_index(a, S_STATUS)] = 1u << 15; meta[a].flags = 0; }
  const uint32_t series_stride = SER_PREFIX + n_ticks * (MAX_N + 2 * SHORT_MDCT) + SER_SLACK;
  std::vector<float> series((size_t)n_lanes * CC * series_stride, 0.f);
  const int disable_inv = CC == 1;
  // phase 1
  for (uint32_t f = 0; f < F; f++)
    parse_cell(arena, offsets, f, rec.data() + rec_index(f, 0), CC, quirks, disable_inv, (int)(pcm_stride / CC) * g_ds, g_flags,
               f >= n_lanes ? f - n_lanes : 0xffffffffu, f >= n_lanes ? nullptr : &st[f].prev_mode, rec.data(), aux);
  if (g_leaf_dump)
    for (uint32_t f = 0; f < F; f++) {
      const uint32_t* r = rec.data() + rec_index(f, 0);
      if (!MB_ST_DECODED(r[S_STATUS * 32])) continue;
      const uint32_t nl = MB_H1_NLEAVES(r[S_HDR1 * 32]);
      for (uint32_t l = 0; l < nl && g_leaf_n + 3 <= g_leaf_cap; l++) {
        const uint32_t g = r[(S_LEAVES + 3 * l + 2) * 32];
        g_leaf_dump[g_leaf_n++] = f; g_leaf_dump[g_leaf_n++] = MB_LF_N(g); g_leaf_dump[g_leaf_n++] = MB_LF_K(g);
      }
    }
  // phase 2a
  Team tm = {0, 1};
  EnergySmem esm;
  EnergyParams eq;
  eq.n_lanes = n_lanes; eq.n_ticks = n_ticks; eq.pkt_map = nullptr; eq.rec = rec.data(); eq.st_all = st.data(); eq.lane_stream = nullptr;
  eq.en = en.data(); eq.meta = meta.data(); eq.samples_out = samples_out; eq.final_range_out = final_range_out; eq.quirks = quirks;
  eq.flags = g_flags; eq.serial_all = getenv("MB_EMUL_PARALLEL_BANDS") ? 0 : 1;   // default: the per-stream band pass of large calls
  eq.force_float_row = g_gain_q8 != 0 || !pcm16; eq.CC = CC; eq.order = order.data(); eq.order_n = order_n.data(); eq.side_of = side_of.data();
  eq.aux = aux; eq.aux_main = aux_main.data();
  for (uint32_t j = 0; j < n_lanes; j++) celt_energy_stream(tm, &esm, j, eq);
  if (getenv("MB_EMUL_DUMP"))   // debug aid: the lanes' frame lists
    for (uint32_t j = 0; j < n_lanes; j++)
      for (uint32_t k = 0; k < order_n[j]; k++) {
        const uint32_t f = order[(size_t)k * n_lanes + j];
        fprintf(stderr, "lane %u frame %u: cell %u flags %08x off %u seed %u hdr0 %08x status %08x\n", j, k, f, meta[f].flags, meta[f].sample_off,
                meta[f].seed_in, rec[rec_index(f, S_HDR0)], rec[rec_index(f, S_STATUS)]);
      }
  // phase 2b
  std::vector<float> xbuf((size_t)FA * XBUF_STRIDE, getenv("MB_EMUL_CLEAN") ? 0.f : 12345.678f), xstage(2 * MAX_N);
  PvqParams pq; pq.rec = rec.data(); pq.meta = meta.data(); pq.xbuf = xbuf.data(); pq.iy_dbg = iy_dbg;
  for (uint32_t f = 0; f < FA; f++) {
    if (f >= F) pq.iy_dbg = nullptr;
    celt_pvq_frame(tm, xstage.data(), f, pq);
  }
  PlanParams pl; pl.rec = rec.data(); pl.en = en.data(); pl.meta = meta.data();
  for (uint32_t f = 0; f < FA; f++) celt_plan_frame(f, pl);
  BandsSmem* bsm = new BandsSmem;
  BandsParams bp; bp.rec = rec.data(); bp.en = en.data(); bp.meta = meta.data(); bp.xbuf = xbuf.data(); bp.x_dbg = x_dbg;
  std::vector<float> fold((size_t)n_lanes * FOLD_WORDS, 0.f);
  bp.fold_all = fold.data();
  for (uint32_t f = 0; f < FA; f++) {
    if (f >= F) bp.x_dbg = nullptr;
    celt_bands_frame(tm, bsm, f, bp);
  }
  // SILK-side frames: decode order per stream, fold buffer carried in stream state (k_bands_hyb)
  for (uint32_t j = 0; j < n_lanes; j++) {
    float* fb = fold.data() + (size_t)j * FOLD_WORDS;
    for (uint32_t k = 0; k < order_n[j]; k++) {
      const uint32_t f = order[(size_t)k * n_lanes + j];
      if (!(meta[f].flags & MB_FL_FOLD)) continue;
      bp.x_dbg = f >= F ? nullptr : x_dbg;
      memcpy(bsm->norm, fb, FOLD_WORDS * sizeof(float));
      celt_bands_frame(tm, bsm, f, bp, true);
      memcpy(fb, bsm->norm, FOLD_WORDS * sizeof(float));
    }
  }
  delete bsm;
  float* ism = new float[IMDCT_ROW];   // the warp-resident form k_imdct runs, its 32 lanes walked phase by phase
  ImdctParams ip; ip.rec = rec.data(); ip.meta = meta.data(); ip.xbuf = xbuf.data(); ip.series = series.data();
  ip.series_stride = series_stride; ip.CC = CC;
  const bool ola = !(quirks & Q2_MDCT_ZERO_FILL);   // overlap-add form of the inverse MDCT (quirk Q2 cleared)
  std::vector<float> tails(ola ? (size_t)FA * CC * TAIL_STRIDE : 0, 0.f);
  if (ola) { ip.ola = 1; ip.tails = tails.data(); }
  ip.ds = g_ds;
  for (uint32_t f = 0; f < FA; f++)
    for (int u = 0; u < CC; u++) celt_imdct_unit_w(tm, ism, imdct_tables_host(), f, u, ip);
  delete[] ism;
  // phase 2c
  std::vector<float> tmp;
  float* out = pcm;
  if (!out) { tmp.assign((size_t)F * pcm_stride, 0.f); out = tmp.data(); }
  TailParams tp;
  tp.rec = rec.data(); tp.meta = meta.data(); tp.pkt_map = nullptr; tp.lane_stream = nullptr; tp.st_all = st.data(); tp.hist = hist.data();
  tp.series = series.data(); tp.series_stride = series_stride; tp.n_lanes = n_lanes; tp.n_ticks = n_ticks; tp.CC = CC;
  tp.gain = g_gain_q8 == 0 ? 1.f : (float)exp(0.6931471805599453094 * (double)((3.32192809488736234787f / 5120.0f) * (float)g_gain_q8));
  tp.apply_gain = g_gain_q8 != 0; tp.pcm_f32 = out; tp.pcm_stride = pcm_stride; tp.xbuf = xbuf.data();
  tp.order = order.data(); tp.order_n = order_n.data(); tp.side_of = side_of.data(); tp.aux_main = aux_main.data(); tp.aux_base = F;
  if (ola) { tp.ola = 1; tp.tails = tails.data(); }
  tp.ds = g_ds;
  // SILK-side cells (k_silk): lane per stream, ticks in order
  {
    std::vector<SilkState> sst(n_lanes);
    std::vector<uint32_t> shist((size_t)n_lanes * 320, 0);
    std::vector<int32_t> sexc((size_t)n_lanes * 1280, 0x5a5a5a5a);   // (garbage: the lanes clear what they use)
    for (auto& x : sst) { memset((void*)&x, 0x5a, sizeof x); silk_state_reset(x, 1); }   // (garbage first: a reset must set every field it relies on)
    SilkWork* W = new SilkWork;
    memset((void*)W, 0x5a, sizeof *W);   // (the device's local memory starts as garbage too)
    SilkLaneParams sq;
    sq.n_lanes = n_lanes; sq.n_ticks = n_ticks; sq.pkt_map = nullptr; sq.lane_stream = nullptr; sq.rec = rec.data(); sq.meta = meta.data();
    sq.st_all = sst.data(); sq.silk_hist = shist.data(); sq.silk_exc = sexc.data(); sq.n_streams = n_lanes;
    sq.pcm_f32 = out; sq.pcm_i16 = pcm16; sq.pcm_stride = pcm_stride; sq.CC = CC; sq.quirks = quirks; sq.gain = tp.gain; sq.apply_gain = tp.apply_gain;
    sq.out_map = nullptr; sq.xbuf = xbuf.data(); sq.side_of = side_of.data();
    int32_t stage[32];
    for (uint32_t j = 0; j < n_lanes; j++) silk_lane(sq, j, *W, stage, silk_resampler_frac_FIR_12);
    delete W;
  }
  CombSmem* csm = new CombSmem;
  for (uint32_t j = 0; j < n_lanes; j++)
    for (int c = 0; c < CC; c++) celt_comb_channel(tm, csm, j, c, tp);
  delete csm;
  PlcSmem* psm = new PlcSmem;
  for (uint32_t j = 0; j < n_lanes; j++) celt_comb_plc_lane(tm, psm, j, tp);
  delete psm;
  for (uint32_t j = 0; j < n_lanes; j++)
    for (int c = 0; c < CC; c++) celt_deemph_lane(j, c, tp);
  for (uint32_t f = 0; f < F; f++)
    for (int c = 0; c < CC; c++) celt_fade_cell(f, c, tp);
  if (pcm16) {
    for (uint32_t j = 0; j < n_lanes; j++)
      for (uint32_t t = 0; t < n_ticks; t++) {
        uint32_t p = t * n_lanes + j;
        if (samples_out[p] <= 0 || !(meta[p].flags & (3u | MB_FL_PLC))) continue;
        if (g_flags && (g_flags[p] & 3u)) {   // a lost packet (or one rebuilt from FEC data) is not soft-clipped and clears the clipper's memory
          st[j].softclip_mem[0] = st[j].softclip_mem[1] = 0.f;
          if (!(meta[p].flags & MB_FL_F32ROW)) continue;   // int16 straight from the SILK lane
          for (int i = 0; i < samples_out[p] * CC; i++) pcm16[(size_t)p * pcm_stride + i] = float2int16(out[(size_t)p * pcm_stride + i]);
          continue;
        }
        if (!(meta[p].flags & MB_FL_F32ROW)) {   // int16 straight from the SILK lane: final unless a non-linearity carries over
          if (st[j].softclip_mem[0] == 0.f && st[j].softclip_mem[1] == 0.f) continue;
          for (int i = 0; i < samples_out[p] * CC; i++) out[(size_t)p * pcm_stride + i] = (float)pcm16[(size_t)p * pcm_stride + i] * (1.0f / 32768.0f);
        }
        soft_clip_packet(out + (size_t)p * pcm_stride, samples_out[p], CC, st[j].softclip_mem);
        for (int i = 0; i < samples_out[p] * CC; i++) pcm16[(size_t)p * pcm_stride + i] = float2int16(out[(size_t)p * pcm_stride + i]);
      }
  }
  return 0;
}

// Operation-list statistics of a CELT grid (tools/op_histogram.py): per op type, how many and over how many samples.
int emu_op_histogram(int CC, uint32_t quirks, const uint8_t* arena, const uint32_t* offsets, uint32_t n_lanes, uint32_t n_ticks,
                     uint64_t* count /*16*/, uint64_t* samples /*16*/, uint64_t* totals /*4: frames, op words, nodes, leaves*/) {
  const uint32_t F = n_lanes * n_ticks;
  std::vector<uint32_t> rec(((size_t)F + 31) / 32 * REC_SLOTS * 32, 0);
  std::vector<StreamState> st(n_lanes);
  for (auto& s : st) stream_state_reset(s);
  std::vector<float> en((size_t)F * EN_STRIDE, 0.f);
  std::vector<FrameMeta> meta(F);
  std::vector<int32_t> so(F);
  std::vector<uint32_t> fr(F);
  for (uint32_t f = 0; f < F; f++)
    parse_cell(arena, offsets, f, rec.data() + rec_index(f, 0), CC, quirks, CC == 1, 960, nullptr, f >= n_lanes ? f - n_lanes : 0xffffffffu,
               f >= n_lanes ? nullptr : &st[f].prev_mode);
  Team tm = {0, 1};
  EnergySmem esm;
  std::vector<uint32_t> order((size_t)2 * F, 0), order_n(n_lanes, 0), side_of(F, 0xffffffffu);
  EnergyParams eq;
  eq.n_lanes = n_lanes; eq.n_ticks = n_ticks; eq.pkt_map = nullptr; eq.rec = rec.data(); eq.st_all = st.data(); eq.lane_stream = nullptr;
  eq.en = en.data(); eq.meta = meta.data(); eq.samples_out = so.data(); eq.final_range_out = fr.data(); eq.quirks = quirks;
  eq.flags = nullptr; eq.serial_all = 1; eq.force_float_row = 1; eq.CC = CC; eq.order = order.data(); eq.order_n = order_n.data();
  eq.side_of = side_of.data(); eq.aux = AuxPool{0, 0, nullptr}; eq.aux_main = nullptr;
  for (uint32_t j = 0; j < n_lanes; j++) celt_energy_stream(tm, &esm, j, eq);
  std::vector<float> xbuf((size_t)F * XBUF_STRIDE, 0.f), xstage(2 * MAX_N);
  PvqParams pq; pq.rec = rec.data(); pq.meta = meta.data(); pq.xbuf = xbuf.data(); pq.iy_dbg = nullptr;
  for (uint32_t f = 0; f < F; f++) celt_pvq_frame(tm, xstage.data(), f, pq);
  PlanParams pl; pl.rec = rec.data(); pl.en = en.data(); pl.meta = meta.data();
  for (uint32_t f = 0; f < F; f++) celt_plan_frame(f, pl);
  for (int i = 0; i < 16; i++) count[i] = samples[i] = 0;
  for (int i = 0; i < 4; i++) totals[i] = 0;
  for (uint32_t f = 0; f < F; f++) {
    if (!(meta[f].flags & 1u)) continue;
    const uint32_t* r = rec.data() + rec_index(f, 0);
    const uint32_t n = r[S_PLAN * 32];
    if (n == 0xffffffffu) continue;
    totals[0]++; totals[1] += n; totals[2] += MB_H1_NNODES(r[S_HDR1 * 32]); totals[3] += MB_H1_NLEAVES(r[S_HDR1 * 32]);
    for (uint32_t p = 0; p < n;) {
      const uint32_t op = r[(S_OPS + p) * 32], type = MB_OP_TYPE(op);
      count[type]++; samples[type] += type == OP_AVG ? (MB_OP_N(op) | (MB_OP_P(op) << 9)) : MB_OP_N(op);
      p++;
      if (type == OP_SKIP) p += MB_OP_N(op);
      else if (type == OP_COPY_LB || type == OP_LBOUT || type == OP_MERGE || type == OP_STEREO2 || type == OP_HYB) p += 1;
      else if (type == OP_FOLD || type == OP_NOISE || type == OP_ANTI) p += 2;
    }
  }
  return 0;
}

// ---- the host front end for multi-frame packets (mb_host_packets.h), exposed for the CPU tests
static mbh::Split g_split;
static std::vector<mbh::StreamTrack> g_track;   // per-stream memory of the front end (frame size / mode of the last packet)
void emu_split_set_streams(uint32_t n_streams) { g_track.assign(n_streams, mbh::StreamTrack()); }

// Cuts the batch into frames; returns the number of virtual packets, *arena_bytes = size of the virtual arena.
uint32_t emu_split_batch(const uint8_t* arena, const uint32_t* offsets, const uint32_t* stream_ids, const uint8_t* flags, uint32_t n_packets,
                         int cap_per_ch, int CC, uint32_t* arena_bytes) {
  mbh::split_batch(arena, offsets, stream_ids, flags, n_packets, cap_per_ch, CC, &g_split, g_track.empty() ? nullptr : g_track.data());
  *arena_bytes = (uint32_t)g_split.arena.size();
  return (uint32_t)g_split.stream_ids.size();
}
// Copies the result of the last emu_split_batch out: offsets[nv+1], stream_ids[nv], flags[nv], out_map[2*nv], first[n_packets+1].
void emu_split_fetch(uint8_t* arena, uint32_t* offsets, uint32_t* stream_ids, uint8_t* flags, uint32_t* out_map, uint32_t* first) {
  memcpy(arena, g_split.arena.data(), g_split.arena.size());
  memcpy(offsets, g_split.offsets.data(), g_split.offsets.size() * 4);
  memcpy(stream_ids, g_split.stream_ids.data(), g_split.stream_ids.size() * 4);
  memcpy(flags, g_split.flags.data(), g_split.flags.size());
  memcpy(out_map, g_split.out_map.data(), g_split.out_map.size() * 4);
  memcpy(first, g_split.first.data(), g_split.first.size() * 4);
}
void emu_merge_results(uint32_t n_packets, const int32_t* v_samples, const uint32_t* v_range, int32_t* samples_out, uint32_t* final_range_out) {
  mbh::merge_results(g_split, n_packets, v_samples, v_range, samples_out, final_range_out);
}

}  // extern "C"
