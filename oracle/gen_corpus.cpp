// ORACLE-SIDE TOOL (test / bench infrastructure only — never linked into the product library).
// P-struct: a structured CELT corpus written THROUGH the restated decoder.
// Built only into oracle/_build/liborcgen.so, from celt_oracle.cpp compiled with -DORC_GEN (namespace renamed to orcg): the
// decoder runs "driven" (ec.h, GenCtl) — every symbol it asks for is chosen and range-ENCODED instead of read — so the finished
// bytes decode along exactly the path the writer took. What is steered, as on real material: no silence flag, intra energy on a
// stream's first frame only, ~10 % transients, the post-filter on ~30 % of the frames with a mild gain, and the coarse band
// energies, which follow a per-stream envelope (level, spectral tilt, slowly wandering loudness, per-band wander) sized so that the
// PCM stays inside +-1. Everything else (TF, spread, dynalloc, trim, intensity / dual stereo, fine energy, PVQ indices, split
// angles) is drawn from each symbol's own probability model. The reference's bench decodes encoder-made packets
// (src/bin/codec_bench.rs:620-702); this is the closest stand-in that needs no encoder.
#include <math.h>
#include <string.h>
#include <thread>
#include <vector>
#include "celt_oracle.h"

using namespace orc;

extern "C" {

// Packets of `payload` bytes behind `toc` (CELT-only, code 0), [tick][lane] order like helpers.fuzz_grid: arena / offsets out.
// pcm (nullable): [packet][frame_size * channels] float, what the writer's decoder produced (the checker must reproduce it);
// final_range (nullable): the coder's range after each frame. Returns 0, or -1 when a frame could not be written.
int orcgen_celt_grid(uint64_t seed, int n_lanes, int n_ticks, int payload, int toc, int channels, float level, uint32_t quirks,
                     uint8_t* arena, uint32_t* offsets, float* pcm, uint32_t* final_range, int threads) {
  const int lm = (toc >> 3) & 3, bw = (toc >> 5) & 3, stereo = (toc >> 2) & 1;
  const int frame_size = SHORT_MDCT << lm;
  static const int end_of_bw[4] = {13, 17, 19, 21};   // NB/WB/SWB/FB (src/opus_decoder.rs, celt end band by bandwidth)
  const size_t n = (size_t)n_lanes * n_ticks;
  for (size_t p = 0; p <= n; p++) offsets[p] = (uint32_t)(p * (size_t)(payload + 1));
  if (threads < 1) threads = 1;
  std::vector<int> bad(threads, 0);
  auto work = [&](int tid) {
    std::vector<float> out((size_t)frame_size * channels);
    for (int j = tid; j < n_lanes; j += threads) {
      CeltDecoder st;
      st.init(channels, 1);
      st.quirks = quirks;
      st.stream_channels = stereo ? 2 : 1;
      st.end_band = end_of_bw[bw];
      GenCtl g;
      g.s = (seed + 0x9e3779b97f4a7c15ull * (uint64_t)(j + 1)) | 1ull;
      for (int k = 0; k < 8; k++) g.next();
      g.level = level + 1.5f * (g.unit() - 0.5f);          // streams differ in loudness ...
      g.tilt = 3.0f * (g.unit() - 0.3f);                     // ... and in spectral tilt (most fall towards the high bands)
      for (int t = 0; t < n_ticks; t++) {
        const size_t p = (size_t)t * n_lanes + j;
        uint8_t* pkt = arena + offsets[p];
        pkt[0] = (uint8_t)toc;
        // envelope: loudness is a slow random walk pulled back to zero, every band wanders a little around it
        g.loud = 0.95f * g.loud + 0.8f * (g.unit() - 0.5f);
        for (int i = 0; i < 2 * NB_EBANDS; i++) g.band_dev[i] = 0.8f * g.band_dev[i] + 1.5f * (g.unit() - 0.5f);
        g.frame = t;
        EcEnc enc;
        enc.init(pkt + 1, (uint32_t)payload);
        g.enc = &enc;
        g.forced = -1;
        EcDec dec;
        dec.gen_open(&g, (uint32_t)payload);
        const int ret = celt_decode_with_ec(&st, pkt + 1, payload, out.data(), frame_size, &dec, 0);
        const uint32_t rng = enc.rng;
        enc.done();
        if (ret != frame_size || enc.error) bad[tid] = 1;
        if (pcm) memcpy(pcm + p * (size_t)frame_size * channels, out.data(), out.size() * sizeof(float));
        if (final_range) final_range[p] = rng;
      }
    }
  };
  std::vector<std::thread> th;
  for (int i = 1; i < threads; i++) th.emplace_back(work, i);
  work(0);
  for (auto& t : th) t.join();
  for (int b : bad) if (b) return -1;
  return 0;
}

}  // extern "C"
