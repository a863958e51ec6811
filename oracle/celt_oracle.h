// ORACLE (test infrastructure only — never linked into the product library).
// CPU restatement of mousiki's *float* CELT frame decoder, quirks included
// (SURVEY.md §0.1 Q1/Q2 and the ones found while restating).
#pragma once
#include <stdint.h>
#include <vector>
#include "ec.h"

namespace orc {

enum {
  NB_EBANDS = 21, OVERLAP = 120, SHORT_MDCT = 120, MAX_LM = 3,
  DECODE_BUFFER_SIZE = 2048, MAX_PERIOD = 1024, COMBFILTER_MINPERIOD = 15,
  SPREAD_NONE = 0, SPREAD_LIGHT = 1, SPREAD_NORMAL = 2, SPREAD_AGGRESSIVE = 3,
};

// Quirk switches (bit set = behave like mousiki, clear = behave like libopus).
enum : uint32_t {
  QUIRK_Q1_DYNALLOC_BUDGET = 1u << 0,  // celt_decoder.rs:3054-3057: post-dynalloc budget to quant_all_bands
  QUIRK_Q2_MDCT_ZERO_FILL  = 1u << 1,  // mdct.rs:373: output.fill(0.0) before IMDCT write (kills TDAC)
  QUIRK_Q5_FOLD_START      = 1u << 2,  // bands.rs:2809-2817: fold_start threshold includes +N
  QUIRK_Q6_HYBRID_LOWBAND  = 1u << 3,  // bands.rs:2539-2569: overlapping lowband windows -> (None, None)
  QUIRK_Q7_EBITS_BUST      = 1u << 4,  // rate.rs:777: "bust" test uses bits>>stereo>>BITRES
  QUIRK_Q8_COARSE_CLAMP    = 1u << 5,  // quant_bands.rs:944-946: clamp +-28 and prev += (q - beta*q)
  QUIRK_Q9_DUAL_STEREO_RSV = 1u << 6,  // rate.rs:703-715: dual-stereo flag read even when intensity<=start
  QUIRKS_MOUSIKI = 0x7f,
};

struct CeltTrace {
  int lm = 0, C = 0, N = 0, silence = 0, transient = 0, intra = 0, spread = 0, anti_collapse_on = 0;
  int intensity = 0, dual_stereo = 0, coded_bands = 0, pf_pitch = 0, pf_tapset = 0;
  float pf_gain = 0;
  int tf_res[NB_EBANDS] = {0}, pulses[NB_EBANDS] = {0}, fine_quant[NB_EBANDS] = {0};
  std::vector<int32_t> iy;     // C*N, decoded PVQ pulses at their X position (0 elsewhere)
  std::vector<float> X;        // C*N normalised spectrum after quant_all_bands + anti_collapse
  std::vector<float> bandE;    // 2*21 oldEBands after finalise
  std::vector<float> syn;      // CC*(N+overlap) IMDCT output (before comb filter)
  std::vector<float> post;     // CC*N after comb filter (before de-emphasis)
  uint32_t seed_out = 0;
};

struct CeltDecoder {
  int channels = 1;         // CC (output)
  int stream_channels = 1;  // C (coded)
  int start_band = 0, end_band = NB_EBANDS;
  int downsample = 1;
  int disable_inv = 0;
  uint32_t quirks = QUIRKS_MOUSIKI;
  uint32_t rng = 0;
  int error = 0;
  int loss_duration = 0;
  int skip_plc = 1, prefilter_and_fold = 0, last_pitch_index = 0;   // types.rs:807-816 (PLC)
  float lpc[2][24];                                                 // PLC synthesis filter per channel
  int postfilter_period = 0, postfilter_period_old = 0;
  float postfilter_gain = 0, postfilter_gain_old = 0;
  int postfilter_tapset = 0, postfilter_tapset_old = 0;
  float preemph_mem[2] = {0, 0};
  float decode_mem[2][DECODE_BUFFER_SIZE + OVERLAP];
  float old_ebands[2 * NB_EBANDS], old_log_e[2 * NB_EBANDS], old_log_e2[2 * NB_EBANDS],
      background_log_e[2 * NB_EBANDS];
  std::vector<float> norm_storage;  // bands.rs:2614-2617 persistent fold buffer (never cleared)
  CeltTrace* trace = nullptr;  // optional stage dump of the last decoded frame

  void init(int channels_, int downsample_);
  void reset();
};

// celt_decode_with_ec (src/celt/celt_decoder.rs:3209-4137). `ec` may be a range
// decoder already positioned inside the frame (hybrid) or nullptr (owns one).
// Returns samples per channel (>=0) or a negative Opus error code.
int celt_decode_with_ec(CeltDecoder* st, const uint8_t* data, int len, float* pcm, int frame_size,
                        EcDec* ec, int accum);

const float* celt_window120();

}  // namespace orc
