// ORACLE (test infrastructure only — never linked into the product library).
// CPU restatement of mousiki's fixed-point SILK decoder as used by the full Opus decoder
// (src/silk/dec_api.rs and the files it pulls in): normal decode, packet-loss concealment (plc.rs), comfort noise
// (cng.rs) and LBRR (in-band FEC) frames (dec_api.rs:443-530, decode_frame.rs:54-153).
// Quirk switches: bit set = behave like mousiki as written, clear = behave like libopus.
#pragma once
#include <stdint.h>
#include "ec.h"

namespace orc {

enum : uint32_t {
  // nlsf_decode.rs:109-111: silk_SMULBB written as a wrapping 16-bit multiply (result truncated to i16)
  QUIRK_S1_NLSF_SMULBB16 = 1u << 8,
  // nlsf2a.rs:37-52: cos table looked up *and stored* at ordering[k] (identity permutation); libopus reads
  // NLSF[k] and stores at ordering[k]
  QUIRK_S2_NLSF2A_ORDER = 1u << 9,
  // code_signs.rs:112-118: sign decode stops at frame_length inside the last (padded) shell block (12 kHz, 10 ms)
  QUIRK_S3_SIGNS_FRAMELEN = 1u << 10,
  // dec_api.rs:150-154: mono -> stereo switch re-initialises BOTH channel states (libopus: only the side channel)
  QUIRK_S4_STEREO_REINIT = 1u << 11,
  QUIRKS_SILK_MOUSIKI = 0xf00,
};

enum { SILK_MAX_FRAME = 320, SILK_MAX_ORDER = 16, SILK_LTP_ORDER = 5, SILK_MAX_SUBFR = 4, SILK_MAX_FRAMES_PER_PACKET = 3,
       SILK_TYPE_INACTIVE = 0, SILK_TYPE_UNVOICED = 1, SILK_TYPE_VOICED = 2,
       SILK_CODE_INDEPENDENTLY = 0, SILK_CODE_INDEPENDENTLY_NO_LTP_SCALING = 1, SILK_CODE_CONDITIONALLY = 2,
       SILK_FLAG_DECODE_NORMAL = 0, SILK_FLAG_PACKET_LOST = 1, SILK_FLAG_DECODE_LBRR = 2 };

struct SilkNlsfCb {
  int n_vectors, order, quant_step_size_q16;
  const uint8_t* cb1_nlsf_q8; const int16_t* cb1_wght_q9; const uint8_t* cb1_icdf; const uint8_t* pred_q8;
  const uint8_t* ec_sel; const uint8_t* ec_icdf; const int16_t* delta_min_q15;
};

struct SilkIndices {  // decode_indices.rs:50-63
  int8_t gains[SILK_MAX_SUBFR] = {0}, ltp[SILK_MAX_SUBFR] = {0}, nlsf[SILK_MAX_ORDER + 1] = {0};
  int16_t lag_index = 0;
  int8_t contour_index = 0, signal_type = 0, quant_offset_type = 0, nlsf_interp_q2 = 4, per_index = 0, ltp_scale_index = 0, seed = 0;
};

struct SilkControl {  // decoder_control.rs
  int32_t pitch_l[SILK_MAX_SUBFR] = {0}, gains_q16[SILK_MAX_SUBFR] = {0};
  int16_t pred_coef_q12[2][SILK_MAX_ORDER] = {{0}}, ltp_coef_q14[SILK_LTP_ORDER * SILK_MAX_SUBFR] = {0};
  int32_t ltp_scale_q14 = 0;
};

struct SilkResampler {  // resampler.rs:40-48
  int fs_in_khz = 0, fs_out_khz = 0, batch_size = 0, input_delay = 0, inv_ratio_q16 = 0, mode = 0;  // 0 copy 1 up2 2 iir_fir 3 down_fir
  int16_t delay_buf[96] = {0};
  int32_t s_iir[6] = {0};
  int16_t s_fir[8] = {0};
  // down-sampling kernel (resampler_private_down_fir.rs:20-80): AR2 state in s_iir[0..2), FIR tail, ratio configuration
  int32_t s_down[36] = {0};
  int down_order = 0, down_fracs = 0;
  const int16_t* down_coefs = nullptr;
  int init(int fs_in_hz, int fs_out_hz);
  int run(int16_t* out, const int16_t* in, int in_len);
};

struct SilkPlc {  // decoder_state.rs:23-75 (Default); reset(): decoder_state.rs:77-95
  int32_t pitch_l_q8 = 0;
  int16_t ltp_coef_q14[SILK_LTP_ORDER] = {0}, prev_lpc_q12[SILK_MAX_ORDER] = {0};
  int last_frame_lost = 0;
  int32_t rand_seed = 0;
  int16_t rand_scale_q14 = 0;
  int32_t conc_energy = 0;
  int conc_energy_shift = 0;
  int16_t prev_ltp_scale_q14 = 0;
  int32_t prev_gain_q16[2] = {65536, 65536};
  int fs_khz = 0, nb_subfr = 2, subfr_length = 20;
  void reset(int frame_length);
};
struct SilkCng {  // cng.rs:31-54
  int32_t exc_buf_q14[SILK_MAX_FRAME] = {0};
  int16_t smth_nlsf_q15[SILK_MAX_ORDER] = {0};
  int32_t synth_state[SILK_MAX_ORDER] = {0};
  int32_t smth_gain_q16 = 0, rand_seed = 3176576;
  int fs_khz = -1;
  void reset(int lpc_order);
};

struct SilkChannel {  // decoder_state.rs:99-131 + decoder_set_fs.rs:63-124
  int fs_khz = 0, fs_api_hz = 0, nb_subfr = 4, subfr_length = 0, frame_length = 0, ltp_mem_length = 0, lpc_order = 16;
  int first_frame_after_reset = 1, lag_prev = 0, last_gain_index = 0, prev_signal_type = 0;
  const uint8_t* pitch_lag_low_bits_icdf = nullptr; const uint8_t* pitch_contour_icdf = nullptr;
  const SilkNlsfCb* cb = nullptr;
  SilkResampler resampler;
  int16_t out_buf[480] = {0};
  int32_t s_lpc_q14[SILK_MAX_ORDER] = {0};
  int32_t prev_gain_q16 = 65536;
  int32_t exc_q14[SILK_MAX_FRAME] = {0};
  int16_t prev_nlsf_q15[SILK_MAX_ORDER] = {0};
  SilkIndices indices;
  int n_frames_decoded = 0, n_frames_per_packet = 0, ec_prev_signal_type = 0, ec_prev_lag_index = 0;
  int vad_flags[SILK_MAX_FRAMES_PER_PACKET] = {0}, lbrr_flag = 0, lbrr_flags[SILK_MAX_FRAMES_PER_PACKET] = {0};
  int loss_count = 0;
  SilkPlc plc;
  SilkCng cng;
  int set_fs(int fs_khz_, int fs_api_hz_);
};

struct SilkDecControl {  // dec_api.rs:74-89
  int n_channels_api = 1, n_channels_internal = 1, api_sample_rate = 16000, internal_sample_rate = 16000, payload_size_ms = 20,
      prev_pitch_lag = 0;
};

struct SilkTrace {  // last decoded frame of channel 0/1 (tests)
  SilkIndices indices[2];
  int16_t pulses[2][SILK_MAX_FRAME];
  int16_t nlsf_q15[2][SILK_MAX_ORDER];
  SilkControl ctrl[2];
  int16_t out[2][SILK_MAX_FRAME];
};

struct SilkDecoder {  // dec_api.rs:35-46
  SilkChannel ch[2];
  int16_t pred_prev_q13[2] = {0, 0}, s_mid[2] = {0, 0}, s_side[2] = {0, 0};
  int n_channels_api = 1, n_channels_internal = 1, prev_decode_only_middle = 0;
  uint32_t quirks = QUIRKS_SILK_MOUSIKI;
  SilkTrace* trace = nullptr;
  void reset(int init);  // silk_ResetDecoder (init=0) / silk_InitDecoder (init=1)
  // dec_api.rs:132-441. Returns samples per channel at the API rate (>0) or a negative error.
  int decode(SilkDecControl* c, int lost_flag, int new_packet, EcDec* ec, int16_t* samples_out);
};

// stand-alone pieces exposed for the known-answer tests
void silk_nlsf_decode(int16_t* nlsf_q15, const int8_t* indices, const SilkNlsfCb* cb, uint32_t quirks);
void silk_nlsf2a(int16_t* a_q12, const int16_t* nlsf_q15, int order, uint32_t quirks);
void silk_nlsf_stabilize(int16_t* nlsf_q15, const int16_t* delta_min_q15, int L);
void silk_gains_dequant(int32_t* gain_q16, const int8_t* ind, int8_t* prev_ind, int conditional, int nb_subfr);
void silk_decode_pitch(int lag_index, int contour_index, int32_t* pitch_lags, int fs_khz, int nb_subfr);
int32_t silk_lpc_inverse_pred_gain(const int16_t* a_q12, int order);
int32_t silk_log2lin(int32_t in_log_q7);
const SilkNlsfCb* silk_cb_wb();
const SilkNlsfCb* silk_cb_nb_mb();

}  // namespace orc
