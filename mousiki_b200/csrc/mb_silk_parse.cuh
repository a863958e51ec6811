// Phase 1 for SILK-only packets: the whole entropy side of a SILK payload, one lane per packet, no stream state.
//
// Replaces (reference, src/silk/): the entropy half of silk_decode (dec_api.rs:132-260: VAD/LBRR flags, LBRR payload
// skip :443-530, stereo predictor + mid-only flag), decode_indices.rs:109-269, decode_pulses.rs:50-158,
// shell_coder.rs:94-126, code_signs.rs:93-125, stereo_decode_pred.rs:21-56, and the SILK-only part of the redundancy
// test in opus_decoder.rs:746-759.
//
// Why this can run for every packet of every tick at once: nothing the range decoder reads depends on decoder state.
// Conditional coding only refers to the previous frame *of the same packet* (dec_api.rs:561-586: frame 0 is always
// coded independently), the sampling rate / subframe count / channel count come from the TOC.
//
// Output: a SILK record in the cell's scratch slots (same tile-of-32 layout as the CELT records, see mb_common.cuh):
//   S_HDR0    fs_khz[0..4] nb_subfr[5..7] n_frames[8..9] channels_internal[10..11]
//   S_RANGE   range decoder `rng` after the SILK payload (= final_range of a SILK-only packet)
//   S_STATUS  return code, bit 14 "parsed", bit 16 "SILK cell"
//   SL_BASE + i*SL_FRAME_WORDS: frame i: [0] ms_pred_q13[0]  [1] ms_pred_q13[1]  [2] decode_only_middle
//       then per channel c: SL_IDX_WORDS packed side-information words and 160 words of pulses (2 x i16 each)
#pragma once
#include "mb_rangedec.cuh"

#define SILK_TABLE MB_TBL
#include "silk_tables_data.h"

namespace mb {

enum {
  S1_NLSF_SMULBB16 = 1u << 8, S2_NLSF2A_ORDER = 1u << 9, S3_SIGNS_FRAMELEN = 1u << 10, S4_STEREO_REINIT = 1u << 11,
  QUIRKS_SILK = 0xf00u,
  SL_BASE = 4, SL_IDX_WORDS = 9, SL_PULSE_WORDS = 160, SL_CH_WORDS = SL_IDX_WORDS + SL_PULSE_WORDS,
  SL_FRAME_WORDS = 3 + 2 * SL_CH_WORDS,
  SILK_T_INACTIVE = 0, SILK_T_UNVOICED = 1, SILK_T_VOICED = 2,
  SILK_COND_INDEP = 0, SILK_COND_INDEP_NO_LTP = 1, SILK_COND_COND = 2,
};
#define MB_ST_SILK(s) (((s) >> 16) & 1u)
#define MB_SH_FS(h) ((h) & 31u)
#define MB_SH_SUBFR(h) (((h) >> 5) & 7u)
#define MB_SH_FRAMES(h) (((h) >> 8) & 3u)
#define MB_SH_NCI(h) (((h) >> 10) & 3u)

// side information of one SILK frame (decode_indices.rs:50-63), unpacked form
struct SilkSide {
  int8_t gains[4], ltp[4], nlsf[17];
  int16_t lag_index;
  int8_t contour, signal_type, quant_offset, interp_q2, per_index, ltp_scale, seed, cond, coded;
};

// words 0..8: gains | ltp | nlsf[0..3] | nlsf[4..7] | nlsf[8..11] | nlsf[12..15] | nlsf[16], contour, per_index, ltp_scale |
//             lag_index (u16), signal_type, quant_offset | interp_q2, seed, cond, coded
MB_DEV uint32_t pack4(int a, int b, int c, int d) {
  return (uint32_t)(uint8_t)a | ((uint32_t)(uint8_t)b << 8) | ((uint32_t)(uint8_t)c << 16) | ((uint32_t)(uint8_t)d << 24);
}
MB_DEV void silk_side_store(const SilkSide& s, uint32_t* w /*stride 32*/) {
  w[0 * 32] = pack4(s.gains[0], s.gains[1], s.gains[2], s.gains[3]);
  w[1 * 32] = pack4(s.ltp[0], s.ltp[1], s.ltp[2], s.ltp[3]);
  for (int k = 0; k < 4; k++) w[(2 + k) * 32] = pack4(s.nlsf[4 * k], s.nlsf[4 * k + 1], s.nlsf[4 * k + 2], s.nlsf[4 * k + 3]);
  w[6 * 32] = pack4(s.nlsf[16], s.contour, s.per_index, s.ltp_scale);
  w[7 * 32] = (uint32_t)(uint16_t)s.lag_index | ((uint32_t)(uint8_t)s.signal_type << 16) | ((uint32_t)(uint8_t)s.quant_offset << 24);
  w[8 * 32] = pack4(s.interp_q2, s.seed, s.cond, s.coded);
}
MB_DEV void silk_side_load(SilkSide& s, const uint32_t* w /*stride 32*/) {
  uint32_t a = w[0], b = w[32];
  for (int k = 0; k < 4; k++) { s.gains[k] = (int8_t)(a >> (8 * k)); s.ltp[k] = (int8_t)(b >> (8 * k)); }
  for (int k = 0; k < 4; k++) {
    uint32_t v = w[(2 + k) * 32];
    for (int j = 0; j < 4; j++) s.nlsf[4 * k + j] = (int8_t)(v >> (8 * j));
  }
  a = w[6 * 32];
  s.nlsf[16] = (int8_t)a; s.contour = (int8_t)(a >> 8); s.per_index = (int8_t)(a >> 16); s.ltp_scale = (int8_t)(a >> 24);
  a = w[7 * 32];
  s.lag_index = (int16_t)(a & 0xffffu); s.signal_type = (int8_t)(a >> 16); s.quant_offset = (int8_t)(a >> 24);
  a = w[8 * 32];
  s.interp_q2 = (int8_t)a; s.seed = (int8_t)(a >> 8); s.cond = (int8_t)(a >> 16); s.coded = (int8_t)(a >> 24);
}

// Every SILK symbol is an 8-bit inverse-CDF read: one out-of-line copy keeps the parse small enough for the
// instruction cache (inlined at ~60 sites the payload parser alone was 330 KB of SASS).
MB_DEV_NOINLINE int silk_icdf(RangeDec& ec, const uint8_t* tab, unsigned ftb) { return ec.icdf(tab, ftb); }

// per-rate constants (decoder_set_fs.rs:137-198): everything the entropy side needs from the sampling rate
struct SilkRate {
  int fs_khz, nb_subfr, frame_length, order;
  const uint8_t *cb1_icdf, *ec_sel, *ec_icdf, *lag_low_icdf, *contour_icdf;
};
MB_DEV SilkRate silk_rate(int fs_khz, int nb_subfr) {
  SilkRate r;
  r.fs_khz = fs_khz; r.nb_subfr = nb_subfr; r.frame_length = 5 * fs_khz * nb_subfr;
  const bool wb = fs_khz == 16;
  r.order = wb ? 16 : 10;
  r.cb1_icdf = wb ? silk_NLSF_CB1_iCDF_WB : silk_NLSF_CB1_iCDF_NB_MB;
  r.ec_sel = wb ? silk_NLSF_CB2_SELECT_WB : silk_NLSF_CB2_SELECT_NB_MB;
  r.ec_icdf = wb ? silk_NLSF_CB2_iCDF_WB : silk_NLSF_CB2_iCDF_NB_MB;
  r.lag_low_icdf = fs_khz == 8 ? silk_uniform4_iCDF : (fs_khz == 12 ? silk_uniform6_iCDF : silk_uniform8_iCDF);
  r.contour_icdf = fs_khz == 8 ? (nb_subfr == 4 ? silk_pitch_contour_NB_iCDF : silk_pitch_contour_10_ms_NB_iCDF)
                               : (nb_subfr == 4 ? silk_pitch_contour_iCDF : silk_pitch_contour_10_ms_iCDF);
  return r;
}

// decode_indices.rs:109-269. prev_type / prev_lag: the ec_prev_* pair of this channel (packet-local, see header).
MB_DEV_NOINLINE void silk_parse_indices(RangeDec& ec, const SilkRate& R, int vad, int lbrr, int cond, int& prev_type, int& prev_lag, SilkSide& s) {
  const int raw = (lbrr || vad) ? silk_icdf(ec, silk_type_offset_VAD_iCDF, 8) + 2 : silk_icdf(ec, silk_type_offset_no_VAD_iCDF, 8);
  s.signal_type = (int8_t)(raw >> 1);
  s.quant_offset = (int8_t)(raw & 1);
  if (cond == SILK_COND_COND) s.gains[0] = (int8_t)silk_icdf(ec, silk_delta_gain_iCDF, 8);
  else {
    const int msb = silk_icdf(ec, silk_gain_iCDF + 8 * s.signal_type, 8);
    s.gains[0] = (int8_t)((msb << 3) | silk_icdf(ec, silk_uniform8_iCDF, 8));
  }
  for (int k = 1; k < 4; k++) s.gains[k] = k < R.nb_subfr ? (int8_t)silk_icdf(ec, silk_delta_gain_iCDF, 8) : 0;
  const int cb1 = silk_icdf(ec, R.cb1_icdf + (s.signal_type >> 1) * 32, 8);
  s.nlsf[0] = (int8_t)cb1;
  const uint8_t* sel = R.ec_sel + cb1 * (R.order >> 1);
  for (int i = 0; i < 16; i++) {
    if (i >= R.order) { s.nlsf[i + 1] = 0; continue; }
    const uint32_t e = sel[i >> 1];
    const int ix = (int)((i & 1) ? ((e >> 5) & 7u) : ((e >> 1) & 7u)) * 9;   // nlsf_unpack.rs:30-47 (entropy table select)
    int v = silk_icdf(ec, R.ec_icdf + ix, 8);
    if (v == 0) v -= silk_icdf(ec, silk_NLSF_EXT_iCDF, 8);
    else if (v == 8) v += silk_icdf(ec, silk_NLSF_EXT_iCDF, 8);
    s.nlsf[i + 1] = (int8_t)(v - 4);
  }
  s.interp_q2 = R.nb_subfr == 4 ? (int8_t)silk_icdf(ec, silk_NLSF_interpolation_factor_iCDF, 8) : 4;
  s.lag_index = 0; s.contour = 0; s.per_index = 0; s.ltp_scale = 0;
  s.ltp[0] = s.ltp[1] = s.ltp[2] = s.ltp[3] = 0;
  if (s.signal_type == SILK_T_VOICED) {
    bool absolute = true;
    if (cond == SILK_COND_COND && prev_type == SILK_T_VOICED) {
      const int delta = silk_icdf(ec, silk_pitch_delta_iCDF, 8);
      if (delta > 0) {
        int v = prev_lag + delta - 9;
        s.lag_index = (int16_t)(v > 32767 ? 32767 : (v < -32768 ? -32768 : v));
        absolute = false;
      }
    }
    if (absolute) {
      const int hi = silk_icdf(ec, silk_pitch_lag_iCDF, 8);
      s.lag_index = (int16_t)(hi * (R.fs_khz >> 1) + silk_icdf(ec, R.lag_low_icdf, 8));
    }
    prev_lag = s.lag_index;
    s.contour = (int8_t)silk_icdf(ec, R.contour_icdf, 8);
    s.per_index = (int8_t)silk_icdf(ec, silk_LTP_per_index_iCDF, 8);
    const uint8_t* g = s.per_index == 0 ? silk_LTP_gain_iCDF_0 : (s.per_index == 1 ? silk_LTP_gain_iCDF_1 : silk_LTP_gain_iCDF_2);
    for (int k = 0; k < R.nb_subfr; k++) s.ltp[k] = (int8_t)silk_icdf(ec, g, 8);
    s.ltp_scale = cond == SILK_COND_INDEP ? (int8_t)silk_icdf(ec, silk_LTPscale_iCDF, 8) : 0;
  }
  s.seed = (int8_t)silk_icdf(ec, silk_uniform4_iCDF, 8);
  prev_type = s.signal_type;
}

// one binary split of the shell code (shell_coder.rs:36-45)
MB_DEV void silk_split(RangeDec& ec, int total, const uint8_t* table, int& a, int& b) {
  if (total > 0) { a = silk_icdf(ec, table + silk_shell_code_table_offsets[total], 8); b = total - a; }
  else { a = 0; b = 0; }
}

// decode_pulses.rs:50-158 + code_signs.rs:93-125. q: frame_length rounded up to 16 entries.
MB_DEV_NOINLINE void silk_parse_pulses(RangeDec& ec, int16_t* q, int signal_type, int quant_offset, int frame_length, uint32_t quirks) {
  const int rate = silk_icdf(ec, silk_rate_levels_iCDF + 9 * (signal_type >> 1), 8);
  const int nblocks = (frame_length + 15) >> 4;
  uint8_t sum[20], nls[20];
  for (int i = 0; i < nblocks; i++) {
    int n = 0;
    int v = silk_icdf(ec, silk_pulses_per_block_iCDF + 18 * rate, 8);
    while (v == 17) { n++; v = silk_icdf(ec, silk_pulses_per_block_iCDF + 18 * 9 + (n == 10), 8); }
    sum[i] = (uint8_t)v; nls[i] = (uint8_t)n;
  }
  for (int i = 0; i < nblocks; i++) {
    int16_t* o = q + 16 * i;
    if (sum[i] == 0) { for (int k = 0; k < 16; k++) o[k] = 0; continue; }
    // 16 -> 8 -> 4 -> 2 -> 1 binary tree, depth first (shell_coder.rs:94-126)
    int h[2], qd[2], e[2], a, b;
    silk_split(ec, sum[i], silk_shell_code_table3, h[0], h[1]);
    for (int x = 0; x < 2; x++) {
      silk_split(ec, h[x], silk_shell_code_table2, qd[0], qd[1]);
      for (int y = 0; y < 2; y++) {
        silk_split(ec, qd[y], silk_shell_code_table1, e[0], e[1]);
        for (int z = 0; z < 2; z++) {
          silk_split(ec, e[z], silk_shell_code_table0, a, b);
          o[8 * x + 4 * y + 2 * z] = (int16_t)a; o[8 * x + 4 * y + 2 * z + 1] = (int16_t)b;
        }
      }
    }
  }
  for (int i = 0; i < nblocks; i++) {
    if (nls[i] == 0) continue;
    for (int k = 0; k < 16; k++) {
      int v = q[16 * i + k];
      for (int j = 0; j < nls[i]; j++) v = (v << 1) + silk_icdf(ec, silk_lsb_iCDF, 8);
      q[16 * i + k] = (int16_t)v;
    }
  }
  const uint8_t* base = silk_sign_iCDF + 7 * (quant_offset + (signal_type << 1));
  const int nsb = (frame_length + 8) >> 4;
  for (int i = 0; i < nsb; i++) {
    if (sum[i] == 0 && nls[i] == 0) continue;   // sum_pulses |= nLshifts << 5 (decode_pulses.rs:143)
    const int ti = imin((int)sum[i], 6);        // (sum | nLshifts << 5) & 0x1f == sum (< 17)
    uint8_t tab[2]; tab[0] = base[ti]; tab[1] = 0;
    int end = 16 * i + 16;
    if ((quirks & S3_SIGNS_FRAMELEN) && end > frame_length) end = frame_length;
    for (int k = 16 * i; k < end; k++)
      if (q[k] > 0) {
        // one binary symbol with P(0) = tab[0]/256 (code_signs.rs:118-121)
        const uint32_t r = ec.rng >> 8, sthr = r * tab[0];
        int sym;
        if (ec.val >= sthr) { ec.val -= sthr; ec.rng -= sthr; sym = 0; } else { ec.rng = sthr; sym = 1; }
        ec.renorm();
        if (!sym) q[k] = (int16_t)-q[k];
      }
  }
}

// stereo_decode_pred.rs:21-56
MB_DEV_NOINLINE void silk_parse_stereo_pred(RangeDec& ec, int32_t pred_q13[2]) {
  int ix[2][3];
  const int n = silk_icdf(ec, silk_stereo_pred_joint_iCDF, 8);
  ix[0][2] = n / 5; ix[1][2] = n - 5 * ix[0][2];
  for (int c = 0; c < 2; c++) { ix[c][0] = silk_icdf(ec, silk_uniform3_iCDF, 8); ix[c][1] = silk_icdf(ec, silk_uniform5_iCDF, 8); }
  for (int c = 0; c < 2; c++) {
    const int k = ix[c][0] + 3 * ix[c][2];
    const int32_t low = silk_stereo_pred_quant_Q13[k];
    const int32_t step = (int32_t)(((int64_t)(silk_stereo_pred_quant_Q13[k + 1] - low) * 6554) >> 16);
    pred_q13[c] = low + (int32_t)(int16_t)step * (int32_t)(int16_t)(2 * ix[c][1] + 1);
  }
  pred_q13[0] -= pred_q13[1];
}

MB_DEV void silk_store_pulses(const int16_t* q, int frame_length, uint32_t* w /*stride 32*/) {
  for (int k = 0; 2 * k < frame_length; k++) w[k * 32] = (uint32_t)(uint16_t)q[2 * k] | ((uint32_t)(uint16_t)q[2 * k + 1] << 16);
}

// The SILK payload behind an open range decoder: flags, LBRR skip, and per frame the stereo predictor and each channel's
// side information + pulses, stored from slot `frames` on (frame i at frames + i * SL_FRAME_WORDS * 32 words).
// fec: the packet is decoded from its LBRR data (lost_flag == FLAG_DECODE_LBRR, dec_api.rs:200-330): the LBRR frames take the
// place of the regular ones in the record, a frame / channel without one is left for the concealment (coded = 0), and the
// regular frames are not read at all.
MB_DEV_NOINLINE void silk_parse_payload(RangeDec& ec, int fs_khz, int nb_subfr, int n_frames, int nci, uint32_t quirks, uint32_t* frames,
                                        int fec = 0) {
  const SilkRate R = silk_rate(fs_khz, nb_subfr);
  int vad[2][3], lbrr_flags[2][3], prev_type[2] = {0, 0}, prev_lag[2] = {0, 0};
  for (int n = 0; n < nci; n++) {                        // dec_api.rs:451-456
    for (int i = 0; i < n_frames; i++) vad[n][i] = ec.bit_logp(1);
    lbrr_flags[n][0] = ec.bit_logp(1);                   // (LBRR_flag, expanded below)
  }
  for (int n = 0; n < nci; n++) {
    const int flag = lbrr_flags[n][0];
    lbrr_flags[n][0] = lbrr_flags[n][1] = lbrr_flags[n][2] = 0;
    if (flag) {
      if (n_frames == 1) lbrr_flags[n][0] = 1;
      else {
        const int sym = silk_icdf(ec, n_frames == 2 ? silk_LBRR_flags_2_iCDF : silk_LBRR_flags_3_iCDF, 8) + 1;
        for (int i = 0; i < n_frames; i++) lbrr_flags[n][i] = (sym >> i) & 1;
      }
    }
  }
  int16_t q[336];
  SilkSide side;
  int decode_only_middle = 0;
  if (fec) {
    for (int i = 0; i < n_frames; i++) {
      uint32_t* fr = frames + (i * SL_FRAME_WORDS) * 32;
      int32_t pred[2] = {0, 0};
      uint32_t w2 = 0;
      if (nci == 2) {
        if (lbrr_flags[0][i]) {
          silk_parse_stereo_pred(ec, pred);
          w2 = lbrr_flags[1][i] == 0 ? (uint32_t)silk_icdf(ec, silk_stereo_only_code_mid_iCDF, 8) : 0u;
        } else w2 = 2u | 4u;   // SILK_FR_DOM_KEPT | SILK_FR_PRED_KEPT (mb_silk_out.cuh)
      }
      fr[0] = (uint32_t)pred[0]; fr[32] = (uint32_t)pred[1]; fr[64] = w2;
      for (int n = 0; n < nci; n++) {
        uint32_t* chw = fr + (3 + n * SL_CH_WORDS) * 32;
        if (lbrr_flags[n][i]) {
          const int cond = (i > 0 && lbrr_flags[n][i - 1]) ? SILK_COND_COND : SILK_COND_INDEP;
          silk_parse_indices(ec, R, 0, 1, cond, prev_type[n], prev_lag[n], side);
          silk_parse_pulses(ec, q, side.signal_type, side.quant_offset, R.frame_length, quirks);
          side.cond = (int8_t)cond; side.coded = 1;
          silk_side_store(side, chw);
          silk_store_pulses(q, R.frame_length, chw + SL_IDX_WORDS * 32);
        } else {
          side = SilkSide(); side.coded = 0;
          silk_side_store(side, chw);
        }
      }
    }
    return;
  }
  // LBRR payloads are decoded and dropped (dec_api.rs:476-529); they do move ec_prev_* of their channel
  for (int i = 0; i < n_frames; i++)
    for (int n = 0; n < nci; n++) {
      if (!lbrr_flags[n][i]) continue;
      if (nci == 2 && n == 0) {
        int32_t skip[2];
        silk_parse_stereo_pred(ec, skip);
        if (!lbrr_flags[1][i]) decode_only_middle = silk_icdf(ec, silk_stereo_only_code_mid_iCDF, 8);
      }
      const int cond = (i > 0 && lbrr_flags[n][i - 1]) ? SILK_COND_COND : SILK_COND_INDEP;
      silk_parse_indices(ec, R, 0, 1, cond, prev_type[n], prev_lag[n], side);
      silk_parse_pulses(ec, q, side.signal_type, side.quant_offset, R.frame_length, quirks);
    }
  // regular frames (dec_api.rs:200-330). prev_decode_only_middle of frame 0 is stream state: it only selects the
  // conditional-coding mode of the side channel for frame_index > 0, where it is the previous frame of this packet.
  int prev_dom = 0;
  for (int i = 0; i < n_frames; i++) {
    uint32_t* fr = frames + (i * SL_FRAME_WORDS) * 32;
    int32_t pred[2] = {0, 0};
    if (nci == 2) {
      silk_parse_stereo_pred(ec, pred);
      decode_only_middle = vad[1][i] == 0 ? silk_icdf(ec, silk_stereo_only_code_mid_iCDF, 8) : 0;
    } else decode_only_middle = 0;
    fr[0] = (uint32_t)pred[0]; fr[32] = (uint32_t)pred[1]; fr[64] = (uint32_t)decode_only_middle;
    for (int n = 0; n < nci; n++) {
      uint32_t* chw = fr + (3 + n * SL_CH_WORDS) * 32;
      if (n == 0 || !decode_only_middle) {
        // frame index seen by channel n: mid counts every frame, side counts from its own n_frames_decoded
        int cond;
        if (i == 0) cond = SILK_COND_INDEP;
        else if (n > 0 && prev_dom) cond = SILK_COND_INDEP_NO_LTP;
        else cond = SILK_COND_COND;
        silk_parse_indices(ec, R, vad[n][i], 0, cond, prev_type[n], prev_lag[n], side);
        silk_parse_pulses(ec, q, side.signal_type, side.quant_offset, R.frame_length, quirks);
        side.cond = (int8_t)cond; side.coded = 1;
        silk_side_store(side, chw);
        silk_store_pulses(q, R.frame_length, chw + SL_IDX_WORDS * 32);
      } else {
        side = SilkSide(); side.coded = 0;
        silk_side_store(side, chw);
      }
    }
    prev_dom = decode_only_middle;
  }
}

// One SILK-only Opus frame. `data`/`len`: the frame without its TOC. cap_per_ch: PCM capacity per channel at 48 kHz.
// One SILK-only Opus frame. `data`/`len`: the frame without its TOC. cap_per_ch: PCM capacity per channel at 48 kHz.
// Packets of one SILK frame (10 / 20 ms) keep their record where a hybrid cell keeps it (S_SILK_HYB: header, frame,
// final range in the last slot), which leaves the CELT slots of the cell free for the CELT frame such a packet may need
// next to it: the redundancy frame it carries (opus_decoder.rs:746-797) or the fade-out frame after a hybrid packet.
// Status bit 18 = that layout. `red` receives the redundancy split (bytes == 0: none).
struct SilkRedundancy { int celt_to_silk, cutoff, bytes; };
MB_DEV_NOINLINE uint32_t silk_parse_frame(const uint8_t* data, int len, uint32_t toc, uint32_t quirks, int cap_per_ch, uint32_t* rec_base,
                                          SilkRedundancy* red, int fec = 0) {
  const int bw = (int)((toc >> 5) & 3u);                 // 0 NB, 1 MB, 2 WB (3 = SWB never appears in SILK-only configs 0..11)
  const int sc = (int)((toc >> 3) & 3u);                 // 10, 20, 40, 60 ms
  const int fs_khz = bw == 0 ? 8 : (bw == 1 ? 12 : 16);
  const int n_frames = sc <= 1 ? 1 : sc;                 // dec_api.rs:351-359
  const int nb_subfr = sc == 0 ? 2 : 4;
  const int nci = (toc & 4u) ? 2 : 1;
  const int samples48 = (sc == 3 ? 2880 : (480 << sc));
  red->bytes = 0;
  if (samples48 > cap_per_ch) { rec_base[S_STATUS * 32] = (uint32_t)(uint8_t)(int8_t)MB_BUFFER_TOO_SMALL | (1u << 16); return rec_base[S_STATUS * 32]; }
  const int hyb_layout = n_frames == 1;
  RangeDec ec;
  ec.open(data, (uint32_t)len);
  silk_parse_payload(ec, fs_khz, nb_subfr, n_frames, nci, quirks, rec_base + (hyb_layout ? S_SILK_HYB + 1 : SL_BASE) * 32, fec);
  // SILK-only redundancy (opus_decoder.rs:746-797): with at least 17 bits left the frame ends in a 5 ms CELT frame; one
  // bit tells which side of the mode switch it belongs to, its size is whatever the range coder has not used.
  int code = MB_OK;
  if (!fec && ec.tell() + 17 <= 8 * len) {   // (a packet decoded from its LBRR data takes no redundancy frame, opus_decoder.rs:746)
    {   // (40 / 60 ms packets: their cell has no room for the CELT record; the caller parses it into an auxiliary cell)
      red->celt_to_silk = ec.bit_logp(1);
      red->bytes = len - ((ec.tell() + 7) >> 3);
      red->cutoff = len - red->bytes;
      if (red->bytes < 0 || red->bytes > len) { code = MB_BAD_ARG; red->bytes = 0; }
      else if (red->cutoff * 8 < ec.tell() || red->cutoff <= 1 || red->bytes < 2) { code = MB_UNIMPLEMENTED; red->bytes = 0; }  // degenerate splits
    }
  }
  const uint32_t h0 = (uint32_t)fs_khz | ((uint32_t)nb_subfr << 5) | ((uint32_t)n_frames << 8) | ((uint32_t)nci << 10);
  rec_base[S_HDR0 * 32] = h0;
  rec_base[S_RANGE * 32] = ec.rng;
  if (hyb_layout) { rec_base[S_SILK_HYB * 32] = h0; rec_base[(S_SILK_HYB + 342) * 32] = ec.rng; }
  const uint32_t status = (uint32_t)(uint8_t)(int8_t)code | (code == MB_OK ? (1u << 14) : 0u) | (1u << 16) | ((uint32_t)hyb_layout << 18) |
                          (fec ? (1u << 30) : 0u);
  rec_base[S_STATUS * 32] = status;
  return status;
}

}  // namespace mb
