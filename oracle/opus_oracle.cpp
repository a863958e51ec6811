// ORACLE (test infrastructure only — never linked into the product library).
// CPU restatement of the Opus decoder front-end:
//   TOC helpers / packet parse    src/packet.rs:30-276, 316-498
//   decode_native_with            src/opus_decoder.rs:1139-1367
//   decode_frame (CELT-only arm)  src/opus_decoder.rs:453-1133
//   opus_decode_float / opus_decode (i16: soft clip + float2int16)
//                                 src/opus_decoder.rs:1632-1715, 1803-1826; src/opus.rs:16-140;
//                                 src/celt/float_cast.rs:38-41
// SILK-only / Hybrid frames return OPUS_UNIMPLEMENTED (-5) until the SILK oracle lands
// (SURVEY.md §8 rows A16, C3/C4). Packet-loss concealment likewise (row F2).
// Exposes a plain C API (orc_*) that tests/ drive through ctypes.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <thread>
#include <vector>

#include "celt_oracle.h"
#include "silk_oracle.h"

namespace orc {

enum { MODE_SILK_ONLY = 1000, MODE_HYBRID = 1001, MODE_CELT_ONLY = 1002 };
enum { BW_NB = 1101, BW_MB = 1102, BW_WB = 1103, BW_SWB = 1104, BW_FB = 1105 };
enum { OPUS_BAD_ARG = -1, OPUS_BUFFER_TOO_SMALL = -2, OPUS_INTERNAL_ERROR = -3, OPUS_INVALID_PACKET = -4, OPUS_UNIMPLEMENTED = -5 };

static int packet_mode(uint8_t toc) { return (toc & 0x80) ? MODE_CELT_ONLY : ((toc & 0x60) == 0x60 ? MODE_HYBRID : MODE_SILK_ONLY); }
static int packet_bandwidth(uint8_t toc) {  // packet.rs:177-204
  if (toc & 0x80) { int b = (toc >> 5) & 3; return b == 0 ? BW_NB : b == 1 ? BW_WB : b == 2 ? BW_SWB : BW_FB; }
  if ((toc & 0x60) == 0x60) return (toc & 0x10) ? BW_FB : BW_SWB;
  int b = (toc >> 5) & 3;
  return b == 0 ? BW_NB : b == 1 ? BW_MB : b == 2 ? BW_WB : BW_SWB;
}
static int packet_samples_per_frame(uint8_t toc, int fs) {  // packet.rs:212-236
  if (toc & 0x80) return (fs << ((toc >> 3) & 3)) / 400;
  if ((toc & 0x60) == 0x60) return (toc & 0x08) ? fs / 50 : fs / 100;
  int sc = (toc >> 3) & 3;
  return sc == 3 ? fs * 60 / 1000 : (fs << sc) / 100;
}
static int parse_size(const uint8_t* d, int len, int* size) {
  if (len < 1) return -1;
  if (d[0] < 252) { *size = d[0]; return 1; }
  if (len < 2) return -1;
  *size = 4 * d[1] + d[0];
  return 2;
}

struct Parsed { int count; const uint8_t* frames[48]; int sizes[48]; int packet_offset; };

// packet.rs:316-498
static int packet_parse(const uint8_t* data, int len, int self_delimited, Parsed* out) {
  if (len == 0) return OPUS_INVALID_PACKET;
  int idx = 1, remaining = len - 1;
  uint8_t toc = data[0];
  int framesize = packet_samples_per_frame(toc, 48000);
  int cbr = 0, count, last_size = remaining, pad = 0;
  int* sz = out->sizes;
  memset(sz, 0, sizeof(out->sizes));
  switch (toc & 3) {
    case 0: count = 1; break;
    case 1:
      count = 2; cbr = 1;
      if (!self_delimited) {
        if (remaining & 1) return OPUS_INVALID_PACKET;
        last_size = remaining / 2;
        sz[0] = last_size;
      }
      break;
    case 2: {
      count = 2;
      int s, nb = parse_size(data + idx, len - idx, &s);
      if (nb < 0) return OPUS_INVALID_PACKET;
      if (s > (remaining >= nb ? remaining - nb : 0)) return OPUS_INVALID_PACKET;
      idx += nb; remaining -= nb;
      sz[0] = s;
      last_size = remaining - s;
      break;
    }
    default: {
      if (remaining == 0) return OPUS_INVALID_PACKET;
      uint8_t ch = data[idx++];
      remaining--;
      count = ch & 0x3F;
      if (count == 0 || count > 48 || framesize * count > 5760) return OPUS_INVALID_PACKET;
      if (ch & 0x40) {
        for (;;) {
          if (remaining == 0) return OPUS_INVALID_PACKET;
          int p = data[idx++];
          remaining--;
          int tmp = p == 255 ? 254 : p;
          pad += tmp;
          if (remaining < tmp) return OPUS_INVALID_PACKET;
          remaining -= tmp;
          if (p != 255) break;
        }
      }
      cbr = !(ch & 0x80);
      if (!cbr) {
        last_size = remaining;
        for (int i = 0; i < count - 1; i++) {
          int s, nb = parse_size(data + idx, len - idx, &s);
          if (nb < 0) return OPUS_INVALID_PACKET;
          if (s > (remaining >= nb ? remaining - nb : 0)) return OPUS_INVALID_PACKET;
          idx += nb; remaining -= nb;
          sz[i] = s;
          last_size -= nb + s;
          if (last_size < 0) return OPUS_INVALID_PACKET;
        }
      } else if (!self_delimited) {
        int per = remaining / count;
        if (per * count != remaining) return OPUS_INVALID_PACKET;
        last_size = per;
        for (int i = 0; i < count - 1; i++) sz[i] = per;
      }
    }
  }
  if (self_delimited) {
    int s, nb = parse_size(data + idx, len - idx, &s);
    if (nb < 0) return OPUS_INVALID_PACKET;
    if (s > (remaining >= nb ? remaining - nb : 0)) return OPUS_INVALID_PACKET;
    idx += nb; remaining -= nb;
    sz[count - 1] = s;
    if (cbr) {
      if (s * count > remaining) return OPUS_INVALID_PACKET;
      for (int i = 0; i < count - 1; i++) sz[i] = s;
    } else if (nb + s > last_size) return OPUS_INVALID_PACKET;
  } else {
    if (last_size < 0 || last_size > 1275) return OPUS_INVALID_PACKET;
    sz[count - 1] = last_size;
  }
  int cur = idx;
  for (int i = 0; i < count; i++) {
    if (cur + sz[i] > len) return OPUS_INVALID_PACKET;
    out->frames[i] = data + cur;
    cur += sz[i];
  }
  if (cur + pad > len) return OPUS_INVALID_PACKET;
  out->count = count;
  out->packet_offset = cur + pad;
  return 0;
}

struct OpusDecoder {
  int self_delimited = 0, last_packet_offset = 0;   // set by the multistream wrapper below
  int fs = 48000, channels = 2;
  CeltDecoder celt;
  SilkDecoder silk;
  SilkDecControl dec_control;
  int mode = 0, prev_mode = 0, bandwidth = 0, frame_size = 0, stream_channels = 0, prev_redundancy = 0;
  int last_packet_duration = 0, decode_gain = 0;
  uint32_t range_final = 0;
  int last_redundancy = 0;  // test hook: the last decoded frame carried a CELT redundancy frame
  float softclip_mem[2] = {0, 0};
  CeltTrace trace;

  int init(int fs_, int ch) {
    if (!(fs_ == 48000 || fs_ == 24000 || fs_ == 16000 || fs_ == 12000 || fs_ == 8000) || (ch != 1 && ch != 2)) return OPUS_BAD_ARG;
    fs = fs_; channels = ch;
    celt.init(ch, 48000 / fs_);
    silk.reset(1);
    dec_control = SilkDecControl();  // opus_decoder.rs:1438-1448
    dec_control.n_channels_api = ch; dec_control.n_channels_internal = 0; dec_control.api_sample_rate = fs_;
    dec_control.internal_sample_rate = 0; dec_control.payload_size_ms = 0;
    stream_channels = ch;
    frame_size = fs_ / 400;
    reset();
    return 0;
  }
  void reset() {  // opus_decoder.rs:1451-1476
    celt.reset();
    silk.reset(0);
    stream_channels = channels;
    bandwidth = 0; mode = 0; prev_mode = 0; frame_size = fs / 400; prev_redundancy = 0; last_packet_duration = 0;
    range_final = 0; softclip_mem[0] = softclip_mem[1] = 0;
  }

  // celt_decoder.rs decode entry used for redundancy / silence frames (no shared range decoder)
  int celt_frame(const uint8_t* data, int len, float* out, int n, int accum) { return celt_decode_with_ec(&celt, data, len, out, n, nullptr, accum); }

  // opus_decoder.rs:100-135
  void smooth_fade(const float* in1, const float* in2, float* out, int overlap) {
    int inc = 48000 / fs;
    for (int c = 0; c < channels; c++)
      for (int i = 0; i < overlap; i++) {
        if (i * inc >= 120) break;
        float w = celt_window120()[i * inc] * celt_window120()[i * inc];
        out[i * channels + c] = w * in2[i * channels + c] + (1.0f - w) * in1[i * channels + c];
      }
  }

  // opus_decoder.rs:453-1133
  int decode_frame(const uint8_t* data, int len, float* pcm, int fsz, int decode_fec) {
    const int F20 = fs / 50, F10 = F20 / 2, F5 = F10 / 2, F2_5 = F5 / 2;
    if (fsz < F2_5) return OPUS_BUFFER_TOO_SMALL;
    fsz = fsz < fs / 25 * 3 ? fsz : fs / 25 * 3;
    if (data == nullptr || len <= 1) { data = nullptr; fsz = fsz < frame_size ? fsz : frame_size; }   // :491-496: concealment
    int audiosize, md, bw, celt_only;
    if (data) { audiosize = frame_size; md = mode; bw = bandwidth; celt_only = md == MODE_CELT_ONLY; }
    else {
      audiosize = fsz;
      md = prev_redundancy ? MODE_CELT_ONLY : prev_mode;
      bw = 0;
      celt_only = md == MODE_CELT_ONLY;
      if (md == 0) {   // nothing decoded yet: silence
        for (int i = 0; i < audiosize * channels; i++) pcm[i] = 0.f;
        prev_mode = 0; prev_redundancy = 0; range_final = 0;
        return audiosize;
      }
      if (audiosize > F20) {   // conceal in chunks of at most 20 ms
        int done = 0;
        while (done < audiosize) {
          const int chunk = audiosize - done < F20 ? audiosize - done : F20;
          const int r = decode_frame(nullptr, 0, pcm + (size_t)done * channels, chunk, 0);
          if (r < 0) return r;
          if (r != chunk) return OPUS_INTERNAL_ERROR;
          done += r;
        }
        prev_mode = md; prev_redundancy = 0; range_final = 0;
        return audiosize;
      } else if (audiosize < F20) {
        if (audiosize > F10) audiosize = F10;
        else if (!celt_only && audiosize > F5 && audiosize < F10) audiosize = F5;
      }
    }
    EcDec dec;
    dec.init(data ? data : (const uint8_t*)"", data ? (uint32_t)len : 0u);
    const int prev_celt_only = prev_mode == MODE_CELT_ONLY;
    int transition = 0;
    std::vector<float> pcm_transition;
    if (data && prev_mode > 0 && ((celt_only && !prev_celt_only && !prev_redundancy) || (!celt_only && prev_celt_only))) {
      transition = 1;
      if (celt_only) {   // SILK / hybrid -> CELT without redundancy: 5 ms of the old mode's concealment to fade out of
        pcm_transition.assign((size_t)F5 * channels, 0.f);
        const int ts = audiosize < F5 ? audiosize : F5;
        const int r = decode_frame(nullptr, 0, pcm_transition.data(), ts, 0);
        if (r < 0) return r;
        if (r != ts) return OPUS_INTERNAL_ERROR;
      }
    }
    if (audiosize > fsz) return OPUS_BAD_ARG;
    int celt_accum = !celt_only, redundancy = 0, celt_to_silk = 0, have_range_final = 0;
    uint32_t redundant_rng = 0, silk_range_final = 0;
    const uint8_t* red_data = nullptr; int red_bytes = 0;
    std::vector<float> redundant_audio;
    if (!celt_only) {
      const int pcm_too_small = audiosize < F10;
      const int silk_len = pcm_too_small ? F10 : audiosize;
      int payload_ms = 1000 * audiosize / fs; if (payload_ms < 10) payload_ms = 10;
      dec_control.n_channels_api = channels; dec_control.api_sample_rate = fs; dec_control.payload_size_ms = payload_ms;
      if (data) {
        dec_control.n_channels_internal = stream_channels;
        dec_control.internal_sample_rate = md == MODE_SILK_ONLY ? (bw == BW_NB ? 8000 : (bw == BW_MB ? 12000 : 16000)) : 16000;
      } else if (dec_control.internal_sample_rate == 0 && silk.ch[0].fs_khz > 0) dec_control.internal_sample_rate = silk.ch[0].fs_khz * 1000;
      if (dec_control.n_channels_internal == 0) dec_control.n_channels_internal = stream_channels;
      if (prev_celt_only) silk.reset(0);
      std::vector<int16_t> so((size_t)silk_len * channels, 0);
      int decoded = 0;
      while (decoded < audiosize) {
        const int flag = data ? (decode_fec ? SILK_FLAG_DECODE_LBRR : SILK_FLAG_DECODE_NORMAL) : SILK_FLAG_PACKET_LOST;
        int r = silk.decode(&dec_control, flag, decoded == 0, &dec, so.data() + (size_t)decoded * channels);
        if (r < 0 || r == 0) {
          if (data) return OPUS_INTERNAL_ERROR;
          for (int i = decoded * channels; i < audiosize * channels; i++) so[i] = 0;   // :699-702: a failing concealment is silence
          r = audiosize - decoded;
        }
        decoded += r;
      }
      for (int i = 0; i < audiosize * channels; i++) pcm[i] = (float)so[i] * (1.0f / 32768.0f);  // silk_int16_to_res
      if (!decode_fec && data) {
        const int tell = dec.tell();
        if (tell + 17 + (md == MODE_HYBRID ? 20 : 0) <= 8 * len) {
          redundancy = md == MODE_HYBRID ? dec.bit_logp(12) : 1;
          if (redundancy) {
            celt_to_silk = dec.bit_logp(1);
            int bytes = md == MODE_HYBRID ? (int)dec.uint(256) + 2 : len - ((dec.tell() + 7) >> 3);
            if (bytes > len || bytes < 0) return OPUS_BAD_ARG;
            red_bytes = bytes;
            int cutoff = len - bytes;
            red_data = data + cutoff;
            len = cutoff;
            if (len * 8 < dec.tell()) { len = 0; redundancy = 0; red_data = nullptr; red_bytes = 0; }
            if (red_bytes > 0) {
              if (dec.storage < (uint32_t)red_bytes) return OPUS_BAD_ARG;
              dec.storage -= (uint32_t)red_bytes;
            }
          }
        }
      }
      if (data && len > 1 && (md == MODE_SILK_ONLY || decode_fec)) { silk_range_final = dec.rng; have_range_final = 1; }
      if (redundancy) transition = 0;
      else if (transition) {   // CELT -> SILK / hybrid without redundancy: 5 ms of CELT concealment to fade out of
        pcm_transition.assign((size_t)F5 * channels, 0.f);
        const int ts = audiosize < F5 ? audiosize : F5;
        const int r = decode_frame(nullptr, 0, pcm_transition.data(), ts, 0);
        if (r < 0) return r;
        if (r != ts) return OPUS_INTERNAL_ERROR;
      }
    }
    int start_band = celt_only ? 0 : 17;
    switch (bw) { case BW_NB: celt.end_band = 13; break; case BW_MB: case BW_WB: celt.end_band = 17; break; case BW_SWB: celt.end_band = 19; break;
                  case BW_FB: celt.end_band = 21; break; default: break; }   // bandwidth 0 (concealment): the end band stays
    celt.stream_channels = stream_channels;
    if (redundancy && celt_to_silk) {
      redundant_audio.assign((size_t)F5 * channels, 0.f);
      celt.start_band = 0;
      int r = celt_frame(red_data, red_bytes, redundant_audio.data(), F5, 0);
      if (r != F5) return r < 0 ? r : OPUS_INTERNAL_ERROR;
      redundant_rng = celt.rng;
    }
    celt.start_band = start_band;
    uint32_t celt_final_range = 0;
    if (md == MODE_SILK_ONLY) {
      if (prev_mode == MODE_HYBRID && !(redundancy && celt_to_silk && prev_redundancy)) {
        celt.start_band = 0;
        static const uint8_t silence[2] = {0xFF, 0xFF};
        int r = celt_frame(silence, 2, pcm, F2_5, celt_accum);
        if (r != F2_5) return r < 0 ? r : OPUS_INTERNAL_ERROR;
      }
    } else {
      if (md != prev_mode && prev_mode > 0 && !prev_redundancy) celt.reset();
      int celt_n = audiosize < F20 ? audiosize : F20;
      const uint8_t* celt_data = decode_fec ? nullptr : data;
      int ret = celt_decode_with_ec(&celt, celt_data, celt_data ? len : 0, pcm, celt_n, celt_data ? &dec : nullptr, celt_accum);
      if (ret < 0) return ret;
      if (ret != celt_n) return OPUS_INTERNAL_ERROR;
      if (data && len > 1 && celt_data) celt_final_range = celt.rng;
    }
    if (redundancy && !celt_to_silk) {
      redundant_audio.assign((size_t)F5 * channels, 0.f);
      celt.reset();
      celt.start_band = 0;
      int r = celt_frame(red_data, red_bytes, redundant_audio.data(), F5, 0);
      if (r != F5) return r < 0 ? r : OPUS_INTERNAL_ERROR;
      redundant_rng = celt.rng;
    }
    int fade = F2_5 * channels;
    if (redundancy) {
      if (!celt_to_silk) {
        float* p = pcm + (size_t)(audiosize - F2_5) * channels;
        std::vector<float> cur(p, p + fade);
        smooth_fade(cur.data(), redundant_audio.data() + fade, p, F2_5);
      } else if (prev_mode != MODE_SILK_ONLY || prev_redundancy) {
        for (int i = 0; i < fade; i++) pcm[i] = redundant_audio[i];
        std::vector<float> tail(pcm + fade, pcm + 2 * fade);
        smooth_fade(redundant_audio.data() + fade, tail.data(), pcm + fade, F2_5);
      }
    } else if (transition && !pcm_transition.empty()) {
      if (audiosize >= F5) {
        for (int i = 0; i < fade; i++) pcm[i] = pcm_transition[i];
        std::vector<float> tail(pcm + fade, pcm + 2 * fade);
        smooth_fade(pcm_transition.data() + fade, tail.data(), pcm + fade, F2_5);
      } else {
        std::vector<float> cur(pcm, pcm + fade);
        smooth_fade(pcm_transition.data(), cur.data(), pcm, F2_5);
      }
    }
    const int have_packet = data != nullptr && len > 1;
    uint32_t final_range = have_packet ? (have_range_final ? silk_range_final : celt_final_range) : 0;
    last_redundancy = redundancy;
    prev_mode = md;
    prev_redundancy = redundancy && !celt_to_silk;
    range_final = have_packet ? final_range ^ redundant_rng : 0;
    return audiosize;
  }

  // opus_decoder.rs:1139-1367
  int decode_native(const uint8_t* data, int len, float* pcm, int fsz, int decode_fec, int soft_clip) {
    if (fsz <= 0) return OPUS_BAD_ARG;
    if ((decode_fec || len == 0 || data == nullptr) && fsz % (fs / 400) != 0) return OPUS_BAD_ARG;
    if (len == 0 || data == nullptr) {   // packet loss (:1186-1222): conceal the whole buffer, no soft clip
      int done = 0;
      while (done < fsz) {
        const int r = decode_frame(nullptr, 0, pcm + (size_t)done * channels, fsz - done, 0);
        if (r < 0) return r;
        if (r == 0) return OPUS_INTERNAL_ERROR;
        done += r;
      }
      last_packet_duration = done;
      apply_gain_clip(pcm, done, 0);
      return done;
    }
    uint8_t toc = data[0];
    int pmode = packet_mode(toc), pbw = packet_bandwidth(toc), pfs = packet_samples_per_frame(toc, fs);
    int pch = (toc & 4) ? 2 : 1;
    Parsed P;
    int r = packet_parse(data, len, self_delimited, &P);   // (self-delimited framing: substreams of a multistream packet, :1139-1160)
    if (r < 0) return r;
    last_packet_offset = P.packet_offset;
    if (decode_fec) {   // :1240-1316
      if (fsz < pfs || pmode == MODE_CELT_ONLY || mode == MODE_CELT_ONLY) return decode_native(nullptr, 0, pcm, fsz, 0, soft_clip);
      const int duration_copy = last_packet_duration;
      if (fsz != pfs) {
        const int lead = fsz - pfs;
        const int rr = decode_native(nullptr, 0, pcm, lead, 0, soft_clip);
        if (rr < 0) { last_packet_duration = duration_copy; return rr; }
        if (rr != lead) { last_packet_duration = duration_copy; return OPUS_INTERNAL_ERROR; }
      }
      mode = pmode; bandwidth = pbw; frame_size = pfs; stream_channels = pch;
      const int ret = decode_frame(P.frames[0], P.sizes[0], pcm + (size_t)(fsz - pfs) * channels, pfs, 1);
      if (ret < 0) return ret;
      last_packet_duration = fsz;
      apply_gain_clip(pcm, fsz, 0);
      return fsz;
    }
    if (P.count * pfs > fsz) return OPUS_BUFFER_TOO_SMALL;
    mode = pmode; bandwidth = pbw; frame_size = pfs; stream_channels = pch;
    int nb = 0;
    for (int i = 0; i < P.count; i++) {
      int ret = decode_frame(P.frames[i], P.sizes[i], pcm + nb * channels, fsz - nb, 0);
      if (ret < 0) return ret;
      nb += ret;
    }
    last_packet_duration = nb;
    apply_gain_clip(pcm, nb, soft_clip);
    return nb;
  }

  void apply_gain_clip(float* pcm, int n, int soft_clip) {  // opus_decoder.rs:1376-1418
    if (decode_gain != 0) {
      const float scale = 3.32192809488736234787f / 5120.0f;  // DECODE_GAIN_SCALE opus_decoder.rs:53
      float g = (float)exp(0.6931471805599453094 * (double)(scale * (float)decode_gain));
      for (int i = 0; i < n * channels; i++) pcm[i] *= g;
    }
    if (soft_clip) pcm_soft_clip(pcm, n, channels, softclip_mem);
    else softclip_mem[0] = softclip_mem[1] = 0;
  }

  // src/opus.rs:16-140 (opus_pcm_soft_clip_impl)
  static void pcm_soft_clip(float* x_, int N, int C, float* declip_mem) {
    if (C < 1 || N < 1) return;
    for (int i = 0; i < N * C; i++) x_[i] = x_[i] < -2.f ? -2.f : (x_[i] > 2.f ? 2.f : x_[i]);
    for (int c = 0; c < C; c++) {
      float* x = x_ + c;
      float a = declip_mem[c];
      for (int i = 0; i < N; i++) {
        if (x[i * C] * a >= 0) break;
        x[i * C] = x[i * C] + a * x[i * C] * x[i * C];
      }
      int curr = 0;
      float x0 = x[0];
      for (;;) {
        int i;
        for (i = curr; i < N; i++) if (!(x[i * C] >= -1.f && x[i * C] <= 1.f)) break;
        if (i == N) { a = 0; break; }
        int peak_pos = i, start = i, end = i;
        float maxval = fabsf(x[i * C]);
        while (start > 0 && x[i * C] * x[(start - 1) * C] >= 0) start--;
        while (end < N && x[i * C] * x[end * C] >= 0) {
          if (fabsf(x[end * C]) > maxval) { maxval = fabsf(x[end * C]); peak_pos = end; }
          end++;
        }
        int special = (start == 0 && x[i * C] * x[0] >= 0);
        a = (maxval - 1) / (maxval * maxval);
        a += a * 2.4e-7f;
        if (x[i * C] > 0) a = -a;
        for (i = start; i < end; i++) x[i * C] = x[i * C] + a * x[i * C] * x[i * C];
        if (special && peak_pos >= 2) {
          float offset = x0 - x[0];
          float delta = offset / (float)peak_pos;
          for (i = curr; i < peak_pos; i++) {
            offset -= delta;
            x[i * C] += offset;
            x[i * C] = x[i * C] < -1.f ? -1.f : (x[i * C] > 1.f ? 1.f : x[i * C]);
          }
        }
        curr = end;
        if (curr == N) break;
      }
      declip_mem[c] = a;
    }
  }
};

static inline int16_t float2int16(float v) {  // float_cast.rs:38-41
  float s = v * 32768.0f;
  s = s < -32768.0f ? -32768.0f : (s > 32767.0f ? 32767.0f : s);
  return (int16_t)lrintf(s);
}

}  // namespace orc

using namespace orc;

extern "C" {

void* orc_create(int fs, int channels, uint32_t quirks) {
  OpusDecoder* d = new OpusDecoder();
  if (d->init(fs, channels) < 0) { delete d; return nullptr; }
  d->celt.quirks = quirks;
  d->silk.quirks = quirks;
  return d;
}
void orc_destroy(void* h) { delete (OpusDecoder*)h; }
void orc_reset(void* h) { ((OpusDecoder*)h)->reset(); }
void orc_set_gain(void* h, int gain_q8) { ((OpusDecoder*)h)->decode_gain = gain_q8; }
uint32_t orc_final_range(void* h) { return ((OpusDecoder*)h)->range_final; }
int orc_last_redundancy(void* h) { return ((OpusDecoder*)h)->last_redundancy; }
int orc_last_packet_duration(void* h) { return ((OpusDecoder*)h)->last_packet_duration; }
int orc_bandwidth(void* h) { return ((OpusDecoder*)h)->bandwidth; }                                  // OPUS_GET_BANDWIDTH (opus_decoder.rs:1832)
int orc_pitch(void* h) {                                                                              // OPUS_GET_PITCH (opus_decoder.rs:1840-1849)
  OpusDecoder* d = (OpusDecoder*)h;
  return d->prev_mode == MODE_CELT_ONLY ? d->celt.postfilter_period : d->dec_control.prev_pitch_lag;
}
// opus_decode24 (opus_decoder.rs:1718-1800): float decode without soft clip, then res_to_int24 (celt_decoder.rs:3165-3176)
int orc_decode_i24(void* h, const uint8_t* data, int len, int32_t* pcm, int frame_size, int fec) {
  OpusDecoder* d = (OpusDecoder*)h;
  if (frame_size <= 0) return OPUS_BAD_ARG;
  std::vector<float> tmp((size_t)frame_size * d->channels, 0.f);
  const int r = d->decode_native(data, len, tmp.data(), frame_size, fec, 0);
  for (int i = 0; i < r * d->channels; i++) {
    float v = tmp[i] * (32768.0f * 256.0f);
    v = v < -8388608.0f ? -8388608.0f : (v > 8388607.0f ? 8388607.0f : v);
    pcm[i] = (int32_t)lrintf(v);
  }
  return r;
}

// opus_decode_float (opus_decoder.rs:1803-1826)
int orc_decode_float(void* h, const uint8_t* data, int len, float* pcm, int frame_size, int fec) {
  return ((OpusDecoder*)h)->decode_native(data, len, pcm, frame_size, fec, 0);
}
// opus_decode (i16; opus_decoder.rs:1632-1715): decode to float scratch with soft clip, then float2int16.
int orc_decode_i16(void* h, const uint8_t* data, int len, int16_t* pcm, int frame_size, int fec) {
  OpusDecoder* d = (OpusDecoder*)h;
  if (frame_size <= 0) return OPUS_BAD_ARG;
  // (the reference limits frame_size to the packet's sample count when a packet is present)
  if (data && len > 0 && !fec) {
    int count = (data[0] & 3) == 0 ? 1 : ((data[0] & 3) != 3 ? 2 : (len < 2 ? -1 : (data[1] & 0x3F)));
    if (count < 0) return OPUS_INVALID_PACKET;
    int ns = count * packet_samples_per_frame(data[0], d->fs);
    if (ns * 25 > d->fs * 3) return OPUS_INVALID_PACKET;
    if (ns > 0 && ns < frame_size) frame_size = ns;
  }
  std::vector<float> tmp((size_t)frame_size * d->channels);
  int ret = d->decode_native(data, len, tmp.data(), frame_size, fec, 1);
  if (ret > 0) for (int i = 0; i < ret * d->channels; i++) pcm[i] = float2int16(tmp[i]);
  return ret;
}


// ---- SILK stand-alone hooks for the known-answer tests (tests/test_silk_oracle_kat.py) ----
void orc_silk_nlsf_decode(int wb, const int8_t* indices, int16_t* nlsf_q15, uint32_t quirks) {
  silk_nlsf_decode(nlsf_q15, indices, wb ? silk_cb_wb() : silk_cb_nb_mb(), quirks);
}
void orc_silk_nlsf2a(const int16_t* nlsf_q15, int order, int16_t* a_q12, uint32_t quirks) { silk_nlsf2a(a_q12, nlsf_q15, order, quirks); }
int32_t orc_silk_inverse_pred_gain(const int16_t* a_q12, int order) { return order == 0 ? 1 << 30 : silk_lpc_inverse_pred_gain(a_q12, order); }
int32_t orc_silk_log2lin(int32_t x) { return silk_log2lin(x); }
// opus_pcm_soft_clip (src/opus.rs:16-150) on its own, for the known answers of src/opus.rs:157-208
void orc_pcm_soft_clip(float* pcm, int frame_size, int channels, float* softclip_mem) { OpusDecoder::pcm_soft_clip(pcm, frame_size, channels, softclip_mem); }
int orc_silk_gains_dequant(const int8_t* ind, int prev, int conditional, int nb_subfr, int32_t* gains) {
  int8_t p = (int8_t)prev;
  silk_gains_dequant(gains, ind, &p, conditional, nb_subfr);
  return p;
}
void orc_silk_decode_pitch(int lag_index, int contour, int fs_khz, int nb_subfr, int32_t* lags) { silk_decode_pitch(lag_index, contour, lags, fs_khz, nb_subfr); }
// resampler: fresh state, one call of silk_resampler on `n` input samples; returns produced samples
int orc_silk_resample(int fs_in, int fs_out, const int16_t* in, int n, int16_t* out, int32_t* iir_state /*6, nullable*/) {
  SilkResampler r;
  if (r.init(fs_in, fs_out) < 0) return -1;
  int p = r.run(out, in, n);
  if (iir_state) memcpy(iir_state, r.s_iir, sizeof(r.s_iir));
  return p;
}
void orc_set_quirks(void* h, uint32_t quirks) { OpusDecoder* d = (OpusDecoder*)h; d->celt.quirks = quirks; d->silk.quirks = quirks; }

// Stage dump of the LAST decoded CELT frame (enable before decoding).
void orc_trace_enable(void* h, int on) { OpusDecoder* d = (OpusDecoder*)h; d->celt.trace = on ? &d->trace : nullptr; }
int orc_trace_info(void* h, int* out /*16 ints*/) {
  CeltTrace& t = ((OpusDecoder*)h)->trace;
  int v[16] = {t.lm, t.C, t.N, t.silence, t.transient, t.intra, t.spread, t.anti_collapse_on, t.intensity, t.dual_stereo,
               t.coded_bands, t.pf_pitch, t.pf_tapset, (int)t.seed_out, 0, 0};
  memcpy(out, v, sizeof(v));
  memcpy(&out[14], &t.pf_gain, 4);
  return 0;
}
int orc_trace_get(void* h, int what, void* dst, int max_elems) {
  CeltTrace& t = ((OpusDecoder*)h)->trace;
  const void* src = nullptr; size_t n = 0, es = 4;
  switch (what) {
    case 0: src = t.iy.data(); n = t.iy.size(); break;
    case 1: src = t.X.data(); n = t.X.size(); break;
    case 2: src = t.bandE.data(); n = t.bandE.size(); break;
    case 3: src = t.syn.data(); n = t.syn.size(); break;
    case 4: src = t.post.data(); n = t.post.size(); break;
    case 5: src = t.tf_res; n = NB_EBANDS; break;
    case 6: src = t.pulses; n = NB_EBANDS; break;
    case 7: src = t.fine_quant; n = NB_EBANDS; break;
    default: return -1;
  }
  if ((int)n > max_elems) n = max_elems;
  memcpy(dst, src, n * es);
  return (int)n;
}

// Batch helper used by tests and by bench.py's cpu_baseline leg: decodes a [n_ticks][n_lanes] packet grid
// (packet p = t*n_lanes + j belongs to lane j) with one fresh decoder per lane, `threads` std::threads,
// lanes statically partitioned. pcm layout: [p][frame_size*channels]. Returns 0.
int orc_decode_grid(int fs, int channels, uint32_t quirks, const uint8_t* arena, const uint32_t* offsets, int n_lanes,
                    int n_ticks, int frame_size, float* pcm /*nullable*/, int16_t* pcm16 /*nullable*/, int32_t* samples_out,
                    uint32_t* final_range_out, int threads) {
  if (threads < 1) threads = 1;
  auto work = [&](int tid) {
    std::vector<float> buf((size_t)frame_size * channels);
    std::vector<int16_t> buf16((size_t)frame_size * channels);
    for (int j = tid; j < n_lanes; j += threads) {
      OpusDecoder d;
      d.init(fs, channels);
      d.celt.quirks = quirks;
      d.silk.quirks = quirks;
      for (int t = 0; t < n_ticks; t++) {
        size_t p = (size_t)t * n_lanes + j;
        const uint8_t* pkt = arena + offsets[p];
        int len = (int)(offsets[p + 1] - offsets[p]);
        int ret;
        if (pcm16) {
          int16_t* o = pcm16 + p * frame_size * channels;
          ret = orc_decode_i16(&d, pkt, len, o, frame_size, 0);
        } else {
          float* o = pcm ? pcm + p * frame_size * channels : buf.data();
          ret = d.decode_native(pkt, len, o, frame_size, 0, 0);
        }
        if (samples_out) samples_out[p] = ret;
        if (final_range_out) final_range_out[p] = d.range_final;
      }
    }
  };
  std::vector<std::thread> th;
  for (int i = 1; i < threads; i++) th.emplace_back(work, i);
  work(0);
  for (auto& t : th) t.join();
  return 0;
}

// ---- multistream decoder (row F4): opus_multistream_decode_native (src/opus_multistream.rs:717-950) — N substream decoders, the
// first N-1 substream packets in self-delimited framing, channel mapping as in get_left/right/mono_channel (:60-140).
struct MsDecoder {
  int fs = 48000, channels = 0, streams = 0, coupled = 0;
  uint8_t mapping[256];
  std::vector<OpusDecoder> dec;
};
void* orc_ms_create(int fs, int channels, int streams, int coupled, const uint8_t* mapping, uint32_t quirks) {
  if (channels < 1 || channels > 255 || streams < 1 || coupled < 0 || coupled > streams || streams + coupled > 255) return nullptr;
  for (int c = 0; c < channels; c++) if (mapping[c] != 255 && mapping[c] >= streams + coupled) return nullptr;
  MsDecoder* m = new MsDecoder;
  m->fs = fs; m->channels = channels; m->streams = streams; m->coupled = coupled;
  memcpy(m->mapping, mapping, channels);
  m->dec.resize(streams);
  for (int s = 0; s < streams; s++) {
    if (m->dec[s].init(fs, s < coupled ? 2 : 1) != 0) { delete m; return nullptr; }
    m->dec[s].celt.quirks = quirks; m->dec[s].silk.quirks = quirks;
  }
  return m;
}
void orc_ms_destroy(void* h) { delete (MsDecoder*)h; }
// pcm (float) or pcm16 (int16: the substreams decode with the soft clipper on, then float2int16): [frame_size][channels]
int orc_ms_decode(void* h, const uint8_t* data, int len, float* pcm, int16_t* pcm16, int frame_size, int fec) {
  MsDecoder* m = (MsDecoder*)h;
  if (frame_size <= 0) return OPUS_BAD_ARG;
  const int nb = m->streams;
  if (frame_size > m->fs / 25 * 3) frame_size = m->fs / 25 * 3;
  const bool do_plc = len == 0 || data == nullptr;
  if (!do_plc) {
    if (len < 2 * nb - 1) return OPUS_INVALID_PACKET;
    const uint8_t* cur = data; int remaining = len, samples = -1;     // opus_multistream_packet_validate (:717-752)
    for (int s = 0; s < nb; s++) {
      if (remaining == 0) return OPUS_INVALID_PACKET;
      Parsed P;
      const int r = packet_parse(cur, remaining, s + 1 != nb, &P);
      if (r < 0) return r;
      const int ns = P.count * packet_samples_per_frame(cur[0], m->fs);
      if (ns * 25 > m->fs * 3) return OPUS_INVALID_PACKET;
      if (samples >= 0 && samples != ns) return OPUS_INVALID_PACKET;
      samples = ns;
      cur += P.packet_offset; remaining -= P.packet_offset;
    }
    if (samples > frame_size) return OPUS_BUFFER_TOO_SMALL;
  }
  std::vector<float> scratch((size_t)2 * frame_size);
  const uint8_t* cur = data; int remaining = len, decoded_fs = frame_size;
  auto put = [&](int chan, const float* src, int stride, int n) {
    for (int i = 0; i < n; i++) {
      const float v = src ? src[(size_t)i * stride] : 0.f;
      if (pcm16) pcm16[(size_t)i * m->channels + chan] = float2int16(v); else pcm[(size_t)i * m->channels + chan] = v;
    }
  };
  for (int s = 0; s < nb; s++) {
    OpusDecoder& d = m->dec[s];
    d.self_delimited = s + 1 != nb;
    int ret;
    if (do_plc) ret = d.decode_native(nullptr, 0, scratch.data(), decoded_fs, fec, pcm16 != nullptr);
    else {
      if (remaining == 0) return OPUS_INTERNAL_ERROR;
      ret = d.decode_native(cur, remaining, scratch.data(), decoded_fs, fec, pcm16 != nullptr);
      if (ret >= 0) { cur += d.last_packet_offset; remaining -= d.last_packet_offset; }
    }
    d.self_delimited = 0;
    if (ret < 0) return ret;
    if (ret == 0) return OPUS_INTERNAL_ERROR;
    decoded_fs = ret;
    for (int c = 0; c < m->channels; c++) {
      if (s < m->coupled) {
        if (m->mapping[c] == 2 * s) put(c, scratch.data(), 2, decoded_fs);
        else if (m->mapping[c] == 2 * s + 1) put(c, scratch.data() + 1, 2, decoded_fs);
      } else if (m->mapping[c] == s + m->coupled) put(c, scratch.data(), 1, decoded_fs);
    }
  }
  for (int c = 0; c < m->channels; c++) if (m->mapping[c] == 255) put(c, nullptr, 0, decoded_fs);   // muted channels
  return decoded_fs;
}
uint32_t orc_ms_final_range(void* h) {   // OPUS_GET_FINAL_RANGE of a multistream decoder: xor over the substreams (:640-660)
  MsDecoder* m = (MsDecoder*)h; uint32_t r = 0;
  for (auto& d : m->dec) r ^= d.range_final;
  return r;
}

// ---- projection decoder (row F4): opus_projection_decode / _decode_float (src/projection.rs:533-706) — a multistream decoder with
// the identity mapping whose decoded channels are accumulated through the demixing matrix (column-major int16, Q15;
// mapping_matrix_multiply_channel_out_float / _short, src/mapping_matrix.rs:188-283).
// x: [frame_size][cols] decoded channels (float; for the int16 form already soft-clipped); out: [frame_size][rows], zeroed here.
void orc_matrix_out_float(const int16_t* m, int rows, int cols, const float* x, int n_in, float* out, int frame_size) {
  for (int i = 0; i < frame_size * rows; i++) out[i] = 0.f;
  for (int c = 0; c < n_in; c++)
    for (int f = 0; f < frame_size; f++) {
      const float v = x[(size_t)f * cols + c];
      for (int r = 0; r < rows; r++) {
        const float coeff = (float)m[(size_t)rows * c + r] / 32768.0f;
        out[(size_t)f * rows + r] += coeff * v;
      }
    }
}
void orc_matrix_out_short(const int16_t* m, int rows, int cols, const float* x, int n_in, int16_t* out, int frame_size) {
  for (int i = 0; i < frame_size * rows; i++) out[i] = 0;
  for (int c = 0; c < n_in; c++)
    for (int f = 0; f < frame_size; f++) {
      const int32_t v = float2int16(x[(size_t)f * cols + c]);
      for (int r = 0; r < rows; r++) {
        const int32_t delta = ((int32_t)m[(size_t)rows * c + r] * v + 16384) >> 15;
        out[(size_t)f * rows + r] = (int16_t)((int32_t)out[(size_t)f * rows + r] + delta);
      }
    }
}
// h: a multistream decoder created with channels output channels and mapping[c] = c; matrix: rows = channels, cols = streams + coupled
int orc_proj_decode(void* h, const int16_t* matrix, const uint8_t* data, int len, float* pcm, int16_t* pcm16, int frame_size, int fec) {
  MsDecoder* m = (MsDecoder*)h;
  const int rows = m->channels, cols = m->streams + m->coupled;
  if (frame_size > m->fs / 25 * 3) frame_size = m->fs / 25 * 3;
  std::vector<float> x((size_t)frame_size * rows, 0.f);
  // (the substreams decode to float — with the soft clipper on for the int16 form — into the identity-mapped channels)
  int ret;
  {
    // decode float with soft clip: run the int16 path's substream decode but keep floats
    std::vector<float> scratch((size_t)2 * frame_size);
    const int nb = m->streams;
    const bool do_plc = len == 0 || data == nullptr;
    if (!do_plc) {
      if (len < 2 * nb - 1) return OPUS_INVALID_PACKET;
      const uint8_t* cur = data; int remaining = len, samples = -1;
      for (int s = 0; s < nb; s++) {
        if (remaining == 0) return OPUS_INVALID_PACKET;
        Parsed P;
        const int r = packet_parse(cur, remaining, s + 1 != nb, &P);
        if (r < 0) return r;
        const int ns = P.count * packet_samples_per_frame(cur[0], m->fs);
        if (ns * 25 > m->fs * 3) return OPUS_INVALID_PACKET;
        if (samples >= 0 && samples != ns) return OPUS_INVALID_PACKET;
        samples = ns;
        cur += P.packet_offset; remaining -= P.packet_offset;
      }
      if (samples > frame_size) return OPUS_BUFFER_TOO_SMALL;
    }
    const uint8_t* cur = data; int remaining = len, dfs = frame_size;
    for (int s = 0; s < nb; s++) {
      OpusDecoder& d = m->dec[s];
      d.self_delimited = s + 1 != nb;
      int r = do_plc ? d.decode_native(nullptr, 0, scratch.data(), dfs, fec, pcm16 != nullptr)
                     : d.decode_native(cur, remaining, scratch.data(), dfs, fec, pcm16 != nullptr);
      if (!do_plc && r >= 0) { cur += d.last_packet_offset; remaining -= d.last_packet_offset; }
      d.self_delimited = 0;
      if (r < 0) return r;
      if (r == 0) return OPUS_INTERNAL_ERROR;
      dfs = r;
      const int width = s < m->coupled ? 2 : 1, first = s < m->coupled ? 2 * s : s + m->coupled;
      for (int k = 0; k < width; k++)
        if (first + k < rows)
          for (int i = 0; i < dfs; i++) x[(size_t)i * rows + first + k] = scratch[(size_t)i * width + k];
    }
    ret = dfs;
  }
  const int n_in = cols < rows ? cols : rows;   // decoded channel c reaches the handler only when an output channel maps to it (c < channels)
  // x is laid out [frame][rows]; the matrix routines take the input stride as `cols`
  if (pcm16) orc_matrix_out_short(matrix, rows, rows, x.data(), n_in, pcm16, ret);
  else orc_matrix_out_float(matrix, rows, rows, x.data(), n_in, pcm, ret);
  return ret;
}

}  // extern "C"
