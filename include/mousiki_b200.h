/* mousiki_b200 — C ABI of the B200 batched Opus decoder.
 *
 * This is the drop-in boundary for ONE path of cijiugechu/mousiki: Opus packet decode.
 * Each entry point names the reference interface it replaces (paths relative to the
 * reference tree). Signatures use plain pointers and sizes only; a Rust / C / Python
 * host binds them directly (see INTEGRATION.md for the `extern "C"` block a mousiki
 * maintainer would add).
 *
 * Semantics per stream are those of `mousiki::Decoder` (src/codec.rs:792-977) /
 * `c_style_api::opus_decoder::opus_decode_float` (src/opus_decoder.rs:1803-1826):
 * one decoder state per stream id, packets of a stream are decoded in the order they
 * appear in the call, output is interleaved PCM, errors are reported per packet with
 * libopus codes (src/opus_decoder.rs:275-286) and never abort the batch.
 *
 * There is no CPU fallback: every decode entry point runs CUDA kernels on the device
 * chosen at bd_create and fails with BD_ERR_CUDA if that is impossible.
 */
#ifndef MOUSIKI_B200_H
#define MOUSIKI_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* API-level return codes (per-packet results use the Opus codes below) */
#define BD_OK 0
#define BD_ERR_BAD_ARG (-1)        /* OpusDecoderInitError::BadArgument, src/opus_decoder.rs:1550-1556 */
#define BD_ERR_CUDA (-100)         /* no usable device / kernel launch or copy failed (see bd_last_error) */
#define BD_ERR_ALLOC (-101)
#define BD_ERR_UNIMPLEMENTED (-5)
#define BD_ERR_TOO_LARGE (-102)    /* bd_parse_noalloc: more packets than its static scratch holds */

/* per-packet codes written to samples_out[] when negative (OpusDecodeError::code, src/opus_decoder.rs:275-286) */
#define BD_OPUS_BAD_ARG (-1)
#define BD_OPUS_BUFFER_TOO_SMALL (-2)
#define BD_OPUS_INTERNAL_ERROR (-3)
#define BD_OPUS_INVALID_PACKET (-4)
#define BD_OPUS_UNIMPLEMENTED (-5)

/* Behaviour switches: bit set = mousiki's decoder as written, clear = libopus (DESIGN.md "Quirks").
 * bits 0..6: CELT Q1,Q2,Q5..Q9; bits 8..11: SILK S1 (nlsf_decode.rs:109 16-bit SMULBB), S2 (nlsf2a.rs:42-60 cosine
 * ordering), S3 (code_signs.rs:112-118 sign decode stops at frame_length), S4 (dec_api.rs:150-154 mono->stereo
 * re-initialises both channels).
 * Q2 (bit 1, mdct.rs:373 zero fill) cleared = libopus' overlap-add between blocks: coded CELT-only frames decode, lost packets,
 * FEC and packets with a SILK part report BD_OPUS_UNIMPLEMENTED per packet. */
#define BD_QUIRKS_MOUSIKI 0xf7fu

typedef struct bd_handle bd_handle;

typedef struct bd_config {
  uint32_t struct_size;              /* sizeof(bd_config), for forward compatibility */
  uint32_t n_streams;                /* decoder states to allocate (stream ids are 0..n_streams-1) */
  int32_t sample_rate_hz;            /* Decoder::new(sample_rate, ..): 48000, 24000, 16000, 12000 or 8000 (CELT's down-sampling,
                                      * celt_decoder.rs:2232-2297; every mode of the SILK resampler, resampler.rs:87-176); sample
                                      * counts and pcm_stride are in output samples */
  int32_t channels;                  /* Channels::Mono = 1 / Channels::Stereo = 2 */
  int32_t gain_q8;                   /* DecoderBuilder::gain / OPUS_SET_GAIN, Q8 dB (src/codec.rs:411-463) */
  int32_t disable_phase_inversion;   /* <0: decoder default (mono decoders disable it), 0/1 explicit */
  uint32_t quirks;                   /* BD_QUIRKS_MOUSIKI */
  int32_t device;                    /* CUDA device ordinal */
  uint32_t max_frames_per_pass;      /* 0 = default; bounds the per-frame scratch (~30 KB per frame) */
  uint32_t reserved;
} bd_config;

/* Decoder::new / DecoderBuilder::build for n_streams decoders (src/codec.rs:792-801, src/opus_decoder.rs:1536-1603). */
int bd_create(const bd_config* cfg, bd_handle** out);
void bd_destroy(bd_handle* h);
/* Decoder::reset / OPUS_RESET_STATE (src/opus_decoder.rs:1451-1476). ids == NULL resets every stream. */
int bd_reset_streams(bd_handle* h, const uint32_t* stream_ids, uint32_t n);
const char* bd_last_error(const bd_handle* h);

/* Decoder::decode_float / opus_decode_float for a batch (src/codec.rs:892-907, src/opus_decoder.rs:1803-1826).
 * HOST buffers. Packet p is arena[offsets[p] .. offsets[p+1]) and belongs to stream_ids[p]; packets of one
 * stream are decoded in array order. flags[p] bit0 = decode_fec (Decoder::decode(.., fec = true): the in-band LBRR copy of
 * the packet BEFORE this one, or concealment when there is none, src/opus_decoder.rs:1186-1316), bit1 = packet lost
 * (Decoder::decode(None) / conceal*: concealment for the row's capacity, which must be a whole number of 2.5 ms frames);
 * a frame of <= 1 byte is concealed too (src/opus_decoder.rs:491-496); flags may be NULL. pcm receives packet p's interleaved samples at
 * pcm + p * pcm_stride (pcm_stride elements >= frame_size * channels; the per-channel capacity
 * pcm_stride / channels plays the role of `frame_size`). samples_out[p] = samples per channel (>= 0) or an
 * Opus error code; final_range_out[p] = Decoder::final_range() after that packet. Returns BD_OK or an API error.
 * Packets with several frames (TOC codes 1-3, padding: opus_packet_parse_impl, src/packet.rs:316-498) are cut into
 * frames on the host; the frames of a packet run in consecutive ticks of the stream and fill the packet's row in
 * order; a frame that fails ends the packet with its code, as the frame loop of src/opus_decoder.rs:1339-1350 does.
 * Mode switches: redundancy frames on SILK-only packets of every duration and on hybrid packets, the hybrid -> SILK-only
 * fade-out and mode switches without a redundancy frame, in both directions, decode (src/opus_decoder.rs:575-640, 746-797,
 * 868-893, 1012-1112). BD_OPUS_UNIMPLEMENTED is left for a hybrid redundancy split that leaves the SILK part no room and for a
 * pass that runs out of auxiliary cells; the stream is untouched then. */
int bd_decode_f32(bd_handle* h, const uint8_t* arena, const uint32_t* offsets, const uint32_t* stream_ids, uint32_t n_packets,
                  const uint8_t* flags, float* pcm, uint32_t pcm_stride, int32_t* samples_out, uint32_t* final_range_out);
/* Decoder::decode / opus_decode (i16: soft clip + round; src/codec.rs:843-859, src/opus_decoder.rs:1632-1715). */
int bd_decode_i16(bd_handle* h, const uint8_t* arena, const uint32_t* offsets, const uint32_t* stream_ids, uint32_t n_packets,
                  const uint8_t* flags, int16_t* pcm, uint32_t pcm_stride, int32_t* samples_out, uint32_t* final_range_out);

/* 24-bit output in int32 containers (opus_decode24, src/opus_decoder.rs:1718-1800; Decoder::decode_24bit, src/codec.rs:867):
 * the float path without soft clip, then clamp(x * 2^23, -8388608, 8388607) rounded to nearest (res_to_int24,
 * src/celt/celt_decoder.rs:3165-3176). Same arguments as bd_decode_f32 with an int32 PCM buffer. */
int bd_decode_i24(bd_handle* h, const uint8_t* arena, const uint32_t* offsets, const uint32_t* stream_ids, uint32_t n_packets,
                  const uint8_t* flags, int32_t* pcm, uint32_t pcm_stride, int32_t* samples_out, uint32_t* final_range_out);

/* Device-resident form of the same call for callers that already hold packets in HBM and keep PCM there.
 * Regular grid: packet (t, j) = index t * n_lanes + j belongs to stream lane_stream[j] (NULL: stream j);
 * a zero-length packet marks "nothing for this lane at this tick". All pointers are DEVICE pointers;
 * cuda_stream is a cudaStream_t (NULL = default stream). The call returns after enqueueing; results are
 * valid once the stream is synchronised. d_pcm_f32 or d_pcm_i16 may be NULL (but not both non-NULL).
 * Every packet must hold ONE frame (TOC code 0) here: the host is not allowed to look at the bytes, so a packet with
 * codes 1-3 comes back as BD_OPUS_UNIMPLEMENTED; cut such packets into frames before uploading them. */
int bd_decode_grid_dev(bd_handle* h, const uint8_t* d_arena, const uint32_t* d_offsets, uint32_t n_lanes, uint32_t n_ticks,
                       const uint32_t* d_lane_stream, float* d_pcm_f32, int16_t* d_pcm_i16, uint32_t pcm_stride,
                       int32_t* d_samples_out, uint32_t* d_final_range_out, void* cuda_stream);

/* Decoder::{bandwidth,last_packet_duration,final_range} style queries (src/codec.rs:838-977); synchronises. */
typedef struct bd_stream_info {
  uint32_t final_range;
  int32_t last_packet_duration;
  int32_t prev_mode;      /* 0 none, 1000 SILK, 1001 hybrid, 1002 CELT */
  int32_t celt_error;
  int32_t bandwidth;      /* Decoder::bandwidth (src/codec.rs:941-953): 0 before the first packet, else 1101 NB .. 1105 FB */
  int32_t pitch;          /* Decoder::pitch (src/opus_decoder.rs:1840-1849): CELT post-filter period after a CELT-only packet,
                             else the SILK pitch lag of the last voiced frame at 48 kHz (0 if unvoiced) */
} bd_stream_info;
int bd_get_stream_info(bd_handle* h, uint32_t stream_id, bd_stream_info* out);

/* Packet queries without a decoder (src/packet.rs:30-56, 177-280; `Decoder::packet_samples`, src/codec.rs:838): host-only
 * byte work on the packet the caller holds. Each returns the value or a negative Opus error code: BD_OPUS_BAD_ARG for an
 * empty packet (or len == 0 / len > avail for the frame count), BD_OPUS_INVALID_PACKET for a code-3 packet without its
 * count byte or a packet longer than 120 ms. `avail` is the slice length, `len` the `len` argument of the reference. */
int bd_packet_get_mode(const uint8_t* packet, uint32_t avail);                 /* 1000 SILK, 1001 hybrid, 1002 CELT */
int bd_packet_get_bandwidth(const uint8_t* packet, uint32_t avail);            /* 1101 NB .. 1105 FB (Bandwidth::to_opus_int) */
int bd_packet_get_nb_channels(const uint8_t* packet, uint32_t avail);
int bd_packet_get_samples_per_frame(const uint8_t* packet, uint32_t avail, uint32_t fs_hz);
int bd_packet_get_nb_frames(const uint8_t* packet, uint32_t avail, uint32_t len);
int bd_packet_get_nb_samples(const uint8_t* packet, uint32_t avail, uint32_t len, uint32_t fs_hz);

/* Pinned host memory helpers so that a host language can stage packets/PCM for full-rate PCIe copies. */
void* bd_alloc_pinned(uint64_t bytes);
void bd_free_pinned(void* p);

/* Number of kernel launches issued by the last decode call (bench.py's gpu_launches). */
uint32_t bd_last_launch_count(const bd_handle* h);

/* Per-kernel device timing for the roofline report: when enabled, every kernel launch of the decode calls is
 * bracketed by CUDA events on the launching stream. bd_profile_read synchronises, returns the accumulated
 * milliseconds and launch counts per kernel id (arrays of BD_K_COUNT) since the last read, and resets them. */
enum { BD_K_PARSE = 0, BD_K_ENERGY = 1, BD_K_PVQ = 2, BD_K_PLAN = 3, BD_K_BANDS = 4, BD_K_IMDCT = 5, BD_K_COMB = 6, BD_K_DEEMPH = 7,
       BD_K_CLIP_I16 = 8, BD_K_SILK = 9, BD_K_COUNT = 10 };
int bd_profile_enable(bd_handle* h, int on);
int bd_profile_read(bd_handle* h, double* ms_out, uint32_t* launches_out);

/* Test hooks (parity suite): optional device-side capture of the decoded PVQ pulse vectors and normalised spectra
 * of the next bd_decode_* call; layout [packet][2][960] int32 / float. Pass 0 to disable. Host buffers are
 * filled when the call returns. */
int bd_debug_capture(bd_handle* h, int32_t* pulses_host, float* spectrum_host, uint32_t max_packets);

/* Test hook: replays one of the reference's unit known-answer cases on the device functions the kernels are built from
 * (one warp on `device`; protocol in mousiki_b200/csrc/mb_kat.cuh: 1 bitexact_cos, 2 bitexact_log2tan, 3 cwrsi,
 * 4 comb_filter, 5 ifft, 6 frame IMDCT, 7 range-decoder script, 8 pulse cache). Reference tests replayed:
 * src/celt/bands.rs:4805-4915, cwrs.rs:681-786, celt.rs:868-1190, kiss_fft.rs:1364-1440, mdct.rs:1746-1843,
 * entdec.rs:327-377, src/range.rs:659-695, rate.rs:1424-1465. Host word buffers in and out. */
int bd_debug_kat(int device, int what, const uint32_t* in, uint32_t n_in, uint32_t* out, uint32_t n_out);

/* Allocation-free entry point. No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Runs the entropy-parse kernel (phase 1: TOC, range decoder over the whole CELT frame) on n_packets stand-alone CELT-only
 * single-frame packets, each taken as the first packet of a fresh decoder of `channels` channels, and reports what that kernel
 * alone determines: samples_out[p] = samples per channel the packet decodes to at 48 kHz (or the negative Opus code bd_decode_*
 * would report; packets of other modes / frame counts: BD_OPUS_UNIMPLEMENTED) and final_range_out[p] = the decoder's final_range
 * after the packet (0 where bd_decode_* reports 0). d_arena / d_offsets (n_packets + 1) / d_samples_out / d_final_range_out are
 * DEVICE pointers owned by the caller; frame_size = the caller's PCM capacity per channel (the reference's `frame_size`);
 * cuda_stream = a cudaStream_t or NULL. The kernel is enqueued on that stream; the call does not synchronise, allocates nothing
 * and frees nothing. Its record scratch (about 24 KB per packet) is a file-scope device array of BD_NOALLOC_MAX_PACKETS records
 * (12.2 MB, held by every GPU that loads the library until the process exits); more packets: BD_ERR_TOO_LARGE. All calls on a
 * GPU share that array: calls that overlap on the GPU (different streams, or graph launches in flight together) race. */
#define BD_NOALLOC_MAX_PACKETS 512u
int bd_parse_noalloc(int device, const uint8_t* d_arena, const uint32_t* d_offsets, uint32_t n_packets, int channels, uint32_t quirks,
                     int frame_size, int32_t* d_samples_out, uint32_t* d_final_range_out, void* cuda_stream);

/* ---- Row A2: mousiki's LIGHTWEIGHT decoder (`mousiki::decoder::Decoder`, src/decoder.rs:142-262, the README's minimal API), batched.
 * It is a different decoder from the one above: the float, pion-derived SILK path of
 * src/silk/decoder.rs:971-1163 — SILK-only, mono, one 20 ms frame per packet, no LBRR — with a zero-order hold x3 to 48 kHz
 * (src/resample.rs:3-15). Every accepted packet yields 960 samples. One decoder per stream: lane j of a call is stream j, and a
 * call decodes n_ticks packets per lane in order (packet p = tick * n_lanes + lane; arena / offsets as in bd_decode_*; host
 * buffers, copies inside the call). status[p] = 960, or the reference's error for that packet (the stream's state is then
 * untouched and its row of `pcm` is cleared): */
#define BD_LITE_ERR_TOO_SHORT (-1)        /* DecoderError::TooShortForTableOfContents, src/decoder.rs:151-153 */
#define BD_LITE_ERR_FRAME_CODE (-2)       /* DecoderError::UnsupportedFrameCode, src/decoder.rs:167-173 */
#define BD_LITE_ERR_CONFIG_MODE (-3)      /* DecoderError::UnsupportedConfigurationMode (hybrid / CELT), src/decoder.rs:175-184 */
#define BD_LITE_ERR_FRAME_DURATION (-4)   /* silk DecodeError::UnsupportedFrameDuration (not 20 ms), src/silk/decoder.rs:987-993 */
#define BD_LITE_ERR_STEREO (-5)           /* silk DecodeError::StereoUnsupported, src/silk/decoder.rs:994-996 */
#define BD_LITE_ERR_LBRR (-7)             /* silk DecodeError::UnsupportedLowBitrateRedundancy, src/silk/decoder.rs:1021-1023 */
typedef struct bd_lite_handle bd_lite_handle;
/* Decoder::new() for n_streams streams on `device` (src/decoder.rs:143-148). */
int bd_lite_create(int device, uint32_t n_streams, bd_lite_handle** out);
void bd_lite_destroy(bd_lite_handle* h);
/* Every stream back to Decoder::new(). */
int bd_lite_reset(bd_lite_handle* h);
/* Decoder::decode_float32 (src/decoder.rs:235-248): pcm[n_packets][960] f32. */
int bd_lite_decode_f32(bd_lite_handle* h, const uint8_t* arena, const uint32_t* offsets, uint32_t n_lanes, uint32_t n_ticks, float* pcm,
                       int32_t* status);
/* Decoder::decode (src/decoder.rs:220-233, floor(x * 32767), src/bitdepth.rs:15-51): pcm[n_packets][960] i16. */
int bd_lite_decode_i16(bd_lite_handle* h, const uint8_t* arena, const uint32_t* offsets, uint32_t n_lanes, uint32_t n_ticks, int16_t* pcm,
                       int32_t* status);
/* Decoder::decode_int24 (src/decoder.rs:252-260, rint(x * 2^23), src/bitdepth.rs:58-90): pcm[n_packets][960] i32. */
int bd_lite_decode_i24(bd_lite_handle* h, const uint8_t* arena, const uint32_t* offsets, uint32_t n_lanes, uint32_t n_ticks, int32_t* pcm,
                       int32_t* status);

#ifdef __cplusplus
}
#endif
#endif
