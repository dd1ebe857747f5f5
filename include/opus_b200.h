/* opus_b200.h -- C ABI of libopus_b200.so, the B200-native batched Opus engine.
 *
 * Two groups of entry points:
 *
 *  1. The libopus decoder ABI that Concentus binds by P/Invoke
 *     (reference: CSharp/Concentus/Native/NativeOpus.cs:85-159; C prototypes
 *     opus-fix/include/opus.h:208-328,423-512).  Same names, argument meaning, return values and
 *     error codes, so the shared library can be dropped in as "opus".  A decoder handle owns a
 *     state slot in GPU memory; opus_decode() is a batch of one.
 *
 *  2. Batch entry points (new; SURVEY section 8b "New exports"): n independent streams decoded by
 *     one call -- semantics are those of n opus_decode() calls, results are bit-identical to the
 *     reference's fixed-point decoder, and every stream gets its own return code.
 *
 * All decoders taking part in one batch call must live on the same CUDA device (the device that
 * was current when they were created).  Calls on the same decoders must not overlap in time
 * (same rule as libopus: one state object, one caller).
 */
#ifndef OPUS_B200_H
#define OPUS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int32_t opus_int32;
typedef uint32_t opus_uint32;
typedef int16_t opus_int16;

/* Error codes (opus-fix/include/opus_defines.h:46-60) */
#define OPUS_OK 0
#define OPUS_BAD_ARG -1
#define OPUS_BUFFER_TOO_SMALL -2
#define OPUS_INTERNAL_ERROR -3
#define OPUS_INVALID_PACKET -4
#define OPUS_UNIMPLEMENTED -5
#define OPUS_INVALID_STATE -6
#define OPUS_ALLOC_FAIL -7

/* ctl requests understood by opus_decoder_ctl (opus_defines.h:130-167; opus_decoder.c:802-912) */
#define OPUS_GET_BANDWIDTH_REQUEST 4009
#define OPUS_RESET_STATE 4028
#define OPUS_GET_SAMPLE_RATE_REQUEST 4029
#define OPUS_GET_FINAL_RANGE_REQUEST 4031
#define OPUS_GET_PITCH_REQUEST 4033
#define OPUS_SET_GAIN_REQUEST 4034
#define OPUS_GET_LAST_PACKET_DURATION_REQUEST 4039
#define OPUS_GET_GAIN_REQUEST 4045

typedef struct OpusDecoder OpusDecoder;

/* ---- libopus ABI ------------------------------------------------------------------------ */

/* replaces opus_get_version_string (celt/celt.c:286-299; NativeOpus.cs:157).  Contains "-fixed". */
const char *opus_get_version_string(void);
/* replaces opus_strerror (celt/celt.c:268-284) */
const char *opus_strerror(int error);

/* replaces opus_decoder_get_size (src/opus_decoder.c:82-94): size of the HOST handle. */
int opus_decoder_get_size(int channels);
/* replaces opus_decoder_create (src/opus_decoder.c:139-166; NativeOpus.cs:88). */
OpusDecoder *opus_decoder_create(opus_int32 Fs, int channels, int *error);
/* replaces opus_decode (src/opus_decoder.c:713-719; NativeOpus.cs:94): data==NULL or len==0
 * requests packet-loss concealment. Returns samples per channel or a negative error code. */
int opus_decode(OpusDecoder *st, const unsigned char *data, opus_int32 len, opus_int16 *pcm,
                int frame_size, int decode_fec);
/* replaces opus_decode_float (src/opus_decoder.c:722-753; NativeOpus.cs:97): int16 result / 32768. */
int opus_decode_float(OpusDecoder *st, const unsigned char *data, opus_int32 len, float *pcm,
                      int frame_size, int decode_fec);
/* replaces opus_decoder_ctl (src/opus_decoder.c:789-912; NativeOpus.cs:100-106): one int or int*. */
int opus_decoder_ctl(OpusDecoder *st, int request, ...);
/* replaces opus_decoder_destroy (src/opus_decoder.c:914-917; NativeOpus.cs:91). */
void opus_decoder_destroy(OpusDecoder *st);

/* packet inspection helpers, host side (src/opus_decoder.c:921-981, src/opus.c:169-188,345-352) */
int opus_packet_get_bandwidth(const unsigned char *data);
int opus_packet_get_nb_channels(const unsigned char *data);
int opus_packet_get_nb_frames(const unsigned char packet[], opus_int32 len);
int opus_packet_get_samples_per_frame(const unsigned char *data, opus_int32 Fs);
int opus_packet_get_nb_samples(const unsigned char packet[], opus_int32 len, opus_int32 Fs);
int opus_decoder_get_nb_samples(const OpusDecoder *dec, const unsigned char packet[], opus_int32 len);
int opus_packet_parse(const unsigned char *data, opus_int32 len, unsigned char *out_toc,
                      const unsigned char *frames[48], opus_int16 size[48], int *payload_offset);

/* ---- repacketizer and padding (opus-fix/include/opus.h:886-981, src/repacketizer.c) ----
 * Host-side byte surgery, identical results to the reference: merge frames of equal configuration into one
 * packet (up to 120 ms), split them again, pad a packet to a fixed size (CBR transport) and strip padding. */
typedef struct OpusRepacketizer OpusRepacketizer;
int opus_repacketizer_get_size(void);
OpusRepacketizer *opus_repacketizer_init(OpusRepacketizer *rp);
OpusRepacketizer *opus_repacketizer_create(void);
void opus_repacketizer_destroy(OpusRepacketizer *rp);
int opus_repacketizer_cat(OpusRepacketizer *rp, const unsigned char *data, opus_int32 len);
opus_int32 opus_repacketizer_out_range(OpusRepacketizer *rp, int begin, int end, unsigned char *data,
                                       opus_int32 maxlen);
int opus_repacketizer_get_nb_frames(OpusRepacketizer *rp);
opus_int32 opus_repacketizer_out(OpusRepacketizer *rp, unsigned char *data, opus_int32 maxlen);
int opus_packet_pad(unsigned char *data, opus_int32 len, opus_int32 new_len);
opus_int32 opus_packet_unpad(unsigned char *data, opus_int32 len);
int opus_multistream_packet_pad(unsigned char *data, opus_int32 len, opus_int32 new_len, int nb_streams);
opus_int32 opus_multistream_packet_unpad(unsigned char *data, opus_int32 len, int nb_streams);

/* Index an opus_demo ".bit" image (opus-fix/src/opus_demo.c:750-760: per packet BE32 length, BE32 encoder
 * final range, payload) so that a recorded stream feeds the flat batch calls without per-packet host work:
 * offsets[i] / lens[i] address packet i inside `file` (-1 / 0 for a lost packet), final_ranges[i] is the check
 * value.  Returns the packet count (at most max_packets entries written) or OPUS_INVALID_PACKET. */
opus_int32 opus_b200_bit_index(const unsigned char *file, opus_int32 file_bytes, opus_int32 *offsets,
                               opus_int32 *lens, opus_uint32 *final_ranges, opus_int32 max_packets);

/* ---- batch ABI (no counterpart in libopus; semantics = n x opus_decode) -------------------- */

/* Pointer-array form.  data[i]==NULL or len[i]==0 => concealment for stream i.
 * pcm[i] must hold frame_size*channels samples.  decode_fec may be NULL (all 0).
 * ret[i] receives what opus_decode() would have returned for stream i.
 * Return value: OPUS_OK if the batch ran (look at ret[] for per-stream results), else an error
 * that prevented the whole batch (bad arguments, mixed devices, CUDA failure). */
int opus_decode_batch(OpusDecoder *const *st, const unsigned char *const *data, const opus_int32 *len,
                      opus_int16 *const *pcm, int frame_size, const int *decode_fec, int n, int *ret);

/* Flat host form: one packet arena in, one PCM arena out (no per-stream host copies).
 * Packet i occupies data[offsets[i] .. offsets[i]+len[i]); offsets[i] < 0 means "lost".
 * Stream i's PCM goes to pcm + i*pcm_stride (pcm_stride in int16 units, >= frame_size*channels).
 * Pinned host buffers make both copies asynchronous DMA. */
int opus_decode_batch_flat(OpusDecoder *const *st, int n, const unsigned char *data, opus_int32 data_bytes,
                           const opus_int32 *offsets, const opus_int32 *len, opus_int16 *pcm,
                           opus_int32 pcm_stride, int frame_size, const int *decode_fec, int *ret);

/* A decoder group pins down a fixed set of decoders (their GPU state pointers are uploaded once and
 * the per-stream scratch is allocated once), so that repeated batch calls cost no set-up. */
typedef struct OpusDecoderGroup OpusDecoderGroup;
OpusDecoderGroup *opus_decoder_group_create(OpusDecoder *const *st, int n, int *error);
void opus_decoder_group_destroy(OpusDecoderGroup *grp);

/* Flat host form on a group (same meaning as opus_decode_batch_flat). */
int opus_decode_group_flat(OpusDecoderGroup *grp, const unsigned char *data, opus_int32 data_bytes,
                           const opus_int32 *offsets, const opus_int32 *len, opus_int16 *pcm,
                           opus_int32 pcm_stride, int frame_size, const int *decode_fec, int *ret);

/* Pipelined host form.  opus_decode_group_submit() queues the host->device copy of the packets, the
 * kernels and the device->host copy of PCM + return values on separate streams and returns; up to three calls
 * per group are in flight (a fourth submit waits for the oldest), so the copies of one call overlap with the
 * kernels of its neighbours.  pcm / ret are complete after opus_decode_group_wait().  All buffers must stay
 * alive until then and should be page-locked (cudaHostRegister / cudaMallocHost) for the overlap to happen.
 * Calls on one group run in submission order, so decoder state advances exactly as with the synchronous
 * form.  frame_size <= 0 is answered with OPUS_BAD_ARG for the whole call. */
int opus_decode_group_submit(OpusDecoderGroup *grp, const unsigned char *data, opus_int32 data_bytes,
                             const opus_int32 *offsets, const opus_int32 *len, opus_int16 *pcm,
                             opus_int32 pcm_stride, int frame_size, const int *decode_fec, int *ret);
int opus_decode_group_wait(OpusDecoderGroup *grp);

/* Device-resident form: every data pointer is a DEVICE pointer on the group's device; nothing is
 * copied and nothing is synchronised -- the work is queued on `cuda_stream` (a cudaStream_t), in order, or, with
 * NULL, on the library's own streams, where consecutive calls are PIPELINED: everything of a call that reads
 * packets (plan, routing, SILK layer, the parse kernel) runs ahead, beside the rebuild / synthesis kernels of the
 * call before it.  Decoder state advances exactly as with in-order calls; but because the SILK layer of call k+1
 * may be written while call k still adds its CELT layer, consecutive NULL-stream calls must be given different
 * d_pcm buffers (two alternating ones are enough) -- or pass a stream.  opus_b200_synchronize() /
 * opus_b200_timer_mark() order behind all of it.  max_frames = the number of passes the call launches: it must cover the
 * frames of the longest packet (1 for ordinary one-frame packets) AND, where packets can be lost or FEC is
 * requested, the concealment pieces of a call -- a lost packet is concealed in pieces no longer than the last
 * decoded frame, so frame_size / (shortest frame duration in use), plus one for an FEC frame.  A stream that would
 * need more passes gets OPUS_BAD_ARG in d_ret and its state is left untouched (the host forms count the passes
 * from the packets themselves).  max_packet_bytes = expected upper bound on len[i] (sizes the shared-memory
 * staging; a longer packet is still decoded, read in place).  pcm_stride >= frame_size * channels, else
 * OPUS_BAD_ARG. */
int opus_decode_group_device(OpusDecoderGroup *grp, const unsigned char *d_data,
                             const opus_int32 *d_offsets, const opus_int32 *d_len, opus_int16 *d_pcm,
                             opus_int32 pcm_stride, int frame_size, const int *d_decode_fec, int *d_ret,
                             int max_frames, int max_packet_bytes, void *cuda_stream);

/* ---- multistream (surround) decoder: libopus ABI (opus-fix/include/opus_multistream.h:257-448,504-650;
 *      NativeOpus.cs:131-155) + batch forms.  A multistream packet carries `streams` elementary Opus packets
 *      (all but the last self-delimited); each is decoded by its own elementary decoder on the GPU and the
 *      channels are then copied out according to `mapping`. ---- */
typedef struct OpusMSDecoder OpusMSDecoder;
typedef struct OpusMSDecoderGroup OpusMSDecoderGroup;
#define OPUS_MULTISTREAM_GET_DECODER_STATE_REQUEST 5122
/* replaces opus_multistream_decoder_get_size (src/opus_multistream_decoder.c:52-63): size of the HOST handle */
opus_int32 opus_multistream_decoder_get_size(int streams, int coupled_streams);
/* replaces opus_multistream_decoder_create (src/opus_multistream_decoder.c:107-142) */
OpusMSDecoder *opus_multistream_decoder_create(opus_int32 Fs, int channels, int streams, int coupled_streams,
                                               const unsigned char *mapping, int *error);
/* replaces opus_multistream_decode (src/opus_multistream_decoder.c:365-376) */
int opus_multistream_decode(OpusMSDecoder *st, const unsigned char *data, opus_int32 len, opus_int16 *pcm,
                            int frame_size, int decode_fec);
/* replaces opus_multistream_decode_float (src/opus_multistream_decoder.c:379-384) */
int opus_multistream_decode_float(OpusMSDecoder *st, const unsigned char *data, opus_int32 len, float *pcm,
                                  int frame_size, int decode_fec);
/* replaces opus_multistream_decoder_ctl (src/opus_multistream_decoder.c:409-531) */
int opus_multistream_decoder_ctl(OpusMSDecoder *st, int request, ...);
/* replaces opus_multistream_decoder_destroy (src/opus_multistream_decoder.c:534-537) */
void opus_multistream_decoder_destroy(OpusMSDecoder *st);
/* batch forms (new): n independent multistream decoders per call, per-decoder return codes */
int opus_multistream_decode_batch(OpusMSDecoder *const *st, const unsigned char *const *data, const opus_int32 *len,
                                  opus_int16 *const *pcm, int frame_size, const int *decode_fec, int n, int *ret);
OpusMSDecoderGroup *opus_multistream_decoder_group_create(OpusMSDecoder *const *st, int n, int *error);
void opus_multistream_decoder_group_destroy(OpusMSDecoderGroup *grp);
int opus_multistream_decode_group_flat(OpusMSDecoderGroup *grp, const unsigned char *data, opus_int32 data_bytes,
                                       const opus_int32 *offsets, const opus_int32 *len, opus_int16 *pcm,
                                       opus_int32 pcm_stride, int frame_size, const int *decode_fec, int *ret);
/* Pipelined form (see opus_decode_group_submit): pcm / ret are complete after the wait. */
int opus_multistream_decode_group_submit(OpusMSDecoderGroup *grp, const unsigned char *data, opus_int32 data_bytes,
                                         const opus_int32 *offsets, const opus_int32 *len, opus_int16 *pcm,
                                         opus_int32 pcm_stride, int frame_size, const int *decode_fec, int *ret);
int opus_multistream_decode_group_wait(OpusMSDecoderGroup *grp);

/* No counterpart in the reference (libopus keeps all scratch on the caller's stack).
 * Device-resident batch decode that allocates nothing: for callers that must not call cudaMalloc --
 * a real-time decode thread (allocation takes driver-wide locks, cudaFree synchronises the whole
 * GPU and stalls every other stream's audio) or a host with a fixed, pre-carved GPU memory budget.
 * It takes the decoders directly (no OpusDecoderGroup, whose creation
 * allocates); the per-stream scratch is a fixed array inside the library image, about 4.1 MB, held
 * by every GPU that loads the library until the process exits.  Limit: n <= 256
 * (OPUS_B200_STATIC_MAX_STREAMS); more is rejected with OPUS_BAD_ARG and a message on stderr.
 * All pointers except `st` are device pointers; the decoding is bit-identical to
 * opus_decode_group_device().  The array is shared by every call on that GPU and still holds the
 * previous call's records, which this call overwrites: calls that could run concurrently on one GPU
 * (different streams or threads) race and must be ordered by the caller.  The first call on a GPU
 * looks up the array's device address (cudaGetSymbolAddress); make it at start-up, before any
 * stream capture.  decode_fec is not offered here (treated as 0).  max_frames, max_packet_bytes, pcm_stride
 * and cuda_stream (NULL = the library's own stream) mean what they mean for opus_decode_group_device(). */
#define OPUS_B200_STATIC_MAX_STREAMS 256
int opus_decode_batch_device_static(OpusDecoder *const *st, int n, const unsigned char *d_data,
                                    const opus_int32 *d_offsets, const opus_int32 *d_len, opus_int16 *d_pcm,
                                    opus_int32 pcm_stride, int frame_size, int *d_ret, int max_frames,
                                    int max_packet_bytes, void *cuda_stream);

/* Batched OPUS_GET_FINAL_RANGE: out[i] = final range of the last packet decoded by st[i]. */
int opus_decoder_get_final_range_batch(OpusDecoder *const *st, int n, opus_uint32 *out);

/* ---- Encoder (CELT-only Opus packets) ---------------------------------------------------- */

/* Applications and special values (opus-fix/include/opus_defines.h:190-214) */
#define OPUS_AUTO -1000
#define OPUS_BITRATE_MAX -1
#define OPUS_APPLICATION_VOIP 2048
#define OPUS_APPLICATION_AUDIO 2049
#define OPUS_APPLICATION_RESTRICTED_LOWDELAY 2051
#define OPUS_SIGNAL_VOICE 3001
#define OPUS_SIGNAL_MUSIC 3002
#define OPUS_BANDWIDTH_NARROWBAND 1101
#define OPUS_BANDWIDTH_MEDIUMBAND 1102
#define OPUS_BANDWIDTH_WIDEBAND 1103
#define OPUS_BANDWIDTH_SUPERWIDEBAND 1104
#define OPUS_BANDWIDTH_FULLBAND 1105
#define OPUS_FRAMESIZE_ARG 5000
#define OPUS_FRAMESIZE_2_5_MS 5001
#define OPUS_FRAMESIZE_5_MS 5002
#define OPUS_FRAMESIZE_10_MS 5003
#define OPUS_FRAMESIZE_20_MS 5004
#define OPUS_FRAMESIZE_40_MS 5005
#define OPUS_FRAMESIZE_60_MS 5006
#define OPUS_FRAMESIZE_VARIABLE 5010
#define OPUS_GET_VOICE_RATIO_REQUEST 11019
/* ctl requests understood by opus_encoder_ctl (opus_defines.h:130-167; opus_encoder.c:2031-2504):
 * setters take an opus_int32, getters an opus_int32* (OPUS_GET_FINAL_RANGE an opus_uint32*). */
#define OPUS_SET_APPLICATION_REQUEST 4000
#define OPUS_GET_APPLICATION_REQUEST 4001
#define OPUS_SET_BITRATE_REQUEST 4002
#define OPUS_GET_BITRATE_REQUEST 4003
#define OPUS_SET_MAX_BANDWIDTH_REQUEST 4004
#define OPUS_GET_MAX_BANDWIDTH_REQUEST 4005
#define OPUS_SET_VBR_REQUEST 4006
#define OPUS_GET_VBR_REQUEST 4007
#define OPUS_SET_BANDWIDTH_REQUEST 4008
#define OPUS_SET_COMPLEXITY_REQUEST 4010
#define OPUS_GET_COMPLEXITY_REQUEST 4011
#define OPUS_SET_INBAND_FEC_REQUEST 4012
#define OPUS_GET_INBAND_FEC_REQUEST 4013
#define OPUS_SET_PACKET_LOSS_PERC_REQUEST 4014
#define OPUS_GET_PACKET_LOSS_PERC_REQUEST 4015
#define OPUS_SET_DTX_REQUEST 4016
#define OPUS_GET_DTX_REQUEST 4017
#define OPUS_SET_VBR_CONSTRAINT_REQUEST 4020
#define OPUS_GET_VBR_CONSTRAINT_REQUEST 4021
#define OPUS_SET_FORCE_CHANNELS_REQUEST 4022
#define OPUS_GET_FORCE_CHANNELS_REQUEST 4023
#define OPUS_SET_SIGNAL_REQUEST 4024
#define OPUS_GET_SIGNAL_REQUEST 4025
#define OPUS_GET_LOOKAHEAD_REQUEST 4027
#define OPUS_SET_LSB_DEPTH_REQUEST 4036
#define OPUS_GET_LSB_DEPTH_REQUEST 4037
#define OPUS_SET_EXPERT_FRAME_DURATION_REQUEST 4040
#define OPUS_GET_EXPERT_FRAME_DURATION_REQUEST 4041
#define OPUS_SET_PREDICTION_DISABLED_REQUEST 4042
#define OPUS_GET_PREDICTION_DISABLED_REQUEST 4043
#define OPUS_SET_FORCE_MODE_REQUEST 11002 /* src/opus_private.h:79; 1002 = CELT-only */
#define OPUS_SET_LFE_REQUEST 10024         /* celt/celt.h:113: this encoder codes an LFE channel */
#define OPUS_SET_ENERGY_MASK_REQUEST 10026 /* celt/celt.h:116: const opus_int16[21 * channels] surround masks (Q10),
                                             copied at the call (the reference keeps the pointer); NULL clears */

typedef struct OpusEncoder OpusEncoder;
typedef struct OpusEncoderGroup OpusEncoderGroup;

/* The libopus encoder ABI that Concentus binds (CSharp/Concentus/Native/NativeOpus.cs:105-150,180-225;
 * C prototypes opus-fix/include/opus.h:165-350).  Packets are byte-identical to the reference's
 * fixed-point encoder whenever its mode decision is CELT-only (always for
 * OPUS_APPLICATION_RESTRICTED_LOWDELAY or OPUS_SET_FORCE_MODE(1002); for OPUS_APPLICATION_AUDIO / _VOIP at
 * the rates and frame sizes where the reference picks CELT).  Frames the reference would code with SILK or
 * hybrid return OPUS_UNIMPLEMENTED.  40 / 60 ms frames are coded as the reference codes
 * them in CELT-only mode: two or three 20 ms frames in one packet (opus_encoder.c:1362-1438). */
/* replaces opus_encoder_get_size (src/opus_encoder.c:150-164): size of the host handle only */
int opus_encoder_get_size(int channels);
/* replaces opus_encoder_create (src/opus_encoder.c:482-510; NativeOpus.cs:105,180) */
OpusEncoder *opus_encoder_create(opus_int32 Fs, int channels, int application, int *error);
/* replaces opus_encode (src/opus_encoder.c:2007-2031; NativeOpus.cs:111,186): a batch of one */
opus_int32 opus_encode(OpusEncoder *st, const opus_int16 *pcm, int frame_size, unsigned char *data,
                       opus_int32 max_data_bytes);
/* replaces the FIXED_POINT opus_encode_float (src/opus_encoder.c:1968-2005; NativeOpus.cs:114,189) */
opus_int32 opus_encode_float(OpusEncoder *st, const float *pcm, int frame_size, unsigned char *data,
                             opus_int32 max_data_bytes);
/* replaces opus_encoder_ctl (src/opus_encoder.c:2031-2489; NativeOpus.cs:129-132,204-207) */
int opus_encoder_ctl(OpusEncoder *st, int request, ...);
/* replaces opus_encoder_destroy (src/opus_encoder.c:2491-2494; NativeOpus.cs:150,225) */
void opus_encoder_destroy(OpusEncoder *st);

/* Batch entry points (new).  n independent streams, one frame each per call; semantics are those
 * of n opus_encode() calls.  out_len[i] is stream i's packet length or negative error code. */
int opus_encode_batch(OpusEncoder *const *st, int n, const opus_int16 *const *pcm, int frame_size,
                      unsigned char *const *data, opus_int32 max_data_bytes, opus_int32 *out_len);
/* A fixed set of encoders (same device; same channel count for the flat/device calls). */
OpusEncoderGroup *opus_encoder_group_create(OpusEncoder *const *st, int n, int *error);
void opus_encoder_group_destroy(OpusEncoderGroup *grp);
/* Host buffers: pcm = n rows of frame_size*channels samples, data = n rows of max_data_bytes. */
int opus_encode_group_flat(OpusEncoderGroup *grp, const opus_int16 *pcm, int frame_size, unsigned char *data,
                           opus_int32 max_data_bytes, opus_int32 *out_len);
/* Device buffers, same layout; asynchronous on the library stream (opus_b200_synchronize waits). */
/* Pipelined host form of opus_encode_group_flat (see opus_decode_group_submit): data / out_len are complete
 * after opus_encode_group_wait(). */
int opus_encode_group_submit(OpusEncoderGroup *grp, const opus_int16 *pcm, int frame_size, unsigned char *data,
                             opus_int32 max_data_bytes, opus_int32 *out_len);
int opus_encode_group_wait(OpusEncoderGroup *grp);
int opus_encode_group_device(OpusEncoderGroup *grp, const opus_int16 *d_pcm, int frame_size,
                             unsigned char *d_data, opus_int32 max_data_bytes, opus_int32 *d_len);
/* Batched OPUS_GET_FINAL_RANGE for encoders. */
int opus_encoder_get_final_range_batch(OpusEncoder *const *st, int n, opus_uint32 *out);

/* ---- multistream encoder (opus-fix/include/opus_multistream.h:257-448, src/opus_multistream_encoder.c) ----
 * The streams of one multistream frame are encoded one after the other (the byte budget of a stream depends
 * on its predecessors), each by an ordinary elementary encoder on the GPU; channel copy-in, the rate split
 * and the packet assembly are host code.  Built: explicit layouts (opus_multistream_encoder_create) and the
 * surround families 0, 1 and 255.  Family 1 with 3..8 channels is the surround encoder proper
 * (opus_multistream_encoder.c:224-360, 525, 816-887): a masking analysis of all input channels runs on the device
 * before the streams are coded and hands every stream its masks (the reference's OPUS_SET_ENERGY_MASK), the
 * sixth channel of 5.1 / 6.1 / 7.1 is an LFE stream (OPUS_SET_LFE).  A frame the elementary encoder refuses
 * (see opus_encode above: a stream for which the reference would pick SILK or hybrid) returns OPUS_UNIMPLEMENTED. */
typedef struct OpusMSEncoder OpusMSEncoder;
#define OPUS_MULTISTREAM_GET_ENCODER_STATE_REQUEST 5120
opus_int32 opus_multistream_encoder_get_size(int streams, int coupled_streams);
opus_int32 opus_multistream_surround_encoder_get_size(int channels, int mapping_family);
OpusMSEncoder *opus_multistream_encoder_create(opus_int32 Fs, int channels, int streams, int coupled_streams,
                                               const unsigned char *mapping, int application, int *error);
OpusMSEncoder *opus_multistream_surround_encoder_create(opus_int32 Fs, int channels, int mapping_family,
                                                        int *streams, int *coupled_streams,
                                                        unsigned char *mapping, int application, int *error);
int opus_multistream_encode(OpusMSEncoder *st, const opus_int16 *pcm, int frame_size, unsigned char *data,
                            opus_int32 max_data_bytes);
int opus_multistream_encode_float(OpusMSEncoder *st, const float *pcm, int frame_size, unsigned char *data,
                                  opus_int32 max_data_bytes);
int opus_multistream_encoder_ctl(OpusMSEncoder *st, int request, ...);
void opus_multistream_encoder_destroy(OpusMSEncoder *st);
/* Device timing of the encode kernel with CUDA events on the launching stream (bench.py). */
int opus_b200_encode_profile_begin(void);
int opus_b200_encode_profile_end(double *ms, long long *count);

/* Blocks until all work queued by this library on the current device has finished. */
int opus_b200_synchronize(void);
/* Number of kernel launches issued by this library so far (process-wide counter). */
long long opus_b200_launch_count(void);

/* Per-kernel device timing with CUDA events on the launching stream (used by bench.py for the
 * roofline line).  Between begin and end every kernel launch is bracketed by an event pair and calls are queued
 * on one stream (no run-ahead of the parse side, see opus_decode_group_device); end synchronises and returns
 * summed milliseconds and launch counts for [0] plan + routing + SILK kernel, [1] parse kernel, [2] rebuild
 * kernel, [3] synthesis kernel. */
int opus_b200_profile_begin(void);
int opus_b200_profile_end(double ms[4], long long counts[4]);
/* Timing marks for work queued on the library's own stream (cuda_stream = NULL), where the caller has no stream
 * to record events on: opus_b200_timer_mark(idx), idx 0..15, records a timing event on the current device's main
 * stream behind everything queued so far; opus_b200_timer_elapsed waits for both marks and returns the
 * milliseconds from mark `first` to mark `second`.  (No counterpart in the reference.) */
int opus_b200_timer_mark(int idx);
int opus_b200_timer_elapsed(int first, int second, double *ms);
/* Diagnostic (tools/pipe_trace.py): between trace_begin and trace_end every decode launch is bracketed by an event
 * pair as with the profile calls, but the calls keep their normal schedule; trace_end returns the number of records
 * (at most cap) of (kind as in opus_b200_profile_end, start ms, end ms) relative to the first recorded launch. */
int opus_b200_trace_begin(void);
int opus_b200_trace_end(int *kinds, double *t0, double *t1, int cap);

/* Test support (tests/test_kat.py): the kernels' PVQ index -> pulse vector routine (one lane per vector) and its inverse
 * (one warp per vector) on arrays of (n, k, index); all pointers are device pointers, y holds count x 176 values.
 * Follows the reference's celt/tests/test_unit_cwrs32.c. */
int opus_b200_selftest_cwrs(const int *d_n, const int *d_k, const unsigned *d_idx, int count, short *d_y, int *d_yy,
                            unsigned *d_back);
/* Test support: the device form of isqrt32 (celt/mathops.c:42-63) on an array (device pointers). */
int opus_b200_selftest_isqrt(const unsigned *d_in, int count, unsigned *d_out);

#ifdef __cplusplus
}
#endif
#endif /* OPUS_B200_H */
