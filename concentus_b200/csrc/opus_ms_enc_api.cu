// opus_ms_enc_api.cu -- libopus multistream ENCODER ABI (opus-fix/include/opus_multistream.h:257-448,
// src/opus_multistream_encoder.c) on top of the elementary GPU encoders.
//
// A multistream packet is the concatenation of its streams' packets (all but the last self-delimited),
// and the byte budget of stream s depends on what streams 0..s-1 produced (opus_multistream_encoder.c:
// 820-905), so the streams of one multistream frame are encoded one after the other; each is an ordinary
// opus_encode() on its own elementary encoder.  Host code here: channel copy-in, rate split, packet
// assembly.  Built: opus_multistream_encoder_create() layouts and every opus_multistream_surround_encoder_create()
// family (0, 1, 255).  Family 1 with three or more channels is the surround case (:224-360, 525, 816-887): a
// masking analysis of all input channels runs first (surround_analysis_kernel, cb_surround_kernels.cu) and writes
// each channel's masks into the CELT state of its stream on the device; the sixth channel of 5.1 and up is an
// LFE stream; coupled streams are forced to stereo CELT and the bandwidth follows the total rate.
// What the elementary encoder does not build (SILK / hybrid mode decisions: a mono stream below ~23 kbit/s)
// comes back from opus_encode() as OPUS_UNIMPLEMENTED and is passed on.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#include <vector>

#include <cuda_runtime.h>

#include "../../include/opus_b200.h"
#include "cb_repack_internal.h"
#include "cb_surround_kernels.h"

#include "cb_host.h"
extern "C" int opusb_internal_encoder_mask_slots(OpusEncoder* st, int side, opus_int16** mask, opus_int32** has_mask);


struct OpusMSEncoder {
  int magic;
  int nb_channels, nb_streams, nb_coupled;
  unsigned char mapping[256];
  opus_int32 Fs;
  int application;
  opus_int32 bitrate_bps;
  int variable_duration;
  OpusEncoder** enc;             // nb_streams elementary encoders: the coupled (stereo) ones first
  int surround, lfe_stream;      // family 1 with >= 3 channels; index of the LFE stream or -1
  // surround analysis (device): SurroundState, work buffers, the frame's PCM, where each channel's masks go
  void* d_surround_state;
  void* d_surround_scratch;
  opus_int16* d_pcm;
  size_t d_pcm_cap;
  SurroundDst dst;
  int device;                    // GPU the encoder (and its elementary encoders) live on
};

namespace {

const int kMSEncMagic = 0x4f4d4532;   // "OME2"
const int MS_FRAME_TMP = 3 * 1275 + 7;

bool valid(const OpusMSEncoder* st) { return st && st->magic == kMSEncMagic; }

// opus_multistream.c:53-91
int left_channel(const OpusMSEncoder* st, int s) {
  for (int i = 0; i < st->nb_channels; i++)
    if (st->mapping[i] == s * 2) return i;
  return -1;
}
int right_channel(const OpusMSEncoder* st, int s) {
  for (int i = 0; i < st->nb_channels; i++)
    if (st->mapping[i] == s * 2 + 1) return i;
  return -1;
}
int mono_channel(const OpusMSEncoder* st, int s) {
  for (int i = 0; i < st->nb_channels; i++)
    if (st->mapping[i] == s + st->nb_coupled) return i;
  return -1;
}

// frame_size_select (opus_encoder.c:811-835)
int select_frame_size(int frame_size, int variable_duration, opus_int32 Fs) {
  int new_size;
  if (frame_size < Fs / 400) return -1;
  if (variable_duration == OPUS_FRAMESIZE_ARG) new_size = frame_size;
  else if (variable_duration == OPUS_FRAMESIZE_VARIABLE) new_size = Fs / 50;
  else if (variable_duration >= OPUS_FRAMESIZE_2_5_MS && variable_duration <= OPUS_FRAMESIZE_60_MS) {
    const int a = 3 * Fs / 50, b = (Fs / 400) << (variable_duration - OPUS_FRAMESIZE_2_5_MS);
    new_size = a < b ? a : b;
  } else return -1;
  if (new_size > frame_size) return -1;
  if (400 * new_size != Fs && 200 * new_size != Fs && 100 * new_size != Fs && 50 * new_size != Fs &&
      25 * new_size != Fs && 50 * new_size != 3 * Fs)
    return -1;
  return new_size;
}

// surround_rate_allocation (opus_multistream_encoder.c:621-695)
opus_int32 split_rate(const OpusMSEncoder* st, opus_int32* rate, int frame_size) {
  const opus_int32 Fs = st->Fs;
  int stream_offset;
  if (st->bitrate_bps > st->nb_channels * 40000) stream_offset = 20000;
  else stream_offset = st->bitrate_bps / st->nb_channels / 2;
  stream_offset += 60 * (Fs / frame_size - 50);
  const int lfe_offset = 3500 + 60 * (Fs / frame_size - 50);
  const int coupled_ratio = 512;                       // Q8: a coupled stream gets twice the mono rate
  const int lfe_ratio = 32;                            // Q8: the LFE stream gets 1/8 of it
  opus_int32 channel_rate;
  if (st->bitrate_bps == OPUS_AUTO) {
    channel_rate = Fs + 60 * Fs / frame_size;
  } else if (st->bitrate_bps == OPUS_BITRATE_MAX) {
    channel_rate = 300000;
  } else {
    const int nb_lfe = st->lfe_stream != -1;
    const int nb_coupled = st->nb_coupled, nb_uncoupled = st->nb_streams - nb_coupled - nb_lfe;
    const int total = (nb_uncoupled << 8) + coupled_ratio * nb_coupled + nb_lfe * lfe_ratio;
    channel_rate = 256 * (st->bitrate_bps - lfe_offset * nb_lfe - stream_offset * (nb_coupled + nb_uncoupled)) / total;
  }
  opus_int32 sum = 0;
  for (int i = 0; i < st->nb_streams; i++) {
    if (i < st->nb_coupled) rate[i] = stream_offset + (channel_rate * coupled_ratio >> 8);
    else if (i != st->lfe_stream) rate[i] = stream_offset + channel_rate;
    else rate[i] = lfe_offset + (channel_rate * lfe_ratio >> 8);
    if (rate[i] < 500) rate[i] = 500;
    sum += rate[i];
  }
  return sum;
}

void free_encoder(OpusMSEncoder* st) {
  if (!st) return;
  if (st->enc) {
    for (int i = 0; i < st->nb_streams; i++)
      if (st->enc[i]) opus_encoder_destroy(st->enc[i]);
    free(st->enc);
  }
  cudaFree(st->d_surround_state);
  cudaFree(st->d_surround_scratch);
  cudaFree(st->d_pcm);
  st->magic = 0;
  free(st);
}

OpusMSEncoder* create_impl(opus_int32 Fs, int channels, int streams, int coupled_streams, const unsigned char* mapping,
                           int application, int* error, int surround = 0, int lfe_stream = -1) {
  int err = OPUS_OK;
  OpusMSEncoder* st = nullptr;
  do {
    // opus_multistream_encoder_init_impl (:412-473)
    if (channels > 255 || channels < 1 || coupled_streams > streams || streams < 1 || coupled_streams < 0 ||
        streams > 255 - coupled_streams || !mapping) { err = OPUS_BAD_ARG; break; }
    st = (OpusMSEncoder*)calloc(1, sizeof(OpusMSEncoder));
    if (!st) { err = OPUS_ALLOC_FAIL; break; }
    st->magic = kMSEncMagic;
    st->nb_channels = channels; st->nb_streams = streams; st->nb_coupled = coupled_streams;
    st->Fs = Fs; st->application = application;
    st->bitrate_bps = OPUS_AUTO;
    st->variable_duration = OPUS_FRAMESIZE_ARG;
    st->surround = surround;
    st->lfe_stream = surround ? lfe_stream : -1;
    const int max_channel = streams + coupled_streams;           // validate_layout (opus_multistream.c:39-51)
    bool ok = max_channel <= 255;
    for (int i = 0; i < channels && ok; i++) {
      st->mapping[i] = mapping[i];
      if (mapping[i] >= max_channel && mapping[i] != 255) ok = false;
    }
    for (int s = 0; s < streams && ok; s++) {                      // validate_encoder_layout (:127-144)
      if (s < coupled_streams) ok = left_channel(st, s) != -1 && right_channel(st, s) != -1;
      else ok = mono_channel(st, s) != -1;
    }
    if (!ok) { err = OPUS_BAD_ARG; break; }
    st->enc = (OpusEncoder**)calloc((size_t)streams, sizeof(OpusEncoder*));
    if (cudaGetDevice(&st->device) != cudaSuccess) st->device = -1;
    if (!st->enc) { err = OPUS_ALLOC_FAIL; break; }
    for (int i = 0; i < streams; i++) {
      st->enc[i] = opus_encoder_create(Fs, i < coupled_streams ? 2 : 1, application, &err);
      if (!st->enc[i]) break;
      if (i == st->lfe_stream) err = opus_encoder_ctl(st->enc[i], OPUS_SET_LFE_REQUEST, 1);
      if (err != OPUS_OK) break;
    }
    if (err != OPUS_OK || !surround) break;
    // surround analysis state (zero: :466-470) and buffers; channel c's masks go to the stream that codes it
    if (cudaMalloc(&st->d_surround_state, cbk_sizeof_surround_state()) != cudaSuccess ||
        cudaMalloc(&st->d_surround_scratch, cbk_sizeof_surround_scratch()) != cudaSuccess ||
        cudaMemset(st->d_surround_state, 0, cbk_sizeof_surround_state()) != cudaSuccess) { err = OPUS_ALLOC_FAIL; break; }
    for (int s = 0; s < streams && err == OPUS_OK; s++) {
      if (s < coupled_streams) {
        const int l = left_channel(st, s), r = right_channel(st, s);
        err = opusb_internal_encoder_mask_slots(st->enc[s], 0, &st->dst.mask[l], &st->dst.has_mask[l]);
        if (err == OPUS_OK) err = opusb_internal_encoder_mask_slots(st->enc[s], 1, &st->dst.mask[r], &st->dst.has_mask[r]);
      } else {
        const int c = mono_channel(st, s);
        err = opusb_internal_encoder_mask_slots(st->enc[s], 0, &st->dst.mask[c], &st->dst.has_mask[c]);
      }
    }
  } while (0);
  if (err != OPUS_OK) {
    free_encoder(st);
    st = nullptr;
  }
  if (error) *error = err;
  return st;
}

// opus_multistream_encode_native (:699-913) for an encoder without surround analysis; pcm16 = the input
// already as 16-bit samples, interleaved by nb_channels
int encode_impl(OpusMSEncoder* st, const opus_int16* pcm16, int analysis_frame_size, unsigned char* data,
                opus_int32 max_data_bytes) {
  const opus_int32 Fs = st->Fs;
  opus_int32 vbr = 0;
  int rc = opus_encoder_ctl(st->enc[0], OPUS_GET_VBR_REQUEST, &vbr);
  if (rc != OPUS_OK) return rc;
  const int frame_size = select_frame_size(analysis_frame_size, st->variable_duration, Fs);
  if (frame_size < 0 || 400 * frame_size < Fs) return OPUS_BAD_ARG;
  const opus_int32 smallest_packet = st->nb_streams * 2 - 1;
  if (max_data_bytes < smallest_packet) return OPUS_BUFFER_TOO_SMALL;
  opus_int32 bitrates[256];
  const opus_int32 rate_sum = split_rate(st, bitrates, frame_size);
  if (!vbr) {
    if (st->bitrate_bps == OPUS_AUTO) {
      const opus_int32 cap = 3 * rate_sum / (3 * 8 * Fs / frame_size);
      if (cap < max_data_bytes) max_data_bytes = cap;
    } else if (st->bitrate_bps != OPUS_BITRATE_MAX) {
      opus_int32 cap = 3 * st->bitrate_bps / (3 * 8 * Fs / frame_size);
      if (cap < smallest_packet) cap = smallest_packet;
      if (cap < max_data_bytes) max_data_bytes = cap;
    }
  }
  if (st->surround) {
    // masks of this frame, written into the streams' states on the device (:787-790, 862-887)
    int err = 0;
    OpusbDevice* dv = opusb_device(st->device, &err);        // the GPU this encoder was created on
    if (!dv) return err;
    cudaStream_t stream = dv->stream;
    const size_t bytes = sizeof(opus_int16) * (size_t)frame_size * st->nb_channels;
    if (bytes > st->d_pcm_cap) {
      cudaFree(st->d_pcm);
      st->d_pcm = nullptr; st->d_pcm_cap = 0;
      if (cudaMalloc((void**)&st->d_pcm, bytes) != cudaSuccess) return OPUS_ALLOC_FAIL;
      st->d_pcm_cap = bytes;
    }
    if (cudaMemcpyAsync(st->d_pcm, pcm16, bytes, cudaMemcpyHostToDevice, stream) != cudaSuccess) return OPUS_INTERNAL_ERROR;
    if (cbk_launch_surround(st->d_pcm, st->nb_channels, frame_size, Fs, st->d_surround_state, st->d_surround_scratch,
                            &st->dst, stream) != 0)
      return OPUS_INTERNAL_ERROR;
    opusb_internal_count_launches(1);
    if (cudaStreamSynchronize(stream) != cudaSuccess) return OPUS_INTERNAL_ERROR;   // pcm16 may be pageable
  }
  for (int s = 0; s < st->nb_streams; s++) {
    rc = opus_encoder_ctl(st->enc[s], OPUS_SET_BITRATE_REQUEST, bitrates[s]);
    if (rc != OPUS_OK) return rc;
    if (st->surround) {
      opus_int32 equiv_rate = st->bitrate_bps;
      if (frame_size * 50 < Fs) equiv_rate -= 60 * (Fs / frame_size - 50) * st->nb_channels;
      const int bw = equiv_rate > 10000 * st->nb_channels ? OPUS_BANDWIDTH_FULLBAND
                     : equiv_rate > 7000 * st->nb_channels ? OPUS_BANDWIDTH_SUPERWIDEBAND
                     : equiv_rate > 5000 * st->nb_channels ? OPUS_BANDWIDTH_WIDEBAND : OPUS_BANDWIDTH_NARROWBAND;
      rc = opus_encoder_ctl(st->enc[s], OPUS_SET_BANDWIDTH_REQUEST, bw);
      if (rc == OPUS_OK && s < st->nb_coupled) {
        // the spatial image needs stereo CELT on the coupled streams (:831-836)
        rc = opus_encoder_ctl(st->enc[s], OPUS_SET_FORCE_MODE_REQUEST, 1002);   // MODE_CELT_ONLY
        if (rc == OPUS_OK) rc = opus_encoder_ctl(st->enc[s], OPUS_SET_FORCE_CHANNELS_REQUEST, 2);
      }
      if (rc != OPUS_OK) return rc;
    }
  }
  std::vector<opus_int16> buf((size_t)2 * frame_size);
  unsigned char tmp_data[MS_FRAME_TMP];
  opus_int32 tot_size = 0;
  const int CH = st->nb_channels;
  for (int s = 0; s < st->nb_streams; s++) {
    if (s < st->nb_coupled) {
      const int l = left_channel(st, s), r = right_channel(st, s);
      for (int i = 0; i < frame_size; i++) {
        buf[2 * i] = pcm16[(size_t)i * CH + l];
        buf[2 * i + 1] = pcm16[(size_t)i * CH + r];
      }
    } else {
      const int c = mono_channel(st, s);
      for (int i = 0; i < frame_size; i++) buf[i] = pcm16[(size_t)i * CH + c];
    }
    // bytes left; one byte is kept for the last stream and two for each of the others (:884-892)
    opus_int32 curr_max = max_data_bytes - tot_size;
    const int reserve = 2 * (st->nb_streams - s - 1) - 1;
    curr_max -= reserve > 0 ? reserve : 0;
    if (curr_max > MS_FRAME_TMP) curr_max = MS_FRAME_TMP;
    if (s != st->nb_streams - 1) curr_max -= curr_max > 253 ? 2 : 1;          // the self-delimiting length
    if (!vbr && s == st->nb_streams - 1) {
      rc = opus_encoder_ctl(st->enc[s], OPUS_SET_BITRATE_REQUEST, (opus_int32)(curr_max * (8 * Fs / frame_size)));
      if (rc != OPUS_OK) return rc;
    }
    opus_int32 len = opus_encode(st->enc[s], buf.data(), frame_size, tmp_data, curr_max);
    if (len < 0) return len;
    OpusRepacketizer rp;
    rp.nb_frames = 0;
    opusb_rp_cat(&rp, tmp_data, len, 0);
    len = opusb_rp_emit(&rp, 0, rp.nb_frames, data, max_data_bytes - tot_size, s != st->nb_streams - 1,
                        !vbr && s == st->nb_streams - 1);
    if (len < 0) return len;
    data += len;
    tot_size += len;
  }
  return tot_size;
}

}  // namespace

extern "C" {

opus_int32 opus_multistream_encoder_get_size(int streams, int coupled_streams) {
  if (streams < 1 || coupled_streams > streams || coupled_streams < 0) return 0;
  return (opus_int32)sizeof(OpusMSEncoder);            // the codec states live in device memory
}

opus_int32 opus_multistream_surround_encoder_get_size(int channels, int mapping_family) {
  if (mapping_family == 0) return (channels == 1 || channels == 2) ? (opus_int32)sizeof(OpusMSEncoder) : 0;
  if (mapping_family == 1 && channels <= 8 && channels >= 1) return (opus_int32)sizeof(OpusMSEncoder);
  if (mapping_family == 255 && channels >= 1 && channels <= 255) return (opus_int32)sizeof(OpusMSEncoder);
  return 0;
}

OpusMSEncoder* opus_multistream_encoder_create(opus_int32 Fs, int channels, int streams, int coupled_streams,
                                               const unsigned char* mapping, int application, int* error) {
  if (channels > 255 || channels < 1 || coupled_streams > streams || streams < 1 || coupled_streams < 0 ||
      streams > 255 - coupled_streams) {
    if (error) *error = OPUS_BAD_ARG;
    return nullptr;
  }
  return create_impl(Fs, channels, streams, coupled_streams, mapping, application, error);
}

OpusMSEncoder* opus_multistream_surround_encoder_create(opus_int32 Fs, int channels, int mapping_family, int* streams,
                                                        int* coupled_streams, unsigned char* mapping, int application,
                                                        int* error) {
  if (channels > 255 || channels < 1 || !streams || !coupled_streams || !mapping) {
    if (error) *error = OPUS_BAD_ARG;
    return nullptr;
  }
  // opus_multistream_surround_encoder_init (:488-537)
  if (mapping_family == 0 && (channels == 1 || channels == 2)) {
    *streams = 1;
    *coupled_streams = channels - 1;
    mapping[0] = 0;
    if (channels == 2) mapping[1] = 1;
  } else if (mapping_family == 1 && channels <= 8) {
    // vorbis_mappings (:49-62)
    static const struct { int nb_streams, nb_coupled; unsigned char mapping[8]; } vorbis[8] = {
        {1, 0, {0}}, {1, 1, {0, 1}}, {2, 1, {0, 2, 1}}, {2, 2, {0, 1, 2, 3}}, {3, 2, {0, 4, 1, 2, 3}},
        {4, 2, {0, 4, 1, 2, 3, 5}}, {4, 3, {0, 4, 1, 2, 3, 5, 6}}, {5, 3, {0, 6, 1, 2, 3, 4, 5, 7}}};
    *streams = vorbis[channels - 1].nb_streams;
    *coupled_streams = vorbis[channels - 1].nb_coupled;
    for (int i = 0; i < channels; i++) mapping[i] = vorbis[channels - 1].mapping[i];
    return create_impl(Fs, channels, *streams, *coupled_streams, mapping, application, error, channels > 2,
                       channels >= 6 ? *streams - 1 : -1);
  } else if (mapping_family == 255) {
    *streams = channels;
    *coupled_streams = 0;
    for (int i = 0; i < channels; i++) mapping[i] = (unsigned char)i;
  } else {
    if (error) *error = OPUS_UNIMPLEMENTED;
    return nullptr;
  }
  return create_impl(Fs, channels, *streams, *coupled_streams, mapping, application, error);
}

void opus_multistream_encoder_destroy(OpusMSEncoder* st) {
  if (valid(st)) free_encoder(st);
}

int opus_multistream_encode(OpusMSEncoder* st, const opus_int16* pcm, int frame_size, unsigned char* data,
                            opus_int32 max_data_bytes) {
  if (!valid(st) || !pcm || !data) return OPUS_BAD_ARG;
  return encode_impl(st, pcm, frame_size, data, max_data_bytes);
}

// FIXED_POINT build: FLOAT2INT16 on the way in (opus_multistream_encoder.c:919-936), then as above
int opus_multistream_encode_float(OpusMSEncoder* st, const float* pcm, int frame_size, unsigned char* data,
                                  opus_int32 max_data_bytes) {
  if (!valid(st) || !pcm || !data || frame_size <= 0) return OPUS_BAD_ARG;
  std::vector<opus_int16> in((size_t)frame_size * st->nb_channels);
  for (size_t i = 0; i < in.size(); i++) {
    float x = pcm[i] * 32768.f;
    x = x > -32768.f ? x : -32768.f;
    x = x < 32767.f ? x : 32767.f;
    in[i] = (opus_int16)lrintf(x);
  }
  return encode_impl(st, in.data(), frame_size, data, max_data_bytes);
}

int opus_multistream_encoder_ctl(OpusMSEncoder* st, int request, ...) {
  if (!valid(st)) return OPUS_BAD_ARG;
  va_list ap;
  va_start(ap, request);
  int ret = OPUS_OK;
  switch (request) {
    case OPUS_SET_BITRATE_REQUEST: {
      const opus_int32 value = va_arg(ap, opus_int32);
      if (value < 0 && value != OPUS_AUTO && value != OPUS_BITRATE_MAX) { ret = OPUS_BAD_ARG; break; }
      st->bitrate_bps = value;
      break;
    }
    case OPUS_GET_BITRATE_REQUEST: {
      opus_int32* value = va_arg(ap, opus_int32*);
      if (!value) { ret = OPUS_BAD_ARG; break; }
      *value = 0;
      for (int s = 0; s < st->nb_streams; s++) {
        opus_int32 rate = 0;
        opus_encoder_ctl(st->enc[s], request, &rate);
        *value += rate;
      }
      break;
    }
    case OPUS_GET_LSB_DEPTH_REQUEST:
    case OPUS_GET_VBR_REQUEST:
    case OPUS_GET_APPLICATION_REQUEST:
    case OPUS_GET_BANDWIDTH_REQUEST:
    case OPUS_GET_COMPLEXITY_REQUEST:
    case OPUS_GET_PACKET_LOSS_PERC_REQUEST:
    case OPUS_GET_DTX_REQUEST:
    case OPUS_GET_VOICE_RATIO_REQUEST:
    case OPUS_GET_VBR_CONSTRAINT_REQUEST:
    case OPUS_GET_SIGNAL_REQUEST:
    case OPUS_GET_LOOKAHEAD_REQUEST:
    case OPUS_GET_SAMPLE_RATE_REQUEST:
    case OPUS_GET_INBAND_FEC_REQUEST:
    case OPUS_GET_FORCE_CHANNELS_REQUEST:
    case OPUS_GET_PREDICTION_DISABLED_REQUEST: {
      opus_int32* value = va_arg(ap, opus_int32*);     // the first stream answers for all
      ret = opus_encoder_ctl(st->enc[0], request, value);
      break;
    }
    case OPUS_GET_FINAL_RANGE_REQUEST: {
      opus_uint32* value = va_arg(ap, opus_uint32*);
      if (!value) { ret = OPUS_BAD_ARG; break; }
      *value = 0;
      for (int s = 0; s < st->nb_streams; s++) {
        opus_uint32 tmp = 0;
        ret = opus_encoder_ctl(st->enc[s], request, &tmp);
        if (ret != OPUS_OK) break;
        *value ^= tmp;
      }
      break;
    }
    case OPUS_SET_LSB_DEPTH_REQUEST:
    case OPUS_SET_COMPLEXITY_REQUEST:
    case OPUS_SET_VBR_REQUEST:
    case OPUS_SET_VBR_CONSTRAINT_REQUEST:
    case OPUS_SET_MAX_BANDWIDTH_REQUEST:
    case OPUS_SET_BANDWIDTH_REQUEST:
    case OPUS_SET_SIGNAL_REQUEST:
    case OPUS_SET_APPLICATION_REQUEST:
    case OPUS_SET_INBAND_FEC_REQUEST:
    case OPUS_SET_PACKET_LOSS_PERC_REQUEST:
    case OPUS_SET_DTX_REQUEST:
    case OPUS_SET_FORCE_MODE_REQUEST:
    case OPUS_SET_FORCE_CHANNELS_REQUEST:
    case OPUS_SET_PREDICTION_DISABLED_REQUEST: {
      const opus_int32 value = va_arg(ap, opus_int32);
      for (int s = 0; s < st->nb_streams; s++) {
        ret = opus_encoder_ctl(st->enc[s], request, value);
        if (ret != OPUS_OK) break;
      }
      break;
    }
    case OPUS_MULTISTREAM_GET_ENCODER_STATE_REQUEST: {
      const opus_int32 stream_id = va_arg(ap, opus_int32);
      OpusEncoder** value = va_arg(ap, OpusEncoder**);
      if (!value || stream_id < 0 || stream_id >= st->nb_streams) { ret = OPUS_BAD_ARG; break; }
      *value = st->enc[stream_id];
      break;
    }
    case OPUS_SET_EXPERT_FRAME_DURATION_REQUEST: {
      st->variable_duration = va_arg(ap, opus_int32);
      break;
    }
    case OPUS_GET_EXPERT_FRAME_DURATION_REQUEST: {
      opus_int32* value = va_arg(ap, opus_int32*);
      if (!value) { ret = OPUS_BAD_ARG; break; }
      *value = st->variable_duration;
      break;
    }
    case OPUS_RESET_STATE: {
      if (st->surround && cudaMemset(st->d_surround_state, 0, cbk_sizeof_surround_state()) != cudaSuccess) {
        ret = OPUS_INTERNAL_ERROR;
        break;
      }
      for (int s = 0; s < st->nb_streams; s++) {
        ret = opus_encoder_ctl(st->enc[s], OPUS_RESET_STATE);
        if (ret != OPUS_OK) break;
      }
      break;
    }
    default:
      ret = OPUS_UNIMPLEMENTED;
      break;
  }
  va_end(ap);
  return ret;
}

}  // extern "C"
