// cb_enc_ctl.h -- host-side opus_encoder_ctl on an OpusEncState image
// (opus-fix/src/opus_encoder.c:2031-2504; celt side celt/celt_encoder.c:2345-2470).
// Setters take the value, getters write *out.  The product API applies it to a host copy of the
// stream's scalars and writes them back to the device; the CPU-side tests apply it directly.
#pragma once
#include "cb_enc_state.h"

#define CB_REQ_SET_APPLICATION 4000
#define CB_REQ_GET_APPLICATION 4001
#define CB_REQ_SET_BITRATE 4002
#define CB_REQ_GET_BITRATE 4003
#define CB_REQ_SET_MAX_BANDWIDTH 4004
#define CB_REQ_GET_MAX_BANDWIDTH 4005
#define CB_REQ_SET_VBR 4006
#define CB_REQ_GET_VBR 4007
#define CB_REQ_SET_BANDWIDTH 4008
#define CB_REQ_GET_BANDWIDTH 4009
#define CB_REQ_SET_COMPLEXITY 4010
#define CB_REQ_GET_COMPLEXITY 4011
#define CB_REQ_SET_INBAND_FEC 4012
#define CB_REQ_GET_INBAND_FEC 4013
#define CB_REQ_SET_PACKET_LOSS_PERC 4014
#define CB_REQ_GET_PACKET_LOSS_PERC 4015
#define CB_REQ_SET_DTX 4016
#define CB_REQ_GET_DTX 4017
#define CB_REQ_SET_VBR_CONSTRAINT 4020
#define CB_REQ_GET_VBR_CONSTRAINT 4021
#define CB_REQ_SET_FORCE_CHANNELS 4022
#define CB_REQ_GET_FORCE_CHANNELS 4023
#define CB_REQ_SET_SIGNAL 4024
#define CB_REQ_GET_SIGNAL 4025
#define CB_REQ_GET_LOOKAHEAD 4027
#define CB_REQ_RESET_STATE 4028
#define CB_REQ_GET_SAMPLE_RATE 4029
#define CB_REQ_GET_FINAL_RANGE 4031
#define CB_REQ_SET_LSB_DEPTH 4036
#define CB_REQ_GET_LSB_DEPTH 4037
#define CB_REQ_SET_EXPERT_FRAME_DURATION 4040
#define CB_REQ_GET_EXPERT_FRAME_DURATION 4041
#define CB_REQ_SET_PREDICTION_DISABLED 4042
#define CB_REQ_GET_PREDICTION_DISABLED 4043
#define CB_REQ_SET_FORCE_MODE 11002
#define CB_REQ_SET_LFE 10024          // celt.h:113 (used by the multistream surround encoder)
#define CB_REQ_SET_ENERGY_MASK 10026  // celt.h:116; takes a pointer, handled by the caller

// True for the requests that only read the scalars.
static inline bool opus_enc_ctl_is_get(int request) {
  switch (request) {
    case CB_REQ_GET_APPLICATION: case CB_REQ_GET_BITRATE: case CB_REQ_GET_MAX_BANDWIDTH: case CB_REQ_GET_VBR:
    case CB_REQ_GET_BANDWIDTH: case CB_REQ_GET_COMPLEXITY: case CB_REQ_GET_INBAND_FEC:
    case CB_REQ_GET_PACKET_LOSS_PERC: case CB_REQ_GET_DTX: case CB_REQ_GET_VBR_CONSTRAINT:
    case CB_REQ_GET_FORCE_CHANNELS: case CB_REQ_GET_SIGNAL: case CB_REQ_GET_LOOKAHEAD: case CB_REQ_GET_SAMPLE_RATE:
    case CB_REQ_GET_FINAL_RANGE: case CB_REQ_GET_LSB_DEPTH: case CB_REQ_GET_EXPERT_FRAME_DURATION:
    case CB_REQ_GET_PREDICTION_DISABLED:
      return true;
    default:
      return false;
  }
}

// Scalars-only part: `o` and `cs` are the Opus and CELT scalar blocks.  RESET_STATE is handled by
// the caller (it clears the arrays too).  Returns OPUS_OK / OPUS_BAD_ARG / OPUS_UNIMPLEMENTED.
static inline int opus_enc_ctl_scalars(OpusEncScalars* o, CeltEncScalars* cs, int request, i32 value, i32* out) {
  switch (request) {
    case CB_REQ_SET_APPLICATION:
      if ((value != CB_APP_VOIP && value != CB_APP_AUDIO && value != CB_APP_LOWDELAY) ||
          (!o->first && o->application != value))
        return CB_BAD_ARG;
      o->application = value;
      return CB_OK;
    case CB_REQ_GET_APPLICATION: *out = o->application; return CB_OK;
    case CB_REQ_SET_BITRATE:
      if (value != CB_OPUS_AUTO && value != CB_OPUS_BITRATE_MAX) {
        if (value <= 0) return CB_BAD_ARG;
        else if (value <= 500) value = 500;
        else if (value > 300000 * o->channels) value = 300000 * o->channels;
      }
      o->user_bitrate_bps = value;
      return CB_OK;
    case CB_REQ_GET_BITRATE: {
      int fs = o->prev_framesize ? o->prev_framesize : o->Fs / 400;
      if (o->user_bitrate_bps == CB_OPUS_AUTO) *out = 60 * o->Fs / fs + o->Fs * o->channels;
      else if (o->user_bitrate_bps == CB_OPUS_BITRATE_MAX) *out = 1276 * 8 * o->Fs / fs;
      else *out = o->user_bitrate_bps;
      return CB_OK;
    }
    case CB_REQ_SET_FORCE_CHANNELS:
      if ((value < 1 || value > o->channels) && value != CB_OPUS_AUTO) return CB_BAD_ARG;
      o->force_channels = value;
      return CB_OK;
    case CB_REQ_GET_FORCE_CHANNELS: *out = o->force_channels; return CB_OK;
    case CB_REQ_SET_MAX_BANDWIDTH:
      if (value < CB_BW_NARROWBAND || value > CB_BW_FULLBAND) return CB_BAD_ARG;
      o->max_bandwidth = value;
      return CB_OK;
    case CB_REQ_GET_MAX_BANDWIDTH: *out = o->max_bandwidth; return CB_OK;
    case CB_REQ_SET_BANDWIDTH:
      if ((value < CB_BW_NARROWBAND || value > CB_BW_FULLBAND) && value != CB_OPUS_AUTO) return CB_BAD_ARG;
      o->user_bandwidth = value;
      return CB_OK;
    case CB_REQ_GET_BANDWIDTH: *out = o->bandwidth; return CB_OK;
    case CB_REQ_SET_DTX:
      if (value < 0 || value > 1) return CB_BAD_ARG;
      o->use_dtx = value;
      return CB_OK;
    case CB_REQ_GET_DTX: *out = o->use_dtx; return CB_OK;
    case CB_REQ_SET_COMPLEXITY:
      if (value < 0 || value > 10) return CB_BAD_ARG;
      o->complexity = value;
      cs->complexity = value;
      return CB_OK;
    case CB_REQ_GET_COMPLEXITY: *out = o->complexity; return CB_OK;
    case CB_REQ_SET_INBAND_FEC:
      if (value < 0 || value > 1) return CB_BAD_ARG;
      o->use_inband_fec = value;
      return CB_OK;
    case CB_REQ_GET_INBAND_FEC: *out = o->use_inband_fec; return CB_OK;
    case CB_REQ_SET_PACKET_LOSS_PERC:
      if (value < 0 || value > 100) return CB_BAD_ARG;
      o->packet_loss_perc = value;
      cs->loss_rate = value;
      return CB_OK;
    case CB_REQ_GET_PACKET_LOSS_PERC: *out = o->packet_loss_perc; return CB_OK;
    case CB_REQ_SET_VBR:
      if (value < 0 || value > 1) return CB_BAD_ARG;
      o->use_vbr = value;
      return CB_OK;
    case CB_REQ_GET_VBR: *out = o->use_vbr; return CB_OK;
    case CB_REQ_SET_VBR_CONSTRAINT:
      if (value < 0 || value > 1) return CB_BAD_ARG;
      o->vbr_constraint = value;
      return CB_OK;
    case CB_REQ_GET_VBR_CONSTRAINT: *out = o->vbr_constraint; return CB_OK;
    case CB_REQ_SET_SIGNAL:
      if (value != CB_OPUS_AUTO && value != 3001 && value != 3002) return CB_BAD_ARG;
      o->signal_type = value;
      return CB_OK;
    case CB_REQ_GET_SIGNAL: *out = o->signal_type; return CB_OK;
    case CB_REQ_GET_LOOKAHEAD:
      *out = o->Fs / 400;
      if (o->application != CB_APP_LOWDELAY) *out += o->delay_compensation;
      return CB_OK;
    case CB_REQ_GET_SAMPLE_RATE: *out = o->Fs; return CB_OK;
    case CB_REQ_GET_FINAL_RANGE: *out = (i32)o->rangeFinal; return CB_OK;
    case CB_REQ_SET_LSB_DEPTH:
      if (value < 8 || value > 24) return CB_BAD_ARG;
      o->lsb_depth = value;
      return CB_OK;
    case CB_REQ_GET_LSB_DEPTH: *out = o->lsb_depth; return CB_OK;
    case CB_REQ_SET_EXPERT_FRAME_DURATION:
      if (value != 5000 && !(value >= 5001 && value <= 5006) && value != 5010) return CB_BAD_ARG;
      o->variable_duration = value;
      cs->variable_duration = value;
      return CB_OK;
    case CB_REQ_GET_EXPERT_FRAME_DURATION: *out = o->variable_duration; return CB_OK;
    case CB_REQ_SET_PREDICTION_DISABLED:
      if (value > 1 || value < 0) return CB_BAD_ARG;
      o->reduced_dependency = value;
      return CB_OK;
    case CB_REQ_GET_PREDICTION_DISABLED: *out = o->reduced_dependency; return CB_OK;
    case CB_REQ_SET_LFE:                     // opus_encoder.c:2455-2460, celt_encoder.c:2500-2504
      o->lfe = value;
      cs->lfe = value;
      return CB_OK;
    case CB_REQ_SET_FORCE_MODE:
      if ((value < CB_MODE_SILK_ONLY || value > CB_MODE_CELT_ONLY) && value != CB_OPUS_AUTO) return CB_BAD_ARG;
      o->user_forced_mode = value;
      return CB_OK;
    default:
      return CB_UNIMPLEMENTED;
  }
}
