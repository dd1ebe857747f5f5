// cb_silk_pass.h -- the SILK branch of opus_decode_frame (opus-fix/src/opus_decoder.c:206-616) for
// one stream and one frame of its plan, run by ONE THREAD.  Covers SILK-only packets (NB/MB/WB,
// mono/stereo, 10-60 ms), their concealment (lost packets, DTX) and in-band FEC (LBRR).
// Not built yet and answered with OPUS_UNIMPLEMENTED: hybrid packets, CELT<->SILK mode transitions
// and SILK packets that carry a redundant CELT frame.
#pragma once
#include "cb_state.h"
#include "cb_opus_plan.h"
#include "cb_silk_dec.h"

CB_FN void decode_silk_pass(OpusDecState* st, StreamPlan* pl, FrameSide* side, const u8* data, i16* row, i32* ret_slot,
                            int pass, SilkCallWork& cw) {
  {
    // first kernel of the pass: decide who decodes this stream's frame
    const int kind = pl->kind;
    int route = -1;
    if (!(pl->ret < 0 || pass >= pl->nframes || kind == CB_PLAN_NONE || kind == CB_PLAN_ZERO)) route = opus_frame_route(st, pl, pass);
    pl->route = route;
    if (route == CB_ROUTE_PLC) {
      // lost CELT frame, pitch-based concealment: the general kernel does it, concurrently with the CELT
      // kernels of the other streams; the Opus-layer bookkeeping of opus_decode_frame(data = NULL) is done here
      int LMp = 0;
      const int fs48 = pl->nsamp[pass] * st->celt.downsample;
      while ((120 << LMp) < fs48) LMp++;
      side->kind = 2;
      side->LM = LMp;
      side->accum = 0;
      st->rangeFinal = 0;
      st->prev_redundancy = 0;
      return;
    }
    if (route != CB_ROUTE_SILK && route != CB_ROUTE_HYBRID) return;
  }
  const bool hybrid = pl->route == CB_ROUTE_HYBRID;
  const int Fs = st->Fs, CC = st->channels;
  const int F20 = Fs / 50, F10 = F20 >> 1;
  const int decode_fec = pl->fec_frame == pass;
  int len = pl->len[pass];
  const bool have_data = len > 1;                  // payloads of 0/1 bytes conceal (opus_decoder.c:241-247)
  const int audiosize = pl->nsamp[pass];
  const int frame_size = audiosize;
  const int mode = have_data ? st->mode : st->prev_mode;
  i16* pcm = row + (size_t)pl->pcm_off[pass] * CC;
  RangeDec rd;
  const u8* pkt = data + pl->off[pass];
  if (have_data) rd_init(rd, pkt, (u32)len);
  SilkDecState& sd = st->silk;
  // (silk_InitDecoder when prev_mode == CELT_ONLY: excluded above)
  sd.ctl_payloadSize_ms = cb_imax(10, 1000 * audiosize / Fs);
  if (have_data) {
    sd.ctl_nChannelsInternal = st->stream_channels;
    if (hybrid) sd.ctl_internalSampleRate = 16000;
    else if (st->bandwidth == CB_BW_NARROWBAND) sd.ctl_internalSampleRate = 8000;
    else if (st->bandwidth == CB_BW_MEDIUMBAND) sd.ctl_internalSampleRate = 12000;
    else sd.ctl_internalSampleRate = 16000;
  }
  const int lost_flag = !have_data ? SK_FLAG_PACKET_LOST : 2 * decode_fec;
  // SILK output: straight into the PCM row when the frame is at least 10 ms (the reference's
  // celt_accum case), else through a 10 ms scratch of which only frame_size samples are kept
  i16 small[2 * 480];
  const bool direct = frame_size >= F10;
  i16* pcm_ptr = direct ? pcm : small;
  int decoded_samples = 0;
  do {
    const int first_frame = decoded_samples == 0;
    i32 silk_frame_size = 0;
    const int silk_ret = sk_decode(sd, Fs, CC, lost_flag, first_frame, rd, pcm_ptr, &silk_frame_size, cw);
    if (silk_ret) {
      if (lost_flag) {
        silk_frame_size = frame_size;
        for (int i = 0; i < frame_size * CC; i++) pcm_ptr[i] = 0;
      } else {
        pl->ret = CB_INTERNAL_ERROR;
        *ret_slot = CB_INTERNAL_ERROR;
        return;
      }
    }
    pcm_ptr += silk_frame_size * CC;
    decoded_samples += silk_frame_size;
  } while (decoded_samples < frame_size);
  if (!direct)
    for (int i = 0; i < frame_size * CC; i++) pcm[i] = small[i];
  if (hybrid) {
    // the CELT kernels continue from this coder state with bands 17.. accumulated onto the SILK output;
    // a hybrid packet flags its redundant CELT frame explicitly (opus_decoder.c:412-440)
    if (have_data) {
      const RangeDec before = rd;
      int red = 0;
      if ((int)rd_tell(rd) + 37 <= 8 * len) red = rd_bit_logp_g(rd, 12);
      const u32* src = (const u32*)(red ? &before : &rd);
      for (int i = 0; i < (int)(sizeof(RangeDec) / 4); i++) pl->rd_save[i] = src[i];
      if (red) pl->route = CB_ROUTE_FINISH;
    }
    return;
  }
  // a SILK-only packet with >= 17 bits left over carries a redundant CELT frame (opus_decoder.c:412-440)
  if (!decode_fec && have_data && (int)rd_tell(rd) + 17 <= 8 * len) {
    // the general kernel (cb_opus_general.h) decodes the CELT frame and does the cross-fade: hand over
    // the coder state; gain and Opus-layer state are applied there
    const u32* src = (const u32*)&rd;
    for (int i = 0; i < (int)(sizeof(RangeDec) / 4); i++) pl->rd_save[i] = src[i];
    pl->route = CB_ROUTE_FINISH;
    return;
  }
  if (st->decode_gain) {
    const i32 gain = cb_exp2(mult16_16_p15(21771, st->decode_gain));
    for (int i = 0; i < frame_size * CC; i++) {
      const i32 x = mult16_32_p16(pcm[i], gain);
      pcm[i] = (i16)(x > 32767 ? 32767 : x < -32767 ? -32767 : x);
    }
  }
  st->rangeFinal = len <= 1 ? 0 : rd.rng;
  st->prev_mode = mode;
  st->prev_redundancy = 0;
}
