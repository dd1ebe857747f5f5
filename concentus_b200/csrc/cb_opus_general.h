// cb_opus_general.h -- opus_decode_frame (opus-fix/src/opus_decoder.c:206-616) in full, for the frames
// the steady-state kernels do not take: hybrid packets, CELT<->SILK mode transitions (with their
// concealment cross-fades and decoder resets), redundant CELT frames inside SILK/hybrid packets, the
// hybrid->SILK fade-out frame, and concealment after hybrid frames.  One warp per stream runs the whole
// frame in order: the SILK layer on lane 0 (scalar code), the CELT layer on all lanes, then the fades.
// These frames are rare (a mode switch, or a hybrid stream), so clarity wins over speed here.
#pragma once
#include "cb_decode_pass.h"
#include "cb_silk_pass.h"

struct alignas(16) GenWork {
  union {
    struct { EntWork ew; u8 pkt[1280]; };
    SynWork sw;
    PlcWork pw;
  };
  i16 trans[2 * 240];            // pcm_transition: 5 ms
  i16 red[2 * 240];              // redundant_audio: 5 ms
  i16 silk_small[2 * 480];       // SILK output when the frame is shorter than 10 ms
  RangeDec rd;                   // the Opus-level range decoder (opus_decoder.c:209)
  i32 silk_err;
  i32 xfer[4];                   // lane 0 -> all lanes hand-off
  u8 sil[4];                     // the two-byte silence packet of the hybrid->SILK fade-out
  FrameTape* tape;               // this stream's tape (scratch arena): every CELT frame decoded here goes through it
};

// smooth_fade (opus_decoder.c:185-204)
CB_FN void gen_smooth_fade(const i16* in1, const i16* in2, i16* out, int overlap, int channels, int Fs) {
  const int inc = 48000 / Fs;
  CB_WSYNC();
  CB_LANE_LOOP(u, overlap * channels) {
    const int i = u / channels;
    const i32 wv = cb_ldg(&ct_window120[i * inc]);
    const i32 w = (i16)mult16_16_q15(wv, wv);
    out[u] = (i16)(mac16_16(mult16_16(w, in2[u]), CB_Q15ONE - w, in1[u]) >> 15);
  }
  CB_WSYNC();
}

// celt_decode_with_ec as one unit for a warp (celt_decoder.c:713-1072): entropy half, then synthesis of
// every channel; data == nullptr (or len <= 1) runs celt_decode_lost.  frame_size at the API rate.
CB_FN_NOINLINE void gen_celt_decode(GenWork& g, OpusDecState* st, FrameSide* side, const u8* data, int len, i16* pcm,
                           int frame_size, const RangeDec* rd0, int accum, int start, int end, int C, u32* final_range) {
  CeltDecState* celt = &st->celt;
  int LM = 0;
  {
    const int fs48 = frame_size * celt->downsample;
    while ((120 << LM) < fs48) LM++;
  }
  CB_WSYNC();
  if (data == nullptr || len <= 1) {
    if (celt->loss_count >= 5 || start != 0) {
      celt_plc_noise(g.ew, celt, side, g.tape, LM, start, end);
      CB_WSYNC();
      side->accum = accum;
      CB_WSYNC();
      for (int c = 0; c < celt->channels; c++) celt_synth_channel(g.sw, celt, side, c, pcm, 0);
    } else {
      celt_plc_pitch(g.pw, celt, 120 << LM, pcm, accum, 0);
    }
    CB_WSYNC();
    return;
  }
  CB_LANE_LOOP(i, len) g.pkt[i] = data[i];
  CB_WSYNC();
  EntFrameCfg cfg;
  cfg.LM = LM; cfg.C = C; cfg.start = start; cfg.end = end; cfg.accum = accum; cfg.rd0 = rd0;
  u32 fr = 0;
  celt_entropy_decode(g.ew, celt, side, g.tape, g.pkt, len, cfg, &fr);
  if (final_range) *final_range = fr;
  CB_WSYNC();
  for (int c = 0; c < celt->channels; c++) celt_synth_channel(g.sw, celt, side, c, pcm, 0);
  CB_WSYNC();
}

// The SILK layer of one frame on lane 0 (opus_decoder.c:352-409).  data == nullptr conceals.
CB_FN_NOINLINE void gen_silk_layer(GenWork& g, OpusDecState* st, const u8* data, int len, i16* pcm_silk, int mode, int audiosize,
                          int decode_fec, SilkCallWork& cw) {
  const int Fs = st->Fs, CC = st->channels;
  CB_WSYNC();
  CB_IF_LANE0 {
    g.silk_err = 0;
    SilkDecState& sd = st->silk;
    if (st->prev_mode == CB_MODE_CELT_ONLY) sk_init_decoder(sd);
    sd.ctl_payloadSize_ms = cb_imax(10, 1000 * audiosize / Fs);
    if (data != nullptr) {
      rd_init(g.rd, data, (u32)len);
      sd.ctl_nChannelsInternal = st->stream_channels;
      if (mode == CB_MODE_SILK_ONLY) {
        if (st->bandwidth == CB_BW_NARROWBAND) sd.ctl_internalSampleRate = 8000;
        else if (st->bandwidth == CB_BW_MEDIUMBAND) sd.ctl_internalSampleRate = 12000;
        else sd.ctl_internalSampleRate = 16000;
      } else {
        sd.ctl_internalSampleRate = 16000;
      }
    }
    const int lost_flag = data == nullptr ? SK_FLAG_PACKET_LOST : 2 * decode_fec;
    i16* pcm_ptr = pcm_silk;
    int decoded_samples = 0;
    do {
      i32 silk_frame_size = 0;
      const int silk_ret = sk_decode(sd, Fs, CC, lost_flag, decoded_samples == 0, g.rd, pcm_ptr, &silk_frame_size, cw);
      if (silk_ret) {
        if (lost_flag) {
          silk_frame_size = audiosize;
          for (int i = 0; i < audiosize * CC; i++) pcm_ptr[i] = 0;
        } else {
          g.silk_err = 1;
          break;
        }
      }
      pcm_ptr += silk_frame_size * CC;
      decoded_samples += silk_frame_size;
    } while (decoded_samples < audiosize);
  }
  CB_WSYNC();
}

// opus_decode_frame(st, NULL, 0, pcm, audiosize, 0) with audiosize <= 20 ms already chosen by the caller
// (the chunking of opus_decoder.c:249-292 is done by the plan): conceal with the previous mode.
CB_FN_NOINLINE void gen_conceal(GenWork& g, OpusDecState* st, FrameSide* side, i16* pcm, int audiosize, SilkCallWork& cw) {
  const int Fs = st->Fs, CC = st->channels;
  const int F10 = Fs / 100;
  const int mode = st->prev_mode;
  if (mode == 0) {
    CB_WSYNC();
    CB_LANE_LOOP(i, audiosize * CC) pcm[i] = 0;
    CB_WSYNC();
    return;
  }
  const int celt_accum = (mode != CB_MODE_CELT_ONLY) && (audiosize >= F10);
  i16* pcm_silk = celt_accum ? pcm : g.silk_small;
  if (mode != CB_MODE_CELT_ONLY) gen_silk_layer(g, st, nullptr, 0, pcm_silk, mode, audiosize, 0, cw);
  const int start_band = mode != CB_MODE_CELT_ONLY ? 17 : 0;
  const int endband = opus_bandwidth_endband(st->bandwidth);
  if (mode != CB_MODE_SILK_ONLY) {
    gen_celt_decode(g, st, side, nullptr, 0, pcm, cb_imin(Fs / 50, audiosize), nullptr, celt_accum, start_band, endband,
                    st->stream_channels, nullptr);
  } else if (!celt_accum) {
    CB_WSYNC();
    CB_LANE_LOOP(i, audiosize * CC) pcm[i] = 0;
    CB_WSYNC();
  }
  if (mode != CB_MODE_CELT_ONLY && !celt_accum) {
    CB_WSYNC();
    CB_LANE_LOOP(i, audiosize * CC) pcm[i] = sat16((i32)pcm[i] + (i32)g.silk_small[i]);
    CB_WSYNC();
  }
  if (st->decode_gain) {
    const i32 gain = cb_exp2(mult16_16_p15(21771, st->decode_gain));
    CB_WSYNC();
    CB_LANE_LOOP(i, audiosize * CC) {
      const i32 x = mult16_32_p16(pcm[i], gain);
      pcm[i] = (i16)(x > 32767 ? 32767 : x < -32767 ? -32767 : x);
    }
  }
  CB_WSYNC();
  CB_IF_LANE0 {
    st->rangeFinal = 0;
    st->prev_mode = mode;
    st->prev_redundancy = 0;
  }
  CB_WSYNC();
}

// opus_decode_frame for one frame of the plan.  data == nullptr conceals.
CB_FN int gen_decode_frame(GenWork& g, OpusDecState* st, FrameSide* side, const u8* data, int len, i16* pcm,
                           int frame_size, int decode_fec, SilkCallWork& cw, const u32* resume_rd = nullptr) {
  const int Fs = st->Fs, CC = st->channels;
  const int F20 = Fs / 50, F10 = F20 >> 1, F5 = F10 >> 1, F2_5 = F5 >> 1;
  // data != nullptr and len > 1 here: 0/1-byte frames are concealed through gen_conceal
  const int audiosize = st->frame_size, mode = st->mode;
  (void)F20;
  const int celt_accum = (mode != CB_MODE_CELT_ONLY) && (audiosize >= F10);
  int transition = 0;
  if (st->prev_mode > 0 &&
      ((mode == CB_MODE_CELT_ONLY && st->prev_mode != CB_MODE_CELT_ONLY && !st->prev_redundancy) ||
       (mode != CB_MODE_CELT_ONLY && st->prev_mode == CB_MODE_CELT_ONLY)))
    transition = 1;
  if (transition && mode == CB_MODE_CELT_ONLY) gen_conceal(g, st, side, g.trans, cb_imin(F5, audiosize), cw);
  frame_size = audiosize;

  // ---- SILK layer (lane 0) ----
  i16* pcm_silk = celt_accum ? pcm : g.silk_small;
  if (resume_rd) {
    // the SILK kernel has decoded the SILK layer into pcm and left its coder state (late redundancy)
    CB_WSYNC();
    CB_IF_LANE0 {
      u32* dst = (u32*)&g.rd;
      for (int i = 0; i < (int)(sizeof(RangeDec) / 4); i++) dst[i] = resume_rd[i];
    }
    CB_WSYNC();
  } else if (mode != CB_MODE_CELT_ONLY) {
    gen_silk_layer(g, st, data, len, pcm_silk, mode, audiosize, decode_fec, cw);
    if (g.silk_err) return CB_INTERNAL_ERROR;
  } else {
    CB_WSYNC();
    CB_IF_LANE0 rd_init(g.rd, data, (u32)len);
    CB_WSYNC();
  }

  // ---- redundancy side information (opus_decoder.c:412-440) ----
  int start_band = 0, redundancy = 0, redundancy_bytes = 0, celt_to_silk = 0;
  if (!decode_fec && mode != CB_MODE_CELT_ONLY) {
    CB_WSYNC();
    CB_IF_LANE0 {
      RangeDec& rd = g.rd;
      int red = 0, c2s = 0, rbytes = 0, ln = len;
      if ((int)rd_tell(rd) + 17 + 20 * (st->mode == CB_MODE_HYBRID) <= 8 * len) {
        if (mode == CB_MODE_HYBRID) red = rd_bit_logp(rd, 12);
        else red = 1;
        if (red) {
          c2s = rd_bit_logp(rd, 1);
          rbytes = mode == CB_MODE_HYBRID ? (int)rd_uint(rd, 256) + 2 : len - (((int)rd_tell(rd) + 7) >> 3);
          ln -= rbytes;
          if (ln * 8 < (int)rd_tell(rd)) { ln = 0; rbytes = 0; red = 0; }
          rd.storage -= (u32)rbytes;
        }
      }
      g.xfer[0] = red; g.xfer[1] = c2s; g.xfer[2] = rbytes; g.xfer[3] = ln;
    }
    CB_WSYNC();
    redundancy = g.xfer[0]; celt_to_silk = g.xfer[1]; redundancy_bytes = g.xfer[2]; len = g.xfer[3];
    CB_WSYNC();
  }
  if (mode != CB_MODE_CELT_ONLY) start_band = 17;
  const int endband = opus_bandwidth_endband(st->bandwidth);
  if (redundancy) transition = 0;
  if (transition && mode != CB_MODE_CELT_ONLY) gen_conceal(g, st, side, g.trans, cb_imin(F5, audiosize), cw);

  u32 redundant_rng = 0;
  if (redundancy && celt_to_silk)
    gen_celt_decode(g, st, side, data + len, redundancy_bytes, g.red, F5, nullptr, 0, 0, endband, st->stream_channels,
                    &redundant_rng);

  // ---- CELT layer ----
  u32 celt_rng = 0;
  int have_celt_rng = 0;
  if (mode != CB_MODE_SILK_ONLY) {
    const int celt_frame_size = cb_imin(F20, frame_size);
    if (mode != st->prev_mode && st->prev_mode > 0 && !st->prev_redundancy) {
      CB_WSYNC();
      CB_IF_LANE0 celt_state_reset(&st->celt, st->celt.channels, st->celt.downsample);
      CB_WSYNC();
    }
    const RangeDec rd_copy = g.rd;
    const bool coded = !decode_fec;
    gen_celt_decode(g, st, side, coded ? data : nullptr, len, pcm, celt_frame_size, coded ? &rd_copy : nullptr, celt_accum,
                    start_band, endband, st->stream_channels, &celt_rng);
    have_celt_rng = coded && len > 1;
  } else {
    if (!celt_accum && !resume_rd) {
      CB_LANE_LOOP(i, frame_size * CC) pcm[i] = 0;
      CB_WSYNC();
    }
    if (st->prev_mode == CB_MODE_HYBRID && !(redundancy && celt_to_silk && st->prev_redundancy)) {
      // hybrid -> SILK: let the CELT MDCT fade out by decoding a silence frame
      CB_WSYNC();
      CB_IF_LANE0 { g.sil[0] = 0xFF; g.sil[1] = 0xFF; }
      CB_WSYNC();
      gen_celt_decode(g, st, side, g.sil, 2, pcm, F2_5, nullptr, celt_accum, 0, endband, st->stream_channels, nullptr);
    }
  }
  if (mode != CB_MODE_CELT_ONLY && !celt_accum && !resume_rd) {
    CB_WSYNC();
    CB_LANE_LOOP(i, frame_size * CC) pcm[i] = sat16((i32)pcm[i] + (i32)g.silk_small[i]);
    CB_WSYNC();
  }
  if (redundancy && !celt_to_silk) {
    CB_WSYNC();
    CB_IF_LANE0 celt_state_reset(&st->celt, st->celt.channels, st->celt.downsample);
    CB_WSYNC();
    gen_celt_decode(g, st, side, data + len, redundancy_bytes, g.red, F5, nullptr, 0, 0, endband, st->stream_channels,
                    &redundant_rng);
    gen_smooth_fade(pcm + CC * (frame_size - F2_5), g.red + CC * F2_5, pcm + CC * (frame_size - F2_5), F2_5, CC, Fs);
  }
  if (redundancy && celt_to_silk) {
    CB_WSYNC();
    CB_LANE_LOOP(i, F2_5 * CC) pcm[i] = g.red[i];
    gen_smooth_fade(g.red + CC * F2_5, pcm + CC * F2_5, pcm + CC * F2_5, F2_5, CC, Fs);
  }
  if (transition) {
    if (audiosize >= F5) {
      CB_WSYNC();
      CB_LANE_LOOP(i, CC * F2_5) pcm[i] = g.trans[i];
      gen_smooth_fade(g.trans + CC * F2_5, pcm + CC * F2_5, pcm + CC * F2_5, F2_5, CC, Fs);
    } else {
      gen_smooth_fade(g.trans, pcm, pcm, F2_5, CC, Fs);
    }
  }
  if (st->decode_gain) {
    const i32 gain = cb_exp2(mult16_16_p15(21771, st->decode_gain));
    CB_WSYNC();
    CB_LANE_LOOP(i, frame_size * CC) {
      const i32 x = mult16_32_p16(pcm[i], gain);
      pcm[i] = (i16)(x > 32767 ? 32767 : x < -32767 ? -32767 : x);
    }
  }
  CB_WSYNC();
  u32 final_rng;
  if (len <= 1) final_rng = 0;
  else final_rng = (have_celt_rng ? celt_rng : g.rd.rng) ^ redundant_rng;
  CB_WSYNC();
  CB_IF_LANE0 {
    st->rangeFinal = final_rng;
    st->prev_mode = mode;
    st->prev_redundancy = redundancy && !celt_to_silk;
  }
  CB_WSYNC();
  return audiosize;
}

// Kernel G, one stream, one pass: frames routed here by kernel A (or handed over by the SILK kernel).
CB_FN void decode_general_pass(GenWork& g, OpusDecState* st, StreamPlan* pl, FrameSide* side, FrameTape* tape, const u8* data,
                               i16* row, i32* ret_slot, int pass, SilkCallWork& cw) {
  const int route = pl->route;
  if (route != CB_ROUTE_GENERAL && route != CB_ROUTE_FINISH) return;
  CB_WSYNC();
  CB_IF_LANE0 g.tape = tape;
  CB_WSYNC();
  const int CC = st->channels;
  i16* pcm = row + (size_t)pl->pcm_off[pass] * CC;
  const int len = pl->len[pass];
  if (pl->kind == CB_PLAN_PLC || len <= 1) {
    gen_conceal(g, st, side, pcm, pl->nsamp[pass], cw);
    return;
  }
  const int ret = gen_decode_frame(g, st, side, data + pl->off[pass], len, pcm, pl->nsamp[pass], pl->fec_frame == pass, cw,
                                   route == CB_ROUTE_FINISH ? pl->rd_save : nullptr);
  if (ret < 0) {
    CB_WSYNC();
    CB_IF_LANE0 { pl->ret = ret; *ret_slot = ret; }
    CB_WSYNC();
  }
}
