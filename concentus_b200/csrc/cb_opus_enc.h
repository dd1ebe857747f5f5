// cb_opus_enc.h -- the Opus layer of the encoder for CELT-only packets: opus_encode /
// opus_encode_native (opus-fix/src/opus_encoder.c:938-1964), FIXED_POINT + DISABLE_FLOAT_API.
// Restated: bit-rate / channel / mode / bandwidth decisions, DC rejection, the delay buffer,
// stereo-width fade, TOC byte, CBR padding.  A frame whose mode decision is SILK or hybrid, the
// VOIP high-pass, and 40/60 ms frames return OPUS_UNIMPLEMENTED (SILK is a later row of the plan).
#pragma once
#include "cb_common.h"
#include "cb_enc_state.h"
#include "cb_celt_enc.h"
#include "cb_opus_plan.h"
#include "cb_repack.h"

#define CB_SIGNAL_VOICE 3001
#define CB_SIGNAL_MUSIC 3002
#define CB_FRAMESIZE_ARG 5000
#define CB_FRAMESIZE_2_5_MS 5001
#define CB_FRAMESIZE_60_MS 5006
#define CB_FRAMESIZE_VARIABLE 5010

CB_UTABLE i32 ct_bw_thresh_mono_voice[8] = {11000, 1000, 14000, 1000, 17000, 1000, 21000, 2000};
CB_UTABLE i32 ct_bw_thresh_mono_music[8] = {12000, 1000, 15000, 1000, 18000, 2000, 22000, 2000};
CB_UTABLE i32 ct_bw_thresh_stereo_voice[8] = {11000, 1000, 14000, 1000, 21000, 2000, 28000, 2000};
CB_UTABLE i32 ct_bw_thresh_stereo_music[8] = {12000, 1000, 18000, 2000, 21000, 2000, 30000, 2000};

// frame_size_select (opus_encoder.c:807-826)
CB_HD int enc_frame_size_select(int frame_size, int variable_duration, int Fs) {
  int new_size;
  if (frame_size < Fs / 400) return -1;
  if (variable_duration == CB_FRAMESIZE_ARG) new_size = frame_size;
  else if (variable_duration == CB_FRAMESIZE_VARIABLE) new_size = Fs / 50;
  else if (variable_duration >= CB_FRAMESIZE_2_5_MS && variable_duration <= CB_FRAMESIZE_60_MS)
    new_size = cb_imin(3 * Fs / 50, (Fs / 400) << (variable_duration - CB_FRAMESIZE_2_5_MS));
  else return -1;
  if (new_size > frame_size) return -1;
  if (400 * new_size != Fs && 200 * new_size != Fs && 100 * new_size != Fs && 50 * new_size != Fs &&
      25 * new_size != Fs && 50 * new_size != 3 * Fs)
    return -1;
  return new_size;
}

// gen_toc (opus_encoder.c:254-284)
CB_HD u8 enc_gen_toc(int mode, int framerate, int bandwidth, int channels) {
  int period = 0;
  while (framerate < 400) { framerate <<= 1; period++; }
  int toc;
  if (mode == CB_MODE_SILK_ONLY) {
    toc = (bandwidth - CB_BW_NARROWBAND) << 5;
    toc |= (period - 2) << 3;
  } else if (mode == CB_MODE_CELT_ONLY) {
    int tmp = bandwidth - CB_BW_MEDIUMBAND;
    if (tmp < 0) tmp = 0;
    toc = 0x80 | (tmp << 5) | (period << 3);
  } else {
    toc = 0x60 | ((bandwidth - CB_BW_SUPERWIDEBAND) << 4) | ((period - 2) << 3);
  }
  toc |= (channels == 2) << 2;
  return (u8)toc;
}

// opus_packet_pad (repacketizer.c:102-258) for the single-frame packets this encoder produces:
// code 0 -> code 3 with padding.  Returns 0 or a negative code.
CB_FN_NOINLINE int enc_packet_pad(u8* data, int len, int new_len) {
  if (len < 1) return CB_BAD_ARG;
  if (len == new_len) return CB_OK;
  if (len > new_len) return CB_BAD_ARG;
  const int L = len - 1;                     // payload bytes after the TOC
  const u8 toc = data[0];
  if (L + 2 > new_len) return CB_BUFFER_TOO_SMALL;
  const int pad_amount = new_len - (L + 2);
  const int nb_255s = pad_amount != 0 ? (pad_amount - 1) / 255 : 0;
  const int hdr = 2 + (pad_amount != 0 ? nb_255s + 1 : 0);
  CB_WSYNC();
  // move the payload up (highest index first so the in-place move is safe), one lane
  CB_IF_LANE0 {
    for (int i = L - 1; i >= 0; i--) data[hdr + i] = data[1 + i];
    data[0] = (u8)((toc & 0xFC) | 0x3);
    data[1] = (u8)(1 | (pad_amount != 0 ? 0x40 : 0));
    if (pad_amount != 0) {
      for (int i = 0; i < nb_255s; i++) data[2 + i] = 255;
      data[2 + nb_255s] = (u8)(pad_amount - 255 * nb_255s - 1);
    }
    for (int i = hdr + L; i < new_len; i++) data[i] = 0;
  }
  CB_WSYNC();
  return CB_OK;
}

// buf[i] = buf[i + shift] for i < n (an OPUS_MOVE toward lower addresses): ascending chunks, each
// read by all lanes before any lane writes it; later chunks never read what earlier ones wrote.
CB_FN void enc_move_down(i16* buf, int shift, int n) {
#if CB_DEVICE_BUILD
  CB_NOUNROLL for (int base = 0; base < n; base += 32) {
    const int i = base + CB_LANE();
    i16 v = 0;
    if (i < n) v = buf[i + shift];
    CB_WSYNC();
    if (i < n) buf[i] = v;
    CB_WSYNC();
  }
#else
  for (int i = 0; i < n; i++) buf[i] = buf[i + shift];
#endif
}

// Host build: both channels in turn.  (The device posts a request instead: enc_post_dc below.)
CB_FN void enc_dc_reject(const i16* in, int cutoff_Hz, i16* out, i32* hp_mem, int len, int channels, int Fs) {
  const int shift = cb_ilog2(Fs / (cutoff_Hz * 3));
  for (int c = 0; c < channels; c++) enc_dc_chain(in, out, hp_mem, c, len, channels, shift);
}

#if CB_DEVICE_BUILD
CB_FN void enc_post_dc(EncWork& w, const i16* in, int cutoff_Hz, i16* out, const i32* hp_mem, int len, int channels, int Fs) {
  CB_WSYNC();
  if (CB_LANE() == 0) {
    EncSerialReq& r = *w.sreq;
    r.in = in; r.out = out; r.len = len; r.channels = channels; r.shift = cb_ilog2(Fs / (cutoff_Hz * 3));
    for (int k = 0; k < 4; k++) r.m[k] = hp_mem[k];
    r.kind = CB_ES_DC;
  }
  CB_WSYNC();
}
#endif

// stereo_fade (opus_encoder.c:411-441), in place
CB_FN void enc_stereo_fade(i16* buf, i32 g1, i32 g2, int frame_size, int Fs) {
  const int inc = 48000 / Fs;
  const int overlap = CB_OVERLAP / inc;
  g1 = (i16)(CB_Q15ONE - g1);
  g2 = (i16)(CB_Q15ONE - g2);
  CB_WSYNC();
  CB_LANE_LOOP(i, frame_size) {
    i32 g = g2;
    if (i < overlap) {
      const i32 wv = cb_ldg(&ct_window120[i * inc]);
      const i32 wq = (i16)mult16_16_q15(wv, wv);
      g = (i16)(mac16_16(mult16_16(wq, g2), CB_Q15ONE - wq, g1) >> 15);
    }
    i32 diff = (i16)(((i32)buf[2 * i] - (i32)buf[2 * i + 1]) >> 1);
    diff = mult16_16_q15(g, diff);
    buf[2 * i] = (i16)(buf[2 * i] - diff);
    buf[2 * i + 1] = (i16)(buf[2 * i + 1] + diff);
  }
  CB_WSYNC();
}

// The few SILK fixed-point primitives the VOIP high-pass needs (silk/macros.h:53,60,92; SigProc_FIX.h:561;
// lin2log.c, log2lin.c), 64-bit forms as the reference compiles on x86-64.
CB_FN i32 esk_smulwb(i32 a, i32 b) { return (i32)(((long long)a * (long long)(i16)b) >> 16); }
CB_FN i32 esk_smlawb(i32 a, i32 b, i32 c) { return (i32)((u32)a + (u32)esk_smulwb(b, c)); }
CB_FN i32 esk_smulww(i32 a, i32 b) { return (i32)(((long long)a * (long long)b) >> 16); }
CB_FN i32 esk_rshift_round(i32 a, int s) { return s == 1 ? (a >> 1) + (a & 1) : ((a >> (s - 1)) + 1) >> 1; }
CB_FN i32 esk_lin2log(i32 inLin) {
  const int lz = cb_clz((u32)inLin);
  const u32 x = (u32)inLin;
  const int rot = 24 - lz;                                    // silk_ROR32
  const u32 r = rot == 0 ? x : rot < 0 ? ((x << (-rot)) | (x >> (32 + rot))) : ((x << (32 - rot)) | (x >> rot));
  const i32 frac_Q7 = (i32)(r & 0x7f);
  return ((31 - lz) << 7) + esk_smlawb(frac_Q7, frac_Q7 * (128 - frac_Q7), 179);
}
CB_FN i32 esk_log2lin(i32 inLog_Q7) {
  if (inLog_Q7 < 0) return 0;
  if (inLog_Q7 >= 3967) return 0x7fffffff;
  i32 out = (i32)(1u << (inLog_Q7 >> 7));
  const i32 frac_Q7 = inLog_Q7 & 0x7F;
  const i32 t = esk_smlawb(frac_Q7, (i32)(i16)frac_Q7 * (i32)(i16)(128 - frac_Q7), -174);
  if (inLog_Q7 < 2048) out = out + ((out * t) >> 7);
  else out = out + (out >> 7) * t;
  return out;
}

// hp_cutoff (opus_encoder.c:325-360) + silk_biquad_alt (silk/biquad_alt.c:41-78): the VOIP application's input
// high-pass, a second-order section per channel (direct form II transposed); one lane per channel.
CB_FN void enc_hp_cutoff(const i16* in, i32 cutoff_Hz, i16* out, i32* hp_mem, int len, int channels, int Fs) {
  const i32 Fc_Q19 = ((i32)(i16)2471 * (i32)(i16)cutoff_Hz) / (i32)(i16)(Fs / 1000);   // 1.5*pi/1000 in Q19
  const i32 r_Q28 = (1 << 28) - 471 * Fc_Q19;                                               // 0.92 in Q9
  const i32 B0 = r_Q28, B1 = (i32)((u32)(-r_Q28) << 1), B2 = r_Q28;
  const i32 r_Q22 = r_Q28 >> 6;
  const i32 A0 = esk_smulww(r_Q22, esk_smulww(Fc_Q19, Fc_Q19) - (2 << 22));
  const i32 A1 = esk_smulww(r_Q22, r_Q22);
  const i32 A0_L = (-A0) & 0x3FFF, A0_U = (-A0) >> 14, A1_L = (-A1) & 0x3FFF, A1_U = (-A1) >> 14;
  CB_WSYNC();
#if CB_DEVICE_BUILD
  const int c = CB_LANE();
  if (c < channels) {
#else
  for (int c = 0; c < channels; c++) {
#endif
    i32 S0 = hp_mem[2 * c], S1 = hp_mem[2 * c + 1];
    CB_NOUNROLL for (int k = 0; k < len; k++) {
      const i32 inval = in[k * channels + c];
      const i32 out32_Q14 = (i32)((u32)esk_smlawb(S0, B0, inval) << 2);
      S0 = S1 + esk_rshift_round(esk_smulwb(out32_Q14, A0_L), 14);
      S0 = esk_smlawb(S0, out32_Q14, A0_U);
      S0 = esk_smlawb(S0, B1, inval);
      S1 = esk_rshift_round(esk_smulwb(out32_Q14, A1_L), 14);
      S1 = esk_smlawb(S1, out32_Q14, A1_U);
      S1 = esk_smlawb(S1, B2, inval);
      i32 v = (out32_Q14 + (1 << 14) - 1) >> 14;
      v = v > 32767 ? 32767 : (v < -32768 ? -32768 : v);
      out[k * channels + c] = (i16)v;
    }
    hp_mem[2 * c] = S0;
    hp_mem[2 * c + 1] = S1;
  }
  CB_WSYNC();
}

// compute_stereo_width (opus_encoder.c:861-936); the StereoWidthState lives in the scalars
CB_FN i32 enc_stereo_width(const i16* pcm, int frame_size, int Fs, OpusEncScalars& o) {
  const int frame_rate = Fs / frame_size;
  const i32 short_alpha = (i16)(CB_Q15ONE - 25 * CB_Q15ONE / cb_imax(50, frame_rate));
  i32 xx = 0, xy = 0, yy = 0;
  CB_WSYNC();
  CB_LANE_LOOP(q, frame_size / 4) {
    const int i = 4 * q;
    i32 pxx = 0, pxy = 0, pyy = 0;
    for (int k = 0; k < 4; k++) {
      const i32 x = pcm[2 * i + 2 * k], y = pcm[2 * i + 2 * k + 1];
      pxx += mult16_16(x, x) >> 2;
      pxy += mult16_16(x, y) >> 2;
      pyy += mult16_16(y, y) >> 2;
    }
    xx += pxx >> 10; xy += pxy >> 10; yy += pyy >> 10;
  }
  xx = cb_wsum(xx); xy = cb_wsum(xy); yy = cb_wsum(yy);
  o.width_XX += mult16_32_q15(short_alpha, xx - o.width_XX);
  o.width_XY += mult16_32_q15(short_alpha, xy - o.width_XY);
  o.width_YY += mult16_32_q15(short_alpha, yy - o.width_YY);
  o.width_XX = max32(0, o.width_XX);
  o.width_XY = max32(0, o.width_XY);
  o.width_YY = max32(0, o.width_YY);
  if (max32(o.width_XX, o.width_YY) > 210) {                  // QCONST16(8e-4f, 18)
    const i32 sqrt_xx = (i16)cb_sqrt(o.width_XX), sqrt_yy = (i16)cb_sqrt(o.width_YY);
    const i32 qrrt_xx = (i16)cb_sqrt(sqrt_xx), qrrt_yy = (i16)cb_sqrt(sqrt_yy);
    o.width_XY = min32(o.width_XY, sqrt_xx * sqrt_yy);
    const i32 corr = (i16)(cb_frac_div32(o.width_XY, 1 + mult16_16(sqrt_xx, sqrt_yy)) >> 16);
    const i32 ad = qrrt_xx - qrrt_yy;
    const i32 ldiff = (i16)(CB_Q15ONE * (ad < 0 ? -ad : ad) / (1 + qrrt_xx + qrrt_yy));
    const i32 width = (i16)mult16_16_q15(cb_sqrt(1073741824 - mult16_16(corr, corr)), ldiff);
    o.width_smoothed = (i16)(o.width_smoothed + (width - o.width_smoothed) / frame_rate);
    o.width_max_follower = (i16)max32(o.width_max_follower - 655 / frame_rate, o.width_smoothed);
  }
  return (i16)min32(CB_Q15ONE, 20 * o.width_max_follower);
}

// opus_encode (opus_encoder.c:2012-2031) + opus_encode_native (:938-1964), CELT-only path.
//
// enc_decide: the decisions opus_encode_native takes before any signal processing -- bit-rate bookkeeping,
// the one-byte packet when there is too little room, channel count, mode, bandwidth (:1014-1359).  `o` is the
// stream's scalar state (in registers), max_data_bytes comes in as min(1276, room) and goes out CBR-adjusted.
// Returns 0 to go on, 1 when the packet is complete (*early_ret = its length), or a negative OPUS_* code.
struct EncDecision {
  int max_data_bytes, total_buffer, lsb_depth;
  i32 equiv_rate;
};
CB_FN int enc_decide(OpusEncState* st, OpusEncScalars& o, const i16* pcm, int frame_size, u8* data, EncDecision& d,
                     int* early_ret) {
  int max_data_bytes = d.max_data_bytes;
  const int delay_compensation = o.application == CB_APP_LOWDELAY ? 0 : o.delay_compensation;
  const int lsb_depth = cb_imin(16, o.lsb_depth);
  o.voice_ratio = -1;
  i32 stereo_width = 0;
  if (o.channels == 2 && o.force_channels != 1) stereo_width = enc_stereo_width(pcm, frame_size, o.Fs, o);
  const int total_buffer = delay_compensation;
  // user_bitrate_to_bitrate (opus_encoder.c:512-521)
  if (o.user_bitrate_bps == CB_OPUS_AUTO) o.bitrate_bps = 60 * o.Fs / frame_size + o.Fs * o.channels;
  else if (o.user_bitrate_bps == CB_OPUS_BITRATE_MAX) o.bitrate_bps = max_data_bytes * 8 * o.Fs / frame_size;
  else o.bitrate_bps = o.user_bitrate_bps;
  const int frame_rate = o.Fs / frame_size;
  if (!o.use_vbr) {
    const int frame_rate3 = 3 * o.Fs / frame_size;
    const int cbrBytes = cb_imin((3 * o.bitrate_bps / 8 + frame_rate3 / 2) / frame_rate3, max_data_bytes);
    o.bitrate_bps = cbrBytes * (i32)frame_rate3 * 8 / 3;
    max_data_bytes = cbrBytes;
  }
  if (max_data_bytes < 3 || o.bitrate_bps < 3 * frame_rate * 8 ||
      (frame_rate < 50 && (max_data_bytes * frame_rate < 300 || o.bitrate_bps < 2400))) {
    // too little room: a one-byte packet the decoder conceals (opus_encoder.c:1079-1106)
    int tocmode = o.mode;
    int bw = o.bandwidth == 0 ? CB_BW_NARROWBAND : o.bandwidth;
    if (tocmode == 0) tocmode = CB_MODE_SILK_ONLY;
    if (frame_rate > 100) tocmode = CB_MODE_CELT_ONLY;
    if (frame_rate < 50) tocmode = CB_MODE_SILK_ONLY;
    if (tocmode == CB_MODE_SILK_ONLY && bw > CB_BW_WIDEBAND) bw = CB_BW_WIDEBAND;
    else if (tocmode == CB_MODE_CELT_ONLY && bw == CB_BW_MEDIUMBAND) bw = CB_BW_NARROWBAND;
    else if (tocmode == CB_MODE_HYBRID && bw <= CB_BW_SUPERWIDEBAND) bw = CB_BW_SUPERWIDEBAND;
    CB_WSYNC();
    data[0] = enc_gen_toc(tocmode, frame_rate, bw, o.stream_channels);
    CB_WSYNC();
    int ret = 1;
    if (!o.use_vbr) {
      ret = enc_packet_pad(data, ret, max_data_bytes);
      if (ret == CB_OK) ret = max_data_bytes;
    }
    *early_ret = ret;
    return 1;
  }
  i32 equiv_rate = o.bitrate_bps - (40 * o.channels + 20) * (o.Fs / frame_size - 50);
  int voice_est;
  if (o.signal_type == CB_SIGNAL_VOICE) voice_est = 127;
  else if (o.signal_type == CB_SIGNAL_MUSIC) voice_est = 0;
  else if (o.voice_ratio >= 0) voice_est = o.voice_ratio * 327 >> 8;     // never: no analysis
  else if (o.application == CB_APP_VOIP) voice_est = 115;
  else voice_est = 48;
  if (o.force_channels != CB_OPUS_AUTO && o.channels == 2) {
    o.stream_channels = o.force_channels;
  } else if (o.channels == 2) {
    i32 stereo_threshold = 30000 + ((voice_est * voice_est * (30000 - 30000)) >> 14);
    if (o.stream_channels == 2) stereo_threshold -= 1000;
    else stereo_threshold += 1000;
    o.stream_channels = (equiv_rate > stereo_threshold) ? 2 : 1;
  } else {
    o.stream_channels = o.channels;
  }
  equiv_rate = o.bitrate_bps - (40 * o.stream_channels + 20) * (o.Fs / frame_size - 50);
  if (o.application == CB_APP_LOWDELAY) {
    o.mode = CB_MODE_CELT_ONLY;
  } else if (o.user_forced_mode == CB_OPUS_AUTO) {
    const i32 mode_voice = mult16_32_q15(CB_Q15ONE - stereo_width, 64000) + mult16_32_q15(stereo_width, 36000);
    const i32 mode_music = mult16_32_q15(CB_Q15ONE - stereo_width, 16000) + mult16_32_q15(stereo_width, 16000);
    i32 threshold = mode_music + ((voice_est * voice_est * (mode_voice - mode_music)) >> 14);
    if (o.application == CB_APP_VOIP) threshold += 8000;
    if (o.prev_mode == CB_MODE_CELT_ONLY) threshold -= 4000;
    else if (o.prev_mode > 0) threshold += 4000;
    o.mode = (equiv_rate >= threshold) ? CB_MODE_CELT_ONLY : CB_MODE_SILK_ONLY;
    if (o.use_inband_fec && o.packet_loss_perc > (128 - voice_est) >> 4) o.mode = CB_MODE_SILK_ONLY;
    if (o.use_dtx && voice_est > 100) o.mode = CB_MODE_SILK_ONLY;
  } else {
    o.mode = o.user_forced_mode;
  }
  if (o.mode != CB_MODE_CELT_ONLY && frame_size < o.Fs / 100) o.mode = CB_MODE_CELT_ONLY;
  if (o.lfe) o.mode = CB_MODE_CELT_ONLY;
  if (max_data_bytes < (frame_rate > 50 ? 12000 : 8000) * frame_size / (o.Fs * 8)) o.mode = CB_MODE_CELT_ONLY;
  if (o.mode != CB_MODE_CELT_ONLY) return CB_UNIMPLEMENTED;   // SILK / hybrid are not built
  // prev_mode is 0 or CELT-only here, so no redundancy / mode transition is ever needed.

  // automatic bandwidth (opus_encoder.c:1253-1304)
  {
    const i32* voice_thr = (o.channels == 2 && o.force_channels != 1) ? ct_bw_thresh_stereo_voice : ct_bw_thresh_mono_voice;
    const i32* music_thr = (o.channels == 2 && o.force_channels != 1) ? ct_bw_thresh_stereo_music : ct_bw_thresh_mono_music;
    int bandwidth = CB_BW_FULLBAND;
    do {
      const int k = 2 * (bandwidth - CB_BW_MEDIUMBAND);
      int threshold = music_thr[k] + ((voice_est * voice_est * (voice_thr[k] - music_thr[k])) >> 14);
      const int hysteresis = music_thr[k + 1] + ((voice_est * voice_est * (voice_thr[k + 1] - music_thr[k + 1])) >> 14);
      if (!o.first) {
        if (o.bandwidth >= bandwidth) threshold -= hysteresis;
        else threshold += hysteresis;
      }
      if (equiv_rate >= threshold) break;
    } while (--bandwidth > CB_BW_NARROWBAND);
    o.bandwidth = bandwidth;
  }
  if (o.bandwidth > o.max_bandwidth) o.bandwidth = o.max_bandwidth;
  if (o.user_bandwidth != CB_OPUS_AUTO) o.bandwidth = o.user_bandwidth;
  if (o.Fs <= 24000 && o.bandwidth > CB_BW_SUPERWIDEBAND) o.bandwidth = CB_BW_SUPERWIDEBAND;
  if (o.Fs <= 16000 && o.bandwidth > CB_BW_WIDEBAND) o.bandwidth = CB_BW_WIDEBAND;
  if (o.Fs <= 12000 && o.bandwidth > CB_BW_MEDIUMBAND) o.bandwidth = CB_BW_MEDIUMBAND;
  if (o.Fs <= 8000 && o.bandwidth > CB_BW_NARROWBAND) o.bandwidth = CB_BW_NARROWBAND;
  if (o.bandwidth == CB_BW_MEDIUMBAND) o.bandwidth = CB_BW_WIDEBAND;        // CELT has no mediumband (:1356-1357)
  if (o.lfe) o.bandwidth = CB_BW_NARROWBAND;
  d.max_data_bytes = max_data_bytes;
  d.total_buffer = total_buffer;
  d.lsb_depth = lsb_depth;
  d.equiv_rate = equiv_rate;
  return 0;
}

// One frame of at most 20 ms: decisions, then the CELT encoder (:1442-1964).  Returns the packet length.
// EXT: the LFE / surround-mask terms are compiled in (cb_celt_enc.h).
template <bool EXT>
CB_FN int enc_one_frame(EncWork& w, EncScratch& sc, OpusEncState* st, OpusEncScalars& o, const i16* pcm, int frame_size,
                        u8* data, int out_data_bytes) {
  CeltEncState* celt = &st->celt;
  EncDecision d;
  d.max_data_bytes = cb_imin(1276, out_data_bytes);
  o.rangeFinal = 0;
  int early = 0;
  const int dr = enc_decide(st, o, pcm, frame_size, data, d, &early);
  if (dr < 0) return dr;
  if (dr == 1) return early;
  const int max_data_bytes = d.max_data_bytes, total_buffer = d.total_buffer, lsb_depth = d.lsb_depth;
  const i32 equiv_rate = d.equiv_rate;
  CeltEncScalars cs = celt->s;
  cs.lsb_depth = lsb_depth;
  const int curr_bandwidth = o.bandwidth;
  const int bytes_target = cb_imin(max_data_bytes, o.bitrate_bps * frame_size / (o.Fs * 8)) - 1;
  u8* payload = data + 1;
  re_init(w.enc, payload, (u32)(max_data_bytes - 1));

  // delay-compensated, DC-rejected input (opus_encoder.c:1454-1476)
  i16* pcm_buf = sc.pcm_buf;
  const int ch = o.channels;
  CB_WSYNC();
  CB_LANE_LOOP(i, total_buffer * ch) pcm_buf[i] = st->delay_buffer[(o.encoder_buffer - total_buffer) * ch + i];
  {
    // variable high-pass smoother (opus_encoder.c:1456-1467): in CELT-only mode its input is the constant minimum
    // cut-off, so it stays where init / reset put it (60 Hz); kept as the reference computes it
    const i32 hp_freq_smth1 = esk_lin2log(60) << 8;
    o.variable_HP_smth2_Q15 = esk_smlawb(o.variable_HP_smth2_Q15, hp_freq_smth1 - o.variable_HP_smth2_Q15, 983);
    const i32 cutoff_Hz = esk_log2lin(o.variable_HP_smth2_Q15 >> 8);
#if CB_DEVICE_BUILD
    // the VOIP high-pass (a biquad) stays with the warp; dc_reject goes to the CTA's serial section
    const bool voip = o.application == CB_APP_VOIP;
    if (voip) {
      enc_hp_cutoff(pcm, cutoff_Hz, &pcm_buf[total_buffer * ch], o.hp_mem, frame_size, ch, o.Fs);
      for (int k = 0; k < 4; k++) o.hp_mem[k] = __shfl_sync(0xffffffffu, o.hp_mem[k], k >> 1);
    } else {
      enc_post_dc(w, pcm, 3, &pcm_buf[total_buffer * ch], o.hp_mem, frame_size, ch, o.Fs);
    }
    CB_ENC_PHASE(w);                                          // 0: serial-section requests posted
    enc_serial_run(w.sreq - (threadIdx.x >> 5));
    CB_ENC_PHASE(w);                                          // 1: input conditioned
    if (!voip) {
      for (int k = 0; k < 4; k++) o.hp_mem[k] = w.sreq->m[k];
    }
    CB_WSYNC();
#else
    if (o.application == CB_APP_VOIP) enc_hp_cutoff(pcm, cutoff_Hz, &pcm_buf[total_buffer * ch], o.hp_mem, frame_size, ch, o.Fs);
    else enc_dc_reject(pcm, 3, &pcm_buf[total_buffer * ch], o.hp_mem, frame_size, ch, o.Fs);
#endif
  }

  // CELT set-up (opus_encoder.c:1697-1761)
  {
    int endband = 21;
    if (curr_bandwidth == CB_BW_NARROWBAND) endband = 13;
    else if (curr_bandwidth == CB_BW_MEDIUMBAND || curr_bandwidth == CB_BW_WIDEBAND) endband = 17;
    else if (curr_bandwidth == CB_BW_SUPERWIDEBAND) endband = 19;
    cs.end = endband;
    cs.stream_channels = o.stream_channels;
  }
  cs.bitrate = CB_OPUS_BITRATE_MAX;
  cs.vbr = 0;
  cs.disable_pf = o.reduced_dependency ? 1 : 0;                // CELT_SET_PREDICTION(2 or 0)
  cs.force_intra = o.reduced_dependency ? 1 : 0;
  int nb_compr_bytes;
  if (o.use_vbr) {
    cs.vbr = 1;
    cs.constrained_vbr = o.vbr_constraint;
    if (o.bitrate_bps > 500) cs.bitrate = cb_imin(o.bitrate_bps, 260000 * cs.channels);   // celt_encoder.c:2400-2406
    nb_compr_bytes = max_data_bytes - 1;
  } else {
    nb_compr_bytes = bytes_target;
  }

  // delay buffer update (opus_encoder.c:1769-1778)
  CB_WSYNC();
  if (ch * (o.encoder_buffer - (frame_size + total_buffer)) > 0) {
    const int keep = ch * (o.encoder_buffer - frame_size - total_buffer);
    enc_move_down(st->delay_buffer, ch * frame_size, keep);
    CB_LANE_LOOP(i, (frame_size + total_buffer) * ch) st->delay_buffer[keep + i] = pcm_buf[i];
  } else {
    const int off = (frame_size + total_buffer - o.encoder_buffer) * ch;
    CB_LANE_LOOP(i, o.encoder_buffer * ch) st->delay_buffer[i] = pcm_buf[off + i];
  }
  CB_WSYNC();
  o.prev_HB_gain = CB_Q15ONE;
  o.silk_stereoWidth_Q14 = cb_imin(1 << 14, 2 * cb_imax(0, equiv_rate - 30000));
  if (!(EXT && cs.has_mask) && o.channels == 2) {                 // no width reduction under a surround mask (:1792)
    if (o.hybrid_stereo_width_Q14 < (1 << 14) || o.silk_stereoWidth_Q14 < (1 << 14)) {
      i32 g1 = (i16)o.hybrid_stereo_width_Q14, g2 = (i16)o.silk_stereoWidth_Q14;
      g1 = g1 == 16384 ? CB_Q15ONE : shl16(g1, 1);
      g2 = g2 == 16384 ? CB_Q15ONE : shl16(g2, 1);
      enc_stereo_fade(pcm_buf, g1, g2, frame_size, o.Fs);
      o.hybrid_stereo_width_Q14 = o.silk_stereoWidth_Q14;
    }
  }
  o.silk_bw_switch = 0;
  nb_compr_bytes = cb_imin(max_data_bytes - 1, nb_compr_bytes);
  re_shrink(w.enc, (u32)nb_compr_bytes);
  cs.start = 0;
  CB_WSYNC();
  CB_IF_LANE0 celt->s = cs;
  CB_WSYNC();
  int ret = 0;
  if (re_tell(w.enc) <= 8 * nb_compr_bytes) {
    ret = celt_encode_frame<EXT>(w, sc, celt, pcm_buf, frame_size, nb_compr_bytes);
    if (ret < 0) return CB_INTERNAL_ERROR;
  }
  CB_WSYNC();
  data[0] = enc_gen_toc(o.mode, o.Fs / frame_size, curr_bandwidth, o.stream_channels);
  o.rangeFinal = w.enc.rng;
  o.prev_mode = o.mode;
  o.prev_channels = o.stream_channels;
  o.prev_framesize = frame_size;
  o.first = 0;
  CB_WSYNC();
  ret += 1;
  if (!o.use_vbr) {
    if (enc_packet_pad(data, ret, max_data_bytes) != CB_OK) return CB_INTERNAL_ERROR;
    ret = max_data_bytes;
  }
  return ret;
}

// opus_encode for one stream.  pcm: analysis_frame_size interleaved samples per channel; data: out_data_bytes
// of room.  Returns the packet length or a negative OPUS_* code.  All lanes run it uniformly.
// A 40 / 60 ms frame in CELT-only mode is two or three 20 ms frames in one packet (opus_encoder.c:1362-1438):
// the decisions are taken once for the long frame and forced on the short ones, each short frame is coded
// into scratch, and the repacketizer's assembly (cb_repack.h) builds the packet.
// MULTI = false leaves the long-frame path and the LFE / surround-mask terms out (the kernel instance for launches
// in which no stream needs them: the plain 20 ms path then compiles exactly as it did before they were supported).
template <bool MULTI>
CB_FN int opus_encode_stream(EncWork& w, EncScratch& sc, OpusEncState* st, const i16* pcm, int analysis_frame_size,
                             u8* data, int out_data_bytes) {
  OpusEncScalars o = st->o;
  const int frame_size = enc_frame_size_select(analysis_frame_size, o.variable_duration, o.Fs);
  if (frame_size < 0 || 400 * frame_size < o.Fs || cb_imin(1276, out_data_bytes) <= 0) return CB_BAD_ARG;
  int nb_frames = 1, bytes_per_frame = 0;
  int bak_mode = 0, bak_bandwidth = 0, bak_channels = 0;
  if (!MULTI && frame_size > o.Fs / 50) return CB_UNIMPLEMENTED;   // (the host picks the other instance for these)
  if (MULTI && frame_size > o.Fs / 50) {
    EncDecision d;
    d.max_data_bytes = cb_imin(1276, out_data_bytes);
    o.rangeFinal = 0;
    int early = 0;
    const int dr = enc_decide(st, o, pcm, frame_size, data, d, &early);
    if (dr < 0) return dr;
    if (dr == 1) {
      CB_WSYNC();
      CB_IF_LANE0 st->o = o;
      CB_WSYNC();
      return early;
    }
    nb_frames = frame_size > o.Fs / 25 ? 3 : 2;
    bytes_per_frame = cb_imin(1276, (out_data_bytes - 3) / nb_frames);
    bak_mode = o.user_forced_mode; bak_bandwidth = o.user_bandwidth; bak_channels = o.force_channels;
    o.user_forced_mode = o.mode;
    o.user_bandwidth = o.bandwidth;
    o.force_channels = o.stream_channels;
    o.prev_channels = o.stream_channels;
  }
  int ret = 0;
  unsigned long long sub_len = 0;      // lengths of the short packets, 16 bits each
  CB_NOUNROLL for (int i = 0; i < (MULTI ? nb_frames : 1); i++) {
    const bool multi = MULTI && nb_frames > 1;
    const int sub = multi ? o.Fs / 50 : frame_size;
    ret = enc_one_frame<MULTI>(w, sc, st, o, multi ? pcm + (size_t)i * o.channels * sub : pcm, sub,
                        multi ? sc.multi_tmp + i * bytes_per_frame : data, multi ? bytes_per_frame : out_data_bytes);
    if (ret < 0) return multi ? CB_INTERNAL_ERROR : ret;
    sub_len |= (unsigned long long)ret << (16 * i);
  }
  if (MULTI && nb_frames > 1) {
    // opus_repacketizer_cat on each short packet (a CBR one is a padded code-3 packet), then out_range
    CB_WSYNC();
    const i32 repacketize_len = o.use_vbr ? out_data_bytes
                                          : cb_imin(3 * o.bitrate_bps / (3 * 8 * 50 / nb_frames), out_data_bytes);
    i32 total = 0;
    CB_IF_LANE0 {
      const u8* frames[3];
      i16 flen[3];
      u8 toc = 0;
      total = 0;
      for (int i = 0; i < nb_frames && total >= 0; i++) {
        const u8* pk = sc.multi_tmp + i * bytes_per_frame;
        int t = 0, poff = 0;
        i16 sizes[CB_MAX_FRAMES];
        const int cnt = opus_parse_packet(pk, (int)((sub_len >> (16 * i)) & 0xFFFF), 0, &t, sizes, &poff, nullptr);
        if (cnt != 1 || (i > 0 && ((t ^ toc) & 0xFC))) { total = CB_INTERNAL_ERROR; break; }
        if (i == 0) toc = (u8)t;
        frames[i] = pk + poff;
        flen[i] = sizes[0];
      }
      if (total >= 0) {
        total = cb_rp_emit(frames, flen, nb_frames, toc, data, repacketize_len, 0, !o.use_vbr);
        if (total < 0) total = CB_INTERNAL_ERROR;
      }
    }
    CB_WSYNC();
    ret = cb_bcast(total, 0);
    if (ret < 0) return ret;
    o.user_forced_mode = bak_mode;
    o.user_bandwidth = bak_bandwidth;
    o.force_channels = bak_channels;
  }
  CB_WSYNC();
  CB_IF_LANE0 st->o = o;
  CB_WSYNC();
  return ret;
}
