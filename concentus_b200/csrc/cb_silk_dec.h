// cb_silk_dec.h -- the SILK decoder API layer: silk_InitDecoder / silk_decoder_set_fs /
// silk_decode_frame / silk_Decode (silk/init_decoder.c, decoder_set_fs.c, decode_frame.c, dec_API.c).
// One thread per stream; `out` receives interleaved PCM at the API rate.
#pragma once
#include "cb_silk_synth.h"

// silk_decoder_set_fs (decoder_set_fs.c:35-108)
CB_FN int sk_decoder_set_fs(SilkChannel& ps, int fs_kHz, i32 fs_API_Hz) {
  int ret = 0;
  ps.subfr_length = sk_smulbb(5, fs_kHz);
  const int frame_length = sk_smulbb(ps.nb_subfr, ps.subfr_length);
  if (ps.fs_kHz != fs_kHz || ps.fs_API_hz != fs_API_Hz) {
    ret += sk_resampler_init(ps.resampler, sk_smulbb(fs_kHz, 1000), fs_API_Hz);
    ps.fs_API_hz = fs_API_Hz;
  }
  if (ps.fs_kHz != fs_kHz || frame_length != ps.frame_length) {
    if (ps.fs_kHz != fs_kHz) {
      ps.ltp_mem_length = sk_smulbb(20, fs_kHz);
      ps.LPC_order = (fs_kHz == 8 || fs_kHz == 12) ? 10 : 16;
      ps.first_frame_after_reset = 1;
      ps.lagPrev = 100;
      ps.LastGainIndex = 10;
      ps.prevSignalType = SK_TYPE_NO_VOICE;
      for (int i = 0; i < SK_MAX_FRAME + 2 * SK_MAX_SUBFR; i++) ps.outBuf[i] = 0;
      for (int i = 0; i < SK_MAX_ORDER; i++) ps.sLPC_Q14_buf[i] = 0;
    }
    ps.fs_kHz = fs_kHz;
    ps.frame_length = frame_length;
  }
  return ret;
}

// silk_decode_frame (decode_frame.c:38-129)
CB_FN void sk_decode_frame(SilkChannel& ps, RangeDec& rd, i16* pOut, int* pN, int lostFlag, int condCoding, SilkWork& w) {
  SilkDecControl ctrl;
  const int L = ps.frame_length;
  ctrl.LTP_scale_Q14 = 0;
  if (lostFlag == SK_FLAG_DECODE_NORMAL || (lostFlag == SK_FLAG_DECODE_LBRR && ps.LBRR_flags[ps.nFramesDecoded] == 1)) {
    sk_decode_indices(ps, rd, ps.nFramesDecoded, lostFlag, condCoding);
    sk_decode_pulses(rd, w.pulses, ps.indices.signalType, ps.indices.quantOffsetType, ps.frame_length);
    sk_decode_parameters(ps, ctrl, condCoding);
    sk_decode_core(ps, ctrl, pOut, w);
    sk_plc(ps, ctrl, pOut, 0, w);
    ps.lossCnt = 0;
    ps.prevSignalType = ps.indices.signalType;
    ps.first_frame_after_reset = 0;
  } else {
    sk_plc(ps, ctrl, pOut, 1, w);
  }
  const int mv_len = ps.ltp_mem_length - ps.frame_length;
  sk_copy_i16(ps.outBuf, &ps.outBuf[ps.frame_length], mv_len);
  sk_copy_i16(&ps.outBuf[mv_len], pOut, ps.frame_length);
  sk_cng(ps, ctrl, pOut, L, w);
  sk_plc_glue_frames(ps, pOut, L);
  ps.lagPrev = ctrl.pitchL[ps.nb_subfr - 1];
  *pN = L;
}

// per-thread buffers of one silk_Decode call
struct SilkCallWork {
  SilkWork w;
  SilkRsWork rs;
  i16 tmp[2][SK_MAX_FRAME + 2];     // samplesOut1_tmp: channel n at tmp[n], 2 samples of history in front
  alignas(16) i16 obuf[2 * 960];    // resampled output of one SILK frame (<= 20 ms at 48 kHz, interleaved)
};

// silk_Decode (dec_API.c:79-380).  API_sampleRate / nChannelsAPI come from the Opus decoder; the
// per-packet control values (nChannelsInternal, internalSampleRate, payloadSize_ms) live in `d`.
// `out`: interleaved by nChannelsAPI.  Returns 0 or a negative SILK error.
CB_FN int sk_decode(SilkDecState& d, i32 API_sampleRate, int nChannelsAPI, int lostFlag, int newPacketFlag,
                    RangeDec& rd, i16* out, i32* nSamplesOut, SilkCallWork& cw) {
  int decode_only_middle = 0, ret = 0;
  i32 nSamplesOutDec = 0;
  i32 MS_pred_Q13[2] = {0, 0};
  SilkChannel* cs = d.ch;
  const int nInt = d.ctl_nChannelsInternal;
  if (newPacketFlag)
    for (int n = 0; n < nInt; n++) cs[n].nFramesDecoded = 0;
  if (nInt > d.nChannelsInternal) sk_init_channel(cs[1]);
  const int stereo_to_mono = nInt == 1 && d.nChannelsInternal == 2 && (d.ctl_internalSampleRate == 1000 * cs[0].fs_kHz);
  if (cs[0].nFramesDecoded == 0) {
    for (int n = 0; n < nInt; n++) {
      const int ms = d.ctl_payloadSize_ms;
      if (ms == 0 || ms == 10) { cs[n].nFramesPerPacket = 1; cs[n].nb_subfr = 2; }
      else if (ms == 20) { cs[n].nFramesPerPacket = 1; cs[n].nb_subfr = 4; }
      else if (ms == 40) { cs[n].nFramesPerPacket = 2; cs[n].nb_subfr = 4; }
      else if (ms == 60) { cs[n].nFramesPerPacket = 3; cs[n].nb_subfr = 4; }
      else return -203;                                           // SILK_DEC_INVALID_FRAME_SIZE
      const int fs_kHz_dec = (d.ctl_internalSampleRate >> 10) + 1;
      if (fs_kHz_dec != 8 && fs_kHz_dec != 12 && fs_kHz_dec != 16) return -200;   // SILK_DEC_INVALID_SAMPLING_FREQUENCY
      ret += sk_decoder_set_fs(cs[n], fs_kHz_dec, API_sampleRate);
    }
  }
  if (nChannelsAPI == 2 && nInt == 2 && (d.nChannelsAPI == 1 || d.nChannelsInternal == 1)) {
    d.sStereo.pred_prev_Q13[0] = d.sStereo.pred_prev_Q13[1] = 0;
    d.sStereo.sSide[0] = d.sStereo.sSide[1] = 0;
    cs[1].resampler = cs[0].resampler;
  }
  d.nChannelsAPI = nChannelsAPI;
  d.nChannelsInternal = nInt;
  if (API_sampleRate > 48000 || API_sampleRate < 8000) return -200;
  if (lostFlag != SK_FLAG_PACKET_LOST && cs[0].nFramesDecoded == 0) {
    for (int n = 0; n < nInt; n++) {
      for (int i = 0; i < cs[n].nFramesPerPacket; i++) cs[n].VAD_flags[i] = rd_bit_logp_g(rd, 1);
      cs[n].LBRR_flag = rd_bit_logp_g(rd, 1);
    }
    for (int n = 0; n < nInt; n++) {
      cs[n].LBRR_flags[0] = cs[n].LBRR_flags[1] = cs[n].LBRR_flags[2] = 0;
      if (cs[n].LBRR_flag) {
        if (cs[n].nFramesPerPacket == 1) {
          cs[n].LBRR_flags[0] = 1;
        } else {
          const i32 sym = rd_icdf_g(rd, cs[n].nFramesPerPacket == 2 ? st_LBRR_flags_2_iCDF : st_LBRR_flags_3_iCDF, 8) + 1;
          for (int i = 0; i < cs[n].nFramesPerPacket; i++) cs[n].LBRR_flags[i] = (sym >> i) & 1;
        }
      }
    }
    if (lostFlag == SK_FLAG_DECODE_NORMAL) {
      for (int i = 0; i < cs[0].nFramesPerPacket; i++) {
        for (int n = 0; n < nInt; n++) {
          if (cs[n].LBRR_flags[i]) {
            if (nInt == 2 && n == 0) {
              sk_stereo_decode_pred(rd, MS_pred_Q13);
              if (cs[1].LBRR_flags[i] == 0) decode_only_middle = rd_icdf_g(rd, st_stereo_only_code_mid_iCDF, 8);
            }
            const int condCoding = (i > 0 && cs[n].LBRR_flags[i - 1]) ? SK_CODE_CONDITIONALLY : SK_CODE_INDEPENDENTLY;
            sk_decode_indices(cs[n], rd, i, 1, condCoding);
            sk_decode_pulses(rd, cw.w.pulses, cs[n].indices.signalType, cs[n].indices.quantOffsetType, cs[n].frame_length);
          }
        }
      }
    }
  }
  if (nInt == 2) {
    if (lostFlag == SK_FLAG_DECODE_NORMAL ||
        (lostFlag == SK_FLAG_DECODE_LBRR && cs[0].LBRR_flags[cs[0].nFramesDecoded] == 1)) {
      sk_stereo_decode_pred(rd, MS_pred_Q13);
      if ((lostFlag == SK_FLAG_DECODE_NORMAL && cs[1].VAD_flags[cs[0].nFramesDecoded] == 0) ||
          (lostFlag == SK_FLAG_DECODE_LBRR && cs[1].LBRR_flags[cs[0].nFramesDecoded] == 0))
        decode_only_middle = rd_icdf_g(rd, st_stereo_only_code_mid_iCDF, 8);
      else
        decode_only_middle = 0;
    } else {
      MS_pred_Q13[0] = d.sStereo.pred_prev_Q13[0];
      MS_pred_Q13[1] = d.sStereo.pred_prev_Q13[1];
    }
  }
  if (nInt == 2 && decode_only_middle == 0 && d.prev_decode_only_middle == 1) {
    for (int i = 0; i < SK_MAX_FRAME + 2 * SK_MAX_SUBFR; i++) cs[1].outBuf[i] = 0;
    for (int i = 0; i < SK_MAX_ORDER; i++) cs[1].sLPC_Q14_buf[i] = 0;
    cs[1].lagPrev = 100;
    cs[1].LastGainIndex = 10;
    cs[1].prevSignalType = SK_TYPE_NO_VOICE;
    cs[1].first_frame_after_reset = 1;
  }
  int has_side;
  if (lostFlag == SK_FLAG_DECODE_NORMAL) has_side = !decode_only_middle;
  else has_side = !d.prev_decode_only_middle ||
                  (nInt == 2 && lostFlag == SK_FLAG_DECODE_LBRR && cs[1].LBRR_flags[cs[1].nFramesDecoded] == 1);
  for (int n = 0; n < nInt; n++) {
    if (n == 0 || has_side) {
      const int FrameIndex = cs[0].nFramesDecoded - n;
      int condCoding;
      if (FrameIndex <= 0) condCoding = SK_CODE_INDEPENDENTLY;
      else if (lostFlag == SK_FLAG_DECODE_LBRR) condCoding = cs[n].LBRR_flags[FrameIndex - 1] ? SK_CODE_CONDITIONALLY : SK_CODE_INDEPENDENTLY;
      else if (n > 0 && d.prev_decode_only_middle) condCoding = SK_CODE_INDEPENDENTLY_NO_LTP_SCALING;
      else condCoding = SK_CODE_CONDITIONALLY;
      int nn = 0;
      sk_decode_frame(cs[n], rd, &cw.tmp[n][2], &nn, lostFlag, condCoding, cw.w);
      nSamplesOutDec = nn;
    } else {
      for (int i = 0; i < nSamplesOutDec; i++) cw.tmp[n][2 + i] = 0;
    }
    cs[n].nFramesDecoded++;
  }
  if (nChannelsAPI == 2 && nInt == 2) {
    sk_stereo_ms_to_lr(d.sStereo, cw.tmp[0], cw.tmp[1], MS_pred_Q13, cs[0].fs_kHz, nSamplesOutDec);
  } else {
    cw.tmp[0][0] = d.sStereo.sMid[0];
    cw.tmp[0][1] = d.sStereo.sMid[1];
    d.sStereo.sMid[0] = cw.tmp[0][nSamplesOutDec];
    d.sStereo.sMid[1] = cw.tmp[0][nSamplesOutDec + 1];
  }
  *nSamplesOut = (nSamplesOutDec * API_sampleRate) / sk_smulbb(cs[0].fs_kHz, 1000);
  // the resamplers write sample by sample into a private buffer; the PCM row gets it in 16-byte pieces
  const int nres = nChannelsAPI < nInt ? nChannelsAPI : nInt;
  i16* const ob = cw.obuf;
  for (int n = 0; n < nres; n++)
    sk_resampler(cs[n].resampler, ob + n, nChannelsAPI, &cw.tmp[n][1], nSamplesOutDec, cw.rs);
  if (nChannelsAPI == 2 && nInt == 1) {
    if (stereo_to_mono) {
      sk_resampler(cs[1].resampler, ob + 1, 2, &cw.tmp[0][1], nSamplesOutDec, cw.rs);
    } else {
      for (int i = 0; i < *nSamplesOut; i++) ob[1 + 2 * i] = ob[2 * i];
    }
  }
  sk_copy_i16(out, ob, *nSamplesOut * nChannelsAPI);
  if (cs[0].prevSignalType == SK_TYPE_VOICED) {
    const int mult = cs[0].fs_kHz == 8 ? 6 : cs[0].fs_kHz == 12 ? 4 : 3;
    d.ctl_prevPitchLag = cs[0].lagPrev * mult;
  } else {
    d.ctl_prevPitchLag = 0;
  }
  if (lostFlag == SK_FLAG_PACKET_LOST) {
    for (int i = 0; i < d.nChannelsInternal; i++) d.ch[i].LastGainIndex = 10;
  } else {
    d.prev_decode_only_middle = decode_only_middle;
  }
  return ret;
}
