// cb_celt_enc.h -- one CELT frame, encoder side: celt_encode_with_ec
// (opus-fix/celt/celt_encoder.c:1398-2274) for the static 48 kHz / 960 mode, FIXED_POINT,
// DISABLE_FLOAT_API (no tonality analysis), start band 0.  EXT = true adds what only the streams of a surround
// multistream encoder use: the LFE restrictions (OPUS_SET_LFE) and the surround masking terms (OPUS_SET_ENERGY_MASK).
// One warp encodes one stream's frame; the range coder lives in w.enc and has been initialised
// by the Opus layer over the packet's payload bytes.
#pragma once
#include "cb_common.h"
#include "cb_enc_state.h"
#include "cb_enc_analysis.h"
#include "cb_enc_mdct.h"
#include "cb_enc_decisions.h"
#include "cb_enc_quant.h"

// celt_maxabs16 over an interleaved PCM run (lane-parallel)
CB_FN i32 enc_maxabs16(const i16* x, int len) {
  i32 mx = 0, mn = 0;
  CB_LANE_LOOP(i, len) { i32 v = x[i]; mx = max32(mx, v); mn = min32(mn, v); }
  return max32(cb_wmax(mx), neg32(cb_wmin(mn)));
}

// Returns the number of payload bytes, or a negative OPUS_* code.
template <bool EXT>
CB_FN int celt_encode_frame(EncWork& w, EncScratch& sc, CeltEncState* st, const i16* pcm, int frame_size,
                            int nbCompressedBytes) {
  RangeEnc& enc = w.enc;
  CeltEncScalars s = st->s;
  const int CC = s.channels, C = s.stream_channels;
  const int start = s.start, end = s.end;
  const int overlap = CB_OVERLAP;
  const int lfe = EXT ? s.lfe : 0, has_mask = EXT ? s.has_mask : 0;
  i32 tf_estimate = 0;
  if (nbCompressedBytes < 2) return CB_BAD_ARG;
  frame_size *= s.upsample;
  int LM;
  for (LM = 0; LM <= CB_MAX_LM; LM++)
    if (CB_SHORT_MDCT << LM == frame_size) break;
  if (LM > CB_MAX_LM) return CB_BAD_ARG;
  const int M = 1 << LM, N = M * CB_SHORT_MDCT;
  i32 tell = re_tell(enc);
  const int nbFilledBytes = (tell + 4) >> 3;
  nbCompressedBytes = cb_imin(nbCompressedBytes, 1275);
  int nbAvailableBytes = nbCompressedBytes - nbFilledBytes;
  i32 vbr_rate;
  int effectiveBytes;
  if (s.vbr && s.bitrate != CB_OPUS_BITRATE_MAX) {
    const i32 den = 48000 >> CB_BITRES;
    vbr_rate = (s.bitrate * frame_size + (den >> 1)) / den;
    effectiveBytes = vbr_rate >> (3 + CB_BITRES);
  } else {
    vbr_rate = 0;
    i32 tmp = s.bitrate * frame_size;
    if (tell > 1) tmp += tell;
    if (s.bitrate != CB_OPUS_BITRATE_MAX)
      nbCompressedBytes = cb_imax(2, cb_imin(nbCompressedBytes, (tmp + 4 * 48000) / (8 * 48000)));
    effectiveBytes = nbCompressedBytes;
  }
  i32 equiv_rate = 510000;
  if (s.bitrate != CB_OPUS_BITRATE_MAX) equiv_rate = s.bitrate - (40 * C + 20) * ((400 >> LM) - 50);
  if (vbr_rate > 0 && s.constrained_vbr) {
    const i32 vbr_bound = vbr_rate;
    const i32 max_allowed = cb_imin(cb_imax(tell == 1 ? 2 : 0, (vbr_rate + vbr_bound - s.vbr_reservoir) >> (CB_BITRES + 3)),
                                    nbAvailableBytes);
    if (max_allowed < nbAvailableBytes) {
      nbCompressedBytes = nbFilledBytes + max_allowed;
      nbAvailableBytes = max_allowed;
      re_shrink(enc, (u32)nbCompressedBytes);
    }
  }
  i32 total_bits = nbCompressedBytes * 8;
  const int effEnd = end;                                     // effEBands = 21 in this mode

  CB_WSYNC();
  i32 sample_max = max32(s.overlap_max, enc_maxabs16(pcm, C * (N - overlap) / s.upsample));
  s.overlap_max = enc_maxabs16(pcm + C * (N - overlap) / s.upsample, C * overlap / s.upsample);
  sample_max = max32(sample_max, s.overlap_max);
  int silence = sample_max == 0;
  if (tell == 1) re_bit_logp(enc, silence, 15);
  else silence = 0;
  if (silence) {
    if (vbr_rate > 0) {
      effectiveBytes = nbCompressedBytes = cb_imin(nbCompressedBytes, nbFilledBytes + 2);
      total_bits = nbCompressedBytes * 8;
      nbAvailableBytes = 2;
      re_shrink(enc, (u32)nbCompressedBytes);
    }
    tell = nbCompressedBytes * 8;
    enc.nbits_total += tell - re_tell(enc);
  }
  for (int c = 0; c < CC; c++)
    enc_preemphasis(pcm + c, sc.in[c] + overlap, N, CC, s.upsample, &s.preemph_memE[c]);

  CB_ENC_PHASE(w);                                            // 2: pre-emphasis done
  // ---- pitch pre-filter ----
  int pitch_index = CB_COMBFILTER_MINPERIOD, pitch_change = 0, pf_on, prefilter_tapset;
  i32 gain1 = 0;
  {
    int qg;
    const int enabled = ((lfe && nbAvailableBytes > 3) || nbAvailableBytes > 12 * C) && start == 0 && !silence && !s.disable_pf && s.complexity >= 5 &&
                        !(s.consec_transient && LM != 3 && s.variable_duration == 5010);
    prefilter_tapset = s.tapset_decision;
    pf_on = enc_run_prefilter(w, s, st, sc, CC, N, prefilter_tapset, &pitch_index, &gain1, &qg, enabled, nbAvailableBytes);
    if ((gain1 > 13107 || s.prefilter_gain > 13107) &&
        ((double)pitch_index > 1.26 * (double)s.prefilter_period || (double)pitch_index < .79 * (double)s.prefilter_period))
      pitch_change = 1;
    if (pf_on == 0) {
      if (start == 0 && tell + 16 <= total_bits) re_bit_logp(enc, 0, 1);
    } else {
      re_bit_logp(enc, 1, 1);
      pitch_index += 1;
      const int octave = cb_ilog((u32)pitch_index) - 5;
      re_uint(enc, (u32)octave, 6);
      re_bits(enc, (u32)(pitch_index - (16 << octave)), (u32)(4 + octave));
      pitch_index -= 1;
      re_bits(enc, (u32)qg, 3);
      re_icdf(enc, prefilter_tapset, ct_tapset_icdf, 2);
    }
  }

  // ---- transient decision, MDCT, band energies ----
  int isTransient = 0, shortBlocks = 0, tf_chan = 0, transient_got_disabled = 0;
  const bool want_ta = s.complexity >= 1 && !lfe;
#if CB_DEVICE_BUILD
  // the detector's recurrences run in the CTA's serial section (cb_enc_state.h: EncSerialReq), its metric here
  CB_WSYNC();
  if (want_ta && CB_LANE() == 0) {
    EncSerialReq& r = *w.sreq;
    r.in = sc.in[0]; r.out = w.ta_tmp[0]; r.len = N + overlap; r.channels = CC;
    r.stride_in = CB_MAXN + CB_OVERLAP; r.stride_out = CB_MAXN + CB_OVERLAP;
    r.kind = CB_ES_TRANSIENT;
  }
  CB_ENC_PHASE(w);                                            // 3: pre-filter done, requests posted
  enc_serial_run(w.sreq - (threadIdx.x >> 5));
  CB_ENC_PHASE(w);                                            // 3b: transient recurrences done
  if (want_ta) {
    i32 mean_c[2], maxE_c[2];
    for (int c = 0; c < 2; c++) { mean_c[c] = w.sreq->m[c]; maxE_c[c] = w.sreq->m[2 + c]; }
    isTransient = enc_transient_finish(N + overlap, CC, mean_c, maxE_c, &tf_estimate, &tf_chan, w.ta_tmp[0], CB_MAXN + CB_OVERLAP);
  }
#else
  CB_ENC_PHASE(w);                                            // 3: pre-filter done
  if (want_ta)
    isTransient = enc_transient_analysis(sc.in[0], CB_MAXN + CB_OVERLAP, N + overlap, CC, &tf_estimate, &tf_chan, w.ta_tmp[0],
                                         CB_MAXN + CB_OVERLAP);
#endif
  if (LM > 0 && re_tell(enc) + 3 <= total_bits) {
    if (isTransient) shortBlocks = M;
  } else {
    isTransient = 0;
    transient_got_disabled = 1;
  }
  CB_ENC_PHASE(w);                                            // 4: transient analysis done
  const int secondMdct = shortBlocks && s.complexity >= 8;
  if (secondMdct) {
    enc_compute_mdcts(0, sc.in[0], sc.in[1], sc.freq[0], sc.freq[1], sc.f2, C, CC, LM, s.upsample);
    enc_band_energies(sc.freq[0], sc.freq[1], w.bandE, effEnd, C, LM);
    enc_amp2log2(effEnd, end, w.bandE, w.bandLogE2, C);
    CB_LANE_LOOP(i, C * CB_NBANDS) w.bandLogE2[i] = (i16)(w.bandLogE2[i] + LM * 512);
    CB_WSYNC();
  }
  enc_compute_mdcts(shortBlocks, sc.in[0], sc.in[1], sc.freq[0], sc.freq[1], sc.f2, C, CC, LM, s.upsample);
  if (CC == 2 && C == 1) tf_chan = 0;
  enc_band_energies(sc.freq[0], sc.freq[1], w.bandE, effEnd, C, LM);
  if (lfe) {                                                  // everything of interest is in the first bands (:1716-1723)
    CB_NOUNROLL for (int i = 2; i < end; i++) {
      const i32 e = max32(min32(w.bandE[i], mult16_32_q15(3, w.bandE[0])), 1);
      CB_WSYNC();
      w.bandE[i] = e;
    }
    CB_WSYNC();
  }
  enc_amp2log2(effEnd, end, w.bandE, w.bandLogE, C);

  CB_ENC_PHASE(w);                                            // 5: MDCT + band energies done
  // how much the other surround channels mask this one (celt_encoder.c:1726-1801)
  i32 surround_masking = 0, surround_trim = 0;
  if (EXT) {
    CB_WSYNC();
    CB_NOUNROLL for (int i = 0; i < CB_NBANDS; i++) w.surround_dynalloc[i] = 0;
    if (start == 0 && has_mask && !lfe) {
      const i16* em = st->energy_mask;
      const int mask_end = cb_imax(2, s.lastCodedBands);
      i32 mask_avg = 0, diff = 0;
      int count = 0;
      for (int c = 0; c < C; c++)
        CB_NOUNROLL for (int i = 0; i < mask_end; i++) {
          i32 mask = max32(min32(em[CB_NBANDS * c + i], 256), -2048);
          if (mask > 0) mask >>= 1;
          mask_avg += mult16_16(mask, ebands(i + 1) - ebands(i));
          count += ebands(i + 1) - ebands(i);
          diff += mult16_16(mask, 1 + 2 * i - mask_end);
        }
      mask_avg = (i16)(mask_avg / (i32)(i16)count);              // DIV32_16
      mask_avg += 205;                                           // QCONST16(.2f, DB_SHIFT)
      diff = diff * 6 / (C * (mask_end - 1) * (mask_end + 1) * mask_end);
      diff = diff >> 1;
      diff = max32(min32(diff, 32), -32);                        // QCONST32(.031f, DB_SHIFT)
      int midband = 0;
      while (ebands(midband + 1) < ebands(mask_end) / 2) midband++;
      int count_dynalloc = 0;
      CB_NOUNROLL for (int i = 0; i < mask_end; i++) {
        const i32 lin = mask_avg + diff * (i - midband);
        i32 unmask = C == 2 ? max32(em[i], em[CB_NBANDS + i]) : em[i];
        unmask = min32(unmask, 0);
        unmask = (i16)(unmask - lin);
        if (unmask > 256) {
          w.surround_dynalloc[i] = (i16)(unmask - 256);
          count_dynalloc++;
        }
      }
      CB_WSYNC();
      if (count_dynalloc >= 3) {
        // dynalloc wanted in many bands: the masking rate was probably too low
        mask_avg += 256;
        if (mask_avg > 0) {
          mask_avg = 0;
          diff = 0;
          CB_NOUNROLL for (int i = 0; i < mask_end; i++) w.surround_dynalloc[i] = 0;
        } else {
          CB_NOUNROLL for (int i = 0; i < mask_end; i++) {
            const i32 v = max32(0, w.surround_dynalloc[i] - 256);
            CB_WSYNC();
            w.surround_dynalloc[i] = (i16)v;
          }
        }
      }
      mask_avg += 205;
      surround_trim = (i16)(64 * diff);
      surround_masking = (i16)mask_avg;
    }
    CB_WSYNC();
  }
  // temporal VBR (celt_encoder.c:1795-1812), not for LFE
  i32 temporal_vbr = 0;
  if (!lfe) {
    i32 follow = -10240;
    i32 frame_avg = 0;
    const i32 offset = shortBlocks ? LM * 512 : 0;
    CB_NOUNROLL for (int i = start; i < end; i++) {
      follow = (i16)max32(follow - 1024, w.bandLogE[i] - offset);
      if (C == 2) follow = (i16)max32(follow, w.bandLogE[i + CB_NBANDS] - offset);
      frame_avg += follow;
    }
    frame_avg /= (end - start);
    temporal_vbr = (i16)sub16(frame_avg, s.spec_avg);
    temporal_vbr = (i16)min32(3072, max32(-1536, temporal_vbr));
    s.spec_avg = (i16)(s.spec_avg + mult16_16_q15(655, temporal_vbr));
  }
  if (!secondMdct) {
    CB_LANE_LOOP(i, C * CB_NBANDS) w.bandLogE2[i] = w.bandLogE[i];
    CB_WSYNC();
  }
  CB_LANE_LOOP(i, 2 * CB_NBANDS) w.oldE[i] = st->oldBandE[i];
  CB_WSYNC();
  if (LM > 0 && re_tell(enc) + 3 <= total_bits && !isTransient && s.complexity >= 5 && !lfe) {
    if (enc_patch_transient(w.bandLogE, w.oldE, start, end, C)) {
      isTransient = 1;
      shortBlocks = M;
      enc_compute_mdcts(shortBlocks, sc.in[0], sc.in[1], sc.freq[0], sc.freq[1], sc.f2, C, CC, LM, s.upsample);
      enc_band_energies(sc.freq[0], sc.freq[1], w.bandE, effEnd, C, LM);
      enc_amp2log2(effEnd, end, w.bandE, w.bandLogE, C);
      CB_LANE_LOOP(i, C * CB_NBANDS) w.bandLogE2[i] = (i16)(w.bandLogE2[i] + LM * 512);
      CB_WSYNC();
      tf_estimate = 3277;
    }
  }
  if (LM > 0 && re_tell(enc) + 3 <= total_bits) re_bit_logp(enc, isTransient, 3);
  enc_normalise_bands(sc.freq[0], sc.freq[1], w.X, N, w.bandE, effEnd, C, M);

  CB_ENC_PHASE(w);                                            // 6: normalised spectrum ready
  // ---- time/frequency resolution, coarse energy, spreading ----
  int tf_select, tf_sum;
  if (effectiveBytes >= 15 * C && start == 0 && s.complexity >= 2 && !lfe) {
    int lambda;
    if (effectiveBytes < 40) lambda = 12;
    else if (effectiveBytes < 60) lambda = 6;
    else if (effectiveBytes < 100) lambda = 4;
    else lambda = 3;
    lambda *= 2;
    tf_select = enc_tf_analysis(w, effEnd, isTransient, lambda, N, LM, &tf_sum, tf_estimate, tf_chan);
    for (int i = effEnd; i < end; i++) w.tf_res[i] = w.tf_res[effEnd - 1];
  } else {
    tf_sum = 0;
    CB_WSYNC();
    for (int i = 0; i < end; i++) w.tf_res[i] = isTransient;
    CB_WSYNC();
    tf_select = 0;
  }
  (void)tf_sum;
  enc_coarse_energy(w, sc, enc, start, end, effEnd, w.oldE, (u32)total_bits, C, LM, nbAvailableBytes, s.force_intra,
                    &s.delayedIntra, s.complexity >= 4, s.loss_rate, lfe);
  enc_tf_encode(w, enc, start, end, isTransient, LM, tf_select);
  if (re_tell(enc) + 4 <= total_bits) {
    if (lfe) {
      s.tapset_decision = 0;
      s.spread_decision = CB_SPREAD_NORMAL;
    } else if (shortBlocks || s.complexity < 3 || nbAvailableBytes < 10 * C || start != 0) {
      s.spread_decision = s.complexity == 0 ? CB_SPREAD_NONE : CB_SPREAD_NORMAL;
    } else {
      s.spread_decision = enc_spreading_decision(w.X, &s.tonal_average, s.spread_decision, &s.hf_average,
                                                 &s.tapset_decision, pf_on && !shortBlocks, effEnd, C, M);
    }
    re_icdf(enc, s.spread_decision, ct_spread_icdf, 5);
  }

  CB_ENC_PHASE(w);                                            // 7: tf + coarse energy coded
  // ---- dynamic allocation boosts ----
  i32 tot_boost;
  const i32 maxDepth = enc_dynalloc_analysis(w, start, end, C, s.lsb_depth, isTransient, s.vbr, s.constrained_vbr, LM,
                                             effectiveBytes, &tot_boost, lfe, EXT ? w.surround_dynalloc : nullptr);
  if (lfe) {
    CB_WSYNC();
    w.offsets[0] = cb_imin(8, effectiveBytes / 3);
    CB_WSYNC();
  }
  CB_NOUNROLL for (int i = 0; i < CB_NBANDS; i++) {              // init_caps (celt.c:255-264)
    const int Nb = (ebands(i + 1) - ebands(i)) << LM;
    w.cap[i] = (ct_cache_caps50[CB_NBANDS * (2 * LM + C - 1) + i] + 64) * C * Nb >> 2;
  }
  CB_WSYNC();
  int dynalloc_logp = 6;
  total_bits <<= CB_BITRES;
  i32 total_boost = 0;
  tell = (i32)re_tell_frac(enc);
  CB_NOUNROLL for (int i = start; i < end; i++) {
    const int width = C * (ebands(i + 1) - ebands(i)) << LM;
    const int quanta = cb_imin(width << CB_BITRES, cb_imax(6 << CB_BITRES, width));
    int dynalloc_loop_logp = dynalloc_logp;
    int boost = 0;
    int j;
    const int off = w.offsets[i];
    CB_NOUNROLL for (j = 0; tell + (dynalloc_loop_logp << CB_BITRES) < total_bits - total_boost && boost < w.cap[i]; j++) {
      const int flag = j < off;
      re_bit_logp(enc, flag, (u32)dynalloc_loop_logp);
      tell = (i32)re_tell_frac(enc);
      if (!flag) break;
      boost += quanta;
      total_boost += quanta;
      dynalloc_loop_logp = 1;
    }
    if (j) dynalloc_logp = cb_imax(2, dynalloc_logp - 1);
    CB_WSYNC();
    w.offsets[i] = boost;
  }
  CB_WSYNC();
  int dual_stereo = 0;
  if (C == 2) {
    if (LM != 0) dual_stereo = enc_stereo_analysis(w.X, LM, N);
    s.intensity = enc_hysteresis((i16)(equiv_rate / 1000), ct_intensity_thresholds, ct_intensity_histeresis, 21, s.intensity);
    s.intensity = cb_imin(end, cb_imax(start, s.intensity));
  }
  int alloc_trim = 5;
  if (tell + (6 << CB_BITRES) <= total_bits - total_boost) {
    if (!lfe)
      alloc_trim = enc_alloc_trim_analysis(w.X, w.bandLogE, end, LM, C, N, &s.stereo_saving, tf_estimate, s.intensity,
                                           surround_trim);
    re_icdf(enc, alloc_trim, ct_trim_icdf, 7);
    tell = (i32)re_tell_frac(enc);
  }

  // ---- variable bitrate (celt_encoder.c:2004-2093) ----
  if (vbr_rate > 0) {
    const int lm_diff = CB_MAX_LM - LM;
    nbCompressedBytes = cb_imin(nbCompressedBytes, 1275 >> (3 - LM));
    i32 base_target = vbr_rate - ((40 * C + 20) << CB_BITRES);
    if (s.constrained_vbr) base_target += (s.vbr_offset >> lm_diff);
    i32 target = enc_compute_vbr(base_target, LM, equiv_rate, s.lastCodedBands, C, s.intensity, s.constrained_vbr,
                                 s.stereo_saving, tot_boost, tf_estimate, maxDepth, s.variable_duration, temporal_vbr,
                                 lfe, has_mask, surround_masking);
    (void)pitch_change;
    target = target + tell;
    const i32 min_allowed = ((tell + total_boost + (1 << (CB_BITRES + 3)) - 1) >> (CB_BITRES + 3)) + 2 - nbFilledBytes;
    nbAvailableBytes = (target + (1 << (CB_BITRES + 2))) >> (CB_BITRES + 3);
    nbAvailableBytes = cb_imax(min_allowed, nbAvailableBytes);
    nbAvailableBytes = cb_imin(nbCompressedBytes, nbAvailableBytes + nbFilledBytes) - nbFilledBytes;
    i32 delta = target - vbr_rate;
    target = nbAvailableBytes << (CB_BITRES + 3);
    if (silence) {
      nbAvailableBytes = 2;
      target = 2 * 8 << CB_BITRES;
      delta = 0;
    }
    i32 alpha;
    if (s.vbr_count < 970) {
      s.vbr_count++;
      alpha = (i16)cb_rcp(shl32(s.vbr_count + 20, 16));
    } else {
      alpha = 33;                                               // QCONST16(.001f,15)
    }
    if (s.constrained_vbr) s.vbr_reservoir += target - vbr_rate;
    if (s.constrained_vbr) {
      s.vbr_drift += mult16_32_q15(alpha, (delta * (1 << lm_diff)) - s.vbr_offset - s.vbr_drift);
      s.vbr_offset = -s.vbr_drift;
    }
    if (s.constrained_vbr && s.vbr_reservoir < 0) {
      const int adjust = (-s.vbr_reservoir) / (8 << CB_BITRES);
      nbAvailableBytes += silence ? 0 : adjust;
      s.vbr_reservoir = 0;
    }
    nbCompressedBytes = cb_imin(nbCompressedBytes, nbAvailableBytes + nbFilledBytes);
    re_shrink(enc, (u32)nbCompressedBytes);
  }

  CB_ENC_PHASE(w);                                            // 8: side decisions coded
  // ---- bit allocation, fine energy, residual ----
  i32 bits = (((i32)nbCompressedBytes * 8) << CB_BITRES) - (i32)re_tell_frac(enc) - 1;
  const int anti_collapse_rsv = isTransient && LM >= 2 && bits >= ((LM + 2) << CB_BITRES) ? (1 << CB_BITRES) : 0;
  bits -= anti_collapse_rsv;
  const int signalBandwidth = lfe ? 1 : end - 1;
  i32 balance;
  int intensity = s.intensity;
  AllocEncPolicy ac{enc, s.lastCodedBands, signalBandwidth};
  const int codedBands = cb_allocation(w, ac, start, end, alloc_trim, intensity, dual_stereo, bits, balance, C, LM);
  s.intensity = intensity;
  if (s.lastCodedBands) s.lastCodedBands = cb_imin(s.lastCodedBands + 1, cb_imax(s.lastCodedBands - 1, codedBands));
  else s.lastCodedBands = codedBands;
  enc_fine_energy(w, enc, start, end, w.oldE, C);
  enc_all_bands(w, enc, start, end, C, N, shortBlocks, s.spread_decision, dual_stereo, s.intensity,
                nbCompressedBytes * (8 << CB_BITRES) - anti_collapse_rsv, balance, LM, codedBands);
  if (anti_collapse_rsv > 0) {
    const int anti_collapse_on = s.consec_transient < 2;
    re_bits(enc, (u32)anti_collapse_on, 1);
  }
  enc_energy_finalise(w, enc, start, end, w.oldE, nbCompressedBytes * 8 - re_tell(enc), C);

  CB_ENC_PHASE(w);                                            // 9: residual coded
  // ---- state update (celt_encoder.c:2156-2241) ----
  CB_WSYNC();
  CB_LANE_LOOP(i, 2 * CB_NBANDS) {
    const int c = i / CB_NBANDS, b = i % CB_NBANDS;
    if (c >= CC) continue;
    i32 e = (c < C) ? w.oldE[i] : w.oldE[b];                   // CC==2 && C==1: channel 1 mirrors channel 0
    if (silence && c < C) e = -28672;
    if (silence && c >= C) e = -28672;
    i32 l1 = st->oldLogE[i], l2 = st->oldLogE2[i];
    if (!isTransient) { l2 = l1; l1 = e; }
    else l1 = min32(l1, e);
    if (b < start || b >= end) { e = 0; l1 = l2 = -28672; }
    st->oldBandE[i] = (i16)e;
    st->oldLogE[i] = (i16)l1;
    st->oldLogE2[i] = (i16)l2;
  }
  s.prefilter_period = pitch_index;
  s.prefilter_gain = gain1;
  s.prefilter_tapset = prefilter_tapset;
  if (isTransient || transient_got_disabled) s.consec_transient++;
  else s.consec_transient = 0;
  s.rng = enc.rng;
  re_done(enc);
  CB_WSYNC();
  CB_IF_LANE0 st->s = s;
  CB_WSYNC();
  if (enc.error) return CB_INTERNAL_ERROR;
  return nbCompressedBytes;
}
