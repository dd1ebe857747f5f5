// cb_enc_decisions.h -- the encoder's per-frame decisions: time/frequency resolution, spreading,
// dynamic allocation boosts, allocation trim, dual stereo, intensity hysteresis and the VBR target.
// Bit-exact with tf_analysis / alloc_trim_analysis / stereo_analysis / dynalloc_analysis /
// compute_vbr (opus-fix/celt/celt_encoder.c:538-1066,1196-1313) and spreading_decision /
// hysteresis_decision (celt/bands.c:48-61,428-519).  No float analysis (DISABLE_FLOAT_API).
// The LFE flag and the surround mask terms are only non-zero for the streams of a surround multistream encoder.
#pragma once
#include "cb_common.h"
#include "cb_celt_entropy.h"   // v_haar1, ct_tf_select, ebands
#include "cb_enc_state.h"

// l1_metric (celt_encoder.c:538-549): lane-parallel |x| sum
CB_FN i32 enc_l1_metric(const i16* tmp, int N, int LM, i32 bias) {
  i32 s = 0;
  CB_LANE_LOOP(i, N) { i32 v = tmp[i]; s += v < 0 ? -v : v; }
  s = cb_wsum(s);
  return add32(s, mult16_32_q15((i16)(LM * bias), s));       // MAC16_32_Q15
}

// tf_analysis (celt_encoder.c:551-712).  Fills w.tf_res[0..len), returns tf_select.
CB_FN int enc_tf_analysis(EncWork& w, int len, int isTransient, int lambda, int N0, int LM, int* tf_sum,
                          i32 tf_estimate, int tf_chan) {
  const i32 bias = (i16)mult16_16_q14(1311, max32(-4096, 8192 - tf_estimate));
  i16* tmp = w.vtmp;
  i16* tmp_1 = w.vtmp1;
  *tf_sum = 0;
  CB_NOUNROLL for (int i = 0; i < len; i++) {
    const int N = (ebands(i + 1) - ebands(i)) << LM;
    const int narrow = (ebands(i + 1) - ebands(i)) == 1;
    const i16* src = &w.X[tf_chan * N0 + (ebands(i) << LM)];
    CB_WSYNC();
    CB_LANE_LOOP(j, N) tmp[j] = src[j];
    CB_WSYNC();
    i32 L1 = enc_l1_metric(tmp, N, isTransient ? LM : 0, bias);
    i32 best_L1 = L1;
    int best_level = 0;
    if (isTransient && !narrow) {
      CB_LANE_LOOP(j, N) tmp_1[j] = tmp[j];
      v_haar1(tmp_1, N >> LM, 1 << LM);
      L1 = enc_l1_metric(tmp_1, N, LM + 1, bias);
      if (L1 < best_L1) { best_L1 = L1; best_level = -1; }
    }
    const int kmax = LM + !(isTransient || narrow);
    CB_NOUNROLL for (int k = 0; k < kmax; k++) {
      const int B = isTransient ? (LM - k - 1) : (k + 1);
      v_haar1(tmp, N >> k, 1 << k);
      L1 = enc_l1_metric(tmp, N, B, bias);
      if (L1 < best_L1) { best_L1 = L1; best_level = k + 1; }
    }
    int metric = isTransient ? 2 * best_level : -2 * best_level;
    *tf_sum += (isTransient ? LM : 0) - metric / 2;
    if (narrow && (metric == 0 || metric == -2 * LM)) metric -= 1;
    w.metric[i] = metric;
  }
  CB_WSYNC();
  // tf_select search + Viterbi: short serial recurrences, run uniformly
  int tf_select = 0;
  int selcost[2];
  for (int sel = 0; sel < 2; sel++) {
    int cost0 = 0, cost1 = isTransient ? 0 : lambda;
    CB_NOUNROLL for (int i = 1; i < len; i++) {
      int curr0 = cb_imin(cost0, cost1 + lambda);
      int curr1 = cb_imin(cost0 + lambda, cost1);
      int d0 = w.metric[i] - 2 * ct_tf_select[LM][4 * isTransient + 2 * sel + 0];
      int d1 = w.metric[i] - 2 * ct_tf_select[LM][4 * isTransient + 2 * sel + 1];
      cost0 = curr0 + (d0 < 0 ? -d0 : d0);
      cost1 = curr1 + (d1 < 0 ? -d1 : d1);
    }
    selcost[sel] = cb_imin(cost0, cost1);
  }
  if (selcost[1] < selcost[0] && isTransient) tf_select = 1;
  int cost0 = 0, cost1 = isTransient ? 0 : lambda;
  CB_NOUNROLL for (int i = 1; i < len; i++) {
    int curr0, curr1;
    int from0 = cost0, from1 = cost1 + lambda;
    if (from0 < from1) { curr0 = from0; w.path0[i] = 0; } else { curr0 = from1; w.path0[i] = 1; }
    from0 = cost0 + lambda; from1 = cost1;
    if (from0 < from1) { curr1 = from0; w.path1[i] = 0; } else { curr1 = from1; w.path1[i] = 1; }
    int d0 = w.metric[i] - 2 * ct_tf_select[LM][4 * isTransient + 2 * tf_select + 0];
    int d1 = w.metric[i] - 2 * ct_tf_select[LM][4 * isTransient + 2 * tf_select + 1];
    cost0 = curr0 + (d0 < 0 ? -d0 : d0);
    cost1 = curr1 + (d1 < 0 ? -d1 : d1);
  }
  CB_WSYNC();
  w.tf_res[len - 1] = cost0 < cost1 ? 0 : 1;
  CB_NOUNROLL for (int i = len - 2; i >= 0; i--) {
    CB_WSYNC();
    w.tf_res[i] = (w.tf_res[i + 1] == 1) ? w.path1[i + 1] : w.path0[i + 1];
  }
  CB_WSYNC();
  return tf_select;
}

// hysteresis_decision (bands.c:48-61)
CB_FN int enc_hysteresis(i32 val, const i16* thresholds, const i16* hysteresis, int N, int prev) {
  int i;
  for (i = 0; i < N; i++)
    if (val < thresholds[i]) break;
  if (i > prev && val < thresholds[prev] + hysteresis[prev]) i = prev;
  if (i < prev && val > thresholds[prev - 1] - hysteresis[prev - 1]) i = prev;
  return i;
}
CB_UTABLE i16 ct_intensity_thresholds[21] = {1, 2, 3, 4, 5, 6, 7, 8, 16, 24, 36, 44, 50, 56, 62, 67, 72, 79, 88, 106, 134};
CB_UTABLE i16 ct_intensity_histeresis[21] = {1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 3, 3, 4, 5, 6, 8, 8};

// spreading_decision (bands.c:428-519)
CB_FN int enc_spreading_decision(const i16* X, i32* average, int last_decision, i32* hf_average,
                                 i32* tapset_decision, int update_hf, int end, int C, int M) {
  const int N0 = M * CB_SHORT_MDCT;
  int sum = 0, nbBands = 0, hf_sum = 0;
  if (M * (ebands(end) - ebands(end - 1)) <= 8) return CB_SPREAD_NONE;
  CB_WSYNC();
  CB_NOUNROLL for (int c = 0; c < C; c++) {
    CB_NOUNROLL for (int i = 0; i < end; i++) {
      const i16* x = X + M * ebands(i) + c * N0;
      const int N = M * (ebands(i + 1) - ebands(i));
      if (N <= 8) continue;
      int t0 = 0, t1 = 0, t2 = 0;
      CB_LANE_LOOP(j, N) {
        i32 x2N = mult16_16(mult16_16_q15(x[j], x[j]), N);
        t0 += x2N < 2048;
        t1 += x2N < 512;
        t2 += x2N < 128;
      }
      t0 = cb_wsum(t0); t1 = cb_wsum(t1); t2 = cb_wsum(t2);
      if (i > CB_NBANDS - 4) hf_sum += (int)((u32)(32 * (t1 + t0)) / (u32)N);
      sum += ((2 * t2 >= N) + (2 * t1 >= N) + (2 * t0 >= N)) * 256;
      nbBands++;
    }
  }
  if (update_hf) {
    if (hf_sum) hf_sum = (int)((u32)hf_sum / (u32)(C * (4 - CB_NBANDS + end)));
    *hf_average = (*hf_average + hf_sum) >> 1;
    hf_sum = *hf_average;
    if (*tapset_decision == 2) hf_sum += 4;
    else if (*tapset_decision == 0) hf_sum -= 4;
    if (hf_sum > 22) *tapset_decision = 2;
    else if (hf_sum > 18) *tapset_decision = 1;
    else *tapset_decision = 0;
  }
  sum = (int)((u32)sum / (u32)nbBands);
  sum = (sum + *average) >> 1;
  *average = sum;
  sum = (3 * sum + (((3 - last_decision) << 7) + 64) + 2) >> 2;
  if (sum < 80) return CB_SPREAD_AGGRESSIVE;
  if (sum < 256) return CB_SPREAD_NORMAL;
  if (sum < 384) return CB_SPREAD_LIGHT;
  return CB_SPREAD_NONE;
}

// median_of_5 / median_of_3 (celt_encoder.c:876-935)
CB_FN i32 enc_median5(const i16* x) {
  i32 t0, t1, t2 = x[2], t3, t4;
  if (x[0] > x[1]) { t0 = x[1]; t1 = x[0]; } else { t0 = x[0]; t1 = x[1]; }
  if (x[3] > x[4]) { t3 = x[4]; t4 = x[3]; } else { t3 = x[3]; t4 = x[4]; }
  if (t0 > t3) { i32 t = t0; t0 = t3; t3 = t; t = t1; t1 = t4; t4 = t; }
  if (t2 > t1) return (t1 < t3) ? min32(t2, t3) : min32(t4, t1);
  return (t2 < t3) ? min32(t1, t3) : min32(t2, t4);
}
CB_FN i32 enc_median3(const i16* x) {
  i32 t0, t1, t2 = x[2];
  if (x[0] > x[1]) { t0 = x[1]; t1 = x[0]; } else { t0 = x[0]; t1 = x[1]; }
  if (t1 < t2) return t1;
  if (t0 < t2) return t2;
  return t0;
}

// dynalloc_analysis (celt_encoder.c:937-1065).  21-band recurrences, run uniformly; every
// opus_val16 assignment of the reference is an explicit (i16) here.  Returns maxDepth.
CB_FN i32 enc_dynalloc_analysis(EncWork& w, int start, int end, int C, int lsb_depth, int isTransient, int vbr,
                                int constrained_vbr, int LM, int effectiveBytes, i32* tot_boost_, int lfe = 0,
                                const i16* surround_dynalloc = nullptr) {
  const i16* bandLogE = w.bandLogE;
  const i16* bandLogE2 = w.bandLogE2;
  i16* follower = w.follower;
  i16* noise_floor = w.noise_floor;
  i32 tot_boost = 0;
  CB_WSYNC();
  for (int i = 0; i < CB_NBANDS; i++) w.offsets[i] = 0;
  i16 maxDepth = -32666;
  CB_NOUNROLL for (int i = 0; i < end; i++)
    noise_floor[i] = (i16)(mult16_16(64, ct_logN400[i]) + 512 + shl16(9 - lsb_depth, CB_DB_SHIFT) -
                           shl16((i16)ct_eMeans[i], 6) + mult16_16(6, (i + 5) * (i + 5)));
  CB_WSYNC();
  for (int c = 0; c < C; c++)
    CB_NOUNROLL for (int i = 0; i < end; i++)
      maxDepth = (i16)max32(maxDepth, bandLogE[c * CB_NBANDS + i] - noise_floor[i]);
  if (effectiveBytes > 50 && LM >= 1 && !lfe) {
    int last = 0;
    for (int c = 0; c < C; c++) {
      i16* f = &follower[c * CB_NBANDS];
      const i16* e2 = &bandLogE2[c * CB_NBANDS];
      i32 fp = e2[0];
      f[0] = (i16)fp;
      CB_NOUNROLL for (int i = 1; i < end; i++) {
        if (e2[i] > e2[i - 1] + 512) last = i;
        fp = (i16)min32(fp + 1536, e2[i]);
        f[i] = (i16)fp;
      }
      CB_WSYNC();
      CB_NOUNROLL for (int i = last - 1; i >= 0; i--) {
        i32 v = (i16)min32(f[i], min32(f[i + 1] + 2048, e2[i]));
        CB_WSYNC();
        f[i] = (i16)v;
        CB_WSYNC();
      }
      const i32 offset = 1024;
      CB_NOUNROLL for (int i = 2; i < end - 2; i++) f[i] = (i16)max32(f[i], enc_median5(&e2[i - 2]) - offset);
      i32 tmp = (i16)(enc_median3(&e2[0]) - offset);
      f[0] = (i16)max32(f[0], tmp);
      f[1] = (i16)max32(f[1], tmp);
      tmp = (i16)(enc_median3(&e2[end - 3]) - offset);
      f[end - 2] = (i16)max32(f[end - 2], tmp);
      f[end - 1] = (i16)max32(f[end - 1], tmp);
      CB_NOUNROLL for (int i = 0; i < end; i++) f[i] = (i16)max32(f[i], noise_floor[i]);
      CB_WSYNC();
    }
    if (C == 2) {
      CB_NOUNROLL for (int i = start; i < end; i++) {
        i32 f1 = (i16)max32(follower[CB_NBANDS + i], follower[i] - 4096);
        i32 f0 = (i16)max32(follower[i], f1 - 4096);
        i32 v = (i16)half32(max32(0, bandLogE[i] - f0) + max32(0, bandLogE[CB_NBANDS + i] - f1));
        CB_WSYNC();
        follower[CB_NBANDS + i] = (i16)f1;
        follower[i] = (i16)v;
      }
    } else {
      CB_NOUNROLL for (int i = start; i < end; i++) {
        i32 v = (i16)max32(0, bandLogE[i] - follower[i]);
        CB_WSYNC();
        follower[i] = (i16)v;
      }
    }
    CB_WSYNC();
    // without a surround mask surround_dynalloc is all zero
    CB_NOUNROLL for (int i = start; i < end; i++) {
      i32 fo = max32(follower[i], surround_dynalloc ? surround_dynalloc[i] : 0);
      if ((!vbr || constrained_vbr) && !isTransient) fo = (i16)(fo >> 1);
      if (i < 8) fo = (i16)(fo * 2);
      if (i >= 12) fo = (i16)(fo >> 1);
      fo = (i16)min32(fo, 4096);
      const int width = C * (ebands(i + 1) - ebands(i)) << LM;
      int boost, boost_bits;
      if (width < 6) {
        boost = (int)(fo >> CB_DB_SHIFT);
        boost_bits = boost * width << CB_BITRES;
      } else if (width > 48) {
        boost = (int)((fo * 8) >> CB_DB_SHIFT);
        boost_bits = (boost * width << CB_BITRES) / 8;
      } else {
        boost = (int)((fo * width / 6) >> CB_DB_SHIFT);
        boost_bits = boost * 6 << CB_BITRES;
      }
      if ((!vbr || (constrained_vbr && !isTransient)) &&
          ((tot_boost + boost_bits) >> CB_BITRES >> 3) > effectiveBytes / 4) {
        i32 cap = ((effectiveBytes / 4) << CB_BITRES << 3);
        w.offsets[i] = cap - tot_boost;
        tot_boost = cap;
        break;
      } else {
        w.offsets[i] = boost;
        tot_boost += boost_bits;
      }
    }
  }
  CB_WSYNC();
  *tot_boost_ = tot_boost;
  return maxDepth;
}

// lane-parallel inner product of two normalised-spectrum bands
CB_FN i32 enc_inner_norm(const i16* a, const i16* b, int n) {
  i32 s = 0;
  CB_LANE_LOOP(j, n) s = mac16_16(s, a[j], b[j]);
  return cb_wsum(s);
}

// alloc_trim_analysis (celt_encoder.c:756-841); *stereo_saving is a caller local.
CB_FN int enc_alloc_trim_analysis(const i16* X, const i16* bandLogE, int end, int LM, int C, int N0,
                                  i32* stereo_saving, i32 tf_estimate, int intensity, i32 surround_trim = 0) {
  i32 diff = 0;
  i32 trim = 1280;                                            // QCONST16(5.f, 8)
  CB_WSYNC();
  if (C == 2) {
    i32 sum = 0;
    CB_NOUNROLL for (int i = 0; i < 8; i++) {
      i32 partial = enc_inner_norm(&X[ebands(i) << LM], &X[N0 + (ebands(i) << LM)], (ebands(i + 1) - ebands(i)) << LM);
      sum = (i16)(sum + (i16)(partial >> 18));
    }
    sum = (i16)mult16_16_q15(4096, sum);
    sum = (i16)min32(1024, sum < 0 ? -sum : sum);
    i32 minXC = sum;
    CB_NOUNROLL for (int i = 8; i < intensity; i++) {
      i32 partial = enc_inner_norm(&X[ebands(i) << LM], &X[N0 + (ebands(i) << LM)], (ebands(i + 1) - ebands(i)) << LM);
      i32 p = (i16)(partial >> 18);
      minXC = (i16)min32(minXC, p < 0 ? -p : p);
    }
    minXC = (i16)min32(1024, minXC < 0 ? -minXC : minXC);
    i32 logXC = cb_log2(1049625 - mult16_16(sum, sum));         // QCONST32(1.001f,20)
    i32 logXC2 = (i16)max32(logXC >> 1, cb_log2(1049625 - mult16_16(minXC, minXC)));
    logXC = (i16)pshr32(logXC - 6144, CB_DB_SHIFT - 8);
    logXC2 = (i16)pshr32(logXC2 - 6144, CB_DB_SHIFT - 8);
    trim = (i16)(trim + max32(-1024, mult16_16_q15(24576, logXC)));
    *stereo_saving = (i16)min32(*stereo_saving + 64, -(logXC2 >> 1));
  }
  for (int c = 0; c < C; c++)
    CB_NOUNROLL for (int i = 0; i < end - 1; i++) diff += bandLogE[i + c * CB_NBANDS] * (i32)(2 + 2 * i - end);
  diff /= C * (end - 1);
  trim = (i16)(trim - max32(-512, min32(512, ((diff + 1024) >> (CB_DB_SHIFT - 8)) / 6)));
  trim = (i16)(trim - ((i16)surround_trim >> (CB_DB_SHIFT - 8)));
  trim = (i16)(trim - 2 * (tf_estimate >> (14 - 8)));
  int trim_index = pshr32(trim, 8);
  return cb_imax(0, cb_imin(10, trim_index));
}

// stereo_analysis (celt_encoder.c:843-873)
CB_FN int enc_stereo_analysis(const i16* X, int LM, int N0) {
  i32 sLR = 0, sMS = 0;
  const int n = ebands(13) << LM;
  CB_WSYNC();
  CB_LANE_LOOP(j, n) {
    i32 L = X[j], R = X[N0 + j];
    i32 M = L + R, S = L - R;
    sLR += (L < 0 ? -L : L) + (R < 0 ? -R : R);
    sMS += (M < 0 ? -M : M) + (S < 0 ? -S : S);
  }
  i32 sumLR = add32(1, cb_wsum(sLR)), sumMS = add32(1, cb_wsum(sMS));
  sumMS = mult16_32_q15(23170, sumMS);
  int thetas = 13;
  if (LM <= 1) thetas -= 8;
  return mult16_32_q15((ebands(13) << (LM + 1)) + thetas, sumMS) > mult16_32_q15(ebands(13) << (LM + 1), sumLR);
}

// compute_vbr (celt_encoder.c:1196-1313) without analysis
CB_FN i32 enc_compute_vbr(i32 base_target, int LM, i32 bitrate, int lastCodedBands, int C, int intensity,
                          int constrained_vbr, i32 stereo_saving, int tot_boost, i32 tf_estimate, i32 maxDepth,
                          int variable_duration, i32 temporal_vbr, int lfe = 0, int has_surround_mask = 0,
                          i32 surround_masking = 0) {
  const int coded_bands = lastCodedBands ? lastCodedBands : CB_NBANDS;
  int coded_bins = ebands(coded_bands) << LM;
  if (C == 2) coded_bins += ebands(cb_imin(intensity, coded_bands)) << LM;
  i32 target = base_target;
  if (C == 2) {
    const int coded_stereo_bands = cb_imin(intensity, coded_bands);
    const int coded_stereo_dof = (ebands(coded_stereo_bands) << LM) - coded_stereo_bands;
    const i32 max_frac = (i16)(mult16_16(26214, coded_stereo_dof) / (i32)(i16)coded_bins);   // DIV32_16
    stereo_saving = (i16)min32(stereo_saving, 256);
    target -= min32(mult16_32_q15(max_frac, target),
                    mult16_16(stereo_saving - 26, (coded_stereo_dof << CB_BITRES)) >> 8);
  }
  target += tot_boost - (16 << LM);
  const i32 tf_calibration = variable_duration == 5010 ? 328 : 655;     // OPUS_FRAMESIZE_VARIABLE
  target += shl32(mult16_32_q15(tf_estimate - tf_calibration, target), 1);
  if (has_surround_mask && !lfe) {
    const i32 surround_target = target + (mult16_16((i16)surround_masking, (i16)(coded_bins << CB_BITRES)) >> CB_DB_SHIFT);
    target = cb_imax(target / 4, surround_target);
  }
  {
    const int bins = ebands(CB_NBANDS - 2) << LM;
    i32 floor_depth = mult16_16((C * bins << CB_BITRES), maxDepth) >> CB_DB_SHIFT;
    floor_depth = max32(floor_depth, target >> 2);
    target = min32(target, floor_depth);
  }
  if ((!has_surround_mask || lfe) && (constrained_vbr || bitrate < 64000)) {
    i32 rate_factor = max32(0, bitrate - 32000);
    if (constrained_vbr) rate_factor = min32(rate_factor, 21955);
    target = base_target + mult16_32_q15(rate_factor, target - base_target);
  }
  if (!has_surround_mask && tf_estimate < 3277) {
    const i32 amount = (i16)mult16_16_q15(3329, cb_imax(0, cb_imin(32000, 96000 - bitrate)));
    const i32 tvbr_factor = (i16)(mult16_16(temporal_vbr, amount) >> CB_DB_SHIFT);
    target += mult16_32_q15(tvbr_factor, target);
  }
  target = min32(2 * base_target, target);
  return target;
}
