// cb_silk_params.h -- SILK side information: entropy decoding of the indices and the excitation pulses,
// and their conversion to gains, LPC/LTP coefficients and pitch lags.  Bit-exact with
// silk/decode_indices.c, decode_pulses.c, shell_coder.c:137-151, code_signs.c:74-115,
// decode_parameters.c, gain_quant.c:107-138, NLSF_decode.c, NLSF_unpack.c, NLSF_VQ_weights_laroia.c,
// NLSF_stabilize.c, NLSF2A.c, bwexpander.c, bwexpander_32.c, LPC_inv_pred_gain.c, decode_pitch.c,
// stereo_decode_pred.c of the reference.  Scalar code: one thread per stream.
#pragma once
#include "cb_silk_common.h"
#include "cb_rangedec.h"

// the two NLSF codebooks (tables_NLSF_CB_WB.c:184-198, tables_NLSF_CB_NB_MB.c:146-160)
struct SilkNLSFCB {
  int nVectors, order;
  i32 quantStepSize_Q16;
  const u8* CB1_NLSF_Q8; const u8* CB1_iCDF; const u8* pred_Q8; const u8* ec_sel; const u8* ec_iCDF;
  const i16* deltaMin_Q15;
};
CB_FN SilkNLSFCB sk_nlsf_cb(int wb) {
  SilkNLSFCB c;
  c.nVectors = 32;
  if (wb) {
    c.order = 16; c.quantStepSize_Q16 = 9830;
    c.CB1_NLSF_Q8 = st_NLSF_CB1_WB_Q8; c.CB1_iCDF = st_NLSF_CB1_iCDF_WB; c.pred_Q8 = st_NLSF_PRED_WB_Q8;
    c.ec_sel = st_NLSF_CB2_SELECT_WB; c.ec_iCDF = st_NLSF_CB2_iCDF_WB; c.deltaMin_Q15 = st_NLSF_DELTA_MIN_WB_Q15;
  } else {
    c.order = 10; c.quantStepSize_Q16 = 11796;
    c.CB1_NLSF_Q8 = st_NLSF_CB1_NB_MB_Q8; c.CB1_iCDF = st_NLSF_CB1_iCDF_NB_MB; c.pred_Q8 = st_NLSF_PRED_NB_MB_Q8;
    c.ec_sel = st_NLSF_CB2_SELECT_NB_MB; c.ec_iCDF = st_NLSF_CB2_iCDF_NB_MB; c.deltaMin_Q15 = st_NLSF_DELTA_MIN_NB_MB_Q15;
  }
  return c;
}
// iCDF tables that depend on rate / frame size (decoder_set_fs.c:57-92)
CB_FN const u8* sk_pitch_contour_icdf(int fs_kHz, int nb_subfr) {
  if (fs_kHz == 8) return nb_subfr == 4 ? st_pitch_contour_NB_iCDF : st_pitch_contour_10ms_NB_iCDF;
  return nb_subfr == 4 ? st_pitch_contour_iCDF : st_pitch_contour_10ms_iCDF;
}
// (indexed per lane like every SILK table, and padded by 4 for rd_icdf_g's read-ahead)
CB_GTABLE u8 st_uniform3_iCDF[3 + 4] = {171, 85, 0, 0, 0, 0, 0};
CB_GTABLE u8 st_uniform4_iCDF[4 + 4] = {192, 128, 64, 0, 0, 0, 0, 0};
CB_GTABLE u8 st_uniform5_iCDF[5 + 4] = {205, 154, 102, 51, 0, 0, 0, 0, 0};
CB_GTABLE u8 st_uniform6_iCDF[6 + 4] = {213, 171, 128, 85, 43, 0, 0, 0, 0, 0};
CB_GTABLE u8 st_uniform8_iCDF[8 + 4] = {224, 192, 160, 128, 96, 64, 32, 0, 0, 0, 0, 0};
CB_GTABLE u8 st_LTPscale_iCDF[3 + 4] = {128, 64, 0, 0, 0, 0, 0};
CB_GTABLE u8 st_stereo_only_code_mid_iCDF[2 + 4] = {64, 0, 0, 0, 0, 0};
CB_FN const u8* sk_pitch_lag_low_bits_icdf(int fs_kHz) {
  return fs_kHz == 16 ? st_uniform8_iCDF : fs_kHz == 12 ? st_uniform6_iCDF : st_uniform4_iCDF;
}

// silk_NLSF_unpack (NLSF_unpack.c:34-55)
CB_FN void sk_nlsf_unpack(i16* ec_ix, u8* pred_Q8, const SilkNLSFCB& cb, int CB1_index) {
  const u8* ec_sel_ptr = &cb.ec_sel[CB1_index * cb.order / 2];
  for (int i = 0; i < cb.order; i += 2) {
    const u8 entry = *ec_sel_ptr++;
    ec_ix[i] = (i16)sk_smulbb((entry >> 1) & 7, 2 * 4 + 1);
    pred_Q8[i] = cb.pred_Q8[i + (entry & 1) * (cb.order - 1)];
    ec_ix[i + 1] = (i16)sk_smulbb((entry >> 5) & 7, 2 * 4 + 1);
    pred_Q8[i + 1] = cb.pred_Q8[i + ((entry >> 4) & 1) * (cb.order - 1) + 1];
  }
}

// silk_decode_indices (decode_indices.c:35-151)
CB_FN void sk_decode_indices(SilkChannel& ps, RangeDec& rd, int FrameIndex, int decode_LBRR, int condCoding) {
  SilkIndices& ix = ps.indices;
  int Ix;
  i16 ec_ix[SK_MAX_ORDER];
  u8 pred_Q8[SK_MAX_ORDER];
  if (decode_LBRR || ps.VAD_flags[FrameIndex]) Ix = rd_icdf_g(rd, st_type_offset_VAD_iCDF, 8) + 2;
  else Ix = rd_icdf_g(rd, st_type_offset_no_VAD_iCDF, 8);
  ix.signalType = (i8)(Ix >> 1);
  ix.quantOffsetType = (i8)(Ix & 1);
  if (condCoding == SK_CODE_CONDITIONALLY) {
    ix.GainsIndices[0] = (i8)rd_icdf_g(rd, st_delta_gain_iCDF, 8);
  } else {
    ix.GainsIndices[0] = (i8)(rd_icdf_g(rd, st_gain_iCDF + 8 * ix.signalType, 8) << 3);
    ix.GainsIndices[0] = (i8)(ix.GainsIndices[0] + (i8)rd_icdf_g(rd, st_uniform8_iCDF, 8));
  }
  for (int i = 1; i < ps.nb_subfr; i++) ix.GainsIndices[i] = (i8)rd_icdf_g(rd, st_delta_gain_iCDF, 8);
  const SilkNLSFCB cb = sk_nlsf_cb(ps.LPC_order == 16);
  ix.NLSFIndices[0] = (i8)rd_icdf_g(rd, &cb.CB1_iCDF[(ix.signalType >> 1) * cb.nVectors], 8);
  sk_nlsf_unpack(ec_ix, pred_Q8, cb, ix.NLSFIndices[0]);
  for (int i = 0; i < cb.order; i++) {
    Ix = rd_icdf_g(rd, &cb.ec_iCDF[ec_ix[i]], 8);
    if (Ix == 0) Ix -= rd_icdf_g(rd, st_NLSF_EXT_iCDF, 8);
    else if (Ix == 2 * 4) Ix += rd_icdf_g(rd, st_NLSF_EXT_iCDF, 8);
    ix.NLSFIndices[i + 1] = (i8)(Ix - 4);
  }
  if (ps.nb_subfr == SK_MAX_NB_SUBFR) ix.NLSFInterpCoef_Q2 = (i8)rd_icdf_g(rd, st_NLSF_interp_iCDF, 8);
  else ix.NLSFInterpCoef_Q2 = 4;
  if (ix.signalType == SK_TYPE_VOICED) {
    int decode_absolute_lagIndex = 1;
    if (condCoding == SK_CODE_CONDITIONALLY && ps.ec_prevSignalType == SK_TYPE_VOICED) {
      int delta_lagIndex = (i16)rd_icdf_g(rd, st_pitch_delta_iCDF, 8);
      if (delta_lagIndex > 0) {
        delta_lagIndex = delta_lagIndex - 9;
        ix.lagIndex = (i16)(ps.ec_prevLagIndex + delta_lagIndex);
        decode_absolute_lagIndex = 0;
      }
    }
    if (decode_absolute_lagIndex) {
      ix.lagIndex = (i16)((i16)rd_icdf_g(rd, st_pitch_lag_iCDF, 8) * (ps.fs_kHz >> 1));
      ix.lagIndex = (i16)(ix.lagIndex + (i16)rd_icdf_g(rd, sk_pitch_lag_low_bits_icdf(ps.fs_kHz), 8));
    }
    ps.ec_prevLagIndex = ix.lagIndex;
    ix.contourIndex = (i8)rd_icdf_g(rd, sk_pitch_contour_icdf(ps.fs_kHz, ps.nb_subfr), 8);
    ix.PERIndex = (i8)rd_icdf_g(rd, st_LTP_per_index_iCDF, 8);
    const u8* gi = ix.PERIndex == 0 ? st_LTP_gain_iCDF_0 : ix.PERIndex == 1 ? st_LTP_gain_iCDF_1 : st_LTP_gain_iCDF_2;
    for (int k = 0; k < ps.nb_subfr; k++) ix.LTPIndex[k] = (i8)rd_icdf_g(rd, gi, 8);
    if (condCoding == SK_CODE_INDEPENDENTLY) ix.LTP_scaleIndex = (i8)rd_icdf_g(rd, st_LTPscale_iCDF, 8);
    else ix.LTP_scaleIndex = 0;
  }
  ps.ec_prevSignalType = ix.signalType;
  ix.Seed = (i8)rd_icdf_g(rd, st_uniform4_iCDF, 8);
}

// decode_split / silk_shell_decoder (shell_coder.c:60-76,137-151)
CB_FN void sk_decode_split(i16* c1, i16* c2, RangeDec& rd, int p, const u8* shell_table) {
  if (p > 0) {
    c1[0] = (i16)rd_icdf_g(rd, &shell_table[st_shell_code_table_offsets[p]], 8);
    c2[0] = (i16)(p - c1[0]);
  } else {
    c1[0] = 0;
    c2[0] = 0;
  }
}
CB_FN void sk_shell_decoder(i16* pulses0, RangeDec& rd, int pulses4) {
  i16 pulses3[2], pulses2[4], pulses1[8];
  sk_decode_split(&pulses3[0], &pulses3[1], rd, pulses4, st_shell_code_table3);
  for (int a = 0; a < 2; a++) {
    sk_decode_split(&pulses2[2 * a], &pulses2[2 * a + 1], rd, pulses3[a], st_shell_code_table2);
    for (int b = 0; b < 2; b++) {
      const int i1 = 2 * a + b;
      sk_decode_split(&pulses1[2 * i1], &pulses1[2 * i1 + 1], rd, pulses2[i1], st_shell_code_table1);
      for (int c = 0; c < 2; c++) {
        const int i0 = 2 * i1 + c;
        sk_decode_split(&pulses0[2 * i0], &pulses0[2 * i0 + 1], rd, pulses1[i0], st_shell_code_table0);
      }
    }
  }
}

// silk_decode_pulses + silk_decode_signs (decode_pulses.c:38-115, code_signs.c:74-115)
CB_FN void sk_decode_pulses(RangeDec& rd, i16* pulses, int signalType, int quantOffsetType, int frame_length) {
  int sum_pulses[SK_MAX_FRAME / 16], nLshifts[SK_MAX_FRAME / 16];
  const int RateLevelIndex = rd_icdf_g(rd, st_rate_levels_iCDF + 9 * (signalType >> 1), 8);
  int iter = frame_length >> 4;
  if (iter * 16 < frame_length) iter++;
  const u8* cdf_ptr = st_pulses_per_block_iCDF + 18 * RateLevelIndex;
  for (int i = 0; i < iter; i++) {
    nLshifts[i] = 0;
    sum_pulses[i] = rd_icdf_g(rd, cdf_ptr, 8);
    while (sum_pulses[i] == 16 + 1) {
      nLshifts[i]++;
      sum_pulses[i] = rd_icdf_g(rd, st_pulses_per_block_iCDF + 18 * 9 + (nLshifts[i] == 10), 8);
    }
  }
  for (int i = 0; i < iter; i++) {
    if (sum_pulses[i] > 0) sk_shell_decoder(&pulses[16 * i], rd, sum_pulses[i]);
    else for (int k = 0; k < 16; k++) pulses[16 * i + k] = 0;
  }
  for (int i = 0; i < iter; i++) {
    if (nLshifts[i] > 0) {
      const int nLS = nLshifts[i];
      i16* pp = &pulses[16 * i];
      for (int k = 0; k < 16; k++) {
        int abs_q = pp[k];
        for (int j = 0; j < nLS; j++) {
          abs_q = abs_q << 1;
          abs_q += rd_icdf2_g(rd, 120, 8);               // silk_lsb_iCDF = {120, 0}
        }
        pp[k] = (i16)abs_q;
      }
      sum_pulses[i] |= nLS << 5;
    }
  }
  // signs
  const u8* icdf_ptr = &st_sign_iCDF[7 * (quantOffsetType + (signalType << 1))];
  const int nblk = (frame_length + 8) >> 4;
  i16* q = pulses;
  for (int i = 0; i < nblk; i++) {
    const int p = sum_pulses[i];
    if (p > 0) {
      const u32 p0 = icdf_ptr[sk_min(p & 0x1F, 6)];            // the table is {p0, 0} (code_signs.c:98-113)
      for (int j = 0; j < 16; j++) {
        if (q[j] > 0) q[j] = (i16)(q[j] * ((rd_icdf2_g(rd, p0, 8) << 1) - 1));
      }
    }
    q += 16;
  }
}

// silk_gains_dequant (gain_quant.c:107-138)
CB_FN void sk_gains_dequant(i32* gain_Q16, const i8* ind, i32* prev_ind, int conditional, int nb_subfr) {
  const int OFFSET = (2 * 128) / 6 + 16 * 128;                                  // 2090
  const int INV_SCALE_Q16 = (65536 * (((88 - 2) * 128) / 6)) / (64 - 1);        // 1907825
  for (int k = 0; k < nb_subfr; k++) {
    i32 p = (i8)*prev_ind;
    if (k == 0 && conditional == 0) {
      p = sk_max(ind[k], p - 16);
    } else {
      const int ind_tmp = ind[k] + (-4);
      const int thr = 2 * 36 - 64 + p;
      if (ind_tmp > thr) p += (ind_tmp << 1) - thr;
      else p += ind_tmp;
    }
    p = (i8)p;                                                                  // *prev_ind is an opus_int8
    p = sk_limit(p, 0, 64 - 1);
    *prev_ind = p;
    gain_Q16[k] = sk_log2lin(sk_min(sk_smulwb(INV_SCALE_Q16, p) + OFFSET, 3967));
  }
}

// silk_NLSF_stabilize (NLSF_stabilize.c:48-142)
CB_FN void sk_nlsf_stabilize(i16* NLSF_Q15, const i16* NDeltaMin_Q15, int L) {
  int I = 0, loops;
  for (loops = 0; loops < 20; loops++) {
    i32 min_diff_Q15 = NLSF_Q15[0] - NDeltaMin_Q15[0];
    I = 0;
    for (int i = 1; i <= L - 1; i++) {
      const i32 diff_Q15 = NLSF_Q15[i] - (NLSF_Q15[i - 1] + NDeltaMin_Q15[i]);
      if (diff_Q15 < min_diff_Q15) { min_diff_Q15 = diff_Q15; I = i; }
    }
    const i32 diff_Q15 = (1 << 15) - (NLSF_Q15[L - 1] + NDeltaMin_Q15[L]);
    if (diff_Q15 < min_diff_Q15) { min_diff_Q15 = diff_Q15; I = L; }
    if (min_diff_Q15 >= 0) return;
    if (I == 0) {
      NLSF_Q15[0] = NDeltaMin_Q15[0];
    } else if (I == L) {
      NLSF_Q15[L - 1] = (i16)((1 << 15) - NDeltaMin_Q15[L]);
    } else {
      i32 min_center_Q15 = 0;
      for (int k = 0; k < I; k++) min_center_Q15 += NDeltaMin_Q15[k];
      min_center_Q15 += NDeltaMin_Q15[I] >> 1;
      i32 max_center_Q15 = 1 << 15;
      for (int k = L; k > I; k--) max_center_Q15 -= NDeltaMin_Q15[k];
      max_center_Q15 -= NDeltaMin_Q15[I] >> 1;
      const i16 center_freq_Q15 = (i16)sk_limit(sk_rshift_round((i32)NLSF_Q15[I - 1] + (i32)NLSF_Q15[I], 1),
                                                min_center_Q15, max_center_Q15);
      NLSF_Q15[I - 1] = (i16)(center_freq_Q15 - (NDeltaMin_Q15[I] >> 1));
      NLSF_Q15[I] = (i16)(NLSF_Q15[I - 1] + NDeltaMin_Q15[I]);
    }
  }
  // fall-back: insertion sort (sort.c silk_insertion_sort_increasing_all_values_int16) + clamps
  for (int i = 1; i < L; i++) {
    const i16 value = NLSF_Q15[i];
    int j;
    for (j = i - 1; j >= 0 && value < NLSF_Q15[j]; j--) NLSF_Q15[j + 1] = NLSF_Q15[j];
    NLSF_Q15[j + 1] = value;
  }
  NLSF_Q15[0] = (i16)sk_max(NLSF_Q15[0], NDeltaMin_Q15[0]);
  for (int i = 1; i < L; i++) NLSF_Q15[i] = (i16)sk_max(NLSF_Q15[i], NLSF_Q15[i - 1] + NDeltaMin_Q15[i]);
  NLSF_Q15[L - 1] = (i16)sk_min(NLSF_Q15[L - 1], (1 << 15) - NDeltaMin_Q15[L]);
  for (int i = L - 2; i >= 0; i--) NLSF_Q15[i] = (i16)sk_min(NLSF_Q15[i], NLSF_Q15[i + 1] - NDeltaMin_Q15[i + 1]);
}

// silk_NLSF_decode (NLSF_decode.c:34-101) with silk_NLSF_VQ_weights_laroia (NLSF_VQ_weights_laroia.c:43-80)
CB_FN void sk_nlsf_decode(i16* pNLSF_Q15, const i8* NLSFIndices, const SilkNLSFCB& cb) {
  u8 pred_Q8[SK_MAX_ORDER];
  i16 ec_ix[SK_MAX_ORDER], res_Q10[SK_MAX_ORDER], W[SK_MAX_ORDER];
  const int D = cb.order;
  const u8* pCB = &cb.CB1_NLSF_Q8[NLSFIndices[0] * D];
  for (int i = 0; i < D; i++) pNLSF_Q15[i] = (i16)((i16)pCB[i] << 7);
  sk_nlsf_unpack(ec_ix, pred_Q8, cb, NLSFIndices[0]);
  {                                                              // silk_NLSF_residual_dequant
    i32 out_Q10 = 0;
    for (int i = D - 1; i >= 0; i--) {
      const i32 pred_Q10 = sk_smulbb(out_Q10, (i16)pred_Q8[i]) >> 8;
      out_Q10 = sk_lshift(NLSFIndices[1 + i], 10);
      if (out_Q10 > 0) out_Q10 = out_Q10 - 102;
      else if (out_Q10 < 0) out_Q10 = out_Q10 + 102;
      out_Q10 = sk_smlawb(pred_Q10, out_Q10, cb.quantStepSize_Q16);
      res_Q10[i] = (i16)out_Q10;
    }
  }
  {                                                              // Laroia weights, NLSF_W_Q = 2
    i32 t1 = sk_max(pNLSF_Q15[0], 1);
    t1 = (1 << 17) / t1;
    i32 t2 = sk_max(pNLSF_Q15[1] - pNLSF_Q15[0], 1);
    t2 = (1 << 17) / t2;
    W[0] = (i16)sk_min(t1 + t2, 32767);
    for (int k = 1; k < D - 1; k += 2) {
      t1 = sk_max(pNLSF_Q15[k + 1] - pNLSF_Q15[k], 1);
      t1 = (1 << 17) / t1;
      W[k] = (i16)sk_min(t1 + t2, 32767);
      t2 = sk_max(pNLSF_Q15[k + 2] - pNLSF_Q15[k + 1], 1);
      t2 = (1 << 17) / t2;
      W[k + 1] = (i16)sk_min(t1 + t2, 32767);
    }
    t1 = sk_max((1 << 15) - pNLSF_Q15[D - 1], 1);
    t1 = (1 << 17) / t1;
    W[D - 1] = (i16)sk_min(t1 + t2, 32767);
  }
  for (int i = 0; i < D; i++) {
    const i32 W_tmp_Q9 = sk_sqrt_approx(sk_lshift((i32)W[i], 18 - 2));
    const i32 t = sk_add(pNLSF_Q15[i], sk_lshift((i32)res_Q10[i], 14) / W_tmp_Q9);
    pNLSF_Q15[i] = (i16)sk_limit(t, 0, 32767);
  }
  sk_nlsf_stabilize(pNLSF_Q15, cb.deltaMin_Q15, D);
}

// silk_bwexpander / silk_bwexpander_32 (bwexpander.c, bwexpander_32.c)
CB_FN void sk_bwexpander(i16* ar, int d, i32 chirp_Q16) {
  const i32 chirp_minus_one_Q16 = chirp_Q16 - 65536;
  for (int i = 0; i < d - 1; i++) {
    ar[i] = (i16)sk_rshift_round(sk_mul(chirp_Q16, ar[i]), 16);
    chirp_Q16 += sk_rshift_round(sk_mul(chirp_Q16, chirp_minus_one_Q16), 16);
  }
  ar[d - 1] = (i16)sk_rshift_round(sk_mul(chirp_Q16, ar[d - 1]), 16);
}
CB_FN void sk_bwexpander_32(i32* ar, int d, i32 chirp_Q16) {
  const i32 chirp_minus_one_Q16 = chirp_Q16 - 65536;
  for (int i = 0; i < d - 1; i++) {
    ar[i] = sk_smulww(chirp_Q16, ar[i]);
    chirp_Q16 += sk_rshift_round(sk_mul(chirp_Q16, chirp_minus_one_Q16), 16);
  }
  ar[d - 1] = sk_smulww(chirp_Q16, ar[d - 1]);
}

// silk_LPC_inverse_pred_gain (LPC_inv_pred_gain.c:43-133), QA = 24
CB_FN i32 sk_mul32_frac_q(i32 a, i32 b, int Q) { return (i32)sk_rshift_round64((i64)a * (i64)b, Q); }
CB_FN i32 sk_lpc_inverse_pred_gain(const i16* A_Q12, int order) {
  i32 A[2][SK_MAX_ORDER];
  i32* Anew = A[order & 1];
  i32 DC_resp = 0;
  for (int k = 0; k < order; k++) {
    DC_resp += (i32)A_Q12[k];
    Anew[k] = sk_lshift((i32)A_Q12[k], 24 - 12);
  }
  if (DC_resp >= 4096) return 0;
  const i32 A_LIMIT = 16773022;                                  // SILK_FIX_CONST(0.99975, 24)
  i32 invGain_Q30 = 1 << 30;
  for (int k = order - 1; k > 0; k--) {
    if (Anew[k] > A_LIMIT || Anew[k] < -A_LIMIT) return 0;
    const i32 rc_Q31 = -sk_lshift(Anew[k], 31 - 24);
    const i32 rc_mult1_Q30 = (1 << 30) - sk_smmul(rc_Q31, rc_Q31);
    const int mult2Q = 32 - sk_clz32(sk_abs(rc_mult1_Q30));
    const i32 rc_mult2 = sk_inverse32_varQ(rc_mult1_Q30, mult2Q + 30);
    invGain_Q30 = sk_lshift(sk_smmul(invGain_Q30, rc_mult1_Q30), 2);
    i32* Aold = Anew;
    Anew = A[k & 1];
    for (int n = 0; n < k; n++) {
      const i32 tmp_QA = Aold[n] - sk_mul32_frac_q(Aold[k - n - 1], rc_Q31, 31);
      Anew[n] = sk_mul32_frac_q(tmp_QA, rc_mult2, mult2Q);
    }
  }
  if (Anew[0] > A_LIMIT || Anew[0] < -A_LIMIT) return 0;
  const i32 rc_Q31 = -sk_lshift(Anew[0], 31 - 24);
  const i32 rc_mult1_Q30 = (1 << 30) - sk_smmul(rc_Q31, rc_Q31);
  invGain_Q30 = sk_lshift(sk_smmul(invGain_Q30, rc_mult1_Q30), 2);
  return invGain_Q30;
}

// silk_NLSF2A (NLSF2A.c:44-178), QA = 16
CB_UTABLE u8 st_nlsf2a_ordering16[16] = {0, 15, 8, 7, 4, 11, 12, 3, 2, 13, 10, 5, 6, 9, 14, 1};
CB_UTABLE u8 st_nlsf2a_ordering10[10] = {0, 9, 6, 3, 4, 5, 8, 1, 2, 7};
CB_FN void sk_nlsf2a_find_poly(i32* out, const i32* cLSF, int dd) {
  out[0] = 1 << 16;
  out[1] = -cLSF[0];
  for (int k = 1; k < dd; k++) {
    const i32 ftmp = cLSF[2 * k];
    out[k + 1] = sk_lshift(out[k - 1], 1) - (i32)sk_rshift_round64((i64)ftmp * (i64)out[k], 16);
    for (int n = k; n > 1; n--) out[n] += out[n - 2] - (i32)sk_rshift_round64((i64)ftmp * (i64)out[n - 1], 16);
    out[1] -= ftmp;
  }
}
CB_FN void sk_nlsf2a(i16* a_Q12, const i16* NLSF, int d) {
  i32 cos_LSF_QA[SK_MAX_ORDER], P[SK_MAX_ORDER / 2 + 1], Q[SK_MAX_ORDER / 2 + 1], a32_QA1[SK_MAX_ORDER];
  const u8* ordering = d == 16 ? st_nlsf2a_ordering16 : st_nlsf2a_ordering10;
  for (int k = 0; k < d; k++) {
    const i32 f_int = NLSF[k] >> (15 - 7);
    const i32 f_frac = NLSF[k] - (f_int << (15 - 7));
    const i32 cos_val = st_LSFCosTab_Q12[f_int];
    const i32 delta = st_LSFCosTab_Q12[f_int + 1] - cos_val;
    cos_LSF_QA[ordering[k]] = sk_rshift_round(sk_lshift(cos_val, 8) + sk_mul(delta, f_frac), 20 - 16);
  }
  const int dd = d >> 1;
  sk_nlsf2a_find_poly(P, &cos_LSF_QA[0], dd);
  sk_nlsf2a_find_poly(Q, &cos_LSF_QA[1], dd);
  for (int k = 0; k < dd; k++) {
    const i32 Ptmp = P[k + 1] + P[k];
    const i32 Qtmp = Q[k + 1] - Q[k];
    a32_QA1[k] = -Qtmp - Ptmp;
    a32_QA1[d - k - 1] = Qtmp - Ptmp;
  }
  int i;
  for (i = 0; i < 10; i++) {
    i32 maxabs = 0, idx = 0;
    for (int k = 0; k < d; k++) {
      const i32 absval = sk_abs(a32_QA1[k]);
      if (absval > maxabs) { maxabs = absval; idx = k; }
    }
    maxabs = sk_rshift_round(maxabs, 16 + 1 - 12);
    if (maxabs > 32767) {
      maxabs = sk_min(maxabs, 163838);
      const i32 sc_Q16 = 65470 - sk_lshift(maxabs - 32767, 14) / (sk_mul(maxabs, idx + 1) >> 2);
      sk_bwexpander_32(a32_QA1, d, sc_Q16);
    } else {
      break;
    }
  }
  if (i == 10) {
    for (int k = 0; k < d; k++) {
      a_Q12[k] = (i16)sk_sat16(sk_rshift_round(a32_QA1[k], 16 + 1 - 12));
      a32_QA1[k] = sk_lshift((i32)a_Q12[k], 16 + 1 - 12);
    }
  } else {
    for (int k = 0; k < d; k++) a_Q12[k] = (i16)sk_rshift_round(a32_QA1[k], 16 + 1 - 12);
  }
  for (i = 0; i < 16; i++) {
    if (sk_lpc_inverse_pred_gain(a_Q12, d) < 107374) {            // SILK_FIX_CONST(1/1e4, 30)
      sk_bwexpander_32(a32_QA1, d, 65536 - sk_lshift(2, i));
      for (int k = 0; k < d; k++) a_Q12[k] = (i16)sk_rshift_round(a32_QA1[k], 16 + 1 - 12);
    } else {
      break;
    }
  }
}

// silk_decode_pitch (decode_pitch.c:36-77)
CB_FN void sk_decode_pitch(int lagIndex, int contourIndex, i32* pitch_lags, int Fs_kHz, int nb_subfr) {
  const i8* Lag_CB;
  int cbk_size;
  if (Fs_kHz == 8) {
    if (nb_subfr == 4) { Lag_CB = st_CB_lags_stage2; cbk_size = 11; }
    else { Lag_CB = st_CB_lags_stage2_10ms; cbk_size = 3; }
  } else {
    if (nb_subfr == 4) { Lag_CB = st_CB_lags_stage3; cbk_size = 34; }
    else { Lag_CB = st_CB_lags_stage3_10ms; cbk_size = 12; }
  }
  const int min_lag = sk_smulbb(2, Fs_kHz), max_lag = sk_smulbb(18, Fs_kHz);
  const int lag = min_lag + lagIndex;
  for (int k = 0; k < nb_subfr; k++) {
    pitch_lags[k] = lag + Lag_CB[k * cbk_size + contourIndex];
    pitch_lags[k] = sk_limit(pitch_lags[k], min_lag, max_lag);
  }
}

// silk_decode_parameters (decode_parameters.c:35-115)
CB_FN void sk_decode_parameters(SilkChannel& ps, SilkDecControl& ctrl, int condCoding) {
  i16 pNLSF_Q15[SK_MAX_ORDER], pNLSF0_Q15[SK_MAX_ORDER];
  sk_gains_dequant(ctrl.Gains_Q16, ps.indices.GainsIndices, &ps.LastGainIndex, condCoding == SK_CODE_CONDITIONALLY,
                   ps.nb_subfr);
  const SilkNLSFCB cb = sk_nlsf_cb(ps.LPC_order == 16);
  sk_nlsf_decode(pNLSF_Q15, ps.indices.NLSFIndices, cb);
  sk_nlsf2a(ctrl.PredCoef_Q12[1], pNLSF_Q15, ps.LPC_order);
  if (ps.first_frame_after_reset == 1) ps.indices.NLSFInterpCoef_Q2 = 4;
  if (ps.indices.NLSFInterpCoef_Q2 < 4) {
    for (int i = 0; i < ps.LPC_order; i++)
      pNLSF0_Q15[i] = (i16)(ps.prevNLSF_Q15[i] +
                            (sk_mul(ps.indices.NLSFInterpCoef_Q2, pNLSF_Q15[i] - ps.prevNLSF_Q15[i]) >> 2));
    sk_nlsf2a(ctrl.PredCoef_Q12[0], pNLSF0_Q15, ps.LPC_order);
  } else {
    for (int i = 0; i < ps.LPC_order; i++) ctrl.PredCoef_Q12[0][i] = ctrl.PredCoef_Q12[1][i];
  }
  for (int i = 0; i < ps.LPC_order; i++) ps.prevNLSF_Q15[i] = pNLSF_Q15[i];
  if (ps.lossCnt) {
    sk_bwexpander(ctrl.PredCoef_Q12[0], ps.LPC_order, 63570);     // BWE_AFTER_LOSS_Q16
    sk_bwexpander(ctrl.PredCoef_Q12[1], ps.LPC_order, 63570);
  }
  if (ps.indices.signalType == SK_TYPE_VOICED) {
    sk_decode_pitch(ps.indices.lagIndex, ps.indices.contourIndex, ctrl.pitchL, ps.fs_kHz, ps.nb_subfr);
    const i8* cbk = ps.indices.PERIndex == 0 ? st_LTP_gain_vq_0 : ps.indices.PERIndex == 1 ? st_LTP_gain_vq_1 : st_LTP_gain_vq_2;
    for (int k = 0; k < ps.nb_subfr; k++) {
      const int Ix = ps.indices.LTPIndex[k];
      for (int i = 0; i < SK_LTP_ORDER; i++) ctrl.LTPCoef_Q14[k * SK_LTP_ORDER + i] = (i16)sk_lshift(cbk[Ix * SK_LTP_ORDER + i], 7);
    }
    ctrl.LTP_scale_Q14 = st_LTPScales_Q14[ps.indices.LTP_scaleIndex];
  } else {
    for (int k = 0; k < ps.nb_subfr; k++) ctrl.pitchL[k] = 0;
    for (int k = 0; k < SK_LTP_ORDER * ps.nb_subfr; k++) ctrl.LTPCoef_Q14[k] = 0;
    ps.indices.PERIndex = 0;
    ctrl.LTP_scale_Q14 = 0;
  }
}

// silk_stereo_decode_pred / silk_stereo_decode_mid_only (stereo_decode_pred.c:35-73)
CB_FN void sk_stereo_decode_pred(RangeDec& rd, i32* pred_Q13) {
  int ix[2][3];
  int n = rd_icdf_g(rd, st_stereo_pred_joint_iCDF, 8);
  ix[0][2] = n / 5;
  ix[1][2] = n - 5 * ix[0][2];
  for (n = 0; n < 2; n++) {
    ix[n][0] = rd_icdf_g(rd, st_uniform3_iCDF, 8);
    ix[n][1] = rd_icdf_g(rd, st_uniform5_iCDF, 8);
  }
  for (n = 0; n < 2; n++) {
    ix[n][0] += 3 * ix[n][2];
    const i32 low_Q13 = st_stereo_pred_quant_Q13[ix[n][0]];
    const i32 step_Q13 = sk_smulwb(st_stereo_pred_quant_Q13[ix[n][0] + 1] - low_Q13, 6554);
    pred_Q13[n] = sk_smlabb(low_Q13, step_Q13, 2 * ix[n][1] + 1);
  }
  pred_Q13[0] -= pred_Q13[1];
}
