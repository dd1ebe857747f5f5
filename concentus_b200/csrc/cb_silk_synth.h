// cb_silk_synth.h -- SILK signal reconstruction: excitation -> LTP/LPC synthesis (decode_core),
// packet-loss concealment, comfort noise, mid/side to left/right, and the output resampler.
// Bit-exact with silk/decode_core.c, LPC_analysis_filter.c (FIXED_POINT form through celt_fir,
// celt/celt_lpc.c:95-149), PLC.c, CNG.c, stereo_MS_to_LR.c, resampler.c, resampler_private_up2_HQ.c,
// resampler_private_IIR_FIR.c, resampler_private_down_FIR.c, resampler_private_AR2.c.
// Scalar code: one thread per stream; the work arrays are per-thread (local memory).
#pragma once
#include "cb_silk_params.h"

// per-thread work arrays of one SILK frame
struct alignas(16) SilkWork {
  i32 exc_Q14[SK_MAX_FRAME];               // the frame's excitation; stored to the channel state in one go
  i16 pulses[SK_MAX_FRAME];
  i16 sLTP[SK_MAX_FRAME];
  i32 sLTP_Q15[2 * SK_MAX_FRAME];          // decode_core: Q15; PLC: Q14
  i32 res_Q14[SK_MAX_SUBFR];
  i32 sLPC_Q14[SK_MAX_SUBFR + SK_MAX_ORDER];
  i32 CNG_sig_Q10[SK_MAX_FRAME + SK_MAX_ORDER];
  i16 exc_buf[2 * SK_MAX_SUBFR];
};

// silk_LPC_analysis_filter (LPC_analysis_filter.c:43-110, FIXED_POINT branch)
// The last D inputs stay in registers (h[ix mod D] = in[ix]; the loops unroll so every index is a
// constant): one load per sample instead of D.  The sum is a wrapping 32-bit one, so its order is free.
template <int D>
CB_FN void sk_lpc_analysis_filter_t(i16* out, const i16* in, const i16* B, int len) {
  i32 c[D], h[D];
#pragma unroll
  for (int j = 0; j < D; j++) {
    c[j] = (i32)(i16)(-B[j]);
    h[j] = in[j];
    out[j] = 0;
  }
  for (int base = D; base < len; base += D) {
#pragma unroll
    for (int u = 0; u < D; u++) {
      const int ix = base + u;
      if (ix < len) {
        const i32 x = in[ix];
        i32 s0 = 0, s1 = 0;
#pragma unroll
        for (int j = 0; j < D; j += 2) {
          s0 = sk_add(s0, sk_mul(c[j], h[(u - 1 - j + 2 * D) % D]));
          s1 = sk_add(s1, sk_mul(c[j + 1], h[(u - 2 - j + 2 * D) % D]));
        }
        out[ix] = (i16)sk_sat16(sk_add(x, pshr32(sk_add(s0, s1), 12)));
        h[u] = x;
      }
    }
  }
}
CB_FN void sk_lpc_analysis_filter(i16* out, const i16* in, const i16* B, int len, int d) {
  if (d == 16) { sk_lpc_analysis_filter_t<16>(out, in, B, len); return; }
  if (d == 10) { sk_lpc_analysis_filter_t<10>(out, in, B, len); return; }
  for (int ix = d; ix < len; ix++) {
    i32 sum = 0;
    for (int j = 0; j < d; j++) sum = sk_add(sum, sk_mul((i32)(i16)(-B[j]), (i32)in[ix - 1 - j]));
    out[ix] = (i16)sk_sat16(sk_add((i32)in[ix], pshr32(sum, 12)));
  }
  for (int j = 0; j < d; j++) out[j] = 0;
}

// LPC synthesis of one subframe (decode_core.c:197-226) with the last D outputs in registers.
// s = sLPC_Q14 (SK_MAX_ORDER history values in front, filled by the caller); len % D == 0.
template <int D>
CB_FN void sk_lpc_synthesis_t(i32* s, const i32* pres_Q14, i16* pxq, const i16* A_Q12, i32 Gain_Q10, int len) {
  i32 a[D], h[D];
#pragma unroll
  for (int j = 0; j < D; j++) {
    a[j] = A_Q12[j];
    h[j] = s[SK_MAX_ORDER - D + j];        // sample index -D + j; slot j = index mod D as len % D == 0
  }
  for (int base = 0; base < len; base += D) {
    i32 pr[D];
#pragma unroll
    for (int u = 0; u < D; u++) pr[u] = pres_Q14[base + u];      // the block's inputs: loads overlap
#pragma unroll
    for (int u = 0; u < D; u++) {
      const int i = base + u;
      i32 p0 = D >> 1, p1 = 0;
#pragma unroll
      for (int j = 0; j < D; j += 2) {
        p0 = sk_smlawb(p0, h[(u - 1 - j + 2 * D) % D], a[j]);
        p1 = sk_smlawb(p1, h[(u - 2 - j + 2 * D) % D], a[j + 1]);
      }
      const i32 v = sk_add(pr[u], sk_lshift(sk_add(p0, p1), 4));
      h[u] = v;
      s[SK_MAX_ORDER + i] = v;
      pxq[i] = (i16)sk_sat16(sk_rshift_round(sk_smulww(v, Gain_Q10), 8));
    }
  }
}

// silk_decode_core (decode_core.c:38-239)
CB_FN void sk_decode_core(SilkChannel& ps, SilkDecControl& ctrl, i16* xq, SilkWork& w) {
  const i16* pulses = w.pulses;
  i16* sLTP = w.sLTP;
  i32* sLTP_Q15 = w.sLTP_Q15;
  i32* sLPC_Q14 = w.sLPC_Q14;
  int lag = 0;
  const i32 offset_Q10 = st_quant_offsets_Q10[(ps.indices.signalType >> 1) * 2 + ps.indices.quantOffsetType];
  const int NLSF_interpolation_flag = ps.indices.NLSFInterpCoef_Q2 < 4;
  i32 rand_seed = ps.indices.Seed;
  for (int i0 = 0; i0 < ps.frame_length; i0 += 4) {            // frame lengths are multiples of 40
    const i32 q[4] = {pulses[i0], pulses[i0 + 1], pulses[i0 + 2], pulses[i0 + 3]};   // loads first: they overlap
#pragma unroll
    for (int u = 0; u < 4; u++) {
      rand_seed = sk_rand(rand_seed);
      i32 e = sk_lshift(q[u], 14);
      if (e > 0) e -= 80 << 4;
      else if (e < 0) e += 80 << 4;
      e += offset_Q10 << 4;
      if (rand_seed < 0) e = -e;
      w.exc_Q14[i0 + u] = e;
      rand_seed = sk_add(rand_seed, q[u]);
    }
  }
  sk_copy_i32(ps.exc_Q14, w.exc_Q14, ps.frame_length);
  for (int i = 0; i < SK_MAX_ORDER; i++) sLPC_Q14[i] = ps.sLPC_Q14_buf[i];
  i32* pexc_Q14 = w.exc_Q14;
  i16* pxq = xq;
  int sLTP_buf_idx = ps.ltp_mem_length;
  for (int k = 0; k < ps.nb_subfr; k++) {
    i32* pres_Q14 = w.res_Q14;
    const i16* A_Q12 = ctrl.PredCoef_Q12[k >> 1];
    i16* B_Q14 = &ctrl.LTPCoef_Q14[k * SK_LTP_ORDER];
    int signalType = ps.indices.signalType;
    const i32 Gain_Q10 = ctrl.Gains_Q16[k] >> 6;
    i32 inv_gain_Q31 = sk_inverse32_varQ(ctrl.Gains_Q16[k], 47);
    i32 gain_adj_Q16;
    if (ctrl.Gains_Q16[k] != ps.prev_gain_Q16) {
      gain_adj_Q16 = sk_div32_varQ(ps.prev_gain_Q16, ctrl.Gains_Q16[k], 16);
      for (int i = 0; i < SK_MAX_ORDER; i++) sLPC_Q14[i] = sk_smulww(gain_adj_Q16, sLPC_Q14[i]);
    } else {
      gain_adj_Q16 = 1 << 16;
    }
    ps.prev_gain_Q16 = ctrl.Gains_Q16[k];
    if (ps.lossCnt && ps.prevSignalType == SK_TYPE_VOICED && ps.indices.signalType != SK_TYPE_VOICED && k < 2) {
      for (int i = 0; i < SK_LTP_ORDER; i++) B_Q14[i] = 0;
      B_Q14[SK_LTP_ORDER / 2] = 4096;
      signalType = SK_TYPE_VOICED;
      ctrl.pitchL[k] = ps.lagPrev;
    }
    if (signalType == SK_TYPE_VOICED) {
      lag = ctrl.pitchL[k];
      if (k == 0 || (k == 2 && NLSF_interpolation_flag)) {
        const int start_idx = ps.ltp_mem_length - lag - ps.LPC_order - SK_LTP_ORDER / 2;
        if (k == 2) {
          sk_copy_i16(&ps.outBuf[ps.ltp_mem_length], xq, 2 * ps.subfr_length);
        }
        sk_lpc_analysis_filter(&sLTP[start_idx], &ps.outBuf[start_idx + k * ps.subfr_length], A_Q12,
                               ps.ltp_mem_length - start_idx, ps.LPC_order);
        if (k == 0) inv_gain_Q31 = sk_lshift(sk_smulwb(inv_gain_Q31, ctrl.LTP_scale_Q14), 2);
        {
          const int n = lag + SK_LTP_ORDER / 2;
          int i = 0;
          for (; i + 4 <= n; i += 4) {
            const i32 v0 = sLTP[ps.ltp_mem_length - i - 1], v1 = sLTP[ps.ltp_mem_length - i - 2];
            const i32 v2 = sLTP[ps.ltp_mem_length - i - 3], v3 = sLTP[ps.ltp_mem_length - i - 4];
            sLTP_Q15[sLTP_buf_idx - i - 1] = sk_smulwb(inv_gain_Q31, v0);
            sLTP_Q15[sLTP_buf_idx - i - 2] = sk_smulwb(inv_gain_Q31, v1);
            sLTP_Q15[sLTP_buf_idx - i - 3] = sk_smulwb(inv_gain_Q31, v2);
            sLTP_Q15[sLTP_buf_idx - i - 4] = sk_smulwb(inv_gain_Q31, v3);
          }
          for (; i < n; i++) sLTP_Q15[sLTP_buf_idx - i - 1] = sk_smulwb(inv_gain_Q31, sLTP[ps.ltp_mem_length - i - 1]);
        }
      } else if (gain_adj_Q16 != 1 << 16) {
        const int n = lag + SK_LTP_ORDER / 2;
        int i = 0;
        for (; i + 4 <= n; i += 4) {
          i32* q = &sLTP_Q15[sLTP_buf_idx - i - 4];
          const i32 v0 = q[0], v1 = q[1], v2 = q[2], v3 = q[3];
          q[0] = sk_smulww(gain_adj_Q16, v0); q[1] = sk_smulww(gain_adj_Q16, v1);
          q[2] = sk_smulww(gain_adj_Q16, v2); q[3] = sk_smulww(gain_adj_Q16, v3);
        }
        for (; i < n; i++) sLTP_Q15[sLTP_buf_idx - i - 1] = sk_smulww(gain_adj_Q16, sLTP_Q15[sLTP_buf_idx - i - 1]);
      }
    }
    if (signalType == SK_TYPE_VOICED) {
      // the five taps slide through registers; the incoming one is loaded an iteration ahead (it was
      // written lag - 3 >= 13 samples ago, so the look-ahead never passes the store below)
      const i32* pred_lag_ptr = &sLTP_Q15[sLTP_buf_idx - lag + SK_LTP_ORDER / 2];
      const i32 b0 = B_Q14[0], b1 = B_Q14[1], b2 = B_Q14[2], b3 = B_Q14[3], b4 = B_Q14[4];
      i32 t0 = pred_lag_ptr[0], t1 = pred_lag_ptr[-1], t2 = pred_lag_ptr[-2], t3 = pred_lag_ptr[-3], t4 = pred_lag_ptr[-4];
      i32 n1 = pred_lag_ptr[1], n2 = pred_lag_ptr[2], n3 = pred_lag_ptr[3], n4 = pred_lag_ptr[4];
      i32 e0 = pexc_Q14[0], e1 = pexc_Q14[1], e2 = pexc_Q14[2], e3 = pexc_Q14[3];
      for (int i = 0; i < ps.subfr_length; i += 4) {            // subframe lengths are multiples of 4; lag >= 16
        // four samples per trip, and the inputs of the next trip are fetched before this one computes: they
        // were written at least lag - 10 > 0 samples ago (past the subframe's end the values are not used)
        const i32 m1 = pred_lag_ptr[i + 5], m2 = pred_lag_ptr[i + 6], m3 = pred_lag_ptr[i + 7], m4 = pred_lag_ptr[i + 8];
        const i32 f0 = pexc_Q14[i + 4], f1 = pexc_Q14[i + 5], f2 = pexc_Q14[i + 6], f3 = pexc_Q14[i + 7];
        i32 p;
        p = sk_smlawb(sk_smlawb(sk_smlawb(sk_smlawb(sk_smlawb(2, t0, b0), t1, b1), t2, b2), t3, b3), t4, b4);
        const i32 r0 = sk_add(e0, sk_lshift(p, 1));
        p = sk_smlawb(sk_smlawb(sk_smlawb(sk_smlawb(sk_smlawb(2, n1, b0), t0, b1), t1, b2), t2, b3), t3, b4);
        const i32 r1 = sk_add(e1, sk_lshift(p, 1));
        p = sk_smlawb(sk_smlawb(sk_smlawb(sk_smlawb(sk_smlawb(2, n2, b0), n1, b1), t0, b2), t1, b3), t2, b4);
        const i32 r2 = sk_add(e2, sk_lshift(p, 1));
        p = sk_smlawb(sk_smlawb(sk_smlawb(sk_smlawb(sk_smlawb(2, n3, b0), n2, b1), n1, b2), t0, b3), t1, b4);
        const i32 r3 = sk_add(e3, sk_lshift(p, 1));
        t4 = t0; t3 = n1; t2 = n2; t1 = n3; t0 = n4;
        n1 = m1; n2 = m2; n3 = m3; n4 = m4;
        e0 = f0; e1 = f1; e2 = f2; e3 = f3;
        pres_Q14[i] = r0; pres_Q14[i + 1] = r1; pres_Q14[i + 2] = r2; pres_Q14[i + 3] = r3;
        sLTP_Q15[sLTP_buf_idx] = sk_lshift(r0, 1); sLTP_Q15[sLTP_buf_idx + 1] = sk_lshift(r1, 1);
        sLTP_Q15[sLTP_buf_idx + 2] = sk_lshift(r2, 1); sLTP_Q15[sLTP_buf_idx + 3] = sk_lshift(r3, 1);
        sLTP_buf_idx += 4;
      }
    } else {
      pres_Q14 = pexc_Q14;
    }
    if (ps.LPC_order == 16 && ps.subfr_length % 16 == 0) {
      sk_lpc_synthesis_t<16>(sLPC_Q14, pres_Q14, pxq, A_Q12, Gain_Q10, ps.subfr_length);
    } else if (ps.LPC_order == 10 && ps.subfr_length % 10 == 0) {
      sk_lpc_synthesis_t<10>(sLPC_Q14, pres_Q14, pxq, A_Q12, Gain_Q10, ps.subfr_length);
    } else {
      for (int i = 0; i < ps.subfr_length; i++) {
        i32 LPC_pred_Q10 = ps.LPC_order >> 1;
        for (int j = 0; j < ps.LPC_order; j++)
          LPC_pred_Q10 = sk_smlawb(LPC_pred_Q10, sLPC_Q14[SK_MAX_ORDER + i - 1 - j], A_Q12[j]);
        sLPC_Q14[SK_MAX_ORDER + i] = sk_add(pres_Q14[i], sk_lshift(LPC_pred_Q10, 4));
        pxq[i] = (i16)sk_sat16(sk_rshift_round(sk_smulww(sLPC_Q14[SK_MAX_ORDER + i], Gain_Q10), 8));
      }
    }
    for (int i = 0; i < SK_MAX_ORDER; i++) sLPC_Q14[i] = sLPC_Q14[ps.subfr_length + i];
    pexc_Q14 += ps.subfr_length;
    pxq += ps.subfr_length;
  }
  for (int i = 0; i < SK_MAX_ORDER; i++) ps.sLPC_Q14_buf[i] = sLPC_Q14[i];
}

// ---- packet-loss concealment (PLC.c) ----------------------------------------------------------
CB_FN void sk_plc_reset(SilkChannel& ps) {                       // PLC.c:54-63
  ps.sPLC.pitchL_Q8 = sk_lshift(ps.frame_length, 8 - 1);
  ps.sPLC.prevGain_Q16[0] = 65536;
  ps.sPLC.prevGain_Q16[1] = 65536;
  ps.sPLC.subfr_length = 20;
  ps.sPLC.nb_subfr = 2;
}
CB_FN void sk_plc_update(SilkChannel& ps, SilkDecControl& ctrl) {   // PLC.c:99-168
  SilkPLC& pl = ps.sPLC;
  ps.prevSignalType = ps.indices.signalType;
  i32 LTP_Gain_Q14 = 0;
  if (ps.indices.signalType == SK_TYPE_VOICED) {
    for (int j = 0; j * ps.subfr_length < ctrl.pitchL[ps.nb_subfr - 1]; j++) {
      if (j == ps.nb_subfr) break;
      i32 temp = 0;
      for (int i = 0; i < SK_LTP_ORDER; i++) temp += ctrl.LTPCoef_Q14[(ps.nb_subfr - 1 - j) * SK_LTP_ORDER + i];
      if (temp > LTP_Gain_Q14) {
        LTP_Gain_Q14 = temp;
        for (int i = 0; i < SK_LTP_ORDER; i++)
          pl.LTPCoef_Q14[i] = ctrl.LTPCoef_Q14[sk_smulbb(ps.nb_subfr - 1 - j, SK_LTP_ORDER) + i];
        pl.pitchL_Q8 = sk_lshift(ctrl.pitchL[ps.nb_subfr - 1 - j], 8);
      }
    }
    for (int i = 0; i < SK_LTP_ORDER; i++) pl.LTPCoef_Q14[i] = 0;
    pl.LTPCoef_Q14[SK_LTP_ORDER / 2] = (i16)LTP_Gain_Q14;
    if (LTP_Gain_Q14 < 11469) {
      const i32 tmp = sk_lshift(11469, 10);
      const i32 scale_Q10 = tmp / sk_max(LTP_Gain_Q14, 1);
      for (int i = 0; i < SK_LTP_ORDER; i++) pl.LTPCoef_Q14[i] = (i16)(sk_smulbb(pl.LTPCoef_Q14[i], scale_Q10) >> 10);
    } else if (LTP_Gain_Q14 > 15565) {
      const i32 tmp = sk_lshift(15565, 14);
      const i32 scale_Q14 = tmp / sk_max(LTP_Gain_Q14, 1);
      for (int i = 0; i < SK_LTP_ORDER; i++) pl.LTPCoef_Q14[i] = (i16)(sk_smulbb(pl.LTPCoef_Q14[i], scale_Q14) >> 14);
    }
  } else {
    pl.pitchL_Q8 = sk_lshift(sk_smulbb(ps.fs_kHz, 18), 8);
    for (int i = 0; i < SK_LTP_ORDER; i++) pl.LTPCoef_Q14[i] = 0;
  }
  for (int i = 0; i < ps.LPC_order; i++) pl.prevLPC_Q12[i] = ctrl.PredCoef_Q12[1][i];
  pl.prevLTP_scale_Q14 = (i16)ctrl.LTP_scale_Q14;
  pl.prevGain_Q16[0] = ctrl.Gains_Q16[ps.nb_subfr - 2];
  pl.prevGain_Q16[1] = ctrl.Gains_Q16[ps.nb_subfr - 1];
  pl.subfr_length = ps.subfr_length;
  pl.nb_subfr = ps.nb_subfr;
}

CB_FN void sk_plc_conceal(SilkChannel& ps, SilkDecControl& ctrl, i16* frame, SilkWork& w) {   // PLC.c:202-395
  SilkPLC& pl = ps.sPLC;
  i32* sLTP_Q14 = w.sLTP_Q15;
  i16* sLTP = w.sLTP;
  i16 A_Q12[SK_MAX_ORDER];
  i32 prevGain_Q10[2];
  prevGain_Q10[0] = pl.prevGain_Q16[0] >> 6;
  prevGain_Q10[1] = pl.prevGain_Q16[1] >> 6;
  if (ps.first_frame_after_reset) {
    for (int i = 0; i < SK_MAX_ORDER; i++) pl.prevLPC_Q12[i] = 0;
  }
  // silk_PLC_energy (PLC.c:171-199)
  i32 energy1, energy2;
  int shift1, shift2;
  {
    i16* exc_buf_ptr = w.exc_buf;
    for (int k = 0; k < 2; k++) {
      for (int i = 0; i < ps.subfr_length; i++)
        exc_buf_ptr[i] = (i16)sk_sat16(sk_smulww(ps.exc_Q14[i + (k + ps.nb_subfr - 2) * ps.subfr_length], prevGain_Q10[k]) >> 8);
      exc_buf_ptr += ps.subfr_length;
    }
    sk_sum_sqr_shift(&energy1, &shift1, w.exc_buf, ps.subfr_length);
    sk_sum_sqr_shift(&energy2, &shift2, &w.exc_buf[ps.subfr_length], ps.subfr_length);
  }
  const i32* rand_ptr;
  if ((energy1 >> shift2) < (energy2 >> shift1))
    rand_ptr = &ps.exc_Q14[sk_max(0, (pl.nb_subfr - 1) * pl.subfr_length - 128)];
  else
    rand_ptr = &ps.exc_Q14[sk_max(0, pl.nb_subfr * pl.subfr_length - 128)];
  i16* B_Q14 = pl.LTPCoef_Q14;
  i32 rand_scale_Q14 = pl.randScale_Q14;
  const int att = sk_min(2 - 1, ps.lossCnt);
  const i32 harm_Gain_Q15 = att == 0 ? 32440 : 31130;
  i32 rand_Gain_Q15;
  if (ps.prevSignalType == SK_TYPE_VOICED) rand_Gain_Q15 = att == 0 ? 31130 : 26214;
  else rand_Gain_Q15 = att == 0 ? 32440 : 29491;
  sk_bwexpander(pl.prevLPC_Q12, ps.LPC_order, 64881);            // SILK_FIX_CONST(BWE_COEF, 16)
  for (int i = 0; i < ps.LPC_order; i++) A_Q12[i] = pl.prevLPC_Q12[i];
  if (ps.lossCnt == 0) {
    rand_scale_Q14 = 1 << 14;
    if (ps.prevSignalType == SK_TYPE_VOICED) {
      for (int i = 0; i < SK_LTP_ORDER; i++) rand_scale_Q14 = (i16)(rand_scale_Q14 - B_Q14[i]);
      rand_scale_Q14 = (i16)sk_max(3277, rand_scale_Q14);
      rand_scale_Q14 = (i16)(sk_smulbb(rand_scale_Q14, pl.prevLTP_scale_Q14) >> 14);
    } else {
      const i32 invGain_Q30 = sk_lpc_inverse_pred_gain(pl.prevLPC_Q12, ps.LPC_order);
      i32 down_scale_Q30 = sk_min((1 << 30) >> 3, invGain_Q30);
      down_scale_Q30 = sk_max((1 << 30) >> 8, down_scale_Q30);
      down_scale_Q30 = sk_lshift(down_scale_Q30, 3);
      rand_Gain_Q15 = sk_smulwb(down_scale_Q30, rand_Gain_Q15) >> 14;
    }
  }
  i32 rand_seed = pl.rand_seed;
  int lag = sk_rshift_round(pl.pitchL_Q8, 8);
  int sLTP_buf_idx = ps.ltp_mem_length;
  int idx = ps.ltp_mem_length - lag - ps.LPC_order - SK_LTP_ORDER / 2;
  sk_lpc_analysis_filter(&sLTP[idx], &ps.outBuf[idx], A_Q12, ps.ltp_mem_length - idx, ps.LPC_order);
  i32 inv_gain_Q30 = sk_inverse32_varQ(pl.prevGain_Q16[1], 46);
  inv_gain_Q30 = sk_min(inv_gain_Q30, 0x7fffffff >> 1);
  for (int i = idx + ps.LPC_order; i < ps.ltp_mem_length; i++) sLTP_Q14[i] = sk_smulwb(inv_gain_Q30, sLTP[i]);
  for (int k = 0; k < ps.nb_subfr; k++) {
    i32* pred_lag_ptr = &sLTP_Q14[sLTP_buf_idx - lag + SK_LTP_ORDER / 2];
    for (int i = 0; i < ps.subfr_length; i++) {
      i32 LTP_pred_Q12 = 2;
      LTP_pred_Q12 = sk_smlawb(LTP_pred_Q12, pred_lag_ptr[0], B_Q14[0]);
      LTP_pred_Q12 = sk_smlawb(LTP_pred_Q12, pred_lag_ptr[-1], B_Q14[1]);
      LTP_pred_Q12 = sk_smlawb(LTP_pred_Q12, pred_lag_ptr[-2], B_Q14[2]);
      LTP_pred_Q12 = sk_smlawb(LTP_pred_Q12, pred_lag_ptr[-3], B_Q14[3]);
      LTP_pred_Q12 = sk_smlawb(LTP_pred_Q12, pred_lag_ptr[-4], B_Q14[4]);
      pred_lag_ptr++;
      rand_seed = sk_rand(rand_seed);
      idx = (rand_seed >> 25) & 127;
      sLTP_Q14[sLTP_buf_idx] = sk_lshift(sk_smlawb(LTP_pred_Q12, rand_ptr[idx], rand_scale_Q14), 2);
      sLTP_buf_idx++;
    }
    for (int j = 0; j < SK_LTP_ORDER; j++) B_Q14[j] = (i16)(sk_smulbb(harm_Gain_Q15, B_Q14[j]) >> 15);
    rand_scale_Q14 = (i16)(sk_smulbb(rand_scale_Q14, rand_Gain_Q15) >> 15);
    pl.pitchL_Q8 = sk_smlawb(pl.pitchL_Q8, pl.pitchL_Q8, 655);
    pl.pitchL_Q8 = sk_min(pl.pitchL_Q8, sk_lshift(sk_smulbb(18, ps.fs_kHz), 8));
    lag = sk_rshift_round(pl.pitchL_Q8, 8);
  }
  i32* sLPC_Q14_ptr = &sLTP_Q14[ps.ltp_mem_length - SK_MAX_ORDER];
  for (int i = 0; i < SK_MAX_ORDER; i++) sLPC_Q14_ptr[i] = ps.sLPC_Q14_buf[i];
  for (int i = 0; i < ps.frame_length; i++) {
    i32 LPC_pred_Q10 = ps.LPC_order >> 1;
    for (int j = 0; j < ps.LPC_order; j++)
      LPC_pred_Q10 = sk_smlawb(LPC_pred_Q10, sLPC_Q14_ptr[SK_MAX_ORDER + i - 1 - j], A_Q12[j]);
    sLPC_Q14_ptr[SK_MAX_ORDER + i] = sk_add(sLPC_Q14_ptr[SK_MAX_ORDER + i], sk_lshift(LPC_pred_Q10, 4));
    frame[i] = (i16)sk_sat16(sk_sat16(sk_rshift_round(sk_smulww(sLPC_Q14_ptr[SK_MAX_ORDER + i], prevGain_Q10[1]), 8)));
  }
  for (int i = 0; i < SK_MAX_ORDER; i++) ps.sLPC_Q14_buf[i] = sLPC_Q14_ptr[ps.frame_length + i];
  pl.rand_seed = rand_seed;
  pl.randScale_Q14 = (i16)rand_scale_Q14;
  for (int i = 0; i < SK_MAX_NB_SUBFR; i++) ctrl.pitchL[i] = lag;
}

CB_FN void sk_plc(SilkChannel& ps, SilkDecControl& ctrl, i16* frame, int lost, SilkWork& w) {   // PLC.c:66-93
  if (ps.fs_kHz != ps.sPLC.fs_kHz) {
    sk_plc_reset(ps);
    ps.sPLC.fs_kHz = ps.fs_kHz;
  }
  if (lost) {
    sk_plc_conceal(ps, ctrl, frame, w);
    ps.lossCnt++;
  } else {
    sk_plc_update(ps, ctrl);
  }
}

CB_FN void sk_plc_glue_frames(SilkChannel& ps, i16* frame, int length) {   // PLC.c:398-445
  SilkPLC& pl = ps.sPLC;
  if (ps.lossCnt) {
    int sh;
    sk_sum_sqr_shift(&pl.conc_energy, &sh, frame, length);
    pl.conc_energy_shift = sh;
    pl.last_frame_lost = 1;
  } else {
    if (pl.last_frame_lost) {
      i32 energy;
      int energy_shift;
      sk_sum_sqr_shift(&energy, &energy_shift, frame, length);
      if (energy_shift > pl.conc_energy_shift) pl.conc_energy = pl.conc_energy >> (energy_shift - pl.conc_energy_shift);
      else if (energy_shift < pl.conc_energy_shift) energy = energy >> (pl.conc_energy_shift - energy_shift);
      if (energy > pl.conc_energy) {
        i32 LZ = sk_clz32(pl.conc_energy);
        LZ = LZ - 1;
        pl.conc_energy = sk_lshift(pl.conc_energy, LZ);
        energy = energy >> sk_max(24 - LZ, 0);
        const i32 frac_Q24 = pl.conc_energy / sk_max(energy, 1);
        i32 gain_Q16 = sk_lshift(sk_sqrt_approx(frac_Q24), 4);
        i32 slope_Q16 = ((1 << 16) - gain_Q16) / length;
        slope_Q16 = sk_lshift(slope_Q16, 2);
        for (int i = 0; i < length; i++) {
          frame[i] = (i16)sk_smulwb(gain_Q16, frame[i]);
          gain_Q16 += slope_Q16;
          if (gain_Q16 > 1 << 16) break;
        }
      }
    }
    pl.last_frame_lost = 0;
  }
}

// ---- comfort noise (CNG.c) --------------------------------------------------------------------
CB_FN void sk_cng_reset(SilkChannel& ps) {                        // CNG.c:63-79
  const int step = 32767 / (ps.LPC_order + 1);
  int acc = 0;
  for (int i = 0; i < ps.LPC_order; i++) {
    acc += step;
    ps.sCNG.CNG_smth_NLSF_Q15[i] = (i16)acc;
  }
  ps.sCNG.CNG_smth_Gain_Q16 = 0;
  ps.sCNG.rand_seed = 3176576;
}
// Put CNG_exc_buf_Q14 back into the reference's linear layout (rotate left by ring_head slots, three reversals).
CB_FN void sk_cng_linearise(SilkCNG& cng) {
  if (cng.ring_head == 0) return;
  i32* b = cng.CNG_exc_buf_Q14;
  const int sl = cng.ring_sub, n = 4 * sl, k = cng.ring_head * sl;
  for (int lo = 0, hi = k - 1; lo < hi; lo++, hi--) { const i32 t = b[lo]; b[lo] = b[hi]; b[hi] = t; }
  for (int lo = k, hi = n - 1; lo < hi; lo++, hi--) { const i32 t = b[lo]; b[lo] = b[hi]; b[hi] = t; }
  for (int lo = 0, hi = n - 1; lo < hi; lo++, hi--) { const i32 t = b[lo]; b[lo] = b[hi]; b[hi] = t; }
  cng.ring_head = 0;
}
CB_FN void sk_cng(SilkChannel& ps, SilkDecControl& ctrl, i16* frame, int length, SilkWork& w) {   // CNG.c:82-181
  SilkCNG& cng = ps.sCNG;
  i16 A_Q12[SK_MAX_ORDER];
  if (ps.fs_kHz != cng.fs_kHz) {
    sk_cng_linearise(cng);
    sk_cng_reset(ps);
    cng.fs_kHz = ps.fs_kHz;
  }
  if (ps.lossCnt == 0 && ps.prevSignalType == SK_TYPE_NO_VOICE) {
    for (int i = 0; i < ps.LPC_order; i++)
      cng.CNG_smth_NLSF_Q15[i] = (i16)(cng.CNG_smth_NLSF_Q15[i] +
                                       sk_smulwb((i32)ps.prevNLSF_Q15[i] - (i32)cng.CNG_smth_NLSF_Q15[i], 16348));
    i32 max_Gain_Q16 = 0;
    int subfr = 0;
    for (int i = 0; i < ps.nb_subfr; i++) {
      if (ctrl.Gains_Q16[i] > max_Gain_Q16) { max_Gain_Q16 = ctrl.Gains_Q16[i]; subfr = i; }
    }
    if (ps.nb_subfr == SK_MAX_NB_SUBFR) {
      // CNG.c:122-123 on the ring: the new subframe takes the slot of the oldest, which becomes logical slot 0
      if (cng.ring_sub != ps.subfr_length) {
        sk_cng_linearise(cng);
        cng.ring_sub = ps.subfr_length;
      }
      cng.ring_head = (cng.ring_head + 3) & 3;
      sk_copy_i32(&cng.CNG_exc_buf_Q14[cng.ring_head * ps.subfr_length], &ps.exc_Q14[subfr * ps.subfr_length], ps.subfr_length);
    } else {
      sk_cng_linearise(cng);
      sk_copy_i32_down(&cng.CNG_exc_buf_Q14[ps.subfr_length], cng.CNG_exc_buf_Q14, (ps.nb_subfr - 1) * ps.subfr_length);
      sk_copy_i32(cng.CNG_exc_buf_Q14, &ps.exc_Q14[subfr * ps.subfr_length], ps.subfr_length);
    }
    for (int i = 0; i < ps.nb_subfr; i++)
      cng.CNG_smth_Gain_Q16 += sk_smulwb(ctrl.Gains_Q16[i] - cng.CNG_smth_Gain_Q16, 4634);
  }
  if (ps.lossCnt) {
    i32* CNG_sig_Q10 = w.CNG_sig_Q10;
    i32 gain_Q16 = sk_smulww(ps.sPLC.randScale_Q14, ps.sPLC.prevGain_Q16[1]);
    if (gain_Q16 >= (1 << 21) || cng.CNG_smth_Gain_Q16 > (1 << 23)) {
      gain_Q16 = sk_smultt(gain_Q16, gain_Q16);
      gain_Q16 = sk_sub(sk_smultt(cng.CNG_smth_Gain_Q16, cng.CNG_smth_Gain_Q16), sk_lshift(gain_Q16, 5));
      gain_Q16 = sk_lshift(sk_sqrt_approx(gain_Q16), 16);
    } else {
      gain_Q16 = sk_smulww(gain_Q16, gain_Q16);
      gain_Q16 = sk_sub(sk_smulww(cng.CNG_smth_Gain_Q16, cng.CNG_smth_Gain_Q16), sk_lshift(gain_Q16, 5));
      gain_Q16 = sk_lshift(sk_sqrt_approx(gain_Q16), 8);
    }
    {                                                             // silk_CNG_exc (CNG.c:36-61)
      int exc_mask = 255;
      while (exc_mask > length) exc_mask >>= 1;
      i32 seed = cng.rand_seed;
      const int ring_head = cng.ring_head, ring_sub = cng.ring_sub;
      for (int i = 0; i < length; i++) {
        seed = sk_rand(seed);
        int idx = (seed >> 24) & exc_mask;
        if (ring_head) {                                          // logical -> ring position
          const int slot = (idx >= ring_sub) + (idx >= 2 * ring_sub) + (idx >= 3 * ring_sub);
          idx = (((ring_head + slot) & 3) - slot) * ring_sub + idx;
        }
        CNG_sig_Q10[SK_MAX_ORDER + i] = (i16)sk_sat16(sk_smulww(cng.CNG_exc_buf_Q14[idx], gain_Q16 >> 4));
      }
      cng.rand_seed = seed;
    }
    sk_nlsf2a(A_Q12, cng.CNG_smth_NLSF_Q15, ps.LPC_order);
    for (int i = 0; i < SK_MAX_ORDER; i++) CNG_sig_Q10[i] = cng.CNG_synth_state[i];
    for (int i = 0; i < length; i++) {
      i32 sum_Q6 = ps.LPC_order >> 1;
      for (int j = 0; j < ps.LPC_order; j++) sum_Q6 = sk_smlawb(sum_Q6, CNG_sig_Q10[SK_MAX_ORDER + i - 1 - j], A_Q12[j]);
      CNG_sig_Q10[SK_MAX_ORDER + i] = sk_add(CNG_sig_Q10[SK_MAX_ORDER + i], sk_lshift(sum_Q6, 4));
      frame[i] = (i16)sk_add_sat16(frame[i], sk_rshift_round(CNG_sig_Q10[SK_MAX_ORDER + i], 10));
    }
    for (int i = 0; i < SK_MAX_ORDER; i++) cng.CNG_synth_state[i] = CNG_sig_Q10[length + i];
  } else {
    for (int i = 0; i < ps.LPC_order; i++) cng.CNG_synth_state[i] = 0;
  }
}

// ---- mid/side to left/right (stereo_MS_to_LR.c:35-85); x1/x2 have 2 samples of history in front ------
CB_FN void sk_stereo_ms_to_lr(SilkStereo& st, i16* x1, i16* x2, const i32* pred_Q13, int fs_kHz, int frame_length) {
  x1[0] = st.sMid[0]; x1[1] = st.sMid[1];
  x2[0] = st.sSide[0]; x2[1] = st.sSide[1];
  st.sMid[0] = x1[frame_length]; st.sMid[1] = x1[frame_length + 1];
  st.sSide[0] = x2[frame_length]; st.sSide[1] = x2[frame_length + 1];
  i32 pred0_Q13 = st.pred_prev_Q13[0], pred1_Q13 = st.pred_prev_Q13[1];
  const int denom_Q16 = (1 << 16) / (8 * fs_kHz);
  const i32 delta0_Q13 = sk_rshift_round(sk_smulbb(pred_Q13[0] - st.pred_prev_Q13[0], denom_Q16), 16);
  const i32 delta1_Q13 = sk_rshift_round(sk_smulbb(pred_Q13[1] - st.pred_prev_Q13[1], denom_Q16), 16);
  for (int n = 0; n < frame_length; n++) {
    if (n < 8 * fs_kHz) { pred0_Q13 += delta0_Q13; pred1_Q13 += delta1_Q13; }
    else { pred0_Q13 = pred_Q13[0]; pred1_Q13 = pred_Q13[1]; }
    i32 sum = sk_lshift((x1[n] + x1[n + 2]) + sk_lshift(x1[n + 1], 1), 9);
    sum = sk_smlawb(sk_lshift((i32)x2[n + 1], 8), sum, pred0_Q13);
    sum = sk_smlawb(sum, sk_lshift((i32)x1[n + 1], 11), pred1_Q13);
    x2[n + 1] = (i16)sk_sat16(sk_rshift_round(sum, 8));
  }
  st.pred_prev_Q13[0] = (i16)pred_Q13[0];
  st.pred_prev_Q13[1] = (i16)pred_Q13[1];
  for (int n = 0; n < frame_length; n++) {
    const i32 sum = x1[n + 1] + (i32)x2[n + 1];
    const i32 diff = x1[n + 1] - (i32)x2[n + 1];
    x1[n + 1] = (i16)sk_sat16(sum);
    x2[n + 1] = (i16)sk_sat16(diff);
  }
}

// ---- resampler, decoder side (resampler.c, resampler_private_*.c) ------------------------------
#define SK_RS_COPY 0
#define SK_RS_UP2 1
#define SK_RS_IIR_FIR 2
#define SK_RS_DOWN_FIR 3
CB_UTABLE i8 st_delay_matrix_dec[3][5] = {{4, 0, 2, 0, 0}, {0, 9, 4, 7, 4}, {0, 3, 12, 7, 7}};   // resampler.c:65-70
CB_FN int sk_rate_id(i32 R) { return ((((R >> 12) - (R > 16000)) >> (R > 24000)) - 1); }          // resampler.c:73

// silk_resampler_init with forEnc = 0 (resampler.c:81-172)
CB_FN int sk_resampler_init(SilkResampler& S, i32 Fs_Hz_in, i32 Fs_Hz_out) {
  {
    unsigned char* p = (unsigned char*)&S;
    for (size_t i = 0; i < sizeof(SilkResampler); i++) p[i] = 0;
  }
  if ((Fs_Hz_in != 8000 && Fs_Hz_in != 12000 && Fs_Hz_in != 16000) ||
      (Fs_Hz_out != 8000 && Fs_Hz_out != 12000 && Fs_Hz_out != 16000 && Fs_Hz_out != 24000 && Fs_Hz_out != 48000))
    return -1;
  S.inputDelay = st_delay_matrix_dec[sk_rate_id(Fs_Hz_in)][sk_rate_id(Fs_Hz_out)];
  S.Fs_in_kHz = Fs_Hz_in / 1000;
  S.Fs_out_kHz = Fs_Hz_out / 1000;
  S.batchSize = S.Fs_in_kHz * 10;
  int up2x = 0;
  if (Fs_Hz_out > Fs_Hz_in) {
    if (Fs_Hz_out == Fs_Hz_in * 2) {
      S.function = SK_RS_UP2;
    } else {
      S.function = SK_RS_IIR_FIR;
      up2x = 1;
    }
  } else if (Fs_Hz_out < Fs_Hz_in) {
    S.function = SK_RS_DOWN_FIR;
    if (Fs_Hz_out * 4 == Fs_Hz_in * 3) { S.FIR_Fracs = 3; S.FIR_Order = 18; S.coefs_id = 1; }
    else if (Fs_Hz_out * 3 == Fs_Hz_in * 2) { S.FIR_Fracs = 2; S.FIR_Order = 18; S.coefs_id = 2; }
    else if (Fs_Hz_out * 2 == Fs_Hz_in) { S.FIR_Fracs = 1; S.FIR_Order = 24; S.coefs_id = 3; }
    else return -1;
  } else {
    S.function = SK_RS_COPY;
  }
  S.invRatio_Q16 = sk_lshift(sk_lshift(Fs_Hz_in, 14 + up2x) / Fs_Hz_out, 2);
  while (sk_smulww(S.invRatio_Q16, Fs_Hz_out) < sk_lshift(Fs_Hz_in, up2x)) S.invRatio_Q16++;
  return 0;
}

// silk_resampler_private_up2_HQ (resampler_private_up2_HQ.c:38-105)
CB_FN void sk_up2_hq(i32* S, i16* out, const i16* in, int len) {
  const i32 c00 = st_resampler_up2_hq_0[0], c01 = st_resampler_up2_hq_0[1], c02 = st_resampler_up2_hq_0[2];
  const i32 c10 = st_resampler_up2_hq_1[0], c11 = st_resampler_up2_hq_1[1], c12 = st_resampler_up2_hq_1[2];
  for (int k = 0; k < len; k++) {
    const i32 in32 = sk_lshift((i32)in[k], 10);
    i32 Y = sk_sub(in32, S[0]);
    i32 X = sk_smulwb(Y, c00);
    i32 out32_1 = sk_add(S[0], X);
    S[0] = sk_add(in32, X);
    Y = sk_sub(out32_1, S[1]);
    X = sk_smulwb(Y, c01);
    i32 out32_2 = sk_add(S[1], X);
    S[1] = sk_add(out32_1, X);
    Y = sk_sub(out32_2, S[2]);
    X = sk_smlawb(Y, Y, c02);
    out32_1 = sk_add(S[2], X);
    S[2] = sk_add(out32_2, X);
    out[2 * k] = (i16)sk_sat16(sk_rshift_round(out32_1, 10));
    Y = sk_sub(in32, S[3]);
    X = sk_smulwb(Y, c10);
    out32_1 = sk_add(S[3], X);
    S[3] = sk_add(in32, X);
    Y = sk_sub(out32_1, S[4]);
    X = sk_smulwb(Y, c11);
    out32_2 = sk_add(S[4], X);
    S[4] = sk_add(out32_1, X);
    Y = sk_sub(out32_2, S[5]);
    X = sk_smlawb(Y, Y, c12);
    out32_1 = sk_add(S[5], X);
    S[5] = sk_add(out32_2, X);
    out[2 * k + 1] = (i16)sk_sat16(sk_rshift_round(out32_1, 10));
  }
}

// per-thread buffers of the resampler
struct SilkRsWork {
  i16 buf16[2 * 160 + 8];          // IIR_FIR: 2*batchSize + RESAMPLER_ORDER_FIR_12
  i32 buf32[160 + 24];             // down_FIR: batchSize + FIR_Order
};

// silk_resampler_private_IIR_FIR (resampler_private_IIR_FIR.c:36-107); returns the advanced out pointer
CB_FN i16* sk_rs_iir_fir(SilkResampler& S, i16* out, int ostride, const i16* in, int inLen, SilkRsWork& rw) {
  i16* buf = rw.buf16;
  i16* sFIR16 = (i16*)S.sFIR;
  for (int i = 0; i < 8; i++) buf[i] = sFIR16[i];
  const i32 index_increment_Q16 = S.invRatio_Q16;
  int nSamplesIn;
  for (;;) {
    nSamplesIn = sk_min(inLen, S.batchSize);
    sk_up2_hq(S.sIIR, &buf[8], in, nSamplesIn);
    const i32 max_index_Q16 = sk_lshift(nSamplesIn, 16 + 1);
    for (i32 index_Q16 = 0; index_Q16 < max_index_Q16; index_Q16 += index_increment_Q16) {
      const int table_index = sk_smulwb(index_Q16 & 0xFFFF, 12);
      const i16* bp = &buf[index_Q16 >> 16];
      const i16* f0 = &st_resampler_frac_FIR_12[4 * table_index];
      const i16* f1 = &st_resampler_frac_FIR_12[4 * (11 - table_index)];
      i32 res_Q15 = sk_smulbb(bp[0], f0[0]);
      res_Q15 = sk_smlabb(res_Q15, bp[1], f0[1]);
      res_Q15 = sk_smlabb(res_Q15, bp[2], f0[2]);
      res_Q15 = sk_smlabb(res_Q15, bp[3], f0[3]);
      res_Q15 = sk_smlabb(res_Q15, bp[4], f1[3]);
      res_Q15 = sk_smlabb(res_Q15, bp[5], f1[2]);
      res_Q15 = sk_smlabb(res_Q15, bp[6], f1[1]);
      res_Q15 = sk_smlabb(res_Q15, bp[7], f1[0]);
      *out = (i16)sk_sat16(sk_rshift_round(res_Q15, 15));
      out += ostride;
    }
    in += nSamplesIn;
    inLen -= nSamplesIn;
    if (inLen > 0) {
      for (int i = 0; i < 8; i++) buf[i] = buf[(nSamplesIn << 1) + i];
    } else {
      break;
    }
  }
  for (int i = 0; i < 8; i++) sFIR16[i] = buf[(nSamplesIn << 1) + i];
  return out;
}

// silk_resampler_private_down_FIR (resampler_private_down_FIR.c:36-194) with AR2 (resampler_private_AR2.c)
CB_FN i16* sk_rs_down_fir(SilkResampler& S, i16* out, int ostride, const i16* in, int inLen, SilkRsWork& rw) {
  i32* buf = rw.buf32;
  const i16* Coefs = S.coefs_id == 1 ? st_Resampler_3_4_COEFS : S.coefs_id == 2 ? st_Resampler_2_3_COEFS : st_Resampler_1_2_COEFS;
  const i16* FIR_Coefs = &Coefs[2];
  for (int i = 0; i < S.FIR_Order; i++) buf[i] = S.sFIR[i];
  const i32 index_increment_Q16 = S.invRatio_Q16;
  int nSamplesIn;
  for (;;) {
    nSamplesIn = sk_min(inLen, S.batchSize);
    {                                                             // AR2, output Q8
      i32* o = &buf[S.FIR_Order];
      for (int k = 0; k < nSamplesIn; k++) {
        i32 out32 = sk_add(S.sIIR[0], sk_lshift((i32)in[k], 8));
        o[k] = out32;
        out32 = sk_lshift(out32, 2);
        S.sIIR[0] = sk_smlawb(S.sIIR[1], out32, Coefs[0]);
        S.sIIR[1] = sk_smulwb(out32, Coefs[1]);
      }
    }
    const i32 max_index_Q16 = sk_lshift(nSamplesIn, 16);
    if (S.FIR_Order == 18) {
      for (i32 index_Q16 = 0; index_Q16 < max_index_Q16; index_Q16 += index_increment_Q16) {
        const i32* bp = buf + (index_Q16 >> 16);
        const int interpol_ind = sk_smulwb(index_Q16 & 0xFFFF, S.FIR_Fracs);
        const i16* ip = &FIR_Coefs[9 * interpol_ind];
        i32 res_Q6 = sk_smulwb(bp[0], ip[0]);
        for (int t = 1; t < 9; t++) res_Q6 = sk_smlawb(res_Q6, bp[t], ip[t]);
        ip = &FIR_Coefs[9 * (S.FIR_Fracs - 1 - interpol_ind)];
        for (int t = 0; t < 9; t++) res_Q6 = sk_smlawb(res_Q6, bp[17 - t], ip[t]);
        *out = (i16)sk_sat16(sk_rshift_round(res_Q6, 6));
        out += ostride;
      }
    } else {                                                      // FIR_Order 24
      for (i32 index_Q16 = 0; index_Q16 < max_index_Q16; index_Q16 += index_increment_Q16) {
        const i32* bp = buf + (index_Q16 >> 16);
        i32 res_Q6 = sk_smulwb(sk_add(bp[0], bp[23]), FIR_Coefs[0]);
        for (int t = 1; t < 12; t++) res_Q6 = sk_smlawb(res_Q6, sk_add(bp[t], bp[23 - t]), FIR_Coefs[t]);
        *out = (i16)sk_sat16(sk_rshift_round(res_Q6, 6));
        out += ostride;
      }
    }
    in += nSamplesIn;
    inLen -= nSamplesIn;
    if (inLen > 1) {
      for (int i = 0; i < S.FIR_Order; i++) buf[i] = buf[nSamplesIn + i];
    } else {
      break;
    }
  }
  for (int i = 0; i < S.FIR_Order; i++) S.sFIR[i] = buf[nSamplesIn + i];
  return out;
}

// silk_resampler (resampler.c:176-215).  Output written with stride `ostride` (interleaving folded in).
CB_FN void sk_resampler(SilkResampler& Sg, i16* out, int ostride, const i16* in, int inLen, SilkRsWork& rw) {
  // the filter memories are touched once per sample: work on a private copy, store the memories back once
  SilkResampler S = Sg;
  const int nSamples = S.Fs_in_kHz - S.inputDelay;
  for (int i = 0; i < nSamples; i++) S.delayBuf[S.inputDelay + i] = in[i];
  if (S.function == SK_RS_UP2) {
    // up2 writes 2 samples per input: go through a small strided loop to keep one code path
    i16 tmp[2];
    i16* o = out;
    for (int k = 0; k < S.Fs_in_kHz; k++) { sk_up2_hq(S.sIIR, tmp, &S.delayBuf[k], 1); o[0] = tmp[0]; o[ostride] = tmp[1]; o += 2 * ostride; }
    const i16* p = &in[nSamples];
    for (int k = 0; k < inLen - S.Fs_in_kHz; k++) { sk_up2_hq(S.sIIR, tmp, &p[k], 1); o[0] = tmp[0]; o[ostride] = tmp[1]; o += 2 * ostride; }
  } else if (S.function == SK_RS_IIR_FIR) {
    i16* o = sk_rs_iir_fir(S, out, ostride, S.delayBuf, S.Fs_in_kHz, rw);
    (void)o;
    sk_rs_iir_fir(S, out + ostride * S.Fs_out_kHz, ostride, &in[nSamples], inLen - S.Fs_in_kHz, rw);
  } else if (S.function == SK_RS_DOWN_FIR) {
    sk_rs_down_fir(S, out, ostride, S.delayBuf, S.Fs_in_kHz, rw);
    sk_rs_down_fir(S, out + ostride * S.Fs_out_kHz, ostride, &in[nSamples], inLen - S.Fs_in_kHz, rw);
  } else {
    for (int i = 0; i < S.Fs_in_kHz; i++) out[ostride * i] = S.delayBuf[i];
    {
      const int n = inLen - S.Fs_in_kHz;
      i16* o = out + ostride * S.Fs_out_kHz;
      const i16* p = in + nSamples;
      int i = 0;
      for (; i + 8 <= n; i += 8) {                                 // loads first: their latencies overlap
        const i16 v0 = p[i], v1 = p[i + 1], v2 = p[i + 2], v3 = p[i + 3], v4 = p[i + 4], v5 = p[i + 5], v6 = p[i + 6], v7 = p[i + 7];
        o[ostride * i] = v0; o[ostride * (i + 1)] = v1; o[ostride * (i + 2)] = v2; o[ostride * (i + 3)] = v3;
        o[ostride * (i + 4)] = v4; o[ostride * (i + 5)] = v5; o[ostride * (i + 6)] = v6; o[ostride * (i + 7)] = v7;
      }
      for (; i < n; i++) o[ostride * i] = p[i];
    }
  }
  for (int i = 0; i < S.inputDelay; i++) S.delayBuf[i] = in[inLen - S.inputDelay + i];
  for (int i = 0; i < 6; i++) Sg.sIIR[i] = S.sIIR[i];
  for (int i = 0; i < 36; i++) Sg.sFIR[i] = S.sFIR[i];
  for (int i = 0; i < 48; i++) Sg.delayBuf[i] = S.delayBuf[i];
}
