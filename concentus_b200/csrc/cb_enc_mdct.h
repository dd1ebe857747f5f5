// cb_enc_mdct.h -- forward MDCT and band-energy front end of the CELT encoder.
// Bit-exact with clt_mdct_forward_c (opus-fix/celt/mdct.c:119-258), compute_mdcts
// (celt/celt_encoder.c:419-462), compute_band_energies / normalise_bands (celt/bands.c:94-169)
// and amp2Log2 (celt/quant_bands.c:551-572).  The FFT stages are the ones the decoder uses.
#pragma once
#include "cb_common.h"
#include "cb_celt_synth.h"

// B interleaved forward MDCTs of NB output bins each.  in: channel base (B*NB + overlap samples);
// out[b + B*k]; f2: B*NB words of scratch.  sh = maxLM - LM (long) or maxLM (short blocks).
CB_FN_NOINLINE void enc_mdct_forward(const i32* in, i32* out, i32* f2, int B, int NB, int sh) {
  const int N2 = NB, N4 = NB >> 1;
  const int ov = CB_OVERLAP, q = (ov + 3) >> 2;
  const int trig_off = sh == 0 ? 0 : sh == 1 ? 960 : sh == 2 ? 1440 : 1680;
  const i16* trig = ct_mdct_tw960 + trig_off;
  const i16* bitrev = sh == 0 ? ct_fft_bitrev480 : sh == 1 ? ct_fft_bitrev240 : sh == 2 ? ct_fft_bitrev120 : ct_fft_bitrev60;
  const int scale_shift = 7 - sh;      // static_modes_fixed.h:434-489, minus one (mdct.c:131)
  CB_WSYNC();
  CB_LANE_LOOPU(u, B * N4) {           // window + fold + pre-rotation, scattered in bit-reversed order
    const int b = u / N4, i = u % N4;
    const i32* xp1 = in + b * NB + (ov >> 1) + 2 * i;
    const i32* xp2 = in + b * NB + N2 - 1 + (ov >> 1) - 2 * i;
    i32 re, im;
    if (i < q) {
      i32 w1 = cb_ldg(&ct_window120[(ov >> 1) + 2 * i]), w2 = cb_ldg(&ct_window120[(ov >> 1) - 1 - 2 * i]);
      re = add32(mult16_32_q15(w2, xp1[N2]), mult16_32_q15(w1, *xp2));
      im = sub32(mult16_32_q15(w1, *xp1), mult16_32_q15(w2, xp2[-N2]));
    } else if (i < N4 - q) {
      re = *xp2;
      im = *xp1;
    } else {
      const int k = i - (N4 - q);
      i32 w1 = cb_ldg(&ct_window120[2 * k]), w2 = cb_ldg(&ct_window120[ov - 1 - 2 * k]);
      re = add32(neg32(mult16_32_q15(w1, xp1[-N2])), mult16_32_q15(w2, *xp2));
      im = add32(mult16_32_q15(w2, *xp1), mult16_32_q15(w1, xp2[N2]));
    }
    i32 t0 = cb_ldg(&trig[i]), t1 = cb_ldg(&trig[N4 + i]);
    i32 yr = sub32(s_mul(re, t0), s_mul(im, t1));
    i32 yi = add32(s_mul(im, t0), s_mul(re, t1));
    yr = pshr32(mult16_32_q16(17476, yr), scale_shift);
    yi = pshr32(mult16_32_q16(17476, yi), scale_shift);
    const int rev = cb_ldg(&bitrev[i]);
    i32* yp = f2 + NB * b;
    yp[2 * rev] = yr;
    yp[2 * rev + 1] = yi;
  }
  fft_run(f2, B, NB, sh);
  CB_LANE_LOOPU(u, B * N4) {           // post-rotation into the interleaved spectrum
    const int b = u / N4, i = u % N4;
    const i32* fp = f2 + NB * b + 2 * i;
    i32 fr = fp[0], fi = fp[1];
    i32 t0 = cb_ldg(&trig[i]), t1 = cb_ldg(&trig[N4 + i]);
    out[b + B * (2 * i)] = sub32(s_mul(fi, t1), s_mul(fr, t0));
    out[b + B * (N2 - 1 - 2 * i)] = add32(s_mul(fr, t1), s_mul(fi, t0));
  }
  CB_WSYNC();
}

// compute_mdcts (celt_encoder.c:419-462).  in0/in1: channel bases (B*NB+overlap each); freq0/freq1 outputs.
CB_FN void enc_compute_mdcts(int shortBlocks, const i32* in0, const i32* in1, i32* freq0, i32* freq1, i32* f2,
                             int C, int CC, int LM, int upsample) {
  int B, NB, sh;
  if (shortBlocks) { B = shortBlocks; NB = CB_SHORT_MDCT; sh = CB_MAX_LM; }
  else { B = 1; NB = CB_SHORT_MDCT << LM; sh = CB_MAX_LM - LM; }
  enc_mdct_forward(in0, freq0, f2, B, NB, sh);
  if (CC == 2) enc_mdct_forward(in1, freq1, f2, B, NB, sh);
  const int BN = B * NB;
  if (CC == 2 && C == 1) {
    CB_LANE_LOOP(i, BN) freq0[i] = add32(half32(freq0[i]), half32(freq1[i]));
    CB_WSYNC();
  }
  if (upsample != 1) {
    const int bound = BN / upsample;
    CB_NOUNROLL for (int c = 0; c < C; c++) {
      i32* o = c ? freq1 : freq0;
      CB_LANE_LOOP(i, BN) o[i] = (i < bound) ? o[i] * upsample : 0;
    }
    CB_WSYNC();
  }
}

// compute_band_energies (bands.c:94-144): one band at a time, samples across lanes.
CB_FN_NOINLINE void enc_band_energies(const i32* freq0, const i32* freq1, i32* bandE, int end, int C, int LM) {
  CB_WSYNC();
  CB_NOUNROLL for (int c = 0; c < C; c++) {
    const i32* X = c ? freq1 : freq0;
    CB_NOUNROLL for (int i = 0; i < end; i++) {
      const int lo = ct_eband5ms[i] << LM, n = (ct_eband5ms[i + 1] - ct_eband5ms[i]) << LM;
      i32 mx = 0, mn = 0;
      CB_LANE_LOOP(j, n) { i32 v = X[lo + j]; mx = max32(mx, v); mn = min32(mn, v); }
      const i32 maxval = max32(cb_wmax(mx), neg32(cb_wmin(mn)));
      i32 e = 1;
      if (maxval > 0) {
        const int shift = cb_ilog2(maxval) - 14 + ((((int)ct_logN400[i] >> CB_BITRES) + LM + 1) >> 1);
        i32 s = 0;
        if (shift > 0) {
          CB_LANE_LOOP(j, n) { i32 v = (i16)(X[lo + j] >> shift); s = mac16_16(s, v, v); }
        } else {
          CB_LANE_LOOP(j, n) { i32 v = (i16)shl32(X[lo + j], -shift); s = mac16_16(s, v, v); }
        }
        s = cb_wsum(s);
        e = add32(1, vshr32(cb_sqrt(s), -shift));
      }
      bandE[i + c * CB_NBANDS] = e;
    }
  }
  CB_WSYNC();
}

// normalise_bands (bands.c:147-167)
CB_FN void enc_normalise_bands(const i32* freq0, const i32* freq1, i16* X, int N, const i32* bandE, int end, int C,
                               int M) {
  CB_WSYNC();
  CB_NOUNROLL for (int c = 0; c < C; c++) {
    const i32* F = c ? freq1 : freq0;
    CB_NOUNROLL for (int i = 0; i < end; i++) {
      const i32 be = bandE[i + c * CB_NBANDS];
      const int shift = cb_zlog2(be) - 13;
      const i32 E = (i16)vshr32(be, shift);
      const i32 g = (i16)cb_rcp(shl32(E, 3));
      const int lo = M * ct_eband5ms[i], n = M * (ct_eband5ms[i + 1] - ct_eband5ms[i]);
      CB_LANE_LOOP(j, n) X[c * N + lo + j] = (i16)mult16_16_q15(vshr32(F[lo + j], shift - 1), g);
    }
  }
  CB_WSYNC();
}

// amp2Log2 (quant_bands.c:551-572)
CB_FN_NOINLINE void enc_amp2log2(int effEnd, int end, const i32* bandE, i16* bandLogE, int C) {
  CB_WSYNC();
  CB_LANE_LOOP(u, C * end) {
    const int c = u / end, i = u % end;
    i16 v;
    if (i < effEnd) v = (i16)(cb_log2(shl32(bandE[i + c * CB_NBANDS], 2)) - shl16((i16)ct_eMeans[i], 6));
    else v = (i16)(-14 * 1024);
    bandLogE[i + c * CB_NBANDS] = v;
  }
  CB_WSYNC();
}
