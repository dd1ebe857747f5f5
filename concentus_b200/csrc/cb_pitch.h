// cb_pitch.h -- pitch analysis shared by the decoder's concealment and the encoder's pre-filter:
// autocorrelation + Levinson-Durbin, the whitening FIR of pitch_downsample, the two-stage
// decimated pitch search and its candidate ranking.  Bit-exact with opus-fix/celt/celt_lpc.c:37-93,
// 232-328 and celt/pitch.c:45-145,225-369.  Inner products are lane loops folded with redux
// (wrap-around int32 sums are order independent); the candidate ranking is serial.
#pragma once
#include "cb_common.h"
#include "celt_tables.h"

#ifndef CB_OVERLAP
#define CB_OVERLAP 120
#endif
#ifndef CB_LPC_ORDER
#define CB_LPC_ORDER 24
#endif
#ifndef CB_SIG_SHIFT
#define CB_SIG_SHIFT 12
#endif

// _celt_autocorr (celt_lpc.c:232-328) without the SIMD split: ac[k] = sum_{i>=k} x[i]x[i-k], then
// the fixed-point normalisation.  x: n samples; xx: scratch of n samples.
CB_FN int pitch_autocorr(const i16* x, i32* ac, i16* xx, int use_window, int lag, int n) {
  const i16* xptr = x;
  CB_WSYNC();
  if (use_window) {
    CB_LANE_LOOP(i, n) xx[i] = x[i];
    CB_WSYNC();
    CB_LANE_LOOP(i, CB_OVERLAP) {
      i32 w = cb_ldg(&ct_window120[i]);
      xx[i] = (i16)mult16_16_q15(x[i], w);
      xx[n - i - 1] = (i16)mult16_16_q15(x[n - i - 1], w);
    }
    CB_WSYNC();
    xptr = xx;
  }
  i32 part = 0;
  CB_LANE_LOOP(i, n) part = add32(part, mult16_16(xptr[i], xptr[i]) >> 9);
  i32 ac0 = add32(1 + (n << 7), cb_wsum(part));
  int shift = cb_ilog2(ac0) - 30 + 10;
  shift = shift / 2;
  if (shift > 0) {
    CB_WSYNC();
    CB_LANE_LOOP(i, n) {
      i32 v = xptr[i];
      xx[i] = (i16)pshr32(v, shift);
    }
    CB_WSYNC();
    xptr = xx;
  } else {
    shift = 0;
  }
  CB_NOUNROLL for (int k = 0; k <= lag; k++) {
    i32 d = 0;
    CB_LANE_LOOP(t, n - k) d = mac16_16(d, xptr[t + k], xptr[t]);
    ac[k] = cb_wsum(d);
  }
  CB_WSYNC();
  shift = 2 * shift;
  i32 a0 = ac[0];
  if (shift <= 0) a0 = add32(a0, shl32(1, -shift));
  CB_WSYNC();
  ac[0] = a0;
  CB_WSYNC();
  if (a0 < 268435456) {
    int shift2 = 29 - cb_ilog((u32)a0);
    CB_LANE_LOOP(i, lag + 1) ac[i] = shl32(ac[i], shift2);
    shift -= shift2;
  } else if (a0 >= 536870912) {
    int shift2 = 1;
    if (a0 >= 1073741824) shift2++;
    CB_LANE_LOOP(i, lag + 1) ac[i] = ac[i] >> shift2;
    shift += shift2;
  }
  CB_WSYNC();
  return shift;
}

// _celt_lpc (celt_lpc.c:37-93): Levinson-Durbin, uniform (p <= 24)
CB_FN void pitch_lpc(i16* out, const i32* ac, i32* lpc, int p) {
  CB_NOUNROLL for (int i = 0; i < p; i++) lpc[i] = 0;
  i32 error = ac[0];
  if (ac[0] != 0) {
    CB_NOUNROLL for (int i = 0; i < p; i++) {
      i32 rr = 0;
      CB_NOUNROLL for (int j = 0; j < i; j++) rr = add32(rr, mult32_32_q31(lpc[j], ac[i - j]));
      rr = add32(rr, ac[i + 1] >> 3);
      i32 r = neg32(cb_frac_div32(shl32(rr, 3), error));
      lpc[i] = r >> 3;
      CB_NOUNROLL for (int j = 0; j < (i + 1) >> 1; j++) {
        i32 t1 = lpc[j], t2 = lpc[i - 1 - j];
        lpc[j] = add32(t1, mult32_32_q31(r, t2));
        lpc[i - 1 - j] = add32(t2, mult32_32_q31(r, t1));
      }
      error = sub32(error, mult32_32_q31(mult32_32_q31(r, r), error));
      if (error < (ac[0] >> 10)) break;
    }
  }
  CB_NOUNROLL for (int i = 0; i < p; i++) out[i] = round16(lpc[i], 16);
}

// find_best_pitch (pitch.c:45-103), uniform
CB_FN void pitch_find_best(const i32* xcorr, const i16* y, int len, int max_pitch, int* best_pitch,
                               int yshift, i32 maxcorr) {
  i32 Syy = 1;
  i32 best_num0 = -1, best_num1 = -1;
  i32 best_den0 = 0, best_den1 = 0;
  const int xshift = cb_ilog2(maxcorr) - 14;
  best_pitch[0] = 0;
  best_pitch[1] = 1;
  {
    i32 part = 0;
    CB_LANE_LOOP(j, len) part = add32(part, mult16_16(y[j], y[j]) >> yshift);
    Syy = add32(Syy, cb_wsum(part));
  }
  CB_NOUNROLL for (int i = 0; i < max_pitch; i++) {
    if (xcorr[i] > 0) {
      i32 xcorr16 = (i16)vshr32(xcorr[i], xshift);
      i32 num = (i16)mult16_16_q15(xcorr16, xcorr16);
      if (mult16_32_q15(num, best_den1) > mult16_32_q15(best_num1, Syy)) {
        if (mult16_32_q15(num, best_den0) > mult16_32_q15(best_num0, Syy)) {
          best_num1 = best_num0; best_den1 = best_den0; best_pitch[1] = best_pitch[0];
          best_num0 = num; best_den0 = Syy; best_pitch[0] = i;
        } else {
          best_num1 = num; best_den1 = Syy; best_pitch[1] = i;
        }
      }
    }
    Syy = add32(Syy, sub32(mult16_16(y[i + len], y[i + len]) >> yshift, mult16_16(y[i], y[i]) >> yshift));
    Syy = max32(1, Syy);
  }
}


// Tail of pitch_downsample (pitch.c:183-221): 4th-order LPC whitening + one zero, as a 5-tap FIR
// on the inputs (computed out of place through xx).  x_lp: n samples.
CB_FN void pitch_whiten(i16* x_lp, int n, i32* ac, i32* lpc32, i16* xx) {
  pitch_autocorr(x_lp, ac, xx, 0, 4, n);
  i16 lpc[4];
  {
    i32 a[5];
    for (int i = 0; i < 5; i++) a[i] = ac[i];
    a[0] = add32(a[0], a[0] >> 13);
    for (int i = 1; i <= 4; i++) a[i] = sub32(a[i], mult16_32_q15(2 * i * i, a[i]));
    CB_WSYNC();
    for (int i = 0; i < 5; i++) ac[i] = a[i];
    CB_WSYNC();
    i16 l4[4];
    pitch_lpc(l4, ac, lpc32, 4);
    i32 tmp = 32767;
    for (int i = 0; i < 4; i++) {
      tmp = (i16)mult16_16_q15(29491, tmp);            // QCONST16(.9f,15)
      lpc[i] = (i16)mult16_16_q15(l4[i], tmp);
    }
  }
  const i32 c1 = 26214;                                // QCONST16(.8f,15)
  i16 n0 = (i16)(lpc[0] + 3277);                       // QCONST16(.8f,SIG_SHIFT)
  i16 n1 = (i16)(lpc[1] + mult16_16_q15(c1, lpc[0]));
  i16 n2 = (i16)(lpc[2] + mult16_16_q15(c1, lpc[1]));
  i16 n3 = (i16)(lpc[3] + mult16_16_q15(c1, lpc[2]));
  i16 n4 = (i16)mult16_16_q15(c1, lpc[3]);
  CB_WSYNC();
  CB_LANE_LOOP(i, n) {                                 // celt_fir5 (pitch.c:105-145), zero memory
    i32 sum = shl32((i32)x_lp[i], CB_SIG_SHIFT);
    sum = mac16_16(sum, n0, i >= 1 ? x_lp[i - 1] : 0);
    sum = mac16_16(sum, n1, i >= 2 ? x_lp[i - 2] : 0);
    sum = mac16_16(sum, n2, i >= 3 ? x_lp[i - 3] : 0);
    sum = mac16_16(sum, n3, i >= 4 ? x_lp[i - 4] : 0);
    sum = mac16_16(sum, n4, i >= 5 ? x_lp[i - 5] : 0);
    xx[i] = round16(sum, CB_SIG_SHIFT);
  }
  CB_WSYNC();
  CB_LANE_LOOP(i, n) x_lp[i] = xx[i];
  CB_WSYNC();
}

// pitch_search (pitch.c:260-369).  xs: len>>1 valid samples (2x decimated target), y: (len+max_pitch)>>1
// samples; x4/y4/xcorr: scratch of len>>2, (len+max_pitch)>>2 and max_pitch>>1 entries.
CB_FN int pitch_search(const i16* xs, const i16* y, int len, int max_pitch, i16* x4, i16* y4, i32* xcorr) {
  const int lag = len + max_pitch;
  i32 xm = 0, xn = 0;
  CB_WSYNC();
  CB_LANE_LOOP(j, len >> 2) { i16 v = xs[2 * j]; x4[j] = v; xm = max32(xm, v); xn = min32(xn, v); }
  i32 ym = 0, yn = 0;
  CB_LANE_LOOP(j, lag >> 2) { i16 v = y[2 * j]; y4[j] = v; ym = max32(ym, v); yn = min32(yn, v); }
  i32 xmax = max32(cb_wmax(xm), neg32(cb_wmin(xn)));
  i32 ymax = max32(cb_wmax(ym), neg32(cb_wmin(yn)));
  int sh = cb_ilog2(max32(1, max32(xmax, ymax))) - 11;
  CB_WSYNC();
  if (sh > 0) {
    CB_LANE_LOOP(j, len >> 2) x4[j] = (i16)(x4[j] >> sh);
    CB_LANE_LOOP(j, lag >> 2) y4[j] = (i16)(y4[j] >> sh);
    sh *= 2;
  } else {
    sh = 0;
  }
  CB_WSYNC();
  i32 maxcorr = 1;                                     // coarse search, 4x decimation (pitch.c:225-258)
  CB_NOUNROLL for (int i = 0; i < (max_pitch >> 2); i++) {
    i32 d = 0;
    CB_LANE_LOOP(j, len >> 2) d = mac16_16(d, x4[j], y4[i + j]);
    d = cb_wsum(d);
    xcorr[i] = d;
    maxcorr = max32(maxcorr, d);
  }
  CB_WSYNC();
  int best_pitch[2] = {0, 0};
  pitch_find_best(xcorr, y4, len >> 2, max_pitch >> 2, best_pitch, 0, maxcorr);
  maxcorr = 1;                                         // finer search, 2x decimation
  CB_WSYNC();
  CB_NOUNROLL for (int i = 0; i < (max_pitch >> 1); i++) {
    int d0 = i - 2 * best_pitch[0], d1 = i - 2 * best_pitch[1];
    if (d0 < 0) d0 = -d0;
    if (d1 < 0) d1 = -d1;
    i32 val = 0;
    if (!(d0 > 2 && d1 > 2)) {
      i32 sum = 0;
      CB_LANE_LOOP(j, len >> 1) sum = add32(sum, mult16_16(xs[j], y[i + j]) >> sh);
      sum = cb_wsum(sum);
      val = max32(-1, sum);
      maxcorr = max32(maxcorr, sum);
    }
    xcorr[i] = val;
  }
  CB_WSYNC();
  pitch_find_best(xcorr, y, len >> 1, max_pitch >> 1, best_pitch, sh + 1, maxcorr);
  int offset = 0;
  if (best_pitch[0] > 0 && best_pitch[0] < (max_pitch >> 1) - 1) {
    i32 a = xcorr[best_pitch[0] - 1], b = xcorr[best_pitch[0]], c = xcorr[best_pitch[0] + 1];
    if (sub32(c, a) > mult16_32_q15(22938, sub32(b, a))) offset = 1;          // QCONST16(.7f,15)
    else if (sub32(a, c) > mult16_32_q15(22938, sub32(b, c))) offset = -1;
  }
  return 2 * best_pitch[0] - offset;
}
