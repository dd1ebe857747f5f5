// cb_celt_synth.h -- kernel B body: spectrum -> PCM for one channel of one stream (one warp).
//
//   denormalise_bands -> inverse MDCT (pre-rotation, mixed-radix fixed-point FFT as shared-memory
//   stages, post-rotation, TDAC mirror) -> pitch post-filter (in-place comb, processed in waves
//   no longer than the lag) -> ring-buffer store -> de-emphasis -> int16 PCM.
// Bit-exact with celt_synthesis()/deemphasis() (opus-fix/celt/celt_decoder.c:185-350),
// clt_mdct_backward_c (celt/mdct.c:263-363), opus_fft_impl + kf_bfly{2,3,4,5}
// (celt/kiss_fft.c:51-330,532-578) and comb_filter (celt/celt.c:156-244).
// All butterflies of one FFT stage are independent, so a stage is one lane loop; the 8 short
// transforms of a transient frame touch disjoint sample ranges and run side by side.
#pragma once
#include "cb_common.h"
#include "cb_state.h"
#include "celt_tables.h"

struct alignas(16) SynWork {
  i32 freq[CB_MAXN];                 // denormalised MDCT coefficients of this channel
  i32 buf[CB_MAXN + CB_OVERLAP / 2]; // out_syn[0 .. N+overlap/2)
};

CB_UTABLE i16 ct_comb_gains[3][3] = {{10048, 7112, 4248}, {15200, 8784, 0}, {26208, 3280, 0}};  // celt.c:191-194

// S_MUL (celt/_kiss_fft_guts.h:57): 32-bit sample times 16-bit twiddle, Q15
CB_FN i32 s_mul(i32 a, i32 b) { return mult16_32_q15(b, a); }

// denormalise_bands (bands.c:169-238) for one channel into dst[0..N)
// denormalise_bands (bands.c:169-238).  The per-band gains are worked out one band per lane and parked in
// `tmp` (2 x 21 words of free workspace); the spectrum is then one flat pass over the bins, whose loads do
// not wait on each other -- a loop over the bands made 21 dependent round trips to the hand-off record.
CB_FN void v_denormalise(const i16* X, i32* dst, const i16* bandLogE, int start, int end, int M,
                         int downsample, int silence, i32* tmp) {
  const int N = M * 120;
  int bound = M * ct_eband5ms[end];
  if (downsample != 1) bound = cb_imin(bound, N / downsample);
  if (silence) { bound = 0; start = end = 0; }
  const int jlo = M * ct_eband5ms[start];
  const int LM = 31 - cb_clz((u32)M);
  CB_WSYNC();
  CB_LANE_LOOP(i, CB_NBANDS) {
    i32 g = 0;
    int shift = 0;
    if (i >= start && i < end) {
      const i16 lg = add16(bandLogE[i], shl16((i16)cb_ldg(&ct_eMeans_g[i]), 6));
      shift = 16 - (lg >> CB_DB_SHIFT);
      if (shift > 31) { shift = 0; g = 0; }
      else g = (i16)cb_exp2_frac(lg & ((1 << CB_DB_SHIFT) - 1));
      if (shift < -2) { g = 32767; shift = -2; }
    }
    tmp[i] = g;
    tmp[CB_NBANDS + i] = shift;
  }
  CB_WSYNC();
  CB_LANE_LOOPU(j, N) {
    i32 v = 0;
    if (j >= jlo && j < bound) {
      const int b = cb_ldg(&ct_band_of_bin_g[j >> LM]);
      const i32 g = tmp[b];
      const int shift = tmp[CB_NBANDS + b];
      const i32 p = mult16_16(X[j], g);
      v = shift < 0 ? shl32(p, -shift) : p >> shift;
    }
    dst[j] = v;
  }
  CB_WSYNC();
}

// ---- FFT stages on interleaved (re,im) int32 data; `nb` transforms of `nfft` points laid out
// `bstride` ints apart.  G groups of m butterflies, group stride mm, twiddle step fs. -------------
CB_FN void fft_radix4(i32* base, int nb, int bstride, int G, int m, int mm, int fs) {
  const int per = G * m;
  CB_WSYNC();
  CB_LANE_LOOPU(u, nb * per) {
    int blk = u / per, r = u % per;
    int g = r / m, j = r % m;
    i32* F0 = base + blk * bstride + 2 * (g * mm + j);
    i32* F1 = F0 + 2 * m; i32* F2 = F0 + 4 * m; i32* F3 = F0 + 6 * m;
    i32 a0r, a0i, a1r, a1i, a2r, a2i;
    if (m == 1) {
      a0r = F1[0]; a0i = F1[1]; a1r = F2[0]; a1i = F2[1]; a2r = F3[0]; a2i = F3[1];
    } else {
      i32 t1r = cb_ldg(&ct_fft_tw480[2 * (j * fs)]), t1i = cb_ldg(&ct_fft_tw480[2 * (j * fs) + 1]);
      i32 t2r = cb_ldg(&ct_fft_tw480[2 * (2 * j * fs)]), t2i = cb_ldg(&ct_fft_tw480[2 * (2 * j * fs) + 1]);
      i32 t3r = cb_ldg(&ct_fft_tw480[2 * (3 * j * fs)]), t3i = cb_ldg(&ct_fft_tw480[2 * (3 * j * fs) + 1]);
      a0r = sub32(s_mul(F1[0], t1r), s_mul(F1[1], t1i)); a0i = add32(s_mul(F1[0], t1i), s_mul(F1[1], t1r));
      a1r = sub32(s_mul(F2[0], t2r), s_mul(F2[1], t2i)); a1i = add32(s_mul(F2[0], t2i), s_mul(F2[1], t2r));
      a2r = sub32(s_mul(F3[0], t3r), s_mul(F3[1], t3i)); a2i = add32(s_mul(F3[0], t3i), s_mul(F3[1], t3r));
    }
    i32 f0r = F0[0], f0i = F0[1];
    i32 s5r = sub32(f0r, a1r), s5i = sub32(f0i, a1i);
    f0r = add32(f0r, a1r); f0i = add32(f0i, a1i);
    i32 s3r = add32(a0r, a2r), s3i = add32(a0i, a2i);
    i32 s4r = sub32(a0r, a2r), s4i = sub32(a0i, a2i);
    F2[0] = sub32(f0r, s3r); F2[1] = sub32(f0i, s3i);
    F0[0] = add32(f0r, s3r); F0[1] = add32(f0i, s3i);
    F1[0] = add32(s5r, s4i); F1[1] = sub32(s5i, s4r);
    F3[0] = sub32(s5r, s4i); F3[1] = add32(s5i, s4r);
  }
  CB_WSYNC();
}
// kf_bfly2 (kiss_fft.c:51-108): only ever follows a radix-4 with m==1, so m==4, group stride 8
CB_FN void fft_radix2_m4(i32* base, int nb, int bstride, int G) {
  CB_WSYNC();
  CB_LANE_LOOPU(u, nb * G * 4) {
    int blk = u / (G * 4), r = u % (G * 4);
    int g = r >> 2, j = r & 3;
    i32* F = base + blk * bstride + 2 * (g * 8 + j);
    i32* F2 = F + 8;
    i32 tr, ti;
    i32 xr = F2[0], xi = F2[1];
    if (j == 0) { tr = xr; ti = xi; }
    else if (j == 1) { tr = s_mul(add32(xr, xi), 23170); ti = s_mul(sub32(xi, xr), 23170); }
    else if (j == 2) { tr = xi; ti = neg32(xr); }
    else { tr = s_mul(sub32(xi, xr), 23170); ti = s_mul(sub32(neg32(xi), xr), 23170); }
    i32 fr = F[0], fi = F[1];
    F2[0] = sub32(fr, tr); F2[1] = sub32(fi, ti);
    F[0] = add32(fr, tr); F[1] = add32(fi, ti);
  }
  CB_WSYNC();
}
CB_FN void fft_radix3(i32* base, int nb, int bstride, int G, int m, int mm, int fs) {
  const int per = G * m;
  CB_WSYNC();
  CB_LANE_LOOPU(u, nb * per) {
    int blk = u / per, r = u % per;
    int g = r / m, j = r % m;
    i32* F0 = base + blk * bstride + 2 * (g * mm + j);
    i32* F1 = F0 + 2 * m; i32* F2 = F0 + 4 * m;
    i32 t1r = cb_ldg(&ct_fft_tw480[2 * (j * fs)]), t1i = cb_ldg(&ct_fft_tw480[2 * (j * fs) + 1]);
    i32 t2r = cb_ldg(&ct_fft_tw480[2 * (2 * j * fs)]), t2i = cb_ldg(&ct_fft_tw480[2 * (2 * j * fs) + 1]);
    i32 s1r = sub32(s_mul(F1[0], t1r), s_mul(F1[1], t1i)), s1i = add32(s_mul(F1[0], t1i), s_mul(F1[1], t1r));
    i32 s2r = sub32(s_mul(F2[0], t2r), s_mul(F2[1], t2i)), s2i = add32(s_mul(F2[0], t2i), s_mul(F2[1], t2r));
    i32 s3r = add32(s1r, s2r), s3i = add32(s1i, s2i);
    i32 s0r = sub32(s1r, s2r), s0i = sub32(s1i, s2i);
    i32 f0r = F0[0], f0i = F0[1];
    i32 f1r = sub32(f0r, s3r >> 1), f1i = sub32(f0i, s3i >> 1);
    s0r = s_mul(s0r, -28378); s0i = s_mul(s0i, -28378);
    F0[0] = add32(f0r, s3r); F0[1] = add32(f0i, s3i);
    F2[0] = add32(f1r, s0i); F2[1] = sub32(f1i, s0r);
    F1[0] = sub32(f1r, s0i); F1[1] = add32(f1i, s0r);
  }
  CB_WSYNC();
}
CB_FN void fft_radix5(i32* base, int nb, int bstride, int G, int m, int mm, int fs) {
  const int per = G * m;
  const i32 yar = 10126, yai = -31164, ybr = -26510, ybi = -19261;
  CB_WSYNC();
  CB_LANE_LOOPU(u, nb * per) {
    int blk = u / per, r = u % per;
    int g = r / m, j = r % m;
    i32* F0 = base + blk * bstride + 2 * (g * mm + j);
    i32* F1 = F0 + 2 * m; i32* F2 = F0 + 4 * m; i32* F3 = F0 + 6 * m; i32* F4 = F0 + 8 * m;
    i32 s0r = F0[0], s0i = F0[1];
    i32 tr, ti;
    tr = cb_ldg(&ct_fft_tw480[2 * (j * fs)]); ti = cb_ldg(&ct_fft_tw480[2 * (j * fs) + 1]);
    i32 s1r = sub32(s_mul(F1[0], tr), s_mul(F1[1], ti)), s1i = add32(s_mul(F1[0], ti), s_mul(F1[1], tr));
    tr = cb_ldg(&ct_fft_tw480[2 * (2 * j * fs)]); ti = cb_ldg(&ct_fft_tw480[2 * (2 * j * fs) + 1]);
    i32 s2r = sub32(s_mul(F2[0], tr), s_mul(F2[1], ti)), s2i = add32(s_mul(F2[0], ti), s_mul(F2[1], tr));
    tr = cb_ldg(&ct_fft_tw480[2 * (3 * j * fs)]); ti = cb_ldg(&ct_fft_tw480[2 * (3 * j * fs) + 1]);
    i32 s3r = sub32(s_mul(F3[0], tr), s_mul(F3[1], ti)), s3i = add32(s_mul(F3[0], ti), s_mul(F3[1], tr));
    tr = cb_ldg(&ct_fft_tw480[2 * (4 * j * fs)]); ti = cb_ldg(&ct_fft_tw480[2 * (4 * j * fs) + 1]);
    i32 s4r = sub32(s_mul(F4[0], tr), s_mul(F4[1], ti)), s4i = add32(s_mul(F4[0], ti), s_mul(F4[1], tr));
    i32 s7r = add32(s1r, s4r), s7i = add32(s1i, s4i);
    i32 s10r = sub32(s1r, s4r), s10i = sub32(s1i, s4i);
    i32 s8r = add32(s2r, s3r), s8i = add32(s2i, s3i);
    i32 s9r = sub32(s2r, s3r), s9i = sub32(s2i, s3i);
    F0[0] = add32(s0r, add32(s7r, s8r)); F0[1] = add32(s0i, add32(s7i, s8i));
    i32 s5r = add32(add32(s0r, s_mul(s7r, yar)), s_mul(s8r, ybr));
    i32 s5i = add32(add32(s0i, s_mul(s7i, yar)), s_mul(s8i, ybr));
    i32 s6r = add32(s_mul(s10i, yai), s_mul(s9i, ybi));
    i32 s6i = sub32(neg32(s_mul(s10r, yai)), s_mul(s9r, ybi));
    F1[0] = sub32(s5r, s6r); F1[1] = sub32(s5i, s6i);
    F4[0] = add32(s5r, s6r); F4[1] = add32(s5i, s6i);
    i32 s11r = add32(add32(s0r, s_mul(s7r, ybr)), s_mul(s8r, yar));
    i32 s11i = add32(add32(s0i, s_mul(s7i, ybr)), s_mul(s8i, yar));
    i32 s12r = add32(neg32(s_mul(s10i, ybi)), s_mul(s9i, yai));
    i32 s12i = sub32(s_mul(s10r, ybi), s_mul(s9r, yai));
    F2[0] = add32(s11r, s12r); F2[1] = add32(s11i, s12i);
    F3[0] = sub32(s11r, s12r); F3[1] = sub32(s11i, s12i);
  }
  CB_WSYNC();
}
// opus_fft_impl (kiss_fft.c:532-578) for the four static states (static_modes_fixed.h:432-500);
// `sh` = mdct shift (0: 480 points, 1: 240, 2: 120, 3: 60).
CB_FN void fft_run(i32* base, int nb, int bstride, int sh) {
  if (sh == 0) {          // 480 = 5*3*4*2*4, executed 4,2,4,3,5; twiddle shift 0
    fft_radix4(base, nb, bstride, 120, 1, 4, 0);
    fft_radix2_m4(base, nb, bstride, 60);
    fft_radix4(base, nb, bstride, 15, 8, 32, 15);
    fft_radix3(base, nb, bstride, 5, 32, 96, 5);
    fft_radix5(base, nb, bstride, 1, 96, 1, 1);
  } else if (sh == 1) {   // 240 = 5*3*4*4; twiddle shift 1
    fft_radix4(base, nb, bstride, 60, 1, 4, 0);
    fft_radix4(base, nb, bstride, 15, 4, 16, 15 << 1);
    fft_radix3(base, nb, bstride, 5, 16, 48, 5 << 1);
    fft_radix5(base, nb, bstride, 1, 48, 1, 1 << 1);
  } else if (sh == 2) {   // 120 = 5*3*2*4; twiddle shift 2
    fft_radix4(base, nb, bstride, 30, 1, 4, 0);
    fft_radix2_m4(base, nb, bstride, 15);
    fft_radix3(base, nb, bstride, 5, 8, 24, 5 << 2);
    fft_radix5(base, nb, bstride, 1, 24, 1, 1 << 2);
  } else {                // 60 = 5*3*4; twiddle shift 3
    fft_radix4(base, nb, bstride, 15, 1, 4, 0);
    fft_radix3(base, nb, bstride, 5, 4, 12, 5 << 3);
    fft_radix5(base, nb, bstride, 1, 12, 1, 1 << 3);
  }
}

// clt_mdct_backward_c (mdct.c:263-363) for B interleaved transforms of frame length NB each.
// freq: B-interleaved input (stride B); out: out_syn base; transform b writes out + NB*b.
CB_FN void v_imdct(const i32* freq, i32* out, int B, int NB, int sh) {
  const int N2 = NB, N4 = NB >> 1;
  const int trig_off = sh == 0 ? 0 : sh == 1 ? 960 : sh == 2 ? 1440 : 1680;
  const i16* trig = ct_mdct_tw960 + trig_off;
  const i16* bitrev = sh == 0 ? ct_fft_bitrev480 : sh == 1 ? ct_fft_bitrev240 : sh == 2 ? ct_fft_bitrev120 : ct_fft_bitrev60;
  i32* ybase = out + (CB_OVERLAP >> 1);
  CB_WSYNC();
  CB_LANE_LOOPU(u, B * N4) {           // pre-rotation, scattered into bit-reversed order
    int b = u / N4, i = u % N4;
    const i32* in = freq + b;
    i32 x1 = in[2 * B * i];
    i32 x2 = in[B * (N2 - 1) - 2 * B * i];
    i32 t0 = cb_ldg(&trig[i]), t1 = cb_ldg(&trig[N4 + i]);
    int rev = cb_ldg(&bitrev[i]);
    i32 yr = add32(s_mul(x2, t0), s_mul(x1, t1));
    i32 yi = sub32(s_mul(x1, t0), s_mul(x2, t1));
    i32* yp = ybase + NB * b;
    yp[2 * rev + 1] = yr;
    yp[2 * rev] = yi;
  }
  fft_run(ybase, B, NB, sh);
  CB_LANE_LOOPU(u, B * (N4 >> 1)) {    // post-rotation from both ends
    int b = u / (N4 >> 1), i = u % (N4 >> 1);
    i32* yp0 = ybase + NB * b + 2 * i;
    i32* yp1 = ybase + NB * b + N2 - 2 - 2 * i;
    i32 re = yp0[1], im = yp0[0];
    i32 t0 = cb_ldg(&trig[i]), t1 = cb_ldg(&trig[N4 + i]);
    i32 yr = add32(s_mul(re, t0), s_mul(im, t1));
    i32 yi = sub32(s_mul(re, t1), s_mul(im, t0));
    i32 re2 = yp1[1], im2 = yp1[0];
    yp0[0] = yr;
    yp1[1] = yi;
    t0 = cb_ldg(&trig[N4 - i - 1]); t1 = cb_ldg(&trig[N2 - i - 1]);
    yr = add32(s_mul(re2, t0), s_mul(im2, t1));
    yi = sub32(s_mul(re2, t1), s_mul(im2, t0));
    yp1[0] = yr;
    yp0[1] = yi;
  }
  CB_WSYNC();
  CB_LANE_LOOPU(u, B * (CB_OVERLAP / 2)) {   // TDAC mirror
    int b = u / (CB_OVERLAP / 2), i = u % (CB_OVERLAP / 2);
    i32* o = out + NB * b;
    i32 x1 = o[CB_OVERLAP - 1 - i], x2 = o[i];
    i32 w1 = cb_ldg(&ct_window120[i]), w2 = cb_ldg(&ct_window120[CB_OVERLAP - 1 - i]);
    o[i] = sub32(mult16_32_q15(w2, x2), mult16_32_q15(w1, x1));
    o[CB_OVERLAP - 1 - i] = add32(mult16_32_q15(w1, x2), mult16_32_q15(w2, x1));
  }
  CB_WSYNC();
}

// comb_filter (celt.c:183-244), in place on x[0..n) where x[k<0] comes from the history ring.
CB_FN i32 comb_at(const i32* buf, const i32* hist, int pos, int k) {
  return k >= 0 ? buf[k] : hist[(pos + k) & (CB_HIST - 1)];
}
CB_FN void v_comb_filter(i32* buf, int off, int n, const i32* hist, int pos, int T0, int T1, i32 g0, i32 g1,
                         int tapset0, int tapset1, int overlap) {
  if (g0 == 0 && g1 == 0) return;
  i32 g00 = mult16_16_p15(g0, ct_comb_gains[tapset0][0]);
  i32 g01 = mult16_16_p15(g0, ct_comb_gains[tapset0][1]);
  i32 g02 = mult16_16_p15(g0, ct_comb_gains[tapset0][2]);
  i32 g10 = mult16_16_p15(g1, ct_comb_gains[tapset1][0]);
  i32 g11 = mult16_16_p15(g1, ct_comb_gains[tapset1][1]);
  i32 g12 = mult16_16_p15(g1, ct_comb_gains[tapset1][2]);
  if (g0 == g1 && T0 == T1 && tapset0 == tapset1) overlap = 0;
  // wave length: every tap that carries weight must point at an already final sample
  int lag = 1 << 30;
  if (g0 != 0) lag = cb_imin(lag, T0);
  if (g1 != 0) lag = cb_imin(lag, T1);
  int wave = cb_imax(1, lag - 2);
  // cross-fade part
  for (int base = 0; base < overlap; base += wave) {
    int cnt = cb_imin(wave, overlap - base);
    CB_WSYNC();
    CB_LANE_LOOPU(t, cnt) {
      int i = off + base + t;
      i32 w = cb_ldg(&ct_window120[base + t]);
      i32 f = (i16)mult16_16_q15(w, w);
      i32 y = buf[i];
      y = add32(y, mult16_32_q15(mult16_16_q15(CB_Q15ONE - f, g00), comb_at(buf, hist, pos, i - T0)));
      y = add32(y, mult16_32_q15(mult16_16_q15(CB_Q15ONE - f, g01),
                                 add32(comb_at(buf, hist, pos, i - T0 + 1), comb_at(buf, hist, pos, i - T0 - 1))));
      y = add32(y, mult16_32_q15(mult16_16_q15(CB_Q15ONE - f, g02),
                                 add32(comb_at(buf, hist, pos, i - T0 + 2), comb_at(buf, hist, pos, i - T0 - 2))));
      y = add32(y, mult16_32_q15(mult16_16_q15(f, g10), comb_at(buf, hist, pos, i - T1)));
      y = add32(y, mult16_32_q15(mult16_16_q15(f, g11),
                                 add32(comb_at(buf, hist, pos, i - T1 + 1), comb_at(buf, hist, pos, i - T1 - 1))));
      y = add32(y, mult16_32_q15(mult16_16_q15(f, g12),
                                 add32(comb_at(buf, hist, pos, i - T1 + 2), comb_at(buf, hist, pos, i - T1 - 2))));
      buf[i] = y;
    }
  }
  if (g1 == 0) { CB_WSYNC(); return; }
  // constant part (comb_filter_const_c, celt.c:156-178)
  int wave1 = cb_imax(1, T1 - 2);
  for (int base = overlap; base < n; base += wave1) {
    int cnt = cb_imin(wave1, n - base);
    CB_WSYNC();
    CB_LANE_LOOPU(t, cnt) {
      int i = off + base + t;
      i32 y = buf[i];
      y = add32(y, mult16_32_q15(g10, comb_at(buf, hist, pos, i - T1)));
      y = add32(y, mult16_32_q15(g11, add32(comb_at(buf, hist, pos, i - T1 + 1), comb_at(buf, hist, pos, i - T1 - 1))));
      y = add32(y, mult16_32_q15(g12, add32(comb_at(buf, hist, pos, i - T1 + 2), comb_at(buf, hist, pos, i - T1 - 2))));
      buf[i] = y;
    }
  }
  CB_WSYNC();
}

// De-emphasis + PCM store + ring update for a finished frame held in buf[0..N) (shared by the
// synthesis kernel and the concealment kernel).  celt_decoder.c:185-275.  Three steps:
//  * emit_ring: the frame overwrites the N oldest samples of the history ring (lane loop);
//  * emit_deemph: de-emphasis is a serial one-pole IIR (m = 0.85*(x+m), truncating multiply), ONE LANE per
//    channel.  A warp on its own gives it lane 0; the synthesis kernel hands the channels of a whole CTA to the
//    lanes of one warp (EmitReq below), so that the 960 serial steps are issued once per CTA, not once per warp;
//  * emit_pcm: SIG2WORD16 + interleaved int16 store (optionally accumulated, hybrid mode) in a lane loop.
CB_FN void emit_ring(const i32* buf, int N, CeltDecState* st, int c) {
  i32* hist = st->hist[c];
  const int pos = st->hist_pos[c];
  CB_WSYNC();
  CB_LANE_LOOPU(j, N) hist[(pos + j) & (CB_HIST - 1)] = buf[j];
  CB_WSYNC();
  CB_IF_LANE0 st->hist_pos[c] = (pos + N) & (CB_HIST - 1);
}
CB_FN void emit_deemph(i32* buf, int N, CeltDecState* st, int c) {     // one lane
  i32 m = st->preemph_mem[c];
  for (int j = 0; j < N; j++) {
    i32 tmp = add32(buf[j], m);
    m = mult16_32_q15(27853, tmp);
    buf[j] = tmp;
  }
  st->preemph_mem[c] = m;
}
CB_FN void emit_pcm(const i32* buf, int N, const CeltDecState* st, int c, i16* pcm, int accum, int decode_gain) {
  const int CC = st->channels, downsample = st->downsample;
  const int Nd = N / downsample;
  if (decode_gain == 0) {
    if (accum) {
      CB_LANE_LOOPU(j, Nd) pcm[j * CC + c] = sat16((i32)pcm[j * CC + c] + (i32)sig2word16(buf[j * downsample]));
    } else {
      CB_LANE_LOOPU(j, Nd) pcm[j * CC + c] = sig2word16(buf[j * downsample]);
    }
  } else {
    // OPUS_SET_GAIN (opus_decoder.c:567-577): Q8 dB -> linear Q16, rounded multiply, clamp to +-32767
    const i32 gain = cb_exp2(mult16_16_p15(21771, decode_gain));     // QCONST16(6.48814081e-4f, 25)
    CB_LANE_LOOPU(j, Nd) {
      i32 v = sig2word16(buf[j * downsample]);
      if (accum) v = sat16((i32)pcm[j * CC + c] + v);
      i32 x = mult16_32_p16(v, gain);
      pcm[j * CC + c] = (i16)(x > 32767 ? 32767 : x < -32767 ? -32767 : x);
    }
  }
  CB_WSYNC();
}
CB_FN void emit_frame(i32* buf, int N, CeltDecState* st, int c, i16* pcm, int accum, int decode_gain) {
  emit_ring(buf, N, st, c);
  CB_IF_LANE0 emit_deemph(buf, N, st, c);
  CB_WSYNC();
  emit_pcm(buf, N, st, c, pcm, accum, decode_gain);
}
// A frame whose ring is written and whose de-emphasis + PCM store are still to do (synthesis kernel: the CTA's
// warps park theirs in shared memory, one warp runs the de-emphasis of all of them, a lane each).
struct EmitReq {
  i32* buf;
  CeltDecState* st;
  i16* pcm;
  i32 N, c, accum, decode_gain, valid;
};

// Synthesis of output channel `c` of one stream for one frame.
// pcm -> first sample of this frame for this stream (interleaved, CC channels).
CB_FN void celt_synth_channel(SynWork& w, CeltDecState* st, const FrameSide* side, int c, i16* pcm,
                               int decode_gain = 0, EmitReq* defer = nullptr) {
  const int LM = side->LM, C = side->C, CC = st->channels;
  const int M = 1 << LM, N = M * 120;
  const int downsample = st->downsample;
  const int transient = side->transient;
  const int B = transient ? M : 1;
  const int NB = transient ? 120 : N;
  const int sh = transient ? 3 : 3 - LM;
  i32* hist = st->hist[c];
  const int pos = st->hist_pos[c];

  // 1. denormalise (celt_synthesis, celt_decoder.c:313-348)
  if (CC == 1 && C == 2) {
    i32* f2 = w.buf + (CB_OVERLAP >> 1);
    v_denormalise(side->X, w.freq, side->bandE, side->start, side->effEnd, M, downsample, side->silence, w.buf);
    v_denormalise(side->X + N, f2, side->bandE + CB_NBANDS, side->start, side->effEnd, M, downsample, side->silence, w.buf);
    CB_LANE_LOOPU(i, N) w.freq[i] = half32(add32(w.freq[i], f2[i]));
    CB_WSYNC();
  } else {
    int sc = (C == 1) ? 0 : c;
    v_denormalise(side->X + sc * N, w.freq, side->bandE + sc * CB_NBANDS, side->start, side->effEnd, M,
                  downsample, side->silence, w.buf);
  }
  // 2. overlap tail of the previous frame becomes out_syn[0..60)
  CB_LANE_LOOPU(i, CB_OVERLAP / 2) w.buf[i] = st->tail[c][i];
  // 3. inverse MDCT(s)
  v_imdct(w.freq, w.buf, B, NB, sh);
  // 4. new overlap tail
  CB_LANE_LOOPU(i, CB_OVERLAP / 2) st->tail[c][i] = w.buf[N + i];
  // 5. pitch post-filter (celt_decoder.c:1002-1013)
  v_comb_filter(w.buf, 0, 120, hist, pos, side->T0, side->T1, side->g0, side->g1, side->tap0, side->tap1, CB_OVERLAP);
  if (LM != 0)
    v_comb_filter(w.buf, 120, N - 120, hist, pos, side->T1, side->T2, side->g1, side->g2, side->tap1, side->tap2, CB_OVERLAP);
  if (defer == nullptr) {
    emit_frame(w.buf, N, st, c, pcm, side->accum, decode_gain);
    return;
  }
  emit_ring(w.buf, N, st, c);
  CB_IF_LANE0 {
    defer->buf = w.buf; defer->st = st; defer->pcm = pcm;
    defer->N = N; defer->c = c; defer->accum = side->accum; defer->decode_gain = decode_gain; defer->valid = 1;
  }
}
