// cb_enc_analysis.h -- time-domain front end of the CELT encoder: pre-emphasis, pitch pre-filter
// (pitch search, period-doubling removal, FIR comb filter) and transient analysis.
// Bit-exact with celt_preemphasis / run_prefilter / transient_analysis / patch_transient_decision
// (opus-fix/celt/celt_encoder.c:227-535,1067-1192), pitch_downsample / remove_doubling
// (celt/pitch.c:147-222,371-505) and comb_filter with x != y (celt/celt.c:183-244).
// One warp per stream: FIR-shaped work is lane loops, the short IIR recurrences run on one lane.
#pragma once
#include "cb_common.h"
#include "cb_enc_state.h"
#include "cb_pitch.h"
#include "cb_celt_synth.h"   // ct_comb_gains, CB_IF_LANE0

// celt_preemphasis (celt_encoder.c:464-535), FIXED_POINT: m depends on the input only, so it is a FIR.
// pcm: interleaved CC channels, this channel at pcm[CC*i]; inp: N samples out; mem: 1 value.
CB_FN void enc_preemphasis(const i16* pcm, i32* inp, int N, int CC, int upsample, i32* mem) {
  const i32 m0 = *mem;
  const int Nu = N / upsample;
  CB_WSYNC();
  CB_LANE_LOOP(i, N) {
    // x[i] = sample feeding position i (zero-stuffed when upsampling); all taken as opus_val16
    i32 x = (i % upsample == 0) ? (i32)pcm[CC * (i / upsample)] : 0;
    i32 xp;
    if (i == 0) xp = 0; else xp = ((i - 1) % upsample == 0) ? (i32)pcm[CC * ((i - 1) / upsample)] : 0;
    i32 m = (i == 0) ? m0 : (mult16_16(27853, xp) >> 3);
    inp[i] = sub32(shl32(x, CB_SIG_SHIFT), m);
  }
  i32 xl = ((N - 1) % upsample == 0) ? (i32)pcm[CC * ((N - 1) / upsample)] : 0;
  (void)Nu;
  CB_WSYNC();
  *mem = mult16_16(27853, xl) >> 3;
  CB_WSYNC();
}

// comb_filter with separate input and output (celt.c:183-244): a pure FIR, every sample independent.
// x may be indexed down to x[-T-2].
CB_FN_NOINLINE void enc_comb_fir(i32* y, const i32* x, int T0, int T1, int N, i32 g0, i32 g1, int tap0, int tap1,
                        int overlap) {
  CB_WSYNC();
  if (g0 == 0 && g1 == 0) {
    CB_LANE_LOOP(i, N) y[i] = x[i];
    CB_WSYNC();
    return;
  }
  const i32 g00 = mult16_16_p15(g0, ct_comb_gains[tap0][0]), g01 = mult16_16_p15(g0, ct_comb_gains[tap0][1]),
            g02 = mult16_16_p15(g0, ct_comb_gains[tap0][2]);
  const i32 g10 = mult16_16_p15(g1, ct_comb_gains[tap1][0]), g11 = mult16_16_p15(g1, ct_comb_gains[tap1][1]),
            g12 = mult16_16_p15(g1, ct_comb_gains[tap1][2]);
  if (g0 == g1 && T0 == T1 && tap0 == tap1) overlap = 0;
  CB_LANE_LOOP(i, N) {
    i32 v = x[i];
    if (i < overlap) {
      i32 w = cb_ldg(&ct_window120[i]);
      i32 f = (i16)mult16_16_q15(w, w);
      v = add32(v, mult16_32_q15(mult16_16_q15(CB_Q15ONE - f, g00), x[i - T0]));
      v = add32(v, mult16_32_q15(mult16_16_q15(CB_Q15ONE - f, g01), add32(x[i - T0 + 1], x[i - T0 - 1])));
      v = add32(v, mult16_32_q15(mult16_16_q15(CB_Q15ONE - f, g02), add32(x[i - T0 + 2], x[i - T0 - 2])));
      v = add32(v, mult16_32_q15(mult16_16_q15(f, g10), x[i - T1]));
      v = add32(v, mult16_32_q15(mult16_16_q15(f, g11), add32(x[i - T1 + 1], x[i - T1 - 1])));
      v = add32(v, mult16_32_q15(mult16_16_q15(f, g12), add32(x[i - T1 + 2], x[i - T1 - 2])));
    } else if (g1 != 0) {
      v = add32(v, mult16_32_q15(g10, x[i - T1]));
      v = add32(v, mult16_32_q15(g11, add32(x[i - T1 + 1], x[i - T1 - 1])));
      v = add32(v, mult16_32_q15(g12, add32(x[i - T1 + 2], x[i - T1 - 2])));
    }
    y[i] = v;
  }
  CB_WSYNC();
}

// pitch_downsample head (pitch.c:147-182) for linear buffers x0/x1 of `len` samples.
CB_FN void enc_pitch_downsample(const i32* x0, const i32* x1, i16* x_lp, int len, int C, EncScratch& sc, i16* xx) {
  i32 mx = 0, mn = 0;
  CB_WSYNC();
  CB_LANE_LOOP(i, len) {
    i32 v = x0[i];
    mx = max32(mx, v); mn = min32(mn, v);
    if (C == 2) { v = x1[i]; mx = max32(mx, v); mn = min32(mn, v); }
  }
  i32 maxabs = max32(cb_wmax(mx), neg32(cb_wmin(mn)));
  if (maxabs < 1) maxabs = 1;
  int shift = cb_ilog2(maxabs) - 10;
  if (shift < 0) shift = 0;
  if (C == 2) shift++;
  CB_LANE_LOOP(i, len >> 1) {
    i32 v = (i == 0) ? (half32(add32(half32(x0[1]), x0[0])) >> shift)
                     : (half32(add32(half32(add32(x0[2 * i - 1], x0[2 * i + 1])), x0[2 * i])) >> shift);
    i32 acc = (i16)v;
    if (C == 2) {
      i32 v1 = (i == 0) ? (half32(add32(half32(x1[1]), x1[0])) >> shift)
                        : (half32(add32(half32(add32(x1[2 * i - 1], x1[2 * i + 1])), x1[2 * i])) >> shift);
      acc = (i16)(acc + v1);
    }
    x_lp[i] = (i16)acc;
  }
  CB_WSYNC();
  pitch_whiten(x_lp, len >> 1, sc.ac, sc.lpc32, xx);
}

// lane-parallel inner product of two int16 vectors
CB_FN_NOINLINE i32 enc_inner(const i16* a, const i16* b, int n) {
  i32 s = 0;
  CB_LANE_LOOP(j, n) s = mac16_16(s, a[j], b[j]);
  return cb_wsum(s);
}

// remove_doubling (pitch.c:371-505).  x: pitch buffer (2x decimated); returns the pitch gain, *T0_ updated.
CB_FN i32 enc_remove_doubling(const i16* xbuf, int maxperiod, int minperiod, int N, int* T0_, int prev_period,
                              i32 prev_gain, i32* yy_lookup) {
  const int second_check[16] = {0, 0, 3, 2, 3, 2, 5, 2, 3, 2, 3, 2, 5, 2, 3, 2};
  const int minperiod0 = minperiod;
  maxperiod /= 2; minperiod /= 2; *T0_ /= 2; prev_period /= 2; N /= 2;
  const i16* x = xbuf + maxperiod;
  if (*T0_ >= maxperiod) *T0_ = maxperiod - 1;
  int T, T0;
  T = T0 = *T0_;
  CB_WSYNC();
  i32 xx = enc_inner(x, x, N);
  i32 xy = enc_inner(x, x - T0, N);
  // yy_lookup[i] = max(0, running energy): a prefix recurrence, serial on every lane's copy (uniform)
  {
    i32 yy = xx;
    yy_lookup[0] = xx;
    CB_NOUNROLL for (int i = 1; i <= maxperiod; i++) {
      yy = sub32(add32(yy, mult16_16(x[-i], x[-i])), mult16_16(x[N - i], x[N - i]));
      yy_lookup[i] = max32(0, yy);
    }
  }
  CB_WSYNC();
  i32 yy = yy_lookup[T0];
  i32 best_xy = xy, best_yy = yy;
  i32 g, g0;
  {
    i32 x2y2 = add32(1, half32(mult32_32_q31(xx, yy)));
    int sh = cb_ilog2(x2y2) >> 1;
    i32 t = vshr32(x2y2, 2 * (sh - 7));
    g = g0 = vshr32(mult16_32_q15(cb_rsqrt_norm(t), xy), sh + 1);
  }
  CB_NOUNROLL for (int k = 2; k <= 15; k++) {
    int T1 = (int)((u32)(2 * T0 + k) / (u32)(2 * k));
    if (T1 < minperiod) break;
    int T1b;
    if (k == 2) T1b = (T1 + T0 > maxperiod) ? T0 : T0 + T1;
    else T1b = (int)((u32)(2 * second_check[k] * T0 + k) / (u32)(2 * k));
    i32 xya = enc_inner(x, x - T1, N), xyb = enc_inner(x, x - T1b, N);
    xy = add32(xya, xyb);
    yy = add32(yy_lookup[T1], yy_lookup[T1b]);
    i32 g1;
    {
      i32 x2y2 = add32(1, mult32_32_q31(xx, yy));
      int sh = cb_ilog2(x2y2) >> 1;
      i32 t = vshr32(x2y2, 2 * (sh - 7));
      g1 = vshr32(mult16_32_q15(cb_rsqrt_norm(t), xy), sh + 1);
    }
    i32 cont;
    int d = T1 - prev_period; if (d < 0) d = -d;
    if (d <= 1) cont = (i16)prev_gain;
    else if (d <= 2 && 5 * k * k < T0) cont = (i16)half32(prev_gain);
    else cont = 0;
    i32 thresh = max32(9830, sub32(mult16_32_q15(22938, g0), cont));          // QCONST32(.3f,15), QCONST16(.7f,15)
    if (T1 < 3 * minperiod) thresh = max32(13107, sub32(mult16_32_q15(27853, g0), cont));   // .4, .85
    else if (T1 < 2 * minperiod) thresh = max32(16384, sub32(mult16_32_q15(29491, g0), cont)); // .5, .9
    if (g1 > thresh) { best_xy = xy; best_yy = yy; T = T1; g = g1; }
  }
  best_xy = max32(0, best_xy);
  i32 pg;
  if (best_yy <= best_xy) pg = CB_Q15ONE;
  else pg = (i16)(cb_frac_div32(best_xy, add32(best_yy, 1)) >> 16);
  i32 xc0 = enc_inner(x, x - (T - 1), N), xc1 = enc_inner(x, x - T, N), xc2 = enc_inner(x, x - (T + 1), N);
  int offset;
  if (sub32(xc2, xc0) > mult16_32_q15(22938, sub32(xc1, xc0))) offset = 1;
  else if (sub32(xc0, xc2) > mult16_32_q15(22938, sub32(xc1, xc2))) offset = -1;
  else offset = 0;
  if (pg > g) pg = (i16)g;
  *T0_ = 2 * T + offset;
  if (*T0_ < minperiod0) *T0_ = minperiod0;
  return pg;
}

// run_prefilter (celt_encoder.c:1067-1192).  Returns pf_on; *pitch, *gain, *qgain as in the reference.
CB_FN int enc_run_prefilter(EncWork& w, CeltEncScalars& s, CeltEncState* st, EncScratch& sc, int CC, int N, int prefilter_tapset, int* pitch,
                            i32* gain, int* qgain, int enabled, int nbAvailableBytes) {
  const int overlap = CB_OVERLAP;
  CB_WSYNC();
  CB_NOUNROLL for (int c = 0; c < CC; c++) {
    CB_LANE_LOOP(i, CB_COMBFILTER_MAXPERIOD) sc.pre[c][i] = st->prefilter_mem[c][i];
    CB_LANE_LOOP(i, N) sc.pre[c][CB_COMBFILTER_MAXPERIOD + i] = sc.in[c][overlap + i];
  }
  CB_WSYNC();
  int pitch_index;
  i32 gain1;
  if (enabled) {
    enc_pitch_downsample(sc.pre[0], sc.pre[1], w.pitch_buf, CB_COMBFILTER_MAXPERIOD + N, CC, sc, w.pitch_xx);
    pitch_index = pitch_search(w.pitch_buf + (CB_COMBFILTER_MAXPERIOD >> 1), w.pitch_buf, N,
                               CB_COMBFILTER_MAXPERIOD - 3 * CB_COMBFILTER_MINPERIOD, w.pitch_x4, w.pitch_y4, sc.xcorr);
    pitch_index = CB_COMBFILTER_MAXPERIOD - pitch_index;
    gain1 = enc_remove_doubling(w.pitch_buf, CB_COMBFILTER_MAXPERIOD, CB_COMBFILTER_MINPERIOD, N, &pitch_index,
                                s.prefilter_period, s.prefilter_gain, sc.yy_lookup);
    if (pitch_index > CB_COMBFILTER_MAXPERIOD - 2) pitch_index = CB_COMBFILTER_MAXPERIOD - 2;
    gain1 = (i16)mult16_16_q15(22938, gain1);
    if (s.loss_rate > 2) gain1 = (i16)half32(gain1);
    if (s.loss_rate > 4) gain1 = (i16)half32(gain1);
    if (s.loss_rate > 8) gain1 = 0;
  } else {
    gain1 = 0;
    pitch_index = CB_COMBFILTER_MINPERIOD;
  }
  i32 pf_threshold = 6554;                                   // QCONST16(.2f,15)
  int dp = pitch_index - s.prefilter_period; if (dp < 0) dp = -dp;
  if (dp * 10 > pitch_index) pf_threshold += 6554;
  if (nbAvailableBytes < 25) pf_threshold += 3277;          // QCONST16(.1f,15)
  if (nbAvailableBytes < 35) pf_threshold += 3277;
  if (s.prefilter_gain > 13107) pf_threshold -= 3277;     // QCONST16(.4f,15)
  if (s.prefilter_gain > 18022) pf_threshold -= 3277;     // QCONST16(.55f,15)
  pf_threshold = (i16)pf_threshold;
  pf_threshold = max32(pf_threshold, 6554);
  int pf_on, qg;
  if (gain1 < pf_threshold) {
    gain1 = 0; pf_on = 0; qg = 0;
  } else {
    int dg = gain1 - s.prefilter_gain; dg = (i16)dg; if (dg < 0) dg = -dg;
    if (dg < 3277) gain1 = s.prefilter_gain;
    qg = ((gain1 + 1536) >> 10) / 3 - 1;
    qg = cb_imax(0, cb_imin(7, qg));
    gain1 = 3072 * (qg + 1);                                 // QCONST16(0.09375f,15)
    pf_on = 1;
  }
  const int offset = 0;                                      // shortMdctSize - overlap = 0 in this mode
  CB_NOUNROLL for (int c = 0; c < CC; c++) {
    s.prefilter_period = cb_imax(s.prefilter_period, CB_COMBFILTER_MINPERIOD);
    CB_WSYNC();
    CB_LANE_LOOP(i, overlap) sc.in[c][i] = st->in_mem[c][i];
    enc_comb_fir(sc.in[c] + overlap + offset, sc.pre[c] + CB_COMBFILTER_MAXPERIOD + offset, s.prefilter_period,
                 pitch_index, N - offset, (i16)(-s.prefilter_gain), (i16)(-gain1), s.prefilter_tapset,
                 prefilter_tapset, overlap);
    CB_LANE_LOOP(i, overlap) st->in_mem[c][i] = sc.in[c][N + i];
    CB_WSYNC();
    // N < MAXPERIOD always: the shifted memory is pre[N .. N+MAXPERIOD) (celt_encoder.c:1177-1184)
    CB_LANE_LOOP(i, CB_COMBFILTER_MAXPERIOD) st->prefilter_mem[c][i] = sc.pre[c][N + i];
    CB_WSYNC();
  }
  *gain = gain1;
  *pitch = pitch_index;
  *qgain = qg;
  return pf_on;
}

// transient_analysis (celt_encoder.c:227-376).  in: C channels of `len` samples, channel stride `cstride`.
CB_GTABLE u8 ct_inv_table[128] = {
    255, 255, 156, 110, 86, 70, 59, 51, 45, 40, 37, 33, 31, 28, 26, 25, 23, 22, 21, 20, 19, 18, 17, 16, 16, 15, 15, 14,
    13, 13, 12, 12, 12, 12, 11, 11, 11, 10, 10, 10, 9, 9, 9, 9, 9, 9, 8, 8, 8, 8, 8, 7, 7, 7, 7, 7, 7, 6, 6, 6, 6, 6, 6,
    6, 6, 6, 6, 6, 6, 6, 6, 6, 6, 5, 5, 5, 5, 5, 5, 5, 5, 5, 5, 5, 5, 4, 4, 4, 4, 4, 4, 4, 4, 4, 4, 4, 4, 4, 4, 4, 4, 4,
    4, 4, 4, 4, 4, 4, 4, 4, 3, 3, 3, 3, 3, 3, 3, 3, 3, 3, 3, 3, 3, 3, 3, 3, 3, 2};

// transient_analysis (celt_encoder.c:227-330) in two parts.
// Part 1, one channel, one thread -- the recurrences: the high-pass filter, the scaling to full range (a maximum
// over the filtered vector, applied on the fly in the next pass), the forward and the backward smoothing.
// Leaves the smoothed energy envelope in tmp[0 .. len/2) and returns its mean / maximum terms.
CB_FN void enc_ta_chain(const i32* x_in, i16* tmp, int len, i32* mean_out, i32* maxE_out) {
  const int len2 = len / 2;
  i32 mem0 = 0, mem1 = 0;
  i32 mx = 0, mn = 0;
  // len is a multiple of 4 (N + overlap).  Sixteen samples per trip while they last: the loads (global memory) go
  // out together, then the chain runs -- in the CTA's serial section only two warps are there to hide their latency.
  int i = 0;
#define CB_TA_STEP(W)                                                                                      \
  {                                                                                                        \
    i32 xs[W];                                                                                             \
    for (int k = 0; k < W; k++) xs[k] = x_in[i + k];                                                       \
    for (int k = 0; k < W; k++) {                                                                          \
      i32 x = xs[k] >> CB_SIG_SHIFT;                                                                       \
      i32 y = add32(mem0, x);                                                                              \
      mem0 = sub32(add32(mem1, y), shl32(x, 1));                                                           \
      mem1 = sub32(x, y >> 1);                                                                             \
      const i16 v = (i16)(y >> 2);                                                                         \
      tmp[i + k] = v;                                                                                      \
      if (i + k >= 12) { mx = max32(mx, v); mn = min32(mn, v); } /* (the first 12 are cleared below) */    \
    }                                                                                                      \
  }
  CB_NOUNROLL for (; i + 16 <= len; i += 16) CB_TA_STEP(16)
  CB_NOUNROLL for (; i < len; i += 4) CB_TA_STEP(4)
#undef CB_TA_STEP
  for (int i = 0; i < 12; i++) tmp[i] = 0;
  const i32 maxabs = max32(mx, neg32(mn));
  const int shift = 14 - cb_ilog2(add32(1, maxabs));         // normalise to full range, applied as the values are read
  i32 mean = 0, maxE = 0;
  mem0 = 0;
  CB_NOUNROLL for (int i = 0; i < len2; i++) {
    i16 a = tmp[2 * i], b = tmp[2 * i + 1];
    if (shift != 0) { a = shl16(a, shift); b = shl16(b, shift); }
    i16 x2 = (i16)pshr32(add32(mult16_16(a, a), mult16_16(b, b)), 16);
    mean = add32(mean, x2);
    tmp[i] = (i16)add32(mem0, pshr32(sub32(x2, mem0), 4));
    mem0 = tmp[i];
  }
  mem0 = 0;
  CB_NOUNROLL for (int i = len2 - 1; i >= 0; i--) {
    tmp[i] = (i16)add32(mem0, pshr32(sub32(tmp[i], mem0), 3));
    mem0 = tmp[i];
    maxE = max32(maxE, (i16)mem0);
  }
  *mean_out = mean;
  *maxE_out = maxE;
}
// Part 2, lane-parallel: the masking metric of every channel from its envelope.
CB_FN_NOINLINE int enc_transient_finish(int len, int C, const i32* mean_c, const i32* maxE_c, i32* tf_estimate, int* tf_chan,
                                        const i16* tmp0, int tstride) {
  i32 mask_metric = 0;
  const int len2 = len / 2;
  CB_WSYNC();
  CB_NOUNROLL for (int c = 0; c < C; c++) {
    const i16* tmp = tmp0 + c * tstride;
    i32 mean = mean_c[c];
    const i32 maxE = maxE_c[c];
    mean = mult16_16(cb_sqrt(mean), cb_sqrt(mult16_16(maxE, len2 >> 1)));
    i32 norm = shl32(len2, 6 + 14) / add32(1, mean >> 1);
    i32 un = 0;
    CB_LANE_LOOP(q, (len2 - 5 - 12 + 3) / 4) {
      int i = 12 + 4 * q;
      int id = max32(0, min32(127, mult16_32_q15(add32(tmp[i], 1), norm)));
      un += cb_ldg(&ct_inv_table[id]);
    }
    i32 unmask = cb_wsum(un);
    unmask = 64 * unmask * 4 / (6 * (len2 - 17));
    if (unmask > mask_metric) { *tf_chan = c; mask_metric = unmask; }
  }
  int is_transient = mask_metric > 200;
  i32 tf_max = max32(0, (i16)(cb_sqrt(27 * mask_metric) - 42));
  *tf_estimate = (i16)cb_sqrt(max32(0, sub32(shl32(mult16_16(113, min32(163, tf_max)), 14), 37312528)));
  return is_transient;
}
// Host build: both parts in turn.  (The device posts part 1 as a serial-section request: cb_celt_enc.h.)
CB_FN int enc_transient_analysis(const i32* in, int cstride, int len, int C, i32* tf_estimate, int* tf_chan,
                                 i16* tmp0, int tstride) {
  i32 mean_c[2] = {0, 0}, maxE_c[2] = {0, 0};
  for (int c = 0; c < C; c++) enc_ta_chain(in + c * cstride, tmp0 + c * tstride, len, &mean_c[c], &maxE_c[c]);
  return enc_transient_finish(len, C, mean_c, maxE_c, tf_estimate, tf_chan, tmp0, tstride);
}

// dc_reject (opus_encoder.c:362-384): two cascaded one-pole sections.  One channel, one thread: the recurrence.
CB_FN void enc_dc_chain(const i16* in, i16* out, i32* mem, int c, int len, int channels, int shift) {
  i32 m0 = mem[2 * c], m1 = mem[2 * c + 1];
  const i16* ip = in + c;
  i16* op = out + c;
  int i = 0;
  CB_NOUNROLL for (; i + 16 <= len; i += 16) {            // sixteen samples per trip: loads first, then the chain
    i32 xs[16];
    for (int k = 0; k < 16; k++) xs[k] = ip[channels * (i + k)];
    for (int k = 0; k < 16; k++) {
      const i32 x = shl32(xs[k], 15);
      const i32 tmp = sub32(x, m0);
      m0 = add32(m0, pshr32(tmp, shift));
      const i32 y = sub32(tmp, m1);
      m1 = add32(m1, pshr32(y, shift));
      i32 v = pshr32(y, 15);
      v = v > 32767 ? 32767 : (v < -32767 ? -32767 : v);
      op[channels * (i + k)] = (i16)v;
    }
  }
  for (; i < len; i++) {
    const i32 x = shl32(ip[channels * i], 15);
    const i32 tmp = sub32(x, m0);
    m0 = add32(m0, pshr32(tmp, shift));
    const i32 y = sub32(tmp, m1);
    m1 = add32(m1, pshr32(y, shift));
    i32 v = pshr32(y, 15);
    v = v > 32767 ? 32767 : (v < -32767 ? -32767 : v);
    op[channels * i] = (i16)v;
  }
  mem[2 * c] = m0;
  mem[2 * c + 1] = m1;
}

#if CB_DEVICE_BUILD
// ---- serial sections of the CTA's streams on the lanes of warps 0 and 1 (cb_enc_state.h: EncSerialReq) ----------
// Called by every warp after a rendezvous (whether it encodes or only makes up rendezvous).
CB_FN void enc_serial_run(EncSerialReq* table) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp >= 2) return;
  const int nw = blockDim.x >> 5;
  const int slot = warp * 16 + (lane >> 1), c = lane & 1;
  if (slot < nw) {
    EncSerialReq& r = table[slot];
    const int kind = r.kind;
    if (c < r.channels) {
      if (kind == CB_ES_DC) {
        enc_dc_chain((const i16*)r.in, (i16*)r.out, r.m, c, r.len, r.channels, r.shift);
      } else if (kind == CB_ES_TRANSIENT) {
        enc_ta_chain((const i32*)r.in + c * r.stride_in, (i16*)r.out + c * r.stride_out, r.len, &r.m[c], &r.m[2 + c]);
      }
    }
  }
  __syncwarp();
  if (slot < nw && c == 0) table[slot].kind = CB_ES_NONE;
  __syncwarp();
}
#endif

// patch_transient_decision (celt_encoder.c:380-415)
CB_FN_NOINLINE int enc_patch_transient(const i16* newE, const i16* oldE, int start, int end, int C) {
  i32 mean_diff = 0;
  i16 spread_old[26];
  if (C == 1) {
    spread_old[start] = oldE[start];
    for (int i = start + 1; i < end; i++) spread_old[i] = (i16)cb_imax(spread_old[i - 1] - 1024, oldE[i]);
  } else {
    spread_old[start] = (i16)cb_imax(oldE[start], oldE[start + CB_NBANDS]);
    for (int i = start + 1; i < end; i++)
      spread_old[i] = (i16)cb_imax(spread_old[i - 1] - 1024, cb_imax(oldE[i], oldE[i + CB_NBANDS]));
  }
  for (int i = end - 2; i >= start; i--) spread_old[i] = (i16)cb_imax(spread_old[i], spread_old[i + 1] - 1024);
  for (int c = 0; c < C; c++)
    for (int i = cb_imax(2, start); i < end - 1; i++) {
      i32 x1 = cb_imax(0, newE[i + c * CB_NBANDS]);
      i32 x2 = cb_imax(0, spread_old[i]);
      mean_diff = add32(mean_diff, cb_imax(0, sub16(x1, x2)));
    }
  mean_diff = mean_diff / (C * (end - 1 - cb_imax(2, start)));
  return mean_diff > 1024;
}
