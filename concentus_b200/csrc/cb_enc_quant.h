// cb_enc_quant.h -- the encoder's bitstream writers: coarse/fine energy, tf flags, PVQ search and
// index (alg_quant + icwrs), theta splits and the band loop.  Bit-exact with the encode branches
// of opus-fix/celt/quant_bands.c:157-425, celt/bands.c:337-373,645-1502, celt/vq.c:70-327,375-412,
// celt/cwrs.c:440-461 and tf_encode (celt/celt_encoder.c:714-753).
// The encoder is built without RESYNTH, so nothing is reconstructed: collapse masks, folding,
// gains and the noise seed never reach the bitstream and are not computed here.
#pragma once
#include "cb_common.h"
#include "cb_celt_entropy.h"
#include "cb_enc_state.h"

struct EncBandCtx {
  int i, intensity, spread, tf_change;
  i32 remaining_bits;
  const i32* bandE;
};

// celt_atan01 / celt_atan2p (mathops.h:219-255)
CB_FN i32 enc_atan01(i32 x) {
  return mult16_16_p15(x, add32(32767, mult16_16_p15(x, add32(-21, mult16_16_p15(x, add32(-11943, mult16_16_p15(4936, x)))))));
}
CB_FN i32 enc_atan2p(i32 y, i32 x) {
  if (y < x) {
    i32 arg = cb_div(shl32(y, 15), x);
    if (arg >= 32767) arg = 32767;
    return (i16)enc_atan01((i16)arg) >> 1;
  }
  i32 arg = cb_div(shl32(x, 15), y);
  if (arg >= 32767) arg = 32767;
  return (i16)(25736 - ((i16)enc_atan01((i16)arg) >> 1));
}
// stereo_itheta (vq.c:375-412)
CB_FN_NOINLINE int enc_stereo_itheta(const i16* X, const i16* Y, int stereo, int N) {
  i32 em = 0, es = 0;
  CB_WSYNC();
  if (stereo) {
    CB_LANE_LOOP(i, N) {
      i32 m = (i16)((X[i] >> 1) + (Y[i] >> 1)), s = (i16)((X[i] >> 1) - (Y[i] >> 1));
      em = mac16_16(em, m, m);
      es = mac16_16(es, s, s);
    }
  } else {
    CB_LANE_LOOP(i, N) {
      em = mac16_16(em, X[i], X[i]);
      es = mac16_16(es, Y[i], Y[i]);
    }
  }
  const i32 Emid = add32(1, cb_wsum(em)), Eside = add32(1, cb_wsum(es));
  const i32 mid = (i16)cb_sqrt(Emid), side = (i16)cb_sqrt(Eside);
  return mult16_16_q15(20861, enc_atan2p(side, mid));          // QCONST16(0.63662f,15)
}
// intensity_stereo (bands.c:337-360)
CB_FN_NOINLINE void enc_intensity_stereo(i16* X, const i16* Y, const i32* bandE, int i, int N) {
  const int shift = cb_zlog2(max32(bandE[i], bandE[i + CB_NBANDS])) - 13;
  const i32 left = (i16)vshr32(bandE[i], shift), right = (i16)vshr32(bandE[i + CB_NBANDS], shift);
  const i32 norm = (i16)(1 + cb_sqrt(add32(add32(1, mult16_16(left, left)), mult16_16(right, right))));
  const i32 a1 = (i16)(shl32(left, 14) / norm), a2 = (i16)(shl32(right, 14) / norm);
  CB_WSYNC();
  CB_LANE_LOOP(j, N) X[j] = (i16)(mac16_16(mult16_16(a1, X[j]), a2, Y[j]) >> 14);
  CB_WSYNC();
}
// stereo_split (bands.c:362-373)
CB_FN void enc_stereo_split(i16* X, i16* Y, int N) {
  CB_WSYNC();
  CB_LANE_LOOP(j, N) {
    i32 l = mult16_16(23170, X[j]), r = mult16_16(23170, Y[j]);
    X[j] = (i16)(add32(l, r) >> 15);
    Y[j] = (i16)(sub32(r, l) >> 15);
  }
  CB_WSYNC();
}
// exp_rotation, encoder direction (vq.c:70-115 with dir=1)
CB_FN void v_exp_rotation_enc(i16* X, int len, int stride, int K, int spread) {
  if (2 * K >= len || spread == CB_SPREAD_NONE) return;
  int factor = spread == 1 ? 15 : spread == 2 ? 10 : 5;
  i16 gain = (i16)cb_div(mult16_16(CB_Q15ONE, len), len + factor * K);
  i16 theta = (i16)(mult16_16_q15(gain, gain) >> 1);
  i16 c = cb_cos_norm(theta);
  i16 s = cb_cos_norm(sub16(CB_Q15ONE, theta));
  int stride2 = 0;
  if (len >= 8 * stride) {
    stride2 = 1;
    CB_NOUNROLL while ((stride2 * stride2 + stride2) * stride + (stride >> 2) < len) stride2++;
  }
  len = len / stride;
  v_rot_pass(X, stride, len, 1, c, (i16)(-s));
  if (stride2) v_rot_pass(X, stride, len, stride2, s, (i16)(-c));
}

// icwrs (cwrs.c:440-456): index of a pulse vector.  The reference walks from the last dimension
// with a running pulse count k; term j only needs the suffix count, so each lane takes a contiguous
// run of dimensions, a warp scan supplies the counts to its right, and the terms add up mod 2^32.
CB_FN u32 enc_pvq_u_g(int n, int k) {
  const int lo = n < k ? n : k, hi = n < k ? k : n;
  return cb_ldg(&ct_pvq_u_data_g[cb_ldg(&ct_pvq_u_row_g[lo]) + hi]);
}
CB_FN u32 enc_icwrs(int n, const i16* y) {
#if CB_DEVICE_BUILD
  const int per = (n + 31) >> 5;
  const int lane = CB_LANE();
  const int j0 = lane * per, j1 = cb_imin(n, j0 + per);
  i32 mine = 0;
  for (int j = j0; j < j1; j++) { const int v = y[j]; mine += v < 0 ? -v : v; }
  const i32 total = cb_wsum(mine);
  int k = total - cb_wscan_excl(mine) - mine;         // pulses strictly to the right of this lane's run
  u32 acc = 0;
  for (int j = j1 - 1; j >= j0; j--) {
    const int v = y[j];
    if (j == n - 1) {
      acc += v < 0;
    } else {
      acc += enc_pvq_u_g(n - j, k);
      if (v < 0) acc += enc_pvq_u_g(n - j, k + (v < 0 ? -v : v) + 1);
    }
    k += v < 0 ? -v : v;
  }
  return cb_wsumu(acc);
#else
  int j = n - 1;
  u32 i = y[j] < 0;
  int k = y[j] < 0 ? -y[j] : y[j];
  do {
    j--;
    i += pvq_u(n - j, k);
    const int v = y[j];
    k += v < 0 ? -v : v;
    if (v < 0) i += pvq_u(n - j, k + 1);
  } while (j > 0);
  return i;
#endif
}

// alg_quant (vq.c:161-327): greedy pulse search.  Each lane owns the positions j = lane (mod 32);
// one warp arg-max per pulse.  y/iy/signx live in the EncWork vectors.
CB_FN_NOINLINE void enc_alg_quant(EncWork& w, RangeEnc& enc, i16* X, int N, int K, int spread, int B) {
  i16* y = w.vtmp;
  i16* iy = w.iy;
  i16* signx = w.signx;
  v_exp_rotation_enc(X, N, B, K, spread);
  CB_WSYNC();
  CB_LANE_LOOP(j, N) {
    i16 v = X[j];
    if (v > 0) signx[j] = 1;
    else { signx[j] = -1; X[j] = (i16)(-v); }
    iy[j] = 0;
    y[j] = 0;
  }
  CB_WSYNC();
  i32 xy = 0;
  i32 yy = 0;                                                  // opus_val16 in the reference
  int pulsesLeft = K;
  if (K > (N >> 1)) {
    i32 s = 0;
    CB_LANE_LOOP(j, N) s += X[j];
    i32 sum = cb_wsum(s);
    if (sum <= K) {
      CB_WSYNC();
      CB_LANE_LOOP(j, N) X[j] = j == 0 ? 16384 : 0;
      CB_WSYNC();
      sum = 16384;
    }
    const i32 rcp = (i16)mult16_32_q16(K - 1, cb_rcp(sum));
    i32 pyy = 0, pxy = 0, pk = 0;
    CB_LANE_LOOP(j, N) {
      const i32 q = mult16_16_q15(X[j], rcp);
      const i32 yv = (i16)q;
      iy[j] = (i16)q;
      pyy = mac16_16(pyy, yv, yv);
      pxy = mac16_16(pxy, X[j], yv);
      y[j] = (i16)(yv * 2);
      pk += q;
    }
    yy = (i16)cb_wsum(pyy);
    xy = cb_wsum(pxy);
    pulsesLeft -= cb_wsum(pk);
  }
  if (pulsesLeft > N + 3) {
    const i32 tmp = (i16)pulsesLeft;
    CB_WSYNC();
    yy = (i16)mac16_16(yy, tmp, tmp);
    yy = (i16)mac16_16(yy, tmp, y[0]);
    CB_WSYNC();
    CB_IF_LANE0 iy[0] = (i16)(iy[0] + pulsesLeft);
    pulsesLeft = 0;
  }
  const int T = (N + 31) >> 5;                                 // positions per lane
  CB_NOUNROLL for (int i = 0; i < pulsesLeft; i++) {
    const int rshift = 1 + cb_ilog2(K - pulsesLeft + i + 1);
    yy = (i16)(yy + 1);
    CbBest best;
    best.num = -32767; best.den = 0; best.id = 0x7fffffff;
#if CB_DEVICE_BUILD
    CB_NOUNROLL for (int t = 0, j = CB_LANE(); t < T; t++, j += 32) {
      if (j < N) {
#else
    (void)T;
    for (int j = 0; j < N; j++) {
      {
#endif
        i32 Rxy = (i16)(add32(xy, X[j]) >> rshift);
        const i32 Ryy = (i16)(yy + y[j]);
        Rxy = (i16)mult16_16_q15(Rxy, Rxy);
        if (mult16_16(best.den, Rxy) > mult16_16(Ryy, best.num)) { best.den = Ryy; best.num = Rxy; best.id = j; }
      }
    }
    best = cb_wbest(best);
    xy = add32(xy, X[best.id]);
    yy = best.den;                                             // = ADD16(yy, y[best_id])
    if ((best.id & 31) == CB_LANE() || !CB_DEVICE_BUILD) {     // the owning lane updates its position
      y[best.id] = (i16)(y[best.id] + 2);
      iy[best.id] = (i16)(iy[best.id] + 1);
    }
  }
  CB_WSYNC();
  CB_LANE_LOOP(j, N) {
    if (signx[j] < 0) iy[j] = (i16)(-iy[j]);
  }
  CB_WSYNC();
  const u32 idx = enc_icwrs(N, iy);
  re_uint(enc, idx, pvq_u(N, K) + pvq_u(N, K + 1));
}

// compute_theta, encoder side (bands.c:645-818).  A real call; results through w.sp_store.
// For the mono split Y = X + N.  May rewrite X/Y (intensity_stereo / stereo_split / inversion).
CB_FN_NOINLINE void enc_theta(EncWork& w, RangeEnc& enc, int band, int intensity, i32 remaining_bits,
                              const i32* bandE, i16* X, i16* Y, int N, int b, int B0, int LM, int stereo) {
  int inv = 0;
  const int pulse_cap = ct_logN400[band] + LM * (1 << CB_BITRES);
  const int offset = (pulse_cap >> 1) - (stereo && N == 2 ? 16 : 4);
  int qn = compute_qn(N, b, offset, pulse_cap, stereo);
  if (stereo && band >= intensity) qn = 1;
  int itheta = enc_stereo_itheta(X, Y, stereo, N);
  const i32 tell = (i32)re_tell_frac(enc);
  if (qn != 1) {
    itheta = (itheta * qn + 8192) >> 14;
    if (stereo && N > 2) {                    // step pdf
      const int p0 = 3, x = itheta, x0 = qn / 2;
      const u32 ft = (u32)(p0 * (x0 + 1) + x0);
      re_encode(enc, (u32)(x <= x0 ? p0 * x : (x - 1 - x0) + (x0 + 1) * p0),
                (u32)(x <= x0 ? p0 * (x + 1) : (x - x0) + (x0 + 1) * p0), ft);
    } else if (B0 > 1 || stereo) {            // uniform pdf
      re_uint(enc, (u32)itheta, (u32)qn + 1);
    } else {                                  // triangular pdf
      const int half = qn >> 1;
      const int ft = (half + 1) * (half + 1);
      const int fs = itheta <= half ? itheta + 1 : qn + 1 - itheta;
      const int fl = itheta <= half ? itheta * (itheta + 1) >> 1 : ft - ((qn + 1 - itheta) * (qn + 2 - itheta) >> 1);
      re_encode(enc, (u32)fl, (u32)(fl + fs), (u32)ft);
    }
    itheta = (int)((u32)(itheta * 16384) / (u32)qn);
    if (stereo) {
      if (itheta == 0) enc_intensity_stereo(X, Y, bandE, band, N);
      else enc_stereo_split(X, Y, N);
    }
  } else if (stereo) {
    inv = itheta > 8192;
    if (inv) {
      CB_WSYNC();
      CB_LANE_LOOP(j, N) Y[j] = (i16)(-Y[j]);
      CB_WSYNC();
    }
    enc_intensity_stereo(X, Y, bandE, band, N);
    if (b > 2 << CB_BITRES && remaining_bits > 2 << CB_BITRES) re_bit_logp(enc, inv, 2);
    else inv = 0;
    itheta = 0;
  }
  const int qalloc = (i32)re_tell_frac(enc) - tell;
  b -= qalloc;
  int delta;
  if (itheta == 0) delta = -16384;
  else if (itheta == 16384) delta = 16384;
  else {
    const int imid = bitexact_cos((i16)itheta), iside = bitexact_cos((i16)(16384 - itheta));
    delta = frac_mul16((N - 1) << 7, bitexact_log2tan(iside, imid));
  }
  EncSplit& sp = w.sp_store;
  sp.inv = inv; sp.delta = delta; sp.itheta = itheta; sp.qalloc = qalloc; sp.b = b;
}

// leaf of quant_partition (bands.c:966-994, encoder side)
CB_FN void enc_part_leaf(EncWork& w, EncBandCtx& cx, RangeEnc& enc, i16* X, int N, int b, int B, int LM) {
  int q = bits2pulses_tab(cx.i, LM, b);
  int curr_bits = pulses2bits(cx.i, LM, q);
  cx.remaining_bits -= curr_bits;
  CB_NOUNROLL while (cx.remaining_bits < 0 && q > 0) {
    cx.remaining_bits += curr_bits;
    q--;
    curr_bits = pulses2bits(cx.i, LM, q);
    cx.remaining_bits -= curr_bits;
  }
  if (q != 0) enc_alg_quant(w, enc, X, N, get_pulses(q), cx.spread, B);
}

// quant_partition (bands.c:864-1040, encoder side) as an explicit stack
CB_FN void enc_partition(EncWork& w, EncBandCtx& cx, RangeEnc& enc, i16* X0, int N0, int b0, int B_0, int LM0) {
  EncPartFrame* st = w.stack;
  int sp = 0;
  st[0].X = X0; st[0].N = N0; st[0].b = b0; st[0].B = B_0; st[0].LM = LM0; st[0].stage = 0;
  CB_NOUNROLL while (sp >= 0) {
    EncPartFrame& f = st[sp];
    if (f.stage == 0) {
      int N = f.N, b = f.b, B = f.B, LM = f.LM;
      const u8* cache = pulse_cache(cx.i, LM);
      if (LM != -1 && b > cache[cache[0]] + 12 && N > 2) {
        const int B0 = B;
        N >>= 1;
        LM -= 1;
        B = (B + 1) >> 1;
        enc_theta(w, enc, cx.i, cx.intensity, cx.remaining_bits, cx.bandE, f.X, f.X + N, N, b, B0, LM, 0);
        const EncSplit s = w.sp_store;
        b = s.b;
        int delta = s.delta;
        const int itheta = s.itheta;
        if (B0 > 1 && (itheta & 0x3fff)) {
          if (itheta > 8192) delta -= delta >> (4 - LM);
          else delta = cb_imin(0, delta + (N << CB_BITRES >> (5 - LM)));
        }
        const int mbits = cb_imax(0, cb_imin(b, (b - delta) / 2));
        const int sbits = b - mbits;
        cx.remaining_bits -= s.qalloc;
        f.N = N; f.B = B; f.LM = LM; f.mbits = mbits; f.sbits = sbits; f.itheta = itheta;
        f.rebal0 = cx.remaining_bits;
        f.stage = (mbits >= sbits) ? 1 : 2;    // 1: mid first, 2: side first
        EncPartFrame& c = st[sp + 1];
        c.N = N; c.B = B; c.LM = LM; c.stage = 0;
        if (mbits >= sbits) { c.X = f.X; c.b = mbits; }
        else { c.X = f.X + N; c.b = sbits; }
        sp++;
      } else {
        enc_part_leaf(w, cx, enc, f.X, N, b, B, LM);
        sp--;
      }
    } else if (f.stage == 1) {                 // mid done, now side
      const i32 rebalance = f.mbits - (f.rebal0 - cx.remaining_bits);
      if (rebalance > 3 << CB_BITRES && f.itheta != 0) f.sbits += rebalance - (3 << CB_BITRES);
      EncPartFrame& c = st[sp + 1];
      c.N = f.N; c.B = f.B; c.LM = f.LM; c.stage = 0; c.X = f.X + f.N; c.b = f.sbits;
      f.stage = 3;
      sp++;
    } else if (f.stage == 2) {                 // side done, now mid
      const i32 rebalance = f.sbits - (f.rebal0 - cx.remaining_bits);
      if (rebalance > 3 << CB_BITRES && f.itheta != 16384) f.mbits += rebalance - (3 << CB_BITRES);
      EncPartFrame& c = st[sp + 1];
      c.N = f.N; c.B = f.B; c.LM = f.LM; c.stage = 0; c.X = f.X; c.b = f.mbits;
      f.stage = 3;
      sp++;
    } else {
      sp--;
    }
  }
}

// quant_band_n1 (bands.c:819-858, encoder side)
CB_FN void enc_band_n1(EncBandCtx& cx, RangeEnc& enc, const i16* X, const i16* Y) {
  const i16* x = X;
  const int nch = Y ? 2 : 1;
  for (int c = 0; c < nch; c++) {
    if (cx.remaining_bits >= 1 << CB_BITRES) {
      re_bits(enc, (u32)(x[0] < 0), 1);
      cx.remaining_bits -= 1 << CB_BITRES;
    }
    x = Y;
  }
}

// quant_band (bands.c:1044-1172, encoder side): time/frequency reshaping of X, then the split tree
CB_FN_NOINLINE void enc_band(EncWork& w, EncBandCtx& cx, RangeEnc& enc, i16* X, int N, int b, int B, int LM) {
  int N_B = N / B;
  int recombine = 0;
  const int longBlocks = B == 1;
  int tf_change = cx.tf_change;
  if (N == 1) { enc_band_n1(cx, enc, X, nullptr); return; }
  if (tf_change > 0) recombine = tf_change;
  CB_NOUNROLL for (int k = 0; k < recombine; k++) v_haar1(X, N >> k, 1 << k);
  B >>= recombine;
  N_B <<= recombine;
  CB_NOUNROLL while ((N_B & 1) == 0 && tf_change < 0) {
    v_haar1(X, N_B, B);
    B <<= 1;
    N_B >>= 1;
    tf_change++;
  }
  if (B > 1) v_deinterleave(X, w.vtmp1, N_B >> recombine, B << recombine, longBlocks);
  enc_partition(w, cx, enc, X, N, b, B, LM);
}

// quant_band_stereo (bands.c:1176-1335, encoder side)
CB_FN void enc_band_stereo(EncWork& w, EncBandCtx& cx, RangeEnc& enc, i16* X, i16* Y, int N, int b, int B, int LM) {
  if (N == 1) { CB_ENC_PHASE(w); enc_band_n1(cx, enc, X, Y); return; }
  enc_theta(w, enc, cx.i, cx.intensity, cx.remaining_bits, cx.bandE, X, Y, N, b, B, LM, 1);
  CB_ENC_PHASE(w);                     // stereo angle coded: the CTA's warps start the band's two halves together
  const EncSplit s = w.sp_store;
  b = s.b;
  const int itheta = s.itheta;
  i16* first; i16* second;
  int fbits, sbits2;
  int nbands = 2;
  int side_first = 0;
  if (N == 2) {
    int mbits = b, sbits = 0;
    if (itheta != 0 && itheta != 16384) sbits = 1 << CB_BITRES;
    mbits -= sbits;
    const int c = itheta > 8192;
    cx.remaining_bits -= s.qalloc + sbits;
    i16* x2 = c ? Y : X;
    const i16* y2 = c ? X : Y;
    if (sbits) {
      CB_WSYNC();
      const int sign = (int)x2[0] * y2[1] - (int)x2[1] * y2[0] < 0;
      re_bits(enc, (u32)sign, 1);
    }
    first = x2; fbits = mbits; second = nullptr; sbits2 = 0; nbands = 1;
  } else {
    const int mbits = cb_imax(0, cb_imin(b, (b - s.delta) / 2));
    const int sbits = b - mbits;
    cx.remaining_bits -= s.qalloc;
    side_first = !(mbits >= sbits);
    if (!side_first) { first = X; fbits = mbits; second = Y; sbits2 = sbits; }
    else { first = Y; fbits = sbits; second = X; sbits2 = mbits; }
  }
  const i32 rebal0 = cx.remaining_bits;
  i16* xb = first;
  int bb = fbits;
  CB_NOUNROLL for (int k = 0; k < nbands; k++) {
    enc_band(w, cx, enc, xb, N, bb, B, LM);
    const i32 rebalance = fbits - (rebal0 - cx.remaining_bits);
    if (rebalance > 3 << CB_BITRES && itheta != (side_first ? 16384 : 0)) sbits2 += rebalance - (3 << CB_BITRES);
    xb = second;
    bb = sbits2;
  }
}

// quant_all_bands (bands.c:1337-1502) with encode=1 and no resynthesis
CB_FN void enc_all_bands(EncWork& w, RangeEnc& enc, int start, int end, int C, int Nfrm, int shortBlocks, int spread,
                         int dual_stereo, int intensity, i32 total_bits, i32 balance, int LM, int codedBands) {
  const int M = 1 << LM;
  const int B = shortBlocks ? M : 1;
  EncBandCtx cx;
  cx.bandE = w.bandE;
  cx.intensity = intensity;
  cx.spread = spread;
  CB_NOUNROLL for (int i = start; i < end; i++) {
    CB_ENC_PHASE(w);                   // the CTA's warps code band i together (same code in the instruction caches)
    cx.i = i;
    i16* X = w.X + M * ebands(i);
    i16* Y = C == 2 ? w.X + Nfrm + M * ebands(i) : nullptr;
    const int N = M * ebands(i + 1) - M * ebands(i);
    const i32 tell = (i32)re_tell_frac(enc);
    if (i != start) balance -= tell;
    const i32 remaining_bits = total_bits - tell - 1;
    cx.remaining_bits = remaining_bits;
    int b = 0;
    if (i <= codedBands - 1) {
      const i32 curr_balance = balance / cb_imin(3, codedBands - i);
      b = cb_imax(0, cb_imin(16383, cb_imin(remaining_bits + 1, w.pulses[i] + curr_balance)));
    }
    cx.tf_change = w.tf_res[i];
    if (dual_stereo && i == intensity) dual_stereo = 0;
    if (dual_stereo) {
      CB_ENC_PHASE(w);
      enc_band(w, cx, enc, X, N, b / 2, B, LM);
      enc_band(w, cx, enc, Y, N, b / 2, B, LM);
    } else if (Y != nullptr) {
      enc_band_stereo(w, cx, enc, X, Y, N, b, B, LM);
    } else {
      CB_ENC_PHASE(w);
      enc_band(w, cx, enc, X, N, b, B, LM);
    }
    balance += w.pulses[i] + tell;
  }
}

// ---- energy quantisation (quant_bands.c:144-425, encoder side) --------------------------------
// loss_distortion (quant_bands.c:144-155)
CB_FN i32 enc_loss_distortion(const i16* eBands, const i16* oldEBands, int start, int end, int C) {
  i32 dist = 0;
  for (int c = 0; c < C; c++)
    CB_NOUNROLL for (int i = start; i < end; i++) {
      const i32 d = (i16)((eBands[i + c * CB_NBANDS] >> 3) - (oldEBands[i + c * CB_NBANDS] >> 3));
      dist = mac16_16(dist, d, d);
    }
  return min32(200, dist >> (2 * CB_DB_SHIFT - 6));
}

// quant_coarse_energy_impl (quant_bands.c:157-261).  Serial Laplace coding, run uniformly.
CB_FN_NOINLINE int enc_coarse_impl(RangeEnc& enc, int start, int end, const i16* eBands, i16* oldEBands, i32 budget,
                                   i32 tell, const u8* prob_model, i16* error, int C, int LM, int intra,
                                   i32 max_decay, int lfe) {
  int badness = 0;
  i32 prev[2] = {0, 0};
  i32 coef, beta;
  if (tell + 3 <= budget) re_bit_logp(enc, intra, 3);
  if (intra) { coef = 0; beta = 4915; }
  else { beta = ct_beta_coef[LM]; coef = ct_pred_coef[LM]; }
  CB_NOUNROLL for (int i = start; i < end; i++) {
    CB_NOUNROLL for (int c = 0; c < C; c++) {
      const i32 x = eBands[i + c * CB_NBANDS];
      const i32 oldv = oldEBands[i + c * CB_NBANDS];
      const i32 oldE = max32(-9216, oldv);
      const i32 f = sub32(sub32(shl32(x, 7), pshr32(mult16_16(coef, oldE), 8)), prev[c]);
      int qi = add32(f, 65536) >> (CB_DB_SHIFT + 7);
      const i32 decay_bound = (i16)max32(-28672, sub32(oldv, max_decay));
      if (qi < 0 && x < decay_bound) {
        qi += (int)(sub16(decay_bound, x) >> CB_DB_SHIFT);
        if (qi > 0) qi = 0;
      }
      const int qi0 = qi;
      tell = re_tell(enc);
      const int bits_left = budget - tell - 3 * C * (end - i);
      if (i != start && bits_left < 30) {
        if (bits_left < 24) qi = cb_imin(1, qi);
        if (bits_left < 16) qi = cb_imax(-1, qi);
      }
      if (lfe && i >= 2) qi = cb_imin(qi, 0);
      if (budget - tell >= 15) {
        const int pi = 2 * cb_imin(i, 20);
        re_laplace(enc, &qi, (u32)prob_model[pi] << 7, (int)prob_model[pi + 1] << 6);
      } else if (budget - tell >= 2) {
        qi = cb_imax(-1, cb_imin(qi, 1));
        re_icdf(enc, (2 * qi) ^ -(qi < 0), ct_small_energy_icdf, 2);
      } else if (budget - tell >= 1) {
        qi = cb_imin(0, qi);
        re_bit_logp(enc, -qi, 1);
      } else {
        qi = -1;
      }
      error[i + c * CB_NBANDS] = (i16)(pshr32(f, 7) - shl16(qi, CB_DB_SHIFT));
      badness += qi0 > qi ? qi0 - qi : qi - qi0;
      const i32 q = shl32(qi, CB_DB_SHIFT);
      i32 tmp = add32(add32(pshr32(mult16_16(coef, oldE), 8), prev[c]), shl32(q, 7));
      tmp = max32(-3670016, tmp);
      oldEBands[i + c * CB_NBANDS] = (i16)pshr32(tmp, 7);
      prev[c] = sub32(add32(prev[c], shl32(q, 7)), mult16_16(beta, pshr32(q, 8)));
    }
  }
  return lfe ? 0 : badness;
}

// quant_coarse_energy (quant_bands.c:263-362): intra and inter passes, keep the cheaper one.
// oldEBands is the state's oldBandE; w.error receives the residual; *delayedIntra is a caller local.
CB_FN void enc_coarse_energy(EncWork& w, EncScratch& sc, RangeEnc& enc, int start, int end, int effEnd,
                             i16* oldEBands, u32 budget, int C, int LM, int nbAvailableBytes, int force_intra,
                             i32* delayedIntra, int two_pass, int loss_rate, int lfe = 0) {
  const i16* eBands = w.bandLogE;
  int intra = force_intra || (!two_pass && *delayedIntra > 2 * C * (end - start) && nbAvailableBytes > (end - start) * C);
  const i32 intra_bias = (i32)((budget * (u32)*delayedIntra * (u32)loss_rate) / (u32)(C * 512));
  CB_WSYNC();
  const i32 new_distortion = enc_loss_distortion(eBands, oldEBands, start, effEnd, C);
  const u32 tell = (u32)re_tell(enc);
  if (tell + 3 > budget) two_pass = intra = 0;
  i32 max_decay = 16384;
  if (end - start > 10) max_decay = (i16)min32(max_decay, shl32(nbAvailableBytes, CB_DB_SHIFT - 3));
  if (lfe) max_decay = 3072;
  w.enc_save0 = enc;                                           // enc_start_state
  CB_LANE_LOOP(i, C * CB_NBANDS) w.oldE_intra[i] = oldEBands[i];
  CB_WSYNC();
  int badness1 = 0;
  if (two_pass || intra)
    badness1 = enc_coarse_impl(enc, start, end, eBands, w.oldE_intra, (i32)budget, (i32)tell,
                               ct_e_prob_model + (LM * 2 + 1) * 42, w.error_intra, C, LM, 1, max_decay, lfe);
  CB_WSYNC();
  if (!intra) {
    const i32 tell_intra = (i32)re_tell_frac(enc);
    w.enc_save1 = enc;                                         // enc_intra_state
    const u32 nstart_bytes = w.enc_save0.offs, nintra_bytes = enc.offs;
    u8* intra_buf = enc.buf + nstart_bytes;
    const u32 save_bytes = nintra_bytes - nstart_bytes;
    CB_WSYNC();
    CB_LANE_LOOP(i, (int)save_bytes) sc.intra_bits[i] = intra_buf[i];
    CB_WSYNC();
    enc = w.enc_save0;
    const int badness2 = enc_coarse_impl(enc, start, end, eBands, oldEBands, (i32)budget, (i32)tell,
                                         ct_e_prob_model + (LM * 2 + 0) * 42, w.error, C, LM, 0, max_decay, lfe);
    CB_WSYNC();
    if (two_pass && (badness1 < badness2 ||
                     (badness1 == badness2 && (i32)re_tell_frac(enc) + intra_bias > tell_intra))) {
      enc = w.enc_save1;
      CB_LANE_LOOP(i, (int)save_bytes) intra_buf[i] = sc.intra_bits[i];
      CB_LANE_LOOP(i, C * CB_NBANDS) { oldEBands[i] = w.oldE_intra[i]; w.error[i] = w.error_intra[i]; }
      CB_WSYNC();
      intra = 1;
    }
  } else {
    CB_LANE_LOOP(i, C * CB_NBANDS) { oldEBands[i] = w.oldE_intra[i]; w.error[i] = w.error_intra[i]; }
    CB_WSYNC();
  }
  if (intra) *delayedIntra = new_distortion;
  else *delayedIntra = add32(mult16_32_q15(mult16_16_q15(ct_pred_coef[LM], ct_pred_coef[LM]), *delayedIntra), new_distortion);
}

// quant_fine_energy (quant_bands.c:364-400)
CB_FN void enc_fine_energy(EncWork& w, RangeEnc& enc, int start, int end, i16* oldEBands, int C) {
  CB_WSYNC();
  CB_NOUNROLL for (int i = start; i < end; i++) {
    const int fq = w.fine_quant[i];
    if (fq <= 0) continue;
    const int frac = (i16)(1 << fq);
    CB_NOUNROLL for (int c = 0; c < C; c++) {
      const i32 err = w.error[i + c * CB_NBANDS];
      int q2 = (err + 512) >> (CB_DB_SHIFT - fq);
      if (q2 > frac - 1) q2 = frac - 1;
      if (q2 < 0) q2 = 0;
      re_bits(enc, (u32)q2, (u32)fq);
      const i32 offset = (i16)sub16((add32(shl32(q2, CB_DB_SHIFT), 512)) >> fq, 512);
      const i32 o = oldEBands[i + c * CB_NBANDS];
      CB_WSYNC();
      oldEBands[i + c * CB_NBANDS] = (i16)(o + offset);
      w.error[i + c * CB_NBANDS] = (i16)(err - offset);
    }
  }
  CB_WSYNC();
}

// quant_energy_finalise (quant_bands.c:402-433)
CB_FN void enc_energy_finalise(EncWork& w, RangeEnc& enc, int start, int end, i16* oldEBands, int bits_left, int C) {
  CB_WSYNC();
  for (int prio = 0; prio < 2; prio++) {
    CB_NOUNROLL for (int i = start; i < end && bits_left >= C; i++) {
      if (w.fine_quant[i] >= CB_MAX_FINE_BITS || w.fine_prio[i] != prio) continue;
      CB_NOUNROLL for (int c = 0; c < C; c++) {
        const int q2 = w.error[i + c * CB_NBANDS] < 0 ? 0 : 1;
        re_bits(enc, (u32)q2, 1);
        const i32 offset = ((i32)shl16(q2, CB_DB_SHIFT) - 512) >> (w.fine_quant[i] + 1);
        const i32 o = oldEBands[i + c * CB_NBANDS];
        CB_WSYNC();
        oldEBands[i + c * CB_NBANDS] = (i16)(o + (i16)offset);
        bits_left--;
      }
    }
  }
  CB_WSYNC();
}

// tf_encode (celt_encoder.c:714-753)
CB_FN void enc_tf_encode(EncWork& w, RangeEnc& enc, int start, int end, int isTransient, int LM, int tf_select) {
  u32 budget = enc.storage * 8;
  u32 tell = (u32)re_tell(enc);
  int logp = isTransient ? 2 : 4;
  const int tf_select_rsv = LM > 0 && tell + logp + 1 <= budget;
  budget -= tf_select_rsv;
  int curr = 0, tf_changed = 0;
  CB_WSYNC();
  CB_NOUNROLL for (int i = start; i < end; i++) {
    const int r = w.tf_res[i];
    if (tell + logp <= budget) {
      re_bit_logp(enc, r ^ curr, (u32)logp);
      tell = (u32)re_tell(enc);
      curr = r;
      tf_changed |= curr;
    }
    CB_WSYNC();
    w.tf_res[i] = curr;
    logp = isTransient ? 4 : 5;
  }
  if (tf_select_rsv && ct_tf_select[LM][4 * isTransient + 0 + tf_changed] != ct_tf_select[LM][4 * isTransient + 2 + tf_changed])
    re_bit_logp(enc, tf_select, 1);
  else
    tf_select = 0;
  CB_WSYNC();
  CB_NOUNROLL for (int i = start; i < end; i++) {
    const int r = ct_tf_select[LM][4 * isTransient + 2 * tf_select + w.tf_res[i]];
    CB_WSYNC();
    w.tf_res[i] = r;
  }
  CB_WSYNC();
}

// compute_allocation with encode=1 (rate.c:356-362,390-394,407-408)
struct AllocEncPolicy {
  RangeEnc& enc;
  int prev, signalBandwidth;
  CB_MFN int skip(int codedBands, int start, int j, int band_bits, int band_width, int LM) {
    if (codedBands <= start + 2 ||
        (band_bits > ((j < prev ? 7 : 9) * band_width << LM << CB_BITRES) >> 4 && j <= signalBandwidth)) {
      re_bit_logp(enc, 1, 1);
      return 1;
    }
    re_bit_logp(enc, 0, 1);
    return 0;
  }
  CB_MFN int intensity(int cur, int start, int codedBands) {
    cur = cb_imin(cur, codedBands);
    re_uint(enc, (u32)(cur - start), (u32)(codedBands + 1 - start));
    return cur;
  }
  CB_MFN int dual(int cur) { re_bit_logp(enc, cur, 1); return cur; }
};
