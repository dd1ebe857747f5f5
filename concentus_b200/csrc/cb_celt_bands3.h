// cb_celt_bands3.h -- quant_all_bands, decoder side (bands.c:1337-1502), as three phases per frame:
//
//   1. PARSE   the range decoder walks bands / stereo splits / partitions exactly as the reference
//              does, but touches no spectrum.  It leaves a TAPE (FrameTape, cb_state.h): one record per
//              LEAF of the split trees in walk order (PVQ index and its parameters, the gain the angles
//              above it give it, where its collapse bits go in the band's mask) and one record per
//              BAND (stereo mode and angle, folding source, leaf counts).  Nothing in this phase has any
//              width, so it runs either warp-uniform or -- in the parse kernel (cb_parse_kernels.cu) --
//              ONE THREAD PER STREAM.
//   2. PULSES  the leaves are independent of each other: ONE LANE PER LEAF turns index -> pulse vector
//              (cwrsi is a serial walk), normalises, applies the spreading rotation and leaves the
//              collapse mask.
//   3. REBUILD band by band from the records, no tree walk: folding / noise fill of the leaves
//              without pulses, Hadamard / Haar recombination, stereo merge, folding memory for the
//              bands to come -- the lane-parallel vector work.
//
// Nothing the decoder reads from the bitstream depends on the spectrum (collapse masks and folding
// only steer reconstruction), which is what makes the split legal.  The tape is cut into segments
// (runs of bands) whose leaf masks fit the rebuild warp's shared memory; phases 2 + 3 go segment by
// segment.
#pragma once

// Path counters of the host emulation build (tests/emu, CB_HOST_EMU): the CPU tests assert that the fixtures they
// compare with the reference really take every branch of the rebuild below.  Nothing in a device build.
#if defined(CB_HOST_EMU)
enum { CB_COV_EMPTY_IN_TREE, CB_COV_EMPTY_WHOLE_CALL, CB_COV_MASKED, CB_COV_ZERO, CB_COV_NOISE, CB_COV_FOLD, CB_COV_SEGMENTS3,
       CB_COV_SWITCH, CB_COV_JOINT2, CB_COV_N1_STEREO, CB_COV_N1_MONO, CB_COV_SIDE_FIRST, CB_COV_DUAL, CB_COV_RECOMBINE,
       CB_COV_TIME_DIVIDE, CB_COV_SCRATCH, CB_COV_INV, CB_COV_COUNT };
static long cb_cov[CB_COV_COUNT];
#define CB_COV(k) (cb_cov[k]++)
#else
#define CB_COV(k) ((void)0)
#endif

// Writer side of the tape (PARSE): cursors in registers, the tape itself in global memory.
struct TapeW {
  FrameTape* t;
  int jobs;                         // leaf records written so far
  int seg_jobs;                     // ... of them in the open segment
  int call_empty;                   // leaves without pulses since the band call began
  int nseg, seg_band0;
};
// Reader side (REBUILD)
struct B3Cur {
  int jobs;                         // REBUILD: next leaf, counted from the segment's first
  int seg_job0;                     // REBUILD: the segment's first leaf record
  TapeW* tw;                        // PARSE only
  const TapeJob* seg_jobs;          // REBUILD only: the segment's leaf records (tape, global memory)
};

CB_FN void b3_push_job(B3Cur& c, int xoff, int N, int K, u32 idx, i32 gain, int B, int spread, int sh) {
  TapeW& tw = *c.tw;
  TapeJob jb;
  jb.idx = idx; jb.xoff = (u16)xoff; jb.N = (i16)N; jb.K = (i16)K; jb.gain = (i16)gain; jb.B = (u8)B; jb.spread = (u8)spread;
  jb.sh = (u8)sh; jb.pad = 0;
  tw.t->jobs[tw.jobs++] = jb;
  tw.seg_jobs++;
}
// close the open segment at band `band` (exclusive end) and start the next one there
CB_FN void b3_close_segment(TapeW& tw, int band) {
  TapeSeg sg;
  sg.band0 = (u8)tw.seg_band0; sg.band1 = (u8)band;
  sg.jobs = (u16)tw.seg_jobs;
  tw.t->seg[tw.nseg++] = sg;
  tw.seg_jobs = 0;
  tw.seg_band0 = band;
}
// Upper bound of what one band can add: L = most leaves of one channel's partition tree (a node splits
// while LM >= 0 and N > 2); per channel L leaves (a record on the tape, a mask in the rebuild warp's buffer).
CB_FN int b3_band_leaves(int N, int LM) {
  int L = 1;
  CB_NOUNROLL while (LM >= 0 && N > 2) { L <<= 1; N >>= 1; LM--; }
  return L;
}

// ---- phase 2: one leaf on one lane --------------------------------------------------------------
// cwrsi with per-lane table reads (global copy of the U(n,k) table; the constant-memory one wants a
// uniform index).  Same walk as pvq_decode above.
CB_FN u32 b3_u(int row, int col) { return cb_ldg(&ct_pvq_u_data_g[cb_ldg(&ct_pvq_u_row_g[row]) + col]); }
CB_FN i32 b3_cwrsi(i16* y, int n, int k, u32 i) {
  i32 yy = 0;
  u32 p, q;
  int s, k0;
  i16 val;
  CB_NOUNROLL while (n > 2) {
    if (k >= n) {
      const int rowoff = cb_ldg(&ct_pvq_u_row_g[n]);
      p = cb_ldg(&ct_pvq_u_data_g[rowoff + k + 1]);
      s = -(int)(i >= p);
      i -= p & (u32)s;
      k0 = k;
      q = cb_ldg(&ct_pvq_u_data_g[rowoff + n]);
      if (q > i) {
        k = n;
        CB_NOUNROLL do p = b3_u(--k, n); while (p > i);
      } else {
        CB_NOUNROLL for (p = cb_ldg(&ct_pvq_u_data_g[rowoff + k]); p > i; p = cb_ldg(&ct_pvq_u_data_g[rowoff + k])) k--;
      }
      i -= p;
      val = (i16)((k0 - k + s) ^ s);
      *y++ = val;
      yy = mac16_16(yy, val, val);
    } else {
      p = b3_u(k, n);
      q = b3_u(k + 1, n);
      if (p <= i && i < q) {
        i -= p;
        *y++ = 0;
      } else {
        s = -(int)(i >= q);
        i -= q & (u32)s;
        k0 = k;
        CB_NOUNROLL do p = b3_u(--k, n); while (p > i);
        i -= p;
        val = (i16)((k0 - k + s) ^ s);
        *y++ = val;
        yy = mac16_16(yy, val, val);
      }
    }
    n--;
  }
  p = 2 * (u32)k + 1;
  s = -(int)(i >= p);
  i -= p & (u32)s;
  k0 = k;
  k = (int)((i + 1) >> 1);
  if (k) i -= 2 * (u32)k - 1;
  val = (i16)((k0 - k + s) ^ s);
  *y++ = val;
  yy = mac16_16(yy, val, val);
  s = -(int)i;
  val = (i16)((k + s) ^ s);
  *y = val;
  yy = mac16_16(yy, val, val);
  return yy;
}
// exp_rotation1 (vq.c:43-67), one lane
CB_FN void b3_rot1(i16* x, int len, int st, i32 c, i32 s) {
  const i32 ms = (i16)(-s);
  CB_NOUNROLL for (int i = 0; i < len - st; i++) {
    const i32 x1 = x[i], x2 = x[i + st];
    x[i + st] = ext16(pshr32(mac16_16(mult16_16(c, x2), s, x1), 15));
    x[i] = ext16(pshr32(mac16_16(mult16_16(c, x1), ms, x2), 15));
  }
  CB_NOUNROLL for (int i = len - 2 * st - 1; i >= 0; i--) {
    const i32 x1 = x[i], x2 = x[i + st];
    x[i + st] = ext16(pshr32(mac16_16(mult16_16(c, x2), s, x1), 15));
    x[i] = ext16(pshr32(mac16_16(mult16_16(c, x1), ms, x2), 15));
  }
}
// alg_unquant after the index has been read (vq.c:329-347): cwrsi, normalise_residual (:117-137),
// extract_collapse_mask (:139-159), exp_rotation with dir = -1 (:70-115).  X = the leaf, in place.
CB_FN u32 b3_leaf(i16* X, int N, int K, u32 idx, i32 gain, int B, int spread) {
  const i32 Ryy = b3_cwrsi(X, N, K, idx);
  const int k = cb_ilog2(Ryy) >> 1;
  const i32 t = vshr32(Ryy, 2 * (k - 7));
  const i16 g = (i16)mult16_16_p15(cb_rsqrt_norm(t), gain);
  u32 mask = 0;
  const int N0 = N / B;
  CB_NOUNROLL for (int b = 0; b < B; b++) {
    i32 any = 0;
    CB_NOUNROLL for (int j = 0; j < N0; j++) {
      const i32 v = X[b * N0 + j];
      any |= v;
      X[b * N0 + j] = ext16(pshr32(mult16_16(g, v), k + 1));
    }
    if (any) mask |= 1u << b;
  }
  if (B <= 1) mask = 1;
  if (2 * K < N && spread != CB_SPREAD_NONE) {
    const int factor = spread == 1 ? 15 : spread == 2 ? 10 : 5;
    const i16 gn = (i16)cb_div(mult16_16(CB_Q15ONE, N), N + factor * K);
    const i16 theta = (i16)(mult16_16_q15(gn, gn) >> 1);
    const i16 c = cb_cos_norm(theta);
    const i16 s = cb_cos_norm(sub16(CB_Q15ONE, theta));
    int stride2 = 0;
    if (N >= 8 * B) {
      stride2 = 1;
      CB_NOUNROLL while ((stride2 * stride2 + stride2) * B + (B >> 2) < N) stride2++;
    }
    const int len = N / B;
    CB_NOUNROLL for (int i = 0; i < B; i++) {
      if (stride2) b3_rot1(X + i * len, len, stride2, s, c);
      b3_rot1(X + i * len, len, 1, c, s);
    }
  }
  return mask;
}
// Phase 2: lanes take the segment's leaves (records in the tape, global memory); the collapse mask of leaf j,
// already at its place in the band's mask, goes to w.b3_buf[j].  Leaves without pulses wait for phase 3.
CB_FN void b3_run_jobs(EntWork& w, const TapeJob* jobs, int njobs) {
  CB_WSYNC();
  CB_LANE_LOOP(j, njobs) {
    const TapeJob jb = jobs[j];
    u32 m = 0;
    if (jb.K > 0) m = b3_leaf(w.X + jb.xoff, jb.N, jb.K, jb.idx, jb.gain, jb.B, jb.spread) << jb.sh;
    w.b3_buf[j] = (u16)m;
  }
  CB_WSYNC();
}

// ---- leaf of the split recursion (bands.c:966-1038), PARSE: the pulse count and the PVQ index -------------
CB_FN void b3_leaf_node(EntWork& w, BandCtx& cx, RangeDec& rd, B3Cur& cur, int xo, int N, int b, int B, int LM,
                        i32 gain, int sh, int fz) {
  int q = bits2pulses_lanes(cx.i, LM, b);
  int curr_bits = pulses2bits(cx.i, LM, q);
  cx.remaining_bits -= curr_bits;
  CB_NOUNROLL while (cx.remaining_bits < 0 && q > 0) {
    cx.remaining_bits += curr_bits;
    q--;
    curr_bits = pulses2bits(cx.i, LM, q);
    cx.remaining_bits -= curr_bits;
  }
  if (q != 0) {
    const int K = get_pulses(q);
    const u32 idx = rd_uint(rd, pvq_u(N, K) + pvq_u(N, K + 1));
    b3_push_job(cur, xo, N, K, idx, gain, B, cx.spread, sh);
  } else {
    b3_push_job(cur, xo, N, 0, (u32)fz, gain, B, 0, sh);
    cur.tw->call_empty++;
  }
}

// ---- quant_partition (bands.c:864-1040) as an explicit stack, PARSE ----------------------------------
// Besides the bitstream work this tracks, for every node, what the rebuild half needs of the recursion:
// the gain (products of the split angles' cosines), where the node's collapse bits sit in the band's mask
// (bands.c:957-960: the side half is shifted by B0 >> 1 -- the same shift its share of `fill` had on the way
// down, bands.c:933,947) and whether an angle of 0 / 16384 above it masked that share off (bands.c:789-799).
CB_FN void b3_partition(EntWork& w, BandCtx& cx, RangeDec& rd, B3Cur& cur, int xo0, int N0, int b0, int B_0, int LM0,
                        i32 gain0) {
  PartFrame* st = w.stack;
  int sp = 0;
  st[0].xo = (u16)xo0; st[0].N = N0; st[0].b = b0; st[0].B = B_0;
  st[0].LM = LM0; st[0].gain = gain0; st[0].sh = 0; st[0].fz = 0; st[0].stage = 0;
  CB_NOUNROLL while (sp >= 0) {
    PartFrame& f = st[sp];
    if (f.stage == 0) {
      int N = f.N, b = f.b, B = f.B, LM = f.LM;
      // (bands.c:893: b > cache[cache[0]] + 12, from the derived table)
      const int split = LM != -1 && N > 2 && b > ct_split_thresh[(LM + 1) * CB_NBANDS + cx.i];
      if (split) {
        const int B0 = B;
        N >>= 1;
        LM -= 1;
        B = (B + 1) >> 1;
        Split& s = w.sp_store;
        dec_theta(rd, s, cx.i, cx.intensity, cx.remaining_bits, N, b, B, B0, LM, 0);
        b = s.b;
        int delta = s.delta;
        const int itheta = s.itheta, imid = s.imid, iside = s.iside;
        const int s_qalloc = s.qalloc;
        if (B0 > 1 && (itheta & 0x3fff)) {
          if (itheta > 8192) delta -= delta >> (4 - LM);
          else delta = cb_imin(0, delta + (N << CB_BITRES >> (5 - LM)));
        }
        const int mbits = cb_imax(0, cb_imin(b, (b - delta) / 2));
        const int sbits = b - mbits;
        cx.remaining_bits -= s_qalloc;
        const int mid_first = mbits >= sbits;
        f.N = N; f.B = B; f.LM = LM; f.dsh = (u8)(B0 >> 1);
        f.mbits = mbits; f.sbits = sbits; f.itheta = itheta;
        f.rebal0 = cx.remaining_bits;
        f.gmid = mult16_16_p15(f.gain, imid);
        f.gside = mult16_16_p15(f.gain, iside);
        f.stage = mid_first ? 1 : 3;             // 1: mid first, 3: side first
        PartFrame& c = st[sp + 1];
        c.N = N; c.B = B; c.LM = LM; c.stage = 0;
        if (mid_first) {
          c.xo = f.xo; c.b = mbits; c.gain = (i16)f.gmid;
          c.sh = f.sh; c.fz = (u8)(f.fz | (itheta == 16384));
        } else {
          c.xo = f.xo + N; c.b = sbits; c.gain = (i16)f.gside;
          c.sh = (u8)(f.sh + f.dsh); c.fz = (u8)(f.fz | (itheta == 0));
        }
        sp++;
      } else {
        b3_leaf_node(w, cx, rd, cur, f.xo, N, b, B, LM, f.gain, f.sh, f.fz);
        sp--;
      }
    } else if (f.stage == 1) {                   // mid done, now side: the node itself becomes its side half
      const i32 rebalance = f.mbits - (f.rebal0 - cx.remaining_bits);
      if (rebalance > 3 << CB_BITRES && f.itheta != 0) f.sbits += rebalance - (3 << CB_BITRES);
      f.xo = (u16)(f.xo + f.N); f.b = f.sbits; f.gain = (i16)f.gside;
      f.sh = (u8)(f.sh + f.dsh); f.fz = (u8)(f.fz | (f.itheta == 0));
      f.stage = 0;
    } else {                                     // side done, now mid: the node itself becomes its mid half
      const i32 rebalance = f.sbits - (f.rebal0 - cx.remaining_bits);
      if (rebalance > 3 << CB_BITRES && f.itheta != 16384) f.mbits += rebalance - (3 << CB_BITRES);
      f.b = f.mbits; f.gain = (i16)f.gmid;
      f.fz = (u8)(f.fz | (f.itheta == 16384));
      f.stage = 0;
    }
  }
}

// ---- REBUILD: the leaves of one band call, in the order the walk reached them ----------------------------
// Leaves with pulses were decoded in phase 2.  What is left of quant_partition is the collapse mask of the call --
// the OR of the leaves' masks, each at its place (bands.c:957-960) -- and the leaves without pulses
// (bands.c:1003-1036): zeros, noise, or the folded spectrum with a +-1/256 dither, renormalised to the leaf's gain.
// `note` = the number of leaves, + 128 when one of them is without pulses.
CB_FN u32 b3_leaves_rebuild(EntWork& w, BandCtx& cx, B3Cur& cur, int note, int xo, const i16* lowband, int fill) {
  const int nl = note & 127;
  const int j0 = cur.jobs;
  cur.jobs += nl;
  u32 cm = 0;
  CB_LANE_LOOP(l, nl) cm |= w.b3_buf[j0 + l];
  cm = cb_wor(cm);
  if (!(note & 128)) return cm;
  CB_NOUNROLL for (int l = 0; l < nl; l++) {
    const TapeJob jb = cur.seg_jobs[j0 + l];
    if (jb.K != 0) continue;
    if (nl > 1) CB_COV(CB_COV_EMPTY_IN_TREE); else CB_COV(CB_COV_EMPTY_WHOLE_CALL);
    if (jb.idx) CB_COV(CB_COV_MASKED);
    const int N = jb.N;
    i16* const X = w.X + jb.xoff;
    const u32 cm_mask = (u32)((1ull << jb.B) - 1);
    const u32 lf = jb.idx ? 0u : (((u32)fill >> jb.sh) & cm_mask);
    CB_WSYNC();
    if (!lf) {
      CB_COV(CB_COV_ZERO);
      CB_LANE_LOOP(j, N) X[j] = 0;
      CB_WSYNC();
      continue;
    }
    if (lowband == nullptr) CB_COV(CB_COV_NOISE); else CB_COV(CB_COV_FOLD);
    const u32 seed0 = cx.seed;
    if (lowband == nullptr) {                                     // noise
      CB_LANE_LOOP(j, N) X[j] = (i16)((i32)cb_lcg_skip(seed0, (u32)j + 1) >> 20);
      cm |= cm_mask << jb.sh;
    } else {                                                      // folded spectrum, +-1/256 dither
      const i16* lo = lowband + ((int)jb.xoff - xo);
      CB_LANE_LOOP(j, N) {
        const u32 sd = cb_lcg_skip(seed0, (u32)j + 1);
        const i32 t = (sd & 0x8000u) ? 4 : -4;
        X[j] = (i16)(lo[j] + t);
      }
      cm |= lf << jb.sh;
    }
    cx.seed = cb_lcg_skip(seed0, (u32)N);
    v_renormalise(X, N, jb.gain);
  }
  return cm & 0xffffu;
}

// ---- quant_band (bands.c:1044-1172), PARSE ----------------------------------------------------------
// What the band call reads from the bitstream: a sign when N == 1 (bands.c:819-858), else the split tree on the
// block layout tf_change leaves (the Haar / interleave steps themselves are the rebuild's).  Returns the note for
// the band record: the number of leaves the call wrote, + 128 when one of them is without pulses; N == 1: the sign.
CB_FN int b3_band_parse(EntWork& w, BandCtx& cx, RangeDec& rd, B3Cur& cur, int xo, int N, int b, int B, int LM,
                        i32 gain) {
  if (N == 1) {
    int sign = 0;
    if (cx.remaining_bits >= 1 << CB_BITRES) {
      sign = (int)rd_bits(rd, 1);
      cx.remaining_bits -= 1 << CB_BITRES;
    }
    return sign;
  }
  int tf_change = cx.tf_change;
  int N_B = N >> cb_ilog2(B);                                    // block counts are powers of two
  if (tf_change > 0) {
    B >>= tf_change;
    N_B <<= tf_change;
  }
  CB_NOUNROLL while ((N_B & 1) == 0 && tf_change < 0) {
    B <<= 1;
    N_B >>= 1;
    tf_change++;
  }
  TapeW& tw = *cur.tw;
  const int j0 = tw.jobs;
  tw.call_empty = 0;
  b3_partition(w, cx, rd, cur, xo, N, b, B, LM, gain);
  return (tw.jobs - j0) | (tw.call_empty ? 128 : 0);
}

// What the parse walk carries from band to band.
struct B3Parse { i32 balance; int dual_stereo, lowband_offset, update_lowband; };

// One band of quant_all_bands (bands.c:1383-1498) with quant_band_stereo (bands.c:1176-1334) folded in, PARSE:
// the band's bit budget, stereo angle and split trees from the bitstream, and the band record for the rebuild.
CB_FN void b3_one_band_parse(EntWork& w, RangeDec& rd, B3Cur& cur, BandCtx& cx, B3Parse& ps, int i, int start, int end,
                             int C, int Nfrm, int B, int spread, int intensity, i32 total_bits, int LM, int codedBands) {
  const int M = 1 << LM;
  const bool stereo = C == 2;
  const int norm_offset = M * ebands(start);
  cx.i = i;
  const int xo = M * ebands(i), yo = Nfrm + xo;                  // the band in X, per channel
  const int N = M * ebands(i + 1) - xo;
  const int tf_change = w.tf_res[i];
  cx.tf_change = tf_change;
  const i32 tell = (i32)rd_tell_frac(rd);
  if (i != start) ps.balance -= tell;
  const i32 remaining_bits = total_bits - tell - 1;
  cx.remaining_bits = remaining_bits;
  int b = 0;
  if (i <= codedBands - 1) {
    const i32 curr_balance = ps.balance / cb_imin(3, codedBands - i);     // celt_sudiv
    b = cb_imax(0, cb_imin(16383, cb_imin(remaining_bits + 1, w.pulses[i] + curr_balance)));
  }
  TapeBand r;
  r.job0 = (u16)cur.tw->jobs;
  r.nl[0] = r.nl[1] = 0;
  r.itheta = 0; r.imid = 0; r.iside = 0;
  int flags = 0;
  // folding source and the bands whose collapse masks it brings (bands.c:1408-1428)
  if (xo - N >= norm_offset && (ps.update_lowband || ps.lowband_offset == 0)) ps.lowband_offset = i;
  int effective_lowband = -1, fold_start = 0, fold_end = 0;
  if (ps.lowband_offset != 0 && (spread != CB_SPREAD_AGGRESSIVE || B > 1 || tf_change < 0)) {
    effective_lowband = cb_imax(0, M * ebands(ps.lowband_offset) - norm_offset - N);
    fold_start = ps.lowband_offset;
    CB_NOUNROLL while (M * ebands(--fold_start) > effective_lowband + norm_offset) {}
    fold_end = ps.lowband_offset - 1;
    CB_NOUNROLL while (M * ebands(++fold_end) < effective_lowband + norm_offset + N) {}
  }
  r.eff_lowband = (i16)effective_lowband;
  r.fold0 = (u8)fold_start; r.fold1 = (u8)fold_end;
  if (ps.dual_stereo && i == intensity) {
    ps.dual_stereo = 0;
    flags |= CB_TBF_SWITCH;
  }
  // The band call is big: it is instantiated at ONE call site (instruction-cache footprint is what bounds the
  // parse kernel), so the cases below only describe up to two calls and a loop makes them.
  int mode, ncalls = 2;
  int cX0 = xo, cX1 = yo;
  i32 cb0 = 0, cb1 = 0, cg0 = CB_Q15ONE, cg1 = CB_Q15ONE;
  int rebal = 0, blocked = 0;                                    // joint stereo: what the first call leaves goes to the second
  i32 rebalance0 = 0;
  if (ps.dual_stereo) {
    mode = CB_TB_DUAL;
    cb0 = cb1 = b / 2;
  } else if (!stereo) {
    mode = CB_TB_MONO;
    ncalls = 1;
    cb0 = b;
  } else if (N == 1) {                                           // quant_band_n1 on both channels (bands.c:819-858)
    mode = CB_TB_N1;
  } else {
    // stereo split (compute_theta with stereo = 1, bands.c:1200-1220)
    Split& sh = w.sp_store;
    dec_theta(rd, sh, cx.i, cx.intensity, cx.remaining_bits, N, b, B, B, LM, 1);
    const Split s = sh;
    const int itheta = s.itheta;
    r.itheta = (i16)itheta; r.imid = (i16)s.imid; r.iside = (i16)s.iside;
    if (s.inv) flags |= CB_TBF_INV;
    if (N == 2) {                                                // bands.c:1222-1262
      mode = CB_TB_JOINT2;
      ncalls = 1;
      int mbits = s.b, sbits = 0;
      if (itheta != 0 && itheta != 16384) sbits = 1 << CB_BITRES;
      mbits -= sbits;
      cx.remaining_bits -= s.qalloc + sbits;
      if (sbits && rd_bits(rd, 1)) flags |= CB_TBF_N2_NEG;
      cX0 = itheta > 8192 ? yo : xo;
      cb0 = mbits;
    } else {                                                     // bands.c:1263-1318
      mode = CB_TB_JOINT;
      const i32 mbits = cb_imax(0, cb_imin(s.b, (s.b - s.delta) / 2));
      const i32 sbits = s.b - mbits;
      cx.remaining_bits -= s.qalloc;
      rebalance0 = cx.remaining_bits;
      rebal = 1;
      if (mbits >= sbits) {                                      // mid first
        flags |= CB_TBF_MID_FIRST;
        cb0 = mbits; cb1 = sbits; cg1 = s.iside;
        blocked = itheta == 0;
      } else {                                                   // side first
        cX0 = yo; cX1 = xo;
        cb0 = sbits; cb1 = mbits; cg0 = s.iside;
        blocked = itheta == 16384;
      }
    }
  }
  int note0 = 0, note1 = 0;
  CB_NOUNROLL for (int jn = 0; jn < ncalls; jn++) {
    i32 jb = jn ? cb1 : cb0;
    if (jn == 1 && rebal) {
      const i32 rebalance = cb0 - (rebalance0 - cx.remaining_bits);
      if (rebalance > 3 << CB_BITRES && !blocked) jb += rebalance - (3 << CB_BITRES);
    }
    const int nt = b3_band_parse(w, cx, rd, cur, jn ? cX1 : cX0, N, jb, B, LM, jn ? cg1 : cg0);
    if (jn) note1 = nt; else note0 = nt;
  }
  if (N == 1) flags |= (note0 ? CB_TBF_SIGN0 : 0) | (note1 ? CB_TBF_SIGN1 : 0);
  else { r.nl[0] = (u8)note0; r.nl[1] = (u8)note1; }
  r.mode = (u8)mode; r.flags = (u8)flags;
  cur.tw->t->band[i] = r;
  ps.balance += w.pulses[i] + tell;
  ps.update_lowband = b > (N << CB_BITRES);
}

// ---- quant_band (bands.c:1044-1172), REBUILD -------------------------------------------------------------
// The call's spectrum from its leaves: the folding source brought to the block layout of the leaves (Haar steps,
// deinterleave), the leaves without pulses filled from it, the layout undone on the result, the collapse mask and
// the folding memory for the bands to come.  `note` as b3_band_parse returned it; `sign`: N == 1.
CB_FN u32 b3_band_rebuild(EntWork& w, BandCtx& cx, B3Cur& cur, int band, int xo, int N, int B, i16* lowband,
                          i16* lowband_out, i16* lowband_scratch, int fill, int note, int sign) {
  i16* const X = w.X + xo;
  if (N == 1) {                                                  // quant_band_n1 (bands.c:819-858), one channel
    CB_COV(CB_COV_N1_MONO);
    CB_WSYNC();
    X[0] = sign ? -CB_NORM_SCALING : CB_NORM_SCALING;
    CB_WSYNC();
    if (lowband_out) lowband_out[0] = (i16)(X[0] >> 4);
    return 1;
  }
  const int N0 = N;
  int N_B = N >> cb_ilog2(B), B0 = B;                            // block counts are powers of two
  int time_divide = 0, recombine = 0;
  const int longBlocks = B0 == 1;
  int tf_change = cx.tf_change;
  i16* tmp = (i16*)w.tmp32;
  if (tf_change > 0) recombine = tf_change;
  if (lowband_scratch && lowband && (recombine || ((N_B & 1) == 0 && tf_change < 0) || B0 > 1)) {
    CB_COV(CB_COV_SCRATCH);
    CB_WSYNC();
    CB_LANE_LOOP(j, N) lowband_scratch[j] = lowband[j];
    CB_WSYNC();
    lowband = lowband_scratch;
  }
  if (recombine) CB_COV(CB_COV_RECOMBINE);
  CB_NOUNROLL for (int k = 0; k < recombine; k++) {
    if (lowband) v_haar1(lowband, N >> k, 1 << k);
    fill = ct_bit_interleave[fill & 0xF] | ct_bit_interleave[fill >> 4] << 2;
  }
  B >>= recombine;
  N_B <<= recombine;
  CB_NOUNROLL while ((N_B & 1) == 0 && tf_change < 0) {
    if (lowband) v_haar1(lowband, N_B, B);
    fill |= fill << B;
    B <<= 1;
    N_B >>= 1;
    time_divide++;
    tf_change++;
  }
  B0 = B;
  const int N_B0 = N_B;
  if (time_divide) CB_COV(CB_COV_TIME_DIVIDE);
  if (B0 > 1 && lowband) v_deinterleave(lowband, tmp, N_B >> recombine, B0 << recombine, longBlocks);

  u32 cm = b3_leaves_rebuild(w, cx, cur, note, xo, lowband, fill);

  if (B0 > 1) v_interleave(X, tmp, N_B >> recombine, B0 << recombine, longBlocks);
  N_B = N_B0;
  B = B0;
  CB_NOUNROLL for (int k = 0; k < time_divide; k++) {
    B >>= 1;
    N_B <<= 1;
    cm |= cm >> B;
    v_haar1(X, N_B, B);
  }
  CB_NOUNROLL for (int k = 0; k < recombine; k++) {
    cm = ct_bit_deinterleave[cm];
    v_haar1(X, N0 >> k, 1 << k);
  }
  B <<= recombine;
  if (lowband_out) {
    const i16 n = w.sqrtN[band];                                 // celt_sqrt(SHL32(EXTEND32(N0), 22))
    CB_WSYNC();
    CB_LANE_LOOP(j, N0) lowband_out[j] = (i16)mult16_16_q15(n, X[j]);
    CB_WSYNC();
  }
  cm &= (u32)((1 << B) - 1);
  return cm;
}

// One band of quant_all_bands (bands.c:1383-1498) with quant_band_stereo (bands.c:1176-1334) folded in, REBUILD,
// from the band's record.
CB_FN void b3_one_band_rebuild(EntWork& w, B3Cur& cur, BandCtx& cx, int i, int start, int end, int C, int Nfrm, int B,
                               int LM) {
  const TapeBand r = w.bands[i];
  const int M = 1 << LM;
  const int norm_offset = M * ebands(start);
  const int last = (i == end - 1);
  const int xo = M * ebands(i), yo = Nfrm + xo;                  // the band in X, per channel
  const int N = M * ebands(i + 1) - xo;
  i16* const norm = w.norm;
  i16* const norm2 = w.norm + CB_NORM_STRIDE;
  i16* const lowband_scratch = last ? nullptr : w.X + M * ebands(CB_NBANDS - 1);
  i16* const X = w.X + xo;
  i16* const Y = w.X + yo;                                       // (stereo modes only)
  cx.i = i;
  cx.tf_change = w.tf_res[i];
  cur.jobs = (int)r.job0 - cur.seg_job0;
  const int mode = r.mode, flags = r.flags;
  // folding source and the collapse masks it brings
  i16 *lb = nullptr, *lb2 = nullptr, *lout = nullptr, *lout2 = nullptr;
  u32 x_cm = 0, y_cm = 0;
  if (r.eff_lowband >= 0) {
    CB_NOUNROLL for (int f = r.fold0; f < r.fold1; f++) {
      x_cm |= w.collapse[f * C + 0];
      y_cm |= w.collapse[f * C + C - 1];
    }
    lb = norm + r.eff_lowband;
    lb2 = norm2 + r.eff_lowband;
  } else {
    x_cm = y_cm = (u32)((1 << B) - 1);
  }
  if (!last) { lout = norm + xo - norm_offset; lout2 = norm2 + xo - norm_offset; }
  if (flags & CB_TBF_SWITCH) {                                   // dual stereo ends here (bands.c:1430-1437)
    CB_COV(CB_COV_SWITCH);
    const int cnt = xo - norm_offset;
    CB_WSYNC();
    CB_LANE_LOOP(j, cnt) norm[j] = (i16)(((i32)norm[j] + (i32)norm2[j]) >> 1);
    CB_WSYNC();
  }
  // up to two band calls, in the order the parse walk made them (the leaves are on the tape in that order);
  // one call site for both
  const int itheta = r.itheta, imid = r.imid, iside = r.iside;
  int ncalls = 2;
  int cX0 = xo, cX1 = yo;
  int cfill0 = (int)x_cm, cfill1 = (int)y_cm;
  bool side0 = false, side1 = false;                             // the call is the side of a mid / side pair: no folding
  bool second0 = false, second1 = true;                          // the call works on the second channel's folding memory
  if (mode == CB_TB_DUAL) CB_COV(CB_COV_DUAL);
  if (mode == CB_TB_MONO) {
    ncalls = 1;
    cfill0 = (int)(x_cm | y_cm);
  } else if (mode == CB_TB_JOINT2) {
    CB_COV(CB_COV_JOINT2);
    ncalls = 1;
    cX0 = itheta > 8192 ? yo : xo;
    cfill0 = (int)(x_cm | y_cm);
    second0 = false;
  } else if (mode == CB_TB_N1) {
    CB_COV(CB_COV_N1_STEREO);
    side1 = true;                                                // (only X leaves something in the folding memory)
  } else if (mode == CB_TB_JOINT) {
    int fill = (int)(x_cm | y_cm);
    if (itheta == 0) fill &= (1 << B) - 1;
    else if (itheta == 16384) fill &= ((1 << B) - 1) << B;
    if (flags & CB_TBF_MID_FIRST) {
      cfill0 = fill; cfill1 = fill >> B;
      side1 = true;
    } else {
      CB_COV(CB_COV_SIDE_FIRST);
      cX0 = yo; cX1 = xo;
      cfill0 = fill >> B; cfill1 = fill;
      side0 = true;
      second0 = false; second1 = false;
    }
  }
  u32 cm0 = 0, cm1 = 0;
  CB_NOUNROLL for (int jn = 0; jn < ncalls; jn++) {
    const bool side = jn ? side1 : side0;
    const bool second = jn ? second1 : second0;
    i16* const lo = side ? nullptr : (second ? lb2 : lb);
    i16* const lo_out = side ? nullptr : (second ? lout2 : lout);
    const u32 cmr = b3_band_rebuild(w, cx, cur, i, jn ? cX1 : cX0, N, B, lo, lo_out, side ? nullptr : lowband_scratch,
                                    jn ? cfill1 : cfill0, r.nl[jn], flags & (jn ? CB_TBF_SIGN1 : CB_TBF_SIGN0));
    if (jn) cm1 = cmr; else cm0 = cmr;
  }
  if (mode == CB_TB_MONO) {
    x_cm = y_cm = cm0;
  } else if (mode == CB_TB_DUAL) {
    x_cm = cm0;
    y_cm = cm1;
  } else if (mode == CB_TB_N1) {
    x_cm = y_cm = 1;
  } else {
    if (mode == CB_TB_JOINT2) {                                  // bands.c:1248-1261
      const int n2_sign = (flags & CB_TBF_N2_NEG) ? -1 : 1;
      i16* x2 = itheta > 8192 ? Y : X;
      i16* y2 = itheta > 8192 ? X : Y;
      CB_WSYNC();
      y2[0] = (i16)(-n2_sign * x2[1]);
      y2[1] = (i16)(n2_sign * x2[0]);
      const i16 X0 = (i16)mult16_16_q15(imid, X[0]);
      const i16 X1 = (i16)mult16_16_q15(imid, X[1]);
      const i16 Y0 = (i16)mult16_16_q15(iside, Y[0]);
      const i16 Y1 = (i16)mult16_16_q15(iside, Y[1]);
      CB_WSYNC();
      X[0] = (i16)sub16(X0, Y0);
      Y[0] = add16(X0, Y0);
      X[1] = (i16)sub16(X1, Y1);
      Y[1] = add16(X1, Y1);
      CB_WSYNC();
      x_cm = y_cm = cm0;
    } else {
      x_cm = y_cm = cm0 | cm1;
      v_stereo_merge(X, Y, imid, N);
    }
    if (flags & CB_TBF_INV) {
      CB_COV(CB_COV_INV);
      CB_WSYNC();
      CB_LANE_LOOP(j, N) Y[j] = (i16)(-Y[j]);
      CB_WSYNC();
    }
  }
  w.collapse[i * C + 0] = (u8)x_cm;
  w.collapse[i * C + C - 1] = (u8)y_cm;
}

// quant_all_bands (bands.c:1337-1502), decoder side, PARSE half: every band of the frame onto the tape.
// A segment is closed when the next band's leaves might not fit the rebuild warp's mask buffer, and before the band
// whose spectrum slot doubles as scratch for the others (the last band of the mode).
CB_FN void dec_bands_parse(EntWork& w, RangeDec& rd, TapeW& tw, int start, int end, int C, int Nfrm, int shortBlocks,
                           int spread, int dual_stereo, int intensity, i32 total_bits, i32 balance, int LM,
                           int codedBands) {
  const int M = 1 << LM;
  const int B = shortBlocks ? M : 1;
  BandCtx cxp;
  cxp.intensity = intensity;
  cxp.spread = spread;
  cxp.seed = 0;
  B3Parse ps;
  ps.balance = balance;
  ps.dual_stereo = dual_stereo;
  ps.lowband_offset = 0;
  ps.update_lowband = 1;
  B3Cur wr;
  wr.jobs = wr.seg_job0 = 0;
  wr.tw = &tw;
  wr.seg_jobs = nullptr;
  tw.seg_band0 = start;
  tw.call_empty = 0;
  CB_NOUNROLL for (int i = start; i < end; i++) {
    const int L = b3_band_leaves(M * (ebands(i + 1) - ebands(i)), LM);
    const bool full = tw.seg_jobs + 2 * L > CB_B3_BUF;
    if (i > tw.seg_band0 && (full || i == CB_NBANDS - 1)) b3_close_segment(tw, i);
    b3_one_band_parse(w, rd, wr, cxp, ps, i, start, end, C, Nfrm, B, spread, intensity, total_bits, LM, codedBands);
  }
  if (end > tw.seg_band0) b3_close_segment(tw, end);
}

// PULSES + REBUILD halves, segment by segment.  `seed` = celt_lcg_rand state, carried on to anti_collapse.
CB_FN void dec_bands_rebuild(EntWork& w, const FrameTape* t, int start, int end, int C, int Nfrm, int B, int LM,
                             u32& seed) {
  BandCtx cxr;
  cxr.intensity = 0;
  cxr.spread = 0;
  cxr.seed = seed;
  cxr.remaining_bits = 0;
  const int M = 1 << LM;
  CB_WSYNC();
  {                                                              // the band records, and sqrt(N) of every band
    const u32* src = (const u32*)t->band;
    u32* dst = (u32*)w.bands;
    CB_LANE_LOOP(k, (int)(sizeof(TapeBand) / 4) * CB_NBANDS) dst[k] = src[k];
    CB_LANE_LOOP(k, CB_NBANDS) w.sqrtN[k] = (i16)cb_sqrt(shl32(M * (ebands(k + 1) - ebands(k)), 22));
  }
  CB_WSYNC();
  const int nseg = t->nseg;
  if (nseg >= 3) CB_COV(CB_COV_SEGMENTS3);
  int job0 = 0;
  CB_NOUNROLL for (int sgi = 0; sgi < nseg; sgi++) {
    const TapeSeg sg = t->seg[sgi];
    b3_run_jobs(w, t->jobs + job0, sg.jobs);
    B3Cur rdc;
    rdc.jobs = 0;
    rdc.seg_job0 = job0;
    rdc.tw = nullptr;
    rdc.seg_jobs = t->jobs + job0;
    CB_NOUNROLL for (int r = sg.band0; r < sg.band1; r++) b3_one_band_rebuild(w, rdc, cxr, r, start, end, C, Nfrm, B, LM);
    job0 += sg.jobs;
  }
  seed = cxr.seed;
}
