// cb_celt_entropy.h -- kernel A body: everything in a CELT frame that hangs off the range decoder.
//
// One warp per stream.  The warp walks the symbol stream in uniform control flow (header flags,
// coarse/fine energy, tf, bit allocation, per-band theta/PVQ indices) and fans out over its 32
// lanes for the per-band vector work (residual normalisation, spreading rotation chains,
// Hadamard/Haar reordering, folding, stereo merge, renormalisation).  Output: the normalised
// spectrum X, the band log-energies and the post-filter parameters for kernel B, plus the updated
// small state.  Bit-exact with celt_decode_with_ec() up to (not including) celt_synthesis()
// (opus-fix/celt/celt_decoder.c:713-1062) and with the decode half of celt/bands.c, celt/vq.c,
// celt/cwrs.c, celt/rate.c, celt/quant_bands.c, celt/laplace.c.
#pragma once
#include "cb_common.h"
#include "cb_rangedec.h"
#include "cb_state.h"
#include "celt_tables.h"

// ---- small constant tables (uniform index) ------------------------------------------------
CB_UTABLE i8 ct_tf_select[4][8] = {   // celt.c:247-252
    {0, -1, 0, -1, 0, -1, 0, -1}, {0, -1, 0, -2, 1, 0, 1, -1},
    {0, -2, 0, -3, 2, 0, 1, -1},  {0, -2, 0, -3, 3, 0, 1, -1}};
CB_UTABLE u8 ct_trim_icdf[11] = {126, 124, 119, 109, 87, 41, 19, 9, 4, 2, 0};   // celt.h:150
CB_UTABLE u8 ct_spread_icdf[4] = {25, 23, 2, 0};                               // celt.h:152
CB_UTABLE u8 ct_tapset_icdf[3] = {2, 1, 0};                                     // celt.h:154
CB_UTABLE u8 ct_small_energy_icdf[3] = {2, 1, 0};                               // quant_bands.c:142
CB_UTABLE i16 ct_pred_coef[4] = {29440, 26112, 21248, 16384};                   // quant_bands.c:66
CB_UTABLE i16 ct_beta_coef[4] = {30147, 22282, 12124, 6554};                    // quant_bands.c:67
CB_UTABLE u8 ct_ordery[30] = {1, 0, 3, 0, 2, 1, 7, 0, 4, 3, 6, 1, 5, 2,          // bands.c:525-530
                              15, 0, 8, 7, 12, 3, 11, 4, 14, 1, 9, 6, 13, 2, 10, 5};
CB_UTABLE i16 ct_exp2_table8[8] = {16384, 17866, 19483, 21247, 23170, 25267, 27554, 30048};
CB_UTABLE u8 ct_bit_interleave[16] = {0, 1, 1, 1, 2, 3, 3, 3, 2, 3, 3, 3, 2, 3, 3, 3};
CB_UTABLE u8 ct_bit_deinterleave[16] = {0x00, 0x03, 0x0C, 0x0F, 0x30, 0x33, 0x3C, 0x3F,
                                        0xC0, 0xC3, 0xCC, 0xCF, 0xF0, 0xF3, 0xFC, 0xFF};

#define CB_SPREAD_NONE 0
#define CB_SPREAD_LIGHT 1
#define CB_SPREAD_NORMAL 2
#define CB_SPREAD_AGGRESSIVE 3
#define CB_MAX_FINE_BITS 8
#define CB_FINE_OFFSET 21
#define CB_ALLOC_STEPS 6

struct Split { int inv, imid, iside, delta, itheta, qalloc, b; };   // b: updated in/out value

// Three-phase band decoder (cb_celt_bands3.h): what one tape segment may hold (the rebuild warp's copy of it)
#ifndef CB_B3_BUF
// (sized so that a 20 ms stereo frame of up to ~128 kbit/s has bands 0..19 in one segment)
#define CB_B3_BUF 224           // 16-bit entries: the collapse masks of the segment's leaves
#endif

// One level of the split recursion (quant_partition), kept in shared memory.
struct PartFrame {
  i32 b, mbits, sbits, rebal0;
  u16 xo;                         // offset of the node in X
  i16 N, B, LM, gain, stage, itheta, gmid, gside;
  u8 sh;                          // where this node's collapse bits (and its share of `fill`) sit in the band's mask
  u8 fz;                          // an angle above (0 or 16384) leaves this node nothing to fold
  u8 dsh;                         // B0 >> 1 of the split at this node: how far its side half is shifted
  u8 pad;
};

// Per-stream workspace of the entropy half (shared memory): one per warp, or -- in the parse kernel, which is
// built with CB_THREAD_MODE and only ever runs the PARSE half -- one per thread without the spectrum.
#ifndef CB_THREAD_MODE
#define CB_ENTWORK_ALIGN 16
#else
#define CB_ENTWORK_ALIGN 8      // per-thread copies sit at a stride of an odd number of 8-byte words (bank spread)
#endif
struct alignas(CB_ENTWORK_ALIGN) EntWork {
#ifndef CB_THREAD_MODE
  i16 X[2 * CB_MAXN];
#else
  i16 X[2];                      // parse-only build: named by shared code, never touched
#endif
  i16* norm;                     // REBUILD: the stream's folding memory, 2 x CB_NORM_STRIDE values in the scratch arena
                                 //   (FrameTape::norm; lane-parallel reads and writes only, so it need not be on chip)
  i32 tmp32[88];                 // Hadamard temp (as i16) or allocation scratch (4x21 i32)
  i32 pulses[CB_NBANDS];
  i16 bandE[2 * CB_NBANDS];
  i16 cap[CB_NBANDS];
  i16 offsets[CB_NBANDS];
  i8 fine_quant[CB_NBANDS];
  i8 fine_prio[CB_NBANDS];
  i8 tf_res[CB_NBANDS];
  u8 collapse[2 * CB_NBANDS];
  PartFrame stack[5];
  Split sp_store;                // result record of dec_theta
#ifndef CB_THREAD_MODE
  TapeBand bands[CB_NBANDS];     // cb_celt_bands3.h, REBUILD: the frame's band records (copy of FrameTape::band)
  u16 b3_buf[CB_B3_BUF];         //   current tape segment: the collapse masks of its leaves
  i16 sqrtN[CB_NBANDS + 3];      //   sqrt(N) of every band, Q15-ish: the scale of the folding memory (bands.c:1160-1168)
#else
  TapeBand bands[1];
  u16 b3_buf[2];
  i16 sqrtN[2];
#endif
  RangeDec rd;                   // coder registers; in shared memory so the rd_* routines can be real calls
};

struct BandCtx {
  int i, intensity, spread, tf_change;
  i32 remaining_bits;
  u32 seed;
};

CB_FN int ebands(int i) { return ct_eband5ms[i]; }

// ---- rate.h helpers (rate.h:48-87) --------------------------------------------------------
CB_FN int get_pulses(int i) { return i < 8 ? i : (8 + (i & 7)) << ((i >> 3) - 1); }
CB_FN const u8* pulse_cache(int band, int LM) { return ct_cache_bits50 + ct_cache_index50[(LM + 1) * CB_NBANDS + band]; }
CB_FN int bits2pulses(int band, int LM, int bits) {
  const u8* cache = pulse_cache(band, LM);
  int lo = 0, hi = cache[0];
  bits--;
  CB_NOUNROLL for (int it = 0; it < 6; it++) {   // LOG_MAX_PSEUDO
    int mid = (lo + hi + 1) >> 1;
    if ((int)cache[mid] >= bits) hi = mid; else lo = mid;
  }
  if (bits - (lo == 0 ? -1 : (int)cache[lo]) <= (int)cache[hi] - bits) return lo;
  return hi;
}
// The same answer with the lanes looking at 64 entries of the row at once (decoder; in the encoder kernel
// the bisection measured faster).  The reference's 6-step bisection over a non-decreasing row ends at
// hi = the first entry >= bits (or the last entry when none is) and lo = hi - 1 -- checked exhaustively
// over every row and every budget.
// The same answer from the derived table (celt_tables.h: ct_bits2pulses, every (row, budget) of the bisection above
// worked out by the generator; tests/test_kat.py compares the two exhaustively).
CB_FN int bits2pulses_tab(int band, int LM, int bits) {
  if (bits < 0) return bits2pulses(band, LM, bits);
  const int row = ct_pulse_row[(LM + 1) * CB_NBANDS + band];
  return ct_bits2pulses[row * 257 + cb_imin(bits, 256)];
}
CB_FN int bits2pulses_lanes(int band, int LM, int bits) {
#if CB_DEVICE_BUILD
  const u8* cache = pulse_cache(band, LM);
  const int n = cache[0];
  bits--;
  const u8* row = ct_cache_bits50_g + ct_cache_index50[(LM + 1) * CB_NBANDS + band];
  const int lane = CB_LANE();
  const int c0 = cb_ldg(&row[1 + lane]), c1 = cb_ldg(&row[33 + lane]);
  const unsigned m0 = __ballot_sync(0xffffffffu, 1 + lane <= n && c0 >= bits);
  const unsigned m1 = __ballot_sync(0xffffffffu, 33 + lane <= n && c1 >= bits);
  if (!(m0 | m1)) return n;
  const int hi = m0 ? __ffs(m0) : 32 + __ffs(m1);
  const int lo = hi - 1;
  if (bits - (lo == 0 ? -1 : (int)cache[lo]) <= (int)cache[hi] - bits) return lo;
  return hi;
#else
  return bits2pulses_tab(band, LM, bits);       // one thread (the parse kernel, the host build)
#endif
}
CB_FN int pulses2bits(int band, int LM, int pulses) {
  return pulses == 0 ? 0 : pulse_cache(band, LM)[pulses] + 1;
}
CB_FN u32 pvq_u(int n, int k) {   // CELT_PVQ_U (cwrs.c:196)
  int lo = n < k ? n : k, hi = n < k ? k : n;
  return ct_pvq_u_data[ct_pvq_u_row[lo] + hi];
}

// ---- lane-parallel vector helpers -----------------------------------------------------------
// haar1 (bands.c:581-594): independent butterflies
CB_FN_NOINLINE void v_haar1(i16* X, int N0, int stride) {
  N0 >>= 1;
  const int sh = 31 - cb_clz((u32)stride);   // stride is a power of two
  CB_WSYNC();
  CB_LANE_LOOP(p, N0 * stride) {
    int i = p & (stride - 1), j = p >> sh;
    int a = stride * 2 * j + i, b = stride * (2 * j + 1) + i;
    i32 t1 = mult16_16(23170, X[a]);
    i32 t2 = mult16_16(23170, X[b]);
    X[a] = ext16(pshr32(add32(t1, t2), 15));
    X[b] = ext16(pshr32(sub32(t1, t2), 15));
  }
  CB_WSYNC();
}
// deinterleave_hadamard / interleave_hadamard (bands.c:532-579), via the i16 temp
CB_FN void v_deinterleave(i16* X, i16* tmp, int N0, int stride, int hadamard) {
  int N = N0 * stride;
  const int sh = 31 - cb_clz((u32)stride);     // stride is a power of two
  CB_WSYNC();
  CB_LANE_LOOP(p, N) {
    int i = p & (stride - 1), j = p >> sh;     // source index p = j*stride+i
    int row = hadamard ? ct_ordery[stride - 2 + i] : i;
    tmp[row * N0 + j] = X[p];
  }
  CB_WSYNC();
  CB_LANE_LOOP(p, N) X[p] = tmp[p];
  CB_WSYNC();
}
CB_FN void v_interleave(i16* X, i16* tmp, int N0, int stride, int hadamard) {
  int N = N0 * stride;
  const int sh = 31 - cb_clz((u32)stride);     // stride is a power of two
  CB_WSYNC();
  CB_LANE_LOOP(p, N) {
    int i = p & (stride - 1), j = p >> sh;     // destination index p = j*stride+i
    int row = hadamard ? ct_ordery[stride - 2 + i] : i;
    tmp[p] = X[row * N0 + j];
  }
  CB_WSYNC();
  CB_LANE_LOOP(p, N) X[p] = tmp[p];
  CB_WSYNC();
}
// renormalise_vector (vq.c:349-374)
CB_FN_NOINLINE void v_renormalise(i16* X, int N, i32 gain) {
  CB_WSYNC();
  i32 acc = 0;
  CB_LANE_LOOP(j, N) acc = mac16_16(acc, X[j], X[j]);
  i32 E = add32(1, cb_wsum(acc));
  int k = cb_ilog2(E) >> 1;
  i32 t = vshr32(E, 2 * (k - 7));
  i16 g = (i16)mult16_16_p15(cb_rsqrt_norm(t), gain);
  CB_LANE_LOOP(j, N) X[j] = ext16(pshr32(mult16_16(g, X[j]), k + 1));
  CB_WSYNC();
}
// one pass of exp_rotation1 (vq.c:43-67) over `nblocks` blocks of `len` samples; the `st`
// residue classes of a block are independent recurrences, so each lane owns whole chains.
CB_FN_NOINLINE void v_rot_pass(i16* X, int nblocks, int len, int st, i32 c, i32 s) {
  i32 ms = (i16)(-s);
  CB_WSYNC();
  CB_LANE_LOOP(ch, nblocks * st) {
    i16* x = X + (ch / st) * len;
    int r = ch % st;
    CB_NOUNROLL for (int i = r; i < len - st; i += st) {
      i32 x1 = x[i], x2 = x[i + st];
      x[i + st] = ext16(pshr32(mac16_16(mult16_16(c, x2), s, x1), 15));
      x[i] = ext16(pshr32(mac16_16(mult16_16(c, x1), ms, x2), 15));
    }
    int last = len - 2 * st - 1;
    if (last >= r) {
      int i0 = last - ((last - r) % st);
      CB_NOUNROLL for (int i = i0; i >= 0; i -= st) {
        i32 x1 = x[i], x2 = x[i + st];
        x[i + st] = ext16(pshr32(mac16_16(mult16_16(c, x2), s, x1), 15));
        x[i] = ext16(pshr32(mac16_16(mult16_16(c, x1), ms, x2), 15));
      }
    }
  }
  CB_WSYNC();
}
// ---- theta (bands.c:596-818, decoder side) --------------------------------------------------
CB_FN int compute_qn(int N, int b, int offset, int pulse_cap, int stereo) {
  int N2 = 2 * N - 1;
  if (stereo && N == 2) N2--;
  int qb = (b + N2 * offset) / N2;            // celt_sudiv
  qb = cb_imin(b - pulse_cap - (4 << CB_BITRES), qb);
  qb = cb_imin(8 << CB_BITRES, qb);
  int qn;
  if (qb < (1 << CB_BITRES >> 1)) qn = 1;
  else {
    qn = ct_exp2_table8[qb & 0x7] >> (14 - (qb >> CB_BITRES));
    qn = (qn + 1) >> 1 << 1;
  }
  return qn;
}
CB_FN i16 bitexact_cos(i16 x) {              // bands.c:70-80
  i32 tmp = (4096 + (i32)x * x) >> 13;
  i16 x2 = (i16)tmp;
  x2 = (i16)((32767 - x2) + frac_mul16(x2, (-7651 + frac_mul16(x2, (8277 + frac_mul16(-626, x2))))));
  return (i16)(1 + x2);
}
CB_FN int bitexact_log2tan(int isin, int icos) {   // bands.c:82-93
  int lc = cb_ilog((u32)icos), ls = cb_ilog((u32)isin);
  icos <<= 15 - lc;
  isin <<= 15 - ls;
  return (ls - lc) * (1 << 11) + frac_mul16(isin, frac_mul16(isin, -2597) + 7932)
         - frac_mul16(icos, frac_mul16(icos, -2597) + 7932);
}

// A real call (one copy of the code): inputs by value, results through the Split record, which
// lives in shared memory (EntWork::sp).
CB_FN_NOINLINE void dec_theta(RangeDec& rd, Split& sp, int band, int intensity, i32 remaining_bits, int N, int b,
                              int B, int B0, int LM, int stereo) {
  int itheta = 0, inv = 0;
  int pulse_cap = ct_logN400[band] + LM * (1 << CB_BITRES);
  int offset = (pulse_cap >> 1) - (stereo && N == 2 ? 16 : 4);
  int qn = compute_qn(N, b, offset, pulse_cap, stereo);
  if (stereo && band >= intensity) qn = 1;
  i32 tell = (i32)rd_tell_frac(rd);
  if (qn != 1) {
    if (stereo && N > 2) {                    // step pdf
      int p0 = 3, x0 = qn / 2;
      u32 ft = (u32)(p0 * (x0 + 1) + x0);
      int fs = (int)rd_decode(rd, ft);
      int x;
      if (fs < (x0 + 1) * p0) x = fs / p0;
      else x = x0 + 1 + (fs - (x0 + 1) * p0);
      u32 fl = (u32)(x <= x0 ? p0 * x : (x - 1 - x0) + (x0 + 1) * p0);
      u32 fh = (u32)(x <= x0 ? p0 * (x + 1) : (x - x0) + (x0 + 1) * p0);
      rd_update(rd, fl, fh, ft);
      itheta = x;
    } else if (B0 > 1 || stereo) {            // uniform pdf
      itheta = (int)rd_uint(rd, (u32)qn + 1);
    } else {                                  // triangular pdf
      int half = qn >> 1;
      int ft = (half + 1) * (half + 1);
      int fm = (int)rd_decode(rd, (u32)ft);
      int fs, fl;
      if (fm < (half * (half + 1) >> 1)) {
        itheta = (int)((cb_isqrt32(8 * (u32)fm + 1) - 1) >> 1);
        fs = itheta + 1;
        fl = itheta * (itheta + 1) >> 1;
      } else {
        itheta = (int)((2 * (u32)(qn + 1) - cb_isqrt32(8 * (u32)(ft - fm - 1) + 1)) >> 1);
        fs = qn + 1 - itheta;
        fl = ft - ((qn + 1 - itheta) * (qn + 2 - itheta) >> 1);
      }
      rd_update(rd, (u32)fl, (u32)(fl + fs), (u32)ft);
    }
    itheta = (int)((u32)(itheta * 16384) / (u32)qn);
  } else if (stereo) {
    if (b > 2 << CB_BITRES && remaining_bits > 2 << CB_BITRES) inv = rd_bit_logp(rd, 2);
    else inv = 0;
    itheta = 0;
  }
  int qalloc = (i32)rd_tell_frac(rd) - tell;
  b -= qalloc;
  int imid, iside, delta;
  if (itheta == 0) {
    imid = 32767; iside = 0;
    delta = -16384;
  } else if (itheta == 16384) {
    imid = 0; iside = 32767;
    delta = 16384;
  } else {
    imid = bitexact_cos((i16)itheta);
    iside = bitexact_cos((i16)(16384 - itheta));
    delta = frac_mul16((N - 1) << 7, bitexact_log2tan(iside, imid));
  }
  sp.inv = inv; sp.imid = imid; sp.iside = iside; sp.delta = delta; sp.itheta = itheta; sp.qalloc = qalloc;
  sp.b = b;
}


// stereo_merge (bands.c:375-425)
CB_FN void v_stereo_merge(i16* X, i16* Y, i32 mid, int N) {
  CB_WSYNC();
  i32 axp = 0, aside = 0;
  CB_LANE_LOOP(j, N) {
    axp = mac16_16(axp, Y[j], X[j]);
    aside = mac16_16(aside, Y[j], Y[j]);
  }
  i32 xp = cb_wsum(axp), side = cb_wsum(aside);
  xp = mult16_32_q15(mid, xp);
  i16 mid2 = (i16)(mid >> 1);
  i32 m2 = mult16_16(mid2, mid2);
  i32 El = sub32(add32(m2, side), shl32(xp, 1));
  i32 Er = add32(add32(m2, side), shl32(xp, 1));
  if (Er < 161061 || El < 161061) {           // QCONST32(6e-4f, 28)
    CB_LANE_LOOP(j, N) Y[j] = X[j];
    CB_WSYNC();
    return;
  }
  int kl = cb_ilog2(El) >> 1, kr = cb_ilog2(Er) >> 1;
  i32 lgain = cb_rsqrt_norm(vshr32(El, (kl - 7) << 1));
  i32 rgain = cb_rsqrt_norm(vshr32(Er, (kr - 7) << 1));
  if (kl < 7) kl = 7;
  if (kr < 7) kr = 7;
  CB_LANE_LOOP(j, N) {
    i16 l = (i16)mult16_16_p15(mid, X[j]);
    i16 r = Y[j];
    X[j] = ext16(pshr32(mult16_16(lgain, sub16(l, r)), kl + 1));
    Y[j] = ext16(pshr32(mult16_16(rgain, add16(l, r)), kr + 1));
  }
  CB_WSYNC();
}


#include "cb_celt_bands3.h"

// anti_collapse (bands.c:241-335)
CB_FN void v_anti_collapse(EntWork& w, int LM, int C, int Nfrm, int start, int end, const i16* logE,
                           const i16* prev1logE, const i16* prev2logE, u32 seed) {
  const u32 all_blocks = (1u << (1 << LM)) - 1;
  CB_NOUNROLL for (int i = start; i < end; i++) {
    // (nothing to do for a band none of whose short blocks collapsed -- the usual case -- and nothing here but the
    // filled blocks advances the seed)
    if ((w.collapse[i * C] & all_blocks) == all_blocks && (w.collapse[i * C + C - 1] & all_blocks) == all_blocks) continue;
    int N0 = ebands(i + 1) - ebands(i);
    int depth = (int)((u32)(1 + w.pulses[i]) / (u32)N0) >> LM;
    i32 thresh32 = cb_exp2(-(i32)shl16(depth, 10 - CB_BITRES)) >> 1;
    i16 thresh = (i16)mult16_32_q15(16384, min32(32767, thresh32));
    i32 t = N0 << LM;
    int shift = cb_ilog2(t) >> 1;
    t = shl32(t, (7 - shift) << 1);
    i16 sqrt_1 = cb_rsqrt_norm(t);
    CB_NOUNROLL for (int c = 0; c < C; c++) {
      i32 prev1 = prev1logE[c * CB_NBANDS + i];
      i32 prev2 = prev2logE[c * CB_NBANDS + i];
      if (C == 1) {
        prev1 = cb_imax(prev1, prev1logE[CB_NBANDS + i]);
        prev2 = cb_imax(prev2, prev2logE[CB_NBANDS + i]);
      }
      i32 Ediff = (i32)logE[c * CB_NBANDS + i] - cb_imin(prev1, prev2);
      Ediff = max32(0, Ediff);
      i16 r;
      if (Ediff < 16384) {
        i32 r32 = cb_exp2(-(i32)(i16)Ediff) >> 1;
        r = (i16)(2 * cb_imin(16383, r32));
      } else {
        r = 0;
      }
      if (LM == 3) r = (i16)mult16_16_q14(23170, min32(23169, r));
      r = (i16)(cb_imin(thresh, r) >> 1);
      r = (i16)(mult16_16_q15(sqrt_1, r) >> shift);
      i16* X = w.X + c * Nfrm + (ebands(i) << LM);
      int renormalize = 0;
      CB_NOUNROLL for (int k = 0; k < (1 << LM); k++) {
        if (!(w.collapse[i * C + c] & (1 << k))) {
          u32 seed0 = seed;
          CB_WSYNC();
          CB_LANE_LOOP(j, N0) {
            u32 sd = cb_lcg_skip(seed0, (u32)j + 1);
            X[(j << LM) + k] = (sd & 0x8000u) ? r : (i16)(-r);
          }
          seed = cb_lcg_skip(seed0, (u32)N0);
          renormalize = 1;
        }
      }
      if (renormalize) v_renormalise(X, N0 << LM, CB_Q15ONE);
    }
  }
}

// ---- energy decoding (quant_bands.c:435-549) -------------------------------------------------
CB_FN void dec_coarse_energy(EntWork& w, RangeDec& rd, int start, int end, int intra, int C, int LM) {
  const u8* prob = ct_e_prob_model + (LM * 2 + intra) * 42;
  i32 prev[2] = {0, 0};
  i32 coef, beta;
  if (intra) { coef = 0; beta = 4915; }
  else { beta = ct_beta_coef[LM]; coef = ct_pred_coef[LM]; }
  i32 budget = (i32)rd.storage * 8;
  CB_NOUNROLL for (int i = start; i < end; i++) {
    CB_NOUNROLL for (int c = 0; c < C; c++) {
      int qi;
      i32 tell = rd_tell(rd);
      if (budget - tell >= 15) {
        int pi = 2 * cb_imin(i, 20);
        qi = rd_laplace(rd, (u32)prob[pi] << 7, (int)prob[pi + 1] << 6);
      } else if (budget - tell >= 2) {
        qi = rd_icdf(rd, ct_small_energy_icdf, 2);
        qi = (qi >> 1) ^ -(qi & 1);
      } else if (budget - tell >= 1) {
        qi = -rd_bit_logp(rd, 1);
      } else {
        qi = -1;
      }
      i32 q = shl32(qi, CB_DB_SHIFT);
      i16& e = w.bandE[i + c * CB_NBANDS];
      e = (i16)cb_imax(-9216, e);              // -QCONST16(9.f, DB_SHIFT)
      i32 tmp = add32(add32(pshr32(mult16_16(coef, e), 8), prev[c]), shl32(q, 7));
      tmp = max32(-3670016, tmp);              // -QCONST32(28.f, DB_SHIFT+7)
      e = (i16)pshr32(tmp, 7);
      prev[c] = sub32(add32(prev[c], shl32(q, 7)), mult16_16(beta, pshr32(q, 8)));
    }
  }
}
CB_FN void dec_fine_energy(EntWork& w, RangeDec& rd, int start, int end, int C) {
  CB_NOUNROLL for (int i = start; i < end; i++) {
    int fq = w.fine_quant[i];
    if (fq <= 0) continue;
    CB_NOUNROLL for (int c = 0; c < C; c++) {
      i32 q2 = (i32)rd_bits(rd, (u32)fq);
      i32 offset = sub16((add32(shl32(q2, CB_DB_SHIFT), 512)) >> fq, 512);
      w.bandE[i + c * CB_NBANDS] = (i16)(w.bandE[i + c * CB_NBANDS] + (i16)offset);
    }
  }
}
CB_FN void dec_energy_finalise(EntWork& w, RangeDec& rd, int start, int end, int bits_left, int C) {
  CB_NOUNROLL for (int prio = 0; prio < 2; prio++) {
    CB_NOUNROLL for (int i = start; i < end && bits_left >= C; i++) {
      if (w.fine_quant[i] >= CB_MAX_FINE_BITS || w.fine_prio[i] != prio) continue;
      CB_NOUNROLL for (int c = 0; c < C; c++) {
        i32 q2 = (i32)rd_bits(rd, 1);
        i32 offset = ((i32)shl16(q2, CB_DB_SHIFT) - 512) >> (w.fine_quant[i] + 1);
        w.bandE[i + c * CB_NBANDS] = (i16)(w.bandE[i + c * CB_NBANDS] + (i16)offset);
        bits_left--;
      }
    }
  }
}

// tf_decode (celt_decoder.c:352-389)
CB_FN void dec_tf(EntWork& w, RangeDec& rd, int start, int end, int isTransient, int LM) {
  u32 budget = rd.storage * 8;
  u32 tell = (u32)rd_tell(rd);
  int logp = isTransient ? 2 : 4;
  int tf_select_rsv = LM > 0 && tell + logp + 1 <= budget;
  budget -= tf_select_rsv;
  int tf_changed = 0, curr = 0;
  CB_NOUNROLL for (int i = start; i < end; i++) {
    if (tell + logp <= budget) {
      curr ^= rd_bit_logp(rd, (u32)logp);
      tell = (u32)rd_tell(rd);
      tf_changed |= curr;
    }
    w.tf_res[i] = (i8)curr;
    logp = isTransient ? 4 : 5;
  }
  int tf_select = 0;
  if (tf_select_rsv && ct_tf_select[LM][4 * isTransient + 0 + tf_changed] != ct_tf_select[LM][4 * isTransient + 2 + tf_changed])
    tf_select = rd_bit_logp(rd, 1);
  CB_NOUNROLL for (int i = start; i < end; i++)
    w.tf_res[i] = ct_tf_select[LM][4 * isTransient + 2 * tf_select + w.tf_res[i]];
}

// ---- bit allocation (rate.c:248-638), decoder side -------------------------------------------
// The three places where the bitstream is touched go through a coder policy `ac`, so the encoder
// (cb_enc_quant.h) runs the same allocation with its own skip decision.
struct AllocDecPolicy {
  RangeDec& rd;
  CB_MFN int skip(int, int, int, int, int, int) { return rd_bit_logp(rd, 1); }
  CB_MFN int intensity(int, int start, int codedBands) { return start + (int)rd_uint(rd, (u32)(codedBands + 1 - start)); }
  CB_MFN int dual(int) { return rd_bit_logp(rd, 1); }
};
template <class W, class AC>
CB_FN int cb_allocation(W& w, AC& ac, int start, int end, int alloc_trim, int& intensity,
                        int& dual_stereo, i32 total, i32& balance_out, int C, int LM) {
  i32* bits1 = w.tmp32;
  i32* bits2 = w.tmp32 + CB_NBANDS;
  i32* thresh = w.tmp32 + 2 * CB_NBANDS;
  i32* trim_offset = w.tmp32 + 3 * CB_NBANDS;
  i32* bits = w.pulses;
  total = max32(total, 0);
  int skip_start = start;
  int skip_rsv = total >= 1 << CB_BITRES ? 1 << CB_BITRES : 0;
  total -= skip_rsv;
  int intensity_rsv = 0, dual_stereo_rsv = 0;
  if (C == 2) {
    intensity_rsv = ct_log2_frac_table[end - start];
    if (intensity_rsv > total) intensity_rsv = 0;
    else {
      total -= intensity_rsv;
      dual_stereo_rsv = total >= 1 << CB_BITRES ? 1 << CB_BITRES : 0;
      total -= dual_stereo_rsv;
    }
  }
  // One band per lane: the thresholds, both bisections and the interpolation are sums over bands.  "done"
  // of the reference's high-to-low scan (every band at or below the highest one that reaches its threshold)
  // becomes a comparison with that band's index `top`.
  const int nb = end - start;
  CB_WSYNC();
  CB_LANE_LOOP(jj, nb) {
    const int j = start + jj;
    const int wdt = cb_ldg(&ct_eband5ms_g[j + 1]) - cb_ldg(&ct_eband5ms_g[j]);
    thresh[j] = cb_imax(C << CB_BITRES, (3 * wdt << LM << CB_BITRES) >> 4);
    i32 to = C * wdt * (alloc_trim - 5 - LM) * (end - j - 1) * (1 << (LM + CB_BITRES)) >> 6;
    if (wdt << LM == 1) to -= C << CB_BITRES;
    trim_offset[j] = to;
  }
  CB_WSYNC();
  int lo = 1, hi = 11 - 1;                    // nbAllocVectors = 11
  CB_NOUNROLL do {
    const int mid = (lo + hi) >> 1;
#if CB_DEVICE_BUILD
    int top_l = -1;
    CB_LANE_LOOP(jj, nb) {
      const int j = start + jj;
      const int wdt = cb_ldg(&ct_eband5ms_g[j + 1]) - cb_ldg(&ct_eband5ms_g[j]);
      int bitsj = C * wdt * cb_ldg(&ct_band_alloc_g[mid * CB_NBANDS + j]) << LM >> 2;
      if (bitsj > 0) bitsj = cb_imax(0, bitsj + trim_offset[j]);
      bitsj += w.offsets[j];
      bits[j] = bitsj;
      if (bitsj >= thresh[j]) top_l = cb_imax(top_l, j);
    }
    const int top = cb_wmax(top_l);
    i32 psum_l = 0;
    CB_LANE_LOOP(jj, nb) {
      const int j = start + jj;
      const int bitsj = bits[j];
      if (j <= top) psum_l += cb_imin(bitsj, w.cap[j]);
      else if (bitsj >= C << CB_BITRES) psum_l += C << CB_BITRES;
    }
    const i32 psum0 = cb_wsum(psum_l);
#else
    // scalar builds (one thread per stream): the reference's single high-to-low scan (rate.c:567-585)
    i32 psum0 = 0;
    int done = 0;
    CB_NOUNROLL for (int j = end; j-- > start;) {
      const int wdt = ebands(j + 1) - ebands(j);
      int bitsj = C * wdt * ct_band_alloc[mid * CB_NBANDS + j] << LM >> 2;
      if (bitsj > 0) bitsj = cb_imax(0, bitsj + trim_offset[j]);
      bitsj += w.offsets[j];
      if (bitsj >= thresh[j] || done) {
        done = 1;
        psum0 += cb_imin(bitsj, w.cap[j]);
      } else if (bitsj >= C << CB_BITRES) {
        psum0 += C << CB_BITRES;
      }
    }
#endif
    if (psum0 > total) hi = mid - 1; else lo = mid + 1;
  } while (lo <= hi);
  hi = lo--;
  {
    int ss_l = start;
    CB_LANE_LOOP(jj, nb) {
      const int j = start + jj;
      const int wdt = cb_ldg(&ct_eband5ms_g[j + 1]) - cb_ldg(&ct_eband5ms_g[j]);
      int bits1j = C * wdt * cb_ldg(&ct_band_alloc_g[lo * CB_NBANDS + j]) << LM >> 2;
      int bits2j = hi >= 11 ? w.cap[j] : C * wdt * cb_ldg(&ct_band_alloc_g[(hi >= 11 ? 0 : hi) * CB_NBANDS + j]) << LM >> 2;
      if (bits1j > 0) bits1j = cb_imax(0, bits1j + trim_offset[j]);
      if (bits2j > 0) bits2j = cb_imax(0, bits2j + trim_offset[j]);
      if (lo > 0) bits1j += w.offsets[j];
      bits2j += w.offsets[j];
      if (w.offsets[j] > 0) ss_l = cb_imax(ss_l, j);
      bits2j = cb_imax(0, bits2j - bits1j);
      bits1[j] = bits1j;
      bits2[j] = bits2j;
    }
    skip_start = cb_wmax(ss_l);
  }
  // interp_bits2pulses (rate.c:248-525)
  const int alloc_floor = C << CB_BITRES;
  const int stereo = C > 1;
  const int logM = LM << CB_BITRES;
  i32 psum;
  lo = 0;
  hi = 1 << CB_ALLOC_STEPS;
  CB_NOUNROLL for (int it = 0; it < CB_ALLOC_STEPS; it++) {
    const int mid = (lo + hi) >> 1;
#if CB_DEVICE_BUILD
    int top_l = -1;
    CB_LANE_LOOP(jj, nb) {
      const int j = start + jj;
      const int tmp = bits1[j] + (int)((i32)mid * bits2[j] >> CB_ALLOC_STEPS);
      if (tmp >= thresh[j]) top_l = cb_imax(top_l, j);
    }
    const int top = cb_wmax(top_l);
    i32 psum_l = 0;
    CB_LANE_LOOP(jj, nb) {
      const int j = start + jj;
      const int tmp = bits1[j] + (int)((i32)mid * bits2[j] >> CB_ALLOC_STEPS);
      if (j <= top) psum_l += cb_imin(tmp, w.cap[j]);
      else if (tmp >= alloc_floor) psum_l += alloc_floor;
    }
    psum = cb_wsum(psum_l);
#else
    psum = 0;                                                  // rate.c:275-293, one scan
    int done = 0;
    CB_NOUNROLL for (int j = end; j-- > start;) {
      const int tmp = bits1[j] + (int)((i32)mid * bits2[j] >> CB_ALLOC_STEPS);
      if (tmp >= thresh[j] || done) {
        done = 1;
        psum += cb_imin(tmp, w.cap[j]);
      } else if (tmp >= alloc_floor) {
        psum += alloc_floor;
      }
    }
#endif
    if (psum > total) hi = mid; else lo = mid;
  }
  {
    int top_l = -1;
    CB_LANE_LOOP(jj, nb) {
      const int j = start + jj;
      const int tmp = bits1[j] + (lo * bits2[j] >> CB_ALLOC_STEPS);
      if (tmp >= thresh[j]) top_l = cb_imax(top_l, j);
    }
    const int top = cb_wmax(top_l);
    i32 psum_l = 0;
    CB_LANE_LOOP(jj, nb) {
      const int j = start + jj;
      int tmp = bits1[j] + (lo * bits2[j] >> CB_ALLOC_STEPS);
      if (j > top) tmp = tmp >= alloc_floor ? alloc_floor : 0;
      tmp = cb_imin(tmp, w.cap[j]);
      bits[j] = tmp;
      psum_l += tmp;
    }
    psum = cb_wsum(psum_l);
    CB_WSYNC();
  }
  int codedBands;
  CB_NOUNROLL for (codedBands = end;; codedBands--) {
    int j = codedBands - 1;
    if (j <= skip_start) {
      total += skip_rsv;
      break;
    }
    i32 left = total - psum;
    int span = ebands(codedBands) - ebands(start);
    i32 percoeff = (i32)((u32)left / (u32)span);
    left -= span * percoeff;
    int rem = cb_imax(left - (ebands(j) - ebands(start)), 0);
    int band_width = ebands(codedBands) - ebands(j);
    int band_bits = (int)(bits[j] + percoeff * band_width + rem);
    if (band_bits >= cb_imax(thresh[j], alloc_floor + (1 << CB_BITRES))) {
      if (ac.skip(codedBands, start, j, band_bits, band_width, LM)) break;
      psum += 1 << CB_BITRES;
      band_bits -= 1 << CB_BITRES;
    }
    psum -= bits[j] + intensity_rsv;
    if (intensity_rsv > 0) intensity_rsv = ct_log2_frac_table[j - start];
    psum += intensity_rsv;
    if (band_bits >= alloc_floor) {
      psum += alloc_floor;
      bits[j] = alloc_floor;
    } else {
      bits[j] = 0;
    }
  }
  if (intensity_rsv > 0) intensity = ac.intensity(intensity, start, codedBands);
  else intensity = 0;
  if (intensity <= start) {
    total += dual_stereo_rsv;
    dual_stereo_rsv = 0;
  }
  if (dual_stereo_rsv > 0) dual_stereo = ac.dual(dual_stereo);
  else dual_stereo = 0;

  i32 left = total - psum;
  {
    int span = ebands(codedBands) - ebands(start);
    i32 percoeff = (i32)((u32)left / (u32)span);
    left -= span * percoeff;
    CB_NOUNROLL for (int j = start; j < codedBands; j++) bits[j] += (int)percoeff * (ebands(j + 1) - ebands(j));
    CB_NOUNROLL for (int j = start; j < codedBands; j++) {
      int tmp = (int)min32(left, ebands(j + 1) - ebands(j));
      bits[j] += tmp;
      left -= tmp;
    }
  }
  i32 balance = 0;
  int j;
  CB_NOUNROLL for (j = start; j < codedBands; j++) {
    int N0 = ebands(j + 1) - ebands(j);
    int N = N0 << LM;
    i32 bit = bits[j] + balance;
    i32 excess;
    int eb;
    if (N > 1) {
      excess = max32(bit - w.cap[j], 0);
      bits[j] = bit - excess;
      int den = C * N + ((C == 2 && N > 2 && !dual_stereo && j < intensity) ? 1 : 0);
      int NClogN = den * (ct_logN400[j] + logM);
      int offset = (NClogN >> 1) - den * CB_FINE_OFFSET;
      if (N == 2) offset += den << CB_BITRES >> 2;
      if (bits[j] + offset < den * 2 << CB_BITRES) offset += NClogN >> 2;
      else if (bits[j] + offset < den * 3 << CB_BITRES) offset += NClogN >> 3;
      eb = cb_imax(0, (bits[j] + offset + (den << (CB_BITRES - 1))));
      eb = (int)((u32)eb / (u32)den) >> CB_BITRES;
      if (C * eb > (bits[j] >> CB_BITRES)) eb = bits[j] >> stereo >> CB_BITRES;
      eb = cb_imin(eb, CB_MAX_FINE_BITS);
      w.fine_prio[j] = (eb * (den << CB_BITRES) >= bits[j] + offset);
      bits[j] -= C * eb << CB_BITRES;
    } else {
      excess = max32(0, bit - (C << CB_BITRES));
      bits[j] = bit - excess;
      eb = 0;
      w.fine_prio[j] = 1;
    }
    if (excess > 0) {
      int extra_fine = cb_imin(excess >> (stereo + CB_BITRES), CB_MAX_FINE_BITS - eb);
      eb += extra_fine;
      int extra_bits = extra_fine * C << CB_BITRES;
      w.fine_prio[j] = (extra_bits >= excess - balance);
      excess -= extra_bits;
    }
    balance = excess;
    w.fine_quant[j] = eb;
  }
  balance_out = balance;
  CB_NOUNROLL for (; j < end; j++) {
    int eb = bits[j] >> stereo >> CB_BITRES;
    w.fine_quant[j] = eb;
    bits[j] = 0;
    w.fine_prio[j] = (eb < 1);
  }
  return codedBands;
}

CB_FN int dec_allocation(EntWork& w, RangeDec& rd, int start, int end, int alloc_trim, int& intensity,
                         int& dual_stereo, i32 total, i32& balance_out, int C, int LM) {
  AllocDecPolicy ac{rd};
  return cb_allocation(w, ac, start, end, alloc_trim, intensity, dual_stereo, total, balance_out, C, LM);
}

// ---- one CELT frame, entropy half (celt_decoder.c:713-1062 minus synthesis/deemphasis) -------
// Two halves.  celt_frame_parse is everything that hangs off the range decoder -- header flags, energies, tf,
// allocation, the band walk (onto the tape), the frame's tail -- plus the small-state update; it has no width and
// runs one thread or one warp per stream.  celt_frame_rebuild turns the tape into the normalised spectrum (one
// warp per stream) and hands it to the synthesis kernel.
// `pkt`/`len`: CELT payload (shared or global memory).  celt_frame_parse returns 0 or a negative OPUS_* code.
struct EntFrameCfg {
  int LM, C, start, end, accum;
  const RangeDec* rd0;   // hybrid: the coder state the SILK layer left (same packet bytes); else nullptr
};

CB_FN int celt_frame_parse(EntWork& w, CeltDecState* stg, FrameSide* side, FrameTape* tape, const u8* pkt, int len,
                           const EntFrameCfg& cfg, u32* final_range) {
  const int LM = cfg.LM, C = cfg.C, start = cfg.start, end = cfg.end;
  const int M = 1 << LM, N = M * 120;
  // load the small state
  CB_LANE_LOOP(j, 2 * CB_NBANDS) w.bandE[j] = stg->oldBandE[j];
  CB_WSYNC();
  int effEnd = end > CB_NBANDS ? CB_NBANDS : end;

#ifdef CB_THREAD_MODE
  RangeDec rd_regs;                  // one thread per stream: the coder state stays in registers
  RangeDec& rd = rd_regs;
#else
  RangeDec& rd = w.rd;
#endif
  if (cfg.rd0) {                     // celt_decode_with_ec(dec != NULL), celt_decoder.c:813-818
    rd = *cfg.rd0;
    rd.buf = pkt;
    rd.storage = (u32)len;
  } else {
    rd_init(rd, pkt, (u32)len);
  }

  if (C == 1) {
    CB_WSYNC();
    CB_LANE_LOOP(i, CB_NBANDS) w.bandE[i] = (i16)cb_imax(w.bandE[i], w.bandE[CB_NBANDS + i]);
    CB_WSYNC();
  }
  i32 total_bits = len * 8;
  i32 tell = rd_tell(rd);
  int silence;
  if (tell >= total_bits) silence = 1;
  else if (tell == 1) silence = rd_bit_logp(rd, 15);
  else silence = 0;
  if (silence) {
    tell = len * 8;
    rd.nbits_total += tell - rd_tell(rd);
  }
  i32 postfilter_gain = 0;
  int postfilter_pitch = 0, postfilter_tapset = 0;
  if (start == 0 && tell + 16 <= total_bits) {
    if (rd_bit_logp(rd, 1)) {
      int octave = (int)rd_uint(rd, 6);
      postfilter_pitch = (16 << octave) + (int)rd_bits(rd, 4 + (u32)octave) - 1;
      int qg = (int)rd_bits(rd, 3);
      if (rd_tell(rd) + 2 <= total_bits) postfilter_tapset = rd_icdf(rd, ct_tapset_icdf, 2);
      postfilter_gain = 3072 * (qg + 1);       // QCONST16(.09375f,15)
    }
    tell = rd_tell(rd);
  }
  int isTransient = 0;
  if (LM > 0 && tell + 3 <= total_bits) {
    isTransient = rd_bit_logp(rd, 3);
    tell = rd_tell(rd);
  }
  int shortBlocks = isTransient ? M : 0;
  int intra_ener = tell + 3 <= total_bits ? rd_bit_logp(rd, 3) : 0;
  dec_coarse_energy(w, rd, start, end, intra_ener, C, LM);
  dec_tf(w, rd, start, end, isTransient, LM);
  tell = rd_tell(rd);
  int spread_decision = CB_SPREAD_NORMAL;
  if (tell + 4 <= total_bits) spread_decision = rd_icdf(rd, ct_spread_icdf, 5);
  // init_caps (celt.c:255-264)
  CB_NOUNROLL for (int i = 0; i < CB_NBANDS; i++) {
    int Nb = (ebands(i + 1) - ebands(i)) << LM;
    w.cap[i] = (i16)((ct_cache_caps50[CB_NBANDS * (2 * LM + C - 1) + i] + 64) * C * Nb >> 2);
  }
  int dynalloc_logp = 6;
  total_bits <<= CB_BITRES;
  tell = (i32)rd_tell_frac(rd);
  CB_NOUNROLL for (int i = start; i < end; i++) {
    int width = C * (ebands(i + 1) - ebands(i)) << LM;
    int quanta = cb_imin(width << CB_BITRES, cb_imax(6 << CB_BITRES, width));
    int loop_logp = dynalloc_logp;
    int boost = 0;
    CB_NOUNROLL while (tell + (loop_logp << CB_BITRES) < total_bits && boost < w.cap[i]) {
      int flag = rd_bit_logp(rd, (u32)loop_logp);
      tell = (i32)rd_tell_frac(rd);
      if (!flag) break;
      boost += quanta;
      total_bits -= quanta;
      loop_logp = 1;
    }
    w.offsets[i] = (i16)boost;
    if (boost > 0) dynalloc_logp = cb_imax(2, dynalloc_logp - 1);
  }
  int alloc_trim = tell + (6 << CB_BITRES) <= total_bits ? rd_icdf(rd, ct_trim_icdf, 7) : 5;
  i32 bits = (((i32)len * 8) << CB_BITRES) - (i32)rd_tell_frac(rd) - 1;
  int anti_collapse_rsv = isTransient && LM >= 2 && bits >= ((LM + 2) << CB_BITRES) ? (1 << CB_BITRES) : 0;
  bits -= anti_collapse_rsv;
  int intensity = 0, dual_stereo = 0;
  i32 balance = 0;
  int codedBands = dec_allocation(w, rd, start, end, alloc_trim, intensity, dual_stereo, bits, balance, C, LM);
  dec_fine_energy(w, rd, start, end, C);

  // ---- the band walk, onto the tape ----
  TapeW tw;
  tw.t = tape;
  tw.jobs = tw.seg_jobs = 0;
  tw.nseg = 0;
  tw.seg_band0 = start;
  dec_bands_parse(w, rd, tw, start, end, C, N, shortBlocks, spread_decision, dual_stereo, intensity,
                  len * (8 << CB_BITRES) - anti_collapse_rsv, balance, LM, codedBands);
  int anti_collapse_on = 0;
  if (anti_collapse_rsv > 0) anti_collapse_on = (int)rd_bits(rd, 1);
  dec_energy_finalise(w, rd, start, end, len * 8 - rd_tell(rd), C);
  CB_WSYNC();
  if (silence) {
    CB_LANE_LOOP(i, C * CB_NBANDS) w.bandE[i] = -28672;     // -QCONST16(28.f, DB_SHIFT)
    CB_WSYNC();
  }

  // ---- hand-off: tape header for the rebuild half, frame record for the synthesis kernel ----
  tape->kind = CB_TAPE_FRAME;
  tape->nseg = tw.nseg;
  tape->seed = stg->rng;
  tape->LM = LM; tape->C = C; tape->start = start; tape->end = end; tape->B = shortBlocks ? M : 1;
  tape->spread = spread_decision; tape->intensity = intensity; tape->dual_stereo = dual_stereo;
  tape->anti_collapse_on = anti_collapse_on; tape->silence = silence;
  CB_LANE_LOOP(j, CB_NBANDS) { tape->pulses[j] = (i16)w.pulses[j]; tape->tf_res[j] = w.tf_res[j]; }
  if (anti_collapse_on) {
    CB_LANE_LOOP(j, 2 * CB_NBANDS) {
      tape->bandE[j] = w.bandE[j];
      tape->prevLogE[0][j] = stg->oldLogE[j];
      tape->prevLogE[1][j] = stg->oldLogE2[j];
    }
  }
  int pf_period = cb_imax(stg->pf_period, 15);               // COMBFILTER_MINPERIOD
  int pf_period_old = cb_imax(stg->pf_period_old, 15);
  CB_LANE_LOOP(j, 2 * CB_NBANDS) side->bandE[j] = w.bandE[j];
  side->accum = cfg.accum;
  side->kind = 1; side->LM = LM; side->C = C; side->transient = isTransient; side->silence = silence;
  side->start = start; side->effEnd = effEnd;
  side->T0 = pf_period_old; side->T1 = pf_period; side->T2 = postfilter_pitch;
  side->g0 = stg->pf_gain_old; side->g1 = stg->pf_gain; side->g2 = postfilter_gain;
  side->tap0 = stg->pf_tapset_old; side->tap1 = stg->pf_tapset; side->tap2 = postfilter_tapset;

  // ---- small-state update (celt_decoder.c:1014-1062) ----
  int npo = pf_period, ngo = stg->pf_gain, nto = stg->pf_tapset;          // *_old <- current
  int np = postfilter_pitch, ng = postfilter_gain, nt = postfilter_tapset;
  if (LM != 0) { npo = np; ngo = ng; nto = nt; }
  CB_WSYNC();
  if (C == 1) {
    CB_LANE_LOOP(i, CB_NBANDS) w.bandE[CB_NBANDS + i] = w.bandE[i];
    CB_WSYNC();
  }
  const int loss_count = stg->loss_count;
  CB_LANE_LOOP(i, 2 * CB_NBANDS) {
    i16 e = w.bandE[i];
    i16 l1 = stg->oldLogE[i], l2 = stg->oldLogE2[i], bg = stg->backgroundLogE[i];
    if (!isTransient) {
      l2 = l1;
      l1 = e;
      i32 inc = loss_count < 10 ? M * 1 : 1024;   // M*QCONST16(0.001f,DB_SHIFT) : QCONST16(1.f,DB_SHIFT)
      bg = (i16)cb_imin(bg + inc, e);
    } else {
      l1 = (i16)cb_imin(l1, e);
    }
    int band = i % CB_NBANDS;
    if (band < start || band >= end) {
      e = 0;
      l1 = l2 = -28672;
    }
    stg->oldBandE[i] = e;
    stg->oldLogE[i] = l1;
    stg->oldLogE2[i] = l2;
    stg->backgroundLogE[i] = bg;
  }
  CB_WSYNC();
  stg->pf_period = np; stg->pf_gain = ng; stg->pf_tapset = nt;
  stg->pf_period_old = npo; stg->pf_gain_old = ngo; stg->pf_tapset_old = nto;
  stg->rng = rd.rng;
  stg->loss_count = 0;
  if (rd.error) stg->error = 1;
  *final_range = rd.rng;
  if (rd_tell(rd) > 8 * len) return -3;        // OPUS_INTERNAL_ERROR
  return 0;
}

#ifndef CB_THREAD_MODE
// REBUILD half of a frame (one warp): tape -> normalised spectrum in side->X.
CB_FN void celt_frame_rebuild(EntWork& w, FrameTape* tape, FrameSide* side) {
  const int LM = tape->LM, C = tape->C, start = tape->start, end = tape->end;
  const int N = 120 << LM;
  CB_WSYNC();
  CB_IF_LANE0 w.norm = tape->norm;
  CB_LANE_LOOP(j, CB_NBANDS) { w.pulses[j] = tape->pulses[j]; w.tf_res[j] = tape->tf_res[j]; }
  CB_WSYNC();
  u32 seed = tape->seed;
  dec_bands_rebuild(w, tape, start, end, C, N, tape->B, LM, seed);
  if (tape->anti_collapse_on) {
    i16* e = (i16*)w.tmp32;                    // [0,42) bandE, [42,84) oldLogE, [84,126) oldLogE2, before this frame
    CB_WSYNC();
    CB_LANE_LOOP(j, 2 * CB_NBANDS) {
      e[j] = tape->bandE[j];
      e[2 * CB_NBANDS + j] = tape->prevLogE[0][j];
      e[4 * CB_NBANDS + j] = tape->prevLogE[1][j];
    }
    CB_WSYNC();
    v_anti_collapse(w, LM, C, N, start, end, e, e + 2 * CB_NBANDS, e + 4 * CB_NBANDS, seed);
  }
  CB_WSYNC();
  {
    // X as 32-bit words, coalesced
    const u32* src = (const u32*)w.X;
    u32* dst = (u32*)side->X;
    CB_LANE_LOOP(j, C * N / 2) dst[j] = src[j];
  }
}

// Both halves on one warp (the general kernel and the host build; the batch kernels run them apart).
CB_FN int celt_entropy_decode(EntWork& w, CeltDecState* stg, FrameSide* side, FrameTape* tape, const u8* pkt, int len,
                              const EntFrameCfg& cfg, u32* final_range) {
  const int rc = celt_frame_parse(w, stg, side, tape, pkt, len, cfg, final_range);
  CB_WSYNC();
  celt_frame_rebuild(w, tape, side);
  CB_WSYNC();
  return rc;
}
#endif

// Noise-based concealment / comfort noise (celt_decoder.c:451-506), in the two halves of the entropy stage: the
// PARSE half (no width: one thread or one warp) decays the band energies, advances the noise generator and the
// loss counter -- the small state -- and fills the frame record; the REBUILD half (one warp) draws the random
// normalised bands from the seed the tape carries.  The synthesis kernel then treats it like a decoded frame.
CB_FN void celt_plc_noise_parse(CeltDecState* stg, FrameSide* side, FrameTape* tape, int LM, int start, int end) {
  const int C = stg->channels;
  const int effEnd = cb_imax(start, cb_imin(end, CB_NBANDS));
  const int loss_count = stg->loss_count;
  const i32 decay = loss_count == 0 ? 1536 : 512;          // QCONST16(1.5f,DB_SHIFT) : QCONST16(.5f,DB_SHIFT)
  CB_WSYNC();
  CB_LANE_LOOP(j, 2 * CB_NBANDS) {
    i16 e = stg->oldBandE[j];
    int band = j % CB_NBANDS, c = j / CB_NBANDS;
    if (c < C && band >= start && band < end) {
      e = (i16)cb_imax(stg->backgroundLogE[j], e - decay);
      stg->oldBandE[j] = e;
    }
    side->bandE[j] = e;
  }
  const u32 seed = stg->rng;
  u32 total = 0;                                            // samples the generator is advanced by
  CB_NOUNROLL for (int i = start; i < effEnd; i++) total += (u32)((ct_eband5ms[i + 1] - ct_eband5ms[i]) << LM);
  tape->kind = CB_TAPE_NOISE;
  tape->seed = seed;
  tape->LM = LM; tape->C = C; tape->start = start; tape->end = effEnd;
  side->kind = 1; side->LM = LM; side->C = C; side->transient = 0; side->silence = 0;
  side->start = start; side->effEnd = effEnd;
  side->T0 = 15; side->T1 = 15; side->T2 = 15;
  side->g0 = 0; side->g1 = 0; side->g2 = 0;
  side->tap0 = 0; side->tap1 = 0; side->tap2 = 0;
  side->accum = 0;
  stg->rng = cb_lcg_skip(seed, total * (u32)C);
  stg->loss_count = loss_count + 1;
}
#ifndef CB_THREAD_MODE
CB_FN void celt_plc_noise_rebuild(EntWork& w, const FrameTape* tape, FrameSide* side) {
  const int LM = tape->LM, C = tape->C, start = tape->start, effEnd = tape->end;
  const int N = 120 << LM;
  u32 seed = tape->seed;
  CB_NOUNROLL for (int c = 0; c < C; c++) {
    CB_NOUNROLL for (int i = start; i < effEnd; i++) {
      const int boffs = N * c + (ct_eband5ms[i] << LM);
      const int blen = (ct_eband5ms[i + 1] - ct_eband5ms[i]) << LM;
      const u32 seed0 = seed;
      CB_WSYNC();
      CB_LANE_LOOP(j, blen) w.X[boffs + j] = (i16)((i32)cb_lcg_skip(seed0, (u32)j + 1) >> 20);
      seed = cb_lcg_skip(seed0, (u32)blen);
      v_renormalise(w.X + boffs, blen, CB_Q15ONE);
    }
  }
  CB_WSYNC();
  {
    const u32* src = (const u32*)w.X;
    u32* dst = (u32*)side->X;
    CB_LANE_LOOP(j, C * N / 2) dst[j] = src[j];
  }
}
// both halves on one warp
CB_FN void celt_plc_noise(EntWork& w, CeltDecState* stg, FrameSide* side, FrameTape* tape, int LM, int start, int end) {
  celt_plc_noise_parse(stg, side, tape, LM, start, end);
  CB_WSYNC();
  celt_plc_noise_rebuild(w, tape, side);
  CB_WSYNC();
}
#endif
