/* oracle_celt.c -- ORACLE (test infrastructure, never shipped, never on a product path).
 *
 * Plain-C, single-threaded, scalar restatement of the reference's CELT decoder
 * (libopus 1.1.2 FIXED_POINT as found under /root/reference/opus-fix, the C that Concentus was
 * ported from).  Written from the algorithm, in our own structure: one flat decoder struct with a
 * linear 2048+120 history like the reference, recursive band splitting, no SIMD, no shared code
 * with the CUDA sources.  Every function names the reference file:line it follows.
 *
 * Parity status: PINNED -- tests/test_oracle_restatement.py holds this file to the committed
 * golden fixtures (reference encoder+decoder output) and, when oracle/_ref is present, to the
 * compiled reference itself, sample by sample and final range by final range.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "oracle_celt.h"
#include "oracle_tables.h"

typedef int16_t s16;
typedef int32_t s32;
typedef uint32_t u32;
typedef int64_t s64;
typedef uint8_t u8;

/* ---------- fixed-point primitives: celt/fixed_generic.h, celt/arch.h ---------- */
static s32 ADD(s32 a, s32 b) { return (s32)((u32)a + (u32)b); }                       /* ADD32 :101 */
static s32 SUB(s32 a, s32 b) { return (s32)((u32)a - (u32)b); }                       /* SUB32 :103 */
static s32 SHL(s32 a, int s) { return (s32)((u32)a << s); }                           /* SHL32 :74 */
static s32 PSHR(s32 a, int s) { return ADD(a, (1 << s) >> 1) >> s; }                  /* PSHR32 :77 */
static s32 VSHR(s32 a, int s) { return s > 0 ? a >> s : SHL(a, -s); }                 /* VSHR32 :79 */
static s32 M16(s32 a, s32 b) { return (s32)(s16)a * (s32)(s16)b; }                    /* MULT16_16 :109 */
static s32 MQ15(s32 a, s32 b) { return M16(a, b) >> 15; }                             /* MULT16_16_Q15 :126 */
static s32 MP15(s32 a, s32 b) { return ADD(16384, M16(a, b)) >> 15; }                 /* MULT16_16_P15 :130 */
static s32 M32Q15(s32 a, s32 b) { return (s32)(((s64)(s16)a * b) >> 15); }            /* MULT16_32_Q15 :46 */
static s32 M32P16(s32 a, s32 b) { return (s32)(((s64)(s16)a * b + 32768) >> 16); }    /* MULT16_32_P16 :43 */
static s32 M3232Q31(s32 a, s32 b) {                                                   /* MULT32_32_Q31 :49 */
   s32 ah = a >> 16, bh = b >> 16, al = a & 0xffff, bl = b & 0xffff;
   return ADD(ADD(SHL(M16(ah, bh), 1), ((s32)(s16)ah * bl) >> 15), ((s32)(s16)bh * al) >> 15);
}
static s16 A16(s32 a, s32 b) { return (s16)((s16)a + (s16)b); }                       /* ADD16 :96 */
static s32 S16(s32 a, s32 b) { return (s32)(s16)a - (s32)(s16)b; }                    /* SUB16 :98 */
static s16 SL16(s32 a, int s) { return (s16)((uint16_t)a << s); }                     /* SHL16 :70 */
static s16 R16(s32 a, int s) { return (s16)PSHR(a, s); }                              /* ROUND16 :90 */
static s32 FM16(s32 a, s32 b) { return (16384 + (s32)(s16)a * (s32)(s16)b) >> 15; }   /* FRAC_MUL16 mathops.h:42 */
static s32 imin(s32 a, s32 b) { return a < b ? a : b; }
static s32 imax(s32 a, s32 b) { return a > b ? a : b; }
static int ilog(u32 v) { int n = 0; while (v) { n++; v >>= 1; } return n; }           /* EC_ILOG ecintrin.h:69 */
static int ilog2(s32 v) { return ilog((u32)v) - 1; }                                   /* celt_ilog2 mathops.h:148 */
static s16 sat16(s32 x) { return (s16)(x > 32767 ? 32767 : x < -32768 ? -32768 : x); }
static s16 sig2word(s32 x) { x = PSHR(x, 12); return (s16)imax(-32768, imin(32767, x)); } /* SIG2WORD16 :142 */

/* ---------- celt/mathops.c, mathops.h ---------- */
static u32 isqrt32(u32 v) {                                                            /* mathops.c:42-63 */
   u32 g = 0; int bs = (ilog(v) - 1) >> 1; u32 b = 1u << bs;
   do { u32 t = ((g << 1) + b) << bs; if (t <= v) { g += b; v -= t; } b >>= 1; bs--; } while (bs >= 0);
   return g;
}
static s16 rsqrt_norm(s32 x) {                                                         /* mathops.c:90-115 */
   s16 n = (s16)(x - 32768);
   s16 r = A16(23557, MQ15(n, A16(-13490, MQ15(n, 6713))));
   s16 r2 = (s16)MQ15(r, r);
   s16 y = SL16(S16(A16(MQ15(r2, n), r2), 16384), 1);
   return A16(r, MQ15(r, MQ15(y, S16(MQ15(y, 12288), 16384))));
}
static s32 celt_sqrt(s32 x) {                                                          /* mathops.c:118-137 */
   int k; s16 n; s32 rt;
   if (x == 0) return 0;
   if (x >= 1073741824) return 32767;
   k = (ilog2(x) >> 1) - 7;
   x = VSHR(x, 2 * k);
   n = (s16)(x - 32768);
   rt = A16(23175, MQ15(n, A16(11561, MQ15(n, A16(-3011, MQ15(n, A16(1699, MQ15(n, -664))))))));
   return VSHR(rt, 7 - k);
}
static s16 cos_pi_2(s16 x) {                                                           /* mathops.c:144-151 */
   s16 x2 = (s16)MP15(x, x);
   s32 t = ADD(S16(32767, x2), MP15(x2, ADD(-7651, MP15(x2, ADD(8277, MP15(-626, x2))))));
   return A16(1, imin(32766, t));
}
static s16 cos_norm(s32 x) {                                                           /* mathops.c:158-178 */
   x &= 0x1ffff;
   if (x > 65536) x = 131072 - x;
   if (x & 0x7fff) return x < 32768 ? cos_pi_2((s16)x) : (s16)(-(s32)cos_pi_2((s16)(65536 - x)));
   if (x & 0xffff) return 0;
   return (x & 0x1ffff) ? -32767 : 32767;
}
static s32 celt_rcp(s32 x) {                                                           /* mathops.c:180-206 */
   int i = ilog2(x);
   s16 n = (s16)(VSHR(x, i - 15) - 32768);
   s16 r = A16(30840, MQ15(-15420, n));
   r = (s16)S16(r, MQ15(r, A16(MQ15(r, n), A16(r, -32768))));
   r = (s16)S16(r, A16(1, MQ15(r, A16(MQ15(r, n), A16(r, -32768)))));
   return VSHR(r, i - 16);
}
static s32 frac_div32(s32 a, s32 b) {                                                  /* mathops.c:69-87 */
   int shift = ilog2(b) - 29; s16 rcp; s32 result, rem;
   a = VSHR(a, shift); b = VSHR(b, shift);
   rcp = R16(celt_rcp(R16(b, 16)), 3);
   result = M32Q15(rcp, a);
   rem = SUB(PSHR(a, 2), M3232Q31(result, b));
   result = ADD(result, SHL(M32Q15(rcp, rem), 2));
   if (result >= 536870912) return 2147483647;
   if (result <= -536870912) return -2147483647;
   return SHL(result, 2);
}
static s32 exp2_frac(s32 x) {                                                          /* mathops.h:195-200 */
   s16 f = SL16(x, 4);
   return A16(16383, MQ15(f, A16(22804, MQ15(f, A16(14819, MQ15(10204, f))))));
}
static s32 celt_exp2(s32 xin) {                                                        /* mathops.h:202-213 */
   s16 x = (s16)xin; int integer = x >> 10;
   if (integer > 14) return 0x7f000000;
   if (integer < -15) return 0;
   return VSHR(exp2_frac((s16)(x - SL16(integer, 10))), -integer - 2);
}
static u32 lcg(u32 s) { return 1664525u * s + 1013904223u; }                           /* bands.c:63 */

/* ---------- range decoder: celt/entdec.c, entcode.c, laplace.c ---------- */
typedef struct {
   const u8 *buf; u32 storage, end_offs, end_window; int nend_bits, nbits_total;
   u32 offs, rng, val, ext; int rem, error;
} RDec;
static int rd_byte(RDec *d) { return d->offs < d->storage ? d->buf[d->offs++] : 0; }              /* entdec.c:93 */
static int rd_byte_end(RDec *d) { return d->end_offs < d->storage ? d->buf[d->storage - ++d->end_offs] : 0; }
static void rd_norm(RDec *d) {                                                                      /* :104-131 */
   while (d->rng <= 0x800000u) {
      int sym;
      d->nbits_total += 8; d->rng <<= 8;
      sym = d->rem; d->rem = rd_byte(d);
      sym = ((sym << 8) | d->rem) >> 1;
      d->val = ((d->val << 8) + (255u & ~(u32)sym)) & 0x7fffffffu;
   }
}
static void rd_init(RDec *d, const u8 *buf, u32 n) {                                                /* :133-153 */
   memset(d, 0, sizeof(*d));
   d->buf = buf; d->storage = n; d->nbits_total = 9; d->rng = 128;
   d->rem = rd_byte(d); d->val = d->rng - 1 - (u32)(d->rem >> 1);
   rd_norm(d);
}
static int rd_tell(const RDec *d) { return d->nbits_total - ilog(d->rng); }                        /* entcode.h:114 */
static u32 rd_tell_frac(const RDec *d) {                                                            /* entcode.c:69-97 */
   static const u32 corr[8] = {35733, 38967, 42495, 46340, 50535, 55109, 60097, 65535};
   int l = ilog(d->rng); u32 r = d->rng >> (l - 16); u32 b = (r >> 12) - 8;
   b += r > corr[b];
   return ((u32)d->nbits_total << 3) - (u32)((l << 3) + (int)b);
}
static u32 rd_decode(RDec *d, u32 ft) {                                                             /* entdec.c:155 */
   u32 s; d->ext = d->rng / ft; s = d->val / d->ext; return ft - (s + 1 < ft ? s + 1 : ft);
}
static u32 rd_decode_bin(RDec *d, u32 bits) {                                                       /* :174 */
   u32 s, top = 1u << bits; d->ext = d->rng >> bits; s = d->val / d->ext; return top - (s + 1 < top ? s + 1 : top);
}
static void rd_update(RDec *d, u32 fl, u32 fh, u32 ft) {                                            /* :181 */
   u32 s = d->ext * (ft - fh);
   d->val -= s; d->rng = fl > 0 ? d->ext * (fh - fl) : d->rng - s; rd_norm(d);
}
static int rd_bit_logp(RDec *d, u32 logp) {                                                         /* :203 */
   u32 r = d->rng, v = d->val, s = r >> logp; int ret = v < s;
   if (!ret) d->val = v - s;
   d->rng = ret ? s : r - s; rd_norm(d); return ret;
}
static int rd_icdf(RDec *d, const u8 *icdf, u32 ftb) {                                              /* :218 */
   u32 s = d->rng, v = d->val, r = s >> ftb, t; int ret = -1;
   do { t = s; s = r * icdf[++ret]; } while (v < s);
   d->val = v - s; d->rng = t - s; rd_norm(d); return ret;
}
static u32 rd_bits(RDec *d, u32 bits) {                                                             /* :284 */
   u32 w = d->end_window, ret; int av = d->nend_bits;
   if ((u32)av < bits) do { w |= (u32)rd_byte_end(d) << av; av += 8; } while (av <= 24);
   ret = w & ((1u << bits) - 1u);
   d->end_window = w >> bits; d->nend_bits = av - (int)bits; d->nbits_total += (int)bits;
   return ret;
}
static u32 rd_uint(RDec *d, u32 ft) {                                                               /* :238 */
   int ftb; u32 s;
   ft--; ftb = ilog(ft);
   if (ftb > 8) {
      u32 top, t;
      ftb -= 8; top = (ft >> ftb) + 1;
      s = rd_decode(d, top); rd_update(d, s, s + 1, top);
      t = (s << ftb) | rd_bits(d, (u32)ftb);
      if (t <= ft) return t;
      d->error = 1; return ft;
   }
   ft++; s = rd_decode(d, ft); rd_update(d, s, s + 1, ft); return s;
}
static int rd_laplace(RDec *d, u32 fs, int decay) {                                                 /* laplace.c:94 */
   int val = 0; u32 fl = 0, fm = rd_decode_bin(d, 15);
   if (fm >= fs) {
      val++; fl = fs;
      fs = (((32768u - 32u - fs) * (u32)(16384 - decay)) >> 15) + 1;                                /* :44-49 */
      while (fs > 1 && fm >= fl + 2 * fs) { fs *= 2; fl += fs; fs = (((fs - 2) * (u32)decay) >> 15) + 1; val++; }
      if (fs <= 1) { int di = (int)((fm - fl) >> 1); val += di; fl += 2 * (u32)di; }
      if (fm < fl + fs) val = -val; else fl += fs;
   }
   rd_update(d, fl, imin((s32)(fl + fs), 32768), 32768);
   return val;
}

/* ---------- decoder state: celt/celt_decoder.c:67-100 ---------- */
#define NB 21
#define OVL 120
#define DBUF 2048
#define LPC_ORD 24
struct OracleCelt {
   int channels, downsample, start, end;
   u32 rng; int error, last_pitch_index, loss_count;
   int pf_period, pf_period_old, pf_tapset, pf_tapset_old; s16 pf_gain, pf_gain_old;
   s32 preemph_mem[2];
   s32 mem[2][DBUF + OVL];
   s16 lpc[2][LPC_ORD];
   s16 oldE[2 * NB], oldLogE[2 * NB], oldLogE2[2 * NB], bgLogE[2 * NB];
};
static int EB(int i) { return ot_eband5ms[i]; }

void oracle_celt_reset(struct OracleCelt *st) {                                       /* celt_decoder.c:1196-1204 */
   int ch = st->channels, ds = st->downsample, i;
   memset(st, 0, sizeof(*st));
   st->channels = ch; st->downsample = ds; st->start = 0; st->end = NB;
   for (i = 0; i < 2 * NB; i++) st->oldLogE[i] = st->oldLogE2[i] = -28672;
}

/* ---------- rate.h / rate.c ---------- */
static int get_pulses(int i) { return i < 8 ? i : (8 + (i & 7)) << ((i >> 3) - 1); }               /* rate.h:48 */
static const u8 *pcache(int band, int LM) { return ot_cache_bits50 + ot_cache_index50[(LM + 1) * NB + band]; }
static int bits2pulses(int band, int LM, int bits) {                                                /* rate.h:53-76 */
   const u8 *c = pcache(band, LM); int lo = 0, hi = c[0], i;
   bits--;
   for (i = 0; i < 6; i++) { int mid = (lo + hi + 1) >> 1; if ((int)c[mid] >= bits) hi = mid; else lo = mid; }
   return (bits - (lo == 0 ? -1 : (int)c[lo]) <= (int)c[hi] - bits) ? lo : hi;
}
static int pulses2bits(int band, int LM, int p) { return p == 0 ? 0 : pcache(band, LM)[p] + 1; }   /* rate.h:78 */

/* compute_allocation + interp_bits2pulses, decoder side (rate.c:248-638) */
static int allocation(RDec *d, int start, int end, const int *offsets, const int *cap, int trim, int *intensity,
                      int *dual, s32 total, s32 *balance_out, int *bits, int *ebits, int *prio, int C, int LM) {
   static const u8 log2frac[24] = {0, 8, 13, 16, 19, 21, 23, 24, 26, 27, 28, 29, 30, 31, 32, 32, 33, 34, 34, 35, 36, 36, 37, 37};
   int bits1[NB], bits2[NB], thresh[NB], trimoff[NB];
   int lo, hi, j, skip_start = start, skip_rsv, irsv = 0, drsv = 0, coded, floor_ = C << 3, stereo = C > 1, logM = LM << 3;
   s32 psum, left, percoeff, balance;
   total = imax(total, 0);
   skip_rsv = total >= 8 ? 8 : 0; total -= skip_rsv;
   if (C == 2) {
      irsv = log2frac[end - start];
      if (irsv > total) irsv = 0;
      else { total -= irsv; drsv = total >= 8 ? 8 : 0; total -= drsv; }
   }
   for (j = start; j < end; j++) {
      int w = EB(j + 1) - EB(j);
      thresh[j] = imax(C << 3, (3 * w << LM << 3) >> 4);
      trimoff[j] = C * w * (trim - 5 - LM) * (end - j - 1) * (1 << (LM + 3)) >> 6;
      if (w << LM == 1) trimoff[j] -= C << 3;
   }
   lo = 1; hi = 10;
   do {
      int done = 0, mid = (lo + hi) >> 1;
      psum = 0;
      for (j = end; j-- > start;) {
         int b = C * (EB(j + 1) - EB(j)) * ot_band_alloc[mid * NB + j] << LM >> 2;
         if (b > 0) b = imax(0, b + trimoff[j]);
         b += offsets[j];
         if (b >= thresh[j] || done) { done = 1; psum += imin(b, cap[j]); }
         else if (b >= C << 3) psum += C << 3;
      }
      if (psum > total) hi = mid - 1; else lo = mid + 1;
   } while (lo <= hi);
   hi = lo--;
   for (j = start; j < end; j++) {
      int w = EB(j + 1) - EB(j);
      int b1 = C * w * ot_band_alloc[lo * NB + j] << LM >> 2;
      int b2 = hi >= 11 ? cap[j] : C * w * ot_band_alloc[hi * NB + j] << LM >> 2;
      if (b1 > 0) b1 = imax(0, b1 + trimoff[j]);
      if (b2 > 0) b2 = imax(0, b2 + trimoff[j]);
      if (lo > 0) b1 += offsets[j];
      b2 += offsets[j];
      if (offsets[j] > 0) skip_start = j;
      bits1[j] = b1; bits2[j] = imax(0, b2 - b1);
   }
   lo = 0; hi = 64;                                                                                 /* rate.c:269-292 */
   for (j = 0; j < 6; j++) {
      int mid = (lo + hi) >> 1, done = 0, k;
      psum = 0;
      for (k = end; k-- > start;) {
         int t = bits1[k] + (int)((s32)mid * bits2[k] >> 6);
         if (t >= thresh[k] || done) { done = 1; psum += imin(t, cap[k]); }
         else if (t >= floor_) psum += floor_;
      }
      if (psum > total) hi = mid; else lo = mid;
   }
   psum = 0;
   { int done = 0;
     for (j = end; j-- > start;) {
        int t = bits1[j] + (lo * bits2[j] >> 6);
        if (t < thresh[j] && !done) t = t >= floor_ ? floor_ : 0; else done = 1;
        t = imin(t, cap[j]); bits[j] = t; psum += t;
     } }
   for (coded = end;; coded--) {                                                                    /* :312-380 */
      int bw, bb, rem, span;
      j = coded - 1;
      if (j <= skip_start) { total += skip_rsv; break; }
      left = total - psum; span = EB(coded) - EB(start);
      percoeff = (s32)((u32)left / (u32)span); left -= span * percoeff;
      rem = imax(left - (EB(j) - EB(start)), 0);
      bw = EB(coded) - EB(j);
      bb = (int)(bits[j] + percoeff * bw + rem);
      if (bb >= imax(thresh[j], floor_ + 8)) {
         if (rd_bit_logp(d, 1)) break;
         psum += 8; bb -= 8;
      }
      psum -= bits[j] + irsv;
      if (irsv > 0) irsv = log2frac[j - start];
      psum += irsv;
      if (bb >= floor_) { psum += floor_; bits[j] = floor_; } else bits[j] = 0;
   }
   *intensity = irsv > 0 ? start + (int)rd_uint(d, (u32)(coded + 1 - start)) : 0;
   if (*intensity <= start) { total += drsv; drsv = 0; }
   *dual = drsv > 0 ? rd_bit_logp(d, 1) : 0;
   left = total - psum;
   { int span = EB(coded) - EB(start);
     percoeff = (s32)((u32)left / (u32)span); left -= span * percoeff; }
   for (j = start; j < coded; j++) bits[j] += (int)percoeff * (EB(j + 1) - EB(j));
   for (j = start; j < coded; j++) { int t = imin(left, EB(j + 1) - EB(j)); bits[j] += t; left -= t; }
   balance = 0;
   for (j = start; j < coded; j++) {                                                                /* :432-509 */
      int N0 = EB(j + 1) - EB(j), N = N0 << LM;
      s32 bit = bits[j] + balance, excess;
      if (N > 1) {
         int den, NClogN, offset;
         excess = imax(bit - cap[j], 0); bits[j] = bit - excess;
         den = C * N + ((C == 2 && N > 2 && !*dual && j < *intensity) ? 1 : 0);
         NClogN = den * (ot_logN400[j] + logM);
         offset = (NClogN >> 1) - den * 21;
         if (N == 2) offset += den << 3 >> 2;
         if (bits[j] + offset < den * 2 << 3) offset += NClogN >> 2;
         else if (bits[j] + offset < den * 3 << 3) offset += NClogN >> 3;
         ebits[j] = imax(0, bits[j] + offset + (den << 2));
         ebits[j] = (int)((u32)ebits[j] / (u32)den) >> 3;
         if (C * ebits[j] > (bits[j] >> 3)) ebits[j] = bits[j] >> stereo >> 3;
         ebits[j] = imin(ebits[j], 8);
         prio[j] = ebits[j] * (den << 3) >= bits[j] + offset;
         bits[j] -= C * ebits[j] << 3;
      } else {
         excess = imax(0, bit - (C << 3)); bits[j] = bit - excess; ebits[j] = 0; prio[j] = 1;
      }
      if (excess > 0) {
         int xf = imin(excess >> (stereo + 3), 8 - ebits[j]), xb;
         ebits[j] += xf; xb = xf * C << 3;
         prio[j] = xb >= excess - balance;
         excess -= xb;
      }
      balance = excess;
   }
   *balance_out = balance;
   for (; j < end; j++) { ebits[j] = bits[j] >> stereo >> 3; bits[j] = 0; prio[j] = ebits[j] < 1; }
   return coded;
}

/* ---------- vector helpers: celt/vq.c, celt/bands.c ---------- */
static void renormalise(s16 *X, int N, s32 gain) {                                                  /* vq.c:349-374 */
   s32 E = 1, t; int i, k; s16 g;
   for (i = 0; i < N; i++) E = ADD(E, M16(X[i], X[i]));
   k = ilog2(E) >> 1; t = VSHR(E, 2 * (k - 7));
   g = (s16)MP15(rsqrt_norm(t), gain);
   for (i = 0; i < N; i++) X[i] = (s16)PSHR(M16(g, X[i]), k + 1);
}
static void rotation1(s16 *X, int len, int stride, s32 c, s32 s) {                                  /* vq.c:43-67 */
   int i; s32 ms = (s16)(-s); s16 *p = X;
   for (i = 0; i < len - stride; i++) {
      s32 x1 = p[0], x2 = p[stride];
      p[stride] = (s16)PSHR(ADD(M16(c, x2), M16(s, x1)), 15);
      *p++ = (s16)PSHR(ADD(M16(c, x1), M16(ms, x2)), 15);
   }
   p = &X[len - 2 * stride - 1];
   for (i = len - 2 * stride - 1; i >= 0; i--) {
      s32 x1 = p[0], x2 = p[stride];
      p[stride] = (s16)PSHR(ADD(M16(c, x2), M16(s, x1)), 15);
      *p-- = (s16)PSHR(ADD(M16(c, x1), M16(ms, x2)), 15);
   }
}
static void exp_rotation_dec(s16 *X, int len, int stride, int K, int spread) {                      /* vq.c:70-115, dir=-1 */
   static const int factor[3] = {15, 10, 5};
   s16 gain, theta, c, s; int stride2 = 0, i;
   if (2 * K >= len || spread == 0) return;
   gain = (s16)M3232Q31(M16(32767, len), celt_rcp(len + factor[spread - 1] * K));
   theta = (s16)(MQ15(gain, gain) >> 1);
   c = cos_norm(theta); s = cos_norm(S16(32767, theta));
   if (len >= 8 * stride) { stride2 = 1; while ((stride2 * stride2 + stride2) * stride + (stride >> 2) < len) stride2++; }
   len /= stride;
   for (i = 0; i < stride; i++) {
      if (stride2) rotation1(X + i * len, len, stride2, s, c);
      rotation1(X + i * len, len, 1, c, s);
   }
}
static u32 pvq_u(int n, int k) { int lo = n < k ? n : k, hi = n < k ? k : n; return ot_pvq_u_data[ot_pvq_u_row[lo] + hi]; }
static s32 cwrsi(int n, int k, u32 i, int *y) {                                                     /* cwrs.c:463-537 */
   s32 yy = 0; u32 p, q; int s, k0; s16 val;
   while (n > 2) {
      if (k >= n) {
         p = pvq_u(n, k + 1); s = -(int)(i >= p); i -= p & (u32)s; k0 = k; q = pvq_u(n, n);
         if (q > i) { k = n; do p = pvq_u(--k, n); while (p > i); }
         else for (p = pvq_u(n, k); p > i; p = pvq_u(n, k)) k--;
         i -= p; val = (s16)((k0 - k + s) ^ s); *y++ = val; yy = ADD(yy, M16(val, val));
      } else {
         p = pvq_u(k, n); q = pvq_u(k + 1, n);
         if (p <= i && i < q) { i -= p; *y++ = 0; }
         else {
            s = -(int)(i >= q); i -= q & (u32)s; k0 = k;
            do p = pvq_u(--k, n); while (p > i);
            i -= p; val = (s16)((k0 - k + s) ^ s); *y++ = val; yy = ADD(yy, M16(val, val));
         }
      }
      n--;
   }
   p = 2 * (u32)k + 1; s = -(int)(i >= p); i -= p & (u32)s; k0 = k; k = (int)((i + 1) >> 1);
   if (k) i -= 2 * (u32)k - 1;
   val = (s16)((k0 - k + s) ^ s); *y++ = val; yy = ADD(yy, M16(val, val));
   s = -(int)i; val = (s16)((k + s) ^ s); *y = val; yy = ADD(yy, M16(val, val));
   return yy;
}
static unsigned alg_unquant(RDec *d, s16 *X, int N, int K, int spread, int B, s32 gain) {           /* vq.c:329-347 */
   int iy[176], i, k; s32 Ryy, t; s16 g; unsigned mask = 0;
   Ryy = cwrsi(N, K, rd_uint(d, pvq_u(N, K) + pvq_u(N, K + 1)), iy);
   k = ilog2(Ryy) >> 1; t = VSHR(Ryy, 2 * (k - 7));
   g = (s16)MP15(rsqrt_norm(t), gain);
   for (i = 0; i < N; i++) X[i] = (s16)PSHR(M16(g, iy[i]), k + 1);
   exp_rotation_dec(X, N, B, K, spread);
   if (B <= 1) return 1;
   { int N0 = N / B, b, j;                                                                          /* vq.c:139-159 */
     for (b = 0; b < B; b++) { unsigned any = 0; for (j = 0; j < N0; j++) any |= (unsigned)iy[b * N0 + j]; mask |= (unsigned)(any != 0) << b; } }
   return mask;
}
static void haar1(s16 *X, int N0, int stride) {                                                     /* bands.c:581-594 */
   int i, j;
   N0 >>= 1;
   for (i = 0; i < stride; i++) for (j = 0; j < N0; j++) {
      s32 a = M16(23170, X[stride * 2 * j + i]), b = M16(23170, X[stride * (2 * j + 1) + i]);
      X[stride * 2 * j + i] = (s16)PSHR(ADD(a, b), 15);
      X[stride * (2 * j + 1) + i] = (s16)PSHR(SUB(a, b), 15);
   }
}
static const u8 ordery[30] = {1, 0, 3, 0, 2, 1, 7, 0, 4, 3, 6, 1, 5, 2, 15, 0, 8, 7, 12, 3, 11, 4, 14, 1, 9, 6, 13, 2, 10, 5};
static void hadamard(s16 *X, int N0, int stride, int had, int inverse) {                            /* bands.c:532-579 */
   s16 tmp[176]; int i, j, N = N0 * stride;
   for (i = 0; i < stride; i++) for (j = 0; j < N0; j++) {
      int row = had ? ordery[stride - 2 + i] : i;
      if (inverse) tmp[j * stride + i] = X[row * N0 + j];      /* interleave_hadamard */
      else tmp[row * N0 + j] = X[j * stride + i];              /* deinterleave_hadamard */
   }
   memcpy(X, tmp, N * sizeof(s16));
}
static void stereo_merge(s16 *X, s16 *Y, s32 mid, int N) {                                          /* bands.c:375-425 */
   s32 xp = 0, side = 0, El, Er, m2, lg, rg; int j, kl, kr; s16 mid2;
   for (j = 0; j < N; j++) { xp = ADD(xp, M16(Y[j], X[j])); side = ADD(side, M16(Y[j], Y[j])); }
   xp = M32Q15(mid, xp); mid2 = (s16)(mid >> 1); m2 = M16(mid2, mid2);
   El = SUB(ADD(m2, side), SHL(xp, 1)); Er = ADD(ADD(m2, side), SHL(xp, 1));
   if (Er < 161061 || El < 161061) { memcpy(Y, X, N * sizeof(s16)); return; }
   kl = ilog2(El) >> 1; kr = ilog2(Er) >> 1;
   lg = rsqrt_norm(VSHR(El, (kl - 7) << 1)); rg = rsqrt_norm(VSHR(Er, (kr - 7) << 1));
   if (kl < 7) kl = 7;
   if (kr < 7) kr = 7;
   for (j = 0; j < N; j++) {
      s16 l = (s16)MP15(mid, X[j]), r = Y[j];
      X[j] = (s16)PSHR(M16(lg, S16(l, r)), kl + 1);
      Y[j] = (s16)PSHR(M16(rg, A16(l, r)), kr + 1);
   }
}

/* ---------- band decoding: celt/bands.c:596-1502 ---------- */
typedef struct { RDec *d; int i, intensity, spread, tf_change; s32 remaining; u32 seed; } BCtx;
typedef struct { int inv, imid, iside, delta, itheta, qalloc; } SplitCtx;

static s16 bitexact_cos(s16 x) {                                                                    /* bands.c:70-80 */
   s16 x2 = (s16)((4096 + (s32)x * x) >> 13);
   x2 = (s16)((32767 - x2) + FM16(x2, (-7651 + FM16(x2, (8277 + FM16(-626, x2))))));
   return (s16)(1 + x2);
}
static int bitexact_log2tan(int isin, int icos) {                                                   /* bands.c:82-93 */
   int lc = ilog((u32)icos), ls = ilog((u32)isin);
   icos <<= 15 - lc; isin <<= 15 - ls;
   return (ls - lc) * 2048 + FM16(isin, FM16(isin, -2597) + 7932) - FM16(icos, FM16(icos, -2597) + 7932);
}
static void theta(BCtx *c, SplitCtx *sc, int N, int *b, int B, int B0, int LM, int stereo, int *fill) { /* :645-818 */
   static const s16 exp2_8[8] = {16384, 17866, 19483, 21247, 23170, 25267, 27554, 30048};
   int itheta = 0, inv = 0, qn, qb, N2 = 2 * N - 1;
   int pulse_cap = ot_logN400[c->i] + LM * 8;
   int offset = (pulse_cap >> 1) - (stereo && N == 2 ? 16 : 4);
   s32 tell;
   if (stereo && N == 2) N2--;                                                                      /* compute_qn :596-623 */
   qb = (*b + N2 * offset) / N2;
   qb = imin(*b - pulse_cap - 32, qb); qb = imin(64, qb);
   if (qb < 4) qn = 1; else { qn = exp2_8[qb & 7] >> (14 - (qb >> 3)); qn = (qn + 1) >> 1 << 1; }
   if (stereo && c->i >= c->intensity) qn = 1;
   tell = (s32)rd_tell_frac(c->d);
   if (qn != 1) {
      if (stereo && N > 2) {
         int x0 = qn / 2, x; u32 ft = (u32)(3 * (x0 + 1) + x0); int fs = (int)rd_decode(c->d, ft);
         x = fs < (x0 + 1) * 3 ? fs / 3 : x0 + 1 + (fs - (x0 + 1) * 3);
         rd_update(c->d, (u32)(x <= x0 ? 3 * x : (x - 1 - x0) + (x0 + 1) * 3), (u32)(x <= x0 ? 3 * (x + 1) : (x - x0) + (x0 + 1) * 3), ft);
         itheta = x;
      } else if (B0 > 1 || stereo) {
         itheta = (int)rd_uint(c->d, (u32)qn + 1);
      } else {
         int h = qn >> 1, ft = (h + 1) * (h + 1), fm = (int)rd_decode(c->d, (u32)ft), fs, fl;
         if (fm < (h * (h + 1) >> 1)) { itheta = (int)((isqrt32(8 * (u32)fm + 1) - 1) >> 1); fs = itheta + 1; fl = itheta * (itheta + 1) >> 1; }
         else { itheta = (int)((2 * (u32)(qn + 1) - isqrt32(8 * (u32)(ft - fm - 1) + 1)) >> 1); fs = qn + 1 - itheta; fl = ft - ((qn + 1 - itheta) * (qn + 2 - itheta) >> 1); }
         rd_update(c->d, (u32)fl, (u32)(fl + fs), (u32)ft);
      }
      itheta = (int)((u32)(itheta * 16384) / (u32)qn);
   } else if (stereo) {
      inv = (*b > 16 && c->remaining > 16) ? rd_bit_logp(c->d, 2) : 0;
      itheta = 0;
   }
   sc->qalloc = (s32)rd_tell_frac(c->d) - tell;
   *b -= sc->qalloc;
   if (itheta == 0) { sc->imid = 32767; sc->iside = 0; *fill &= (1 << B) - 1; sc->delta = -16384; }
   else if (itheta == 16384) { sc->imid = 0; sc->iside = 32767; *fill &= ((1 << B) - 1) << B; sc->delta = 16384; }
   else {
      sc->imid = bitexact_cos((s16)itheta); sc->iside = bitexact_cos((s16)(16384 - itheta));
      sc->delta = FM16((N - 1) << 7, bitexact_log2tan(sc->iside, sc->imid));
   }
   sc->inv = inv; sc->itheta = itheta;
}
static unsigned band_n1(BCtx *c, s16 *X, s16 *Y, s16 *lowband_out) {                                /* bands.c:819-858 */
   s16 *x = X; int k;
   for (k = 0; k < (Y ? 2 : 1); k++) {
      int sign = 0;
      if (c->remaining >= 8) { sign = (int)rd_bits(c->d, 1); c->remaining -= 8; }
      x[0] = sign ? -16384 : 16384;
      x = Y;
   }
   if (lowband_out) lowband_out[0] = (s16)(X[0] >> 4);
   return 1;
}
static unsigned partition(BCtx *c, s16 *X, int N, int b, int B, s16 *lowband, int LM, s32 gain, int fill) { /* :864-1040 */
   const u8 *cache = pcache(c->i, LM);
   unsigned cm = 0;
   if (LM != -1 && b > cache[cache[0]] + 12 && N > 2) {
      SplitCtx sc; int B0 = B, mbits, sbits, delta; s16 *Y; s16 *lb2 = NULL; s32 rebal;
      N >>= 1; Y = X + N; LM--;
      if (B == 1) fill = (fill & 1) | (fill << 1);
      B = (B + 1) >> 1;
      theta(c, &sc, N, &b, B, B0, LM, 0, &fill);
      delta = sc.delta;
      if (B0 > 1 && (sc.itheta & 0x3fff)) {
         if (sc.itheta > 8192) delta -= delta >> (4 - LM);
         else delta = imin(0, delta + (N << 3 >> (5 - LM)));
      }
      mbits = imax(0, imin(b, (b - delta) / 2)); sbits = b - mbits;
      c->remaining -= sc.qalloc;
      if (lowband) lb2 = lowband + N;
      rebal = c->remaining;
      if (mbits >= sbits) {
         cm = partition(c, X, N, mbits, B, lowband, LM, (s16)MP15(gain, sc.imid), fill);
         rebal = mbits - (rebal - c->remaining);
         if (rebal > 24 && sc.itheta != 0) sbits += rebal - 24;
         cm |= partition(c, Y, N, sbits, B, lb2, LM, (s16)MP15(gain, sc.iside), fill >> B) << (B0 >> 1);
      } else {
         cm = partition(c, Y, N, sbits, B, lb2, LM, (s16)MP15(gain, sc.iside), fill >> B) << (B0 >> 1);
         rebal = sbits - (rebal - c->remaining);
         if (rebal > 24 && sc.itheta != 16384) mbits += rebal - 24;
         cm |= partition(c, X, N, mbits, B, lowband, LM, (s16)MP15(gain, sc.imid), fill);
      }
   } else {
      int q = bits2pulses(c->i, LM, b), curr = pulses2bits(c->i, LM, q), j;
      c->remaining -= curr;
      while (c->remaining < 0 && q > 0) { c->remaining += curr; q--; curr = pulses2bits(c->i, LM, q); c->remaining -= curr; }
      if (q != 0) cm = alg_unquant(c->d, X, N, get_pulses(q), c->spread, B, gain);
      else {
         unsigned cm_mask = (unsigned)((1UL << B) - 1);
         fill &= (int)cm_mask;
         if (!fill) memset(X, 0, N * sizeof(s16));
         else {
            if (lowband == NULL) { for (j = 0; j < N; j++) { c->seed = lcg(c->seed); X[j] = (s16)((s32)c->seed >> 20); } cm = cm_mask; }
            else { for (j = 0; j < N; j++) { c->seed = lcg(c->seed); X[j] = (s16)(lowband[j] + ((c->seed & 0x8000) ? 4 : -4)); } cm = (unsigned)fill; }
            renormalise(X, N, gain);
         }
      }
   }
   return cm;
}
static unsigned band(BCtx *c, s16 *X, int N, int b, int B, s16 *lowband, int LM, s16 *lowband_out, s32 gain,
                     s16 *scratch, int fill) {                                                      /* bands.c:1044-1172 */
   static const u8 inter[16] = {0, 1, 1, 1, 2, 3, 3, 3, 2, 3, 3, 3, 2, 3, 3, 3};
   static const u8 deinter[16] = {0x00, 0x03, 0x0C, 0x0F, 0x30, 0x33, 0x3C, 0x3F, 0xC0, 0xC3, 0xCC, 0xCF, 0xF0, 0xF3, 0xFC, 0xFF};
   int N0 = N, N_B = N / B, B0 = B, time_divide = 0, recombine = 0, longBlocks = B == 1, tf = c->tf_change, k, N_B0;
   unsigned cm;
   if (N == 1) return band_n1(c, X, NULL, lowband_out);
   if (tf > 0) recombine = tf;
   if (scratch && lowband && (recombine || ((N_B & 1) == 0 && tf < 0) || B0 > 1)) { memcpy(scratch, lowband, N * sizeof(s16)); lowband = scratch; }
   for (k = 0; k < recombine; k++) { if (lowband) haar1(lowband, N >> k, 1 << k); fill = inter[fill & 0xF] | inter[fill >> 4] << 2; }
   B >>= recombine; N_B <<= recombine;
   while ((N_B & 1) == 0 && tf < 0) { if (lowband) haar1(lowband, N_B, B); fill |= fill << B; B <<= 1; N_B >>= 1; time_divide++; tf++; }
   B0 = B; N_B0 = N_B;
   if (B0 > 1 && lowband) hadamard(lowband, N_B >> recombine, B0 << recombine, longBlocks, 0);
   cm = partition(c, X, N, b, B, lowband, LM, gain, fill);
   if (B0 > 1) hadamard(X, N_B >> recombine, B0 << recombine, longBlocks, 1);
   N_B = N_B0; B = B0;
   for (k = 0; k < time_divide; k++) { B >>= 1; N_B <<= 1; cm |= cm >> B; haar1(X, N_B, B); }
   for (k = 0; k < recombine; k++) { cm = deinter[cm]; haar1(X, N0 >> k, 1 << k); }
   B <<= recombine;
   if (lowband_out) { s16 n = (s16)celt_sqrt(SHL(N0, 22)); int j; for (j = 0; j < N0; j++) lowband_out[j] = (s16)MQ15(n, X[j]); }
   return cm & (unsigned)((1 << B) - 1);
}
static unsigned band_stereo(BCtx *c, s16 *X, s16 *Y, int N, int b, int B, s16 *lowband, int LM, s16 *lowband_out,
                            s16 *scratch, int fill) {                                               /* bands.c:1176-1334 */
   SplitCtx sc; unsigned cm; int orig_fill = fill, j;
   if (N == 1) return band_n1(c, X, Y, lowband_out);
   theta(c, &sc, N, &b, B, B, LM, 1, &fill);
   if (N == 2) {
      int mbits = b, sbits = 0, sw, sign = 0; s16 *x2, *y2; s16 X0, X1, Y0, Y1;
      if (sc.itheta != 0 && sc.itheta != 16384) sbits = 8;
      mbits -= sbits; sw = sc.itheta > 8192;
      c->remaining -= sc.qalloc + sbits;
      x2 = sw ? Y : X; y2 = sw ? X : Y;
      if (sbits) sign = (int)rd_bits(c->d, 1);
      sign = 1 - 2 * sign;
      cm = band(c, x2, N, mbits, B, lowband, LM, lowband_out, 32767, scratch, orig_fill);
      y2[0] = (s16)(-sign * x2[1]); y2[1] = (s16)(sign * x2[0]);
      X0 = (s16)MQ15(sc.imid, X[0]); X1 = (s16)MQ15(sc.imid, X[1]); Y0 = (s16)MQ15(sc.iside, Y[0]); Y1 = (s16)MQ15(sc.iside, Y[1]);
      X[0] = (s16)S16(X0, Y0); Y[0] = A16(X0, Y0); X[1] = (s16)S16(X1, Y1); Y[1] = A16(X1, Y1);
   } else {
      int mbits = imax(0, imin(b, (b - sc.delta) / 2)), sbits = b - mbits; s32 rebal;
      c->remaining -= sc.qalloc; rebal = c->remaining;
      if (mbits >= sbits) {
         cm = band(c, X, N, mbits, B, lowband, LM, lowband_out, 32767, scratch, fill);
         rebal = mbits - (rebal - c->remaining);
         if (rebal > 24 && sc.itheta != 0) sbits += rebal - 24;
         cm |= band(c, Y, N, sbits, B, NULL, LM, NULL, sc.iside, NULL, fill >> B);
      } else {
         cm = band(c, Y, N, sbits, B, NULL, LM, NULL, sc.iside, NULL, fill >> B);
         rebal = sbits - (rebal - c->remaining);
         if (rebal > 24 && sc.itheta != 16384) mbits += rebal - 24;
         cm |= band(c, X, N, mbits, B, lowband, LM, lowband_out, 32767, scratch, fill);
      }
      stereo_merge(X, Y, sc.imid, N);
   }
   if (sc.inv) for (j = 0; j < N; j++) Y[j] = (s16)(-Y[j]);
   return cm;
}
static void all_bands(RDec *d, int start, int end, s16 *X_, s16 *Y_, u8 *masks, const int *pulses, int shortBlocks,
                      int spread, int dual, int intensity, const int *tf_res, s32 total_bits, s32 balance, int LM,
                      int coded, u32 *seed) {                                                        /* bands.c:1337-1502 */
   int M = 1 << LM, B = shortBlocks ? M : 1, C = Y_ ? 2 : 1, i, lowband_offset = 0, update_lowband = 1;
   int norm_offset = M * EB(start);
   s16 normbuf[2 * 8 * 78];
   s16 *norm = normbuf, *norm2 = normbuf + M * EB(NB - 1) - norm_offset;
   s16 *scratch = X_ + M * EB(NB - 1);
   BCtx c; c.d = d; c.intensity = intensity; c.spread = spread; c.seed = *seed;
   for (i = start; i < end; i++) {
      int last = i == end - 1, N = M * EB(i + 1) - M * EB(i), b, eff = -1; unsigned x_cm, y_cm;
      s16 *X = X_ + M * EB(i), *Y = Y_ ? Y_ + M * EB(i) : NULL;
      s32 tell = (s32)rd_tell_frac(d);
      c.i = i;
      if (i != start) balance -= tell;
      c.remaining = total_bits - tell - 1;
      if (i <= coded - 1) { s32 cb = balance / imin(3, coded - i); b = imax(0, imin(16383, imin(c.remaining + 1, pulses[i] + cb))); }
      else b = 0;
      if (M * EB(i) - N >= M * EB(start) && (update_lowband || lowband_offset == 0)) lowband_offset = i;
      c.tf_change = tf_res[i];
      if (last) scratch = NULL;
      if (lowband_offset != 0 && (spread != 3 || B > 1 || c.tf_change < 0)) {
         int fs, fe, fi;
         eff = imax(0, M * EB(lowband_offset) - norm_offset - N);
         fs = lowband_offset; while (M * EB(--fs) > eff + norm_offset) {}
         fe = lowband_offset - 1; while (M * EB(++fe) < eff + norm_offset + N) {}
         x_cm = y_cm = 0; fi = fs;
         do { x_cm |= masks[fi * C]; y_cm |= masks[fi * C + C - 1]; } while (++fi < fe);
      } else x_cm = y_cm = (unsigned)((1 << B) - 1);
      if (dual && i == intensity) { int j; dual = 0; for (j = 0; j < M * EB(i) - norm_offset; j++) norm[j] = (s16)(((s32)norm[j] + norm2[j]) >> 1); }
      {
         s16 *lb = eff != -1 ? norm + eff : NULL, *lb2 = eff != -1 ? norm2 + eff : NULL;
         s16 *lo = last ? NULL : norm + M * EB(i) - norm_offset, *lo2 = last ? NULL : norm2 + M * EB(i) - norm_offset;
         if (dual) {
            x_cm = band(&c, X, N, b / 2, B, lb, LM, lo, 32767, scratch, (int)x_cm);
            y_cm = band(&c, Y, N, b / 2, B, lb2, LM, lo2, 32767, scratch, (int)y_cm);
         } else {
            x_cm = Y ? band_stereo(&c, X, Y, N, b, B, lb, LM, lo, scratch, (int)(x_cm | y_cm))
                     : band(&c, X, N, b, B, lb, LM, lo, 32767, scratch, (int)(x_cm | y_cm));
            y_cm = x_cm;
         }
      }
      masks[i * C] = (u8)x_cm; masks[i * C + C - 1] = (u8)y_cm;
      balance += pulses[i] + tell;
      update_lowband = b > (N << 3);
   }
   *seed = c.seed;
}
static void anti_collapse(s16 *X_, const u8 *masks, int LM, int C, int size, int start, int end, const s16 *logE,
                          const s16 *p1, const s16 *p2, const int *pulses, u32 seed) {              /* bands.c:241-335 */
   int i, c, j, k;
   for (i = start; i < end; i++) {
      int N0 = EB(i + 1) - EB(i), depth = (int)((u32)(1 + pulses[i]) / (u32)N0) >> LM, shift;
      s32 t32 = celt_exp2(-(s32)SL16(depth, 7)) >> 1, t;
      s16 thresh = (s16)M32Q15(16384, imin(32767, t32)), sqrt_1;
      t = N0 << LM; shift = ilog2(t) >> 1; t = SHL(t, (7 - shift) << 1); sqrt_1 = rsqrt_norm(t);
      for (c = 0; c < C; c++) {
         s32 a = p1[c * NB + i], b = p2[c * NB + i], Ediff; s16 r; int renorm = 0; s16 *X;
         if (C == 1) { a = imax(a, p1[NB + i]); b = imax(b, p2[NB + i]); }
         Ediff = imax(0, (s32)logE[c * NB + i] - imin(a, b));
         if (Ediff < 16384) { s32 r32 = celt_exp2(-(s32)(s16)Ediff) >> 1; r = (s16)(2 * imin(16383, r32)); } else r = 0;
         if (LM == 3) r = (s16)(M16(23170, imin(23169, r)) >> 14);
         r = (s16)(imin(thresh, r) >> 1);
         r = (s16)(MQ15(sqrt_1, r) >> shift);
         X = X_ + c * size + (EB(i) << LM);
         for (k = 0; k < 1 << LM; k++) if (!(masks[i * C + c] & 1 << k)) {
            for (j = 0; j < N0; j++) { seed = lcg(seed); X[(j << LM) + k] = (seed & 0x8000) ? r : (s16)(-r); }
            renorm = 1;
         }
         if (renorm) renormalise(X, N0 << LM, 32767);
      }
   }
}

/* ---------- synthesis: celt/bands.c:169-238, celt/mdct.c:263-363, celt/kiss_fft.c ---------- */
static void denormalise(const s16 *X, s32 *freq, const s16 *logE, int start, int end, int M, int downsample, int silence) {
   int N = M * 120, bound = M * EB(end), i, j;
   if (downsample != 1) bound = imin(bound, N / downsample);
   if (silence) { bound = 0; start = end = 0; }
   for (j = 0; j < M * EB(start); j++) freq[j] = 0;
   for (i = start; i < end; i++) {
      s16 lg = A16(logE[i], SL16(ot_eMeans[i], 6)); int shift = 16 - (lg >> 10); s32 g;
      if (shift > 31) { shift = 0; g = 0; } else g = (s16)exp2_frac(lg & 1023);
      if (shift < 0) { if (shift < -2) { g = 32767; shift = -2; } for (j = M * EB(i); j < M * EB(i + 1); j++) freq[j] = SHL(M16(X[j], g), -shift); }
      else for (j = M * EB(i); j < M * EB(i + 1); j++) freq[j] = M16(X[j], g) >> shift;
   }
   for (j = bound; j < N; j++) freq[j] = 0;
}
typedef struct { s32 r, i; } cpx;
static s32 SM(s32 a, s32 b) { return M32Q15(b, a); }                                                /* S_MUL _kiss_fft_guts.h:57 */
static cpx cmul(cpx a, int twi) {                                                                   /* C_MUL :59 */
   s32 tr = ot_fft_tw480[2 * twi], ti = ot_fft_tw480[2 * twi + 1]; cpx m;
   m.r = SUB(SM(a.r, tr), SM(a.i, ti)); m.i = ADD(SM(a.r, ti), SM(a.i, tr)); return m;
}
static cpx cadd(cpx a, cpx b) { cpx r; r.r = ADD(a.r, b.r); r.i = ADD(a.i, b.i); return r; }
static cpx csub(cpx a, cpx b) { cpx r; r.r = SUB(a.r, b.r); r.i = SUB(a.i, b.i); return r; }
static void bfly2(cpx *F, int N) {                                                                  /* kiss_fft.c:51-108 (m==4) */
   int i;
   for (i = 0; i < N; i++, F += 8) {
      cpx t, *F2 = F + 4;
      t = F2[0]; F2[0] = csub(F[0], t); F[0] = cadd(F[0], t);
      t.r = SM(ADD(F2[1].r, F2[1].i), 23170); t.i = SM(SUB(F2[1].i, F2[1].r), 23170); F2[1] = csub(F[1], t); F[1] = cadd(F[1], t);
      t.r = F2[2].i; t.i = SUB(0, F2[2].r); F2[2] = csub(F[2], t); F[2] = cadd(F[2], t);
      t.r = SM(SUB(F2[3].i, F2[3].r), 23170); t.i = SM(SUB(SUB(0, F2[3].i), F2[3].r), 23170); F2[3] = csub(F[3], t); F[3] = cadd(F[3], t);
   }
}
static void bfly4(cpx *Fb, int fs, int m, int N, int mm) {                                          /* kiss_fft.c:111-182 */
   int i, j;
   for (i = 0; i < N; i++) {
      cpx *F = Fb + i * mm;
      for (j = 0; j < m; j++, F++) {
         cpx s0 = m == 1 ? F[m] : cmul(F[m], j * fs), s1 = m == 1 ? F[2 * m] : cmul(F[2 * m], 2 * j * fs),
             s2 = m == 1 ? F[3 * m] : cmul(F[3 * m], 3 * j * fs), s5, s3, s4;
         s5 = csub(F[0], s1); F[0] = cadd(F[0], s1); s3 = cadd(s0, s2); s4 = csub(s0, s2);
         F[2 * m] = csub(F[0], s3); F[0] = cadd(F[0], s3);
         F[m].r = ADD(s5.r, s4.i); F[m].i = SUB(s5.i, s4.r); F[3 * m].r = SUB(s5.r, s4.i); F[3 * m].i = ADD(s5.i, s4.r);
      }
   }
}
static void bfly3(cpx *Fb, int fs, int m, int N, int mm) {                                          /* kiss_fft.c:185-242 */
   int i, j;
   for (i = 0; i < N; i++) {
      cpx *F = Fb + i * mm;
      for (j = 0; j < m; j++, F++) {
         cpx s1 = cmul(F[m], j * fs), s2 = cmul(F[2 * m], 2 * j * fs), s3 = cadd(s1, s2), s0 = csub(s1, s2);
         F[m].r = SUB(F[0].r, s3.r >> 1); F[m].i = SUB(F[0].i, s3.i >> 1);
         s0.r = SM(s0.r, -28378); s0.i = SM(s0.i, -28378);
         F[0] = cadd(F[0], s3);
         F[2 * m].r = ADD(F[m].r, s0.i); F[2 * m].i = SUB(F[m].i, s0.r);
         F[m].r = SUB(F[m].r, s0.i); F[m].i = ADD(F[m].i, s0.r);
      }
   }
}
static void bfly5(cpx *Fb, int fs, int m, int N, int mm) {                                          /* kiss_fft.c:245-330 */
   const s32 yar = 10126, yai = -31164, ybr = -26510, ybi = -19261; int i, u;
   for (i = 0; i < N; i++) {
      cpx *F0 = Fb + i * mm, *F1 = F0 + m, *F2 = F0 + 2 * m, *F3 = F0 + 3 * m, *F4 = F0 + 4 * m;
      for (u = 0; u < m; u++, F0++, F1++, F2++, F3++, F4++) {
         cpx s0 = *F0, s1 = cmul(*F1, u * fs), s2 = cmul(*F2, 2 * u * fs), s3 = cmul(*F3, 3 * u * fs), s4 = cmul(*F4, 4 * u * fs);
         cpx s7 = cadd(s1, s4), s10 = csub(s1, s4), s8 = cadd(s2, s3), s9 = csub(s2, s3), s5, s6, s11, s12;
         F0->r = ADD(s0.r, ADD(s7.r, s8.r)); F0->i = ADD(s0.i, ADD(s7.i, s8.i));
         s5.r = ADD(ADD(s0.r, SM(s7.r, yar)), SM(s8.r, ybr)); s5.i = ADD(ADD(s0.i, SM(s7.i, yar)), SM(s8.i, ybr));
         s6.r = ADD(SM(s10.i, yai), SM(s9.i, ybi)); s6.i = SUB(SUB(0, SM(s10.r, yai)), SM(s9.r, ybi));
         *F1 = csub(s5, s6); *F4 = cadd(s5, s6);
         s11.r = ADD(ADD(s0.r, SM(s7.r, ybr)), SM(s8.r, yar)); s11.i = ADD(ADD(s0.i, SM(s7.i, ybr)), SM(s8.i, yar));
         s12.r = ADD(SUB(0, SM(s10.i, ybi)), SM(s9.i, yai)); s12.i = SUB(SM(s10.r, ybi), SM(s9.r, yai));
         *F2 = cadd(s11, s12); *F3 = csub(s11, s12);
      }
   }
}
static void fft(cpx *f, int sh) {                                                                   /* opus_fft_impl kiss_fft.c:532-578 */
   static const int fac[4][10] = {{5, 96, 3, 32, 4, 8, 2, 4, 4, 1}, {5, 48, 3, 16, 4, 4, 4, 1, 0, 0},
                                  {5, 24, 3, 8, 2, 4, 4, 1, 0, 0}, {5, 12, 3, 4, 4, 1, 0, 0, 0, 0}};   /* static_modes_fixed.h:432-500 */
   int fstride[6], L = 0, i, m, m2, p;
   fstride[0] = 1;
   do { p = fac[sh][2 * L]; m = fac[sh][2 * L + 1]; fstride[L + 1] = fstride[L] * p; L++; } while (m != 1);
   m = fac[sh][2 * L - 1];
   for (i = L - 1; i >= 0; i--) {
      m2 = i ? fac[sh][2 * i - 1] : 1;
      switch (fac[sh][2 * i]) {
         case 2: bfly2(f, fstride[i]); break;
         case 4: bfly4(f, fstride[i] << sh, m, fstride[i], m2); break;
         case 3: bfly3(f, fstride[i] << sh, m, fstride[i], m2); break;
         default: bfly5(f, fstride[i] << sh, m, fstride[i], m2); break;
      }
      m = m2;
   }
}
static void imdct(const s32 *in, s32 *out, int shift, int stride) {                                 /* mdct.c:263-363 */
   static const int trig_off[4] = {0, 960, 1440, 1680};
   const int16_t *trig = ot_mdct_tw960 + trig_off[shift];
   const int16_t *rev = shift == 0 ? ot_fft_bitrev480 : shift == 1 ? ot_fft_bitrev240 : shift == 2 ? ot_fft_bitrev120 : ot_fft_bitrev60;
   int N = 1920 >> shift, N2 = N >> 1, N4 = N >> 2, i;
   s32 *yp = out + 60;
   for (i = 0; i < N4; i++) {
      s32 x1 = in[2 * stride * i], x2 = in[stride * (N2 - 1) - 2 * stride * i];
      yp[2 * rev[i] + 1] = ADD(SM(x2, trig[i]), SM(x1, trig[N4 + i]));
      yp[2 * rev[i]] = SUB(SM(x1, trig[i]), SM(x2, trig[N4 + i]));
   }
   fft((cpx *)yp, shift);
   for (i = 0; i < (N4 + 1) >> 1; i++) {
      s32 *y0 = yp + 2 * i, *y1 = yp + N2 - 2 - 2 * i;
      s32 re = y0[1], im = y0[0], t0 = trig[i], t1 = trig[N4 + i], yr, yi, re2, im2;
      yr = ADD(SM(re, t0), SM(im, t1)); yi = SUB(SM(re, t1), SM(im, t0));
      re2 = y1[1]; im2 = y1[0]; y0[0] = yr; y1[1] = yi;
      t0 = trig[N4 - i - 1]; t1 = trig[N2 - i - 1];
      yr = ADD(SM(re2, t0), SM(im2, t1)); yi = SUB(SM(re2, t1), SM(im2, t0));
      y1[0] = yr; y0[1] = yi;
   }
   for (i = 0; i < 60; i++) {
      s32 x1 = out[119 - i], x2 = out[i], w1 = ot_window120[i], w2 = ot_window120[119 - i];
      out[i] = SUB(M32Q15(w2, x2), M32Q15(w1, x1));
      out[119 - i] = ADD(M32Q15(w1, x2), M32Q15(w2, x1));
   }
}
static void synthesis(struct OracleCelt *st, const s16 *X, s32 *out_syn[2], const s16 *bandE, int start, int effEnd,
                      int C, int CC, int transient, int LM, int silence) {                          /* celt_decoder.c:280-350 */
   s32 freq[960], freq2[960]; int M = 1 << LM, N = M * 120, B = transient ? M : 1, NBk = transient ? 120 : N;
   int shift = transient ? 3 : 3 - LM, b, c, i;
   if (CC == 2 && C == 1) {
      denormalise(X, freq, bandE, start, effEnd, M, st->downsample, silence);
      memcpy(freq2, freq, N * sizeof(s32));
      for (b = 0; b < B; b++) imdct(&freq2[b], out_syn[0] + NBk * b, shift, B);
      for (b = 0; b < B; b++) imdct(&freq[b], out_syn[1] + NBk * b, shift, B);
   } else if (CC == 1 && C == 2) {
      denormalise(X, freq, bandE, start, effEnd, M, st->downsample, silence);
      denormalise(X + N, freq2, bandE + NB, start, effEnd, M, st->downsample, silence);
      for (i = 0; i < N; i++) freq[i] = ADD(freq[i], freq2[i]) >> 1;
      for (b = 0; b < B; b++) imdct(&freq[b], out_syn[0] + NBk * b, shift, B);
   } else for (c = 0; c < CC; c++) {
      denormalise(X + c * N, freq, bandE + c * NB, start, effEnd, M, st->downsample, silence);
      for (b = 0; b < B; b++) imdct(&freq[b], out_syn[c] + NBk * b, shift, B);
   }
}
static void comb_filter(s32 *y, s32 *x, int T0, int T1, int N, s32 g0, s32 g1, int tap0, int tap1, int overlap) { /* celt.c:183-244 */
   static const s16 gains[3][3] = {{10048, 7112, 4248}, {15200, 8784, 0}, {26208, 3280, 0}};
   s32 g00, g01, g02, g10, g11, g12; int i;
   if (g0 == 0 && g1 == 0) { if (x != y) memmove(y, x, N * sizeof(s32)); return; }
   g00 = MP15(g0, gains[tap0][0]); g01 = MP15(g0, gains[tap0][1]); g02 = MP15(g0, gains[tap0][2]);
   g10 = MP15(g1, gains[tap1][0]); g11 = MP15(g1, gains[tap1][1]); g12 = MP15(g1, gains[tap1][2]);
   if (g0 == g1 && T0 == T1 && tap0 == tap1) overlap = 0;
   for (i = 0; i < overlap; i++) {
      s32 f = (s16)MQ15(ot_window120[i], ot_window120[i]), v = x[i];
      v = ADD(v, M32Q15(MQ15(32767 - f, g00), x[i - T0]));
      v = ADD(v, M32Q15(MQ15(32767 - f, g01), ADD(x[i - T0 + 1], x[i - T0 - 1])));
      v = ADD(v, M32Q15(MQ15(32767 - f, g02), ADD(x[i - T0 + 2], x[i - T0 - 2])));
      v = ADD(v, M32Q15(MQ15(f, g10), x[i - T1]));
      v = ADD(v, M32Q15(MQ15(f, g11), ADD(x[i - T1 + 1], x[i - T1 - 1])));
      v = ADD(v, M32Q15(MQ15(f, g12), ADD(x[i - T1 + 2], x[i - T1 - 2])));
      y[i] = v;
   }
   if (g1 == 0) { if (x != y) memmove(y + overlap, x + overlap, (N - overlap) * sizeof(s32)); return; }
   for (; i < N; i++) {                                                                             /* comb_filter_const_c :156-178 */
      s32 v = x[i];
      v = ADD(v, M32Q15(g10, x[i - T1]));
      v = ADD(v, M32Q15(g11, ADD(x[i - T1 + 1], x[i - T1 - 1])));
      v = ADD(v, M32Q15(g12, ADD(x[i - T1 + 2], x[i - T1 - 2])));
      y[i] = v;
   }
}
static void deemphasis(struct OracleCelt *st, s32 *in[2], s16 *pcm, int N, int accum) {             /* celt_decoder.c:185-275 */
   int CC = st->channels, ds = st->downsample, c, j;
   for (c = 0; c < CC; c++) {
      s32 m = st->preemph_mem[c], *x = in[c]; s16 *y = pcm + c;
      for (j = 0; j < N; j++) {
         s32 t = ADD(x[j], m);
         m = M32Q15(27853, t);
         if (j % ds == 0) {
            s16 v = sig2word(t);
            y[(j / ds) * CC] = accum ? sat16((s32)y[(j / ds) * CC] + v) : v;
         }
      }
      st->preemph_mem[c] = m;
   }
}

/* ---------- packet-loss concealment: celt_decoder.c:399-711, pitch.c, celt_lpc.c ---------- */
static int autocorr(const s16 *x, s32 *ac, int use_window, int lag, int n) {                        /* celt_lpc.c:232-328 */
   s16 xx[1024]; const s16 *xp = x; int i, k, shift; s32 ac0;
   if (use_window) {
      memcpy(xx, x, n * sizeof(s16));
      for (i = 0; i < OVL; i++) { xx[i] = (s16)MQ15(x[i], ot_window120[i]); xx[n - i - 1] = (s16)MQ15(x[n - i - 1], ot_window120[i]); }
      xp = xx;
   }
   ac0 = 1 + (n << 7);
   for (i = 0; i < n; i++) ac0 = ADD(ac0, M16(xp[i], xp[i]) >> 9);
   shift = (ilog2(ac0) - 30 + 10) / 2;
   if (shift > 0) { for (i = 0; i < n; i++) xx[i] = (s16)PSHR(xp[i], shift); xp = xx; } else shift = 0;
   for (k = 0; k <= lag; k++) { s32 d = 0; for (i = k; i < n; i++) d = ADD(d, M16(xp[i], xp[i - k])); ac[k] = d; }
   shift *= 2;
   if (shift <= 0) ac[0] = ADD(ac[0], SHL(1, -shift));
   if (ac[0] < 268435456) { int s2 = 29 - ilog((u32)ac[0]); for (i = 0; i <= lag; i++) ac[i] = SHL(ac[i], s2); shift -= s2; }
   else if (ac[0] >= 536870912) { int s2 = ac[0] >= 1073741824 ? 2 : 1; for (i = 0; i <= lag; i++) ac[i] >>= s2; shift += s2; }
   return shift;
}
static void lpc_from_ac(s16 *out, const s32 *ac, int p) {                                           /* celt_lpc.c:37-93 */
   s32 lpc[LPC_ORD], error = ac[0]; int i, j;
   memset(lpc, 0, sizeof(lpc));
   if (ac[0] != 0) for (i = 0; i < p; i++) {
      s32 rr = 0, r;
      for (j = 0; j < i; j++) rr = ADD(rr, M3232Q31(lpc[j], ac[i - j]));
      rr = ADD(rr, ac[i + 1] >> 3);
      r = SUB(0, frac_div32(SHL(rr, 3), error));
      lpc[i] = r >> 3;
      for (j = 0; j < (i + 1) >> 1; j++) { s32 t1 = lpc[j], t2 = lpc[i - 1 - j]; lpc[j] = ADD(t1, M3232Q31(r, t2)); lpc[i - 1 - j] = ADD(t2, M3232Q31(r, t1)); }
      error = SUB(error, M3232Q31(M3232Q31(r, r), error));
      if (error < (ac[0] >> 10)) break;
   }
   for (i = 0; i < p; i++) out[i] = R16(lpc[i], 16);
}
static void best_pitch(const s32 *xcorr, const s16 *y, int len, int max_pitch, int *bp, int yshift, s32 maxcorr) { /* pitch.c:45-103 */
   s32 Syy = 1, bn0 = -1, bn1 = -1, bd0 = 0, bd1 = 0; int i, j, xshift = ilog2(maxcorr) - 14;
   bp[0] = 0; bp[1] = 1;
   for (j = 0; j < len; j++) Syy = ADD(Syy, M16(y[j], y[j]) >> yshift);
   for (i = 0; i < max_pitch; i++) {
      if (xcorr[i] > 0) {
         s32 x16 = (s16)VSHR(xcorr[i], xshift), num = (s16)MQ15(x16, x16);
         if (M32Q15(num, bd1) > M32Q15(bn1, Syy)) {
            if (M32Q15(num, bd0) > M32Q15(bn0, Syy)) { bn1 = bn0; bd1 = bd0; bp[1] = bp[0]; bn0 = num; bd0 = Syy; bp[0] = i; }
            else { bn1 = num; bd1 = Syy; bp[1] = i; }
         }
      }
      Syy = ADD(Syy, SUB(M16(y[i + len], y[i + len]) >> yshift, M16(y[i], y[i]) >> yshift));
      Syy = imax(1, Syy);
   }
}
static int plc_pitch_search(struct OracleCelt *st) {                                                /* celt_decoder.c:399-413 */
   s16 xlp[1024], tmp[1024], x4[332], y4[487], lpc[4], l4[4]; s32 ac[5], xcorr[310], maxabs = 0, maxcorr, t;
   int C = st->channels, c, i, j, shift, bp[2] = {0, 0}, offset = 0;
   const int len = 1328, max_pitch = 620, lag = len + max_pitch;
   const s16 *xs = xlp + 360, *y = xlp;
   for (c = 0; c < C; c++) for (i = 0; i < DBUF; i++) { s32 v = st->mem[c][i]; maxabs = imax(maxabs, v < 0 ? SUB(0, v) : v); }   /* pitch.c:147-222 */
   if (maxabs < 1) maxabs = 1;
   shift = imax(0, ilog2(maxabs) - 10); if (C == 2) shift++;
   for (i = 0; i < 1024; i++) {
      s32 acc = 0;
      for (c = 0; c < C; c++) {
         s32 *m = st->mem[c];
         s32 v = i == 0 ? (ADD(m[1] >> 1, m[0]) >> 1) >> shift : (ADD(ADD(m[2 * i - 1], m[2 * i + 1]) >> 1, m[2 * i]) >> 1) >> shift;
         acc = c == 0 ? (s16)v : (s16)(acc + v);
      }
      xlp[i] = (s16)acc;
   }
   autocorr(xlp, ac, 0, 4, 1024);
   ac[0] = ADD(ac[0], ac[0] >> 13);
   for (i = 1; i <= 4; i++) ac[i] = SUB(ac[i], M32Q15(2 * i * i, ac[i]));
   lpc_from_ac(l4, ac, 4);
   t = 32767;
   for (i = 0; i < 4; i++) { t = (s16)MQ15(29491, t); lpc[i] = (s16)MQ15(l4[i], t); }
   { s16 n0 = (s16)(lpc[0] + 3277), n1 = (s16)(lpc[1] + MQ15(26214, lpc[0])), n2 = (s16)(lpc[2] + MQ15(26214, lpc[1])),
         n3 = (s16)(lpc[3] + MQ15(26214, lpc[2])), n4 = (s16)MQ15(26214, lpc[3]);                   /* celt_fir5 pitch.c:105-145 */
     for (i = 0; i < 1024; i++) {
        s32 sum = SHL(xlp[i], 12);
        sum = ADD(sum, M16(n0, i >= 1 ? xlp[i - 1] : 0)); sum = ADD(sum, M16(n1, i >= 2 ? xlp[i - 2] : 0));
        sum = ADD(sum, M16(n2, i >= 3 ? xlp[i - 3] : 0)); sum = ADD(sum, M16(n3, i >= 4 ? xlp[i - 4] : 0));
        sum = ADD(sum, M16(n4, i >= 5 ? xlp[i - 5] : 0));
        tmp[i] = R16(sum, 12);
     }
     memcpy(xlp, tmp, sizeof(xlp)); }
   { s32 xmax = 0, ymax = 0; int sh;                                                                /* pitch_search pitch.c:260-369 */
     for (j = 0; j < len >> 2; j++) { x4[j] = xs[2 * j]; xmax = imax(xmax, x4[j] < 0 ? -x4[j] : x4[j]); }
     for (j = 0; j < lag >> 2; j++) { y4[j] = y[2 * j]; ymax = imax(ymax, y4[j] < 0 ? -y4[j] : y4[j]); }
     sh = ilog2(imax(1, imax(xmax, ymax))) - 11;
     if (sh > 0) { for (j = 0; j < len >> 2; j++) x4[j] >>= sh; for (j = 0; j < lag >> 2; j++) y4[j] >>= sh; sh *= 2; } else sh = 0;
     maxcorr = 1;
     for (i = 0; i < max_pitch >> 2; i++) { s32 d = 0; for (j = 0; j < len >> 2; j++) d = ADD(d, M16(x4[j], y4[i + j])); xcorr[i] = d; maxcorr = imax(maxcorr, d); }
     best_pitch(xcorr, y4, len >> 2, max_pitch >> 2, bp, 0, maxcorr);
     maxcorr = 1;
     for (i = 0; i < max_pitch >> 1; i++) {
        s32 sum = 0;
        xcorr[i] = 0;
        if (abs(i - 2 * bp[0]) > 2 && abs(i - 2 * bp[1]) > 2) continue;
        for (j = 0; j < len >> 1; j++) sum = ADD(sum, M16(xs[j], y[i + j]) >> sh);
        xcorr[i] = imax(-1, sum); maxcorr = imax(maxcorr, sum);
     }
     best_pitch(xcorr, y, len >> 1, max_pitch >> 1, bp, sh + 1, maxcorr); }
   if (bp[0] > 0 && bp[0] < (max_pitch >> 1) - 1) {
      s32 a = xcorr[bp[0] - 1], b = xcorr[bp[0]], cc = xcorr[bp[0] + 1];
      if (SUB(cc, a) > M32Q15(22938, SUB(b, a))) offset = 1;
      else if (SUB(a, cc) > M32Q15(22938, SUB(b, cc))) offset = -1;
   }
   return 720 - (2 * bp[0] - offset);
}
static void decode_lost(struct OracleCelt *st, int N, int LM) {                                     /* celt_decoder.c:415-711 */
   int C = st->channels, c, i, start = st->start, loss_count = st->loss_count;
   if (loss_count >= 5 || start != 0) {
      s16 X[1920]; int end = st->end, effEnd = imax(start, imin(end, NB)); u32 seed = st->rng; s32 *out_syn[2];
      s32 decay = loss_count == 0 ? 1536 : 512;
      for (c = 0; c < C; c++) for (i = start; i < end; i++) st->oldE[c * NB + i] = (s16)imax(st->bgLogE[c * NB + i], st->oldE[c * NB + i] - decay);
      for (c = 0; c < C; c++) for (i = start; i < effEnd; i++) {
         int boffs = N * c + (EB(i) << LM), blen = (EB(i + 1) - EB(i)) << LM, j;
         for (j = 0; j < blen; j++) { seed = lcg(seed); X[boffs + j] = (s16)((s32)seed >> 20); }
         renormalise(X + boffs, blen, 32767);
      }
      st->rng = seed;
      for (c = 0; c < C; c++) { memmove(st->mem[c], st->mem[c] + N, (DBUF - N + 60) * sizeof(s32)); out_syn[c] = st->mem[c] + DBUF - N; }
      synthesis(st, X, out_syn, st->oldE, start, effEnd, C, C, 0, LM, 0);
   } else {
      s32 fade = 32767, etmp[OVL]; int pitch; s16 exc[1024];
      if (loss_count == 0) st->last_pitch_index = pitch = plc_pitch_search(st);
      else { pitch = st->last_pitch_index; fade = 26214; }
      for (c = 0; c < C; c++) {
         s32 *buf = st->mem[c], S1 = 0, S2 = 0, E1 = 1, E2 = 1, decay, att; s16 *lpc = st->lpc[c], lmem[LPC_ORD], yh[LPC_ORD + 1080];
         int exc_len, shift, dl, j, elen = N + OVL, eoff;
         for (i = 0; i < 1024; i++) exc[i] = R16(buf[DBUF - 1024 + i], 12);
         if (loss_count == 0) {
            s32 ac[LPC_ORD + 1];
            autocorr(exc, ac, 1, LPC_ORD, 1024);
            ac[0] = ADD(ac[0], ac[0] >> 13);
            for (i = 1; i <= LPC_ORD; i++) ac[i] = SUB(ac[i], M32Q15(2 * i * i, ac[i]));
            lpc_from_ac(lpc, ac, LPC_ORD);
         }
         exc_len = imin(2 * pitch, 1024);
         for (i = 0; i < LPC_ORD; i++) lmem[i] = R16(buf[DBUF - exc_len - 1 - i], 12);
         { s16 y[1024], *x = exc + 1024 - exc_len;                                                  /* celt_fir celt_lpc.c:95-149 */
           for (i = 0; i < exc_len; i++) {
              s32 sum = 0;
              for (j = 0; j < LPC_ORD; j++) { int k = i - 1 - j; sum = ADD(sum, M16(lpc[j], k >= 0 ? x[k] : lmem[-k - 1])); }
              y[i] = sat16(ADD(x[i], PSHR(sum, 12)));
           }
           memcpy(x, y, exc_len * sizeof(s16)); }
         { s32 mx = 0; for (i = 0; i < exc_len; i++) { s32 v = exc[1024 - exc_len + i]; mx = imax(mx, v < 0 ? -v : v); }
           shift = imax(0, 2 * (mx <= 0 ? 0 : ilog2(mx)) - 20); }
         dl = exc_len >> 1;
         for (i = 0; i < dl; i++) { s32 e = exc[1024 - dl + i]; E1 = ADD(E1, M16(e, e) >> shift); e = exc[1024 - 2 * dl + i]; E2 = ADD(E2, M16(e, e) >> shift); }
         E1 = imin(E1, E2);
         decay = (s16)celt_sqrt(frac_div32(E1 >> 1, E2));
         memmove(buf, buf + N, (DBUF - N) * sizeof(s32));
         eoff = 1024 - pitch; att = (s16)MQ15(fade, decay);
         for (i = j = 0; i < elen; i++, j++) {
            s32 t;
            if (j >= pitch) { j -= pitch; att = (s16)MQ15(att, decay); }
            buf[DBUF - N + i] = SHL(MQ15(att, exc[eoff + j]), 12);
            t = R16(buf[DBUF - 1024 - N + eoff + j], 12);
            S1 = ADD(S1, M16(t, t) >> 8);
         }
         for (i = 0; i < LPC_ORD; i++) yh[LPC_ORD - 1 - i] = R16(buf[DBUF - N - 1 - i], 12);         /* celt_iir celt_lpc.c:151-230 */
         for (i = 0; i < elen; i++) {
            s32 sum = buf[DBUF - N + i];
            for (j = 0; j < LPC_ORD; j++) sum = SUB(sum, M16(lpc[j], yh[LPC_ORD + i - 1 - j]));
            buf[DBUF - N + i] = sum; yh[LPC_ORD + i] = R16(sum, 12);
         }
         for (i = 0; i < elen; i++) { s32 t = R16(buf[DBUF - N + i], 12); S2 = ADD(S2, M16(t, t) >> 8); }
         if (!(S1 > (S2 >> 2))) for (i = 0; i < elen; i++) buf[DBUF - N + i] = 0;
         else if (S1 < S2) {
            s32 ratio = (s16)celt_sqrt(frac_div32(ADD(S1 >> 1, 1), ADD(S2, 1)));
            for (i = 0; i < OVL; i++) { s32 g = (s16)(32767 - MQ15(ot_window120[i], 32767 - ratio)); buf[DBUF - N + i] = M32Q15(g, buf[DBUF - N + i]); }
            for (i = OVL; i < elen; i++) buf[DBUF - N + i] = M32Q15(ratio, buf[DBUF - N + i]);
         }
         comb_filter(etmp, buf + DBUF, st->pf_period, st->pf_period, OVL, (s16)(-st->pf_gain), (s16)(-st->pf_gain), st->pf_tapset, st->pf_tapset, 0);
         for (i = 0; i < 60; i++) buf[DBUF + i] = ADD(M32Q15(ot_window120[i], etmp[119 - i]), M32Q15(ot_window120[119 - i], etmp[i]));
      }
   }
   st->loss_count = loss_count + 1;
}

/* ---------- one CELT frame: celt_decode_with_ec, celt_decoder.c:713-1072 ---------- */
int oracle_celt_decode(struct OracleCelt *st, const u8 *data, int len, s16 *pcm, int frame_size48, int C, int accum, u32 *final_range) {
   static const u8 trim_icdf[11] = {126, 124, 119, 109, 87, 41, 19, 9, 4, 2, 0}, spread_icdf[4] = {25, 23, 2, 0}, tapset_icdf[3] = {2, 1, 0}, small_icdf[3] = {2, 1, 0};
   static const int8_t tf_select[4][8] = {{0, -1, 0, -1, 0, -1, 0, -1}, {0, -1, 0, -2, 1, 0, 1, -1}, {0, -2, 0, -3, 2, 0, 1, -1}, {0, -2, 0, -3, 3, 0, 1, -1}};
   static const s16 pred_coef[4] = {29440, 26112, 21248, 16384}, beta_coef[4] = {30147, 22282, 12124, 6554};
   RDec d; int CC = st->channels, LM, M, N, start = st->start, end = st->end, effEnd, c, i, silence, transient = 0, intra, spread = 2, trim, coded;
   int intensity = 0, dual = 0, pf_pitch = 0, pf_tapset = 0, ac_rsv, ac_on = 0, dyn_logp = 6; s32 pf_gain = 0, total_bits, tell, bits, balance;
   int tf_res[NB], cap[NB], offsets[NB], pulses[NB], fine_q[NB], fine_p[NB]; u8 masks[2 * NB]; s16 X[1920]; s32 *out_syn[2];
   for (LM = 0; LM <= 3; LM++) if (120 << LM == frame_size48) break;
   if (LM > 3) return -1;
   if (len < 0 || len > 1275 || pcm == NULL) return -1;
   M = 1 << LM; N = M * 120;
   for (c = 0; c < CC; c++) out_syn[c] = st->mem[c] + DBUF - N;
   effEnd = imin(end, NB);
   if (data == NULL || len <= 1) { decode_lost(st, N, LM); deemphasis(st, out_syn, pcm, N, accum); if (final_range) *final_range = 0; return frame_size48 / st->downsample; }
   rd_init(&d, data, (u32)len);
   if (C == 1) for (i = 0; i < NB; i++) st->oldE[i] = (s16)imax(st->oldE[i], st->oldE[NB + i]);
   total_bits = len * 8; tell = rd_tell(&d);
   silence = tell >= total_bits ? 1 : tell == 1 ? rd_bit_logp(&d, 15) : 0;
   if (silence) { tell = len * 8; d.nbits_total += tell - rd_tell(&d); }
   if (start == 0 && tell + 16 <= total_bits) {
      if (rd_bit_logp(&d, 1)) {
         int octave = (int)rd_uint(&d, 6), qg;
         pf_pitch = (16 << octave) + (int)rd_bits(&d, 4 + (u32)octave) - 1;
         qg = (int)rd_bits(&d, 3);
         if (rd_tell(&d) + 2 <= total_bits) pf_tapset = rd_icdf(&d, tapset_icdf, 2);
         pf_gain = 3072 * (qg + 1);
      }
      tell = rd_tell(&d);
   }
   if (LM > 0 && tell + 3 <= total_bits) { transient = rd_bit_logp(&d, 3); tell = rd_tell(&d); }
   intra = tell + 3 <= total_bits ? rd_bit_logp(&d, 3) : 0;
   {  /* unquant_coarse_energy quant_bands.c:435-498 */
      const u8 *pm = ot_e_prob_model + (LM * 2 + intra) * 42; s32 prev[2] = {0, 0}, coef = intra ? 0 : pred_coef[LM], beta = intra ? 4915 : beta_coef[LM], budget = len * 8;
      for (i = start; i < end; i++) for (c = 0; c < C; c++) {
         int qi; s32 q, tmp; s16 *e = &st->oldE[i + c * NB];
         tell = rd_tell(&d);
         if (budget - tell >= 15) { int pi = 2 * imin(i, 20); qi = rd_laplace(&d, (u32)pm[pi] << 7, (int)pm[pi + 1] << 6); }
         else if (budget - tell >= 2) { qi = rd_icdf(&d, small_icdf, 2); qi = (qi >> 1) ^ -(qi & 1); }
         else if (budget - tell >= 1) qi = -rd_bit_logp(&d, 1);
         else qi = -1;
         q = SHL(qi, 10);
         *e = (s16)imax(-9216, *e);
         tmp = ADD(ADD(PSHR(M16(coef, *e), 8), prev[c]), SHL(q, 7));
         tmp = imax(-3670016, tmp);
         *e = (s16)PSHR(tmp, 7);
         prev[c] = SUB(ADD(prev[c], SHL(q, 7)), M16(beta, PSHR(q, 8)));
      }
   }
   {  /* tf_decode celt_decoder.c:352-389 */
      u32 budget = (u32)len * 8, t = (u32)rd_tell(&d); int logp = transient ? 2 : 4, rsv = LM > 0 && t + logp + 1 <= budget, changed = 0, curr = 0, sel = 0;
      budget -= rsv;
      for (i = start; i < end; i++) {
         if (t + logp <= budget) { curr ^= rd_bit_logp(&d, (u32)logp); t = (u32)rd_tell(&d); changed |= curr; }
         tf_res[i] = curr; logp = transient ? 4 : 5;
      }
      if (rsv && tf_select[LM][4 * transient + changed] != tf_select[LM][4 * transient + 2 + changed]) sel = rd_bit_logp(&d, 1);
      for (i = start; i < end; i++) tf_res[i] = tf_select[LM][4 * transient + 2 * sel + tf_res[i]];
   }
   tell = rd_tell(&d);
   if (tell + 4 <= total_bits) spread = rd_icdf(&d, spread_icdf, 5);
   for (i = 0; i < NB; i++) cap[i] = (ot_cache_caps50[NB * (2 * LM + C - 1) + i] + 64) * C * ((EB(i + 1) - EB(i)) << LM) >> 2;   /* init_caps celt.c:255 */
   total_bits <<= 3; tell = (s32)rd_tell_frac(&d);
   for (i = start; i < end; i++) {                                                                  /* dynalloc :917-943 */
      int width = C * (EB(i + 1) - EB(i)) << LM, quanta = imin(width << 3, imax(48, width)), lp = dyn_logp, boost = 0;
      while (tell + (lp << 3) < total_bits && boost < cap[i]) {
         int flag = rd_bit_logp(&d, (u32)lp);
         tell = (s32)rd_tell_frac(&d);
         if (!flag) break;
         boost += quanta; total_bits -= quanta; lp = 1;
      }
      offsets[i] = boost;
      if (boost > 0) dyn_logp = imax(2, dyn_logp - 1);
   }
   trim = tell + 48 <= total_bits ? rd_icdf(&d, trim_icdf, 7) : 5;
   bits = ((s32)len * 8 << 3) - (s32)rd_tell_frac(&d) - 1;
   ac_rsv = transient && LM >= 2 && bits >= ((LM + 2) << 3) ? 8 : 0;
   bits -= ac_rsv;
   coded = allocation(&d, start, end, offsets, cap, trim, &intensity, &dual, bits, &balance, pulses, fine_q, fine_p, C, LM);
   for (i = start; i < end; i++) if (fine_q[i] > 0) for (c = 0; c < C; c++) {                      /* unquant_fine_energy quant_bands.c:500 */
      s32 q2 = (s32)rd_bits(&d, (u32)fine_q[i]);
      st->oldE[i + c * NB] = (s16)(st->oldE[i + c * NB] + (s16)S16(ADD(SHL(q2, 10), 512) >> fine_q[i], 512));
   }
   for (c = 0; c < CC; c++) memmove(st->mem[c], st->mem[c] + N, (DBUF - N + 60) * sizeof(s32));
   all_bands(&d, start, end, X, C == 2 ? X + N : NULL, masks, pulses, transient ? M : 0, spread, dual, intensity, tf_res,
             len * 64 - ac_rsv, balance, LM, coded, &st->rng);
   if (ac_rsv > 0) ac_on = (int)rd_bits(&d, 1);
   {  /* unquant_energy_finalise quant_bands.c:523-549 */
      int left = len * 8 - rd_tell(&d), prio;
      for (prio = 0; prio < 2; prio++) for (i = start; i < end && left >= C; i++) {
         if (fine_q[i] >= 8 || fine_p[i] != prio) continue;
         for (c = 0; c < C; c++) { s32 q2 = (s32)rd_bits(&d, 1); st->oldE[i + c * NB] = (s16)(st->oldE[i + c * NB] + (s16)(((s32)SL16(q2, 10) - 512) >> (fine_q[i] + 1))); left--; }
      }
   }
   if (ac_on) anti_collapse(X, masks, LM, C, N, start, end, st->oldE, st->oldLogE, st->oldLogE2, pulses, st->rng);
   if (silence) for (i = 0; i < C * NB; i++) st->oldE[i] = -28672;
   synthesis(st, X, out_syn, st->oldE, start, effEnd, C, CC, transient, LM, silence);
   for (c = 0; c < CC; c++) {
      st->pf_period = imax(st->pf_period, 15); st->pf_period_old = imax(st->pf_period_old, 15);
      comb_filter(out_syn[c], out_syn[c], st->pf_period_old, st->pf_period, 120, st->pf_gain_old, st->pf_gain, st->pf_tapset_old, st->pf_tapset, OVL);
      if (LM != 0) comb_filter(out_syn[c] + 120, out_syn[c] + 120, st->pf_period, pf_pitch, N - 120, st->pf_gain, pf_gain, st->pf_tapset, pf_tapset, OVL);
   }
   st->pf_period_old = st->pf_period; st->pf_gain_old = st->pf_gain; st->pf_tapset_old = st->pf_tapset;
   st->pf_period = pf_pitch; st->pf_gain = (s16)pf_gain; st->pf_tapset = pf_tapset;
   if (LM != 0) { st->pf_period_old = st->pf_period; st->pf_gain_old = st->pf_gain; st->pf_tapset_old = st->pf_tapset; }
   if (C == 1) memcpy(&st->oldE[NB], st->oldE, NB * sizeof(s16));
   if (!transient) {
      s32 inc = st->loss_count < 10 ? M : 1024;
      memcpy(st->oldLogE2, st->oldLogE, sizeof(st->oldLogE)); memcpy(st->oldLogE, st->oldE, sizeof(st->oldE));
      for (i = 0; i < 2 * NB; i++) st->bgLogE[i] = (s16)imin(st->bgLogE[i] + inc, st->oldE[i]);
   } else for (i = 0; i < 2 * NB; i++) st->oldLogE[i] = (s16)imin(st->oldLogE[i], st->oldE[i]);
   for (c = 0; c < 2; c++) for (i = 0; i < NB; i++) if (i < start || i >= end) { st->oldE[c * NB + i] = 0; st->oldLogE[c * NB + i] = st->oldLogE2[c * NB + i] = -28672; }
   st->rng = d.rng;
   deemphasis(st, out_syn, pcm, N, accum);
   st->loss_count = 0;
   if (final_range) *final_range = d.rng;
   if (rd_tell(&d) > 8 * len) return -3;
   if (d.error) st->error = 1;
   return frame_size48 / st->downsample;
}

/* decoder gain: celt_exp2(MULT16_16_P15(QCONST16(6.48814081e-4f,25), gain)) (opus_decoder.c:570) */
int32_t oracle_gain_q16(int decode_gain) { return celt_exp2(MP15(21771, decode_gain)); }

struct OracleCelt *oracle_celt_create(int channels, int downsample) {
   struct OracleCelt *st = (struct OracleCelt *)calloc(1, sizeof(*st));
   st->channels = channels; st->downsample = downsample;
   oracle_celt_reset(st);
   return st;
}
void oracle_celt_set_bands(struct OracleCelt *st, int start, int end) { st->start = start; st->end = end; }
int oracle_celt_pitch(const struct OracleCelt *st) { return st->pf_period; }
void oracle_celt_destroy(struct OracleCelt *st) { free(st); }
