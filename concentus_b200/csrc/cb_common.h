// cb_common.h -- execution-model shim + fixed-point primitives for the batched Opus engine.
//
// Execution model.  Every codec routine in csrc/ is written for ONE WARP PER STREAM:
//   * "uniform" code: all 32 lanes execute the same scalar control flow on identical values
//     (range decoder, bit allocation, band recursion).  No divergence, no synchronisation.
//   * "lane loops": CB_LANE_LOOP(i,n) distributes independent iterations over the lanes;
//     cb_wsum/cb_wor/cb_wmax fold per-lane partials; CB_WSYNC() orders shared-memory traffic.
// The same sources also compile with scalar semantics, where a lane loop is an ordinary loop and the
// folds are identities:
//   * CB_THREAD_MODE (device code): ONE THREAD PER STREAM.  The parse kernel (cb_parse_kernels.cu) builds
//     the bitstream-only half of the CELT decoder this way -- that half has no width, so running it
//     warp-uniform computed every scalar 32 times over.
//   * CB_HOST_EMU (plain C++): exists only for the CPU-side unit tests under tests/emu (host-logic
//     checks without a GPU); it is never linked into the shipped library.
//
// Fixed-point semantics follow libopus 1.1.2 FIXED_POINT (reference: opus-fix/celt/fixed_generic.h
// and celt/arch.h); each helper cites the macro it stands for.  All adds wrap mod 2^32.
#pragma once
#include <stdint.h>

typedef int16_t i16;
typedef uint16_t u16;
typedef int32_t i32;
typedef uint32_t u32;
typedef int64_t i64;
typedef uint64_t u64;
typedef uint8_t u8;
typedef int8_t i8;

#if defined(__CUDACC__) && !defined(CB_HOST_EMU) && !defined(CB_THREAD_MODE)
#define CB_DEVICE_BUILD 1
#define CB_FN __device__ __forceinline__
#define CB_MFN __device__ __forceinline__   /* member functions */
#define CB_HD __host__ __device__ __forceinline__
#define CB_FN_NOINLINE static __device__ __noinline__
#define CB_UTABLE static __constant__ const
#define CB_GTABLE static __device__ const
#define CB_LANE() ((int)(threadIdx.x & 31))
#define CB_LANE_LOOP(i, n) _Pragma("unroll 1") for (int i = CB_LANE(); i < (n); i += 32)
#define CB_LANE_LOOPU(i, n) for (int i = CB_LANE(); i < (n); i += 32)   /* compiler may unroll */
#define CB_WSYNC() __syncwarp()
#define CB_NOUNROLL _Pragma("unroll 1")
CB_FN i32 cb_wsum(i32 v) { return (i32)__reduce_add_sync(0xffffffffu, (unsigned)v); }
CB_FN u32 cb_wor(u32 v) { return __reduce_or_sync(0xffffffffu, v); }
CB_FN i32 cb_wmax(i32 v) { return __reduce_max_sync(0xffffffffu, v); }
CB_FN i32 cb_wmin(i32 v) { return __reduce_min_sync(0xffffffffu, v); }
CB_FN int cb_clz(u32 v) { return __clz((int)v); }
template <typename T> CB_FN T cb_ldg(const T* p) { return __ldg(p); }
// arg-max of a ratio num/den (num >= 0, den > 0; an empty lane carries num < 0) over the warp, ties
// to the lowest id (PVQ greedy search).  cb_wbest_exact is the plain butterfly on the exact
// cross-multiplied comparison.  cb_wbest first ranks lanes by a float quotient, which is only used
// to discard lanes that cannot win: fast division is good to 2 ulp, so every lane whose true ratio
// equals the maximum lies within 2^-19 of the largest quotient.  One survivor wins outright;
// survivors with identical (num, den) are resolved by the smallest id; anything else goes through
// the exact butterfly.
struct CbBest { i32 num, den, id; };
CB_FN CbBest cb_wbest_exact(CbBest b) {
  for (int o = 16; o > 0; o >>= 1) {
    CbBest t;
    t.num = __shfl_xor_sync(0xffffffffu, b.num, o);
    t.den = __shfl_xor_sync(0xffffffffu, b.den, o);
    t.id = __shfl_xor_sync(0xffffffffu, b.id, o);
    i32 l = b.den * t.num, r = t.den * b.num;      // t better than b  <=>  b.den*t.num > t.den*b.num
    if (l > r || (l == r && t.id < b.id)) b = t;
  }
  return b;
}
CB_FN CbBest cb_wbest(CbBest b) {
  const bool have = b.num >= 0 && b.den > 0;
  const float key = have ? __fdividef((float)b.num, (float)b.den) : -1.0f;
  const int ikey = have ? __float_as_int(key) : -1;
  const int mx = __reduce_max_sync(0xffffffffu, ikey);
  if (mx > 0) {
    const float thr = __int_as_float(mx) * (1.0f - 1.9073486e-6f);
    const unsigned cand = __ballot_sync(0xffffffffu, have && key >= thr);
    const int first = __ffs((int)cand) - 1;
    const i32 n0 = __shfl_sync(0xffffffffu, b.num, first), d0 = __shfl_sync(0xffffffffu, b.den, first);
    const unsigned same = __ballot_sync(0xffffffffu, b.num == n0 && b.den == d0);
    if ((cand & ~same) == 0) {                     // all survivors are the same fraction
      CbBest r;
      r.num = n0; r.den = d0;
      r.id = __reduce_min_sync(0xffffffffu, ((cand >> CB_LANE()) & 1u) ? b.id : 0x7fffffff);
      return r;
    }
    if (!((cand >> CB_LANE()) & 1u)) { b.num = -32767; b.den = 0; b.id = 0x7fffffff; }
  }
  return cb_wbest_exact(b);
}
#else
#define CB_DEVICE_BUILD 0
#if defined(__CUDACC__) && defined(CB_THREAD_MODE) && !defined(CB_HOST_EMU)
// one thread per stream: device code, scalar semantics; every table is indexed per thread (global, L1-cached)
#define CB_FN __device__ __forceinline__
#define CB_MFN __device__ __forceinline__
#define CB_HD __host__ __device__ __forceinline__
#ifdef CB_THREAD_NOINLINE
#define CB_FN_NOINLINE static __device__ __noinline__
#else
#define CB_FN_NOINLINE static __device__ __forceinline__   /* a lone thread wants its coder state in registers */
#endif
#ifdef CB_THREAD_GLOBAL_TABLES
#define CB_UTABLE static __device__ const
#else
#define CB_UTABLE static __constant__ const   /* few streams to a warp: few distinct addresses per read */
#endif
#define CB_GTABLE static __device__ const
#else
#define CB_FN static inline
#define CB_MFN inline
#define CB_HD static inline
#define CB_FN_NOINLINE static
#define CB_UTABLE static const
#define CB_GTABLE static const
#endif
#define CB_LANE() 0
#define CB_LANE_LOOP(i, n) for (int i = 0; i < (n); ++i)
#define CB_LANE_LOOPU(i, n) for (int i = 0; i < (n); ++i)
#define CB_WSYNC() ((void)0)
#define CB_NOUNROLL
CB_FN i32 cb_wsum(i32 v) { return v; }
CB_FN u32 cb_wor(u32 v) { return v; }
CB_FN i32 cb_wmax(i32 v) { return v; }
CB_FN i32 cb_wmin(i32 v) { return v; }
#if defined(__CUDA_ARCH__)
CB_FN int cb_clz(u32 v) { return __clz((int)v); }
template <typename T> CB_FN T cb_ldg(const T* p) { return __ldg(p); }
#else
CB_FN int cb_clz(u32 v) { return v ? __builtin_clz(v) : 32; }
template <typename T> CB_FN T cb_ldg(const T* p) { return *p; }
#endif
struct CbBest { i32 num, den, id; };
CB_FN CbBest cb_wbest(CbBest b) { return b; }
#endif

// Per-channel serial work: lane c runs channel c on the device; the host walks the channels.
// cb_bcast(v, lane) hands a lane's result to the whole warp.
#if CB_DEVICE_BUILD
#define CB_CHANNEL_LANES(c, C) for (int c = CB_LANE(), once_ = 1; once_ && c < (C); once_ = 0)
CB_FN i32 cb_bcast(i32 v, int lane) { return __shfl_sync(0xffffffffu, v, lane); }
CB_FN u32 cb_wsumu(u32 v) { return __reduce_add_sync(0xffffffffu, v); }
// exclusive prefix sum over lanes (lane 0 gets 0)
CB_FN i32 cb_wscan_excl(i32 v) {
  i32 x = v;
  for (int o = 1; o < 32; o <<= 1) {
    i32 t = __shfl_up_sync(0xffffffffu, x, o);
    if (CB_LANE() >= o) x += t;
  }
  return x - v;
}
#else
#define CB_CHANNEL_LANES(c, C) for (int c = 0; c < (C); c++)
CB_FN i32 cb_bcast(i32 v, int) { return v; }
CB_FN u32 cb_wsumu(u32 v) { return v; }
CB_FN i32 cb_wscan_excl(i32) { return 0; }
#endif

// serial work of one stream: lane 0 on the device, the only "lane" on the host
#if CB_DEVICE_BUILD
#define CB_IF_LANE0 if (CB_LANE() == 0)
#else
#define CB_IF_LANE0 if (1)
#endif

// ---- integer helpers -------------------------------------------------------------------
CB_HD int cb_imin(int a, int b) { return a < b ? a : b; }
CB_HD int cb_imax(int a, int b) { return a > b ? a : b; }
// EC_ILOG (celt/ecintrin.h:69-82): position of the highest set bit, 1-based; 0 for 0.
CB_FN int cb_ilog(u32 v) { return 32 - cb_clz(v); }
// celt_ilog2 / celt_zlog2 (celt/mathops.h:148-160)
CB_FN int cb_ilog2(i32 x) { return cb_ilog((u32)x) - 1; }
CB_FN int cb_zlog2(i32 x) { return x <= 0 ? 0 : cb_ilog2(x); }

// ---- fixed_generic.h equivalents -------------------------------------------------------
CB_FN i32 add32(i32 a, i32 b) { return (i32)((u32)a + (u32)b); }                 // ADD32 :101
CB_FN i32 sub32(i32 a, i32 b) { return (i32)((u32)a - (u32)b); }                 // SUB32 :103
CB_FN i32 neg32(i32 a) { return (i32)(0u - (u32)a); }
CB_FN i32 shl32(i32 a, int s) { return (i32)((u32)a << s); }                      // SHL32 :74
CB_FN i32 shr32(i32 a, int s) { return a >> s; }                                  // SHR32 :72
CB_FN i32 pshr32(i32 a, int s) { return add32(a, (i32)((1 << s) >> 1)) >> s; }    // PSHR32 :77
CB_FN i32 vshr32(i32 a, int s) { return s > 0 ? (a >> s) : shl32(a, -s); }        // VSHR32 :79
CB_FN i16 ext16(i32 a) { return (i16)a; }                                         // EXTRACT16 :63
CB_FN i16 round16(i32 a, int s) { return (i16)pshr32(a, s); }                     // ROUND16 :90 (wraps)
CB_FN i16 shl16(i32 a, int s) { return (i16)((uint16_t)a << s); }                 // SHL16 :70
CB_FN i16 add16(i32 a, i32 b) { return (i16)((i16)a + (i16)b); }                  // ADD16 :96
CB_FN i32 sub16(i32 a, i32 b) { return (i32)(i16)a - (i32)(i16)b; }               // SUB16 :98 (int result)
CB_FN i32 mult16_16(i32 a, i32 b) { return (i32)(i16)a * (i32)(i16)b; }           // MULT16_16 :109
CB_FN i32 mac16_16(i32 c, i32 a, i32 b) { return add32(c, mult16_16(a, b)); }     // MAC16_16 :112
CB_FN i32 mult16_16_q15(i32 a, i32 b) { return mult16_16(a, b) >> 15; }           // :126
CB_FN i32 mult16_16_p15(i32 a, i32 b) { return add32(16384, mult16_16(a, b)) >> 15; } // :130
CB_FN i32 mult16_16_q14(i32 a, i32 b) { return mult16_16(a, b) >> 14; }
CB_FN i32 mult16_16_q13(i32 a, i32 b) { return mult16_16(a, b) >> 13; }
CB_FN i32 mult16_16_q11(i32 a, i32 b) { return mult16_16(a, b) >> 11; }
// MULT16_32_Q15 :46 == low 32 bits of (a*b)>>15 with a taken as int16
CB_FN i32 mult16_32_q15(i32 a, i32 b) { return (i32)(((i64)(i16)a * (i64)b) >> 15); }
// MULT16_32_Q16 :40
CB_FN i32 mult16_32_q16(i32 a, i32 b) { return (i32)(((i64)(i16)a * (i64)b) >> 16); }
// MULT16_32_P16 :43 -- only the low partial product is rounded, which equals rounding the total
CB_FN i32 mult16_32_p16(i32 a, i32 b) { return (i32)(((i64)(i16)a * (i64)b + 32768) >> 16); }
// MULT32_32_Q31 :49 -- three partial products, lo*lo dropped (NOT a plain 64-bit product)
CB_FN i32 mult32_32_q31(i32 a, i32 b) {
  i32 ah = a >> 16, bh = b >> 16;
  i32 al = a & 0xffff, bl = b & 0xffff;
  i32 t0 = shl32(mult16_16(ah, bh), 1);
  i32 t1 = ((i32)(i16)ah * bl) >> 15;
  i32 t2 = ((i32)(i16)bh * al) >> 15;
  return add32(add32(t0, t1), t2);
}
// MAC16_32_Q15 :116 (splits b at bit 15)
CB_FN i32 mac16_32_q15(i32 c, i32 a, i32 b) {
  return add32(c, add32(mult16_16(a, b >> 15), mult16_16(a, b & 0x7fff) >> 15));
}
CB_FN i32 half32(i32 a) { return a >> 1; }
CB_FN i32 min32(i32 a, i32 b) { return a < b ? a : b; }
CB_FN i32 max32(i32 a, i32 b) { return a > b ? a : b; }
CB_FN i16 sat16(i32 x) { return (i16)(x > 32767 ? 32767 : x < -32768 ? -32768 : x); }  // arch.h:111
// SIG2WORD16 (fixed_generic.h:142-148): round Q12 -> int16 with saturation
CB_FN i16 sig2word16(i32 x) {
  x = pshr32(x, 12);
  x = max32(x, -32768);
  x = min32(x, 32767);
  return (i16)x;
}
// FRAC_MUL16 (celt/mathops.h:42)
CB_FN i32 frac_mul16(i32 a, i32 b) { return (16384 + (i32)(i16)a * (i32)(i16)b) >> 15; }

#define CB_Q15ONE 32767
#define CB_SIG_SHIFT 12
#define CB_DB_SHIFT 10
#define CB_BITRES 3
#define CB_NORM_SCALING 16384

// ---- celt/mathops.c ---------------------------------------------------------------------
// isqrt32 (mathops.c:42-63): exact floor(sqrt(v)).  On the device: a float estimate (good to +-1 over the whole
// 32-bit range) and an exact correction in 64-bit arithmetic; the reference's bit-by-bit loop is 16 dependent rounds.
#if defined(__CUDA_ARCH__)
CB_FN u32 cb_isqrt32(u32 v) {
  u32 g = __float2uint_rz(__fsqrt_rn(__uint2float_rn(v)));
  if ((u64)g * g > (u64)v) g--;
  if ((u64)(g + 1) * (g + 1) <= (u64)v) g++;
  return g;
}
#else
CB_FN_NOINLINE u32 cb_isqrt32(u32 v) {
  u32 g = 0;
  int bshift = (cb_ilog(v) - 1) >> 1;
  u32 b = 1u << bshift;
  CB_NOUNROLL do {
    u32 t = ((g << 1) + b) << bshift;
    if (t <= v) { g += b; v -= t; }
    b >>= 1;
    bshift--;
  } while (bshift >= 0);
  return g;
}
#endif
// celt_rsqrt_norm (mathops.c:90-115): Q16 in [0.25,1) -> Q14 reciprocal sqrt
CB_FN i16 cb_rsqrt_norm(i32 x) {
  i16 n = (i16)(x - 32768);
  i16 r = add16(23557, mult16_16_q15(n, add16(-13490, mult16_16_q15(n, 6713))));
  i16 r2 = (i16)mult16_16_q15(r, r);
  i16 y = shl16(sub16(add16(mult16_16_q15(r2, n), r2), 16384), 1);
  return add16(r, mult16_16_q15(r, mult16_16_q15(y, sub16(mult16_16_q15(y, 12288), 16384))));
}
// celt_sqrt (mathops.c:118-137)
CB_FN i32 cb_sqrt(i32 x) {
  if (x == 0) return 0;
  if (x >= 1073741824) return 32767;
  int k = (cb_ilog2(x) >> 1) - 7;
  x = vshr32(x, 2 * k);
  i16 n = (i16)(x - 32768);
  i32 rt = add16(23175, mult16_16_q15(n, add16(11561, mult16_16_q15(n, add16(-3011,
               mult16_16_q15(n, add16(1699, mult16_16_q15(n, -664))))))));
  return vshr32(rt, 7 - k);
}
// _celt_cos_pi_2 / celt_cos_norm (mathops.c:139-178)
CB_FN i16 cb_cos_pi_2(i16 x) {
  i16 x2 = (i16)mult16_16_p15(x, x);
  i32 t = add32(sub16(32767, x2),
                mult16_16_p15(x2, add32(-7651, mult16_16_p15(x2, add32(8277, mult16_16_p15(-626, x2))))));
  i32 m = 32766 < t ? 32766 : t;  // MIN16(32766, .)
  return add16(1, m);
}
CB_FN i16 cb_cos_norm(i32 x) {
  x = x & 0x0001ffff;
  if (x > (1 << 16)) x = (1 << 17) - x;
  if (x & 0x00007fff) {
    if (x < (1 << 15)) return cb_cos_pi_2((i16)x);
    return (i16)(-(i32)cb_cos_pi_2((i16)(65536 - x)));
  }
  if (x & 0x0000ffff) return 0;
  if (x & 0x0001ffff) return -32767;
  return 32767;
}
// celt_rcp (mathops.c:180-206): Q15 in, Q16 out
CB_FN i32 cb_rcp(i32 x) {
  int i = cb_ilog2(x);
  i16 n = (i16)(vshr32(x, i - 15) - 32768);
  i16 r = add16(30840, mult16_16_q15(-15420, n));
  r = (i16)sub16(r, mult16_16_q15(r, add16(mult16_16_q15(r, n), add16(r, -32768))));
  r = (i16)sub16(r, add16(1, mult16_16_q15(r, add16(mult16_16_q15(r, n), add16(r, -32768)))));
  return vshr32((i32)r, i - 16);
}
// celt_div (mathops.h:222)
CB_FN i32 cb_div(i32 a, i32 b) { return mult32_32_q31(a, cb_rcp(b)); }
// frac_div32 (mathops.c:69-87)
CB_FN i32 cb_frac_div32(i32 a, i32 b) {
  int shift = cb_ilog2(b) - 29;
  a = vshr32(a, shift);
  b = vshr32(b, shift);
  i16 rcp = round16(cb_rcp(round16(b, 16)), 3);
  i32 result = mult16_32_q15(rcp, a);
  i32 rem = sub32(pshr32(a, 2), mult32_32_q31(result, b));
  result = add32(result, shl32(mult16_32_q15(rcp, rem), 2));
  if (result >= 536870912) return 2147483647;
  if (result <= -536870912) return -2147483647;
  return shl32(result, 2);
}
// celt_exp2_frac / celt_exp2 (mathops.h:195-213)
CB_FN i32 cb_exp2_frac(i32 x) {
  i16 frac = shl16(x, 4);
  return add16(16383, mult16_16_q15(frac, add16(22804, mult16_16_q15(frac, add16(14819, mult16_16_q15(10204, frac))))));
}
CB_FN i32 cb_exp2(i32 x16) {
  i16 x = (i16)x16;
  int integer = x >> 10;
  if (integer > 14) return 0x7f000000;
  if (integer < -15) return 0;
  i32 frac = cb_exp2_frac((i16)(x - shl16(integer, 10)));
  return vshr32(frac, -integer - 2);
}
// celt_log2 (mathops.h:169-182): Q14 in, Q10 out
CB_FN i16 cb_log2(i32 x) {
  if (x == 0) return -32767;
  int i = cb_ilog2(x);
  i16 n = (i16)(vshr32(x, i - 15) - 32768 - 16384);
  i16 frac = add16(-6801 + (1 << (13 - CB_DB_SHIFT)),
                   mult16_16_q15(n, add16(15746, mult16_16_q15(n, add16(-5217, mult16_16_q15(n, add16(2545, mult16_16_q15(n, -1401))))))));
  return (i16)(shl16(i - 13, CB_DB_SHIFT) + (frac >> (14 - CB_DB_SHIFT)));
}
// celt_lcg_rand (bands.c:63) and an O(log j) skip-ahead so lanes can draw element j directly.
CB_FN u32 cb_lcg(u32 s) { return 1664525u * s + 1013904223u; }
CB_FN_NOINLINE u32 cb_lcg_skip(u32 s, u32 steps) {
  u32 a = 1664525u, c = 1013904223u;   // x -> a*x + c, composed by repeated squaring
  u32 A = 1u, C = 0u;
  CB_NOUNROLL while (steps) {
    if (steps & 1u) { A = A * a; C = C * a + c; }
    c = (a + 1u) * c;
    a = a * a;
    steps >>= 1;
  }
  return A * s + C;
}
