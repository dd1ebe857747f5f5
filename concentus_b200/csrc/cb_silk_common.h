// cb_silk_common.h -- SILK fixed-point primitives and the SILK decoder state.
//
// One SILK stream is decoded by ONE THREAD (32 streams per warp): SILK is a chain of per-sample
// recurrences (LTP + LPC synthesis, all-pass resampler) with no useful width inside a stream, while a
// batch has thousands of streams.  Control flow is mostly stream-independent (same frame length,
// same filter orders), so lanes stay together except inside the entropy decoder.  Per-thread
// work arrays are local memory, which the hardware interleaves across lanes (coalesced for equal
// indices).
//
// The primitives restate silk/macros.h, SigProc_FIX.h and Inlines.h of the reference
// (opus-fix/silk) in their 64-bit forms (OPUS_FAST_INT64, as the reference compiles on x86-64).
#pragma once
#include "cb_common.h"
#include "silk_tables.h"

#define SK_MAX_FRAME 320          // MAX_FRAME_LENGTH (20 ms at 16 kHz)
#define SK_MAX_SUBFR 80           // MAX_SUB_FRAME_LENGTH
#define SK_MAX_ORDER 16           // MAX_LPC_ORDER
#define SK_LTP_ORDER 5
#define SK_MAX_NB_SUBFR 4
#define SK_TYPE_NO_VOICE 0
#define SK_TYPE_UNVOICED 1
#define SK_TYPE_VOICED 2
#define SK_CODE_INDEPENDENTLY 0
#define SK_CODE_INDEPENDENTLY_NO_LTP_SCALING 1
#define SK_CODE_CONDITIONALLY 2
#define SK_FLAG_DECODE_NORMAL 0
#define SK_FLAG_PACKET_LOST 1
#define SK_FLAG_DECODE_LBRR 2

// ---- macros.h / SigProc_FIX.h ------------------------------------------------------------------
#if CB_DEVICE_BUILD
// (a * (i16)b) >> 16 == high word of a * ((i16)b << 16): one IMAD.HI instead of a wide multiply and a funnel shift
CB_FN i32 sk_smulwb(i32 a, i32 b) { return __mulhi(a, (i32)((u32)b << 16)); }                // macros.h:53
#else
CB_FN i32 sk_smulwb(i32 a, i32 b) { return (i32)(((i64)a * (i64)(i16)b) >> 16); }          // macros.h:53
#endif
CB_FN i32 sk_smlawb(i32 a, i32 b, i32 c) { return (i32)((u32)a + (u32)sk_smulwb(b, c)); }  // macros.h:60
CB_FN i32 sk_smulbb(i32 a, i32 b) { return (i32)(i16)a * (i32)(i16)b; }                    // macros.h:76
CB_FN i32 sk_smlabb(i32 a, i32 b, i32 c) { return (i32)((u32)a + (u32)sk_smulbb(b, c)); }  // macros.h:79
CB_FN i32 sk_smulww(i32 a, i32 b) { return (i32)(((i64)a * (i64)b) >> 16); }               // macros.h:92
CB_FN i32 sk_smlaww(i32 a, i32 b, i32 c) { return (i32)((u32)a + (u32)sk_smulww(b, c)); }  // macros.h:99
CB_FN i32 sk_smultt(i32 a, i32 b) { return (a >> 16) * (b >> 16); }                        // SigProc_FIX.h:428
CB_FN i32 sk_smmul(i32 a, i32 b) { return (i32)(((i64)a * (i64)b) >> 32); }                // SigProc_FIX.h:598
CB_FN i32 sk_mul(i32 a, i32 b) { return (i32)((u32)a * (u32)b); }
CB_FN i32 sk_add(i32 a, i32 b) { return (i32)((u32)a + (u32)b); }
CB_FN i32 sk_sub(i32 a, i32 b) { return (i32)((u32)a - (u32)b); }
CB_FN i32 sk_lshift(i32 a, int s) { return (i32)((u32)a << s); }
CB_FN i32 sk_rshift_round(i32 a, int s) { return s == 1 ? (a >> 1) + (a & 1) : ((a >> (s - 1)) + 1) >> 1; }   // :521
CB_FN i64 sk_rshift_round64(i64 a, int s) { return s == 1 ? (a >> 1) + (a & 1) : ((a >> (s - 1)) + 1) >> 1; }
CB_FN i32 sk_sat16(i32 a) { return a > 32767 ? 32767 : (a < -32768 ? -32768 : a); }        // :463
CB_FN i32 sk_limit(i32 a, i32 l1, i32 l2) {                                                 // silk_LIMIT :571
  return l1 > l2 ? (a > l1 ? l1 : (a < l2 ? l2 : a)) : (a > l2 ? l2 : (a < l1 ? l1 : a));
}
CB_FN i32 sk_min(i32 a, i32 b) { return a < b ? a : b; }
CB_FN i32 sk_max(i32 a, i32 b) { return a > b ? a : b; }
CB_FN i32 sk_abs(i32 a) { return a > 0 ? a : (i32)(0u - (u32)a); }
CB_FN int sk_clz32(i32 x) { return x ? cb_clz((u32)x) : 32; }                               // macros.h:126
CB_FN i32 sk_ror32(i32 a, int rot) {                                                        // SigProc_FIX.h:398
  u32 x = (u32)a, r = (u32)rot, m = (u32)(-rot);
  if (rot == 0) return a;
  if (rot < 0) return (i32)((x << m) | (x >> (32 - m)));
  return (i32)((x << (32 - r)) | (x >> r));
}
CB_FN i32 sk_add_sat16(i32 a, i32 b) { return sk_sat16((i32)((u32)(i32)(i16)a + (u32)b)); } // silk_ADD_SAT16 :472
CB_FN i32 sk_lshift_sat32(i32 a, int s) {                                                   // :504
  return sk_lshift(sk_limit(a, (i32)0x80000000 >> s, 0x7fffffff >> s), s);
}
CB_FN i32 sk_rand(i32 seed) { return (i32)(907633515u + (u32)seed * 196314165u); }          // silk_RAND :589

// ---- Inlines.h ---------------------------------------------------------------------------------
CB_FN void sk_clz_frac(i32 in, i32* lz, i32* frac_Q7) {                                      // :56-66
  const int lzeros = sk_clz32(in);
  *lz = lzeros;
  *frac_Q7 = sk_ror32(in, 24 - lzeros) & 0x7f;
}
CB_FN i32 sk_sqrt_approx(i32 x) {                                                            // :71-94
  if (x <= 0) return 0;
  i32 lz, frac_Q7;
  sk_clz_frac(x, &lz, &frac_Q7);
  i32 y = (lz & 1) ? 32768 : 46214;
  y >>= (lz >> 1);
  return sk_smlawb(y, y, sk_smulbb(213, frac_Q7));
}
CB_FN i32 sk_div32_varQ(i32 a32, i32 b32, int Qres) {                                        // :97-141
  const int a_headrm = sk_clz32(sk_abs(a32)) - 1;
  i32 a32_nrm = sk_lshift(a32, a_headrm);
  const int b_headrm = sk_clz32(sk_abs(b32)) - 1;
  const i32 b32_nrm = sk_lshift(b32, b_headrm);
  const i32 b32_inv = (0x7fffffff >> 2) / (b32_nrm >> 16);
  i32 result = sk_smulwb(a32_nrm, b32_inv);
  a32_nrm = sk_sub(a32_nrm, sk_lshift(sk_smmul(b32_nrm, result), 3));
  result = sk_smlawb(result, a32_nrm, b32_inv);
  const int lshift = 29 + a_headrm - b_headrm - Qres;
  if (lshift < 0) return sk_lshift_sat32(result, -lshift);
  if (lshift < 32) return result >> lshift;
  return 0;
}
CB_FN i32 sk_inverse32_varQ(i32 b32, int Qres) {                                             // :143-182
  const int b_headrm = sk_clz32(sk_abs(b32)) - 1;
  const i32 b32_nrm = sk_lshift(b32, b_headrm);
  const i32 b32_inv = (0x7fffffff >> 2) / (b32_nrm >> 16);
  i32 result = sk_lshift(b32_inv, 16);
  const i32 err_Q32 = sk_lshift((1 << 29) - sk_smulwb(b32_nrm, b32_inv), 3);
  result = sk_smlaww(result, err_Q32, b32_inv);
  const int lshift = 61 - b_headrm - Qres;
  if (lshift <= 0) return sk_lshift_sat32(result, -lshift);
  if (lshift < 32) return result >> lshift;
  return 0;
}
// silk_lin2log / silk_log2lin (lin2log.c, log2lin.c)
CB_FN i32 sk_lin2log(i32 inLin) {
  i32 lz, frac_Q7;
  sk_clz_frac(inLin, &lz, &frac_Q7);
  return sk_lshift(31 - lz, 7) + sk_smlawb(frac_Q7, sk_mul(frac_Q7, 128 - frac_Q7), 179);
}
CB_FN i32 sk_log2lin(i32 inLog_Q7) {
  if (inLog_Q7 < 0) return 0;
  if (inLog_Q7 >= 3967) return 0x7fffffff;
  i32 out = sk_lshift(1, inLog_Q7 >> 7);
  const i32 frac_Q7 = inLog_Q7 & 0x7F;
  if (inLog_Q7 < 2048) out = sk_add(out, sk_mul(out, sk_smlawb(frac_Q7, sk_smulbb(frac_Q7, 128 - frac_Q7), -174)) >> 7);
  else out = sk_add(out, sk_mul(out >> 7, sk_smlawb(frac_Q7, sk_smulbb(frac_Q7, 128 - frac_Q7), -174)));
  return out;
}
// silk_sum_sqr_shift (sum_sqr_shift.c)
CB_FN void sk_sum_sqr_shift(i32* energy, int* shift, const i16* x, int len) {
  int i, shft = 0;
  i32 nrg = 0, nrg_tmp;
  len--;
  for (i = 0; i < len; i += 2) {
    nrg = sk_smlabb(nrg, x[i], x[i]);
    nrg = sk_smlabb(nrg, x[i + 1], x[i + 1]);
    if (nrg < 0) {
      nrg = (i32)((u32)nrg >> 2);
      shft = 2;
      i += 2;
      break;
    }
  }
  for (; i < len; i += 2) {
    nrg_tmp = sk_smulbb(x[i], x[i]);
    nrg_tmp = sk_smlabb(nrg_tmp, x[i + 1], x[i + 1]);
    nrg = (i32)((u32)nrg + ((u32)nrg_tmp >> shft));
    if (nrg < 0) {
      nrg = (i32)((u32)nrg >> 2);
      shft += 2;
    }
  }
  if (i == len) {
    nrg_tmp = sk_smulbb(x[i], x[i]);
    nrg = (i32)((u32)nrg + ((u32)nrg_tmp >> shft));
  }
  if (nrg & 0xC0000000) {
    nrg = (i32)((u32)nrg >> 2);
    shft += 2;
  }
  *shift = shft;
  *energy = nrg;
}

// ---- state (silk/structs.h:230-310, resampler_structs.h, dec_API.c:38-45) -----------------------
struct SilkResampler {            // silk_resampler_state_struct
  i32 sIIR[6];
  i32 sFIR[36];                   // union of i32[36] / i16[36]: the i16 view uses the low half-words' storage
  i16 delayBuf[48];
  i32 function, batchSize, invRatio_Q16, FIR_Order, FIR_Fracs, Fs_in_kHz, Fs_out_kHz, inputDelay;
  i32 coefs_id;                   // 0 none, 1: 3_4, 2: 2_3, 3: 1_2
};
struct SilkIndices {              // SideInfoIndices (structs.h:118-131)
  i8 GainsIndices[SK_MAX_NB_SUBFR];
  i8 LTPIndex[SK_MAX_NB_SUBFR];
  i8 NLSFIndices[SK_MAX_ORDER + 1];
  i16 lagIndex;
  i8 contourIndex, signalType, quantOffsetType, NLSFInterpCoef_Q2, PERIndex, LTP_scaleIndex, Seed;
};
struct SilkPLC {                  // silk_PLC_struct
  i32 pitchL_Q8;
  i16 LTPCoef_Q14[SK_LTP_ORDER];
  i16 prevLPC_Q12[SK_MAX_ORDER];
  i32 last_frame_lost, rand_seed;
  i16 randScale_Q14;
  i32 conc_energy, conc_energy_shift;
  i16 prevLTP_scale_Q14;
  i32 prevGain_Q16[2];
  i32 fs_kHz, nb_subfr, subfr_length;
};
// ---- bulk copies ----------------------------------------------------------------------------
// One thread owns a stream, so its state arrays are visited by one lane each: a warp's element-wise
// access touches 32 different sectors.  Bulk moves therefore go 16 bytes at a time when the
// destination is 16-byte aligned (state arrays and PCM rows are); 8x / 4x fewer memory instructions.
struct alignas(16) SkV4 { u32 x, y, z, w; };
// n 16-bit elements, ascending; dst may overlap src when it lies at least 8 elements below it
CB_FN void sk_copy_i16(i16* dst, const i16* src, int n) {
#if CB_DEVICE_BUILD
  if ((((size_t)dst & 15) | ((size_t)src & 3) | (size_t)(n & 7)) == 0) {
    const u32* s = (const u32*)src;
    SkV4* d = (SkV4*)dst;
    for (int i = 0; i < (n >> 3); i++) {
      SkV4 v;
      v.x = s[4 * i]; v.y = s[4 * i + 1]; v.z = s[4 * i + 2]; v.w = s[4 * i + 3];
      d[i] = v;
    }
    return;
  }
#endif
  for (int i = 0; i < n; i++) dst[i] = src[i];
}
// n 32-bit elements, ascending; dst may overlap src when it lies at least 4 elements below it
CB_FN void sk_copy_i32(i32* dst, const i32* src, int n) {
#if CB_DEVICE_BUILD
  if ((((size_t)dst & 15) | ((size_t)src & 15) | (size_t)(n & 3)) == 0) {
    const SkV4* s = (const SkV4*)src;
    SkV4* d = (SkV4*)dst;
    for (int i = 0; i < (n >> 2); i++) d[i] = s[i];
    return;
  }
#endif
  for (int i = 0; i < n; i++) dst[i] = src[i];
}
// n 32-bit elements, descending (memmove towards higher addresses; dst at least 4 elements above src)
CB_FN void sk_copy_i32_down(i32* dst, const i32* src, int n) {
#if CB_DEVICE_BUILD
  if ((((size_t)dst & 15) | ((size_t)src & 15) | (size_t)(n & 3)) == 0) {
    const SkV4* s = (const SkV4*)src;
    SkV4* d = (SkV4*)dst;
    for (int i = (n >> 2) - 1; i >= 0; i--) d[i] = s[i];
    return;
  }
#endif
  for (int i = n - 1; i >= 0; i--) dst[i] = src[i];
}

struct SilkCNG {                  // silk_CNG_struct
  alignas(16) i32 CNG_exc_buf_Q14[SK_MAX_FRAME];
  i16 CNG_smth_NLSF_Q15[SK_MAX_ORDER];
  i32 CNG_synth_state[SK_MAX_ORDER];
  i32 CNG_smth_Gain_Q16, rand_seed, fs_kHz;
  // CNG_exc_buf_Q14 is kept as a ring of four subframe slots while frames have four subframes: the reference's
  // "shift everything down one subframe" (CNG.c:122) is then a change of ring_head instead of a 1 KB move per
  // inactive frame.  Logical subframe k lives in slot (ring_head + k) & 3 of ring_sub samples; ring_head = 0 is
  // the reference's layout, and anything that does not fit the ring (two-subframe frames, a rate change) first
  // puts the buffer back in that layout (sk_cng_linearise).
  i32 ring_head, ring_sub;
};
struct alignas(16) SilkChannel {  // silk_decoder_state
  alignas(16) i32 exc_Q14[SK_MAX_FRAME];
  alignas(16) i32 sLPC_Q14_buf[SK_MAX_ORDER];
  alignas(16) i16 outBuf[SK_MAX_FRAME + 2 * SK_MAX_SUBFR];
  i32 prev_gain_Q16;
  i32 lagPrev;
  i32 LastGainIndex;              // opus_int8 in the reference
  i32 fs_kHz, fs_API_hz, nb_subfr, frame_length, subfr_length, ltp_mem_length, LPC_order;
  i16 prevNLSF_Q15[SK_MAX_ORDER];
  i32 first_frame_after_reset;
  i32 nFramesDecoded, nFramesPerPacket;
  i32 ec_prevSignalType;
  i32 ec_prevLagIndex;            // opus_int16 in the reference
  i32 VAD_flags[3], LBRR_flag, LBRR_flags[3];
  SilkResampler resampler;
  SilkIndices indices;
  SilkCNG sCNG;
  i32 lossCnt, prevSignalType;
  SilkPLC sPLC;
};
struct SilkStereo {               // stereo_dec_state
  i16 pred_prev_Q13[2];
  i16 sMid[2], sSide[2];
};
struct alignas(16) SilkDecState { // silk_decoder (dec_API.c:38-45) + silk_DecControlStruct fields that persist
  SilkChannel ch[2];
  SilkStereo sStereo;
  i32 nChannelsAPI, nChannelsInternal, prev_decode_only_middle;
  i32 ctl_nChannelsInternal, ctl_internalSampleRate, ctl_payloadSize_ms, ctl_prevPitchLag;
};
struct SilkDecControl {           // silk_decoder_control (per frame)
  i32 pitchL[SK_MAX_NB_SUBFR];
  i32 Gains_Q16[SK_MAX_NB_SUBFR];
  i16 PredCoef_Q12[2][SK_MAX_ORDER];
  i16 LTPCoef_Q14[SK_LTP_ORDER * SK_MAX_NB_SUBFR];
  i32 LTP_scale_Q14;
};

// silk_init_decoder (init_decoder.c:38-56) with silk_CNG_Reset (CNG.c:63-79) and silk_PLC_Reset (PLC.c:54-63)
// evaluated on the cleared state (LPC_order = frame_length = 0); usable from host and device.
CB_HD void sk_init_channel(SilkChannel& ps) {
  unsigned char* p = (unsigned char*)&ps;
  for (size_t i = 0; i < sizeof(SilkChannel); i++) p[i] = 0;
  ps.first_frame_after_reset = 1;
  ps.prev_gain_Q16 = 65536;
  ps.sCNG.CNG_smth_Gain_Q16 = 0;
  ps.sCNG.rand_seed = 3176576;
  ps.sPLC.pitchL_Q8 = 0;
  ps.sPLC.prevGain_Q16[0] = 65536;
  ps.sPLC.prevGain_Q16[1] = 65536;
  ps.sPLC.subfr_length = 20;
  ps.sPLC.nb_subfr = 2;
}
// silk_InitDecoder (dec_API.c:61-76)
CB_HD void sk_init_decoder(SilkDecState& d) {
  sk_init_channel(d.ch[0]);
  sk_init_channel(d.ch[1]);
  d.sStereo.pred_prev_Q13[0] = d.sStereo.pred_prev_Q13[1] = 0;
  d.sStereo.sMid[0] = d.sStereo.sMid[1] = 0;
  d.sStereo.sSide[0] = d.sStereo.sSide[1] = 0;
  d.prev_decode_only_middle = 0;
}
