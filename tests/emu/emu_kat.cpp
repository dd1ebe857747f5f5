// tests/emu/emu_kat.cpp -- TEST-ONLY host build (CB_HOST_EMU) of leaf routines of the kernel sources, one C entry per
// routine, for the component known-answer tests (tests/test_kat.py) that follow the reference's own unit tests
// (opus-fix/celt/tests/test_unit_{cwrs32,mathops,entropy,laplace,rotation}.c).  Never linked into libopus_b200.so.
#define CB_HOST_EMU 1
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include "../../concentus_b200/csrc/cb_celt_entropy.h"
#include "../../concentus_b200/csrc/cb_rangeenc.h"
#include "../../concentus_b200/csrc/cb_enc_state.h"
#include "../../concentus_b200/csrc/cb_enc_quant.h"

extern "C" {

// ---- cwrs.c ----
unsigned kat_pvq_v(int n, int k) { return pvq_u(n, k) + pvq_u(n, k + 1); }            // CELT_PVQ_V (cwrs.c:200)
int kat_cwrsi(int n, int k, unsigned idx, short* y) { return b3_cwrsi(y, n, k, idx); }   // returns yy
unsigned kat_icwrs(int n, const short* y) { return enc_icwrs(n, y); }
int kat_get_pulses(int i) { return get_pulses(i); }
// bits2pulses (rate.h:48-70): the reference's bisection and the table-driven form the scalar builds use
int kat_bits2pulses_bisect(int band, int LM, int bits) { return bits2pulses(band, LM, bits); }
int kat_bits2pulses_table(int band, int LM, int bits) { return bits2pulses_lanes(band, LM, bits); }
int kat_compute_qn(int N, int b, int offset, int pulse_cap, int stereo) { return compute_qn(N, b, offset, pulse_cap, stereo); }
int kat_split_thresh(int band, int LM) { return ct_split_thresh[(LM + 1) * CB_NBANDS + band]; }
int kat_split_thresh_ref(int band, int LM) { const u8* c = pulse_cache(band, LM); return c[c[0]] + 12; }

// ---- mathops.c / mathops.h ----
unsigned kat_isqrt32(unsigned v) { return cb_isqrt32(v); }
int kat_sqrt(int x) { return cb_sqrt(x); }
int kat_rcp(int x) { return cb_rcp(x); }
int kat_rsqrt_norm(int x) { return cb_rsqrt_norm(x); }
int kat_cos_norm(int x) { return cb_cos_norm(x); }
int kat_log2(int x) { return cb_log2(x); }
int kat_exp2(int x) { return cb_exp2(x); }
int kat_frac_div32(int a, int b) { return cb_frac_div32(a, b); }
int kat_bitexact_cos(int x) { return bitexact_cos((i16)x); }
int kat_bitexact_log2tan(int isin, int icos) { return bitexact_log2tan(isin, icos); }
unsigned kat_lcg_skip(unsigned s, unsigned n) { return cb_lcg_skip(s, n); }

// ---- entenc.c / entdec.c / laplace.c: a script of symbols through OUR encoder and OUR decoder ----
// ops[i]: 0 = uint (a = value, b = ft), 1 = raw bits (a = value, b = nbits), 2 = bit_logp (a = bit, b = logp),
//         3 = laplace (a = value, b = fs, c = decay), 4 = icdf (a = symbol, table {b, c, 0} with ftb = 8)
int kat_encode(const int* kind, const int* a, const unsigned* b, const int* c, int count, unsigned char* buf, int size,
               int* a_coded, unsigned* tell_frac, unsigned* final_rng) {
  RangeEnc e;
  re_init(e, buf, (u32)size);
  for (int i = 0; i < count; i++) {
    a_coded[i] = a[i];
    switch (kind[i]) {
      case 0: re_uint(e, (u32)a[i], b[i]); break;
      case 1: re_bits(e, (u32)a[i], b[i]); break;
      case 2: re_bit_logp(e, a[i], b[i]); break;
      case 3: { int v = a[i]; re_laplace(e, &v, b[i], c[i]); a_coded[i] = v; break; }
      default: { u8 t[3] = {(u8)b[i], (u8)c[i], 0}; re_icdf(e, a[i], t, 8); break; }
    }
    tell_frac[i] = re_tell_frac(e);
  }
  re_done(e);
  *final_rng = e.rng;
  return e.error ? -1 : (int)(e.offs + e.end_offs);
}
int kat_decode(const int* kind, const unsigned* b, const int* c, int count, const unsigned char* buf, int size, int* out,
               unsigned* tell_frac, unsigned* final_rng) {
  RangeDec d;
  rd_init(d, buf, (u32)size);
  for (int i = 0; i < count; i++) {
    switch (kind[i]) {
      case 0: out[i] = (int)rd_uint(d, b[i]); break;
      case 1: out[i] = (int)rd_bits(d, b[i]); break;
      case 2: out[i] = rd_bit_logp(d, b[i]); break;
      case 3: out[i] = rd_laplace(d, b[i], c[i]); break;
      default: { u8 t[3] = {(u8)b[i], (u8)c[i], 0}; out[i] = rd_icdf(d, t, 8); break; }
    }
    tell_frac[i] = rd_tell_frac(d);
  }
  *final_rng = d.rng;
  return d.error;
}

// ---- vq.c: alg_unquant after the index has been read = b3_leaf (cwrsi + normalise_residual + exp_rotation(-1)) ----
unsigned kat_alg_unquant(short* X, int N, int K, unsigned idx, int gain, int B, int spread) {
  return b3_leaf(X, N, K, idx, gain, B, spread);
}

}  // extern "C"
