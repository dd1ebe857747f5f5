// cb_rangeenc.h -- Opus range encoder as a warp-uniform state machine (all lanes hold the same
// coder state in shared memory and write the same bytes).  Bit-exact counterpart of
// opus-fix/celt/entenc.c (carry propagation, raw bits at the tail, shrink, patch, done) and the
// encoder half of celt/laplace.c.
#pragma once
#include "cb_common.h"

struct RangeEnc {
  u8* buf;            // output buffer
  u32 storage;        // bytes available
  u32 end_offs;       // raw-bit bytes written at the end
  u32 end_window;
  int nend_bits;
  int nbits_total;
  u32 offs;           // range bytes written at the front
  u32 rng, val, ext;
  int rem;
  int error;
};

CB_FN int re_tell(const RangeEnc& e) { return e.nbits_total - cb_ilog(e.rng); }
CB_UTABLE u32 re_tell_correction[8] = {35733, 38967, 42495, 46340, 50535, 55109, 60097, 65535};
CB_FN_NOINLINE u32 re_tell_frac(const RangeEnc& e) {          // entcode.c:69-97
  u32 nbits = (u32)e.nbits_total << 3;
  int l = cb_ilog(e.rng);
  u32 r = e.rng >> (l - 16);
  u32 b = (r >> 12) - 8;
  b += r > re_tell_correction[b];
  return nbits - (u32)((l << 3) + (int)b);
}
CB_FN int re_write_byte(RangeEnc& e, unsigned v) {            // entenc.c:61-67
  if (e.offs + e.end_offs >= e.storage) return -1;
  e.buf[e.offs++] = (u8)v;
  return 0;
}
CB_FN int re_write_byte_at_end(RangeEnc& e, unsigned v) {     // entenc.c:69-98
  if (e.offs + e.end_offs >= e.storage) return -1;
  e.buf[e.storage - ++e.end_offs] = (u8)v;
  return 0;
}
CB_FN void re_carry_out(RangeEnc& e, int c) {                 // entenc.c:111-128
  if (c != 255) {
    int carry = c >> 8;
    if (e.rem >= 0) e.error |= re_write_byte(e, (unsigned)(e.rem + carry));
    if (e.ext > 0) {
      unsigned sym = (255 + carry) & 255;
      CB_NOUNROLL do e.error |= re_write_byte(e, sym); while (--e.ext > 0);
    }
    e.rem = c & 255;
  } else {
    e.ext++;
  }
}
CB_FN void re_normalize(RangeEnc& e) {                        // entenc.c:130-137
  CB_NOUNROLL while (e.rng <= 0x800000u) {
    re_carry_out(e, (int)(e.val >> 23));
    e.val = (e.val << 8) & 0x7fffffffu;
    e.rng <<= 8;
    e.nbits_total += 8;
  }
}
CB_FN void re_init(RangeEnc& e, u8* buf, u32 size) {          // entenc.c:139-153
  e.buf = buf; e.end_offs = 0; e.end_window = 0; e.nend_bits = 0; e.nbits_total = 33;
  e.offs = 0; e.rng = 0x80000000u; e.rem = -1; e.val = 0; e.ext = 0; e.storage = size; e.error = 0;
}
CB_FN_NOINLINE void re_encode(RangeEnc& e, u32 fl, u32 fh, u32 ft) {      // entenc.c:155-164
  u32 r = e.rng / ft;
  if (fl > 0) {
    e.val += e.rng - r * (ft - fl);
    e.rng = r * (fh - fl);
  } else {
    e.rng -= r * (ft - fh);
  }
  re_normalize(e);
}
CB_FN_NOINLINE void re_encode_bin(RangeEnc& e, u32 fl, u32 fh, u32 bits) { // entenc.c:166-175
  u32 r = e.rng >> bits;
  if (fl > 0) {
    e.val += e.rng - r * ((1u << bits) - fl);
    e.rng = r * (fh - fl);
  } else {
    e.rng -= r * ((1u << bits) - fh);
  }
  re_normalize(e);
}
CB_FN_NOINLINE void re_bit_logp(RangeEnc& e, int val, u32 logp) {          // entenc.c:177-188
  u32 r = e.rng, l = e.val, s = r >> logp;
  r -= s;
  if (val) e.val = l + r;
  e.rng = val ? s : r;
  re_normalize(e);
}
CB_FN_NOINLINE void re_icdf(RangeEnc& e, int s, const u8* icdf, u32 ftb) { // entenc.c:190-202
  u32 r = e.rng >> ftb;
  if (s > 0) {
    e.val += e.rng - r * icdf[s - 1];
    e.rng = r * (u32)(icdf[s - 1] - icdf[s]);
  } else {
    e.rng -= r * icdf[s];
  }
  re_normalize(e);
}
CB_FN_NOINLINE void re_bits(RangeEnc& e, u32 fl, u32 bits) {               // entenc.c:222-241
  u32 window = e.end_window;
  int used = e.nend_bits;
  if (used + (int)bits > 32) {
    CB_NOUNROLL do {
      e.error |= re_write_byte_at_end(e, window & 255u);
      window >>= 8;
      used -= 8;
    } while (used >= 8);
  }
  window |= fl << used;
  used += (int)bits;
  e.end_window = window;
  e.nend_bits = used;
  e.nbits_total += (int)bits;
}
CB_FN_NOINLINE void re_uint(RangeEnc& e, u32 fl, u32 ft) {                 // entenc.c:204-220
  ft--;
  int ftb = cb_ilog(ft);
  if (ftb > 8) {
    ftb -= 8;
    u32 top = (ft >> ftb) + 1;
    u32 f = fl >> ftb;
    re_encode(e, f, f + 1, top);
    re_bits(e, fl & ((1u << ftb) - 1u), (u32)ftb);
  } else {
    re_encode(e, fl, fl + 1, ft + 1);
  }
}
// ec_enc_shrink (entenc.c:263-268): move the raw-bit tail to the new end
CB_FN_NOINLINE void re_shrink(RangeEnc& e, u32 size) {
  // memmove of end_offs bytes from old tail to new tail; size <= storage so the move is downward
  const u32 n = e.end_offs;
  u8* dst = e.buf + size - n;
  const u8* src = e.buf + e.storage - n;
  CB_NOUNROLL for (u32 i = 0; i < n; i++) dst[i] = src[i];
  e.storage = size;
}
// ec_enc_done (entenc.c:270-318)
CB_FN_NOINLINE void re_done(RangeEnc& e) {
  int l = 32 - cb_ilog(e.rng);
  u32 msk = 0x7fffffffu >> l;
  u32 end = (e.val + msk) & ~msk;
  if ((end | msk) >= e.val + e.rng) {
    l++;
    msk >>= 1;
    end = (e.val + msk) & ~msk;
  }
  CB_NOUNROLL while (l > 0) {
    re_carry_out(e, (int)(end >> 23));
    end = (end << 8) & 0x7fffffffu;
    l -= 8;
  }
  if (e.rem >= 0 || e.ext > 0) re_carry_out(e, 0);
  u32 window = e.end_window;
  int used = e.nend_bits;
  CB_NOUNROLL while (used >= 8) {
    e.error |= re_write_byte_at_end(e, window & 255u);
    window >>= 8;
    used -= 8;
  }
  if (!e.error) {
    const u32 n = e.storage - e.offs - e.end_offs;
    CB_NOUNROLL for (u32 i = 0; i < n; i++) e.buf[e.offs + i] = 0;
    if (used > 0) {
      if (e.end_offs >= e.storage) e.error = -1;
      else {
        l = -l;
        if (e.offs + e.end_offs >= e.storage && l < used) {
          window &= (1u << l) - 1;
          e.error = -1;
        }
        e.buf[e.storage - e.end_offs - 1] |= (u8)window;
      }
    }
  }
}
// ec_laplace_encode (laplace.c:51-92); *value may be clamped
CB_FN_NOINLINE void re_laplace(RangeEnc& e, int* value, u32 fs, int decay) {
  u32 fl = 0;
  int val = *value;
  if (val) {
    int s = -(val < 0);
    val = (val + s) ^ s;
    fl = fs;
    fs = ((32768u - 32u - fs) * (u32)(16384 - decay)) >> 15;     // ec_laplace_get_freq1
    int i;
    CB_NOUNROLL for (i = 1; fs > 0 && i < val; i++) {
      fs *= 2;
      fl += fs + 2;
      fs = (fs * (u32)decay) >> 15;
    }
    if (!fs) {
      int ndi_max = (int)((32768u - fl + 1 - 1) >> 0);
      ndi_max = (ndi_max - s) >> 1;
      int di = cb_imin(val - i, ndi_max - 1);
      fl += (u32)((2 * di + 1 + s) * 1);
      fs = (1u < 32768u - fl) ? 1u : 32768u - fl;
      *value = (i + di + s) ^ s;
    } else {
      fs += 1;
      fl += fs & (u32)~s;
    }
  }
  re_encode_bin(e, fl, fl + fs, 15);
}
