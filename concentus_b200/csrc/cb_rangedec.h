// cb_rangedec.h -- Opus range decoder as a warp-uniform state machine.
//
// All 32 lanes of the stream's warp hold identical copies of the coder registers and read the
// packet bytes (staged in shared memory) with broadcast loads, so there is no divergence and no
// cross-lane traffic.  Bit-exact counterpart of opus-fix/celt/entdec.c + entcode.c/h.
#pragma once
#include "cb_common.h"

struct RangeDec {
  const u8* buf;      // packet payload (shared memory on device)
  u32 storage;        // bytes available                         (ec_ctx.storage)
  u32 end_offs;       // raw-bit bytes consumed from the end     (ec_ctx.end_offs)
  u32 end_window;     // raw-bit window                          (ec_ctx.end_window)
  int nend_bits;      // valid bits in end_window                (ec_ctx.nend_bits)
  int nbits_total;    // (ec_ctx.nbits_total)
  u32 offs;           // next front byte                         (ec_ctx.offs)
  u32 rng, val, ext;  // range, difference, saved divisor        (ec_ctx.rng/val/ext)
  int rem;            // last byte read                          (ec_ctx.rem)
  int error;
};

// entdec.c:93-95 / :97-100 -- reads past the end yield zeros
CB_FN int rd_byte(RangeDec& d) { return d.offs < d.storage ? d.buf[d.offs++] : 0; }
CB_FN int rd_byte_end(RangeDec& d) { return d.end_offs < d.storage ? d.buf[d.storage - ++d.end_offs] : 0; }

// ec_dec_normalize (entdec.c:104-131): EC_SYM_BITS=8, EC_CODE_BOT=2^23, EC_CODE_EXTRA=7
CB_FN void rd_normalize(RangeDec& d) {
  CB_NOUNROLL while (d.rng <= 0x800000u) {
    d.nbits_total += 8;
    d.rng <<= 8;
    int sym = d.rem;
    d.rem = rd_byte(d);
    sym = ((sym << 8) | d.rem) >> 1;
    d.val = ((d.val << 8) + (255u & ~(u32)sym)) & 0x7fffffffu;
  }
}
// ec_dec_init (entdec.c:133-153)
CB_FN void rd_init(RangeDec& d, const u8* buf, u32 storage) {
  d.buf = buf;
  d.storage = storage;
  d.end_offs = 0;
  d.end_window = 0;
  d.nend_bits = 0;
  d.nbits_total = 32 + 1 - ((32 - 7) / 8) * 8;  // = 9
  d.offs = 0;
  d.rng = 1u << 7;
  d.rem = rd_byte(d);
  d.val = d.rng - 1 - (u32)(d.rem >> 1);
  d.error = 0;
  d.ext = 0;
  rd_normalize(d);
}
// ec_tell (entcode.h:114-120)
CB_FN int rd_tell(const RangeDec& d) { return d.nbits_total - cb_ilog(d.rng); }
// ec_tell_frac (entcode.c:69-97), 1/8-bit resolution
CB_UTABLE u32 rd_tell_correction[8] = {35733, 38967, 42495, 46340, 50535, 55109, 60097, 65535};
CB_FN_NOINLINE u32 rd_tell_frac(const RangeDec& d) {
  u32 nbits = (u32)d.nbits_total << 3;
  int l = cb_ilog(d.rng);
  u32 r = d.rng >> (l - 16);
  u32 b = (r >> 12) - 8;
  b += r > rd_tell_correction[b];
  return nbits - (u32)((l << 3) + (int)b);
}
// ec_decode (entdec.c:155-172): two unsigned divides
CB_FN_NOINLINE u32 rd_decode(RangeDec& d, u32 ft) {
  d.ext = d.rng / ft;
  u32 s = d.val / d.ext;
  return ft - (s + 1 < ft ? s + 1 : ft);
}
// ec_decode_bin (entdec.c:174-179)
CB_FN_NOINLINE u32 rd_decode_bin(RangeDec& d, u32 bits) {
  d.ext = d.rng >> bits;
  u32 s = d.val / d.ext;
  u32 top = 1u << bits;
  return top - (s + 1u < top ? s + 1u : top);
}
// ec_dec_update (entdec.c:181-200)
CB_FN_NOINLINE void rd_update(RangeDec& d, u32 fl, u32 fh, u32 ft) {
  u32 s = d.ext * (ft - fh);
  d.val -= s;
  d.rng = fl > 0 ? d.ext * (fh - fl) : d.rng - s;
  rd_normalize(d);
}
// ec_dec_bit_logp (entdec.c:203-216)
CB_FN_NOINLINE int rd_bit_logp(RangeDec& d, u32 logp) {
  u32 r = d.rng, v = d.val;
  u32 s = r >> logp;
  int ret = v < s;
  if (!ret) d.val = v - s;
  d.rng = ret ? s : r - s;
  rd_normalize(d);
  return ret;
}
// ec_dec_icdf (entdec.c:218-236); icdf lives in constant memory (uniform index)
CB_FN_NOINLINE int rd_icdf(RangeDec& d, const u8* icdf, u32 ftb) {
  u32 s = d.rng, v = d.val, t;
  u32 r = s >> ftb;
  int ret = -1;
  CB_NOUNROLL do {
    t = s;
    s = r * icdf[++ret];
  } while (v < s);
  d.val = v - s;
  d.rng = t - s;
  rd_normalize(d);
  return ret;
}
// ec_dec_icdf for per-lane tables in global memory (SILK): the linear search is a chain of dependent
// loads, so four entries are fetched at once (the tables are padded by 4 bytes, tools/gen_silk_tables.py).
// Inlined, like everything else the one-thread-per-stream SILK decoder calls, so that the coder state
// stays in registers from one symbol to the next.
CB_FN int rd_icdf_g(RangeDec& d, const u8* icdf, u32 ftb) {
  u32 s = d.rng, v = d.val, t;
  const u32 r = s >> ftb;
  int ret = 0;
  CB_NOUNROLL for (;;) {
    const u32 b0 = icdf[ret], b1 = icdf[ret + 1], b2 = icdf[ret + 2], b3 = icdf[ret + 3];
    t = s; s = r * b0; if (v >= s) break;
    ret++;
    t = s; s = r * b1; if (v >= s) break;
    ret++;
    t = s; s = r * b2; if (v >= s) break;
    ret++;
    t = s; s = r * b3; if (v >= s) break;
    ret++;
  }
  d.val = v - s;
  d.rng = t - s;
  rd_normalize(d);
  return ret;
}
// ec_dec_icdf with the two-entry table {p0, 0} (SILK pulse signs and LSBs): no table at all
CB_FN int rd_icdf2_g(RangeDec& d, u32 p0, u32 ftb) {
  const u32 rng = d.rng, v = d.val;
  const u32 s = (rng >> ftb) * p0;
  const int ret = v < s;
  d.val = ret ? v : v - s;
  d.rng = ret ? s : rng - s;
  rd_normalize(d);
  return ret;
}
// ec_dec_bit_logp, inlined form for the SILK decoder
CB_FN int rd_bit_logp_g(RangeDec& d, u32 logp) {
  u32 r = d.rng, v = d.val;
  u32 s = r >> logp;
  int ret = v < s;
  if (!ret) d.val = v - s;
  d.rng = ret ? s : r - s;
  rd_normalize(d);
  return ret;
}
// ec_dec_bits (entdec.c:284-319): raw bits from the tail of the packet
CB_FN_NOINLINE u32 rd_bits(RangeDec& d, u32 bits) {
  u32 window = d.end_window;
  int available = d.nend_bits;
  if ((u32)available < bits) {
    CB_NOUNROLL do {
      window |= (u32)rd_byte_end(d) << available;
      available += 8;
    } while (available <= 32 - 8);
  }
  u32 ret = window & ((1u << bits) - 1u);
  window >>= bits;
  available -= (int)bits;
  d.end_window = window;
  d.nend_bits = available;
  d.nbits_total += (int)bits;
  return ret;
}
// ec_dec_uint (entdec.c:238-282), EC_UINT_BITS=8
CB_FN_NOINLINE u32 rd_uint(RangeDec& d, u32 ft_in) {
  u32 ft = ft_in - 1;
  int ftb = cb_ilog(ft);
  if (ftb > 8) {
    ftb -= 8;
    u32 top = (ft >> ftb) + 1;
    u32 s = rd_decode(d, top);
    rd_update(d, s, s + 1, top);
    u32 t = (s << ftb) | rd_bits(d, (u32)ftb);
    if (t <= ft) return t;
    d.error = 1;
    return ft;
  }
  ft++;
  u32 s = rd_decode(d, ft);
  rd_update(d, s, s + 1, ft);
  return s;
}
// ec_laplace_decode (laplace.c:94-134), LAPLACE_MINP=1, LAPLACE_NMIN=16
CB_FN_NOINLINE int rd_laplace(RangeDec& d, u32 fs, int decay) {
  int val = 0;
  u32 fm = rd_decode_bin(d, 15);
  u32 fl = 0;
  if (fm >= fs) {
    val++;
    fl = fs;
    // ec_laplace_get_freq1 (laplace.c:44-49)
    fs = (((32768u - 32u - fs) * (u32)(16384 - decay)) >> 15) + 1;
    CB_NOUNROLL while (fs > 1 && fm >= fl + 2 * fs) {
      fs *= 2;
      fl += fs;
      fs = (u32)(((i32)(fs - 2) * (i32)decay) >> 15);
      fs += 1;
      val++;
    }
    if (fs <= 1) {
      int di = (int)((fm - fl) >> 1);
      val += di;
      fl += 2 * (u32)di;
    }
    if (fm < fl + fs) val = -val;
    else fl += fs;
  }
  u32 fh = fl + fs < 32768u ? fl + fs : 32768u;
  rd_update(d, fl, fh, 32768u);
  return val;
}
