// cb_repack.h -- assembling one Opus packet from frames of one configuration (RFC 6716 section 3.2;
// opus_repacketizer_out_range_impl, opus-fix/src/repacketizer.c:103-248).  Host and device: the host
// repacketizer entry points use it, and so does the encoder kernel when a 40 / 60 ms frame has been coded
// as two or three 20 ms frames (opus_encoder.c:1362-1438).
#pragma once
#include "cb_common.h"

// frame-length field of RFC 6716 3.2.1: one byte below 252, else two
CB_HD int cb_rp_put_size(int size, u8* p) {
  if (size < 252) { p[0] = (u8)size; return 1; }
  p[0] = (u8)(252 + (size & 3));
  p[1] = (u8)((size - (int)p[0]) >> 2);
  return 2;
}
CB_HD int cb_rp_size_bytes(int size) { return size < 252 ? 1 : 2; }
CB_HD void cb_rp_move(u8* dst, const u8* src, int n) {       // memmove
  if (dst == src || n <= 0) return;
  if (dst < src) for (int i = 0; i < n; i++) dst[i] = src[i];
  else for (int i = n - 1; i >= 0; i--) dst[i] = src[i];
}

// Emit `count` frames (pointers + lengths) as one packet of at most maxlen bytes; with `pad` the packet is
// filled up to exactly maxlen.  The header is planned first (code, VBR flag, padding bytes), then written,
// then the frames are moved behind it (they may live in the output buffer itself: pad / unpad work in place).
// Returns the packet length, OPUS_BUFFER_TOO_SMALL (-2) or OPUS_BAD_ARG (-1).
CB_HD i32 cb_rp_emit(const u8* const* frames, const i16* len, int count, u8 toc, u8* data, i32 maxlen,
                     int self_delimited, int pad) {
  if (count < 1) return -1;
  const u8 cfg = toc & 0xFC;
  const i32 sd_bytes = self_delimited ? cb_rp_size_bytes(len[count - 1]) : 0;
  i32 payload = 0;
  bool equal = true;
  for (int i = 0; i < count; i++) {
    payload += len[i];
    if (len[i] != len[0]) equal = false;
  }
  int code;
  i32 total;
  if (count == 1) { code = 0; total = sd_bytes + 1 + payload; }
  else if (count == 2 && equal) { code = 1; total = sd_bytes + 1 + payload; }
  else if (count == 2) { code = 2; total = sd_bytes + 1 + cb_rp_size_bytes(len[0]) + payload; }
  else { code = 3; total = 0; }
  if (code != 3 && total > maxlen) return -2;
  if (code != 3 && pad && total < maxlen) code = 3;           // only code 3 can carry padding
  u8* p = data;
  if (code == 3) {
    total = sd_bytes + 2 + payload;
    if (!equal)
      for (int i = 0; i < count - 1; i++) total += cb_rp_size_bytes(len[i]);
    if (total > maxlen) return -2;
    const i32 pad_amount = pad ? maxlen - total : 0;
    *p++ = cfg | 3;
    *p++ = (u8)(count | (equal ? 0 : 0x80) | (pad_amount ? 0x40 : 0));
    if (pad_amount) {
      const int n255 = (pad_amount - 1) / 255;                 // the length bytes count as padding themselves
      for (int i = 0; i < n255; i++) *p++ = 255;
      *p++ = (u8)(pad_amount - 255 * n255 - 1);
      total += pad_amount;
    }
    if (!equal)
      for (int i = 0; i < count - 1; i++) p += cb_rp_put_size(len[i], p);
  } else {
    *p++ = cfg | (u8)code;
    if (code == 2) p += cb_rp_put_size(len[0], p);
  }
  if (self_delimited) p += cb_rp_put_size(len[count - 1], p);
  for (int i = 0; i < count; i++) {
    cb_rp_move(p, frames[i], len[i]);
    p += len[i];
  }
  if (pad)
    while (p < data + maxlen) *p++ = 0;
  return total;
}
