// opus_repacketizer.cu -- host-side packet surgery of the libopus ABI: merging / splitting packets
// (opus_repacketizer_*), CBR padding and its removal (opus_packet_pad / _unpad) and the multistream forms.
// Byte shuffling on packets of at most a few hundred bytes: it runs on the calling thread, no GPU work.
// Behaviour follows opus-fix/src/repacketizer.c (cat :59-91, out_range_impl :103-248, pad/unpad :250-349);
// the framing rules are RFC 6716 section 3.2.  The packet parser is the one the decoder uses
// (cb_opus_plan.h:opus_parse_packet = opus_packet_parse_impl).
#include <stdlib.h>
#include <string.h>

#include "../../include/opus_b200.h"
#include "cb_opus_plan.h"
#include "cb_repack_internal.h"
#include "cb_repack.h"



int opusb_rp_cat(OpusRepacketizer* rp, const unsigned char* data, opus_int32 len, int self_delimited) {
  if (len < 1) return OPUS_INVALID_PACKET;
  if (rp->nb_frames == 0) {
    rp->toc = data[0];
    rp->framesize = opus_toc_samples_per_frame(data[0], 8000);
  } else if ((rp->toc & 0xFC) != (data[0] & 0xFC)) {
    return OPUS_INVALID_PACKET;                     // only frames of one configuration can share a packet
  }
  const int incoming = opus_packet_get_nb_frames(data, len);
  if (incoming < 1) return OPUS_INVALID_PACKET;
  if ((incoming + rp->nb_frames) * rp->framesize > 960) return OPUS_INVALID_PACKET;   // 120 ms
  int toc = 0, poff = 0;
  const int count = opus_parse_packet(data, len, self_delimited, &toc, &rp->len[rp->nb_frames], &poff, nullptr);
  if (count < 1) return count;
  const unsigned char* p = data + poff;
  for (int i = 0; i < count; i++) {
    rp->frames[rp->nb_frames + i] = p;
    p += rp->len[rp->nb_frames + i];
  }
  rp->nb_frames += incoming;
  return OPUS_OK;
}

// Emit frames [begin, end) as one packet of at most maxlen bytes; with `pad` the packet is filled up to
// exactly maxlen.  The header is planned first (code, VBR flag, padding bytes), then written, then the
// frames are moved behind it (memmove: pad/unpad work in place).
opus_int32 opusb_rp_emit(OpusRepacketizer* rp, int begin, int end, unsigned char* data, opus_int32 maxlen,
                         int self_delimited, int pad) {
  if (begin < 0 || begin >= end || end > rp->nb_frames) return OPUS_BAD_ARG;
  return cb_rp_emit(rp->frames + begin, rp->len + begin, end - begin, rp->toc, data, maxlen, self_delimited, pad);
}

#define rp_cat opusb_rp_cat
#define rp_emit opusb_rp_emit

extern "C" {

int opus_repacketizer_get_size(void) { return (int)sizeof(OpusRepacketizer); }

OpusRepacketizer* opus_repacketizer_init(OpusRepacketizer* rp) {
  if (rp) rp->nb_frames = 0;
  return rp;
}

OpusRepacketizer* opus_repacketizer_create(void) {
  OpusRepacketizer* rp = (OpusRepacketizer*)malloc(sizeof(OpusRepacketizer));
  return rp ? opus_repacketizer_init(rp) : nullptr;
}

void opus_repacketizer_destroy(OpusRepacketizer* rp) { free(rp); }

int opus_repacketizer_cat(OpusRepacketizer* rp, const unsigned char* data, opus_int32 len) {
  if (!rp || (!data && len > 0)) return OPUS_BAD_ARG;
  return rp_cat(rp, data, len, 0);
}

int opus_repacketizer_get_nb_frames(OpusRepacketizer* rp) { return rp ? rp->nb_frames : 0; }

opus_int32 opus_repacketizer_out_range(OpusRepacketizer* rp, int begin, int end, unsigned char* data,
                                       opus_int32 maxlen) {
  if (!rp || !data) return OPUS_BAD_ARG;
  return rp_emit(rp, begin, end, data, maxlen, 0, 0);
}

opus_int32 opus_repacketizer_out(OpusRepacketizer* rp, unsigned char* data, opus_int32 maxlen) {
  if (!rp || !data) return OPUS_BAD_ARG;
  return rp_emit(rp, 0, rp->nb_frames, data, maxlen, 0, 0);
}

int opus_packet_pad(unsigned char* data, opus_int32 len, opus_int32 new_len) {
  if (!data || len < 1) return OPUS_BAD_ARG;
  if (len == new_len) return OPUS_OK;
  if (len > new_len) return OPUS_BAD_ARG;
  OpusRepacketizer rp;
  rp.nb_frames = 0;
  // park the payload at the end of the buffer, then rebuild the packet from the front
  memmove(data + new_len - len, data, (size_t)len);
  rp_cat(&rp, data + new_len - len, len, 0);
  const opus_int32 ret = rp_emit(&rp, 0, rp.nb_frames, data, new_len, 0, 1);
  return ret > 0 ? OPUS_OK : ret;
}

opus_int32 opus_packet_unpad(unsigned char* data, opus_int32 len) {
  if (!data || len < 1) return OPUS_BAD_ARG;
  OpusRepacketizer rp;
  rp.nb_frames = 0;
  const int rc = rp_cat(&rp, data, len, 0);
  if (rc < 0) return rc;
  return rp_emit(&rp, 0, rp.nb_frames, data, len, 0, 0);
}

int opus_multistream_packet_pad(unsigned char* data, opus_int32 len, opus_int32 new_len, int nb_streams) {
  if (!data || len < 1) return OPUS_BAD_ARG;
  if (len == new_len) return OPUS_OK;
  if (len > new_len) return OPUS_BAD_ARG;
  const opus_int32 amount = new_len - len;
  // the padding goes into the last (not self-delimited) stream
  for (int s = 0; s < nb_streams - 1; s++) {
    if (len <= 0) return OPUS_INVALID_PACKET;
    int toc = 0;
    opus_int16 size[48];
    i32 packet_offset = 0;
    const int count = opus_parse_packet(data, len, 1, &toc, size, nullptr, &packet_offset);
    if (count < 0) return count;
    data += packet_offset;
    len -= packet_offset;
  }
  return opus_packet_pad(data, len, len + amount);
}

opus_int32 opus_multistream_packet_unpad(unsigned char* data, opus_int32 len, int nb_streams) {
  if (!data || len < 1) return OPUS_BAD_ARG;
  unsigned char* dst = data;
  opus_int32 dst_len = 0;
  for (int s = 0; s < nb_streams; s++) {
    const int self_delimited = s != nb_streams - 1;
    if (len <= 0) return OPUS_INVALID_PACKET;
    int toc = 0;
    opus_int16 size[48];
    i32 packet_offset = 0;
    opus_int32 ret = opus_parse_packet(data, len, self_delimited, &toc, size, nullptr, &packet_offset);
    if (ret < 0) return ret;
    OpusRepacketizer rp;
    rp.nb_frames = 0;
    ret = rp_cat(&rp, data, packet_offset, self_delimited);
    if (ret < 0) return ret;
    ret = rp_emit(&rp, 0, rp.nb_frames, dst, len, self_delimited, 0);
    if (ret < 0) return ret;
    dst_len += ret;
    dst += ret;
    data += packet_offset;
    len -= packet_offset;
  }
  return dst_len;
}

// Index an opus_demo ".bit" image (per packet: big-endian u32 length, big-endian u32 encoder final range,
// payload; opus-fix/src/opus_demo.c:750-760) for the flat batch calls: offsets[i] / lens[i] address packet i
// inside `file` itself, final_ranges[i] is the recorded check value (any of the three may be NULL).
// Returns the number of packets in the image, or OPUS_INVALID_PACKET when it is truncated / a length is
// above 1500 (opus_demo's own limit); at most max_packets entries are written.
opus_int32 opus_b200_bit_index(const unsigned char* file, opus_int32 file_bytes, opus_int32* offsets,
                               opus_int32* lens, opus_uint32* final_ranges, opus_int32 max_packets) {
  if (!file || file_bytes < 0) return OPUS_BAD_ARG;
  opus_int32 pos = 0, n = 0;
  while (pos < file_bytes) {
    if (file_bytes - pos < 8) return OPUS_INVALID_PACKET;
    const opus_uint32 len = ((opus_uint32)file[pos] << 24) | ((opus_uint32)file[pos + 1] << 16) |
                            ((opus_uint32)file[pos + 2] << 8) | file[pos + 3];
    const opus_uint32 rng = ((opus_uint32)file[pos + 4] << 24) | ((opus_uint32)file[pos + 5] << 16) |
                            ((opus_uint32)file[pos + 6] << 8) | file[pos + 7];
    pos += 8;
    if (len > 1500 || (opus_int32)len > file_bytes - pos) return OPUS_INVALID_PACKET;
    if (n < max_packets) {
      if (offsets) offsets[n] = len ? pos : -1;       // a zero-length record is a lost packet
      if (lens) lens[n] = (opus_int32)len;
      if (final_ranges) final_ranges[n] = rng;
    }
    pos += (opus_int32)len;
    n++;
  }
  return n;
}

}  // extern "C"
