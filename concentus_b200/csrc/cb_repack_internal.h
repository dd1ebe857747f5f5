// cb_repack_internal.h -- the repacketizer's two workers, shared with the multistream encoder
// (opus_repacketizer_cat_impl / opus_repacketizer_out_range_impl of opus-fix/src/repacketizer.c).
#pragma once
#include "../../include/opus_b200.h"

struct OpusRepacketizer {
  unsigned char toc;
  int nb_frames;
  const unsigned char* frames[48];
  opus_int16 len[48];
  int framesize;                 // samples per frame at 8 kHz (the 120 ms limit is 960 of them)
};

int opusb_rp_cat(OpusRepacketizer* rp, const unsigned char* data, opus_int32 len, int self_delimited);
opus_int32 opusb_rp_emit(OpusRepacketizer* rp, int begin, int end, unsigned char* data, opus_int32 maxlen,
                         int self_delimited, int pad);
