// cb_opus_plan.h -- plan kernel body: the Opus layer in front of the codec kernels.
//
// One thread per stream parses the TOC and the frame packing of its packet, applies the
// argument/packet checks of opus_decode_native(), updates the Opus-layer state and writes a
// StreamPlan that the entropy/synthesis passes follow.  Restates, for the decode direction,
// opus_packet_parse_impl (opus-fix/src/opus.c:190-343), opus_packet_get_samples_per_frame
// (opus.c:169-188), opus_packet_get_bandwidth/nb_channels (src/opus_decoder.c:921-947) and
// opus_decode_native (opus_decoder.c:598-709) incl. the PLC chunking of opus_decode_frame
// (opus_decoder.c:239-292).
#pragma once
#include "cb_common.h"
#include "cb_state.h"

CB_HD int opus_toc_mode(int toc) {                       // opus_decoder.c:184-198
  if (toc & 0x80) return CB_MODE_CELT_ONLY;
  if ((toc & 0x60) == 0x60) return CB_MODE_HYBRID;
  return CB_MODE_SILK_ONLY;
}
CB_HD int opus_toc_bandwidth(int toc) {                  // opus_decoder.c:921-937
  int bw;
  if (toc & 0x80) {
    bw = CB_BW_MEDIUMBAND + ((toc >> 5) & 0x3);
    if (bw == CB_BW_MEDIUMBAND) bw = CB_BW_NARROWBAND;
  } else if ((toc & 0x60) == 0x60) {
    bw = (toc & 0x10) ? CB_BW_FULLBAND : CB_BW_SUPERWIDEBAND;
  } else {
    bw = CB_BW_NARROWBAND + ((toc >> 5) & 0x3);
  }
  return bw;
}
CB_HD int opus_toc_samples_per_frame(int toc, int Fs) {  // opus.c:169-188
  int audiosize;
  if (toc & 0x80) {
    audiosize = (toc >> 3) & 0x3;
    audiosize = (Fs << audiosize) / 400;
  } else if ((toc & 0x60) == 0x60) {
    audiosize = (toc & 0x08) ? Fs / 50 : Fs / 100;
  } else {
    audiosize = (toc >> 3) & 0x3;
    if (audiosize == 3) audiosize = Fs * 60 / 1000;
    else audiosize = (Fs << audiosize) / 100;
  }
  return audiosize;
}
CB_HD int opus_bandwidth_endband(int bw) {               // opus_decoder.c:432-449
  switch (bw) {
    case CB_BW_NARROWBAND: return 13;
    case CB_BW_MEDIUMBAND:
    case CB_BW_WIDEBAND: return 17;
    case CB_BW_SUPERWIDEBAND: return 19;
    default: return 21;
  }
}
CB_HD int parse_size(const u8* data, i32 len, int* size) {   // opus.c:148-167
  if (len < 1) { *size = -1; return -1; }
  if (data[0] < 252) { *size = data[0]; return 1; }
  if (len < 2) { *size = -1; return -1; }
  *size = 4 * data[1] + data[0];
  return 2;
}
// opus_packet_parse_impl.  sizes[] gets the frame sizes, *payload_offset the offset of frame 0,
// *packet_offset the total bytes consumed (self-delimited streams).  Returns count or <0.
CB_HD int opus_parse_packet(const u8* data, i32 len, int self_delimited, int* out_toc, i16* sizes,
                            int* payload_offset, i32* packet_offset) {
  if (len < 0) return CB_BAD_ARG;
  if (len == 0) return CB_INVALID_PACKET;
  const u8* data0 = data;
  int framesize = opus_toc_samples_per_frame(data[0], 48000);
  int cbr = 0, count, bytes, sz;
  int toc = *data++;
  len--;
  i32 last_size = len;
  i32 pad = 0;
  switch (toc & 0x3) {
    case 0:
      count = 1;
      break;
    case 1:
      count = 2;
      cbr = 1;
      if (!self_delimited) {
        if (len & 0x1) return CB_INVALID_PACKET;
        last_size = len / 2;
        sizes[0] = (i16)last_size;
      }
      break;
    case 2:
      count = 2;
      bytes = parse_size(data, len, &sz);
      sizes[0] = (i16)sz;
      len -= bytes;
      if (sz < 0 || sz > len) return CB_INVALID_PACKET;
      data += bytes;
      last_size = len - sz;
      break;
    default: {
      if (len < 1) return CB_INVALID_PACKET;
      int ch = *data++;
      count = ch & 0x3F;
      if (count <= 0 || framesize * count > 5760) return CB_INVALID_PACKET;
      len--;
      if (ch & 0x40) {
        int p;
        do {
          if (len <= 0) return CB_INVALID_PACKET;
          p = *data++;
          len--;
          int tmp = p == 255 ? 254 : p;
          len -= tmp;
          pad += tmp;
        } while (p == 255);
      }
      if (len < 0) return CB_INVALID_PACKET;
      cbr = !(ch & 0x80);
      if (!cbr) {
        last_size = len;
        for (int i = 0; i < count - 1; i++) {
          bytes = parse_size(data, len, &sz);
          sizes[i] = (i16)sz;
          len -= bytes;
          if (sz < 0 || sz > len) return CB_INVALID_PACKET;
          data += bytes;
          last_size -= bytes + sz;
        }
        if (last_size < 0) return CB_INVALID_PACKET;
      } else if (!self_delimited) {
        last_size = len / count;
        if (last_size * count != len) return CB_INVALID_PACKET;
        for (int i = 0; i < count - 1; i++) sizes[i] = (i16)last_size;
      }
      break;
    }
  }
  if (self_delimited) {
    bytes = parse_size(data, len, &sz);
    sizes[count - 1] = (i16)sz;
    len -= bytes;
    if (sz < 0 || sz > len) return CB_INVALID_PACKET;
    data += bytes;
    if (cbr) {
      if (sz * count > len) return CB_INVALID_PACKET;
      for (int i = 0; i < count - 1; i++) sizes[i] = (i16)sz;
    } else if (bytes + sz > last_size) {
      return CB_INVALID_PACKET;
    }
  } else {
    if (last_size > 1275) return CB_INVALID_PACKET;
    sizes[count - 1] = (i16)last_size;
  }
  if (payload_offset) *payload_offset = (int)(data - data0);
  for (int i = 0; i < count; i++) data += sizes[i];
  if (packet_offset) *packet_offset = pad + (i32)(data - data0);
  if (out_toc) *out_toc = toc;
  return count;
}

// Concealment plan for `frame_size` samples: the chunk sizes opus_decode_frame(data=NULL) picks
// when called in the loop of opus_decode_native (opus_decoder.c:613-627, 239-292).
CB_HD void opus_plan_plc(const OpusDecState* st, StreamPlan* pl, int frame_size, int max_passes = CB_MAX_FRAMES) {
  const int Fs = st->Fs;
  const int F20 = Fs / 50, F10 = F20 >> 1, F5 = F10 >> 1, F2_5 = F5 >> 1;
  const int mode = st->prev_mode;
  pl->total_samples = frame_size;
  pl->ret = frame_size;
  if (mode == 0) {
    pl->kind = CB_PLAN_ZERO;
    pl->nframes = 1;
    return;
  }
  pl->kind = CB_PLAN_PLC;
  int done = 0, nf = 0;
  while (done < frame_size) {
    int fs = frame_size - done;
    if (fs < F2_5) { pl->ret = CB_BUFFER_TOO_SMALL; pl->kind = CB_PLAN_NONE; pl->nframes = 0; return; }
    fs = cb_imin(fs, Fs / 25 * 3);
    fs = cb_imin(fs, st->frame_size);
    int audiosize = fs;
    if (audiosize > F20) audiosize = F20;   // (the reference recurses in F20 pieces)
    else if (audiosize < F20) {
      if (audiosize > F10) audiosize = F10;
      else if (mode != CB_MODE_SILK_ONLY && audiosize > F5 && audiosize < F10) audiosize = F5;
    }
    if (nf >= max_passes) { pl->ret = CB_BAD_ARG; pl->kind = CB_PLAN_NONE; pl->nframes = 0; return; }
    pl->off[nf] = 0;
    pl->len[nf] = 0;
    pl->nsamp[nf] = (i16)audiosize;
    pl->pcm_off[nf] = done;
    nf++;
    done += audiosize;
  }
  pl->nframes = nf;
}

// opus_decode_native for one stream.  `data` = start of the packet arena, `off` = this packet's
// offset or <0 for "no packet" (NULL), `len` its length.  `max_passes` = how many passes (frames of a packet,
// concealment pieces, the pieces in front of an FEC frame) the caller is going to launch: a stream that needs
// more fails with OPUS_BAD_ARG before any of its state is touched (the host entry points compute the count from
// the packets; the device-resident ones take it from their max_frames argument).
CB_HD void opus_plan_stream(OpusDecState* st, StreamPlan* pl, const u8* data, i32 off, i32 len,
                            int frame_size, int decode_fec, int self_delimited, i32* packet_offset,
                            int max_passes = CB_MAX_FRAMES) {
  if (max_passes > CB_MAX_FRAMES) max_passes = CB_MAX_FRAMES;
  pl->kind = CB_PLAN_NONE;
  pl->nframes = 0;
  pl->total_samples = 0;
  pl->fec_frame = -1;
  const int Fs = st->Fs;
  if (frame_size <= 0) { pl->ret = CB_BAD_ARG; return; }
  if (decode_fec < 0 || decode_fec > 1) { pl->ret = CB_BAD_ARG; return; }
  const int lost = (len == 0 || off < 0);
  if ((decode_fec || lost) && frame_size % (Fs / 400) != 0) { pl->ret = CB_BAD_ARG; return; }
  if (lost) {
    opus_plan_plc(st, pl, frame_size, max_passes);
    if (pl->ret >= 0) st->last_packet_duration = pl->ret;
    return;
  }
  if (len < 0) { pl->ret = CB_BAD_ARG; return; }
  const u8* pkt = data + off;
  const int toc0 = pkt[0];
  const int packet_mode = opus_toc_mode(toc0);
  const int packet_bandwidth = opus_toc_bandwidth(toc0);
  const int packet_frame_size = opus_toc_samples_per_frame(toc0, Fs);
  const int packet_stream_channels = (toc0 & 0x4) ? 2 : 1;
  i16 sizes[CB_MAX_FRAMES];
  int payload_offset = 0, toc = 0;
  int count = opus_parse_packet(pkt, len, self_delimited, &toc, sizes, &payload_offset, packet_offset);
  if (count < 0) { pl->ret = count; return; }
  if (decode_fec) {
    // No FEC in CELT-only packets: conceal instead (opus_decoder.c:647-648)
    if (frame_size < packet_frame_size || packet_mode == CB_MODE_CELT_ONLY || st->mode == CB_MODE_CELT_ONLY) {
      opus_plan_plc(st, pl, frame_size, max_passes);
      if (pl->ret >= 0) st->last_packet_duration = pl->ret;
      return;
    }
    // SILK in-band FEC (opus_decoder.c:649-672): conceal everything but the last packet_frame_size
    // samples, then decode the first frame's LBRR data into those
    const int duration_copy = st->last_packet_duration;
    int nf = 0, plc_samples = frame_size - packet_frame_size;
    if (plc_samples != 0) {
      opus_plan_plc(st, pl, plc_samples, max_passes);
      if (pl->ret < 0) { st->last_packet_duration = duration_copy; return; }
      if (pl->kind == CB_PLAN_ZERO) {           // nothing decoded yet: one all-zero frame
        pl->off[0] = 0; pl->len[0] = 0; pl->nsamp[0] = (i16)plc_samples; pl->pcm_off[0] = 0;
        nf = 1;
      } else {
        nf = pl->nframes;
      }
    }
    if (nf >= max_passes) { pl->ret = CB_BAD_ARG; pl->kind = CB_PLAN_NONE; pl->nframes = 0; return; }
    st->mode = packet_mode;
    st->bandwidth = packet_bandwidth;
    st->frame_size = packet_frame_size;
    st->stream_channels = packet_stream_channels;
    pl->off[nf] = off + payload_offset;
    pl->len[nf] = sizes[0];
    pl->nsamp[nf] = (i16)packet_frame_size;
    pl->pcm_off[nf] = plc_samples;
    pl->fec_frame = sizes[0] > 1 ? nf : -1;     // a 0/1-byte first frame is concealed, not FEC-decoded
    pl->kind = CB_PLAN_DECODE;
    pl->nframes = nf + 1;
    pl->total_samples = frame_size;
    pl->ret = frame_size;
    st->last_packet_duration = frame_size;
    return;
  }
  if (count * packet_frame_size > frame_size) { pl->ret = CB_BUFFER_TOO_SMALL; return; }
  if (count > max_passes) { pl->ret = CB_BAD_ARG; return; }
  st->mode = packet_mode;
  st->bandwidth = packet_bandwidth;
  st->frame_size = packet_frame_size;
  st->stream_channels = packet_stream_channels;
  i32 o = off + payload_offset;
  for (int i = 0; i < count; i++) {
    pl->off[i] = o;
    pl->len[i] = sizes[i];
    pl->nsamp[i] = (i16)packet_frame_size;
    pl->pcm_off[i] = i * packet_frame_size;
    o += sizes[i];
  }
  pl->kind = CB_PLAN_DECODE;
  pl->nframes = count;
  pl->total_samples = count * packet_frame_size;
  pl->ret = count * packet_frame_size;
  st->last_packet_duration = pl->ret;
}

// Where frame `pass` of the plan is decoded.  The SILK kernel runs first in every pass and stores the
// decision in the plan; the CELT kernels and the general kernel read it.
#define CB_ROUTE_CELT 0        // CELT kernels (A/B/C)
#define CB_ROUTE_SILK 1        // SILK kernel alone
#define CB_ROUTE_GENERAL 2     // general kernel (cb_opus_general.h): transitions, FEC of hybrid frames, odd sizes
#define CB_ROUTE_FINISH 3      // SILK kernel decoded the SILK layer and found a redundant CELT frame: general kernel finishes
#define CB_ROUTE_HYBRID 4      // SILK kernel decodes the SILK layer, then the CELT kernels add bands 17.. on top
#define CB_ROUTE_PLC 5         // lost CELT frame, pitch-based concealment (celt_decode_lost, loss_count < 5): general kernel
CB_HD int opus_frame_route(const OpusDecState* st, const StreamPlan* pl, int pass) {
  const bool conceal = pl->kind == CB_PLAN_PLC || (pl->len[pass] <= 1);
  const int prev = st->prev_mode;
  // the two-kernel hybrid path needs the SILK layer written straight into the PCM row (frames of >= 10 ms,
  // the reference's celt_accum case) and no output gain
  const bool split_ok = pl->nsamp[pass] >= st->Fs / 100 && st->decode_gain == 0;
  if (conceal) {
    if (prev == CB_MODE_SILK_ONLY) return CB_ROUTE_SILK;
    if (prev == CB_MODE_HYBRID) return split_ok ? CB_ROUTE_HYBRID : CB_ROUTE_GENERAL;
    if (prev == CB_MODE_CELT_ONLY && st->celt.loss_count < 5) return CB_ROUTE_PLC;
    return CB_ROUTE_CELT;                      // nothing decoded yet (zeros) or noise-based concealment
  }
  const int mode = st->mode;
  if (mode == CB_MODE_CELT_ONLY) {
    if (prev == 0 || prev == CB_MODE_CELT_ONLY || st->prev_redundancy) return CB_ROUTE_CELT;
    return CB_ROUTE_GENERAL;
  }
  if (prev == CB_MODE_CELT_ONLY) return CB_ROUTE_GENERAL;
  if (mode == CB_MODE_SILK_ONLY) return prev == CB_MODE_HYBRID ? CB_ROUTE_GENERAL : CB_ROUTE_SILK;
  if (pl->fec_frame == pass || !split_ok) return CB_ROUTE_GENERAL;
  return CB_ROUTE_HYBRID;
}

// opus_multistream_packet_validate (opus-fix/src/opus_multistream_decoder.c:155-183)
CB_HD int opus_ms_validate(const u8* data, i32 len, int nb_streams, i32 Fs) {
  int samples = 0;
  i16 size[CB_MAX_FRAMES];
  for (int s = 0; s < nb_streams; s++) {
    if (len <= 0) return CB_INVALID_PACKET;
    int toc = 0;
    i32 packet_offset = 0;
    int count = opus_parse_packet(data, len, s != nb_streams - 1, &toc, size, nullptr, &packet_offset);
    if (count < 0) return count;
    // opus_packet_get_nb_samples(data, packet_offset, Fs) (opus_decoder.c:957-973)
    int cnt;
    {
      int c3 = data[0] & 0x3;
      if (packet_offset < 1) return CB_BAD_ARG;
      if (c3 == 0) cnt = 1;
      else if (c3 != 3) cnt = 2;
      else if (packet_offset < 2) return CB_INVALID_PACKET;
      else cnt = data[1] & 0x3F;
    }
    int tmp_samples = cnt * opus_toc_samples_per_frame(data[0], Fs);
    if (tmp_samples * 25 > Fs * 3) return CB_INVALID_PACKET;
    if (s != 0 && samples != tmp_samples) return CB_INVALID_PACKET;
    samples = tmp_samples;
    data += packet_offset;
    len -= packet_offset;
  }
  return samples;
}

// opus_multistream_decode_native (opus_multistream_decoder.c:184-302), planning half: walks the
// sub-packets of one multistream packet and plans every elementary stream.  states/plans are the
// group's arrays; the layout says which slice belongs to this multistream decoder.
CB_HD int opus_ms_plan(OpusDecState* const* states, StreamPlan* plans, const MSLayout* L, const u8* arena,
                       i32 off, i32 len, int frame_size, int decode_fec, int max_passes = CB_MAX_FRAMES) {
  const int S = L->nb_streams;
  const i32 Fs = L->Fs;
  for (int s = 0; s < S; s++) {
    StreamPlan* pl = &plans[L->first + s];
    pl->kind = CB_PLAN_NONE; pl->nframes = 0; pl->total_samples = 0; pl->ret = CB_INVALID_STATE;
  }
  if (frame_size <= 0) return CB_BAD_ARG;
  frame_size = cb_imin(frame_size, Fs / 25 * 3);
  if (off < 0) len = 0;
  const int do_plc = (len == 0);
  if (len < 0) return CB_BAD_ARG;
  if (!do_plc && len < 2 * S - 1) return CB_INVALID_PACKET;
  const u8* data = arena + (off < 0 ? 0 : off);
  if (!do_plc) {
    int r = opus_ms_validate(data, len, S, Fs);
    if (r < 0) return r;
    if (r > frame_size) return CB_BUFFER_TOO_SMALL;
  }
  i32 cur = off;
  for (int s = 0; s < S; s++) {
    if (!do_plc && len <= 0) return CB_INTERNAL_ERROR;
    i32 packet_offset = 0;
    StreamPlan* pl = &plans[L->first + s];
    opus_plan_stream(states[L->first + s], pl, arena, do_plc ? -1 : cur, do_plc ? 0 : len, frame_size, decode_fec,
                     s != S - 1, &packet_offset, max_passes);
    int r = pl->ret;
    if (r <= 0) return r;
    cur += packet_offset;
    len -= packet_offset;
    frame_size = r;
  }
  return frame_size;
}
