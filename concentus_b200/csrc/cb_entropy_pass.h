// cb_entropy_pass.h -- per-stream bodies of the entropy stage of a decode pass: the PARSE half (parse kernel: one
// thread per stream; general kernel and host build: one warp / the only "lane") and the REBUILD half (rebuild
// kernel: one warp per stream).
#pragma once
#include "cb_celt_entropy.h"
#include "cb_opus_plan.h"

// The entropy stage of one stream and one pass, in two halves (cb_celt_entropy.h).
//
// PARSE half.  `pkt` = staging area for the payload bytes (`pkt_cap` bytes; a longer payload -- the device-resident
// entries cannot check the caller's max_packet_bytes -- is read in place), or nullptr: the range decoder then reads
// the packet where it lies in global memory.  Runs one thread per stream in the parse kernel and one warp per
// stream in the general kernel / host build.  Two instances: HYB = false takes the CELT-only frames (start band 0,
// fresh range decoder: the constants fold into the decoder), HYB = true the CELT layer of hybrid frames whose SILK
// layer the SILK kernel has already written to the PCM row (bands 17.., coder state continued, output accumulated).
// Leaves tape->kind = what the REBUILD half has to do for this stream.
template <bool HYB>
CB_FN void decode_entropy_parse(EntWork& w, u8* pkt, int pkt_cap, OpusDecState* st, StreamPlan* pl, FrameSide* side,
                                FrameTape* tape, const u8* data, i32* ret_slot, int pass) {
  const int kind = pl->kind;
  const int route = pl->route;                       // decided by the SILK kernel at the start of this pass
  if (HYB != (route == CB_ROUTE_HYBRID)) return;     // the other instance's stream (or nobody's, handled by HYB = false)
  if (pl->ret < 0 || pass >= pl->nframes || kind == CB_PLAN_NONE) {
    side->kind = 0;
    return;
  }
  if (kind == CB_PLAN_ZERO) {
    side->kind = 3;
    return;
  }
  if (route != CB_ROUTE_CELT && route != CB_ROUTE_HYBRID) return;   // SILK / general kernel (which may be running now,
                                                                    // with this stream's tape: hands off)
  tape->kind = CB_TAPE_NONE;
  constexpr bool hybrid = HYB;                     // SILK layer already in the PCM row: bands 17.. accumulate
  const int len = pl->len[pass];
  const int nsamp = pl->nsamp[pass];
  if (hybrid && (kind == CB_PLAN_PLC || len <= 1)) {
    // concealment of a hybrid frame: celt_decode_lost with start != 0 is always noise-based
    int LMp = 0;
    {
      int fs48 = nsamp * st->celt.downsample;
      while ((120 << LMp) < fs48) LMp++;
    }
    celt_plc_noise_parse(&st->celt, side, tape, LMp, 17, opus_bandwidth_endband(st->bandwidth));
    CB_WSYNC();
    side->accum = 1;
    st->rangeFinal = 0;
    st->prev_redundancy = 0;
    return;
  }
  if (!hybrid && (kind == CB_PLAN_PLC || len <= 1)) {
    // lost / DTX frame: opus_decode_frame(data=NULL) (opus_decoder.c:246-292, 579-586)
    if (st->prev_mode == 0) {                       // nothing decoded yet: zeros for this frame
      side->kind = 4;
      return;
    }
    int LMp = 0;
    {
      int fs48 = nsamp * st->celt.downsample;
      while ((120 << LMp) < fs48) LMp++;
    }
    // noise-based concealment (five or more losses in a row; the pitch-based kind is the PLC route):
    // synthesised by kernel B
    celt_plc_noise_parse(&st->celt, side, tape, LMp, 0, opus_bandwidth_endband(st->bandwidth));
    st->rangeFinal = 0;
    st->prev_redundancy = 0;
    return;
  }
  const i32 off = pl->off[pass];
  const u8* src = data + off;
  if (pkt && len <= pkt_cap) {
    if ((((size_t)src | (size_t)pkt) & 3) == 0) {        // four bytes at a time when both ends are word aligned
      const u32* s4 = (const u32*)src;
      u32* d4 = (u32*)pkt;
      const int nw = len >> 2;
      CB_LANE_LOOP(i, nw) d4[i] = s4[i];
      CB_LANE_LOOP(i, len & 3) pkt[4 * nw + i] = src[4 * nw + i];
    } else {
      CB_LANE_LOOP(i, len) pkt[i] = src[i];
    }
    src = pkt;
  }
  CB_WSYNC();
  int LM = 0;
  {
    int fs48 = nsamp * st->celt.downsample;
    while ((120 << LM) < fs48) LM++;
  }
  EntFrameCfg cfg;
  cfg.LM = LM;
  cfg.C = st->stream_channels;
  cfg.start = hybrid ? 17 : 0;
  cfg.end = opus_bandwidth_endband(st->bandwidth);
  cfg.accum = hybrid ? 1 : 0;
  cfg.rd0 = hybrid ? reinterpret_cast<const RangeDec*>(pl->rd_save) : nullptr;
  if (hybrid && st->prev_mode != CB_MODE_HYBRID && st->prev_mode > 0 && !st->prev_redundancy) {
    CB_WSYNC();
    CB_IF_LANE0 celt_state_reset(&st->celt, st->celt.channels, st->celt.downsample);   // opus_decoder.c:463-464
    CB_WSYNC();
  }
  u32 fr = 0;
  int rc = celt_frame_parse(w, &st->celt, side, tape, src, len, cfg, &fr);
  CB_WSYNC();
  st->rangeFinal = fr;
  st->prev_mode = hybrid ? CB_MODE_HYBRID : CB_MODE_CELT_ONLY;
  st->prev_redundancy = 0;
  if (rc < 0) {
    pl->ret = rc;
    *ret_slot = rc;
  }
}

#ifndef CB_THREAD_MODE
// REBUILD half: one warp turns the stream's tape into the spectrum the synthesis kernel reads.
CB_FN void decode_entropy_rebuild(EntWork& w, FrameTape* tape, FrameSide* side) {
  const int route = side->route;
  if (route != CB_ROUTE_CELT && route != CB_ROUTE_HYBRID) return;   // only these routes' tapes are this pass's
  const int kind = tape->kind;
  if (kind == CB_TAPE_FRAME) celt_frame_rebuild(w, tape, side);
  else if (kind == CB_TAPE_NOISE) celt_plc_noise_rebuild(w, tape, side);
}
// both halves on one warp (host build)
template <bool HYB>
CB_FN void decode_entropy_pass(EntWork& w, u8* pkt, int pkt_cap, OpusDecState* st, StreamPlan* pl, FrameSide* side,
                               FrameTape* tape, const u8* data, i32* ret_slot, int pass) {
  if (HYB != (pl->route == CB_ROUTE_HYBRID)) return;
  decode_entropy_parse<HYB>(w, pkt, pkt_cap, st, pl, side, tape, data, ret_slot, pass);
  CB_WSYNC();
  decode_entropy_rebuild(w, tape, side);
  CB_WSYNC();
}
#endif
