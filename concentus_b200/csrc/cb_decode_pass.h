// cb_decode_pass.h -- per-stream bodies of the two decode kernels (shared by the __global__
// wrappers in cb_kernels.cu and by the test-only lane-loop emulation).
#pragma once
#include "cb_celt_entropy.h"
#include "cb_celt_synth.h"
#include "cb_celt_plc.h"
#include "cb_opus_plan.h"

// Kernel A, one stream, one pass.  `pkt` = per-warp staging area for the payload bytes (`pkt_cap` bytes; a longer
// payload -- the device-resident entries cannot check the caller's max_packet_bytes -- is read in place), or nullptr: the range
// decoder then reads the packet where it lies in global memory (a few dozen dependent byte loads per frame are
// nothing next to the frame's decode time; staging large packets would cost a resident CTA per SM).
// Two instances: HYB = false takes the CELT-only frames (start band 0, fresh range decoder: the constants
// fold into the decoder), HYB = true the CELT layer of hybrid frames whose SILK layer the SILK kernel has
// already written to the PCM row (bands 17.., coder state continued, output accumulated).
template <bool HYB>
CB_FN void decode_entropy_pass(EntWork& w, u8* pkt, int pkt_cap, OpusDecState* st, StreamPlan* pl, FrameSide* side,
                               const u8* data, i32* ret_slot, int pass) {
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
  if (route != CB_ROUTE_CELT && route != CB_ROUTE_HYBRID) return;   // SILK / general kernel (which may be running now)
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
    celt_plc_noise(w, &st->celt, side, LMp, 17, opus_bandwidth_endband(st->bandwidth));
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
    celt_plc_noise(w, &st->celt, side, LMp, 0, opus_bandwidth_endband(st->bandwidth));
    st->rangeFinal = 0;
    st->prev_redundancy = 0;
    return;
  }
  const i32 off = pl->off[pass];
  const u8* src = data + off;
  if (pkt && len <= pkt_cap) {
    CB_LANE_LOOP(i, len) pkt[i] = src[i];
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
  int rc = celt_entropy_decode(w, &st->celt, side, src, len, cfg, &fr);
  CB_WSYNC();
  st->rangeFinal = fr;
  st->prev_mode = hybrid ? CB_MODE_HYBRID : CB_MODE_CELT_ONLY;
  st->prev_redundancy = 0;
  if (rc < 0) {
    pl->ret = rc;
    *ret_slot = rc;
  }
}

// Kernel B, one (stream, channel), one pass.  `row` = this stream's PCM row.
CB_FN void decode_synth_pass(SynWork& w, OpusDecState* st, const StreamPlan* pl, const FrameSide* side,
                             i16* row, int c, int pass) {
  const int CC = st->channels;
  const int route = pl->route;
  if (route == CB_ROUTE_SILK || route == CB_ROUTE_GENERAL || route == CB_ROUTE_FINISH || route == CB_ROUTE_PLC) return;
  if (side->kind == 1) {
    celt_synth_channel(w, &st->celt, side, c, row + (size_t)pl->pcm_off[pass] * CC, st->decode_gain);
  } else if (side->kind == 3 && pass == 0) {
    const int total = pl->total_samples;
    CB_LANE_LOOP(j, total) row[j * CC + c] = 0;
  } else if (side->kind == 4) {
    i16* p = row + (size_t)pl->pcm_off[pass] * CC;
    const int cnt = pl->nsamp[pass];
    CB_LANE_LOOP(j, cnt) p[j * CC + c] = 0;
  }
}

// Kernel C, one stream, one pass: pitch-based concealment (both channels) when kernel A asked for it.
CB_FN void decode_plc_pass(PlcWork& w, OpusDecState* st, const StreamPlan* pl, const FrameSide* side,
                           i16* row, int pass) {
  if (pl->route != CB_ROUTE_PLC || side->kind != 2) return;
  const int N = 120 << side->LM;
  celt_plc_pitch(w, &st->celt, N, row + (size_t)pl->pcm_off[pass] * st->channels, side->accum, st->decode_gain);
}
