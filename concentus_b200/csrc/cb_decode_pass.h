// cb_decode_pass.h -- per-stream bodies of the decode kernels (shared by the __global__ wrappers in cb_kernels.cu
// and by the test-only lane-loop emulation).  The entropy stage is in cb_entropy_pass.h.
#pragma once
#include "cb_entropy_pass.h"
#include "cb_celt_synth.h"
#include "cb_celt_plc.h"
#include "cb_opus_plan.h"

// Kernel B, one (stream, channel), one pass.  `row` = this stream's PCM row.
CB_FN void decode_synth_pass(SynWork& w, OpusDecState* st, const StreamPlan* pl, const FrameSide* side,
                             i16* row, int c, int pass, EmitReq* defer = nullptr) {
  const int CC = st->channels;
  const int route = side->route;
  if (route == CB_ROUTE_SILK || route == CB_ROUTE_GENERAL || route == CB_ROUTE_FINISH || route == CB_ROUTE_PLC) return;
  if (side->kind == 1) {
    celt_synth_channel(w, &st->celt, side, c, row + (size_t)pl->pcm_off[pass] * CC, st->decode_gain, defer);
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
