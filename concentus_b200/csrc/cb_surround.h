// cb_surround.h -- surround masking analysis of the multistream surround encoder: surround_analysis
// (opus-fix/src/opus_multistream_encoder.c:224-360), logSum (:186-218), channel_pos (:146-183).
// For every input channel: pre-emphasis, one long MDCT of the frame's first 20 ms (shorter frames: one MDCT of
// the frame), band energies in the log domain with a spreading slope; then per band the energy of the channels
// that sit left / right / centre is summed and each channel gets "own level minus its side's total" as the
// mask the CELT encoder of its stream reads (celt_encoder.c:1726-1801).  FIXED_POINT, bit-exact.
// One warp per channel (surround_channel), then one warp over the 21 bands (surround_combine).
#pragma once
#include "cb_common.h"
#include "cb_enc_analysis.h"
#include "cb_enc_mdct.h"

#define CB_SURROUND_MAX_CH 8
#define CB_SURROUND_MAX_FRAME 2880        // 60 ms at 48 kHz

// State the multistream encoder keeps between frames (opus_multistream_encoder.c:84-85)
struct SurroundState {
  i32 window_mem[CB_SURROUND_MAX_CH][CB_OVERLAP];
  i32 preemph_mem[CB_SURROUND_MAX_CH];
};
// Per-channel work buffers (global memory)
struct SurroundScratch {
  i32 in[CB_SURROUND_MAX_FRAME + CB_OVERLAP];
  i32 freq[CB_MAXN];
  i32 f2[CB_MAXN];
  i32 bandE[2 * CB_NBANDS];
};

// Position in the mix: 0 don't mix, 1 left, 2 centre, 3 right (:146-183)
CB_FN int surround_channel_pos(int channels, int c) {
  if (channels == 4) return (c & 1) ? 3 : 1;
  if (channels == 3 || channels == 5 || channels == 6) return c == 0 ? 1 : c == 1 ? 2 : c == 2 ? 3 : c == 3 ? 1 : c == 4 ? 3 : 0;
  if (channels == 7) return c == 0 ? 1 : c == 1 ? 2 : c == 2 ? 3 : c == 3 ? 1 : c == 4 ? 3 : c == 5 ? 2 : 0;
  if (channels == 8) return c == 0 ? 1 : c == 1 ? 2 : c == 2 ? 3 : c == 3 ? 1 : c == 4 ? 3 : c == 5 ? 1 : c == 6 ? 3 : 0;
  return 0;
}

// rough log2(2^a + 2^b), Q10 (:186-218)
CB_FN i32 surround_logsum(i32 a, i32 b) {
  const i32 mx = a > b ? a : b;
  const i32 diff = a > b ? a - b : b - a;
  if (!(diff < 8192)) return mx;
  const int low = diff >> (CB_DB_SHIFT - 1);
  const i32 frac = (i16)shl32((i16)(diff - (low << (CB_DB_SHIFT - 1))), 16 - CB_DB_SHIFT);
  // diff_table[17] = {.5, .2924813, .1609640, .0849625, .0437314, .0221971, .0111839, .0056136, .0028123, 0...} in Q10
  const i32 t0 = low == 0 ? 512 : low == 1 ? 300 : low == 2 ? 165 : low == 3 ? 87 : low == 4 ? 45 : low == 5 ? 23
               : low == 6 ? 11 : low == 7 ? 6 : low == 8 ? 3 : 0;
  const i32 t1 = low == 0 ? 300 : low == 1 ? 165 : low == 2 ? 87 : low == 3 ? 45 : low == 4 ? 23 : low == 5 ? 11
               : low == 6 ? 6 : low == 7 ? 3 : 0;
  return (i16)(mx + t0 + mult16_16_q15(frac, (i16)(t1 - t0)));
}

// One channel: pcm interleaved by `channels`, len samples per channel at rate Fs.  Writes the 21 spread band
// levels to bandLogE and updates this channel's window / pre-emphasis memory.  All lanes of one warp.
CB_FN void surround_channel(const i16* pcm, int channels, int c, int len, int Fs, SurroundState* st, SurroundScratch& sc,
                            i16* bandLogE) {
  const int upsample = 48000 / Fs;
  const int frame_size = len * upsample;
  int LM;
  for (LM = 0; LM < CB_MAX_LM; LM++)
    if (CB_SHORT_MDCT << LM == frame_size) break;          // no match (>= 20 ms): the 20 ms transform
  const int NB = CB_SHORT_MDCT << LM;
  CB_WSYNC();
  CB_LANE_LOOP(i, CB_OVERLAP) sc.in[i] = st->window_mem[c][i];
  i32 mem = st->preemph_mem[c];
  CB_WSYNC();
  enc_preemphasis(pcm + c, sc.in + CB_OVERLAP, frame_size, channels, upsample, &mem);
  enc_mdct_forward(sc.in, sc.freq, sc.f2, 1, NB, CB_MAX_LM - LM);
  if (upsample != 1) {
    CB_LANE_LOOP(i, NB) sc.freq[i] = (i < len) ? sc.freq[i] * upsample : 0;
    CB_WSYNC();
  }
  enc_band_energies(sc.freq, sc.freq, sc.bandE, CB_NBANDS, 1, LM);
  enc_amp2log2(CB_NBANDS, CB_NBANDS, sc.bandE, bandLogE, 1);
  CB_IF_LANE0 {
    // spreading: -6 dB per band going up, -12 dB per band going down (:290-294)
    for (int i = 1; i < CB_NBANDS; i++) bandLogE[i] = (i16)max32(bandLogE[i], (i16)(bandLogE[i - 1] - 1024));
    for (int i = CB_NBANDS - 2; i >= 0; i--) bandLogE[i] = (i16)max32(bandLogE[i], (i16)(bandLogE[i + 1] - 2048));
    st->preemph_mem[c] = mem;
  }
  CB_LANE_LOOP(i, CB_OVERLAP) st->window_mem[c][i] = sc.in[frame_size + i];
  CB_WSYNC();
}

// bandLogE: [channels][21] levels from surround_channel; replaced by the masks (signal-to-mask ratios).
CB_FN void surround_combine(i16* bandLogE, int channels) {
  const i32 channel_offset = (i16)((i16)cb_log2(32768 / (channels - 1)) >> 1);
  CB_WSYNC();
  CB_LANE_LOOP(i, CB_NBANDS) {
    i32 mL = -28672, mR = -28672;
    for (int c = 0; c < channels; c++) {
      const int pos = surround_channel_pos(channels, c);
      const i32 e = bandLogE[CB_NBANDS * c + i];
      if (pos == 1) mL = surround_logsum(mL, e);
      else if (pos == 3) mR = surround_logsum(mR, e);
      else if (pos == 2) {
        mL = surround_logsum(mL, (i16)(e - 512));
        mR = surround_logsum(mR, (i16)(e - 512));
      }
    }
    i32 mC = min32(mL, mR);
    mL = (i16)(mL + channel_offset);
    mC = (i16)(mC + channel_offset);
    mR = (i16)(mR + channel_offset);
    for (int c = 0; c < channels; c++) {
      const int pos = surround_channel_pos(channels, c);
      const i32 m = pos == 1 ? mL : pos == 2 ? mC : mR;
      bandLogE[CB_NBANDS * c + i] = pos ? (i16)(bandLogE[CB_NBANDS * c + i] - m) : (i16)0;
    }
  }
  CB_WSYNC();
}
