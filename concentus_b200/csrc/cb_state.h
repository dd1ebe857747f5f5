// cb_state.h -- per-stream decoder state as laid out in HBM, and the per-frame hand-off record
// between the entropy kernel (A) and the synthesis kernel (B).
//
// Counterpart of struct OpusCustomDecoder (opus-fix/celt/celt_decoder.c:67-100), re-laid-out:
//   * the "small" state (coder seed, post-filter parameters, band energies) is owned by kernel A;
//   * the "signal" state (output history, overlap tail, de-emphasis memory) is owned by kernel B.
// The reference keeps a linear 2048+120 sample buffer per channel and memmoves it by N every
// frame (celt_decoder.c:963).  Here the history is a 2048-sample ring (hist_pos = index of the
// oldest sample) plus a 60-sample overlap tail, so a frame costs N stores instead of ~2048.
#pragma once
#include "cb_common.h"
#include "cb_silk_common.h"
#include <stddef.h>

#define CB_NBANDS 21
#define CB_MAXN 960          // samples per channel per frame at LM=3
#define CB_OVERLAP 120
#define CB_SHORT_MDCT 120   // shortMdctSize
#define CB_MAX_LM 3
#define CB_HIST 2048         // DECODE_BUFFER_SIZE
#define CB_LPC_ORDER 24
#define CB_NORM_STRIDE 624   // 8*eBands[20]

struct alignas(16) CeltDecState {
  // ---- small state (kernel A) ----
  u32 rng;                   // celt_decoder.c:81
  i32 error;
  i32 last_pitch_index;
  i32 loss_count;
  i32 pf_period, pf_period_old;
  i32 pf_gain, pf_gain_old;          // opus_val16 values
  i32 pf_tapset, pf_tapset_old;
  i32 channels;              // CC (output channels), fixed at create
  i32 downsample;            // 48000/Fs
  i16 oldBandE[2 * CB_NBANDS];
  i16 oldLogE[2 * CB_NBANDS];
  i16 oldLogE2[2 * CB_NBANDS];
  i16 backgroundLogE[2 * CB_NBANDS];
  i16 lpc[2 * CB_LPC_ORDER];
  // ---- signal state (kernel B) ----
  i32 preemph_mem[2];
  i32 hist_pos[2];           // ring index of the oldest sample, per channel (each channel warp owns one)
  i32 tail[2][CB_OVERLAP / 2];        // un-windowed IMDCT overlap, decode_mem[2048..2048+60)
  i32 hist[2][CB_HIST];               // decode_mem[0..2048), ring
};

// What kernel A tells kernel B about one frame (lives in the scratch arena, one per stream).
struct alignas(16) FrameSide {
  i32 kind;                  // 0 = nothing to synthesise, 1 = normal frame
  i32 LM, C, transient, silence, start, effEnd;
  i32 T0, T1, T2;            // post-filter periods: old, current, new (celt_decoder.c:1003-1011)
  i32 g0, g1, g2;
  i32 tap0, tap1, tap2;
  i32 accum;                 // hybrid: add into pcm (celt_decoder.c:236-245)
  i32 route;                 // copy of StreamPlan::route for this pass (the rebuild and synthesis kernels read it here:
                             //   the plan's field is rewritten by the next pass's routing while they may still run)
  i32 pad[2];
  i16 bandE[2 * CB_NBANDS];  // log energies used by denormalise_bands
  i16 pad2[6];
  i16 X[2 * CB_MAXN];        // normalised spectrum, channel-major
};

// The tape: what the PARSE half of a CELT frame (everything that reads the bitstream; one thread or one warp per
// stream) leaves for the REBUILD half (one warp per stream: pulse vectors, folding, recombination, stereo merge,
// anti-collapse).  It lives in the scratch arena, one per stream (cb_celt_bands3.h describes the entries).
// Sizes are the worst case of a 20 ms stereo frame: a band's channel splits into at most 16 leaves (bands of
// 32+ bins at LM = 3), 8 for 16 bins, 4 for 8 bins: 208 leaves per channel.
#define CB_TAPE_MAX_JOBS 416
#define CB_TAPE_MAX_SEG 24
#define CB_TAPE_NONE 0             // nothing for the rebuild half
#define CB_TAPE_FRAME 1            // a decoded frame
#define CB_TAPE_NOISE 2            // noise-based concealment: random normalised bands at the energies in bandE
struct TapeJob {                   // a leaf of a band's split tree, in the order the walk reaches them
  u32 idx;                         // K > 0: PVQ index (alg_unquant after the index has been read)
                                   // K == 0, a leaf without pulses (zeros / noise / folded spectrum): 1 = the
                                   //   angles above it leave it nothing to fold (its share of `fill` is masked off)
  u16 xoff;                        // where the leaf lies in X
  i16 N, K, gain;
  u8 B, spread;
  u8 sh;                           // where the leaf's collapse bits sit in the band's mask (and in `fill`)
  u8 pad;
};
// What the rebuild half needs to know of one band (quant_all_bands / quant_band_stereo, bands.c:1176-1502) beyond
// its leaves: everything here is decided by the bitstream alone, so the parse half works it out once.
#define CB_TB_MONO 0               // one call on X
#define CB_TB_DUAL 1               // dual stereo: X and Y coded apart
#define CB_TB_JOINT 2              // mid / side, N > 2
#define CB_TB_JOINT2 3             // mid / side, N == 2 (bands.c:1222-1262)
#define CB_TB_N1 4                 // stereo, N == 1: two signs
#define CB_TBF_INV 1               // side channel inverted
#define CB_TBF_MID_FIRST 2         // joint stereo: the mid call came first in the walk (else the side call)
#define CB_TBF_N2_NEG 4            // N == 2: sign of the side
#define CB_TBF_SWITCH 8            // the band where dual stereo ends: the two folding memories are averaged (bands.c:1430-1437)
#define CB_TBF_SIGN0 16            // N == 1: sign of the first / second channel
#define CB_TBF_SIGN1 32
struct alignas(8) TapeBand {
  u16 job0;                        // the band's first leaf record (index into FrameTape::jobs)
  u8 nl[2];                        // leaves of the band's first and second call, in walk order; + 128: one of them
                                   //   is without pulses
  i16 eff_lowband;                 // folding source (offset into the folding memory), -1: none
  u8 fold0, fold1;                 // the bands [fold0, fold1) whose collapse masks the folding source brings
  u8 mode, flags;                  // CB_TB_*, CB_TBF_*
  i16 itheta, imid, iside;         // joint stereo: the split angle and its cosines
};
struct TapeSeg {                   // a run of bands whose leaf masks fit the rebuild warp's shared memory
  u8 band0, band1;
  u16 jobs;                        // leaf records of the segment (they follow the previous segment's)
};
struct alignas(16) FrameTape {
  i32 kind;                        // CB_TAPE_*
  i32 nseg;
  u32 seed;                        // celt_lcg_rand state at the start of the frame
  i32 LM, C, start, end, B, spread, intensity, dual_stereo, anti_collapse_on, silence;
  i16 pulses[CB_NBANDS + 1];       // allocation per band (anti-collapse depth)
  i8 tf_res[CB_NBANDS + 3];
  i16 bandE[2 * CB_NBANDS];        // final log energies of the frame
  i16 prevLogE[2][2 * CB_NBANDS];  // oldLogE / oldLogE2 as they were before this frame (anti_collapse)
  TapeSeg seg[CB_TAPE_MAX_SEG];
  TapeBand band[CB_NBANDS];
  TapeJob jobs[CB_TAPE_MAX_JOBS];
  i16 norm[2 * CB_NORM_STRIDE];    // REBUILD's own scratch: the normalised lower bands kept for folding (bands.c:1380)
};

// celt_decoder_init + OPUS_RESET_STATE (celt_decoder.c:134-172,1196-1204): all-zero except the
// two log-energy memories, which start at -28 dB.  Host-side helper (the blob is then copied to HBM).
CB_HD void celt_state_reset(CeltDecState* s, int channels, int downsample) {
  unsigned char* p = (unsigned char*)s;
  for (size_t i = 0; i < sizeof(CeltDecState); i++) p[i] = 0;
  s->channels = channels;
  s->downsample = downsample;
  for (int i = 0; i < 2 * CB_NBANDS; i++) s->oldLogE[i] = s->oldLogE2[i] = -28672;
}

// ---- Opus layer ---------------------------------------------------------------------------
#define CB_MODE_SILK_ONLY 1000
#define CB_MODE_HYBRID 1001
#define CB_MODE_CELT_ONLY 1002
#define CB_BW_NARROWBAND 1101
#define CB_BW_MEDIUMBAND 1102
#define CB_BW_WIDEBAND 1103
#define CB_BW_SUPERWIDEBAND 1104
#define CB_BW_FULLBAND 1105
#define CB_OK 0
#define CB_BAD_ARG -1
#define CB_BUFFER_TOO_SMALL -2
#define CB_INTERNAL_ERROR -3
#define CB_INVALID_PACKET -4
#define CB_UNIMPLEMENTED -5
#define CB_INVALID_STATE -6
#define CB_ALLOC_FAIL -7

// Per-stream decoder blob in HBM: struct OpusDecoder (opus-fix/src/opus_decoder.c:55-79) followed
// by the CELT state.  The Opus-layer fields live on the device so that a batch call needs no
// per-packet work on the host (TOC/framing is parsed by the plan kernel).
struct alignas(16) OpusDecState {
  i32 Fs, channels, decode_gain;
  i32 stream_channels, bandwidth, mode, prev_mode, frame_size, prev_redundancy, last_packet_duration;
  u32 rangeFinal;
  i32 pad;
  CeltDecState celt;
  SilkDecState silk;            // silk_dec + the persistent part of DecControl (opus_decoder.c:57,63)
};
// opus_decoder_init + OPUS_RESET_STATE (opus_decoder.c:96-137,833-843)
static inline void opus_state_reset(OpusDecState* s, int Fs, int channels, int keep_gain) {
  int gain = keep_gain ? s->decode_gain : 0;
  celt_state_reset(&s->celt, channels, 48000 / Fs);
  s->Fs = Fs;
  s->channels = channels;
  s->decode_gain = gain;
  s->stream_channels = channels;
  s->bandwidth = 0;
  s->mode = 0;
  s->prev_mode = 0;
  s->frame_size = Fs / 400;
  s->prev_redundancy = 0;
  s->last_packet_duration = 0;
  s->rangeFinal = 0;
  s->pad = 0;
  sk_init_decoder(s->silk);
  if (!keep_gain) {               // opus_decoder_init clears the whole struct; OPUS_RESET_STATE keeps DecControl
    s->silk.nChannelsAPI = s->silk.nChannelsInternal = 0;
    s->silk.ctl_nChannelsInternal = s->silk.ctl_internalSampleRate = s->silk.ctl_payloadSize_ms = s->silk.ctl_prevPitchLag = 0;
  }
}

// What the plan kernel decides for one stream of a batch call: which frames each later pass
// decodes (offset/length in the packet arena), or which concealment chunks it synthesises.
#define CB_MAX_FRAMES 48
#define CB_PLAN_NONE 0      // error / nothing to do
#define CB_PLAN_DECODE 1    // frames of a packet
#define CB_PLAN_PLC 2       // concealment chunks (packet loss / DTX / FEC request on CELT)
#define CB_PLAN_ZERO 3      // loss before any packet was seen: zeros (opus_decoder.c:261-268)
struct alignas(16) StreamPlan {
  i32 ret;                  // value opus_decode() returns for this stream
  i32 nframes;
  i32 kind;
  i32 total_samples;        // samples per channel covered by the plan
  i32 fec_frame;            // index of the frame decoded from in-band FEC (lost_flag 2), or -1
  i32 route;                // who decodes the current pass's frame (CB_ROUTE_*), set by the SILK kernel at pass start
  alignas(8) u32 rd_save[16];  // range-decoder state after the SILK layer, handed to the CELT / general kernel
  i32 off[CB_MAX_FRAMES];   // payload offset in the arena
  i16 len[CB_MAX_FRAMES];   // payload bytes (<=1 means "conceal this frame")
  i16 nsamp[CB_MAX_FRAMES]; // samples per channel produced (API rate)
  i32 pcm_off[CB_MAX_FRAMES];  // first sample of the frame in the stream's PCM row (48 concealment pieces of
                               //   20 ms reach 46080 samples: wider than 16 bits)
};

// ---- multistream (surround) ------------------------------------------------------------------
// struct OpusMSDecoder / ChannelLayout (opus-fix/src/opus_multistream_decoder.c:45-48,
// src/opus_private.h:41-46) as a device record: which elementary decoders belong to the
// multistream decoder and how their channels map to the output channels.
struct alignas(16) MSLayout {
  i32 nb_channels, nb_streams, nb_coupled;
  i32 first;                 // index of the first elementary stream in the group
  i32 Fs;
  i32 pad[3];
  u8 mapping[256];
};
