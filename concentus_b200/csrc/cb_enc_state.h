// cb_enc_state.h -- CELT / Opus encoder state in HBM and the per-stream work areas.
//
// CeltEncState restates struct OpusCustomEncoder (opus-fix/celt/celt_encoder.c:60-128) with the
// trailing variable-size arrays made fixed (in_mem, prefilter_mem, oldBandE/oldLogE/oldLogE2).
// OpusEncState restates the fields of struct OpusEncoder (src/opus_encoder.c:63-109) that the
// CELT-only path uses.  EncScratch holds the large time-domain buffers of one frame in global
// memory (L2 resident); EncWork is the per-warp shared-memory workspace.
#pragma once
#include "cb_common.h"
#include "cb_state.h"
#include "cb_rangeenc.h"

#define CB_COMBFILTER_MAXPERIOD 1024
#define CB_COMBFILTER_MINPERIOD 15
#define CB_MAX_ENCODER_BUFFER 480
#define CB_OPUS_AUTO (-1000)
#define CB_OPUS_BITRATE_MAX (-1)
#define CB_APP_VOIP 2048
#define CB_APP_AUDIO 2049
#define CB_APP_LOWDELAY 2051

// Scalar part of the CELT encoder state.  A frame works on a register/local copy of it (every lane
// holds the same values) and lane 0 stores it back at the end, so no lane ever re-reads a field that
// another lane has already updated.
struct CeltEncScalars {
  i32 channels, stream_channels, force_intra, clip, disable_pf, complexity, upsample, start, end;
  i32 bitrate, vbr, constrained_vbr, loss_rate, lsb_depth, variable_duration, lfe;
  // ---- cleared by OPUS_RESET_STATE (celt_encoder.c:2447-2467) ----
  u32 rng;
  i32 spread_decision, delayedIntra, tonal_average, lastCodedBands, hf_average, tapset_decision;
  i32 prefilter_period, prefilter_gain, prefilter_tapset, consec_transient;
  i32 preemph_memE[2];
  i32 vbr_reservoir, vbr_drift, vbr_offset, vbr_count;
  i32 overlap_max, stereo_saving, intensity, spec_avg;
  i32 has_mask;                   // energy_mask != NULL (celt_encoder.c:115, opus_encoder.c:103): energy_mask[] below is valid
};
struct alignas(16) CeltEncState {
  CeltEncScalars s;
  i32 in_mem[2][CB_OVERLAP];
  i32 prefilter_mem[2][CB_COMBFILTER_MAXPERIOD];
  i16 oldBandE[2 * CB_NBANDS], oldLogE[2 * CB_NBANDS], oldLogE2[2 * CB_NBANDS];
  // Surround masking of this stream's channels, [channel][band], Q10 (celt_encoder.c:115): the reference keeps a
  // pointer to the caller's array, here the values are copied in (by the ctl, or by the surround analysis
  // kernel of the multistream encoder).  OPUS_RESET_STATE drops the mask (has_mask = 0), as it clears the pointer.
  i16 energy_mask[2 * CB_NBANDS];
};

struct OpusEncScalars {
  i32 application, channels, delay_compensation, force_channels, signal_type, user_bandwidth, max_bandwidth;
  i32 user_forced_mode, voice_ratio, Fs, use_vbr, vbr_constraint, variable_duration, bitrate_bps;
  i32 user_bitrate_bps, lsb_depth, encoder_buffer, lfe, complexity, packet_loss_perc;
  i32 use_inband_fec, use_dtx, reduced_dependency;      // silk_mode fields the mode decision reads
  // ---- cleared by OPUS_RESET_STATE (opus_encoder.c:2440-2465) ----
  i32 stream_channels, hybrid_stereo_width_Q14, variable_HP_smth2_Q15, prev_HB_gain;
  i32 hp_mem[4];
  i32 mode, prev_mode, prev_channels, prev_framesize, bandwidth, silk_bw_switch, first;
  i32 width_XX, width_XY, width_YY, width_smoothed, width_max_follower;   // StereoWidthState
  i32 silk_stereoWidth_Q14;
  u32 rangeFinal;
};
struct alignas(16) OpusEncState {
  OpusEncScalars o;
  i16 delay_buffer[CB_MAX_ENCODER_BUFFER * 2];
  CeltEncState celt;
};

// celt OPUS_RESET_STATE + init defaults (celt_encoder.c:165-225, 2447-2467)
static inline void celt_enc_reset(CeltEncState* st) {
  const size_t keep = offsetof(CeltEncScalars, rng);
  unsigned char* p = (unsigned char*)st;
  for (size_t i = keep; i < sizeof(CeltEncState); i++) p[i] = 0;
  for (int i = 0; i < 2 * CB_NBANDS; i++) st->oldLogE[i] = st->oldLogE2[i] = -28672;
  st->s.vbr_offset = 0;
  st->s.delayedIntra = 1;
  st->s.spread_decision = 2;
  st->s.tonal_average = 256;
  st->s.hf_average = 0;
  st->s.tapset_decision = 0;
}
static inline void celt_enc_init(CeltEncState* st, int Fs, int channels) {
  unsigned char* p = (unsigned char*)st;
  for (size_t i = 0; i < sizeof(CeltEncState); i++) p[i] = 0;
  CeltEncScalars* s = &st->s;
  s->channels = s->stream_channels = channels;
  s->upsample = 48000 / Fs;
  s->start = 0;
  s->end = CB_NBANDS;
  s->constrained_vbr = 1;
  s->clip = 1;
  s->bitrate = CB_OPUS_BITRATE_MAX;
  s->vbr = 0;
  s->force_intra = 0;
  s->complexity = 5;
  s->lsb_depth = 24;
  celt_enc_reset(st);
}
// opus_encoder_init (opus_encoder.c:176-252) / OPUS_RESET_STATE (:2440-2465)
static inline void opus_enc_reset(OpusEncState* st) {
  const size_t a = offsetof(OpusEncScalars, stream_channels), b = offsetof(OpusEncState, celt);
  unsigned char* p = (unsigned char*)st;
  for (size_t i = a; i < b; i++) p[i] = 0;
  celt_enc_reset(&st->celt);
  OpusEncScalars* s = &st->o;
  s->stream_channels = s->channels;
  s->hybrid_stereo_width_Q14 = 1 << 14;
  s->prev_HB_gain = 32767;
  s->first = 1;
  s->mode = CB_MODE_HYBRID;
  s->bandwidth = CB_BW_FULLBAND;
  s->variable_HP_smth2_Q15 = 756 << 8;   // silk_lin2log(VARIABLE_HP_MIN_CUTOFF_HZ = 60) << 8 (opus_encoder.c:242, 2442)
  s->silk_stereoWidth_Q14 = 0;
}
static inline void opus_enc_init(OpusEncState* st, int Fs, int channels, int application) {
  unsigned char* p = (unsigned char*)st;
  for (size_t i = 0; i < sizeof(OpusEncState); i++) p[i] = 0;
  OpusEncScalars* s = &st->o;
  s->channels = channels;
  s->Fs = Fs;
  s->complexity = 9;
  celt_enc_init(&st->celt, Fs, channels);
  st->celt.s.complexity = 9;
  s->use_vbr = 1;
  s->vbr_constraint = 1;
  s->user_bitrate_bps = CB_OPUS_AUTO;
  s->bitrate_bps = 3000 + Fs * channels;
  s->application = application;
  s->signal_type = CB_OPUS_AUTO;
  s->user_bandwidth = CB_OPUS_AUTO;
  s->max_bandwidth = CB_BW_FULLBAND;
  s->force_channels = CB_OPUS_AUTO;
  s->user_forced_mode = CB_OPUS_AUTO;
  s->voice_ratio = -1;
  s->encoder_buffer = Fs / 100;
  s->lsb_depth = 24;
  s->variable_duration = 5000;    // OPUS_FRAMESIZE_ARG
  s->delay_compensation = Fs / 250;
  opus_enc_reset(st);
}

// Large per-frame buffers of one stream (global memory).
struct alignas(16) EncScratch {
  i32 in[2][CB_MAXN + CB_OVERLAP];                       // pre-emphasised (then pre-filtered) signal
  // The pre-filter's input is dead once run_prefilter has written `in` and the filter memory back, and the
  // MDCT buffers are first written after that: they share storage (11.5 KB less scratch per stream; the
  // scratch of a 4096-stream batch is larger than the L2, so its size is DRAM traffic).
  union {
    i32 pre[2][CB_MAXN + CB_COMBFILTER_MAXPERIOD];       // pre-filter input with history
    struct {
      i32 freq[2][CB_MAXN];                              // MDCT coefficients
      i32 f2[CB_MAXN];                                   // forward-MDCT FFT buffer
    };
  };
  i32 xcorr[(CB_COMBFILTER_MAXPERIOD >> 1) + 8];
  i32 yy_lookup[(CB_COMBFILTER_MAXPERIOD >> 1) + 8];
  i32 ac[CB_LPC_ORDER + 1];
  i32 lpc32[CB_LPC_ORDER];
  i16 pcm_buf[(CB_MAX_ENCODER_BUFFER + CB_MAXN) * 2];    // delay-compensated, dc-rejected input
  u8 intra_bits[256];                                    // coarse-energy two-pass byte save
  u8 multi_tmp[3 * 1276 + 4];                            // the 20 ms packets of a 40 / 60 ms frame before assembly
};

// Phase rendezvous.  All warps of a CTA run the same stage of the frame at the same time, so the
// SM's instruction caches hold one stage's code instead of every stage's: the encoder is ~18k SASS
// instructions and instruction-fetch stalls dominated when every warp sat in a different stage.
// No data crosses warps; the barrier is only pacing.  Every warp must reach CB_ENC_PHASES barriers
// per frame: a warp that leaves early (error, silence) makes up the remainder in the kernel wrapper.
#define CB_ENC_PHASES (11 + 2 * CB_NBANDS)  // eleven stage boundaries + two per band of the residual loop
#if CB_DEVICE_BUILD
#define CB_ENC_PHASE(w) do { __syncthreads(); if (CB_LANE() == 0) (w).nsync++; __syncwarp(); } while (0)
#else
#define CB_ENC_PHASE(w) ((void)0)
#endif

// Serial sections.  Two parts of a frame are recurrences over its samples that cannot be reassociated bit-exactly:
// the input high-pass (dc_reject, opus_encoder.c:362-384) and the transient detector's filter and smoothing passes
// (celt_encoder.c:227-330).  Run by each warp for itself they keep one lane per channel busy for thousands of dependent
// steps -- 12 % of the kernel's warp instructions.  Since the CTA's warps are at the same stage anyway (phase
// rendezvous above), a warp instead posts a request, and after the next rendezvous the lanes of warps 0 and 1 run the
// chains of all the CTA's streams side by side (lane = stream slot x channel); the rendezvous after that hands the
// results back.  The warps that run the chains look for requests after EVERY rendezvous they pass, also while they are
// only making up rendezvous for a stream that ended early (cb_enc_kernels.cu).
#define CB_ES_NONE 0
#define CB_ES_DC 1                 // dc_reject
#define CB_ES_TRANSIENT 2          // transient_analysis: high-pass, forward / backward smoothing
struct EncSerialReq {
  const void* in;                  // DC: int16 PCM in (interleaved); transient: i32 signal, `stride_in` per channel
  void* out;                       // DC: int16 out (interleaved); transient: i16 work vector, `stride_out` per channel
  i32 kind, len, channels, shift;  // (shift: DC only)
  i32 stride_in, stride_out;
  i32 m[4];                        // DC: the filter memories, in and out; transient: out, mean[0..1] then maxE[0..1]
};

// One level of the encoder's split recursion (quant_partition).
struct EncPartFrame {
  i16* X;
  i32 N, b, B, LM, stage, mbits, sbits, itheta, rebal0;
};
struct EncSplit { int inv, delta, itheta, qalloc, b; };

// Per-warp shared-memory workspace of the encode kernel.  The first block is a union over the
// stages of a frame: the pitch analysis and the transient detector run before anything of the band
// stage exists, and their vectors are re-read many times, so they borrow the band stage's room
// instead of living in global memory.
struct alignas(16) EncWork {
  union {
    struct {                       // MDCT onwards
      i16 X[2 * CB_MAXN];          // normalised spectrum
      i32 bandE[2 * CB_NBANDS];
      i32 tmp32[4 * CB_NBANDS + 4];  // allocation scratch
      i32 pulses[CB_NBANDS];
      i32 offsets[CB_NBANDS];
      i32 cap[CB_NBANDS];
      i32 fine_quant[CB_NBANDS];
      i32 fine_prio[CB_NBANDS];
      i32 tf_res[CB_NBANDS];
      i32 metric[CB_NBANDS], path0[CB_NBANDS], path1[CB_NBANDS];
    };
    struct {                       // pitch pre-filter analysis
      i16 pitch_buf[(CB_COMBFILTER_MAXPERIOD + CB_MAXN) >> 1];
      union {
        i16 pitch_xx[(CB_COMBFILTER_MAXPERIOD + CB_MAXN) >> 1];   // whitening temp, dead before the search
        struct {
          i16 pitch_x4[CB_MAXN >> 2];
          i16 pitch_y4[((CB_MAXN + CB_COMBFILTER_MAXPERIOD) >> 2) + 8];
        };
      };
    };
    i16 ta_tmp[2][CB_MAXN + CB_OVERLAP];   // transient detector, per channel
  };
  i16 bandLogE[2 * CB_NBANDS], bandLogE2[2 * CB_NBANDS], error[2 * CB_NBANDS];
  i16 oldE[2 * CB_NBANDS];      // working copy of the state's oldBandE
  i16 oldE_intra[2 * CB_NBANDS], error_intra[2 * CB_NBANDS];
  i16 follower[2 * CB_NBANDS], noise_floor[2 * CB_NBANDS], surround_dynalloc[2 * CB_NBANDS];
  i16 vtmp[176], vtmp1[176];     // tf_analysis copies / alg_quant y
  i16 iy[176], signx[176];
  EncPartFrame stack[5];
  EncSplit sp_store;
  i32 nsync;                     // phase rendezvous points passed so far (CB_ENC_PHASE)
  EncSerialReq* sreq;            // device: this warp's slot in the CTA's table of serial-section requests
  RangeEnc enc, enc_save0, enc_save1;
};
