// cb_kernels.cu -- sm_100a kernels of the batched Opus decoder and their launchers.
//
//   (the plan -- TOC + framing parse, argument checks, StreamPlan -- is fused with the SILK kernel: cb_silk_kernels.cu)
//   celt_parse_kernel   : 1 thread / stream  -- everything that reads the bitstream (cb_parse_kernels.cu, kernel A1)
//   celt_rebuild_kernel : 1 warp   / stream  -- PVQ pulse vectors + band reconstruction from the tape (kernel A2)
//   celt_synth_kernel   : 1 warp   / (stream, channel) -- IMDCT, post-filter, de-emphasis (kernel B)
//
// They are launched once per frame index ("pass") of the packets in the batch; single-frame
// packets (the normal case) need one pass.  All kernels are integer / latency bound (no tensor
// cores): the design levers are warps resident per SM (28, sized from the 227 KB shared memory) and keeping
// the per-stream arrays the warp works on in shared memory.
#include <cuda_runtime.h>
#include <stdint.h>
#include "cb_decode_pass.h"
#include "cb_kernels.h"

extern "C" int cbk_sm_count(void);   // SMs of the current device (opus_api.cu)

#define A_WARPS 4           // kernel A, small batches: 4 warps per CTA, 7 CTAs per SM
#define A_WARPS_SM 28       // rebuild kernel: warps an SM holds
#define B_WARPS 7           // synthesis kernel: 7 warps per CTA, 4 CTAs per SM (shared memory: 7.9 KB per warp)
#define B_CTAS_SM 4

// Kernel A2: the REBUILD half of the entropy stage (cb_entropy_pass.h), one warp per stream: tape -> pulse vectors
// (a lane per leaf) -> folding / recombination / stereo merge -> normalised spectrum for the synthesis kernel.
// CELT-only frames and the CELT layer of hybrid frames alike (the tape says where the bands start).
// Registers and shared memory are held to 64 per thread / ~5.4 KB per warp on purpose: the parse kernel of the NEXT
// call runs beside this kernel (opus_api.cu: the control side runs a pass ahead), and its CTAs only fit on an SM
// whose rebuild CTAs leave them room.
__global__ void __launch_bounds__(1024, 1)
celt_rebuild_kernel(FrameSide* sides, FrameTape* tapes, int n) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int warp = threadIdx.x >> 5;
  const int s = blockIdx.x * (blockDim.x >> 5) + warp;
  if (s >= n) return;
  const int route = sides[s].route;
  if ((route != CB_ROUTE_CELT && route != CB_ROUTE_HYBRID) || tapes[s].kind == CB_TAPE_NONE) return;
  EntWork& w = *reinterpret_cast<EntWork*>(smem + (size_t)warp * sizeof(EntWork));
  decode_entropy_rebuild(w, &tapes[s], &sides[s]);
}

// Kernel B.  The de-emphasis of a frame is a serial recurrence over its samples (one lane busy for 960 steps), so
// the warps of a CTA do not each run their own: they park the finished frame (EmitReq), and the lanes of warp 0 run
// the de-emphasis of all of them side by side; each warp then converts and stores its own PCM.
__global__ void __launch_bounds__(B_WARPS * 32, B_CTAS_SM)
celt_synth_kernel(OpusDecState* const* states, const StreamPlan* plans, const FrameSide* sides, i16* pcm,
                  i32 pcm_stride, int n, int pass, int max_ch) {
  extern __shared__ __align__(16) unsigned char smem[];
  EmitReq* reqs = reinterpret_cast<EmitReq*>(smem + (size_t)B_WARPS * sizeof(SynWork));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gw = blockIdx.x * B_WARPS + warp;
  const int s = max_ch == 2 ? gw >> 1 : gw, c = max_ch == 2 ? gw & 1 : 0;   // a warp per channel the group can have
  if (lane == 0) reqs[warp].valid = 0;
  __syncwarp();
  if (s < n) {
    OpusDecState* st = states[s];
    if (c < st->channels) {
      SynWork& w = *reinterpret_cast<SynWork*>(smem + (size_t)warp * sizeof(SynWork));
      decode_synth_pass(w, st, &plans[s], &sides[s], pcm + (size_t)s * pcm_stride, c, pass, &reqs[warp]);
    }
  }
  __syncthreads();
  if (warp == 0 && lane < B_WARPS && reqs[lane].valid) {
    const EmitReq r = reqs[lane];
    emit_deemph(r.buf, r.N, r.st, r.c);
  }
  __syncthreads();
  const EmitReq r = reqs[warp];
  if (r.valid) emit_pcm(r.buf, r.N, r.st, r.c, r.pcm, r.accum, r.decode_gain);
}

// Multistream: one thread per multistream decoder plans its elementary streams.
__global__ void opus_ms_plan_kernel(OpusDecState* const* states, StreamPlan* plans, const MSLayout* layouts,
                                    const u8* data, const i32* offs, const i32* lens, const i32* fec,
                                    int frame_size, i32* ms_ret, int n_ms, int max_passes) {
  int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= n_ms) return;
  ms_ret[m] = opus_ms_plan(states, plans, &layouts[m], data, offs[m], lens[m], frame_size, fec ? fec[m] : 0, max_passes);
}

// Multistream: channel mapping copy-out (opus_multistream_decoder.c:262-299).  One CTA per
// multistream decoder; elementary PCM rows (interleaved, 1 or 2 channels) -> output row.
__global__ void opus_ms_scatter_kernel(const MSLayout* layouts, const i32* ms_ret, const i16* epcm,
                                       i32 epcm_stride, i16* pcm, i32 pcm_stride, int n_ms) {
  const int m = blockIdx.x;
  if (m >= n_ms) return;
  const MSLayout* L = &layouts[m];
  const int frames = ms_ret[m];
  if (frames <= 0) return;
  const int CH = L->nb_channels;
  i16* dst = pcm + (size_t)m * pcm_stride;
  for (int idx = threadIdx.x; idx < frames * CH; idx += blockDim.x) {
    const int t = idx / CH, c = idx - t * CH;
    const int mp = L->mapping[c];
    i16 v = 0;
    if (mp != 255) {
      if (mp < 2 * L->nb_coupled) {
        const i16* src = epcm + (size_t)(L->first + (mp >> 1)) * epcm_stride;
        v = src[t * 2 + (mp & 1)];
      } else {
        const i16* src = epcm + (size_t)(L->first + mp - L->nb_coupled) * epcm_stride;
        v = src[t];
      }
    }
    dst[idx] = v;
  }
}

// ---- launchers (host) ------------------------------------------------------------------------
static int g_attr_done[64];

// Warps per CTA of the rebuild kernel.  An SM holds 28 of its warps (shared memory).  A batch that gives every SM at
// least a few warps runs as CTAs that each take an equal share of it -- two CTAs per SM, ceil(n / (2 #SMs)) warps
// each (the same over two, three ... waves when the batch is larger than #SMs x 28) -- so that every SM is busy for
// the whole kernel.  Smaller batches keep 4-warp CTAs and spread over the SMs.
// (The one-kernel entropy stage of round 1 had to pace its warps through the frame's stages with CTA-wide barriers
// to keep 7 k instructions in the instruction caches; the rebuild half alone is small enough to go without.)
static int entropy_cta_warps(int n) {
  const int sms = cbk_sm_count();
  if (n < sms * (A_WARPS + 1)) return A_WARPS;
  const int waves = (n + sms * A_WARPS_SM - 1) / (sms * A_WARPS_SM);
  const int w = (n + 2 * sms * waves - 1) / (2 * sms * waves);
  return w > A_WARPS_SM / 2 ? A_WARPS_SM / 2 : (w < 1 ? 1 : w);
}

static int ensure_attrs() {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 64 && g_attr_done[dev]) return 0;
  e = cudaFuncSetAttribute(celt_rebuild_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if (e != cudaSuccess) return (int)e;
  e = cudaFuncSetAttribute(celt_synth_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if (e != cudaSuccess) return (int)e;
  // Kernels of neighbouring passes share SMs (the parse kernel of the next call runs beside rebuild / synthesis of
  // this one): all of them ask for the largest shared-memory carve-out, so that an SM configured for one of them
  // has room for the others' CTAs.
  e = cudaFuncSetAttribute(celt_rebuild_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (e != cudaSuccess) return (int)e;
  e = cudaFuncSetAttribute(celt_synth_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (e != cudaSuccess) return (int)e;
  if (dev < 64) g_attr_done[dev] = 1;
  return 0;
}

extern "C" int cbk_launch_rebuild(FrameSide* d_sides, FrameTape* d_tapes, int n, cudaStream_t stream) {
  if (n <= 0) return 0;
  int rc = ensure_attrs();
  if (rc) return rc;
  const int aw = entropy_cta_warps(n);
  celt_rebuild_kernel<<<(n + aw - 1) / aw, aw * 32, sizeof(EntWork) * (size_t)aw, stream>>>(d_sides, d_tapes, n);
  return (int)cudaGetLastError();
}

// The entropy stage of a pass: the parse kernel (cb_parse_kernels.cu, a thread per stream), then the rebuild kernel.
extern "C" int cbk_launch_entropy(OpusDecState* const* d_states, const u8* d_data, StreamPlan* d_plans,
                                  FrameSide* d_sides, FrameTape* d_tapes, i32* d_ret, int n, int pass, int max_payload,
                                  cudaStream_t stream) {
  if (n <= 0) return 0;
  int rc = ensure_attrs();
  if (rc) return rc;
  rc = cbk_launch_parse(d_states, d_data, d_plans, d_sides, d_tapes, d_ret, n, pass, max_payload, 0, stream);
  if (rc) return rc;
  return cbk_launch_rebuild(d_sides, d_tapes, n, stream);
}

extern "C" int cbk_launch_synth(OpusDecState* const* d_states, const StreamPlan* d_plans, const FrameSide* d_sides,
                                i16* d_pcm, i32 pcm_stride, int n, int pass, int max_ch, cudaStream_t stream) {
  if (n <= 0) return 0;
  int rc = ensure_attrs();
  if (rc) return rc;
  const size_t smemB = (sizeof(SynWork) + sizeof(EmitReq)) * B_WARPS;
  if (max_ch != 1) max_ch = 2;
  const int warpsB = n * max_ch;
  celt_synth_kernel<<<(warpsB + B_WARPS - 1) / B_WARPS, B_WARPS * 32, smemB, stream>>>(
      d_states, d_plans, d_sides, d_pcm, pcm_stride, n, pass, max_ch);
  return (int)cudaGetLastError();
}

extern "C" int cbk_launch_ms_plan(OpusDecState* const* d_states, StreamPlan* d_plans, const MSLayout* d_layouts,
                                  const u8* d_data, const i32* d_off, const i32* d_len, const i32* d_fec,
                                  int frame_size, i32* d_ms_ret, int n_ms, int max_passes, cudaStream_t stream) {
  if (n_ms <= 0) return 0;
  opus_ms_plan_kernel<<<(n_ms + 63) / 64, 64, 0, stream>>>(d_states, d_plans, d_layouts, d_data, d_off, d_len, d_fec,
                                                          frame_size, d_ms_ret, n_ms, max_passes);
  return (int)cudaGetLastError();
}

extern "C" int cbk_launch_ms_scatter(const MSLayout* d_layouts, const i32* d_ms_ret, const i16* d_epcm,
                                     i32 epcm_stride, i16* d_pcm, i32 pcm_stride, int n_ms, cudaStream_t stream) {
  if (n_ms <= 0) return 0;
  opus_ms_scatter_kernel<<<n_ms, 256, 0, stream>>>(d_layouts, d_ms_ret, d_epcm, epcm_stride, d_pcm, pcm_stride, n_ms);
  return (int)cudaGetLastError();
}

extern "C" size_t cbk_sizeof_state(void) { return sizeof(OpusDecState); }
extern "C" size_t cbk_sizeof_plan(void) { return sizeof(StreamPlan); }
extern "C" size_t cbk_sizeof_side(void) { return sizeof(FrameSide); }
extern "C" size_t cbk_sizeof_tape(void) { return sizeof(FrameTape); }
