// cb_silk_kernels.cu -- sm_100a kernel of the batched SILK decoder and its launcher.
//
//   silk_decode_kernel : 1 THREAD / stream -- the SILK branch of one Opus frame: range decoding of the
//   side information and pulses, NLSF -> LPC, LTP + LPC synthesis, concealment / comfort noise,
//   stereo unmixing, resampling to the API rate, PCM store.
//
// SILK has no width inside a stream (per-sample recurrences of order 16 and pitch-lag feedback), so
// the batch is the parallel axis: 32 streams per warp, small blocks so that the streams spread over
// all SMs.  Work arrays are per-thread local memory (interleaved across lanes by the hardware).
// Streams whose current frame is not SILK leave at once; a CELT-only batch pays one empty launch.
#include <cuda_runtime.h>
#include <stdint.h>
#include "cb_silk_pass.h"
#include "cb_kernels.h"

#ifndef S_THREADS
#define S_THREADS 32
#endif

__global__ void __launch_bounds__(S_THREADS)
silk_decode_kernel(OpusDecState* const* states, StreamPlan* plans, FrameSide* sides, const u8* data, i16* pcm,
                   i32 pcm_stride, i32* ret, int n, int pass) {
  const int s = blockIdx.x * S_THREADS + threadIdx.x;
  if (s >= n) return;
  OpusDecState* st = states[s];
  StreamPlan* pl = &plans[s];
  SilkCallWork cw;
  decode_silk_pass(st, pl, &sides[s], data, pcm + (size_t)s * pcm_stride, &ret[s], pass, cw);
  sides[s].route = pl->route;
}

// The first pass of a batch call: packet parsing / frame plan (opus_plan_stream) and the SILK pass of
// frame 0 in one launch (both are one thread per stream).
__global__ void __launch_bounds__(S_THREADS)
plan_silk_kernel(OpusDecState* const* states, const u8* data, const i32* offs, const i32* lens, const i32* fec,
                 int frame_size, StreamPlan* plans, FrameSide* sides, i16* pcm, i32 pcm_stride, i32* ret, int n,
                 int max_passes) {
  const int s = blockIdx.x * S_THREADS + threadIdx.x;
  if (s >= n) return;
  OpusDecState* st = states[s];
  StreamPlan* pl = &plans[s];
  opus_plan_stream(st, pl, data, offs[s], lens[s], frame_size, fec ? fec[s] : 0, 0, nullptr, max_passes);
  ret[s] = pl->ret;
  SilkCallWork cw;
  decode_silk_pass(st, pl, &sides[s], data, pcm + (size_t)s * pcm_stride, &ret[s], 0, cw);
  sides[s].route = pl->route;
}

extern "C" int cbk_launch_plan_silk(OpusDecState* const* d_states, const u8* d_data, const i32* d_off, const i32* d_len,
                                    const i32* d_fec, int frame_size, StreamPlan* d_plans, FrameSide* d_sides, i16* d_pcm,
                                    i32 pcm_stride, i32* d_ret, int n, int max_passes, cudaStream_t stream) {
  if (n <= 0) return 0;
  plan_silk_kernel<<<(n + S_THREADS - 1) / S_THREADS, S_THREADS, 0, stream>>>(d_states, d_data, d_off, d_len, d_fec,
                                                                               frame_size, d_plans, d_sides, d_pcm,
                                                                               pcm_stride, d_ret, n, max_passes);
  return (int)cudaGetLastError();
}

extern "C" int cbk_launch_silk(OpusDecState* const* d_states, StreamPlan* d_plans, FrameSide* d_sides, const u8* d_data,
                               i16* d_pcm, i32 pcm_stride, i32* d_ret, int n, int pass, cudaStream_t stream) {
  if (n <= 0) return 0;
  silk_decode_kernel<<<(n + S_THREADS - 1) / S_THREADS, S_THREADS, 0, stream>>>(d_states, d_plans, d_sides, d_data,
                                                                                 d_pcm, pcm_stride, d_ret, n, pass);
  return (int)cudaGetLastError();
}
