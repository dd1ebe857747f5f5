// cb_general_kernels.cu -- sm_100a kernel for the Opus frames the fast kernels do not take.
//
//   opus_general_kernel : 1 warp / stream -- opus_decode_frame in full (cb_opus_general.h): hybrid
//   packets, CELT<->SILK/hybrid transitions and packets with a redundant CELT frame.  The SILK layer
//   of such a frame runs on lane 0; the CELT layer, the fades and the gain on the whole warp.
// The same launch runs the pitch-based concealment of CELT frames (cb_celt_plc.h), the other rare
// warp-per-stream path.  Streams with nothing to do here leave at once.
#include <cuda_runtime.h>
#include <stdint.h>
#include "cb_opus_general.h"
#include "cb_kernels.h"

extern "C" int cbk_sm_count(void);

#define G_WARPS 4

__global__ void __launch_bounds__(G_WARPS * 32)
opus_general_kernel(OpusDecState* const* states, StreamPlan* plans, FrameSide* sides, FrameTape* tapes, const u8* data,
                    i16* pcm, i32 pcm_stride, i32* ret, int n, int pass) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int warp = threadIdx.x >> 5;
  GenWork& g = *reinterpret_cast<GenWork*>(smem + (size_t)warp * sizeof(GenWork));
  SilkCallWork cw;
  for (int s = blockIdx.x * G_WARPS + warp; s < n; s += gridDim.x * G_WARPS) {
    StreamPlan* pl = &plans[s];
    const int route = pl->route;
    if (route == CB_ROUTE_PLC) {
      // pitch-based concealment of a CELT frame (celt_decode_lost, loss_count < 5), set up by the SILK kernel
      decode_plc_pass(g.pw, states[s], pl, &sides[s], pcm + (size_t)s * pcm_stride, pass);
    } else if (route == CB_ROUTE_GENERAL || route == CB_ROUTE_FINISH) {
      decode_general_pass(g, states[s], pl, &sides[s], &tapes[s], data, pcm + (size_t)s * pcm_stride, &ret[s], pass, cw);
    }
    __syncwarp();
  }
}

extern "C" int cbk_launch_general(OpusDecState* const* d_states, StreamPlan* d_plans, FrameSide* d_sides, FrameTape* d_tapes,
                                  const u8* d_data, i16* d_pcm, i32 pcm_stride, i32* d_ret, int n, int pass,
                                  cudaStream_t stream) {
  if (n <= 0) return 0;
  static bool attr_set[64];
  const int smem = G_WARPS * (int)sizeof(GenWork);
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev >= 64 || !attr_set[dev]) {
    e = cudaFuncSetAttribute(opus_general_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return (int)e;
    if (dev < 64) attr_set[dev] = true;
  }
  int grid = (n + G_WARPS - 1) / G_WARPS;
  const int cap = cbk_sm_count() * 4;
  if (grid > cap) grid = cap;               // a few CTAs per SM; streams beyond are walked by a grid-stride loop
  opus_general_kernel<<<grid, G_WARPS * 32, smem, stream>>>(d_states, d_plans, d_sides, d_tapes, d_data, d_pcm,
                                                            pcm_stride, d_ret, n, pass);
  return (int)cudaGetLastError();
}
