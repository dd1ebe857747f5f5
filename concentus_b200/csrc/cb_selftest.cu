// cb_selftest.cu -- device-side known-answer hooks (tests/test_kat.py, -m gpu): the PVQ index <-> pulse-vector routines
// exactly as the kernels use them (b3_cwrsi, one lane per vector, as in the rebuild kernel's pulse phase; enc_icwrs,
// one warp per vector, as in the encode kernel), callable on arrays of (n, k, index).  Test support only: nothing in
// the decode / encode paths calls this file.
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/opus_b200.h"
#include "cb_celt_entropy.h"
#include "cb_rangeenc.h"
#include "cb_enc_state.h"
#include "cb_enc_quant.h"

__global__ void selftest_cwrs_kernel(const int* n, const int* k, const u32* idx, int count, i16* y, i32* yy, u32* back) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int base = warp * 32;
  const int mine = base + lane;
  if (mine < count) yy[mine] = b3_cwrsi(y + (size_t)mine * 176, n[mine], k[mine], idx[mine]);     // a lane per vector
  __syncwarp();
  for (int v = base; v < base + 32 && v < count; v++) {                                          // a warp per vector
    const u32 r = enc_icwrs(n[v], y + (size_t)v * 176);
    if (lane == 0) back[v] = r;
    __syncwarp();
  }
}

// d_* are device pointers: n, k, idx [count] in; y [count x 176] pulse vectors, yy [count] sum of squares, back [count]
// re-encoded indices out.  Queued on the default stream and waited for.
extern "C" int opus_b200_selftest_cwrs(const int* d_n, const int* d_k, const unsigned* d_idx, int count, short* d_y, int* d_yy,
                                       unsigned* d_back) {
  if (count <= 0) return OPUS_BAD_ARG;
  const int warps = (count + 31) / 32;
  selftest_cwrs_kernel<<<(warps + 3) / 4, 128>>>(d_n, d_k, d_idx, count, d_y, d_yy, d_back);
  if (cudaGetLastError() != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) return OPUS_INTERNAL_ERROR;
  return OPUS_OK;
}

__global__ void selftest_isqrt_kernel(const u32* in, int count, u32* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) out[i] = cb_isqrt32(in[i]);
}
// the device form of isqrt32 (float estimate + exact correction, cb_common.h) on an array of inputs
extern "C" int opus_b200_selftest_isqrt(const unsigned* d_in, int count, unsigned* d_out) {
  if (count <= 0) return OPUS_BAD_ARG;
  selftest_isqrt_kernel<<<(count + 255) / 256, 256>>>(d_in, count, d_out);
  if (cudaGetLastError() != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) return OPUS_INTERNAL_ERROR;
  return OPUS_OK;
}
