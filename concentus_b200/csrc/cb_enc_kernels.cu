// cb_enc_kernels.cu -- sm_100a kernel of the batched Opus (CELT-only) encoder and its launcher.
//
//   opus_encode_kernel : 1 warp / stream -- the whole frame: DC reject, pre-emphasis, pitch
//   pre-filter, MDCT, band analysis, bit allocation, PVQ search, range coder, TOC / padding.
//
// Integer, latency-bound work (no tensor cores).  Per stream the kernel reads frame_size*channels
// int16 samples and writes one packet; the large per-frame vectors (EncScratch, ~49 KB/stream)
// sit in global memory and stay L2-resident, the band-level working set (EncWork) in shared memory.
#include <cuda_runtime.h>
#include <stdint.h>
#include "cb_opus_enc.h"
#include "cb_enc_kernels.h"

#define E_WARPS 28

template <bool MULTI>
__global__ void __launch_bounds__(E_WARPS * 32, 1)
opus_encode_kernel(OpusEncState* const* states, EncScratch* scratch, const i16* pcm, i32 pcm_stride, int frame_size,
                   u8* data, i32 data_stride, int max_data_bytes, i32* out_len, int n, int phases) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int warp = threadIdx.x >> 5;
  const int s = blockIdx.x * (blockDim.x >> 5) + warp;
  EncWork& w = *reinterpret_cast<EncWork*>(smem + (size_t)warp * sizeof(EncWork));
  // the CTA's table of serial-section requests sits behind the warps' workspaces
  EncSerialReq* table = reinterpret_cast<EncSerialReq*>(smem + (size_t)(blockDim.x >> 5) * sizeof(EncWork));
  if ((threadIdx.x & 31) == 0) {
    w.nsync = 0;
    w.sreq = table + warp;
    table[warp].kind = CB_ES_NONE;
    table[warp].channels = 0;
  }
  __syncthreads();
  if (s < n) {
    const int ret = opus_encode_stream<MULTI>(w, scratch[s], states[s], pcm + (size_t)s * pcm_stride, frame_size,
                                       data + (size_t)s * data_stride, max_data_bytes);
    if ((threadIdx.x & 31) == 0) out_len[s] = ret;
  }
  __syncwarp();
  // warps that left early (or have no stream) still owe the CTA their remaining rendezvous
  // (`phases` = CB_ENC_PHASES times the most 20 ms frames any stream of this launch codes for one call)
  // -- and warps 0 and 1 still run the serial sections of the streams that go on (cb_enc_state.h: EncSerialReq)
  for (int k = w.nsync; k < phases; k++) {
    __syncthreads();
    enc_serial_run(table);
  }
}

extern "C" int cbk_sm_count(void);   // SMs of the current device (opus_api.cu)

extern "C" {

size_t cbk_sizeof_enc_state(void) { return sizeof(OpusEncState); }
size_t cbk_sizeof_enc_scratch(void) { return sizeof(EncScratch); }

int cbk_launch_encode(OpusEncState* const* d_states, void* d_scratch, const i16* d_pcm, i32 pcm_stride,
                      int frame_size, u8* d_data, i32 data_stride, int max_data_bytes, i32* d_len, int n,
                      int max_subframes, int general, cudaStream_t stream) {
  static bool attr_set[16] = {false};
  const size_t smem = (sizeof(EncWork) + sizeof(EncSerialReq)) * E_WARPS;
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 16 && !attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(opus_encode_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(opus_encode_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    attr_set[dev] = true;
  }
  // CTA shape: an equal share of the batch for every SM (ceil(n / #SMs) warps per CTA, at most 28 = what an SM
  // holds; the same over two or more waves for larger batches), so that a small batch still uses the whole GPU
  const int sms = cbk_sm_count();
  const int waves = (n + sms * E_WARPS - 1) / (sms * E_WARPS);
  int ew = (n + sms * waves - 1) / (sms * waves);
  if (ew > E_WARPS) ew = E_WARPS;
  if (ew < 1) ew = 1;
  const int blocks = (n + ew - 1) / ew;
  const size_t smem_l = (sizeof(EncWork) + sizeof(EncSerialReq)) * ew;
  if (max_subframes > 1 || general)
    opus_encode_kernel<true><<<blocks, ew * 32, smem_l, stream>>>(d_states, (EncScratch*)d_scratch, d_pcm, pcm_stride,
                                                                    frame_size, d_data, data_stride, max_data_bytes, d_len, n,
                                                                    CB_ENC_PHASES * max_subframes);
  else
    opus_encode_kernel<false><<<blocks, ew * 32, smem_l, stream>>>(d_states, (EncScratch*)d_scratch, d_pcm, pcm_stride,
                                                                     frame_size, d_data, data_stride, max_data_bytes, d_len, n,
                                                                     CB_ENC_PHASES);
  return (int)cudaGetLastError();
}

}  // extern "C"
