// cb_surround_kernels.cu -- the surround masking analysis of one multistream frame (cb_surround.h) as a kernel:
// one CTA, one warp per input channel for the transform and band levels, then warp 0 combines the channels band by
// band and writes each channel's 21 masks straight into the CELT state of the elementary encoder that codes that
// channel (CeltEncState::energy_mask), so the encode kernels launched behind it on the same stream read them
// without a host round trip.
#include <cuda_runtime.h>
#include "cb_surround.h"
#include "cb_surround_kernels.h"

__global__ void __launch_bounds__(CB_SURROUND_MAX_CH * 32)
surround_analysis_kernel(const i16* pcm, int channels, int len, int Fs, SurroundState* st, SurroundScratch* scratch,
                         SurroundDst dst) {
  __shared__ i16 levels[CB_SURROUND_MAX_CH * CB_NBANDS];
  const int c = threadIdx.x >> 5;
  if (c < channels) surround_channel(pcm, channels, c, len, Fs, st, scratch[c], levels + CB_NBANDS * c);
  __syncthreads();
  if (c == 0) {
    surround_combine(levels, channels);
    for (int ch = 0; ch < channels; ch++) {
      CB_LANE_LOOP(i, CB_NBANDS) dst.mask[ch][i] = levels[CB_NBANDS * ch + i];
      CB_IF_LANE0 *dst.has_mask[ch] = 1;
    }
  }
}

extern "C" {

size_t cbk_sizeof_surround_state(void) { return sizeof(SurroundState); }
size_t cbk_sizeof_surround_scratch(void) { return sizeof(SurroundScratch) * CB_SURROUND_MAX_CH; }

int cbk_launch_surround(const i16* d_pcm, int channels, int len, int Fs, void* d_state, void* d_scratch,
                        const SurroundDst* dst, cudaStream_t stream) {
  if (channels < 3 || channels > CB_SURROUND_MAX_CH || len * (48000 / Fs) > CB_SURROUND_MAX_FRAME) return (int)cudaErrorInvalidValue;
  surround_analysis_kernel<<<1, channels * 32, 0, stream>>>(d_pcm, channels, len, Fs, (SurroundState*)d_state,
                                                            (SurroundScratch*)d_scratch, *dst);
  return (int)cudaGetLastError();
}

}  // extern "C"
