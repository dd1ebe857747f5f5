// cb_surround_kernels.h -- launch interface of the surround masking analysis (cb_surround_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include "cb_common.h"

// Where channel c's masks go: 21 int16 inside the CELT state of its elementary encoder, and that state's
// "a mask is set" flag.
struct SurroundDst {
  i16* mask[8];
  i32* has_mask[8];
};

extern "C" {
size_t cbk_sizeof_surround_state(void);
size_t cbk_sizeof_surround_scratch(void);
// d_pcm: len samples per channel, interleaved by `channels` (3..8), at rate Fs; d_state: SurroundState (zeroed at
// creation / reset); d_scratch: cbk_sizeof_surround_scratch() bytes.
int cbk_launch_surround(const i16* d_pcm, int channels, int len, int Fs, void* d_state, void* d_scratch,
                        const SurroundDst* dst, cudaStream_t stream);
}
