// cb_enc_kernels.h -- launch interface between the encoder host API (opus_enc_api.cu) and the kernel.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include "cb_enc_state.h"

extern "C" {
size_t cbk_sizeof_enc_state(void);
size_t cbk_sizeof_enc_scratch(void);
// One frame of `frame_size` samples per channel for each of n streams.  Stream s reads
// d_pcm + s*pcm_stride (interleaved int16) and writes its packet at d_data + s*data_stride
// (max_data_bytes of room); d_len[s] = packet length or a negative OPUS_* code.
// max_subframes: the most 20 ms frames any stream codes for this call (40 / 60 ms frames); general != 0 when a
// stream has the LFE flag or a surround mask -- either picks the kernel instance that has those paths.
int cbk_launch_encode(OpusEncState* const* d_states, void* d_scratch, const i16* d_pcm, i32 pcm_stride,
                      int frame_size, u8* d_data, i32 data_stride, int max_data_bytes, i32* d_len, int n,
                      int max_subframes, int general, cudaStream_t stream);
}
