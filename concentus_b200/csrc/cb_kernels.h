// cb_kernels.h -- launch interface between the host API (opus_api.cu) and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include "cb_state.h"

extern "C" {
int cbk_launch_entropy(OpusDecState* const* d_states, const u8* d_data, StreamPlan* d_plans, FrameSide* d_sides,
                       FrameTape* d_tapes, i32* d_ret, int n, int pass, int max_payload, cudaStream_t stream);
int cbk_launch_rebuild(FrameSide* d_sides, FrameTape* d_tapes, int n, cudaStream_t stream);
// shared_sms: the kernel will run beside the rebuild / synthesis kernels of the call before (run-ahead scheduling)
int cbk_launch_parse(OpusDecState* const* d_states, const u8* d_data, StreamPlan* d_plans, FrameSide* d_sides,
                     FrameTape* d_tapes, i32* d_ret, int n, int pass, int max_payload, int shared_sms,
                     cudaStream_t stream);
int cbk_launch_synth(OpusDecState* const* d_states, const StreamPlan* d_plans, const FrameSide* d_sides,
                     i16* d_pcm, i32 pcm_stride, int n, int pass, int max_ch, cudaStream_t stream);
// SILK frames of pass `pass` (one thread per stream); streams whose frame is CELT return at once
int cbk_launch_plan_silk(OpusDecState* const* d_states, const u8* d_data, const i32* d_off, const i32* d_len,
                         const i32* d_fec, int frame_size, StreamPlan* d_plans, FrameSide* d_sides, i16* d_pcm,
                         i32 pcm_stride, i32* d_ret, int n, int max_passes, cudaStream_t stream);
int cbk_launch_silk(OpusDecState* const* d_states, StreamPlan* d_plans, FrameSide* d_sides, const u8* d_data,
                    i16* d_pcm, i32 pcm_stride, i32* d_ret, int n, int pass, cudaStream_t stream);
int cbk_launch_general(OpusDecState* const* d_states, StreamPlan* d_plans, FrameSide* d_sides, FrameTape* d_tapes,
                       const u8* d_data, i16* d_pcm, i32 pcm_stride, i32* d_ret, int n, int pass, cudaStream_t stream);
int cbk_launch_ms_plan(OpusDecState* const* d_states, StreamPlan* d_plans, const MSLayout* d_layouts,
                       const u8* d_data, const i32* d_off, const i32* d_len, const i32* d_fec, int frame_size,
                       i32* d_ms_ret, int n_ms, int max_passes, cudaStream_t stream);
int cbk_launch_ms_scatter(const MSLayout* d_layouts, const i32* d_ms_ret, const i16* d_epcm, i32 epcm_stride,
                          i16* d_pcm, i32 pcm_stride, int n_ms, cudaStream_t stream);
size_t cbk_sizeof_state(void);
size_t cbk_sizeof_plan(void);
size_t cbk_sizeof_side(void);
size_t cbk_sizeof_tape(void);
}
