// cb_noalloc.cu -- opus_decode_batch_device_static(): the device-resident batch decode for callers
// that are not allowed to allocate GPU memory.
//
// Who needs this.  A media server decodes on a deadline: one thread per GPU feeds 20 ms frames and
// must hand PCM back before the next tick.  cudaMalloc/cudaFree are off limits on that thread --
// they take driver-wide locks and cudaFree synchronises the whole device, so one allocation stalls
// every other stream's audio.  The same goes for hosts that hand the library a fixed GPU memory
// budget up front (all memory is carved out at start-up, nothing may be requested later).
// The ordinary batch entry points need an OpusDecoderGroup, whose creation
// allocates the per-stream scratch (StreamPlan, FrameSide and FrameTape records).  This entry point needs no
// group: its scratch is the fixed array below, which is part of the library image.
//
// Cost and limit.  The array is OPUS_B200_STATIC_MAX_STREAMS (256) records long, about 4.1 MB, and
// every GPU that loads the library keeps it until the process exits.  Batches larger than that are
// rejected; use a group for them.
//
// Sharing.  There is ONE array per GPU.  Every call on that GPU uses it, and it still holds the
// previous call's records when the next call starts, so each call first rewrites the headers it
// relies on (fill_static_scratch below).  Two calls whose kernels can run at the same time on one
// GPU -- different streams, or different host threads -- would overwrite each other's records:
// callers must order them (same stream, or an event between them).
//
// First call.  The device address of the array is looked up with cudaGetSymbolAddress the first
// time a GPU is used through this entry point.  Make that first call during start-up, before any
// stream capture begins.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/opus_b200.h"
#include "cb_kernels.h"
#include "cb_state.h"

#define STATIC_MAX OPUS_B200_STATIC_MAX_STREAMS

struct StaticScratch {
  OpusDecState* states[STATIC_MAX];   // device addresses of the decoders' state blobs
  StreamPlan plans[STATIC_MAX];
  FrameSide sides[STATIC_MAX];
  FrameTape tapes[STATIC_MAX];
};

// The scratch itself.  In host code this name is only a stand-in for the symbol, not a device
// address: the address comes from cudaGetSymbolAddress below and is what the kernels receive.
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

struct StatePtrPack { OpusDecState* p[STATIC_MAX]; };   // 2 KB, travels as a kernel argument

// Writes the state-pointer table and resets the record headers that the decode kernels read
// before they write them (the array still holds the previous call's contents).
__global__ void fill_static_scratch(StatePtrPack pack, StaticScratch* s, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  s->states[i] = pack.p[i];
  s->plans[i].ret = 0;
  s->plans[i].nframes = 0;
  s->plans[i].kind = CB_PLAN_NONE;
  s->plans[i].total_samples = 0;
  s->sides[i].kind = 0;
  s->sides[i].accum = 0;
  s->tapes[i].kind = CB_TAPE_NONE;
  s->tapes[i].nseg = 0;
}

static StaticScratch* g_scratch_addr[16];   // per device, filled on first use

extern "C" OpusDecState* opusb_internal_device_state(const OpusDecoder* st, int* device, int* channels);
extern "C" cudaStream_t opusb_internal_stream(int* err);
extern "C" void opusb_internal_serial_call_done(int device);

extern "C" int opus_decode_batch_device_static(OpusDecoder* const* st, int n, const unsigned char* d_data,
                                               const opus_int32* d_offsets, const opus_int32* d_len,
                                               opus_int16* d_pcm, opus_int32 pcm_stride, int frame_size,
                                               int* d_ret, int max_frames, int max_packet_bytes,
                                               void* cuda_stream) {
  if (!st || !d_data || !d_offsets || !d_len || !d_pcm || !d_ret) return OPUS_BAD_ARG;
  if (n <= 0 || frame_size <= 0 || max_frames < 1 || max_frames > CB_MAX_FRAMES || max_packet_bytes < 0)
    return OPUS_BAD_ARG;
  if (n > STATIC_MAX) {
    fprintf(stderr, "[opus_b200] opus_decode_batch_device_static: %d streams requested, the static scratch "
                    "holds at most %d (OPUS_B200_STATIC_MAX_STREAMS); use an OpusDecoderGroup for larger batches\n",
            n, STATIC_MAX);
    return OPUS_BAD_ARG;
  }
  StatePtrPack pack;
  int device = -1, maxch = 1;
  for (int i = 0; i < n; i++) {
    int dev = -1, ch = 1;
    pack.p[i] = opusb_internal_device_state(st[i], &dev, &ch);
    if (!pack.p[i]) return OPUS_BAD_ARG;
    if (ch > maxch) maxch = ch;
    if (i == 0) device = dev;
    else if (dev != device) {
      fprintf(stderr, "[opus_b200] opus_decode_batch_device_static: decoders live on different GPUs\n");
      return OPUS_BAD_ARG;
    }
  }
  if (device < 0 || device >= 16) return OPUS_BAD_ARG;
  if ((long long)pcm_stride < (long long)frame_size * maxch) return OPUS_BAD_ARG;   // rows would overlap
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) {
    fprintf(stderr, "[opus_b200] cudaSetDevice(%d) failed: %s\n", device, cudaGetErrorString(e));
    return OPUS_INTERNAL_ERROR;
  }
  if (!g_scratch_addr[device]) {
    // one-time set-up for this GPU (it is the current device here)
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      fprintf(stderr, "[opus_b200] cudaGetSymbolAddress(static scratch) failed: %s\n", cudaGetErrorString(e));
      return OPUS_INTERNAL_ERROR;
    }
    g_scratch_addr[device] = (StaticScratch*)p;
  }
  StaticScratch* s = g_scratch_addr[device];
  // NULL = the library's own stream, as in opus_decode_group_device: the call is then ordered with opus_decode /
  // ctl work on the same decoders and opus_b200_synchronize() waits for it
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  const bool own_stream = !stream;
  if (!stream) {
    int serr = 0;
    stream = opusb_internal_stream(&serr);
    if (!stream) return serr ? serr : OPUS_INTERNAL_ERROR;
  }
  fill_static_scratch<<<(n + 127) / 128, 128, 0, stream>>>(pack, s, n);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    fprintf(stderr, "[opus_b200] static scratch fill launch failed: %s\n", cudaGetErrorString(e));
    return OPUS_INTERNAL_ERROR;
  }
  int rc = cbk_launch_plan_silk(s->states, d_data, d_offsets, d_len, nullptr, frame_size, s->plans, s->sides, d_pcm,
                                pcm_stride, d_ret, n, max_frames, stream);
  for (int p = 0; p < max_frames && !rc; p++) {
    if (p > 0) rc = cbk_launch_silk(s->states, s->plans, s->sides, d_data, d_pcm, pcm_stride, d_ret, n, p, stream);
    if (!rc) rc = cbk_launch_entropy(s->states, d_data, s->plans, s->sides, s->tapes, d_ret, n, p, max_packet_bytes, stream);
    if (!rc) rc = cbk_launch_synth(s->states, s->plans, s->sides, d_pcm, pcm_stride, n, p, 2, stream);
    if (!rc) rc = cbk_launch_general(s->states, s->plans, s->sides, s->tapes, d_data, d_pcm, pcm_stride, d_ret, n, p, stream);
  }
  if (own_stream) opusb_internal_serial_call_done(device);
  if (rc) {
    fprintf(stderr, "[opus_b200] opus_decode_batch_device_static: launch failed: %s\n",
            cudaGetErrorString((cudaError_t)rc));
    return OPUS_INTERNAL_ERROR;
  }
  return OPUS_OK;
}
