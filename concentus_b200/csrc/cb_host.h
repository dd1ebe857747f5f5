// cb_host.h -- per-GPU host context shared by the API translation units (opus_api.cu, opus_enc_api.cu,
// opus_ms_enc_api.cu, cb_noalloc.cu) and the kernel launchers.
//
// Threading model of the library (the reference has no globals, src/opus_decoder.c / opus_encoder.c; here the only
// shared things are per GPU): one mutex PER DEVICE guards that device's allocator, its lazily created streams /
// events and the order in which a call's launches are queued.  It is never held while the host waits for the GPU,
// so threads of one process can drive different GPUs -- or different groups on one GPU -- at the same time.
// One decoder / encoder / group object must not be used by two threads at once (as with libopus).
#pragma once
#include <cuda_runtime.h>
#include <mutex>

#define OPUSB_MAX_DEVICES 64

struct OpusbDevice {
  std::mutex mu;
  int device = -1;             // -1 until first use
  int sm_count = 0;            // cudaDevAttrMultiProcessorCount
  cudaStream_t stream = nullptr;   // the library's own (non-blocking) stream on this GPU
};

// Makes `device` current for the calling thread (device < 0: keeps the thread's current device) and returns its
// context, creating stream and attributes on first use.  nullptr + *err on failure.
OpusbDevice* opusb_device(int device, int* err);
// SM count of the current device (cached); what the launchers size their grids by.
extern "C" int cbk_sm_count(void);
extern "C" void opusb_internal_count_launches(int n);
