// opus_enc_api.cu -- host side of the encoder half of libopus_b200.so: the libopus encoder ABI
// (opus_encoder_create / opus_encode / opus_encode_float / opus_encoder_ctl / opus_encoder_destroy,
// opus-fix/src/opus_encoder.c) plus the batch entry points.  Holds no codec arithmetic: packets
// are produced by opus_encode_kernel (cb_enc_kernels.cu).  There is NO CPU encode path.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <mutex>
#include <vector>
#include "../../include/opus_b200.h"
#include "cb_host.h"
#include "cb_enc_ctl.h"
#include "cb_enc_kernels.h"
#include "cb_enc_state.h"

#define CUDA_OK(call)                                                                            \
  do {                                                                                           \
    cudaError_t e_ = (call);                                                                     \
    if (e_ != cudaSuccess) {                                                                     \
      fprintf(stderr, "[opus_b200] CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e_), __FILE__, \
              __LINE__, cudaGetErrorString(e_));                                                 \
      return OPUS_INTERNAL_ERROR;                                                                \
    }                                                                                            \
  } while (0)


namespace {
const int kEncMagic = 0x4f453230;   // "OE20"

struct EncProf { cudaEvent_t a, b; };
std::atomic<bool> g_enc_prof_on{false};
std::mutex g_enc_prof_mu;              // guards g_enc_prof only
std::vector<EncProf> g_enc_prof;
}  // namespace

struct OpusEncoderGroup;
struct OpusEncoder {
  int magic;
  int Fs, channels, application;
  int device;
  OpusEncState* d_state;
  OpusEncoderGroup* self_group;   // group of one, created on first opus_encode()
  bool ext;                       // has had OPUS_SET_LFE / OPUS_SET_ENERGY_MASK: launches use the general kernel instance
};

// Two-deep pipeline behind opus_encode_group_submit/_wait (copies on their own streams, see opus_api.cu)
struct EncPipeSlot {
  i16* d_pcm; size_t d_pcm_cap;
  u8* d_data; size_t d_data_cap;
  i32* d_len;
  cudaEvent_t in_done, k_done, out_done;
  bool busy;
};
struct EncPipe {
  cudaStream_t s_in, s_out;
  EncPipeSlot slot[2];
  unsigned long long seq;
};

struct OpusEncoderGroup {
  int magic;
  int n, device;
  int Fs, channels;               // common values, or -1 when the members differ
  OpusEncoder** members;
  OpusEncState** d_states;        // device array of n state pointers
  void* d_scratch;                // n EncScratch
  i32* d_len;                     // n
  // staging for the host-buffer entry point (grow-only)
  i16* d_pcm; size_t d_pcm_cap;
  u8* d_data; size_t d_data_cap;
  i16* h_pcm; size_t h_pcm_cap;   // pinned
  u8* h_data; size_t h_data_cap;  // pinned
  i32* h_len;                     // pinned, n
  EncPipe* pipe;                  // created by the first opus_encode_group_submit
};

namespace {

int grow_dev(void** p, size_t* cap, size_t bytes) {
  if (bytes <= *cap) return OPUS_OK;
  if (*p) cudaFree(*p);
  *p = nullptr; *cap = 0;
  if (cudaMalloc(p, bytes) != cudaSuccess) {
    fprintf(stderr, "[opus_b200] cudaMalloc of %zu bytes failed\n", bytes);
    return OPUS_ALLOC_FAIL;
  }
  *cap = bytes;
  return OPUS_OK;
}
int grow_pinned(void** p, size_t* cap, size_t bytes) {
  if (bytes <= *cap) return OPUS_OK;
  if (*p) cudaFreeHost(*p);
  *p = nullptr; *cap = 0;
  if (cudaMallocHost(p, bytes) != cudaSuccess) {
    fprintf(stderr, "[opus_b200] cudaMallocHost of %zu bytes failed\n", bytes);
    return OPUS_ALLOC_FAIL;
  }
  *cap = bytes;
  return OPUS_OK;
}

bool valid(const OpusEncoder* st) { return st && st->magic == kEncMagic; }
bool valid(const OpusEncoderGroup* g) { return g && g->magic == kEncMagic + 1; }

// launch one frame for every member; device buffers as described in cb_enc_kernels.h
int group_launch(OpusEncoderGroup* g, const i16* d_pcm, i32 pcm_stride, int frame_size, u8* d_data, i32 data_stride,
                 int max_data_bytes, i32* d_len, cudaStream_t stream) {
  EncProf pr{nullptr, nullptr};
  const bool prof = g_enc_prof_on.load();
  if (prof) { cudaEventCreate(&pr.a); cudaEventCreate(&pr.b); cudaEventRecord(pr.a, stream); }
  // a frame longer than 20 ms is coded as two or three 20 ms frames: the kernel's stage rendezvous has to
  // know the most any member will do (members may run at different rates)
  int max_sub = 1, general = 0;
  for (int i = 0; i < g->n; i++) {
    const int Fs = g->members[i]->Fs;
    const int nb = frame_size * 50 > Fs ? (frame_size * 25 > Fs ? 3 : 2) : 1;
    if (nb > max_sub) max_sub = nb;
    if (g->members[i]->ext) general = 1;
  }
  int rc = cbk_launch_encode(g->d_states, g->d_scratch, d_pcm, pcm_stride, frame_size, d_data, data_stride,
                             max_data_bytes, d_len, g->n, max_sub, general, stream);
  if (prof) { cudaEventRecord(pr.b, stream); std::lock_guard<std::mutex> lk(g_enc_prof_mu); g_enc_prof.push_back(pr); }
  opusb_internal_count_launches(1);
  if (rc != 0) {
    fprintf(stderr, "[opus_b200] encode kernel launch failed: %s\n", cudaGetErrorString((cudaError_t)rc));
    return OPUS_INTERNAL_ERROR;
  }
  return OPUS_OK;
}

}  // namespace

extern "C" {

int opus_encoder_get_size(int channels) {
  if (channels < 1 || channels > 2) return 0;
  return (int)sizeof(OpusEncoder);           // the codec state itself lives in device memory
}

OpusEncoder* opus_encoder_create(opus_int32 Fs, int channels, int application, int* error) {
  if ((Fs != 48000 && Fs != 24000 && Fs != 16000 && Fs != 12000 && Fs != 8000) || (channels != 1 && channels != 2) ||
      (application != OPUS_APPLICATION_VOIP && application != OPUS_APPLICATION_AUDIO &&
       application != OPUS_APPLICATION_RESTRICTED_LOWDELAY)) {
    if (error) *error = OPUS_BAD_ARG;
    return nullptr;
  }
  int err = 0;
  OpusbDevice* dv = opusb_device(-1, &err);          // the thread's current GPU
  if (!dv) { if (error) *error = err; return nullptr; }
  OpusEncoder* st = (OpusEncoder*)calloc(1, sizeof(OpusEncoder));
  OpusEncState* h = (OpusEncState*)malloc(sizeof(OpusEncState));
  if (!st || !h) { free(st); free(h); if (error) *error = OPUS_ALLOC_FAIL; return nullptr; }
  opus_enc_init(h, Fs, channels, application);
  st->device = dv->device;
  if (cudaMalloc((void**)&st->d_state, sizeof(OpusEncState)) != cudaSuccess ||
      cudaMemcpy(st->d_state, h, sizeof(OpusEncState), cudaMemcpyHostToDevice) != cudaSuccess) {
    fprintf(stderr, "[opus_b200] encoder state allocation failed: %s\n", cudaGetErrorString(cudaGetLastError()));
    if (st->d_state) cudaFree(st->d_state);
    free(st); free(h);
    if (error) *error = OPUS_ALLOC_FAIL;
    return nullptr;
  }
  free(h);
  st->magic = kEncMagic;
  st->Fs = Fs; st->channels = channels; st->application = application;
  if (error) *error = OPUS_OK;
  return st;
}

void opus_encoder_group_destroy(OpusEncoderGroup* g);

void opus_encoder_destroy(OpusEncoder* st) {
  if (!valid(st)) return;
  if (st->self_group) opus_encoder_group_destroy(st->self_group);
  {
    int err = 0;
    OpusbDevice* dv = opusb_device(st->device, &err);
    if (dv) cudaStreamSynchronize(dv->stream);
  }
  cudaFree(st->d_state);
  st->magic = 0;
  free(st);
}

OpusEncoderGroup* opus_encoder_group_create(OpusEncoder* const* st, int n, int* error) {
  if (!st || n <= 0) { if (error) *error = OPUS_BAD_ARG; return nullptr; }
  for (int i = 0; i < n; i++)
    if (!valid(st[i]) || st[i]->device != st[0]->device) { if (error) *error = OPUS_BAD_ARG; return nullptr; }
  {
    int err = 0;
    if (!opusb_device(st[0]->device, &err)) { if (error) *error = err; return nullptr; }   // makes the group's GPU current
  }
  OpusEncoderGroup* g = (OpusEncoderGroup*)calloc(1, sizeof(OpusEncoderGroup));
  if (!g) { if (error) *error = OPUS_ALLOC_FAIL; return nullptr; }
  g->n = n; g->device = st[0]->device;
  g->Fs = st[0]->Fs; g->channels = st[0]->channels;
  g->members = (OpusEncoder**)malloc(sizeof(OpusEncoder*) * n);
  std::vector<OpusEncState*> ptrs(n);
  for (int i = 0; i < n; i++) {
    g->members[i] = st[i];
    ptrs[i] = st[i]->d_state;
    if (st[i]->Fs != g->Fs) g->Fs = -1;
    if (st[i]->channels != g->channels) g->channels = -1;
  }
  bool ok = cudaMalloc((void**)&g->d_states, sizeof(OpusEncState*) * n) == cudaSuccess &&
            cudaMalloc(&g->d_scratch, cbk_sizeof_enc_scratch() * (size_t)n) == cudaSuccess &&
            cudaMalloc((void**)&g->d_len, sizeof(i32) * n) == cudaSuccess &&
            cudaMallocHost((void**)&g->h_len, sizeof(i32) * n) == cudaSuccess &&
            cudaMemcpy(g->d_states, ptrs.data(), sizeof(OpusEncState*) * n, cudaMemcpyHostToDevice) == cudaSuccess;
  if (!ok) {
    fprintf(stderr, "[opus_b200] encoder group allocation failed: %s\n", cudaGetErrorString(cudaGetLastError()));
    g->magic = kEncMagic + 1;
    opus_encoder_group_destroy(g);
    if (error) *error = OPUS_ALLOC_FAIL;
    return nullptr;
  }
  g->magic = kEncMagic + 1;
  if (error) *error = OPUS_OK;
  return g;
}

static void enc_pipe_free(EncPipe* p) {
  if (!p) return;
  for (EncPipeSlot& sl : p->slot) {
    if (sl.busy && sl.out_done) cudaEventSynchronize(sl.out_done);
    cudaFree(sl.d_pcm); cudaFree(sl.d_data); cudaFree(sl.d_len);
    if (sl.in_done) cudaEventDestroy(sl.in_done);
    if (sl.k_done) cudaEventDestroy(sl.k_done);
    if (sl.out_done) cudaEventDestroy(sl.out_done);
  }
  if (p->s_in) cudaStreamDestroy(p->s_in);
  if (p->s_out) cudaStreamDestroy(p->s_out);
  free(p);
}

void opus_encoder_group_destroy(OpusEncoderGroup* g) {
  if (!valid(g)) return;
  {
    int err = 0;
    OpusbDevice* dv = opusb_device(g->device, &err);
    if (dv) cudaStreamSynchronize(dv->stream);
  }
  enc_pipe_free(g->pipe);
  g->pipe = nullptr;
  cudaFree(g->d_states); cudaFree(g->d_scratch); cudaFree(g->d_len);
  cudaFree(g->d_pcm); cudaFree(g->d_data);
  cudaFreeHost(g->h_pcm); cudaFreeHost(g->h_data); cudaFreeHost(g->h_len);
  free(g->members);
  g->magic = 0;
  free(g);
}

// Device-resident batch: d_pcm holds n rows of frame_size*channels int16 (row stride = that count),
// d_data n rows of max_data_bytes, d_len n lengths.  Asynchronous on the library stream
// (opus_b200_synchronize() waits for it).
int opus_encode_group_device(OpusEncoderGroup* g, const opus_int16* d_pcm, int frame_size, unsigned char* d_data,
                             opus_int32 max_data_bytes, opus_int32* d_len) {
  if (!valid(g) || !d_pcm || !d_data || !d_len || frame_size <= 0 || max_data_bytes <= 0) return OPUS_BAD_ARG;
  if (g->channels < 0) return OPUS_BAD_ARG;
  int err = 0;
  OpusbDevice* dv = opusb_device(g->device, &err);
  if (!dv) return err;
  std::lock_guard<std::mutex> lk(dv->mu);
  return group_launch(g, d_pcm, frame_size * g->channels, frame_size, d_data, max_data_bytes, max_data_bytes, d_len,
                      dv->stream);
}

// Host-buffer batch: pcm = n rows of frame_size*channels int16, data = n rows of max_data_bytes,
// out_len[n].  Copies in, encodes, copies packets and lengths back; synchronous.
int opus_encode_group_flat(OpusEncoderGroup* g, const opus_int16* pcm, int frame_size, unsigned char* data,
                           opus_int32 max_data_bytes, opus_int32* out_len) {
  if (!valid(g) || !pcm || !data || !out_len || frame_size <= 0 || max_data_bytes <= 0) return OPUS_BAD_ARG;
  if (g->channels < 0) return OPUS_BAD_ARG;
  int err = 0;
  OpusbDevice* dv = opusb_device(g->device, &err);
  if (!dv) return err;
  cudaStream_t stream = dv->stream;
  const size_t row = (size_t)frame_size * g->channels;
  const size_t pcm_bytes = row * sizeof(i16) * g->n, data_bytes = (size_t)max_data_bytes * g->n;
  int rc;
  if ((rc = grow_dev((void**)&g->d_pcm, &g->d_pcm_cap, pcm_bytes)) != OPUS_OK) return rc;
  if ((rc = grow_dev((void**)&g->d_data, &g->d_data_cap, data_bytes)) != OPUS_OK) return rc;
  cudaEvent_t done = nullptr;
  {
    std::lock_guard<std::mutex> lk(dv->mu);          // queueing order only; released before the wait
    // straight from / to the caller's buffers: page-locked ones go by DMA, pageable ones through the driver's staging
    CUDA_OK(cudaMemcpyAsync(g->d_pcm, pcm, pcm_bytes, cudaMemcpyHostToDevice, stream));
    rc = group_launch(g, g->d_pcm, (i32)row, frame_size, g->d_data, max_data_bytes, max_data_bytes, g->d_len, stream);
    if (rc != OPUS_OK) return rc;
    CUDA_OK(cudaMemcpyAsync(data, g->d_data, data_bytes, cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaMemcpyAsync(out_len, g->d_len, sizeof(i32) * g->n, cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
    CUDA_OK(cudaEventRecord(done, stream));
  }
  cudaError_t e = cudaEventSynchronize(done);
  cudaEventDestroy(done);
  CUDA_OK(e);
  return OPUS_OK;
}

// Asynchronous form of opus_encode_group_flat: returns once copies and kernel are queued; `data` and
// `out_len` are complete when opus_encode_group_wait() returns.  Keep `pcm`, `data`, `out_len` alive (and,
// for the copies to overlap with the kernel, page-locked) until then.  At most two calls are in flight per
// group; a third submit first waits for the oldest.  Calls on one group execute in submission order.
int opus_encode_group_submit(OpusEncoderGroup* g, const opus_int16* pcm, int frame_size, unsigned char* data,
                             opus_int32 max_data_bytes, opus_int32* out_len) {
  if (!valid(g) || !pcm || !data || !out_len || frame_size <= 0 || max_data_bytes <= 0) return OPUS_BAD_ARG;
  if (g->channels < 0) return OPUS_BAD_ARG;
  int err = 0;
  OpusbDevice* dv = opusb_device(g->device, &err);
  if (!dv) return err;
  cudaStream_t stream = dv->stream;
  if (!g->pipe) {
    EncPipe* p = (EncPipe*)calloc(1, sizeof(EncPipe));
    if (!p) return OPUS_ALLOC_FAIL;
    bool ok = cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking) == cudaSuccess;
    for (EncPipeSlot& sl : p->slot)
      ok = ok && cudaEventCreateWithFlags(&sl.in_done, cudaEventDisableTiming) == cudaSuccess &&
           cudaEventCreateWithFlags(&sl.k_done, cudaEventDisableTiming) == cudaSuccess &&
           cudaEventCreateWithFlags(&sl.out_done, cudaEventDisableTiming) == cudaSuccess &&
           cudaMalloc((void**)&sl.d_len, sizeof(i32) * g->n) == cudaSuccess;
    if (!ok) { enc_pipe_free(p); return OPUS_ALLOC_FAIL; }
    g->pipe = p;
  }
  EncPipe* p = g->pipe;
  EncPipeSlot& sl = p->slot[p->seq & 1];
  if (sl.busy) { CUDA_OK(cudaEventSynchronize(sl.out_done)); sl.busy = false; }
  const size_t row = (size_t)frame_size * g->channels;
  const size_t pcm_bytes = row * sizeof(i16) * g->n, data_bytes = (size_t)max_data_bytes * g->n;
  int rc;
  if ((rc = grow_dev((void**)&sl.d_pcm, &sl.d_pcm_cap, pcm_bytes)) != OPUS_OK) return rc;
  if ((rc = grow_dev((void**)&sl.d_data, &sl.d_data_cap, data_bytes)) != OPUS_OK) return rc;
  CUDA_OK(cudaMemcpyAsync(sl.d_pcm, pcm, pcm_bytes, cudaMemcpyHostToDevice, p->s_in));
  CUDA_OK(cudaEventRecord(sl.in_done, p->s_in));
  std::lock_guard<std::mutex> lk(dv->mu);
  CUDA_OK(cudaStreamWaitEvent(stream, sl.in_done, 0));
  rc = group_launch(g, sl.d_pcm, (i32)row, frame_size, sl.d_data, max_data_bytes, max_data_bytes, sl.d_len, stream);
  if (rc != OPUS_OK) return rc;
  CUDA_OK(cudaEventRecord(sl.k_done, stream));
  CUDA_OK(cudaStreamWaitEvent(p->s_out, sl.k_done, 0));
  CUDA_OK(cudaMemcpyAsync(data, sl.d_data, data_bytes, cudaMemcpyDeviceToHost, p->s_out));
  CUDA_OK(cudaMemcpyAsync(out_len, sl.d_len, sizeof(i32) * g->n, cudaMemcpyDeviceToHost, p->s_out));
  CUDA_OK(cudaEventRecord(sl.out_done, p->s_out));
  sl.busy = true;
  p->seq++;
  return OPUS_OK;
}

// Wait for every call submitted on this group.
int opus_encode_group_wait(OpusEncoderGroup* g) {
  if (!valid(g)) return OPUS_BAD_ARG;
  if (!g->pipe) return OPUS_OK;
  CUDA_OK(cudaSetDevice(g->device));
  for (int k = 0; k < 2; k++) {
    EncPipeSlot& sl = g->pipe->slot[(g->pipe->seq + k) & 1];     // oldest first
    if (sl.busy) { CUDA_OK(cudaEventSynchronize(sl.out_done)); sl.busy = false; }
  }
  return OPUS_OK;
}

// The reference call, one stream (opus_encoder.c:2012-2031): a group of one.
opus_int32 opus_encode(OpusEncoder* st, const opus_int16* pcm, int frame_size, unsigned char* data,
                       opus_int32 max_data_bytes) {
  if (!valid(st) || !pcm || !data) return OPUS_BAD_ARG;
  if (frame_size <= 0 || max_data_bytes <= 0) return OPUS_BAD_ARG;
  if (!st->self_group) {
    int err = 0;
    st->self_group = opus_encoder_group_create(&st, 1, &err);
    if (!st->self_group) return err;
  }
  const int mb = max_data_bytes > 3 * 1276 + 3 ? 3 * 1276 + 3 : max_data_bytes;   // three 20 ms frames at most (opus_encoder.c:980, 1380)
  opus_int32 len = 0;
  int rc = opus_encode_group_flat(st->self_group, pcm, frame_size, data, mb, &len);
  if (rc != OPUS_OK) return rc;
  return len;
}

// FIXED_POINT opus_encode_float (opus_encoder.c:1968-2008): FLOAT2INT16 on the host, then opus_encode.
opus_int32 opus_encode_float(OpusEncoder* st, const float* pcm, int frame_size, unsigned char* data,
                             opus_int32 max_data_bytes) {
  if (!valid(st) || !pcm || !data || frame_size <= 0) return OPUS_BAD_ARG;
  std::vector<opus_int16> in((size_t)frame_size * st->channels);
  for (size_t i = 0; i < in.size(); i++) {
    float x = pcm[i] * 32768.f;
    x = x > -32768.f ? x : -32768.f;
    x = x < 32767.f ? x : 32767.f;
    in[i] = (opus_int16)lrintf(x);
  }
  return opus_encode(st, in.data(), frame_size, data, max_data_bytes);
}

// Pointer-array batch over arbitrary encoders (mixed rates / channel counts allowed): streams are
// bucketed by nothing -- each call builds a transient group; for steady-state use make a group once.
int opus_encode_batch(OpusEncoder* const* st, int n, const opus_int16* const* pcm, int frame_size,
                      unsigned char* const* data, opus_int32 max_data_bytes, opus_int32* out_len) {
  if (!st || !pcm || !data || !out_len || n <= 0 || frame_size <= 0 || max_data_bytes <= 0) return OPUS_BAD_ARG;
  int err = 0;
  OpusEncoderGroup* g = opus_encoder_group_create(st, n, &err);
  if (!g) return err;
  int rc = OPUS_OK;
  if (g->channels < 0) rc = OPUS_BAD_ARG;   // a flat batch needs one channel count; mixed groups go one by one
  if (rc == OPUS_OK) {
    const int mb = max_data_bytes > 3 * 1276 + 3 ? 3 * 1276 + 3 : max_data_bytes;
    const size_t row = (size_t)frame_size * g->channels;
    std::vector<opus_int16> flat(row * n);
    std::vector<unsigned char> out((size_t)mb * n);
    for (int i = 0; i < n; i++) memcpy(&flat[row * i], pcm[i], row * sizeof(opus_int16));
    rc = opus_encode_group_flat(g, flat.data(), frame_size, out.data(), mb, out_len);
    if (rc == OPUS_OK)
      for (int i = 0; i < n; i++)
        if (out_len[i] > 0) memcpy(data[i], &out[(size_t)mb * i], (size_t)out_len[i]);
  }
  opus_encoder_group_destroy(g);
  return rc;
}

int opus_encoder_ctl(OpusEncoder* st, int request, ...) {
  if (!valid(st)) return OPUS_BAD_ARG;
  va_list ap;
  va_start(ap, request);
  int ret = OPUS_OK;
  int err = 0;
  OpusbDevice* dv = opusb_device(st->device, &err);
  if (!dv) { va_end(ap); return err; }
  cudaStreamSynchronize(dv->stream);
  if (request == OPUS_RESET_STATE) {
    OpusEncState* h = (OpusEncState*)malloc(sizeof(OpusEncState));
    if (!h) { va_end(ap); return OPUS_ALLOC_FAIL; }
    if (cudaMemcpy(h, st->d_state, sizeof(OpusEncState), cudaMemcpyDeviceToHost) != cudaSuccess) ret = OPUS_INTERNAL_ERROR;
    if (ret == OPUS_OK) {
      opus_enc_reset(h);
      if (cudaMemcpy(st->d_state, h, sizeof(OpusEncState), cudaMemcpyHostToDevice) != cudaSuccess) ret = OPUS_INTERNAL_ERROR;
    }
    free(h);
    va_end(ap);
    return ret;
  }
  if (request == CB_REQ_SET_ENERGY_MASK) {
    // opus_encoder.c:2462-2467: the reference keeps the pointer and reads 21 values per channel at the next
    // opus_encode(); here they are copied into the device state now.  NULL drops the mask.
    const opus_int16* mask = va_arg(ap, const opus_int16*);
    va_end(ap);
    char* d_celt = (char*)st->d_state + offsetof(OpusEncState, celt);
    const i32 has = mask ? 1 : 0;
    if (mask && cudaMemcpy(d_celt + offsetof(CeltEncState, energy_mask), mask, sizeof(i16) * CB_NBANDS * st->channels,
                           cudaMemcpyHostToDevice) != cudaSuccess)
      return OPUS_INTERNAL_ERROR;
    if (cudaMemcpy(d_celt + offsetof(CeltEncState, s) + offsetof(CeltEncScalars, has_mask), &has, sizeof(has),
                   cudaMemcpyHostToDevice) != cudaSuccess)
      return OPUS_INTERNAL_ERROR;
    if (mask) st->ext = true;
    return OPUS_OK;
  }
  // scalars only: read both scalar blocks, apply, write back when it was a setter
  OpusEncScalars o;
  CeltEncScalars cs;
  CeltEncScalars* d_cs = (CeltEncScalars*)((char*)st->d_state + offsetof(OpusEncState, celt));
  if (cudaMemcpy(&o, st->d_state, sizeof(o), cudaMemcpyDeviceToHost) != cudaSuccess ||
      cudaMemcpy(&cs, d_cs, sizeof(cs), cudaMemcpyDeviceToHost) != cudaSuccess) {
    va_end(ap);
    return OPUS_INTERNAL_ERROR;
  }
  if (opus_enc_ctl_is_get(request)) {
    opus_int32* out = va_arg(ap, opus_int32*);
    if (!out) { va_end(ap); return OPUS_BAD_ARG; }
    ret = opus_enc_ctl_scalars(&o, &cs, request, 0, out);
  } else {
    opus_int32 value = va_arg(ap, opus_int32);
    i32 dummy = 0;
    ret = opus_enc_ctl_scalars(&o, &cs, request, value, &dummy);
    if (ret == OPUS_OK && request == CB_REQ_SET_LFE && value) st->ext = true;
    if (ret == OPUS_OK) {
      if (cudaMemcpy(st->d_state, &o, sizeof(o), cudaMemcpyHostToDevice) != cudaSuccess ||
          cudaMemcpy(d_cs, &cs, sizeof(cs), cudaMemcpyHostToDevice) != cudaSuccess)
        ret = OPUS_INTERNAL_ERROR;
    }
  }
  va_end(ap);
  return ret;
}

// For the multistream surround encoder (opus_ms_enc_api.cu): where this encoder's CELT state keeps the surround
// mask of channel `side` (0 / 1) and the flag that says one is set; the surround analysis kernel writes both.
// Launches of this encoder use the kernel instance with the surround terms from now on.
int opusb_internal_encoder_mask_slots(OpusEncoder* st, int side, opus_int16** mask, opus_int32** has_mask) {
  if (!valid(st) || side < 0 || side >= st->channels || !mask || !has_mask) return OPUS_BAD_ARG;
  char* d_celt = (char*)st->d_state + offsetof(OpusEncState, celt);
  *mask = (opus_int16*)(d_celt + offsetof(CeltEncState, energy_mask)) + CB_NBANDS * side;
  *has_mask = (opus_int32*)(d_celt + offsetof(CeltEncState, s) + offsetof(CeltEncScalars, has_mask));
  st->ext = true;
  return OPUS_OK;
}

int opus_encoder_get_final_range_batch(OpusEncoder* const* st, int n, opus_uint32* out) {
  if (!st || !out || n <= 0 || !valid(st[0])) return OPUS_BAD_ARG;
  int err = 0;
  OpusbDevice* dv = opusb_device(st[0]->device, &err);
  if (!dv) return err;
  CUDA_OK(cudaStreamSynchronize(dv->stream));
  for (int i = 0; i < n; i++) {
    if (!valid(st[i])) return OPUS_BAD_ARG;
    CUDA_OK(cudaMemcpy(&out[i], (char*)st[i]->d_state + offsetof(OpusEncScalars, rangeFinal), sizeof(opus_uint32),
                       cudaMemcpyDeviceToHost));
  }
  return OPUS_OK;
}

// per-launch timing of the encode kernel (CUDA events on the launching stream), for bench.py
int opus_b200_encode_profile_begin(void) { g_enc_prof_on = true; return OPUS_OK; }   // (one GPU at a time: bench use)
int opus_b200_encode_profile_end(double* ms, long long* count) {
  g_enc_prof_on = false;
  cudaDeviceSynchronize();
  std::lock_guard<std::mutex> lk(g_enc_prof_mu);
  double total = 0; long long c = 0;
  for (auto& r : g_enc_prof) {
    float t = 0;
    if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) { total += t; c++; }
    cudaEventDestroy(r.a); cudaEventDestroy(r.b);
  }
  g_enc_prof.clear();
  if (ms) *ms = total;
  if (count) *count = c;
  return OPUS_OK;
}

}  // extern "C"
