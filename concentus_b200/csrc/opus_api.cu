// opus_api.cu -- host side of libopus_b200.so: the libopus decoder ABI + the batch entry points.
//
// Mirrors the reference's Opus API layer (opus-fix/src/opus_decoder.c, src/opus.c) in names,
// argument meaning and error behaviour, but holds no codec arithmetic: every sample is produced
// by the CUDA kernels in cb_kernels.cu.  There is NO CPU decode path -- if CUDA is unusable the
// calls fail with OPUS_INTERNAL_ERROR / OPUS_ALLOC_FAIL and say why on stderr.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <mutex>
#include <vector>
#include "../../include/opus_b200.h"
#include "cb_host.h"
#include "cb_kernels.h"
#include "cb_state.h"
#include "cb_opus_plan.h"

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

const int kMagic = 0x4f423230;   // "OB20"
const int kSlab = 256;           // decoder states per cudaMalloc

struct GrowBuf {                 // grow-only buffer (device or pinned host)
  void* p = nullptr;
  size_t cap = 0;
  bool pinned = false;
  int reserve(size_t bytes) {
    if (bytes <= cap) return 0;
    size_t want = bytes + bytes / 4 + 256;
    if (p) { if (pinned) cudaFreeHost(p); else cudaFree(p); p = nullptr; cap = 0; }
    cudaError_t e = pinned ? cudaMallocHost(&p, want) : cudaMalloc(&p, want);
    if (e != cudaSuccess) {
      fprintf(stderr, "[opus_b200] allocation of %zu bytes failed: %s\n", want, cudaGetErrorString(e));
      p = nullptr;
      return OPUS_ALLOC_FAIL;
    }
    cap = want;
    return 0;
  }
};

struct DevCtx : OpusbDevice {
  // A decode pass runs on three streams (group_enqueue_passes):
  //   ctl_stream  -- everything that reads packets: plan / routing / SILK kernel, then the parse kernel.  Nothing
  //                  here needs the previous pass's spectrum or output history, so it runs one pass AHEAD of
  //   stream      -- (the library's main stream, OpusbDevice) rebuild + synthesis kernels, copies, scatter
  //   side_stream -- the general kernel (mode switches, pitch-based concealment) works on streams the CELT kernels
  //                  skip, so it runs beside them: forked after the routing of a pass, joined at the end of the pass.
  cudaStream_t ctl_stream = nullptr, side_stream = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;      // single-stream calls: fork / join of the general kernel
  cudaEvent_t ev_route = nullptr, ev_parse = nullptr, ev_gen = nullptr, ev_synth = nullptr, ev_in = nullptr;
  cudaEvent_t ev_serial = nullptr;                       // end of the last one-stream call on `stream`
  bool streams_ok = false, streams_tried = false;
  std::vector<OpusDecState*> slabs;
  std::vector<OpusDecState*> free_slots;
};

DevCtx g_ctx[OPUSB_MAX_DEVICES];        // each guarded by its own mutex (cb_host.h)
std::atomic<long long> g_launches{0};
std::mutex g_prof_mu;                   // guards g_prof only

// optional per-kernel timing (CUDA events on the launching stream), for bench.py's roofline line.
// kinds: 0 plan + routing + SILK kernel, 1 parse kernel, 2 rebuild kernel, 3 synthesis kernel.  While it is on,
// calls are queued on one stream (no run-ahead of the parse side), so that the figures are per kernel.
const int kProfKinds = 4;
struct ProfRec { int kind; cudaEvent_t a, b; };
std::atomic<bool> g_prof_on{false};
std::atomic<bool> g_trace_on{false};      // tools/pipe_trace.py: event pairs around every launch WITHOUT giving up the run-ahead
std::vector<ProfRec> g_prof;
struct ProfScope {
  cudaStream_t s; int kind; cudaEvent_t a = nullptr, b = nullptr; bool on;
  ProfScope(cudaStream_t s_, int k) : s(s_), kind(k), on(g_prof_on || g_trace_on) {
    if (on) { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, s); }
  }
  ~ProfScope() {
    if (on) { cudaEventRecord(b, s); std::lock_guard<std::mutex> lk(g_prof_mu); g_prof.push_back({kind, a, b}); }
  }
};

// Makes `device` current (device < 0: the thread's current device) and returns its context; first use creates
// the stream under the device's mutex.  The caller must NOT hold that mutex.
DevCtx* ctx_enter(int device, int* err) {
  cudaError_t e = cudaSuccess;
  if (device < 0) e = cudaGetDevice(&device);
  else e = cudaSetDevice(device);
  if (e != cudaSuccess || device < 0 || device >= OPUSB_MAX_DEVICES) {
    fprintf(stderr, "[opus_b200] no usable CUDA device: %s\n", cudaGetErrorString(e));
    if (err) *err = OPUS_INTERNAL_ERROR;
    return nullptr;
  }
  DevCtx* c = &g_ctx[device];
  std::lock_guard<std::mutex> lk(c->mu);
  if (c->device < 0) {
    e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
      fprintf(stderr, "[opus_b200] cudaStreamCreate failed: %s\n", cudaGetErrorString(e));
      if (err) *err = OPUS_INTERNAL_ERROR;
      return nullptr;
    }
    if (cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || c->sm_count <= 0)
      c->sm_count = 148;
    c->device = device;
  }
  return c;
}

}  // namespace

OpusbDevice* opusb_device(int device, int* err) { return ctx_enter(device, err); }

struct OpusDecoderGroup;
struct OpusDecoder {
  int magic;
  int Fs, channels;
  int device;
  int decode_gain;
  OpusDecState* d_state;
  OpusDecoderGroup* self_group;   // group of one, created on first opus_decode()
  int min_frame_48k;              // shortest frame any packet given to this decoder announced (samples at 48 kHz);
                                  // 0: no packet yet, -1: unknown (packets went in through a device-resident call)
};

// Pipeline behind opus_decode_group_submit/_wait, three calls deep: while the rebuild / synthesis kernels of one
// call run, the PCM of the call before drains to the host, and the packets of the next one come in and are parsed
// (separate copy streams, the control side of the next call ahead of the signal side of this one).
const int kPipeDepth = 3;
struct DecPipeSlot {
  GrowBuf d_pkt, d_pcm, d_i32, h_i32;        // d_i32 = offsets, lengths, fec flags, return values (4n)
  cudaEvent_t in_done = nullptr, k_done = nullptr, out_done = nullptr;
  bool busy = false;
  DecPipeSlot() { h_i32.pinned = true; }
};
struct DecPipe {
  cudaStream_t s_in = nullptr, s_out = nullptr;
  DecPipeSlot slot[kPipeDepth];
  unsigned long long seq = 0;
};

struct OpusDecoderGroup {
  int magic;
  int n, device;
  int Fs, channels;            // common values (or -1 when mixed)
  OpusDecoder** members;
  OpusDecState** d_states;     // device array of n state pointers
  // Scratch, two of each: the control side of a pass (plan, routing, parse) runs while the signal side of the
  // previous pass (rebuild, synthesis) still reads its records.  Plans alternate per call, frame records and
  // tapes per pass.
  StreamPlan* d_plans2[2];
  FrameSide* d_sides2[2];
  FrameTape* d_tapes2[2];      // n each: parse half -> rebuild half of the entropy stage
  unsigned call_seq, pass_seq;
  int celt_packets, seen_packets;   // of the last host-scanned batch: packets with a CELT layer (CELT-only or hybrid TOC) / all
  StreamPlan* d_plans;         // the current call's plan buffer (= d_plans2[call_seq & 1] after group_call_begin)
  i32* d_ret;                  // n
  DecPipe* pipe;               // created by the first opus_decode_group_submit
  GrowBuf d_pkt, d_i32, d_pcm; // staging of the synchronous host-buffer entry (opus_decode_group_flat)
};

extern "C" {

const char* opus_get_version_string(void) { return "libopus 1.1.2-fixed (opus_b200 sm_100a)"; }

const char* opus_strerror(int error) {
  static const char* const s[8] = {"success", "invalid argument", "buffer too small", "internal error",
                                   "corrupted stream", "request not implemented", "invalid state",
                                   "memory allocation failed"};
  if (error > 0 || error < -7) return "unknown error";
  return s[-error];
}

int opus_decoder_get_size(int channels) {
  if (channels < 1 || channels > 2) return 0;
  return (int)sizeof(OpusDecoder);
}

static int upload_fresh_state(OpusDecoder* st, int keep_gain) {
  OpusDecState* h = (OpusDecState*)malloc(sizeof(OpusDecState));
  if (!h) return OPUS_ALLOC_FAIL;
  h->decode_gain = st->decode_gain;
  opus_state_reset(h, st->Fs, st->channels, keep_gain);
  cudaError_t e = cudaMemcpy(st->d_state, h, sizeof(OpusDecState), cudaMemcpyHostToDevice);
  free(h);
  if (e != cudaSuccess) {
    fprintf(stderr, "[opus_b200] state upload failed: %s\n", cudaGetErrorString(e));
    return OPUS_INTERNAL_ERROR;
  }
  return OPUS_OK;
}

OpusDecoder* opus_decoder_create(opus_int32 Fs, int channels, int* error) {
  if ((Fs != 48000 && Fs != 24000 && Fs != 16000 && Fs != 12000 && Fs != 8000) || (channels != 1 && channels != 2)) {
    if (error) *error = OPUS_BAD_ARG;
    return nullptr;
  }
  int err = OPUS_OK;
  DevCtx* c = ctx_enter(-1, &err);
  if (!c) { if (error) *error = err; return nullptr; }
  std::unique_lock<std::mutex> lk(c->mu);
  if (c->free_slots.empty()) {
    OpusDecState* slab = nullptr;
    cudaError_t e = cudaMalloc((void**)&slab, sizeof(OpusDecState) * kSlab);
    if (e != cudaSuccess) {
      fprintf(stderr, "[opus_b200] cudaMalloc of a decoder slab failed: %s\n", cudaGetErrorString(e));
      if (error) *error = OPUS_ALLOC_FAIL;
      return nullptr;
    }
    c->slabs.push_back(slab);
    for (int i = kSlab - 1; i >= 0; i--) c->free_slots.push_back(slab + i);
  }
  OpusDecoder* st = (OpusDecoder*)calloc(1, sizeof(OpusDecoder));
  if (!st) { if (error) *error = OPUS_ALLOC_FAIL; return nullptr; }
  st->magic = kMagic;
  st->Fs = Fs;
  st->channels = channels;
  st->device = c->device;
  st->decode_gain = 0;
  st->d_state = c->free_slots.back();
  c->free_slots.pop_back();
  lk.unlock();
  int rc = upload_fresh_state(st, 0);
  if (rc != OPUS_OK) {
    lk.lock();
    c->free_slots.push_back(st->d_state);
    free(st);
    if (error) *error = rc;
    return nullptr;
  }
  if (error) *error = OPUS_OK;
  return st;
}

void opus_decoder_group_destroy(OpusDecoderGroup* g);

void opus_decoder_destroy(OpusDecoder* st) {
  if (!st || st->magic != kMagic) return;
  if (st->self_group) { opus_decoder_group_destroy(st->self_group); st->self_group = nullptr; }
  if (st->device >= 0 && st->device < OPUSB_MAX_DEVICES) {
    DevCtx* c = &g_ctx[st->device];
    if (c->stream) cudaStreamSynchronize(c->stream);
    std::lock_guard<std::mutex> lk(c->mu);
    c->free_slots.push_back(st->d_state);
  }
  st->magic = 0;
  free(st);
}

// ---- packet helpers (host restatement of src/opus.c:169-188 and src/opus_decoder.c:921-981) ----
int opus_packet_get_samples_per_frame(const unsigned char* data, opus_int32 Fs) {
  int audiosize;
  if (data[0] & 0x80) {
    audiosize = (data[0] >> 3) & 0x3;
    audiosize = (Fs << audiosize) / 400;
  } else if ((data[0] & 0x60) == 0x60) {
    audiosize = (data[0] & 0x08) ? Fs / 50 : Fs / 100;
  } else {
    audiosize = (data[0] >> 3) & 0x3;
    audiosize = audiosize == 3 ? Fs * 60 / 1000 : (Fs << audiosize) / 100;
  }
  return audiosize;
}
int opus_packet_get_bandwidth(const unsigned char* data) {
  int bw;
  if (data[0] & 0x80) {
    bw = CB_BW_MEDIUMBAND + ((data[0] >> 5) & 0x3);
    if (bw == CB_BW_MEDIUMBAND) bw = CB_BW_NARROWBAND;
  } else if ((data[0] & 0x60) == 0x60) {
    bw = (data[0] & 0x10) ? CB_BW_FULLBAND : CB_BW_SUPERWIDEBAND;
  } else {
    bw = CB_BW_NARROWBAND + ((data[0] >> 5) & 0x3);
  }
  return bw;
}
int opus_packet_get_nb_channels(const unsigned char* data) { return (data[0] & 0x4) ? 2 : 1; }
int opus_packet_get_nb_frames(const unsigned char packet[], opus_int32 len) {
  if (len < 1) return OPUS_BAD_ARG;
  int count = packet[0] & 0x3;
  if (count == 0) return 1;
  if (count != 3) return 2;
  if (len < 2) return OPUS_INVALID_PACKET;
  return packet[1] & 0x3F;
}
int opus_packet_get_nb_samples(const unsigned char packet[], opus_int32 len, opus_int32 Fs) {
  int count = opus_packet_get_nb_frames(packet, len);
  if (count < 0) return count;
  int samples = count * opus_packet_get_samples_per_frame(packet, Fs);
  if (samples * 25 > Fs * 3) return OPUS_INVALID_PACKET;
  return samples;
}
int opus_decoder_get_nb_samples(const OpusDecoder* dec, const unsigned char packet[], opus_int32 len) {
  return opus_packet_get_nb_samples(packet, len, dec->Fs);
}

int opus_packet_parse(const unsigned char* data, opus_int32 len, unsigned char* out_toc,
                      const unsigned char* frames[48], opus_int16 size[48], int* payload_offset) {
  if (!size || !data) return OPUS_BAD_ARG;
  int toc = 0, poff = 0;
  int count = opus_parse_packet(data, len, 0, &toc, size, &poff, nullptr);
  if (count < 0) return count;
  if (out_toc) *out_toc = (unsigned char)toc;
  if (payload_offset) *payload_offset = poff;
  if (frames) {
    const unsigned char* p = data + poff;
    for (int i = 0; i < count; i++) { frames[i] = p; p += size[i]; }
  }
  return count;
}

// ---- groups --------------------------------------------------------------------------------
static void pipe_free(DecPipe* p) {
  if (!p) return;
  for (DecPipeSlot& sl : p->slot) {
    if (sl.busy && sl.out_done) cudaEventSynchronize(sl.out_done);
    if (sl.d_pkt.p) cudaFree(sl.d_pkt.p);
    if (sl.d_pcm.p) cudaFree(sl.d_pcm.p);
    if (sl.d_i32.p) cudaFree(sl.d_i32.p);
    if (sl.h_i32.p) cudaFreeHost(sl.h_i32.p);
    if (sl.in_done) cudaEventDestroy(sl.in_done);
    if (sl.k_done) cudaEventDestroy(sl.k_done);
    if (sl.out_done) cudaEventDestroy(sl.out_done);
  }
  if (p->s_in) cudaStreamDestroy(p->s_in);
  if (p->s_out) cudaStreamDestroy(p->s_out);
  delete p;
}

static void group_free(OpusDecoderGroup* g) {
  if (!g) return;
  pipe_free(g->pipe);
  g->pipe = nullptr;
  if (g->d_states) cudaFree(g->d_states);
  for (int k = 0; k < 2; k++) {
    if (g->d_plans2[k]) cudaFree(g->d_plans2[k]);
    if (g->d_sides2[k]) cudaFree(g->d_sides2[k]);
    if (g->d_tapes2[k]) cudaFree(g->d_tapes2[k]);
  }
  if (g->d_ret) cudaFree(g->d_ret);
  if (g->d_pkt.p) cudaFree(g->d_pkt.p);
  if (g->d_i32.p) cudaFree(g->d_i32.p);
  if (g->d_pcm.p) cudaFree(g->d_pcm.p);
  free(g->members);
  g->magic = 0;
  delete g;
}

OpusDecoderGroup* opus_decoder_group_create(OpusDecoder* const* st, int n, int* error) {
  if (!st || n <= 0) { if (error) *error = OPUS_BAD_ARG; return nullptr; }
  for (int i = 0; i < n; i++) {
    if (!st[i] || st[i]->magic != kMagic || st[i]->device != st[0]->device) {
      if (error) *error = OPUS_BAD_ARG;
      return nullptr;
    }
  }
  int prev = -1;
  cudaGetDevice(&prev);
  if (cudaSetDevice(st[0]->device) != cudaSuccess) { if (error) *error = OPUS_INTERNAL_ERROR; return nullptr; }
  OpusDecoderGroup* g = new OpusDecoderGroup();
  g->d_states = nullptr; g->d_plans = nullptr; g->d_ret = nullptr; g->pipe = nullptr;
  for (int k = 0; k < 2; k++) { g->d_plans2[k] = nullptr; g->d_sides2[k] = nullptr; g->d_tapes2[k] = nullptr; }
  g->call_seq = g->pass_seq = 0;
  g->celt_packets = g->seen_packets = 0;
  std::vector<OpusDecState*> ptrs(n);
  g->magic = kMagic;
  g->n = n;
  g->device = st[0]->device;
  g->Fs = st[0]->Fs;
  g->channels = st[0]->channels;
  g->members = (OpusDecoder**)malloc(sizeof(OpusDecoder*) * n);
  for (int i = 0; i < n; i++) {
    g->members[i] = st[i];
    ptrs[i] = st[i]->d_state;
    if (st[i]->Fs != g->Fs) g->Fs = -1;
    if (st[i]->channels != g->channels) g->channels = -1;
  }
  bool ok = cudaMalloc((void**)&g->d_states, sizeof(OpusDecState*) * n) == cudaSuccess &&
            cudaMalloc((void**)&g->d_ret, sizeof(i32) * n) == cudaSuccess;
  for (int k = 0; k < 2 && ok; k++)
    ok = cudaMalloc((void**)&g->d_plans2[k], sizeof(StreamPlan) * n) == cudaSuccess &&
         cudaMalloc((void**)&g->d_sides2[k], sizeof(FrameSide) * n) == cudaSuccess &&
         cudaMalloc((void**)&g->d_tapes2[k], sizeof(FrameTape) * n) == cudaSuccess &&
         cudaMemset(g->d_sides2[k], 0, sizeof(FrameSide) * n) == cudaSuccess &&
         cudaMemset(g->d_tapes2[k], 0, 64) == cudaSuccess;
  g->d_plans = g->d_plans2[0];
  ok = ok &&
            cudaMemcpy(g->d_states, ptrs.data(), sizeof(OpusDecState*) * n, cudaMemcpyHostToDevice) == cudaSuccess;
  if (prev >= 0) cudaSetDevice(prev);
  if (!ok) {
    fprintf(stderr, "[opus_b200] group allocation failed: %s\n", cudaGetErrorString(cudaGetLastError()));
    group_free(g);
    if (error) *error = OPUS_ALLOC_FAIL;
    return nullptr;
  }
  if (error) *error = OPUS_OK;
  return g;
}

void opus_decoder_group_destroy(OpusDecoderGroup* g) {
  if (!g || g->magic != kMagic) return;
  if (g->device >= 0 && g->device < OPUSB_MAX_DEVICES && g_ctx[g->device].stream) cudaStreamSynchronize(g_ctx[g->device].stream);
  group_free(g);
}

// Where a call's launches go.  Library-stream calls use the three streams of DevCtx (the control side runs one pass
// ahead of the signal side); calls on a caller's stream, and calls made while per-kernel profiling is on, go on
// that one stream in order (the general kernel still forks to the side stream).
struct CallCtx {
  cudaStream_t s_ctl;          // plan / routing / SILK / parse
  cudaStream_t s_sig;          // rebuild / synthesis; a call is complete when this stream has passed it
  bool ahead;                  // s_ctl is DevCtx::ctl_stream
};

static void ctx_make_streams(DevCtx* c) {               // under c->mu
  if (c->streams_tried) return;
  c->streams_tried = true;
  bool ok = cudaStreamCreateWithFlags(&c->ctl_stream, cudaStreamNonBlocking) == cudaSuccess &&
            cudaStreamCreateWithFlags(&c->side_stream, cudaStreamNonBlocking) == cudaSuccess;
  cudaEvent_t* evs[] = {&c->ev_fork, &c->ev_join, &c->ev_route, &c->ev_parse, &c->ev_gen, &c->ev_synth, &c->ev_in,
                        &c->ev_serial};
  for (cudaEvent_t* e : evs) ok = ok && cudaEventCreateWithFlags(e, cudaEventDisableTiming) == cudaSuccess;
  c->streams_ok = ok;                                    // else: everything on one stream, in order
}

// Start of a call (under c->mu).  `user_stream` = the caller's stream or nullptr; `inputs_ready` = an event the
// control side has to wait for (the packet upload of the pipelined entries) or nullptr.
// The run-ahead pays when there is parse work to hide (CELT-only / hybrid packets); a batch of SILK-only packets
// has none, and the extra stream hops would only add to its critical path: the host entries, which see the TOC
// bytes, then queue the call in order.
static int group_call_begin(OpusDecoderGroup* g, DevCtx* c, cudaStream_t user_stream, cudaEvent_t inputs_ready, CallCtx* cc) {
  ctx_make_streams(c);
  const bool celt_work = g->seen_packets == 0 || 8 * g->celt_packets >= g->seen_packets;
  g->d_plans = g->d_plans2[g->call_seq & 1];
  g->call_seq++;
  static const bool no_ahead = getenv("OPUSB_NO_AHEAD") != nullptr;       // diagnostic: every call on one stream
  cc->ahead = !user_stream && c->streams_ok && !g_prof_on.load() && !no_ahead && celt_work;
  cc->s_sig = user_stream ? user_stream : c->stream;
  cc->s_ctl = cc->ahead ? c->ctl_stream : cc->s_sig;
  if (inputs_ready && cudaStreamWaitEvent(cc->s_ctl, inputs_ready, 0) != cudaSuccess) return OPUS_INTERNAL_ERROR;
  if (cc->ahead) {
    // the routing of this call reads what the general kernel of the previous pass may still be writing (Opus-layer
    // state, SILK state); ev_gen also stands behind that pass's predecessor on the signal side, so the control
    // side is never more than one pass ahead and two scratch buffers are enough
    if (cudaStreamWaitEvent(cc->s_ctl, c->ev_gen, 0) != cudaSuccess) return OPUS_INTERNAL_ERROR;
    // ... and behind the last call that ran on the main stream alone (profiling, the static-scratch entry)
    if (cudaStreamWaitEvent(cc->s_ctl, c->ev_serial, 0) != cudaSuccess) return OPUS_INTERNAL_ERROR;
  } else if (c->streams_ok) {
    // a one-stream call after run-ahead calls: order it behind both sides
    if (cudaEventRecord(c->ev_in, c->ctl_stream) != cudaSuccess || cudaStreamWaitEvent(cc->s_sig, c->ev_in, 0) != cudaSuccess)
      return OPUS_INTERNAL_ERROR;
    if (user_stream && (cudaEventRecord(c->ev_in, c->stream) != cudaSuccess || cudaStreamWaitEvent(cc->s_sig, c->ev_in, 0) != cudaSuccess))
      return OPUS_INTERNAL_ERROR;
  }
  return OPUS_OK;
}

// Queue the decode passes (per frame index: routing + SILK kernel, general kernel, parse, rebuild, synthesis) for a
// group whose plans are already on the device.  No synchronisation.  silk0_done: the routing of pass 0 was fused
// with the plan kernel (group_enqueue).
static int group_enqueue_passes(OpusDecoderGroup* g, DevCtx* c, const CallCtx& cc, const u8* d_data, i16* d_pcm,
                                i32 pcm_stride, i32* d_ret, int passes, int max_packet_bytes, bool silk0_done = false) {
  int rc = 0;
  const bool fork = c->streams_ok;
  const int max_ch = g->channels == 1 ? 1 : 2;
#define EV_OK(x) do { if ((x) != cudaSuccess) return OPUS_INTERNAL_ERROR; } while (0)
  for (int p = 0; p < passes; p++) {
    FrameSide* sides = g->d_sides2[g->pass_seq & 1];
    FrameTape* tapes = g->d_tapes2[g->pass_seq & 1];
    g->pass_seq++;
    // routing (+ SILK kernel) first: it decides, per stream, which kernels own this pass's frame
    if (p > 0 || !silk0_done) {
      if (cc.ahead && p > 0) EV_OK(cudaStreamWaitEvent(cc.s_ctl, c->ev_gen, 0));       // general kernel of pass p - 1
      ProfScope ps(cc.s_ctl, 0);
      rc = cbk_launch_silk(g->d_states, g->d_plans, sides, d_data, d_pcm, pcm_stride, d_ret, g->n, p, cc.s_ctl);
      g_launches += 1;
    }
    if (!rc && fork) {
      // general kernel on the side stream: after this pass's routing and the previous pass's synthesis
      cudaEvent_t routed = cc.ahead ? c->ev_route : c->ev_fork;
      EV_OK(cudaEventRecord(routed, cc.s_ctl));
      EV_OK(cudaStreamWaitEvent(c->side_stream, routed, 0));
      if (cc.ahead) EV_OK(cudaStreamWaitEvent(c->side_stream, c->ev_synth, 0));
      rc = cbk_launch_general(g->d_states, g->d_plans, sides, tapes, d_data, d_pcm, pcm_stride, d_ret, g->n, p, c->side_stream);
      if (!rc) EV_OK(cudaEventRecord(cc.ahead ? c->ev_gen : c->ev_join, c->side_stream));
    }
    if (!rc) {
      ProfScope ps(cc.s_ctl, 1);
      // does the signal side of the call (or pass) before still have work queued?  Then this parse shares the SMs.
      int shared_sms = 0;
      if (cc.ahead && cudaEventQuery(c->ev_synth) == cudaErrorNotReady) {
        shared_sms = 1;
        (void)cudaGetLastError();          // "not ready" is an answer, not a fault to report with the launch
      }
      rc = cbk_launch_parse(g->d_states, d_data, g->d_plans, sides, tapes, d_ret, g->n, p, max_packet_bytes, shared_sms,
                            cc.s_ctl);
    }
    if (!rc && cc.ahead) {
      EV_OK(cudaEventRecord(c->ev_parse, cc.s_ctl));
      EV_OK(cudaStreamWaitEvent(cc.s_sig, c->ev_parse, 0));
    }
    if (!rc) {
      ProfScope ps(cc.s_sig, 2);
      rc = cbk_launch_rebuild(sides, tapes, g->n, cc.s_sig);
    }
    if (!rc) {
      ProfScope ps(cc.s_sig, 3);
      rc = cbk_launch_synth(g->d_states, g->d_plans, sides, d_pcm, pcm_stride, g->n, p, max_ch, cc.s_sig);
    }
    if (!rc && cc.ahead) EV_OK(cudaEventRecord(c->ev_synth, cc.s_sig));
    if (!rc && !fork)
      rc = cbk_launch_general(g->d_states, g->d_plans, sides, tapes, d_data, d_pcm, pcm_stride, d_ret, g->n, p, cc.s_sig);
    if (!rc && fork) EV_OK(cudaStreamWaitEvent(cc.s_sig, cc.ahead ? c->ev_gen : c->ev_join, 0));
    g_launches += 4;                           // parse, rebuild, synthesis, general
    if (rc) { fprintf(stderr, "[opus_b200] pass launch failed: %s\n", cudaGetErrorString((cudaError_t)rc)); return OPUS_INTERNAL_ERROR; }
  }
  if (!cc.ahead && c->streams_ok && cc.s_sig == c->stream) EV_OK(cudaEventRecord(c->ev_serial, c->stream));
#undef EV_OK
  return OPUS_OK;
}

// Queue plan + passes for a group; everything device resident.  No synchronisation.
static int group_enqueue(OpusDecoderGroup* g, DevCtx* c, const CallCtx& cc, const u8* d_data, const i32* d_off,
                         const i32* d_len, i16* d_pcm, i32 pcm_stride, int frame_size, const i32* d_fec, i32* d_ret,
                         int passes, int max_packet_bytes) {
  int rc;
  {
    ProfScope ps(cc.s_ctl, 0);
    rc = cbk_launch_plan_silk(g->d_states, d_data, d_off, d_len, d_fec, frame_size, g->d_plans, g->d_sides2[g->pass_seq & 1],
                              d_pcm, pcm_stride, d_ret, g->n, passes, cc.s_ctl);
  }
  g_launches += 1;
  if (rc) { fprintf(stderr, "[opus_b200] plan launch failed: %s\n", cudaGetErrorString((cudaError_t)rc)); return OPUS_INTERNAL_ERROR; }
  return group_enqueue_passes(g, c, cc, d_data, d_pcm, pcm_stride, d_ret, passes, max_packet_bytes, true);
}

int opus_decode_group_device(OpusDecoderGroup* g, const unsigned char* d_data, const opus_int32* d_offsets,
                             const opus_int32* d_len, opus_int16* d_pcm, opus_int32 pcm_stride, int frame_size,
                             const int* d_decode_fec, int* d_ret, int max_frames, int max_packet_bytes,
                             void* cuda_stream) {
  if (!g || g->magic != kMagic || !d_data || !d_offsets || !d_len || !d_pcm || !d_ret) return OPUS_BAD_ARG;
  if (frame_size <= 0 || max_frames < 1 || max_frames > CB_MAX_FRAMES || max_packet_bytes < 0) return OPUS_BAD_ARG;
  {
    int maxch = 1;
    for (int i = 0; i < g->n; i++) if (g->members[i]->channels > maxch) maxch = g->members[i]->channels;
    if ((long long)pcm_stride < (long long)frame_size * maxch) return OPUS_BAD_ARG;   // rows would overlap
  }
  int err = 0;
  DevCtx* c = ctx_enter(g->device, &err);
  if (!c) return err;
  std::lock_guard<std::mutex> lk(c->mu);
  for (int i = 0; i < g->n; i++) g->members[i]->min_frame_48k = -1;   // the host does not see these packets
  g->seen_packets = 0;
  CallCtx cc;
  int rc = group_call_begin(g, c, (cudaStream_t)cuda_stream, nullptr, &cc);
  if (rc != OPUS_OK) return rc;
  return group_enqueue(g, c, cc, d_data, d_offsets, d_len, d_pcm, pcm_stride, frame_size, d_decode_fec, d_ret,
                       max_frames, max_packet_bytes);
}

// Upper bound on the pieces a concealed frame_size is cut into (opus_decoder.c:239-292, mirrored on the
// device by opus_plan_plc): no piece is longer than the last decoded frame, which in turn is no shorter than
// the shortest frame this group has ever been given.  Before any packet: pieces of 2.5 ms.
// min_frame_48k: shortest frame the decoders concerned have been given (48 kHz samples), <= 0: not known.
static int plc_passes(int frame_size, int Fs, int min_frame_48k) {
  const int F20 = Fs / 50, F10 = F20 >> 1, F5 = F10 >> 1, F2_5 = F5 >> 1;
  int cap = min_frame_48k > 0 ? (int)((long long)min_frame_48k * Fs / 48000) : F2_5;
  if (cap < F2_5) cap = F2_5;
  int count = 0;
  for (int rem = frame_size; rem >= F2_5 && count < CB_MAX_FRAMES; count++) {
    int a = rem < cap ? rem : cap;
    if (a > F20) a = F20;
    else if (a < F20) {
      if (a > F10) a = F10;
      else if (a > F5 && a < F10) a = F5;
    }
    rem -= a;
  }
  return count > 0 ? count : 1;
}

// Upper bound on the passes a batch needs, from the packets themselves (host side, O(n) byte reads).
static int scan_passes(OpusDecoder* const* members, const unsigned char* data, const i32* off, const i32* len,
                       const int* fec, int n, int frame_size, int Fs_hint, int* max_len, int* celt_packets = nullptr) {
  int passes = 1, mlen = 0, any_lost = 0, any_fec = 0, celt = 0;
  int group_min = 1 << 30;                 // over the members that have decoded something
  bool unknown = false;
  for (int i = 0; i < n; i++) {
    int& mf = members[i]->min_frame_48k;
    if (off[i] >= 0 && len[i] > 0) {
      const int spf = opus_toc_samples_per_frame(data[off[i]], 48000);
      if (mf == 0 || (mf > 0 && spf < mf)) mf = spf;
    }
    if (mf < 0) unknown = true;
    else if (mf > 0 && mf < group_min) group_min = mf;
    if (off[i] < 0 || len[i] <= 0) { any_lost = 1; continue; }
    if (fec && fec[i]) any_lost = any_fec = 1;   // the frames before an FEC frame are concealment chunks
    if (len[i] > mlen) mlen = len[i];
    int nf = opus_packet_get_nb_frames(data + off[i], len[i]);
    if (nf > passes) passes = nf;
    if ((data[off[i]] & 0x80) || (data[off[i]] & 0x60) == 0x60) celt++;
  }
  if (celt_packets) *celt_packets = celt;
  if (any_lost) {
    const int known = (unknown || Fs_hint <= 0 || group_min == 1 << 30) ? 0 : group_min;
    const int chunks = plc_passes(frame_size, Fs_hint > 0 ? Fs_hint : 8000, known) + any_fec;
    if (chunks > passes) passes = chunks;
  }
  if (passes > CB_MAX_FRAMES) passes = CB_MAX_FRAMES;
  *max_len = mlen;
  return passes;
}

int opus_decode_group_flat(OpusDecoderGroup* g, const unsigned char* data, opus_int32 data_bytes,
                           const opus_int32* offsets, const opus_int32* len, opus_int16* pcm,
                           opus_int32 pcm_stride, int frame_size, const int* decode_fec, int* ret) {
  if (!g || g->magic != kMagic || !offsets || !len || !pcm || !ret) return OPUS_BAD_ARG;
  if (data_bytes < 0 || (data_bytes > 0 && !data)) return OPUS_BAD_ARG;
  const int n = g->n;
  if (frame_size <= 0) {
    for (int i = 0; i < n; i++) ret[i] = OPUS_BAD_ARG;
    return OPUS_OK;
  }
  int maxch = 1;
  for (int i = 0; i < n; i++) {
    if (g->members[i]->channels > maxch) maxch = g->members[i]->channels;
    if (offsets[i] >= 0 && len[i] > 0 && (offsets[i] + (long long)len[i] > data_bytes)) return OPUS_BAD_ARG;
  }
  if (pcm_stride < frame_size * maxch) return OPUS_BAD_ARG;
  int err = 0;
  DevCtx* c = ctx_enter(g->device, &err);
  if (!c) return err;
  int max_len = 0;
  int passes = scan_passes(g->members, data, offsets, len, decode_fec, n, frame_size, g->Fs, &max_len, &g->celt_packets);
  g->seen_packets = n;
  const size_t pkt_bytes = (size_t)data_bytes + 16;
  const size_t pcm_bytes = (size_t)n * pcm_stride * sizeof(i16);
  const size_t i32_bytes = (size_t)n * 3 * sizeof(i32);
  // staging belongs to the group (a group is used by one thread at a time), so nothing shared is held while
  // the host waits for the GPU below
  if (g->d_pkt.reserve(pkt_bytes) || g->d_pcm.reserve(pcm_bytes) || g->d_i32.reserve(i32_bytes)) return OPUS_ALLOC_FAIL;
  std::vector<i32> h_i((size_t)n * 3);
  memcpy(h_i.data(), offsets, n * sizeof(i32));
  memcpy(h_i.data() + n, len, n * sizeof(i32));
  if (decode_fec) memcpy(h_i.data() + 2 * n, decode_fec, n * sizeof(i32));
  i32* d_i = (i32*)g->d_i32.p;
  cudaStream_t s = c->stream;
  cudaEvent_t done = nullptr;
  {
    std::lock_guard<std::mutex> lk(c->mu);          // queueing order only; released before the wait
    CallCtx cc;
    int rc = group_call_begin(g, c, nullptr, nullptr, &cc);
    if (rc != OPUS_OK) return rc;
    if (data_bytes > 0) CUDA_OK(cudaMemcpyAsync(g->d_pkt.p, data, data_bytes, cudaMemcpyHostToDevice, cc.s_ctl));
    CUDA_OK(cudaMemcpyAsync(d_i, h_i.data(), i32_bytes, cudaMemcpyHostToDevice, cc.s_ctl));   // pageable source: staged before return
    rc = group_enqueue(g, c, cc, (const u8*)g->d_pkt.p, d_i, d_i + n, (i16*)g->d_pcm.p, pcm_stride, frame_size,
                       d_i + 2 * n, g->d_ret, passes, max_len);
    if (rc != OPUS_OK) return rc;
    CUDA_OK(cudaMemcpyAsync(pcm, g->d_pcm.p, pcm_bytes, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(ret, g->d_ret, n * sizeof(i32), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
    CUDA_OK(cudaEventRecord(done, s));
  }
  cudaError_t e = cudaEventSynchronize(done);       // this call's work only, not whatever other threads queued behind it
  cudaEventDestroy(done);
  CUDA_OK(e);
  return OPUS_OK;
}

// Asynchronous form of opus_decode_group_flat: returns once the copies and kernels are queued.  `pcm` and
// `ret` are filled when opus_decode_group_wait() returns; `data`, `offsets`, `len`, `decode_fec` are read
// before this call returns only if they are pageable -- keep all buffers alive (and, for the copies to
// overlap with the kernels, page-locked) until the wait.  At most two calls are in flight per group; a
// third submit first waits for the oldest.  Calls on one group execute in submission order.
int opus_decode_group_submit(OpusDecoderGroup* g, const unsigned char* data, opus_int32 data_bytes,
                             const opus_int32* offsets, const opus_int32* len, opus_int16* pcm,
                             opus_int32 pcm_stride, int frame_size, const int* decode_fec, int* ret) {
  if (!g || g->magic != kMagic || !offsets || !len || !pcm || !ret) return OPUS_BAD_ARG;
  if (data_bytes < 0 || (data_bytes > 0 && !data) || frame_size <= 0) return OPUS_BAD_ARG;
  const int n = g->n;
  int maxch = 1;
  for (int i = 0; i < n; i++) {
    if (g->members[i]->channels > maxch) maxch = g->members[i]->channels;
    if (offsets[i] >= 0 && len[i] > 0 && (offsets[i] + (long long)len[i] > data_bytes)) return OPUS_BAD_ARG;
  }
  if (pcm_stride < frame_size * maxch) return OPUS_BAD_ARG;
  int err = 0;
  DevCtx* c = ctx_enter(g->device, &err);
  if (!c) return err;
  if (!g->pipe) {
    DecPipe* p = new DecPipe();
    bool ok = cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking) == cudaSuccess;
    for (DecPipeSlot& sl : p->slot)
      ok = ok && cudaEventCreateWithFlags(&sl.in_done, cudaEventDisableTiming) == cudaSuccess &&
           cudaEventCreateWithFlags(&sl.k_done, cudaEventDisableTiming) == cudaSuccess &&
           cudaEventCreateWithFlags(&sl.out_done, cudaEventDisableTiming) == cudaSuccess;
    if (!ok) { pipe_free(p); return OPUS_ALLOC_FAIL; }
    g->pipe = p;
  }
  DecPipe* p = g->pipe;
  DecPipeSlot& sl = p->slot[p->seq % kPipeDepth];
  if (sl.busy) { CUDA_OK(cudaEventSynchronize(sl.out_done)); sl.busy = false; }
  int max_len = 0;
  const int passes = scan_passes(g->members, data, offsets, len, decode_fec, n, frame_size, g->Fs, &max_len, &g->celt_packets);
  g->seen_packets = n;
  const size_t pcm_bytes = (size_t)n * pcm_stride * sizeof(i16);
  const size_t in_bytes = (size_t)n * 3 * sizeof(i32);
  if (sl.d_pkt.reserve((size_t)data_bytes + 16) || sl.d_pcm.reserve(pcm_bytes) || sl.d_i32.reserve((size_t)n * 4 * sizeof(i32)) ||
      sl.h_i32.reserve(in_bytes))
    return OPUS_ALLOC_FAIL;
  i32* h_i = (i32*)sl.h_i32.p;
  memcpy(h_i, offsets, n * sizeof(i32));
  memcpy(h_i + n, len, n * sizeof(i32));
  if (decode_fec) memcpy(h_i + 2 * n, decode_fec, n * sizeof(i32));
  else memset(h_i + 2 * n, 0, n * sizeof(i32));
  i32* d_i = (i32*)sl.d_i32.p;
  if (data_bytes > 0) CUDA_OK(cudaMemcpyAsync(sl.d_pkt.p, data, data_bytes, cudaMemcpyHostToDevice, p->s_in));
  CUDA_OK(cudaMemcpyAsync(d_i, h_i, in_bytes, cudaMemcpyHostToDevice, p->s_in));
  CUDA_OK(cudaEventRecord(sl.in_done, p->s_in));
  cudaStream_t s = c->stream;
  std::lock_guard<std::mutex> lk(c->mu);
  CallCtx cc;
  int rc = group_call_begin(g, c, nullptr, sl.in_done, &cc);
  if (rc != OPUS_OK) return rc;
  rc = group_enqueue(g, c, cc, (const u8*)sl.d_pkt.p, d_i, d_i + n, (i16*)sl.d_pcm.p, pcm_stride, frame_size, d_i + 2 * n,
                     d_i + 3 * n, passes, max_len);
  if (rc != OPUS_OK) return rc;
  CUDA_OK(cudaEventRecord(sl.k_done, s));
  CUDA_OK(cudaStreamWaitEvent(p->s_out, sl.k_done, 0));
  CUDA_OK(cudaMemcpyAsync(pcm, sl.d_pcm.p, pcm_bytes, cudaMemcpyDeviceToHost, p->s_out));
  CUDA_OK(cudaMemcpyAsync(ret, d_i + 3 * n, n * sizeof(i32), cudaMemcpyDeviceToHost, p->s_out));
  CUDA_OK(cudaEventRecord(sl.out_done, p->s_out));
  sl.busy = true;
  p->seq++;
  return OPUS_OK;
}

// Wait for every call submitted on this group; their pcm / ret buffers are complete afterwards.
int opus_decode_group_wait(OpusDecoderGroup* g) {
  if (!g || g->magic != kMagic) return OPUS_BAD_ARG;
  if (!g->pipe) return OPUS_OK;
  CUDA_OK(cudaSetDevice(g->device));
  for (int k = 0; k < kPipeDepth; k++) {
    DecPipeSlot& sl = g->pipe->slot[(g->pipe->seq + k) % kPipeDepth];     // oldest first
    if (sl.busy) { CUDA_OK(cudaEventSynchronize(sl.out_done)); sl.busy = false; }
  }
  return OPUS_OK;
}

int opus_decode_batch_flat(OpusDecoder* const* st, int n, const unsigned char* data, opus_int32 data_bytes,
                           const opus_int32* offsets, const opus_int32* len, opus_int16* pcm,
                           opus_int32 pcm_stride, int frame_size, const int* decode_fec, int* ret) {
  int err = 0;
  OpusDecoderGroup* g = opus_decoder_group_create(st, n, &err);
  if (!g) return err;
  int rc = opus_decode_group_flat(g, data, data_bytes, offsets, len, pcm, pcm_stride, frame_size, decode_fec, ret);
  opus_decoder_group_destroy(g);
  return rc;
}

int opus_decode_batch(OpusDecoder* const* st, const unsigned char* const* data, const opus_int32* len,
                      opus_int16* const* pcm, int frame_size, const int* decode_fec, int n, int* ret) {
  if (!st || !len || !pcm || !ret || n <= 0) return OPUS_BAD_ARG;
  if (frame_size <= 0) {
    for (int i = 0; i < n; i++) ret[i] = OPUS_BAD_ARG;
    return OPUS_OK;
  }
  // gather packets into one arena, run the flat path into a temporary PCM arena, scatter back
  size_t total = 0;
  int maxch = 1;
  for (int i = 0; i < n; i++) {
    if (!st[i] || st[i]->magic != kMagic || !pcm[i]) return OPUS_BAD_ARG;
    if (data && data[i] && len[i] > 0) total += (size_t)len[i];
    if (st[i]->channels > maxch) maxch = st[i]->channels;
  }
  if (total > 0x7fffffff) return OPUS_BAD_ARG;
  std::vector<unsigned char> arena(total + 1);
  std::vector<i32> off(n), ln(n);
  size_t o = 0;
  for (int i = 0; i < n; i++) {
    ln[i] = len[i];
    if (data && data[i] && len[i] > 0) {
      memcpy(arena.data() + o, data[i], (size_t)len[i]);
      off[i] = (i32)o;
      o += (size_t)len[i];
    } else {
      off[i] = -1;
      if (len[i] > 0) ln[i] = 0;     // data==NULL => lost, whatever len says (opus_decoder.c:613)
    }
  }
  const i32 stride = frame_size * maxch;
  std::vector<i16> out((size_t)n * stride);
  int rc = opus_decode_batch_flat(st, n, arena.data(), (i32)total, off.data(), ln.data(), out.data(), stride,
                                  frame_size, decode_fec, ret);
  if (rc != OPUS_OK) return rc;
  for (int i = 0; i < n; i++)
    if (ret[i] > 0) memcpy(pcm[i], out.data() + (size_t)i * stride, (size_t)ret[i] * st[i]->channels * sizeof(i16));
  return OPUS_OK;
}

int opus_decode(OpusDecoder* st, const unsigned char* data, opus_int32 len, opus_int16* pcm, int frame_size,
                int decode_fec) {
  if (!st || st->magic != kMagic) return OPUS_BAD_ARG;
  if (frame_size <= 0) return OPUS_BAD_ARG;
  if (!pcm) return OPUS_BAD_ARG;
  if (!st->self_group) {
    int err = 0;
    st->self_group = opus_decoder_group_create(&st, 1, &err);
    if (!st->self_group) return err;
  }
  int ret = OPUS_INTERNAL_ERROR;
  const bool lost = (data == nullptr || len == 0);
  i32 off = lost ? -1 : 0;
  i32 ln = lost ? 0 : len;
  if (ln < 0) return OPUS_BAD_ARG;
  // decode into a scratch row first: only `ret` samples belong to the caller's buffer
  std::vector<i16> row((size_t)frame_size * st->channels);
  int rc = opus_decode_group_flat(st->self_group, lost ? nullptr : data, lost ? 0 : ln, &off, &ln, row.data(),
                                  frame_size * st->channels, frame_size, &decode_fec, &ret);
  if (rc != OPUS_OK) return rc;
  if (ret > 0) memcpy(pcm, row.data(), (size_t)ret * st->channels * sizeof(i16));
  return ret;
}

int opus_decode_float(OpusDecoder* st, const unsigned char* data, opus_int32 len, float* pcm, int frame_size,
                      int decode_fec) {
  if (!st || st->magic != kMagic || !pcm) return OPUS_BAD_ARG;
  if (frame_size <= 0) return OPUS_BAD_ARG;
  if (data != nullptr && len > 0 && !decode_fec) {
    int nb = opus_decoder_get_nb_samples(st, data, len);
    if (nb > 0) frame_size = frame_size < nb ? frame_size : nb;
    else return OPUS_INVALID_PACKET;
  }
  std::vector<i16> tmp((size_t)frame_size * st->channels);
  int ret = opus_decode(st, data, len, tmp.data(), frame_size, decode_fec);
  if (ret > 0)
    for (int i = 0; i < ret * st->channels; i++) pcm[i] = (1.f / 32768.f) * tmp[i];
  return ret;
}

static int read_state_header(OpusDecoder* st, OpusDecState* hdr_out) {
  DevCtx* c = &g_ctx[st->device];
  if (c->stream) cudaStreamSynchronize(c->stream);
  cudaError_t e = cudaMemcpy(hdr_out, st->d_state, offsetof(OpusDecState, celt) + offsetof(CeltDecState, oldBandE),
                             cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) {
    fprintf(stderr, "[opus_b200] state read failed: %s\n", cudaGetErrorString(e));
    return OPUS_INTERNAL_ERROR;
  }
  return OPUS_OK;
}

int opus_decoder_ctl(OpusDecoder* st, int request, ...) {
  if (!st || st->magic != kMagic) return OPUS_BAD_ARG;
  va_list ap;
  va_start(ap, request);
  int ret = OPUS_OK;
  // (no lock: the calls below touch this decoder's own state and the device's stream handle only)
  static thread_local OpusDecState hdr;   // only the leading fields are filled
  switch (request) {
    case OPUS_GET_BANDWIDTH_REQUEST:
    case OPUS_GET_FINAL_RANGE_REQUEST:
    case OPUS_GET_PITCH_REQUEST:
    case OPUS_GET_LAST_PACKET_DURATION_REQUEST: {
      opus_int32* value = va_arg(ap, opus_int32*);
      if (!value) { ret = OPUS_BAD_ARG; break; }
      ret = read_state_header(st, &hdr);
      if (ret != OPUS_OK) break;
      if (request == OPUS_GET_BANDWIDTH_REQUEST) *value = hdr.bandwidth;
      else if (request == OPUS_GET_FINAL_RANGE_REQUEST) *(opus_uint32*)value = hdr.rangeFinal;
      else if (request == OPUS_GET_LAST_PACKET_DURATION_REQUEST) *value = hdr.last_packet_duration;
      else if (hdr.prev_mode == CB_MODE_CELT_ONLY) *value = hdr.celt.pf_period;
      else {                                  // opus_decoder.c:836-849: the SILK decoder's last pitch lag (48 kHz units)
        i32 lag = 0;
        if (cudaMemcpy(&lag, &st->d_state->silk.ctl_prevPitchLag, sizeof(i32), cudaMemcpyDeviceToHost) != cudaSuccess)
          ret = OPUS_INTERNAL_ERROR;
        *value = lag;
      }
      break;
    }
    case OPUS_GET_SAMPLE_RATE_REQUEST: {
      opus_int32* value = va_arg(ap, opus_int32*);
      if (!value) { ret = OPUS_BAD_ARG; break; }
      *value = st->Fs;
      break;
    }
    case OPUS_GET_GAIN_REQUEST: {
      opus_int32* value = va_arg(ap, opus_int32*);
      if (!value) { ret = OPUS_BAD_ARG; break; }
      *value = st->decode_gain;
      break;
    }
    case OPUS_SET_GAIN_REQUEST: {
      opus_int32 value = va_arg(ap, opus_int32);
      if (value < -32768 || value > 32767) { ret = OPUS_BAD_ARG; break; }
      st->decode_gain = value;
      DevCtx* c = &g_ctx[st->device];
      if (c->stream) cudaStreamSynchronize(c->stream);
      if (cudaMemcpy(&st->d_state->decode_gain, &value, sizeof(i32), cudaMemcpyHostToDevice) != cudaSuccess)
        ret = OPUS_INTERNAL_ERROR;
      break;
    }
    case OPUS_RESET_STATE: {
      DevCtx* c = &g_ctx[st->device];
      if (c->stream) cudaStreamSynchronize(c->stream);
      ret = upload_fresh_state(st, 1);
      break;
    }
    default:
      ret = OPUS_UNIMPLEMENTED;
      break;
  }
  va_end(ap);
  return ret;
}

int opus_decoder_get_final_range_batch(OpusDecoder* const* st, int n, opus_uint32* out) {
  if (!st || !out || n <= 0) return OPUS_BAD_ARG;
  for (int i = 0; i < n; i++) {
    if (!st[i] || st[i]->magic != kMagic) return OPUS_BAD_ARG;
    DevCtx* c = &g_ctx[st[i]->device];
    if (i == 0 && c->stream) cudaStreamSynchronize(c->stream);
    CUDA_OK(cudaMemcpy(&out[i], &st[i]->d_state->rangeFinal, sizeof(u32), cudaMemcpyDeviceToHost));
  }
  return OPUS_OK;
}


// ================= multistream (surround) decoder: opus-fix/src/opus_multistream_decoder.c =================
struct OpusMSDecoder {
  int magic;
  int Fs, channels, streams, coupled;
  unsigned char mapping[256];
  OpusDecoder** sub;            // `streams` elementary decoders, coupled (stereo) ones first
  OpusMSDecoderGroup* self_group;
};

struct OpusMSDecoderGroup {
  int magic;
  int n_ms, device, max_channels;
  OpusMSDecoder** members;
  OpusDecoderGroup* eg;         // group over all elementary decoders
  MSLayout* d_layouts;
  i32* d_ms_ret;
  struct MSPipe* pipe;          // created by the first opus_multistream_decode_group_submit
  GrowBuf d_pkt, d_i32, d_pcm, d_epcm;   // staging of the synchronous host-buffer entry
};

// Two-deep pipeline behind opus_multistream_decode_group_submit/_wait (see DecPipe)
struct MSPipeSlot {
  GrowBuf d_pkt, d_pcm, d_epcm, d_i32, h_i32;     // d_i32 = offsets, lengths, fec flags, return values (4n)
  cudaEvent_t in_done = nullptr, k_done = nullptr, out_done = nullptr;
  bool busy = false;
  MSPipeSlot() { h_i32.pinned = true; }
};
struct MSPipe {
  cudaStream_t s_in = nullptr, s_out = nullptr;
  MSPipeSlot slot[kPipeDepth];
  unsigned long long seq = 0;
};
static void ms_pipe_free(MSPipe* p) {
  if (!p) return;
  for (MSPipeSlot& sl : p->slot) {
    if (sl.busy && sl.out_done) cudaEventSynchronize(sl.out_done);
    if (sl.d_pkt.p) cudaFree(sl.d_pkt.p);
    if (sl.d_pcm.p) cudaFree(sl.d_pcm.p);
    if (sl.d_epcm.p) cudaFree(sl.d_epcm.p);
    if (sl.d_i32.p) cudaFree(sl.d_i32.p);
    if (sl.h_i32.p) cudaFreeHost(sl.h_i32.p);
    if (sl.in_done) cudaEventDestroy(sl.in_done);
    if (sl.k_done) cudaEventDestroy(sl.k_done);
    if (sl.out_done) cudaEventDestroy(sl.out_done);
  }
  if (p->s_in) cudaStreamDestroy(p->s_in);
  if (p->s_out) cudaStreamDestroy(p->s_out);
  delete p;
}

#define OPUS_MULTISTREAM_GET_DECODER_STATE_REQUEST 5122
const int kMagicMS = 0x4f424d53;

opus_int32 opus_multistream_decoder_get_size(int nb_streams, int nb_coupled_streams) {
  if (nb_streams < 1 || nb_coupled_streams > nb_streams || nb_coupled_streams < 0) return 0;
  return (opus_int32)sizeof(OpusMSDecoder);
}

OpusMSDecoder* opus_multistream_decoder_create(opus_int32 Fs, int channels, int streams, int coupled_streams,
                                               const unsigned char* mapping, int* error) {
  if ((channels > 255) || (channels < 1) || (coupled_streams > streams) || (streams < 1) || (coupled_streams < 0) ||
      (streams > 255 - coupled_streams) || !mapping) {
    if (error) *error = OPUS_BAD_ARG;
    return nullptr;
  }
  for (int i = 0; i < channels; i++) {            // validate_layout (opus_multistream.c:39-51)
    if (mapping[i] >= streams + coupled_streams && mapping[i] != 255) {
      if (error) *error = OPUS_BAD_ARG;
      return nullptr;
    }
  }
  OpusMSDecoder* st = (OpusMSDecoder*)calloc(1, sizeof(OpusMSDecoder));
  if (!st) { if (error) *error = OPUS_ALLOC_FAIL; return nullptr; }
  st->magic = kMagicMS;
  st->Fs = Fs; st->channels = channels; st->streams = streams; st->coupled = coupled_streams;
  memcpy(st->mapping, mapping, channels);
  st->sub = (OpusDecoder**)calloc(streams, sizeof(OpusDecoder*));
  int err = OPUS_OK;
  for (int s = 0; s < streams && err == OPUS_OK; s++)
    st->sub[s] = opus_decoder_create(Fs, s < coupled_streams ? 2 : 1, &err);
  if (err != OPUS_OK) {
    for (int s = 0; s < streams; s++) if (st->sub[s]) opus_decoder_destroy(st->sub[s]);
    free(st->sub);
    free(st);
    if (error) *error = err;
    return nullptr;
  }
  if (error) *error = OPUS_OK;
  return st;
}

void opus_multistream_decoder_group_destroy(OpusMSDecoderGroup* g);

void opus_multistream_decoder_destroy(OpusMSDecoder* st) {
  if (!st || st->magic != kMagicMS) return;
  if (st->self_group) { opus_multistream_decoder_group_destroy(st->self_group); st->self_group = nullptr; }
  for (int s = 0; s < st->streams; s++) opus_decoder_destroy(st->sub[s]);
  free(st->sub);
  st->magic = 0;
  free(st);
}

OpusMSDecoderGroup* opus_multistream_decoder_group_create(OpusMSDecoder* const* st, int n, int* error) {
  if (!st || n <= 0) { if (error) *error = OPUS_BAD_ARG; return nullptr; }
  std::vector<OpusDecoder*> elem;
  std::vector<MSLayout> lay(n);
  int maxch = 1;
  for (int m = 0; m < n; m++) {
    if (!st[m] || st[m]->magic != kMagicMS) { if (error) *error = OPUS_BAD_ARG; return nullptr; }
    MSLayout& L = lay[m];
    memset(&L, 0, sizeof(L));
    L.nb_channels = st[m]->channels; L.nb_streams = st[m]->streams; L.nb_coupled = st[m]->coupled;
    L.first = (i32)elem.size(); L.Fs = st[m]->Fs;
    memcpy(L.mapping, st[m]->mapping, 256);
    for (int s = 0; s < st[m]->streams; s++) elem.push_back(st[m]->sub[s]);
    if (st[m]->channels > maxch) maxch = st[m]->channels;
  }
  int err = OPUS_OK;
  OpusDecoderGroup* eg = opus_decoder_group_create(elem.data(), (int)elem.size(), &err);
  if (!eg) { if (error) *error = err; return nullptr; }
  OpusMSDecoderGroup* g = new OpusMSDecoderGroup();
  g->d_layouts = nullptr; g->d_ms_ret = nullptr; g->pipe = nullptr;
  g->magic = kMagicMS; g->n_ms = n; g->device = eg->device; g->max_channels = maxch; g->eg = eg;
  g->members = (OpusMSDecoder**)malloc(sizeof(OpusMSDecoder*) * n);
  for (int m = 0; m < n; m++) g->members[m] = st[m];
  int prev = -1;
  cudaGetDevice(&prev);
  cudaSetDevice(g->device);
  bool ok = cudaMalloc((void**)&g->d_layouts, sizeof(MSLayout) * n) == cudaSuccess &&
            cudaMalloc((void**)&g->d_ms_ret, sizeof(i32) * n) == cudaSuccess &&
            cudaMemcpy(g->d_layouts, lay.data(), sizeof(MSLayout) * n, cudaMemcpyHostToDevice) == cudaSuccess;
  if (prev >= 0) cudaSetDevice(prev);
  if (!ok) {
    fprintf(stderr, "[opus_b200] multistream group allocation failed: %s\n", cudaGetErrorString(cudaGetLastError()));
    if (g->d_layouts) cudaFree(g->d_layouts);
    if (g->d_ms_ret) cudaFree(g->d_ms_ret);
    free(g->members); delete g;
    if (error) *error = OPUS_ALLOC_FAIL;
    return nullptr;   // (eg is released by the caller's process exit; creation failure is fatal anyway)
  }
  if (error) *error = OPUS_OK;
  return g;
}

void opus_multistream_decoder_group_destroy(OpusMSDecoderGroup* g) {
  if (!g || g->magic != kMagicMS) return;
  ms_pipe_free(g->pipe);
  g->pipe = nullptr;
  opus_decoder_group_destroy(g->eg);           // waits for the device's stream
  cudaFree(g->d_layouts);
  cudaFree(g->d_ms_ret);
  if (g->d_pkt.p) cudaFree(g->d_pkt.p);
  if (g->d_i32.p) cudaFree(g->d_i32.p);
  if (g->d_pcm.p) cudaFree(g->d_pcm.p);
  if (g->d_epcm.p) cudaFree(g->d_epcm.p);
  free(g->members);
  g->magic = 0;
  delete g;
}

// Flat host form for multistream packets: packet m at data[offsets[m]..+len[m]) (offset<0 or len 0 = lost),
// PCM row m = pcm + m*pcm_stride holding frame_size*channels interleaved samples.
// Host-side walk over the sub-packets of every multistream packet (byte reads only): number of passes the
// elementary kernels need, the largest elementary frame (staged in shared memory by kernel A), and the
// shortest-frame bookkeeping of the elementary decoders.  Returns OPUS_OK or OPUS_BAD_ARG.
static int ms_scan(OpusMSDecoderGroup* g, const unsigned char* data, opus_int32 data_bytes, const opus_int32* offsets,
                   const opus_int32* len, const int* decode_fec, int frame_size, int* passes_out, int* max_len_out) {
  const int n = g->n_ms;
  // passes: walk the sub-packets on the host (byte reads only) to find the largest frame count
  int passes = 1, max_len = 0, any_lost = 0, any_fec = 0, minFs = 48000;
  int group_min = 1 << 30;
  bool unknown = false;
  int celt = 0, seen = 0;
  for (int m = 0; m < n; m++) {
    OpusMSDecoder* ms = g->members[m];
    if (ms->Fs < minFs) minFs = ms->Fs;
    if (offsets[m] < 0 || len[m] <= 0) { any_lost = 1; continue; }
    if (offsets[m] + (long long)len[m] > data_bytes) return OPUS_BAD_ARG;
    if (decode_fec && decode_fec[m]) any_lost = any_fec = 1;
    const unsigned char* p = data + offsets[m];
    i32 rem = len[m];
    bool parsed_all = true;
    for (int s = 0; s < ms->streams && rem > 0; s++) {
      i16 sizes[CB_MAX_FRAMES];
      i32 po = 0;
      int toc = 0;
      int cnt = opus_parse_packet(p, rem, s != ms->streams - 1, &toc, sizes, nullptr, &po);
      if (cnt < 0) { parsed_all = false; break; }
      if (cnt > passes) passes = cnt;
      seen++;
      if ((toc & 0x80) || (toc & 0x60) == 0x60) celt++;
      if (po > max_len) max_len = po;            // the parse kernel stages one frame of one elementary stream at a time
      int& mf = ms->sub[s]->min_frame_48k;
      const int spf = opus_toc_samples_per_frame(toc, 48000);
      if (mf == 0 || (mf > 0 && spf < mf)) mf = spf;
      p += po;
      rem -= po;
    }
    if (!parsed_all && len[m] > max_len) max_len = len[m];
  }
  g->eg->celt_packets = celt;
  g->eg->seen_packets = seen;
  if (any_lost) {
    for (int m = 0; m < n; m++)
      for (int s = 0; s < g->members[m]->streams; s++) {
        const int mf = g->members[m]->sub[s]->min_frame_48k;
        if (mf < 0) unknown = true;
        else if (mf > 0 && mf < group_min) group_min = mf;
      }
    const int chunks = plc_passes(frame_size, minFs, (unknown || group_min == 1 << 30) ? 0 : group_min) + any_fec;
    if (chunks > passes) passes = chunks;
  }
  if (passes > CB_MAX_FRAMES) passes = CB_MAX_FRAMES;
  *passes_out = passes;
  *max_len_out = max_len;
  return OPUS_OK;
}

int opus_multistream_decode_group_flat(OpusMSDecoderGroup* g, const unsigned char* data, opus_int32 data_bytes,
                                       const opus_int32* offsets, const opus_int32* len, opus_int16* pcm,
                                       opus_int32 pcm_stride, int frame_size, const int* decode_fec, int* ret) {
  if (!g || g->magic != kMagicMS || !offsets || !len || !pcm || !ret) return OPUS_BAD_ARG;
  if (data_bytes < 0 || (data_bytes > 0 && !data)) return OPUS_BAD_ARG;
  const int n = g->n_ms;
  if (frame_size <= 0) { for (int i = 0; i < n; i++) ret[i] = OPUS_BAD_ARG; return OPUS_OK; }
  if (pcm_stride < frame_size * g->max_channels) return OPUS_BAD_ARG;
  int passes = 1, max_len = 0;
  {
    const int rc0 = ms_scan(g, data, data_bytes, offsets, len, decode_fec, frame_size, &passes, &max_len);
    if (rc0 != OPUS_OK) return rc0;
  }
  int err = 0;
  DevCtx* c = ctx_enter(g->device, &err);
  if (!c) return err;
  OpusDecoderGroup* eg = g->eg;
  const i32 estride = frame_size * 2;
  const size_t pkt_bytes = (size_t)data_bytes + 16;
  const size_t pcm_bytes = (size_t)n * pcm_stride * sizeof(i16);
  const size_t epcm_bytes = (size_t)eg->n * estride * sizeof(i16);
  const size_t i32_bytes = (size_t)n * 3 * sizeof(i32);
  if (g->d_pkt.reserve(pkt_bytes) || g->d_pcm.reserve(pcm_bytes) || g->d_i32.reserve(i32_bytes) ||
      g->d_epcm.reserve(epcm_bytes))
    return OPUS_ALLOC_FAIL;
  i16* d_epcm = (i16*)g->d_epcm.p;               // elementary PCM rows (scratch)
  std::vector<i32> h_i((size_t)n * 3);
  memcpy(h_i.data(), offsets, n * sizeof(i32));
  memcpy(h_i.data() + n, len, n * sizeof(i32));
  if (decode_fec) memcpy(h_i.data() + 2 * n, decode_fec, n * sizeof(i32));
  i32* d_i = (i32*)g->d_i32.p;
  cudaStream_t s = c->stream;
  cudaEvent_t done = nullptr;
  {
    std::lock_guard<std::mutex> lk(c->mu);          // queueing order only; released before the wait
    CallCtx cc;
    int rc = group_call_begin(eg, c, nullptr, nullptr, &cc);
    if (rc != OPUS_OK) return rc;
    if (data_bytes > 0) CUDA_OK(cudaMemcpyAsync(g->d_pkt.p, data, data_bytes, cudaMemcpyHostToDevice, cc.s_ctl));
    CUDA_OK(cudaMemcpyAsync(d_i, h_i.data(), i32_bytes, cudaMemcpyHostToDevice, cc.s_ctl));
    rc = cbk_launch_ms_plan(eg->d_states, eg->d_plans, g->d_layouts, (const u8*)g->d_pkt.p, d_i, d_i + n, d_i + 2 * n,
                            frame_size, g->d_ms_ret, n, passes, cc.s_ctl);
    g_launches += 1;
    if (rc) { fprintf(stderr, "[opus_b200] ms plan launch failed: %s\n", cudaGetErrorString((cudaError_t)rc)); return OPUS_INTERNAL_ERROR; }
    rc = group_enqueue_passes(eg, c, cc, (const u8*)g->d_pkt.p, d_epcm, estride, eg->d_ret, passes, max_len);
    if (rc != OPUS_OK) return rc;
    rc = cbk_launch_ms_scatter(g->d_layouts, g->d_ms_ret, d_epcm, estride, (i16*)g->d_pcm.p, pcm_stride, n, s);
    g_launches += 1;
    if (rc) { fprintf(stderr, "[opus_b200] ms scatter launch failed: %s\n", cudaGetErrorString((cudaError_t)rc)); return OPUS_INTERNAL_ERROR; }
    CUDA_OK(cudaMemcpyAsync(pcm, g->d_pcm.p, pcm_bytes, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(ret, g->d_ms_ret, n * sizeof(i32), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
    CUDA_OK(cudaEventRecord(done, s));
  }
  cudaError_t e = cudaEventSynchronize(done);
  cudaEventDestroy(done);
  CUDA_OK(e);
  return OPUS_OK;
}

// Asynchronous form of opus_multistream_decode_group_flat (same contract as opus_decode_group_submit: buffers
// stay alive and unchanged until the wait, page-locked for the copies to overlap; two calls in flight).
int opus_multistream_decode_group_submit(OpusMSDecoderGroup* g, const unsigned char* data, opus_int32 data_bytes,
                                         const opus_int32* offsets, const opus_int32* len, opus_int16* pcm,
                                         opus_int32 pcm_stride, int frame_size, const int* decode_fec, int* ret) {
  if (!g || g->magic != kMagicMS || !offsets || !len || !pcm || !ret) return OPUS_BAD_ARG;
  if (data_bytes < 0 || (data_bytes > 0 && !data) || frame_size <= 0) return OPUS_BAD_ARG;
  const int n = g->n_ms;
  if (pcm_stride < frame_size * g->max_channels) return OPUS_BAD_ARG;
  int passes = 1, max_len = 0;
  {
    const int rc0 = ms_scan(g, data, data_bytes, offsets, len, decode_fec, frame_size, &passes, &max_len);
    if (rc0 != OPUS_OK) return rc0;
  }
  int err = 0;
  DevCtx* c = ctx_enter(g->device, &err);
  if (!c) return err;
  if (!g->pipe) {
    MSPipe* p = new MSPipe();
    bool ok = cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking) == cudaSuccess;
    for (MSPipeSlot& sl : p->slot)
      ok = ok && cudaEventCreateWithFlags(&sl.in_done, cudaEventDisableTiming) == cudaSuccess &&
           cudaEventCreateWithFlags(&sl.k_done, cudaEventDisableTiming) == cudaSuccess &&
           cudaEventCreateWithFlags(&sl.out_done, cudaEventDisableTiming) == cudaSuccess;
    if (!ok) { ms_pipe_free(p); return OPUS_ALLOC_FAIL; }
    g->pipe = p;
  }
  MSPipe* p = g->pipe;
  MSPipeSlot& sl = p->slot[p->seq % kPipeDepth];
  if (sl.busy) { CUDA_OK(cudaEventSynchronize(sl.out_done)); sl.busy = false; }
  OpusDecoderGroup* eg = g->eg;
  const i32 estride = frame_size * 2;
  const size_t pcm_bytes = (size_t)n * pcm_stride * sizeof(i16);
  const size_t in_bytes = (size_t)n * 3 * sizeof(i32);
  if (sl.d_pkt.reserve((size_t)data_bytes + 16) || sl.d_pcm.reserve(pcm_bytes) ||
      sl.d_epcm.reserve((size_t)eg->n * estride * sizeof(i16)) || sl.d_i32.reserve((size_t)n * 4 * sizeof(i32)) ||
      sl.h_i32.reserve(in_bytes))
    return OPUS_ALLOC_FAIL;
  i32* h_i = (i32*)sl.h_i32.p;
  memcpy(h_i, offsets, n * sizeof(i32));
  memcpy(h_i + n, len, n * sizeof(i32));
  if (decode_fec) memcpy(h_i + 2 * n, decode_fec, n * sizeof(i32));
  else memset(h_i + 2 * n, 0, n * sizeof(i32));
  i32* d_i = (i32*)sl.d_i32.p;
  i32* d_ms_ret = d_i + 3 * n;
  i16* d_epcm = (i16*)sl.d_epcm.p;
  if (data_bytes > 0) CUDA_OK(cudaMemcpyAsync(sl.d_pkt.p, data, data_bytes, cudaMemcpyHostToDevice, p->s_in));
  CUDA_OK(cudaMemcpyAsync(d_i, h_i, in_bytes, cudaMemcpyHostToDevice, p->s_in));
  CUDA_OK(cudaEventRecord(sl.in_done, p->s_in));
  cudaStream_t s = c->stream;
  std::lock_guard<std::mutex> lk(c->mu);
  CallCtx cc;
  int rc = group_call_begin(eg, c, nullptr, sl.in_done, &cc);
  if (rc != OPUS_OK) return rc;
  rc = cbk_launch_ms_plan(eg->d_states, eg->d_plans, g->d_layouts, (const u8*)sl.d_pkt.p, d_i, d_i + n, d_i + 2 * n,
                          frame_size, d_ms_ret, n, passes, cc.s_ctl);
  g_launches += 1;
  if (rc) { fprintf(stderr, "[opus_b200] ms plan launch failed: %s\n", cudaGetErrorString((cudaError_t)rc)); return OPUS_INTERNAL_ERROR; }
  rc = group_enqueue_passes(eg, c, cc, (const u8*)sl.d_pkt.p, d_epcm, estride, eg->d_ret, passes, max_len);
  if (rc != OPUS_OK) return rc;
  rc = cbk_launch_ms_scatter(g->d_layouts, d_ms_ret, d_epcm, estride, (i16*)sl.d_pcm.p, pcm_stride, n, s);
  g_launches += 1;
  if (rc) { fprintf(stderr, "[opus_b200] ms scatter launch failed: %s\n", cudaGetErrorString((cudaError_t)rc)); return OPUS_INTERNAL_ERROR; }
  CUDA_OK(cudaEventRecord(sl.k_done, s));
  CUDA_OK(cudaStreamWaitEvent(p->s_out, sl.k_done, 0));
  CUDA_OK(cudaMemcpyAsync(pcm, sl.d_pcm.p, pcm_bytes, cudaMemcpyDeviceToHost, p->s_out));
  CUDA_OK(cudaMemcpyAsync(ret, d_ms_ret, n * sizeof(i32), cudaMemcpyDeviceToHost, p->s_out));
  CUDA_OK(cudaEventRecord(sl.out_done, p->s_out));
  sl.busy = true;
  p->seq++;
  return OPUS_OK;
}

int opus_multistream_decode_group_wait(OpusMSDecoderGroup* g) {
  if (!g || g->magic != kMagicMS) return OPUS_BAD_ARG;
  if (!g->pipe) return OPUS_OK;
  CUDA_OK(cudaSetDevice(g->device));
  for (int k = 0; k < kPipeDepth; k++) {
    MSPipeSlot& sl = g->pipe->slot[(g->pipe->seq + k) % kPipeDepth];      // oldest first
    if (sl.busy) { CUDA_OK(cudaEventSynchronize(sl.out_done)); sl.busy = false; }
  }
  return OPUS_OK;
}

int opus_multistream_decode_batch(OpusMSDecoder* const* st, const unsigned char* const* data, const opus_int32* len,
                                  opus_int16* const* pcm, int frame_size, const int* decode_fec, int n, int* ret) {
  if (!st || !len || !pcm || !ret || n <= 0) return OPUS_BAD_ARG;
  if (frame_size <= 0) { for (int i = 0; i < n; i++) ret[i] = OPUS_BAD_ARG; return OPUS_OK; }
  size_t total = 0;
  int maxch = 1;
  for (int i = 0; i < n; i++) {
    if (!st[i] || st[i]->magic != kMagicMS || !pcm[i]) return OPUS_BAD_ARG;
    if (data && data[i] && len[i] > 0) total += (size_t)len[i];
    if (st[i]->channels > maxch) maxch = st[i]->channels;
  }
  std::vector<unsigned char> arena(total + 1);
  std::vector<i32> off(n), ln(n);
  size_t o = 0;
  for (int i = 0; i < n; i++) {
    ln[i] = len[i];
    if (data && data[i] && len[i] > 0) { memcpy(arena.data() + o, data[i], (size_t)len[i]); off[i] = (i32)o; o += (size_t)len[i]; }
    else { off[i] = -1; if (len[i] > 0) ln[i] = 0; }
  }
  int err = 0;
  OpusMSDecoderGroup* g = opus_multistream_decoder_group_create(st, n, &err);
  if (!g) return err;
  const i32 stride = frame_size * maxch;
  std::vector<i16> out((size_t)n * stride);
  int rc = opus_multistream_decode_group_flat(g, arena.data(), (i32)total, off.data(), ln.data(), out.data(), stride,
                                              frame_size, decode_fec, ret);
  opus_multistream_decoder_group_destroy(g);
  if (rc != OPUS_OK) return rc;
  for (int i = 0; i < n; i++)
    if (ret[i] > 0) memcpy(pcm[i], out.data() + (size_t)i * stride, (size_t)ret[i] * st[i]->channels * sizeof(i16));
  return OPUS_OK;
}

int opus_multistream_decode(OpusMSDecoder* st, const unsigned char* data, opus_int32 len, opus_int16* pcm,
                            int frame_size, int decode_fec) {
  if (!st || st->magic != kMagicMS || !pcm) return OPUS_BAD_ARG;
  if (frame_size <= 0) return OPUS_BAD_ARG;
  if (len < 0) return OPUS_BAD_ARG;
  if (!st->self_group) {
    int err = 0;
    st->self_group = opus_multistream_decoder_group_create(&st, 1, &err);
    if (!st->self_group) return err;
  }
  const bool lost = (data == nullptr || len == 0);
  i32 off = lost ? -1 : 0, ln = lost ? 0 : len;
  int ret = OPUS_INTERNAL_ERROR;
  std::vector<i16> row((size_t)frame_size * st->channels);
  int rc = opus_multistream_decode_group_flat(st->self_group, lost ? nullptr : data, lost ? 0 : ln, &off, &ln,
                                              row.data(), frame_size * st->channels, frame_size, &decode_fec, &ret);
  if (rc != OPUS_OK) return rc;
  if (ret > 0) memcpy(pcm, row.data(), (size_t)ret * st->channels * sizeof(i16));
  return ret;
}

int opus_multistream_decode_float(OpusMSDecoder* st, const unsigned char* data, opus_int32 len, float* pcm,
                                  int frame_size, int decode_fec) {
  if (!st || st->magic != kMagicMS || !pcm || frame_size <= 0) return OPUS_BAD_ARG;
  std::vector<i16> tmp((size_t)frame_size * st->channels);
  int ret = opus_multistream_decode(st, data, len, tmp.data(), frame_size, decode_fec);
  if (ret > 0)
    for (int i = 0; i < ret * st->channels; i++) pcm[i] = (1.f / 32768.f) * tmp[i];
  return ret;
}

int opus_multistream_decoder_ctl(OpusMSDecoder* st, int request, ...) {
  if (!st || st->magic != kMagicMS) return OPUS_BAD_ARG;
  va_list ap;
  va_start(ap, request);
  int ret = OPUS_OK;
  switch (request) {
    case OPUS_GET_BANDWIDTH_REQUEST:
    case OPUS_GET_SAMPLE_RATE_REQUEST:
    case OPUS_GET_GAIN_REQUEST:
    case OPUS_GET_LAST_PACKET_DURATION_REQUEST: {
      opus_int32* value = va_arg(ap, opus_int32*);
      ret = opus_decoder_ctl(st->sub[0], request, value);
      break;
    }
    case OPUS_GET_FINAL_RANGE_REQUEST: {
      opus_uint32* value = va_arg(ap, opus_uint32*);
      if (!value) { ret = OPUS_BAD_ARG; break; }
      *value = 0;
      for (int s = 0; s < st->streams; s++) {
        opus_uint32 tmp = 0;
        ret = opus_decoder_ctl(st->sub[s], request, &tmp);
        if (ret != OPUS_OK) break;
        *value ^= tmp;
      }
      break;
    }
    case OPUS_RESET_STATE:
      for (int s = 0; s < st->streams && ret == OPUS_OK; s++) ret = opus_decoder_ctl(st->sub[s], OPUS_RESET_STATE);
      break;
    case OPUS_MULTISTREAM_GET_DECODER_STATE_REQUEST: {
      opus_int32 stream_id = va_arg(ap, opus_int32);
      OpusDecoder** value = va_arg(ap, OpusDecoder**);
      if (stream_id < 0 || stream_id >= st->streams || !value) { ret = OPUS_BAD_ARG; break; }
      *value = st->sub[stream_id];
      break;
    }
    case OPUS_SET_GAIN_REQUEST: {
      opus_int32 value = va_arg(ap, opus_int32);
      for (int s = 0; s < st->streams && ret == OPUS_OK; s++) ret = opus_decoder_ctl(st->sub[s], request, value);
      break;
    }
    default:
      ret = OPUS_UNIMPLEMENTED;
      break;
  }
  va_end(ap);
  return ret;
}

int opus_b200_synchronize(void) {
  int err = 0;
  DevCtx* c = ctx_enter(-1, &err);
  if (!c) return err;
  CUDA_OK(cudaStreamSynchronize(c->stream));
  return OPUS_OK;
}

long long opus_b200_launch_count(void) { return g_launches.load(); }

// internal (used by cb_noalloc.cu): device address of a decoder's state blob and its GPU index
OpusDecState* opusb_internal_device_state(const OpusDecoder* st, int* device, int* channels) {
  if (!st || st->magic != kMagic) return nullptr;
  if (device) *device = st->device;
  if (channels) *channels = st->channels;
  const_cast<OpusDecoder*>(st)->min_frame_48k = -1;    // its caller feeds device-resident packets the host never sees
  return st->d_state;
}

// internal (used by opus_enc_api.cu): the per-device stream all launches of this library go on,
// so that opus_b200_synchronize() covers encoder launches too; and the launch counter.
cudaStream_t opusb_internal_stream(int* err) {
  DevCtx* c = ctx_enter(-1, err);
  return c ? c->stream : nullptr;
}
void opusb_internal_count_launches(int n) { g_launches += n; }
// internal (cb_noalloc.cu): a call that queued its kernels on the device's main stream alone is over; later
// run-ahead calls order themselves behind it
void opusb_internal_serial_call_done(int device) {
  if (device < 0 || device >= OPUSB_MAX_DEVICES) return;
  DevCtx* c = &g_ctx[device];
  std::lock_guard<std::mutex> lk(c->mu);
  ctx_make_streams(c);
  if (c->streams_ok) cudaEventRecord(c->ev_serial, c->stream);
}
int cbk_sm_count(void) {
  int dev = -1;
  if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < OPUSB_MAX_DEVICES && g_ctx[dev].sm_count > 0) return g_ctx[dev].sm_count;
  int sm = 0;
  if (dev >= 0 && cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sm > 0) return sm;
  return 148;
}

// Timing marks for callers that queue work on the library's own stream (cuda_stream = NULL) and so have no stream
// of their own to record events on: mark `idx` (0..15) is a timing event recorded on the current device's main
// stream, behind everything queued so far; opus_b200_timer_elapsed waits for both marks and returns the milliseconds
// between them.
static cudaEvent_t g_marks[OPUSB_MAX_DEVICES][16];
int opus_b200_timer_mark(int idx) {
  if (idx < 0 || idx >= 16) return OPUS_BAD_ARG;
  int err = 0;
  DevCtx* c = ctx_enter(-1, &err);
  if (!c) return err;
  std::lock_guard<std::mutex> lk(c->mu);
  cudaEvent_t& e = g_marks[c->device][idx];
  if (!e) CUDA_OK(cudaEventCreate(&e));
  CUDA_OK(cudaEventRecord(e, c->stream));
  return OPUS_OK;
}
int opus_b200_timer_elapsed(int first, int second, double* ms) {
  if (first < 0 || first >= 16 || second < 0 || second >= 16 || !ms) return OPUS_BAD_ARG;
  int err = 0;
  DevCtx* c = ctx_enter(-1, &err);
  if (!c) return err;
  cudaEvent_t a = g_marks[c->device][first], b = g_marks[c->device][second];
  if (!a || !b) return OPUS_BAD_ARG;
  CUDA_OK(cudaEventSynchronize(b));
  float t = 0;
  CUDA_OK(cudaEventElapsedTime(&t, a, b));
  *ms = t;
  return OPUS_OK;
}

// Diagnostic (tools/pipe_trace.py): between trace_begin and trace_end every launch is bracketed by an event pair as
// with the profile calls, but calls keep their normal schedule; trace_end returns up to `cap` records of
// (kind, start ms, end ms) relative to the first recorded event.
int opus_b200_trace_begin(void) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_trace_on = true;
  return OPUS_OK;
}
int opus_b200_trace_end(int* kinds, double* t0, double* t1, int cap) {
  g_trace_on = false;
  cudaDeviceSynchronize();
  std::lock_guard<std::mutex> lk(g_prof_mu);
  int n = 0;
  cudaEvent_t ref = g_prof.empty() ? nullptr : g_prof[0].a;
  for (auto& r : g_prof) {
    float a = 0, b = 0;
    if (n < cap && ref && cudaEventElapsedTime(&a, ref, r.a) == cudaSuccess && cudaEventElapsedTime(&b, ref, r.b) == cudaSuccess) {
      kinds[n] = r.kind; t0[n] = a; t1[n] = b; n++;
    }
  }
  for (auto& r : g_prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  g_prof.clear();
  return n;
}

int opus_b200_profile_begin(void) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = true;
  return OPUS_OK;
}

int opus_b200_profile_end(double* ms, long long* counts) {
  g_prof_on = false;
  for (int k = 0; k < kProfKinds; k++) { if (ms) ms[k] = 0; if (counts) counts[k] = 0; }
  cudaDeviceSynchronize();
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (auto& r : g_prof) {
    float t = 0;
    if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess && r.kind >= 0 && r.kind < kProfKinds) {
      if (ms) ms[r.kind] += t;
      if (counts) counts[r.kind] += 1;
    }
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
  }
  g_prof.clear();
  return OPUS_OK;
}

}  // extern "C"
