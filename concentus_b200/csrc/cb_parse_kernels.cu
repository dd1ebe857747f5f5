// cb_parse_kernels.cu -- the PARSE half of the CELT entropy stage, ONE THREAD PER STREAM.
//
//   celt_parse_kernel : everything of a CELT frame that reads the bitstream -- silence / post-filter / transient
//   flags, coarse and fine energies, tf, dynalloc and trim, the bit allocation, the walk over bands / stereo splits /
//   partitions (range decoder only, no spectrum), the anti-collapse bit, the energy finalisation -- and the
//   small-state update.  It writes the frame's tape (cb_state.h: FrameTape) and the frame record for the synthesis
//   kernel; the rebuild kernel (cb_kernels.cu, one warp per stream) then turns the tape into the spectrum.
//
// Why a thread per stream.  This half has no width: the warp-per-stream decoder executed it as uniform code, every
// lane computing the same scalars -- about two thirds of all warp instructions of the old entropy kernel.  Here a
// warp carries up to 32 streams; lanes diverge where their streams do (the SILK kernel runs the same way).
// This translation unit builds the shared decoder sources with CB_THREAD_MODE (cb_common.h): lane loops are plain
// loops, warp folds are identities, tables are indexed per thread.
//
// The per-thread work area (EntWork without the spectrum, ~0.9 KB, plus the staged packet) lives in shared memory
// at a stride of an odd number of 8-byte words, so the lanes of a warp spread over the banks when they touch the
// same field (two lanes per bank at worst).
#define CB_THREAD_MODE 1
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include "cb_entropy_pass.h"
#include "cb_kernels.h"

extern "C" int cbk_sm_count(void);   // SMs of the current device (opus_api.cu)

#ifndef P_SPW
#define P_SPW 2             // streams per warp (active lanes): fewer streams to a warp = less divergence, more warps
#endif
#define P_WARPS 1           // warps per CTA: small CTAs so that a 4096-stream batch spreads over all SMs

__global__ void __launch_bounds__(P_WARPS * 32)
celt_parse_kernel(OpusDecState* const* states, const u8* data, StreamPlan* plans, FrameSide* sides, FrameTape* tapes,
                  i32* ret, int n, int pass, int per_thread_bytes, int spw) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane >= spw) return;
  const int s = (blockIdx.x * P_WARPS + warp) * spw + lane;
  if (s >= n) return;
  unsigned char* mine = smem + (size_t)(warp * spw + lane) * per_thread_bytes;
  EntWork& w = *reinterpret_cast<EntWork*>(mine);
  const int pkt_cap = per_thread_bytes - (int)sizeof(EntWork);
  u8* pkt = pkt_cap > 0 ? mine + sizeof(EntWork) : nullptr;
  OpusDecState* st = states[s];
  StreamPlan* pl = &plans[s];
  const int route = pl->route;
  if (route == CB_ROUTE_HYBRID)
    decode_entropy_parse<true>(w, pkt, pkt_cap, st, pl, &sides[s], &tapes[s], data, &ret[s], pass);
  else
    decode_entropy_parse<false>(w, pkt, pkt_cap, st, pl, &sides[s], &tapes[s], data, &ret[s], pass);
}

// shared memory of one thread: the work area plus a packet staging area, rounded so that the stride is an odd
// number of 8-byte words
static int parse_bytes_per_thread(int max_payload, int spw) {
  const int budget = (200 * 1024) / (P_WARPS * spw) - 8;          // what one CTA may take
  int b = (int)sizeof(EntWork) + ((max_payload + 15) & ~15);
  if (b > budget) b = (int)sizeof(EntWork);                        // large packets are read in place
  b = (b + 7) & ~7;
  if (((b >> 3) & 1) == 0) b += 8;
  return b;
}

extern "C" int cbk_launch_parse(OpusDecState* const* d_states, const u8* d_data, StreamPlan* d_plans, FrameSide* d_sides,
                                FrameTape* d_tapes, i32* d_ret, int n, int pass, int max_payload, int shared_sms,
                                cudaStream_t stream) {
  if (n <= 0) return 0;
  static int attr_done[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev >= 64 || !attr_done[dev]) {
    e = cudaFuncSetAttribute(celt_parse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return (int)e;
    // shares SMs with the rebuild / synthesis kernels of the call before (see cb_kernels.cu: ensure_attrs)
    e = cudaFuncSetAttribute(celt_parse_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return (int)e;
    if (dev < 64) attr_done[dev] = 1;
  }
  static int spw_env = -1;
  if (spw_env < 0) { const char* e2 = getenv("OPUSB_PARSE_SPW"); spw_env = e2 ? atoi(e2) : 0; }
  // few streams to a warp while that still fits the GPU in one wave (about 16 of these warps per SM: registers),
  // more for larger batches.  A lone parse kernel is a latency chain and is fastest with 2 streams to a warp; when
  // it shares the SMs with the rebuild / synthesis kernels of the call before (`shared_sms`), issue slots are what
  // runs out, and 4 streams to a warp -- two thirds of the warp instructions -- measured faster (124 k -> 127 k).
  int spw = shared_sms ? 2 * P_SPW : P_SPW;
  while (spw < 32 && (n + spw - 1) / spw > cbk_sm_count() * 16) spw <<= 1;
  if (spw_env >= 1 && spw_env <= 32) spw = spw_env;
  const int per_thread = parse_bytes_per_thread(max_payload, spw);
  const int per_cta = P_WARPS * spw;
  celt_parse_kernel<<<(n + per_cta - 1) / per_cta, P_WARPS * 32, (size_t)per_thread * per_cta, stream>>>(
      d_states, d_data, d_plans, d_sides, d_tapes, d_ret, n, pass, per_thread, spw);
  return (int)cudaGetLastError();
}
