#!/usr/bin/env python3
"""bench_silk.py -- measurement of the batched SILK decode path (BASELINE.json config 4): 16384 independent
16 kHz mono voice streams at 32 kbps, 20 ms frames.  Same conventions and JSON shape as bench.py (which
stays the driver's headline line).  A "step" = one batch call = one 20 ms frame of every stream.

  value : packets already in HBM, PCM left in HBM (opus_decode_group_device), device time from the
          library's CUDA events around the launches, L2 flushed between steps.
  e2e   : opus_decode_group_flat with host packets in and host PCM out, copies inside the timed region.
  cpu_baseline : the reference decoder (oracle/_ref/ref_bench) on all host cores, same packets.

Data: stream i plays the committed fixture tests/golden/silk_wb_mono_32k.bit (200 packets the reference
encoder produced from AudioData/16Khz Mono.raw) starting at packet (37*i mod 200).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from bench import ClockSampler, peaks, read_bit  # noqa: E402

FS, CH, SPF = 16000, 1, 320
FIXTURE = os.path.join(ROOT, "tests", "golden", "silk_wb_mono_32k.bit")
METRIC = "aggregate realtime stream-sec/sec (16k mono SILK decode)"
UNIT = "stream-seconds/s"
WORKLOAD = "batched SILK decode: 16384 x 16 kHz mono 32 kbps 20 ms streams (config 4)"
KERNEL = "silk_decode_kernel"
NOTE = "step = plan + SILK kernel + (empty) CELT/general kernels; serial per-sample recurrences, latency bound"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--streams", type=int, default=16384)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--hybrid", action="store_true",
                    help="config 4's hybrid set: 48 kHz mono 32 kbps VOIP streams (every packet hybrid fullband 20 ms)")
    args = ap.parse_args()
    global FS, SPF, FIXTURE, METRIC, WORKLOAD, KERNEL, NOTE
    if args.hybrid:
        FS, SPF = 48000, 960
        FIXTURE = os.path.join(ROOT, "tests", "golden", "hyb_fb_mono_32k.bit")
        METRIC = "aggregate realtime stream-sec/sec (48k mono hybrid decode)"
        WORKLOAD = "batched hybrid decode: 16384 x 48 kHz mono 32 kbps 20 ms streams (config 4, hybrid set)"
        KERNEL = "silk_decode_kernel + celt_entropy_kernel + celt_synth_kernel"
        NOTE = "step = plan + SILK kernel (SILK layer, thread per stream) + CELT kernels (bands 17.., warp per stream)"
    import torch
    import concentus_b200 as cb
    lib = cb.load()
    n, K, W = args.streams, args.steps, max(3, args.warmup)
    pk = [p for p, _ in read_bit(FIXTURE)]
    npk = len(pk)
    maxlen = max(len(p) for p in pk)
    table = np.zeros((npk, maxlen), dtype=np.uint8)
    plen = np.zeros(npk, dtype=np.int32)
    for i, p in enumerate(pk):
        table[i, :len(p)] = np.frombuffer(p, dtype=np.uint8)
        plen[i] = len(p)
    rot = (37 * np.arange(n)) % npk
    offs = (np.arange(n, dtype=np.int32) * maxlen).astype(np.int32)

    def arena_for(t):
        idx = (rot + t) % npk
        return np.ascontiguousarray(table[idx]).reshape(-1), plen[idx].copy()

    bd = cb.OpusBatchDecoder(FS, CH, n)
    stream = torch.cuda.Stream()
    d_off = torch.from_numpy(offs).cuda()
    d_ret = torch.zeros(n, dtype=torch.int32, device="cuda")
    d_pcm = torch.zeros(n * SPF * CH, dtype=torch.int16, device="cuda")
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    steps_data = [arena_for(t) for t in range(W + 2 * K + 8)]
    d_arena = [(torch.from_numpy(a).cuda(), torch.from_numpy(l).cuda()) for a, l in steps_data]
    torch.cuda.synchronize()
    t = 0

    def step_device(t):
        a, l = d_arena[t]
        rc = lib.opus_decode_group_device(bd._g, a.data_ptr(), d_off.data_ptr(), l.data_ptr(), d_pcm.data_ptr(),
                                          SPF * CH, SPF, None, d_ret.data_ptr(), 1, maxlen,
                                          ctypes.c_void_p(stream.cuda_stream))
        assert rc == 0

    for _ in range(W):
        step_device(t)
        t += 1
    torch.cuda.synchronize()
    assert int((d_ret != SPF).sum().item()) == 0
    sampler = ClockSampler(0)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    launches0 = lib.opus_b200_launch_count()
    with torch.cuda.stream(stream):
        for k in range(K):
            flush.fill_(k & 0xFF)
            ev[k][0].record(stream)
            step_device(t)
            ev[k][1].record(stream)
            t += 1
    torch.cuda.synchronize()
    launches = lib.opus_b200_launch_count() - launches0
    dev_ms = float(sum(a.elapsed_time(b) for a, b in ev))
    assert int((d_ret != SPF).sum().item()) == 0
    # per-kernel timing (library events around every launch), separate loop
    lib.opus_b200_profile_begin()
    with torch.cuda.stream(stream):
        for k in range(min(K, 20)):
            flush.fill_(k & 0xFF)
            step_device(t % len(d_arena))
            t += 1
    ms3 = (ctypes.c_double * 3)()
    cnt3 = (ctypes.c_longlong * 3)()
    lib.opus_b200_profile_end(ms3, cnt3)
    kern_ms = {"plan+silk": ms3[0] / max(1, cnt3[0]), "celt_entropy": ms3[1] / max(1, cnt3[1]),
               "celt_synth": ms3[2] / max(1, cnt3[2])}
    # e2e: page-locked host packets in, page-locked host PCM out, every step; (a) one synchronous call per step
    # (opus_decode_group_flat), (b) the pipelined form (opus_decode_group_submit/_wait, two calls in flight)
    h_pcm = [torch.zeros(n * SPF * CH, dtype=torch.int16).pin_memory() for _ in range(3)]
    h_ret = [torch.zeros(n, dtype=torch.int32).pin_memory() for _ in range(3)]
    h_off = torch.from_numpy(offs).pin_memory()
    h_in = [(torch.from_numpy(a).pin_memory(), torch.from_numpy(l).pin_memory()) for a, l in steps_data]

    def call(fn, t, slot):
        a, l = h_in[t % len(h_in)]
        return fn(bd._g, a.data_ptr(), a.numel(), h_off.data_ptr(), l.data_ptr(), h_pcm[slot].data_ptr(), SPF * CH, SPF,
                  None, h_ret[slot].data_ptr())

    for k in range(3):
        assert call(lib.opus_decode_group_flat, t, 0) == 0
        t += 1
    torch.cuda.synchronize()
    s0 = time.perf_counter()
    for k in range(K):
        assert call(lib.opus_decode_group_flat, t, 0) == 0 and bool((h_ret[0] == SPF).all())
        t += 1
    sync_s = time.perf_counter() - s0
    for k in range(3):                                   # warm-up of the pipelined entry (it allocates its two slots)
        assert call(lib.opus_decode_group_submit, t, k % 3) == 0
        t += 1
    assert lib.opus_decode_group_wait(bd._g) == 0
    e0 = time.perf_counter()
    for k in range(K):
        assert call(lib.opus_decode_group_submit, t, k % 3) == 0
        if k >= 2:
            assert bool((h_ret[(k - 2) % 3] == SPF).all())
        t += 1
    assert lib.opus_decode_group_wait(bd._g) == 0
    e1 = time.perf_counter()
    assert all(bool((r == SPF).all()) for r in h_ret)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    units = n * SPF / FS
    value = units * K / (dev_ms * 1e-3)
    e2e_value = units * K / (e1 - e0)
    peak, peak_kind = peaks()
    alg_bytes = n * (int(plen.mean()) + SPF * CH * 2)
    achieved = alg_bytes / (dev_ms / K * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": K, "warmup": W,
        "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int32/int16 fixed-point",
        "data": "synthetic: rotations of one reference-encoded stream (tests/golden/%s)" % os.path.basename(FIXTURE),
        "config": {"workload": WORKLOAD,
                   "streams_per_gpu": n, "frame_ms": 20, "mean_packet_bytes": float(plen.mean()),
                   "l2": "flushed between timed steps (256 MiB write, outside the event pair)"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(n * maxlen + 8 * n),
                "d2h_bytes_per_step": int(n * SPF * CH * 2 + 4 * n), "ms_per_step": 1e3 * (e1 - e0) / K,
                "api": "opus_decode_group_submit/_wait (pinned host buffers, two calls in flight)",
                "synchronous_call_value": units * K / sync_s, "synchronous_call_api": "opus_decode_group_flat"},
        "gpu_launches": int(launches), "clocks": sampler.summary(),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "peak_kind": peak_kind, "kernel": KERNEL,
                     "step_ms": dev_ms / K, "kernel_ms": kern_ms, "algorithmic_bytes_per_launch": alg_bytes,
                     "note": NOTE},
    }
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_bench")
    if not args.no_cpu_baseline and os.path.exists(exe):
        cores = os.cpu_count() or 1
        streams_s = 256 * cores
        out = subprocess.run([exe, FIXTURE, str(FS), str(CH), str(streams_s), str(npk), str(cores)],
                             stdout=subprocess.PIPE, text=True, check=True).stdout.split()
        line["cpu_baseline"] = {"value": float(out[0]) / float(out[1]), "unit": UNIT, "cores": cores, "kind": "reference",
                                "sample": "%d of 16384 streams x %d frames, one pass" % (streams_s, npk)}
    print(json.dumps(line))
    bd.close()


if __name__ == "__main__":
    main()
