#!/usr/bin/env python3
"""bench_ms.py -- measurement of the batched 5.1 multistream decode (BASELINE.json config 5): 1024
six-channel 48 kHz decoders (4 elementary streams each, 2 coupled), 20 ms frames, packets from the
reference surround encoder at 256 kbps with 4 % of the packets lost (pitch-based concealment) and the
transient frames the material has.  Same JSON shape as bench.py (which stays the driver's headline line).
A "step" = one batch call = one 20 ms frame of every decoder.

  value / e2e : opus_multistream_decode_group_submit/_wait with page-locked host packets in and page-locked host
                PCM out -- the multistream entry has no device-resident form, so both figures are end to end
                (the synchronous opus_multistream_decode_group_flat is timed beside it);
                kernel_ms gives the device share (library events around the elementary entropy / synthesis
                launches of the 4096 elementary streams).
  cpu_baseline : the reference opus_multistream_decode (oracle/_ref/ref_bench) on all host cores.

Data: decoder i plays the committed fixture tests/golden/ms51_256k.bit (200 packets, 8 of them lost) from
packet (37*i mod 200).
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

FS, CH, SPF = 48000, 6, 960
FIXTURE = os.path.join(ROOT, "tests", "golden", "ms51_256k.bit")
METRIC = "aggregate realtime stream-sec/sec (48k 5.1 multistream decode)"
UNIT = "stream-seconds/s"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--streams", type=int, default=1024)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    import torch
    import concentus_b200 as cb
    lib = cb.load()
    meta = json.load(open(os.path.join(ROOT, "tests", "golden", "index_ms.json")))["ms51_256k"]
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
    base_off = (np.arange(n, dtype=np.int32) * maxlen).astype(np.int32)

    def step_input(t):
        idx = (rot + t) % npk
        lens = plen[idx].copy()
        offs = np.where(lens > 0, base_off, -1).astype(np.int32)
        return (torch.from_numpy(np.ascontiguousarray(table[idx]).reshape(-1)).pin_memory(),
                torch.from_numpy(offs).pin_memory(), torch.from_numpy(lens).pin_memory(), int((lens == 0).sum()))

    inputs = [step_input(t) for t in range(W + K + 4)]
    bd = cb.OpusMSBatchDecoder(FS, CH, meta["streams"], meta["coupled"], meta["mapping"], n)
    h_pcms = [torch.zeros(n * SPF * CH, dtype=torch.int16).pin_memory() for _ in range(3)]
    h_rets = [torch.zeros(n, dtype=torch.int32).pin_memory() for _ in range(3)]
    h_pcm, h_ret = h_pcms[0], h_rets[0]

    def call(t):
        a, o, l, _ = inputs[t]
        rc = lib.opus_multistream_decode_group_flat(bd._g, a.data_ptr(), a.numel(), o.data_ptr(), l.data_ptr(),
                                                    h_pcm.data_ptr(), SPF * CH, SPF, None, h_ret.data_ptr())
        assert rc == 0 and bool((h_ret == SPF).all())

    def submit(t, slot):
        a, o, l, _ = inputs[t % len(inputs)]
        assert lib.opus_multistream_decode_group_submit(bd._g, a.data_ptr(), a.numel(), o.data_ptr(), l.data_ptr(),
                                                        h_pcms[slot].data_ptr(), SPF * CH, SPF, None,
                                                        h_rets[slot].data_ptr()) == 0

    t = 0
    for _ in range(W):
        call(t)
        t += 1
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    launches0 = lib.opus_b200_launch_count()
    lib.opus_b200_profile_begin()
    e0 = time.perf_counter()
    lost = 0
    for _ in range(K):
        call(t)
        lost += inputs[t][3]
        t += 1
    torch.cuda.synchronize()
    e1 = time.perf_counter()
    ms3 = (ctypes.c_double * 3)()
    cnt3 = (ctypes.c_longlong * 3)()
    lib.opus_b200_profile_end(ms3, cnt3)
    launches = lib.opus_b200_launch_count() - launches0
    sampler.stop_flag = True
    sampler.join(timeout=2)
    kern_ms = {"celt_entropy": ms3[1] / K, "celt_synth": ms3[2] / K}
    units = n * SPF / FS
    sync_value = units * K / (e1 - e0)
    # the pipelined form of the same call (two calls in flight: copies overlap the neighbours' kernels)
    for k in range(3):
        submit(t + k, k % 3)
    assert lib.opus_multistream_decode_group_wait(bd._g) == 0
    t += 3
    p0 = time.perf_counter()
    for k in range(K):
        submit(t + k, k % 3)
        if k >= 2:
            assert bool((h_rets[(k - 2) % 3] == SPF).all())
    assert lib.opus_multistream_decode_group_wait(bd._g) == 0
    p1 = time.perf_counter()
    assert all(bool((r == SPF).all()) for r in h_rets)
    value = units * K / (p1 - p0)
    peak, peak_kind = peaks()
    mean_len = float(plen[plen > 0].mean())
    alg_bytes = n * (int(mean_len) + SPF * CH * 2)
    dom = max(kern_ms, key=kern_ms.get)
    achieved = alg_bytes / (kern_ms[dom] * 1e-3) / 1e9 if kern_ms[dom] > 0 else 0.0
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": K, "warmup": W,
        "ms_per_step": 1e3 * (p1 - p0) / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int32/int16 fixed-point",
        "data": "synthetic: rotations of one reference-encoded 5.1 stream (tests/golden/ms51_256k.bit)",
        "config": {"workload": "batched 5.1 multistream decode: 1024 x 48 kHz six-channel 256 kbps 20 ms (config 5)",
                   "streams_per_gpu": n, "elementary_streams": n * meta["streams"], "frame_ms": 20,
                   "mean_packet_bytes": mean_len, "lost_packet_fraction": lost / float(n * K),
                   "l2": "inputs and outputs larger than L2 across steps (12 MB PCM per step, host round trip each step)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": int(n * maxlen + 12 * n),
                "d2h_bytes_per_step": int(n * SPF * CH * 2 + 4 * n), "ms_per_step": 1e3 * (p1 - p0) / K,
                "api": "opus_multistream_decode_group_submit/_wait (pinned host buffers, two calls in flight)",
                "synchronous_call_value": sync_value, "synchronous_call_api": "opus_multistream_decode_group_flat"},
        "gpu_launches": int(launches), "clocks": sampler.summary(),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "peak_kind": peak_kind, "kernel": dom + "_kernel", "kernel_ms": kern_ms,
                     "algorithmic_bytes_per_launch": alg_bytes,
                     "note": "integer/latency-bound path: HBM fraction is small by construction"},
    }
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_bench")
    if not args.no_cpu_baseline and os.path.exists(exe):
        cores = os.cpu_count() or 1
        streams_s = 8 * cores
        out = subprocess.run([exe, FIXTURE, str(FS), str(CH), str(streams_s), str(npk), str(cores), str(meta["streams"]),
                              str(meta["coupled"]), ",".join(str(m) for m in meta["mapping"])],
                             stdout=subprocess.PIPE, text=True, check=True).stdout.split()
        line["cpu_baseline"] = {"value": float(out[0]) / float(out[1]), "unit": UNIT, "cores": cores, "kind": "reference",
                                "sample": "%d of 1024 decoders x %d frames, one pass" % (streams_s, npk)}
    print(json.dumps(line))
    bd.close()


if __name__ == "__main__":
    main()
