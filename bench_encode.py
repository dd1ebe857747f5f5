#!/usr/bin/env python3
"""bench_encode.py -- measurement of the batched CELT ENCODE path (BASELINE.json config 3).

Not the driver's headline line (that is bench.py, config 2); same conventions and JSON shape so the
numbers can be read side by side.  Metric: aggregate realtime stream-seconds per second, 48 kHz stereo,
128 kbps VBR, complexity 10, 20 ms frames, 4096 independent streams per GPU, RESTRICTED_LOWDELAY
(CELT-only).  A "step" = one batch call = one 20 ms frame of every stream.

  value : PCM already in HBM, packets left in HBM (opus_encode_group_device), kernel timed by the
          library's CUDA events on its launching stream, L2 flushed between steps.
  e2e   : opus_encode_group_flat with host PCM in and host packets out, copies inside the timed region.
  cpu_baseline : the reference encoder (oracle/_ref/ref_enc_bench) on all host cores, same PCM.

Data: synthetic -- 64 seeded signals (tones, noise, clicks, chirps, gated mixes; tests/enc_cases.py)
of 1 s each; stream i plays signal i mod 64 starting at frame (37*i mod 50).
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
import enc_cases  # noqa: E402
from bench import ClockSampler, peaks  # noqa: E402

FS, CH, SPF, BITRATE, CX = 48000, 2, 960, 128000, 10
MAXB = 1275
METRIC = "aggregate realtime stream-sec/sec (48k stereo CELT encode)"
UNIT = "stream-seconds/s"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--streams", type=int, default=4096)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    import torch
    import concentus_b200 as cb
    lib = cb.load()
    n, K, W = args.streams, args.steps, max(3, args.warmup)
    nsig, nfr = 64, 50
    sigs = np.stack([enc_cases.signal(enc_cases.KINDS[s % len(enc_cases.KINDS)], SPF * nfr, CH, 500 + s)
                     for s in range(nsig)]).reshape(nsig, nfr, SPF * CH)
    rot = (37 * np.arange(n)) % nfr
    sid = np.arange(n) % nsig

    def rows_for(t):
        return np.ascontiguousarray(sigs[sid, (rot + t) % nfr])

    be = cb.OpusBatchEncoder(n, FS, CH, enc_cases.LOWDELAY)
    be.ctl(enc_cases.SET_BITRATE, BITRATE)
    be.ctl(enc_cases.SET_COMPLEXITY, CX)
    d_rows = [torch.from_numpy(rows_for(t)).cuda() for t in range(nfr)]
    d_data = torch.zeros((n, MAXB), dtype=torch.uint8, device="cuda")
    d_len = torch.zeros(n, dtype=torch.int32, device="cuda")
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    t = 0
    for _ in range(W):
        be.encode_device(d_rows[t % nfr].data_ptr(), SPF, d_data.data_ptr(), MAXB, d_len.data_ptr())
        t += 1
    lib.opus_b200_synchronize()
    assert int((d_len <= 0).sum().item()) == 0
    sampler = ClockSampler(0)
    sampler.start()
    launches0 = lib.opus_b200_launch_count()
    lib.opus_b200_encode_profile_begin()
    w0 = time.perf_counter()
    for k in range(K):
        flush.fill_(k & 0xFF)
        torch.cuda.synchronize()
        be.encode_device(d_rows[t % nfr].data_ptr(), SPF, d_data.data_ptr(), MAXB, d_len.data_ptr())
        t += 1
    lib.opus_b200_synchronize()
    w1 = time.perf_counter()
    ms = ctypes.c_double(0)
    cnt = ctypes.c_longlong(0)
    lib.opus_b200_encode_profile_end(ctypes.byref(ms), ctypes.byref(cnt))
    launches = lib.opus_b200_launch_count() - launches0
    dev_ms = ms.value
    lens = d_len.cpu().numpy()
    assert (lens > 0).all()
    # e2e: page-locked host PCM in, page-locked host packets out, every step; (a) one synchronous call per step
    # (opus_encode_group_flat), (b) the pipelined form (opus_encode_group_submit/_wait, two calls in flight)
    import torch as _t
    h_rows = [_t.from_numpy(rows_for(k)).pin_memory() for k in range(8)]
    h_data = [_t.zeros((n, MAXB), dtype=_t.uint8).pin_memory() for _ in range(3)]
    h_len = [_t.zeros(n, dtype=_t.int32).pin_memory() for _ in range(3)]
    lib = cb.load()
    for k in range(3):
        assert lib.opus_encode_group_flat(be._g, h_rows[k % 8].data_ptr(), SPF, h_data[0].data_ptr(), MAXB, h_len[0].data_ptr()) == 0
    s0 = time.perf_counter()
    for k in range(K):
        rc = lib.opus_encode_group_flat(be._g, h_rows[k % 8].data_ptr(), SPF, h_data[0].data_ptr(), MAXB, h_len[0].data_ptr())
        assert rc == 0 and bool((h_len[0] > 0).all())
    sync_s = time.perf_counter() - s0
    for k in range(3):                                   # warm-up of the pipelined entry (it allocates its two slots)
        assert lib.opus_encode_group_submit(be._g, h_rows[k % 8].data_ptr(), SPF, h_data[k % 3].data_ptr(), MAXB,
                                            h_len[k % 3].data_ptr()) == 0
    assert lib.opus_encode_group_wait(be._g) == 0
    e0 = time.perf_counter()
    for k in range(K):
        rc = lib.opus_encode_group_submit(be._g, h_rows[k % 8].data_ptr(), SPF, h_data[k % 3].data_ptr(), MAXB,
                                          h_len[k % 3].data_ptr())
        assert rc == 0
        if k >= 2:
            assert bool((h_len[(k - 2) % 3] > 0).all())
    assert lib.opus_encode_group_wait(be._g) == 0
    e1 = time.perf_counter()
    assert all(bool((l > 0).all()) for l in h_len)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    units = n * SPF / FS
    value = units * K / (dev_ms * 1e-3)
    e2e_value = units * K / (e1 - e0)
    peak, peak_kind = peaks()
    alg_bytes = n * (SPF * CH * 2 + int(lens.mean()))
    achieved = alg_bytes / (dev_ms / max(1, cnt.value) * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": K, "warmup": W,
        "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int32/int16 fixed-point", "data": "synthetic: 64 seeded signals x 1 s (tests/enc_cases.py), rotated per stream",
        "config": {"workload": "batched CELT encode: 4096 x 48 kHz stereo 128 kbps VBR cx10 20 ms streams (config 3)",
                   "streams_per_gpu": n, "frame_ms": 20, "mean_packet_bytes": float(lens.mean()),
                   "l2": "flushed between timed steps (256 MiB write, outside the event pair)"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(n * SPF * CH * 2),
                "d2h_bytes_per_step": int(n * MAXB + 4 * n), "ms_per_step": 1e3 * (e1 - e0) / K,
                "api": "opus_encode_group_submit/_wait (pinned host buffers, two calls in flight)",
                "synchronous_call_value": units * K / sync_s, "synchronous_call_api": "opus_encode_group_flat"},
        "gpu_launches": int(launches), "clocks": sampler.summary(),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "peak_kind": peak_kind, "kernel": "opus_encode_kernel",
                     "kernel_ms": dev_ms / max(1, cnt.value), "algorithmic_bytes_per_launch": alg_bytes,
                     "note": "integer/latency-bound path: HBM fraction is small by construction"},
        "wall_ms_per_step": 1e3 * (w1 - w0) / K,
    }
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_enc_bench")
    if not args.no_cpu_baseline and os.path.exists(exe):
        cores = os.cpu_count() or 1
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        raw = os.path.join(ROOT, "gpurun_out", "enc_bench_pcm.raw")
        sigs[:8].reshape(-1).tofile(raw)
        streams_s = 8 * cores
        out = subprocess.run([exe, raw, str(FS), str(CH), str(BITRATE), str(CX), str(SPF), str(streams_s), "100", str(cores)],
                             stdout=subprocess.PIPE, text=True, check=True).stdout.split()
        os.remove(raw)
        line["cpu_baseline"] = {"value": float(out[0]) / float(out[1]), "unit": UNIT, "cores": cores, "kind": "reference",
                                "sample": "%d of 4096 streams x 100 frames (2 s each), one pass" % streams_s}
    print(json.dumps(line))
    be.close()


if __name__ == "__main__":
    main()
