#!/usr/bin/env python3
"""bench.py -- headline benchmark of the batched Opus decode path (BASELINE.json config 2).

Metric: aggregate realtime stream-seconds per second, 48 kHz stereo CELT decode, 96 kbps, 20 ms
frames, 4096 independent streams per GPU (weak scaling: every rank decodes its own 4096 streams;
there is no cross-GPU exchange, so no collective on the data path).

A "step" = one batch call = one 20 ms frame of every stream (81.92 stream-seconds per GPU).

  value : device-resident inputs (packets already in HBM), timed with CUDA events on the stream
          the kernels are launched on, L2 flushed between steps.
  e2e   : the same call through the host-buffer C-ABI entry (opus_decode_group_flat) with pinned
          host packets in and pinned host PCM out, copies inside the timed region.
  --impl reference : the reference's own CPU decoder (oracle/_ref, libopus 1.1.2 FIXED_POINT built
          from /root/reference) on all host cores, same packets.

Data: synthetic -- stream i plays the committed fixture tests/golden/c2_96k_cbr_st.bit (250 packets
the reference encoder produced from AudioData/48Khz Stereo.raw at 96 kbps CBR) starting at packet
(37*i mod 250), so the streams are at different places of real audio.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

STREAMS = 4096
FS = 48000
CH = 2
SPF = 960
FIXTURE = os.path.join(ROOT, "tests", "golden", "c2_96k_cbr_st.bit")
BYTES_PER_FRAME = 240 + SPF * CH * 2          # algorithmic bytes per stream-frame: packet + PCM
METRIC = "aggregate realtime stream-sec/sec (48k stereo CELT decode)"
UNIT = "stream-seconds/s"


def read_bit(path):
    import struct
    d = open(path, "rb").read()
    o, out = 0, []
    while o < len(d):
        ln, rng = struct.unpack(">II", d[o:o + 8])
        o += 8
        out.append((d[o:o + ln], rng))
        o += ln
    return out


def ncu_traffic():
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture
    (profiles/r01_entropy_ncu_raw.txt: dram__bytes_read.sum + dram__bytes_write.sum), or None."""
    tot, unit = 0.0, {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    try:
        for line in open(os.path.join(ROOT, "profiles", "r01_entropy_ncu_raw.txt")):
            f = line.split()
            if len(f) == 3 and f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(f[2]) * unit[f[1]]
    except (OSError, KeyError, ValueError):
        return None
    return tot or None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self.stop_flag = False

    def run(self):
        # NVML in-process (nvidia_ml_py) when available: a query costs microseconds.  Spawning nvidia-smi takes
        # hundreds of milliseconds per sample and its start-up can stall the CUDA calls of the process under test.
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
            get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            while not self.stop_flag:
                self.samples.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                r = int(get_reasons(h))
                for nm, bit in bits.items():
                    if r & bit:
                        self.reasons.add(nm)
                time.sleep(0.02)
            return
        except Exception:
            pass
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                      "--format=csv,noheader,nounits"], stdout=subprocess.PIPE,
                                     stderr=subprocess.DEVNULL, text=True, timeout=5).stdout.strip()
                f = [x.strip() for x in out.split(",")]
                self.samples.append(float(f[0]))
                self.max_mhz = float(f[1])
                for nm, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def ref_bench(streams, frames, threads):
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_bench")
    if not os.path.exists(exe):
        return None
    out = subprocess.run([exe, FIXTURE, str(FS), str(CH), str(streams), str(frames), str(threads)],
                         stdout=subprocess.PIPE, text=True, check=True).stdout.split()
    audio, wall = float(out[0]), float(out[1])
    return audio, wall


def run_reference(args, rank, world):
    """Reference arm: libopus 1.1.2 FIXED_POINT (oracle/_ref) on the host cores, all threads."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    streams = 16 * cores                       # bounded sample of the 4096-stream workload
    per_step_frames = 50                       # 1 s of audio per stream per step
    times = []
    audio_total = 0.0
    for it in range(args.warmup + args.steps):
        r = ref_bench(streams, per_step_frames, cores)
        if r is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_bench not built (needs /root/reference at build time)"}))
            return
        if it >= args.warmup:
            audio_total += r[0]
            times.append(r[1])
    wall = sum(times)
    value = audio_total / wall
    sample = "%d of 4096 streams x %d frames per step, %d steps" % (streams, per_step_frames, args.steps)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32/int16 fixed-point",
        "data": "synthetic: rotations of one reference-encoded 250-packet stream (tests/golden/c2_96k_cbr_st.bit)",
        "config": {"workload": "batched CELT decode: 4096 x 48 kHz stereo 96 kbps CBR 20 ms streams (config 2)",
                   "streams_per_gpu": STREAMS, "frame_ms": 20, "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--streams", type=int, default=STREAMS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import concentus_b200 as cb

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))
    lib = cb.load()
    n = args.streams
    pk = read_bit(FIXTURE)
    npk = len(pk)
    K, W = args.steps, args.warmup
    total_steps = K + W

    # ---- workload: per-step packet arenas (stream i, step t -> packet (37*i + t) mod 250) ----
    plen = len(pk[0][0])
    assert all(len(p) == plen for p, _ in pk), "fixture is CBR"
    table = np.frombuffer(b"".join(p for p, _ in pk), dtype=np.uint8).reshape(npk, plen)
    rot = (37 * np.arange(n)) % npk
    offs = (np.arange(n, dtype=np.int32) * plen).astype(np.int32)
    lens = np.full(n, plen, dtype=np.int32)
    NBUF = 8                                            # distinct arenas cycled through the run

    def arena_for(t):
        return np.ascontiguousarray(table[(rot + t) % npk]).reshape(-1)

    bd = cb.OpusBatchDecoder(FS, CH, n)                 # n decoders + one group (state in HBM)
    stream = torch.cuda.Stream()
    d_off = torch.from_numpy(offs).cuda()
    d_len = torch.from_numpy(lens).cuda()
    d_ret = torch.zeros(n, dtype=torch.int32, device="cuda")
    d_pcm = [torch.zeros(n * SPF * CH, dtype=torch.int16, device="cuda") for _ in range(2)]
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    h_arena = [torch.from_numpy(arena_for(t)).pin_memory() for t in range(total_steps)] if total_steps <= 400 else None
    if h_arena is None:
        h_arena = [torch.from_numpy(arena_for(t)).pin_memory() for t in range(npk)]
    d_arena = [a.cuda() for a in h_arena[:max(NBUF, 1)]]
    # all steps' packets resident in HBM before the timed region starts
    d_arena_all = [a.cuda() for a in h_arena]
    torch.cuda.synchronize()

    def step_device(t):
        a = d_arena_all[t % len(d_arena_all)]
        rc = lib.opus_decode_group_device(bd._g, a.data_ptr(), d_off.data_ptr(), d_len.data_ptr(),
                                          d_pcm[t & 1].data_ptr(), SPF * CH, SPF, None, d_ret.data_ptr(), 1, plen,
                                          ctypes.c_void_p(stream.cuda_stream))
        if rc != 0:
            raise RuntimeError("opus_decode_group_device failed: %d" % rc)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up ----
    t_global = 0
    for _ in range(W):
        step_device(t_global)
        t_global += 1
    barrier()
    assert int((d_ret != SPF).sum().item()) == 0, "decode errors in warm-up"

    sampler = ClockSampler(local_rank)
    sampler.start()
    # ---- timed: K steps, device-resident inputs, CUDA events on the launching stream ----
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    launches0 = lib.opus_b200_launch_count()
    barrier()
    wall0 = time.perf_counter()
    with torch.cuda.stream(stream):
        for k in range(K):
            flush.fill_(k & 0xFF)                       # L2 flush between timed iterations (not timed)
            ev[k][0].record(stream)
            step_device(t_global)
            ev[k][1].record(stream)
            t_global += 1
    barrier()
    wall1 = time.perf_counter()
    launches = lib.opus_b200_launch_count() - launches0
    step_ms = [a.elapsed_time(b) for a, b in ev]
    dev_ms = float(sum(step_ms))
    assert int((d_ret != SPF).sum().item()) == 0, "decode errors in the timed region"

    # ---- per-kernel timing for the roofline (events around every launch, separate loop) ----
    lib.opus_b200_profile_begin()
    with torch.cuda.stream(stream):
        for k in range(K):
            flush.fill_(k & 0xFF)
            step_device(t_global)
            t_global += 1
    ms3 = (ctypes.c_double * 3)()
    cnt3 = (ctypes.c_longlong * 3)()
    lib.opus_b200_profile_end(ms3, cnt3)
    kern_ms = [ms3[i] / max(1, cnt3[i]) for i in range(3)]

    # ---- e2e: host pinned packets -> C-ABI -> host pinned PCM, wall clock around the calls ----
    # (a) synchronous call per step (opus_decode_group_flat); (b) the pipelined form of the same call
    # (opus_decode_group_submit/_wait, two calls in flight so copies overlap the neighbours' kernels): the
    # e2e figure.  Both copy every step's packets in and every step's PCM + return values out.
    h_pcm = [torch.zeros(n * SPF * CH, dtype=torch.int16).pin_memory() for _ in range(3)]
    h_ret = [torch.zeros(n, dtype=torch.int32).pin_memory() for _ in range(3)]
    h_off = torch.from_numpy(offs).pin_memory()
    h_len = torch.from_numpy(lens).pin_memory()
    e2e_steps = K
    for _ in range(3):
        a = h_arena[t_global % len(h_arena)]
        lib.opus_decode_group_flat(bd._g, a.data_ptr(), a.numel(), h_off.data_ptr(), h_len.data_ptr(),
                                   h_pcm[0].data_ptr(), SPF * CH, SPF, None, h_ret[0].data_ptr())
        t_global += 1
    barrier()
    s0 = time.perf_counter()
    for _ in range(e2e_steps):
        a = h_arena[t_global % len(h_arena)]
        rc = lib.opus_decode_group_flat(bd._g, a.data_ptr(), a.numel(), h_off.data_ptr(), h_len.data_ptr(),
                                        h_pcm[0].data_ptr(), SPF * CH, SPF, None, h_ret[0].data_ptr())
        assert rc == 0 and bool((h_ret[0] == SPF).all())
        t_global += 1
    torch.cuda.synchronize()
    e2e_sync_s = time.perf_counter() - s0
    for k in range(3):                                   # warm-up of the pipelined entry (it allocates its two slots)
        a = h_arena[t_global % len(h_arena)]
        assert lib.opus_decode_group_submit(bd._g, a.data_ptr(), a.numel(), h_off.data_ptr(), h_len.data_ptr(),
                                            h_pcm[k % 3].data_ptr(), SPF * CH, SPF, None, h_ret[k % 3].data_ptr()) == 0
        t_global += 1
    assert lib.opus_decode_group_wait(bd._g) == 0
    barrier()
    e0 = time.perf_counter()
    for k in range(e2e_steps):
        a = h_arena[t_global % len(h_arena)]
        rc = lib.opus_decode_group_submit(bd._g, a.data_ptr(), a.numel(), h_off.data_ptr(), h_len.data_ptr(),
                                          h_pcm[k % 3].data_ptr(), SPF * CH, SPF, None, h_ret[k % 3].data_ptr())
        assert rc == 0
        if k >= 2:
            assert bool((h_ret[(k - 2) % 3] == SPF).all())      # call k-2 completed before submit k returned
        t_global += 1
    assert lib.opus_decode_group_wait(bd._g) == 0
    e1 = time.perf_counter()
    assert all(bool((r == SPF).all()) for r in h_ret)
    e2e_s = e1 - e0
    sampler.stop_flag = True
    sampler.join(timeout=2)

    # ---- reduce over ranks: max time ----
    if world > 1:
        tt = torch.tensor([dev_ms, e2e_s, e2e_sync_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dev_ms, e2e_s, e2e_sync_s = float(tt[0].item()), float(tt[1].item()), float(tt[2].item())

    units_per_step = n * SPF / FS                        # stream-seconds per step per GPU
    value = world * units_per_step * K / (dev_ms * 1e-3)
    e2e_value = world * units_per_step * e2e_steps / e2e_s
    peak, peak_kind = peaks()
    dom = 1 if kern_ms[1] >= kern_ms[2] else 2
    alg_bytes = n * BYTES_PER_FRAME
    achieved = alg_bytes / (kern_ms[dom] * 1e-3) / 1e9 if kern_ms[dom] > 0 else 0.0

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int32/int16 fixed-point", "data": "synthetic: rotations of one reference-encoded 250-packet stream (tests/golden/c2_96k_cbr_st.bit)",
        "config": {"workload": "batched CELT decode: 4096 x 48 kHz stereo 96 kbps CBR 20 ms streams (config 2)",
                   "streams_per_gpu": n, "frame_ms": 20, "packet_bytes": plen,
                   "l2": "flushed between timed steps (256 MiB write, outside the event pair)",
                   "step": "one 20 ms frame of every stream"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(n * plen + 3 * 4 * n),
                "d2h_bytes_per_step": int(n * SPF * CH * 2 + 4 * n), "ms_per_step": 1e3 * e2e_s / e2e_steps,
                "api": "opus_decode_group_submit/_wait (pinned host buffers, two calls in flight)",
                "synchronous_call_value": world * units_per_step * e2e_steps / e2e_sync_s,
                "synchronous_call_api": "opus_decode_group_flat"},
        "gpu_launches": int(launches),
        "clocks": sampler.summary(),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": ncu_traffic(), "peak_kind": peak_kind,
                     "kernel": ["plan_silk_kernel", "celt_entropy_kernel", "celt_synth_kernel"][dom],
                     "kernel_ms": {"plan": kern_ms[0], "entropy": kern_ms[1], "synth": kern_ms[2]},
                     "algorithmic_bytes_per_launch": alg_bytes,
                     "note": "integer/latency-bound path: HBM fraction is tiny by construction, see profiles/ for issue-slot utilisation"},
        "wall_ms_per_step": 1e3 * (wall1 - wall0) / K,
    }
    if rank == 0 and not args.no_cpu_baseline and world == 1:
        cores = os.cpu_count() or 1
        streams_s = 70 * cores
        r = ref_bench(streams_s, 250, cores)
        if r is not None:
            line["cpu_baseline"] = {"value": r[0] / r[1], "unit": UNIT, "cores": cores, "kind": "reference",
                                    "sample": "%d of 4096 streams x 250 frames (5 s each), one pass" % streams_s}
        else:
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": cores, "kind": "reference",
                                    "sample": "oracle/_ref/ref_bench missing"}
    if rank == 0:
        print(json.dumps(line))
    bd.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
