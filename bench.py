#!/usr/bin/env python3
"""bench.py -- one benchmark for every BASELINE.json configuration of the batched Opus engine.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config C] [--impl b200|reference]

  --config 2     (default, the driver's line) batched CELT decode: 4096 x 48 kHz stereo 96 kbit/s CBR 20 ms per GPU
  --config 2het  the same batch size with independently encoded streams of mixed rates and frame sizes
                 (64 / 96 / 128 kbit/s, CBR and VBR, a tenth of the streams at 10 ms frames)
  --config 3     batched CELT encode: 4096 x 48 kHz stereo 128 kbit/s CBR complexity 10 20 ms, real audio
  --config 4     batched SILK decode: 16384 x 16 kHz mono 32 kbit/s 20 ms
  --config 4h    the hybrid set of config 4: 16384 x 48 kHz mono 32 kbit/s hybrid fullband 20 ms
  --config 5     5.1 multistream decode: 1024 six-channel decoders, 256 kbit/s, 4 % of the packets lost

Metric: aggregate realtime stream-seconds per second.  Weak scaling: every rank (one process per GPU, launched by
torch.distributed.run for N > 1) works on its own streams; there is no exchange step, so no collective is on the
data path (NCCL carries only the barrier and the max-over-ranks of the timings).  A "step" = one batch call = one
packet (20 ms of audio, 10 ms for the short-frame streams of 2het) of every stream.

  value : inputs already in HBM, outputs left in HBM (the *_group_device entry), timed with CUDA events on the
          stream the kernels are launched on, L2 flushed between steps.  (Config 5 has no device-resident entry:
          its value is the end-to-end figure.)
  e2e   : the host-buffer C-ABI entry with page-locked host inputs and outputs, copies inside the timed region,
          two calls in flight (*_group_submit / _wait); the synchronous call is timed beside it.
  parity gate : before anything is timed, the last warm-up step's output is held to the committed golden vectors
          (the stream that plays each fixture from its start) and -- when oracle/_ref is present -- to the compiled
          reference for one stream of every distinct (fixture, rotation), and every copy must equal its twin.
          A mismatch aborts the run.
  --impl reference : the reference's own CPU implementation (oracle/_ref, libopus 1.1.2 FIXED_POINT built from
          /root/reference) on all host cores, on a bounded sample of the same workload.
  --threads-per-process N : one process, N host threads, thread d drives GPU d through the host-buffer entry
          (what a single-process P/Invoke host does); prints its own JSON line ("mode": "threads").
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
UNIT = "stream-seconds/s"
DTYPE = "int32/int16 fixed-point"


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


def crc32(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


def ncu_traffic(fname):
    """DRAM bytes per launch of the dominant kernel from a committed ncu --set full summary
    (dram__bytes_read.sum + dram__bytes_write.sum), or None."""
    tot, unit = 0.0, {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    try:
        for line in open(os.path.join(ROOT, "profiles", fname)):
            f = line.split()
            if len(f) == 3 and f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(f[2]) * unit[f[1]]
    except (OSError, KeyError, ValueError):
        return None
    return tot or None


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


# ---------------------------------------------------------------------------------------------------------------
# Workload descriptions
# ---------------------------------------------------------------------------------------------------------------
DECODE_SPECS = {
    "2": dict(fs=48000, ch=2, n=4096, fixtures=[("c2_96k_cbr_st", 1.0)], index="index.json",
              metric="aggregate realtime stream-sec/sec (48k stereo CELT decode)",
              workload="batched CELT decode: 4096 x 48 kHz stereo 96 kbps CBR 20 ms streams (config 2)",
              data="synthetic: rotations of one reference-encoded 250-packet stream (tests/golden/c2_96k_cbr_st.bit)",
              ncu=True),
    "2het": dict(fs=48000, ch=2, n=4096, index="index.json",
                 fixtures=[("het_64k_cbr_st", 0.20), ("c2_96k_cbr_st", 0.20), ("het_96k_vbr_st", 0.15),
                           ("c1_128k_cbr_st", 0.20), ("vbr_128k_st", 0.15), ("f10_64k_st", 0.10)],
                 metric="aggregate realtime stream-sec/sec (48k stereo CELT decode, mixed streams)",
                 workload="batched CELT decode: 4096 x 48 kHz stereo streams, independently encoded at 64 / 96 / 128 kbps, "
                          "CBR and VBR, a tenth of them with 10 ms frames (config 2, heterogeneous variant)",
                 data="synthetic: rotations of six reference-encoded streams of different rates (tests/golden)", ncu=None),
    "4": dict(fs=16000, ch=1, n=16384, fixtures=[("silk_wb_mono_32k", 1.0)], index="index_silk.json", in_order=True,
              metric="aggregate realtime stream-sec/sec (16k mono SILK decode)",
              workload="batched SILK decode: 16384 x 16 kHz mono 32 kbps 20 ms streams (config 4)",
              data="synthetic: rotations of one reference-encoded stream (tests/golden/silk_wb_mono_32k.bit)",
              ncu=True),
    "4h": dict(fs=48000, ch=1, n=16384, fixtures=[("hyb_fb_mono_32k", 1.0)], index="index_hybrid.json",
               metric="aggregate realtime stream-sec/sec (48k mono hybrid decode)",
               workload="batched hybrid decode: 16384 x 48 kHz mono 32 kbps 20 ms streams (config 4, hybrid set)",
               data="synthetic: rotations of one reference-encoded stream (tests/golden/hyb_fb_mono_32k.bit)", ncu=None),
}
ENC_METRIC = "aggregate realtime stream-sec/sec (48k stereo CELT encode)"
ENC_WORKLOAD = "batched CELT encode: 4096 x 48 kHz stereo 128 kbps CBR cx10 20 ms streams (config 3)"
MS_METRIC = "aggregate realtime stream-sec/sec (48k 5.1 multistream decode)"
MS_WORKLOAD = "batched 5.1 multistream decode: 1024 x 48 kHz six-channel 256 kbps 20 ms, 4 % lost (config 5)"


# committed ncu --set full summaries (profiles/) of the kernels a roofline line can name: DRAM bytes per launch
NCU_FILES = {"celt_parse": "r02_parse_ncu_raw.txt", "celt_rebuild": "r02_rebuild_ncu_raw.txt",
             "celt_synth": "r02_synth_ncu_raw.txt", "plan_silk": "r02_silk_ncu_raw.txt", "encode": "r02_encode_ncu_raw.txt"}


class ParityError(RuntimeError):
    pass


def _ref_lib():
    so = os.path.join(REF_DIR, "libopus_ref.so")
    return ctypes.CDLL(so) if os.path.exists(so) else None


class DecodeWorkload(object):
    """Configs 2, 2het, 4, 4h: n elementary decoders in one group."""

    kind = "decode"

    def __init__(self, key, n, total_steps, use_ref, gpu=True):
        spec = DECODE_SPECS[key]
        self.spec, self.key, self.n = spec, key, n
        self.fs, self.ch = spec["fs"], spec["ch"]
        self.fsz = self.fs // 50                               # frame_size argument of every call: room for 20 ms
        index = json.load(open(os.path.join(GOLDEN, spec["index"])))
        names = [f for f, _ in spec["fixtures"]]
        w = np.cumsum([x for _, x in spec["fixtures"]])
        packs = [read_bit(os.path.join(GOLDEN, nm + ".bit")) for nm in names]
        self.names, self.meta = names, [index[nm] for nm in names]
        self.npk = np.array([len(p) for p in packs])
        self.base = np.concatenate([[0], np.cumsum(self.npk)[:-1]])
        self.maxlen = max(len(p) for pk in packs for p, _ in pk)
        tot = int(self.npk.sum())
        self.table = np.zeros((tot, self.maxlen), dtype=np.uint8)
        self.plen = np.zeros(tot, dtype=np.int32)
        self.prng = np.zeros(tot, dtype=np.uint32)
        r = 0
        for pk in packs:
            for p, rng in pk:
                self.table[r, :len(p)] = np.frombuffer(p, dtype=np.uint8)
                self.plen[r] = len(p)
                self.prng[r] = rng
                r += 1
        # which fixture each stream plays (fixed pseudo-random draw by the weights) and where it starts
        u = ((np.arange(n, dtype=np.uint64) * np.uint64(2654435761)) % np.uint64(1 << 32)).astype(np.float64) / float(1 << 32)
        self.fix = np.minimum(np.searchsorted(w, u, side="right"), len(names) - 1).astype(np.int64)
        self.rot = (37 * np.arange(n)) % self.npk[self.fix]
        for k in range(min(len(names), n)):                   # stream k plays fixture k from its start (golden gate)
            self.fix[k], self.rot[k] = k, 0
        spf_fix = np.array([self._spf(m) for m in self.meta])
        self.expect = spf_fix[self.fix].astype(np.int32)       # samples every call returns per stream
        self.units_per_step = float(self.expect.sum()) / self.fs
        self.offs = (np.arange(n, dtype=np.int32) * self.maxlen).astype(np.int32)
        self.total_steps = total_steps
        self.use_ref = use_ref
        # distinct packet arenas the steps cycle through: one per packet of the longest fixture, at least 140 MB in all
        self.nbuf = max(int(self.npk.max()), -(-140000000 // (n * self.maxlen)))
        if not gpu:                                            # the reference arm only wants describe()
            return
        import torch
        import concentus_b200 as cb
        self.torch, self.cb, self.lib = torch, cb, cb.load()
        self.bd = cb.OpusBatchDecoder(self.fs, self.ch, n)
        self.stream = torch.cuda.Stream()
        self.d_off = torch.from_numpy(self.offs).cuda()
        self.d_ret = torch.zeros(n, dtype=torch.int32, device="cuda")
        self.d_pcm = [torch.zeros(n * self.fsz * self.ch, dtype=torch.int16, device="cuda") for _ in range(2)]
        nbuf = self.nbuf
        self.h_in = []
        for t in range(nbuf):
            a, l = self.inputs(t)
            self.h_in.append((torch.from_numpy(a).pin_memory(), torch.from_numpy(l).pin_memory()))
        self.d_in = [(a.cuda(), l.cuda()) for a, l in self.h_in]   # every step's packets resident in HBM up front
        self.d_expect = torch.from_numpy(self.expect).cuda()
        self.h_off = torch.from_numpy(self.offs).pin_memory()
        self.t = 0
        torch.cuda.synchronize()

    def _spf(self, meta):
        return int(meta["samples_per_frame"])

    def row_index(self, t):
        return self.base[self.fix] + (self.rot + t) % self.npk[self.fix]

    def inputs(self, t):
        idx = self.row_index(t)
        return np.ascontiguousarray(self.table[idx]).reshape(-1), self.plen[idx].copy()

    # ---- device-resident call ----
    def step_device(self, t, own_stream=True):
        """own_stream: queue on this bench's CUDA stream (calls run in order); else on the library's streams, where
        consecutive calls are pipelined (the parse side of a call runs beside the rebuild / synthesis of the one before)."""
        a, l = self.d_in[t % len(self.d_in)]
        rc = self.lib.opus_decode_group_device(self.bd._g, a.data_ptr(), self.d_off.data_ptr(), l.data_ptr(),
                                               self.d_pcm[t & 1].data_ptr(), self.fsz * self.ch, self.fsz, None,
                                               self.d_ret.data_ptr(), 1, self.maxlen,
                                               ctypes.c_void_p(self.stream.cuda_stream) if own_stream else None)
        if rc != 0:
            raise RuntimeError("opus_decode_group_device failed: %d" % rc)

    def timed_pipelined(self, K):
        """K device-resident steps on the library's own streams, one timing mark before and one after.  Every step
        reads its own packet arena (len(self.d_in) distinct ones, larger than L2 together)."""
        self.torch.cuda.synchronize()
        for _ in range(3):
            self.step_device(self.t, own_stream=False)
            self.t += 1
        assert self.lib.opus_b200_synchronize() == 0
        assert self.lib.opus_b200_timer_mark(0) == 0
        for _ in range(K):
            self.step_device(self.t, own_stream=False)
            self.t += 1
        assert self.lib.opus_b200_timer_mark(1) == 0
        ms = ctypes.c_double(0)
        assert self.lib.opus_b200_timer_elapsed(0, 1, ctypes.byref(ms)) == 0
        assert self.lib.opus_b200_synchronize() == 0
        self.check_ret("pipelined device-resident steps")
        return ms.value

    def check_ret(self, where):
        bad = int((self.d_ret != self.d_expect).sum().item())
        if bad:
            raise ParityError("%d streams returned an unexpected sample count (%s)" % (bad, where))

    def warmup(self, W):
        for _ in range(W):
            self.step_device(self.t)
            self.t += 1
        self.torch.cuda.synchronize()
        self.check_ret("warm-up")

    def gate(self):
        """Output of the last warm-up step against the golden vectors / the compiled reference / the twins."""
        t = self.t - 1
        pcm = self.d_pcm[t & 1].cpu().numpy().reshape(self.n, self.fsz, self.ch)
        fr = self.bd.final_ranges()
        res = {"step": t, "golden_streams": 0, "reference_streams": 0, "twins_equal": True}
        if t < int(self.npk.min()):
            for k in range(min(len(self.names), self.n)):
                want = self.meta[k]["pcm_crc32"][t]
                got = crc32(pcm[k, :self.expect[k]])
                if got != want:
                    raise ParityError("stream %d (%s from its start): PCM CRC %08x != golden %08x at packet %d"
                                      % (k, self.names[k], got, want, t))
                if int(fr[k]) != int(self.prng[self.base[k] + t]):
                    raise ParityError("stream %d (%s): final range differs from the encoder's at packet %d" % (k, self.names[k], t))
                res["golden_streams"] += 1
        key = self.fix * 100000 + self.rot
        order = np.argsort(key, kind="stable")
        ks = key[order]
        starts = np.nonzero(np.concatenate([[True], ks[1:] != ks[:-1]]))[0]
        reps = order[starts]
        rep_of = np.repeat(reps, np.diff(np.concatenate([starts, [self.n]])))
        same = (pcm[order] == pcm[rep_of]).reshape(self.n, -1).all(axis=1) & (fr[order] == fr[rep_of])
        if not bool(same.all()):
            raise ParityError("streams with the same packets disagree: %s ..." % order[~same][:8].tolist())
        res["distinct_streams"] = int(len(reps))
        ref = _ref_lib() if self.use_ref else None
        if ref is not None:
            ref.opus_decoder_create.restype = ctypes.c_void_p
            ref.opus_decoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
            ref.opus_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
            ref.opus_decoder_destroy.argtypes = [ctypes.c_void_p]
            out = np.zeros((self.fsz, self.ch), dtype=np.int16)
            v = ctypes.c_uint32(0)
            for i in reps:
                err = ctypes.c_int(0)
                h = ref.opus_decoder_create(self.fs, self.ch, ctypes.byref(err))
                for s in range(t + 1):
                    r = int(self.base[self.fix[i]] + (self.rot[i] + s) % self.npk[self.fix[i]])
                    ret = ref.opus_decode(h, self.table[r, :self.plen[r]].tobytes(), int(self.plen[r]), out.ctypes.data, self.fsz, 0)
                ref.opus_decoder_ctl(ctypes.c_void_p(h), 4031, ctypes.byref(v))
                ref.opus_decoder_destroy(h)
                if ret != self.expect[i] or not np.array_equal(out[:ret], pcm[i, :ret]) or v.value != int(fr[i]):
                    raise ParityError("stream %d (%s, start %d): output differs from the compiled reference at step %d"
                                      % (i, self.names[self.fix[i]], self.rot[i], t))
                res["reference_streams"] += 1
        return res

    def timed_device(self, K, flush):
        torch = self.torch
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        with torch.cuda.stream(self.stream):
            for k in range(K):
                flush.fill_(k & 0xFF)                       # L2 flush between timed iterations (not timed)
                ev[k][0].record(self.stream)
                self.step_device(self.t)
                ev[k][1].record(self.stream)
                self.t += 1
        return ev

    def profile(self, K, flush):
        torch = self.torch
        self.lib.opus_b200_profile_begin()
        with torch.cuda.stream(self.stream):
            for k in range(K):
                flush.fill_(k & 0xFF)
                self.step_device(self.t)
                self.t += 1
        ms4 = (ctypes.c_double * 4)()
        cnt4 = (ctypes.c_longlong * 4)()
        self.lib.opus_b200_profile_end(ms4, cnt4)
        return {"plan_silk": ms4[0] / max(1, cnt4[0]), "celt_parse": ms4[1] / max(1, cnt4[1]),
                "celt_rebuild": ms4[2] / max(1, cnt4[2]), "celt_synth": ms4[3] / max(1, cnt4[3])}

    # ---- host-buffer calls ----
    def _host_bufs(self):
        if not hasattr(self, "h_pcm"):
            torch = self.torch
            self.h_pcm = [torch.zeros(self.n * self.fsz * self.ch, dtype=torch.int16).pin_memory() for _ in range(4)]
            self.h_ret = [torch.zeros(self.n, dtype=torch.int32).pin_memory() for _ in range(4)]
            self.h_expect = torch.from_numpy(self.expect)

    def _call(self, fn, t, slot):
        a, l = self.h_in[t % len(self.h_in)]
        return fn(self.bd._g, a.data_ptr(), a.numel(), self.h_off.data_ptr(), l.data_ptr(), self.h_pcm[slot].data_ptr(),
                  self.fsz * self.ch, self.fsz, None, self.h_ret[slot].data_ptr())

    def e2e_sync(self, K):
        self._host_bufs()
        for _ in range(3):
            assert self._call(self.lib.opus_decode_group_flat, self.t, 0) == 0
            self.t += 1
        s0 = time.perf_counter()
        for _ in range(K):
            rc = self._call(self.lib.opus_decode_group_flat, self.t, 0)
            assert rc == 0 and bool((self.h_ret[0] == self.h_expect).all())
            self.t += 1
        return time.perf_counter() - s0

    def e2e_pipelined(self, K, barrier=None):
        self._host_bufs()
        for k in range(4):                                   # warm-up of the pipelined entry (it allocates its slots)
            assert self._call(self.lib.opus_decode_group_submit, self.t, k % 4) == 0
            self.t += 1
        assert self.lib.opus_decode_group_wait(self.bd._g) == 0
        if barrier:
            barrier()
        e0 = time.perf_counter()
        for k in range(K):
            assert self._call(self.lib.opus_decode_group_submit, self.t, k % 4) == 0
            if k >= 3:
                assert bool((self.h_ret[(k - 3) % 4] == self.h_expect).all())   # call k-3 completed before submit k returned
            self.t += 1
        assert self.lib.opus_decode_group_wait(self.bd._g) == 0
        e1 = time.perf_counter()
        assert all(bool((r == self.h_expect).all()) for r in self.h_ret)
        return e1 - e0

    def describe(self):
        mean_len = float(self.plen[self.row_index(0)].mean())
        pcm_bytes = float(self.expect.mean()) * self.ch * 2
        cfg = {"workload": self.spec["workload"], "streams_per_gpu": self.n, "frame_ms": 20,
               "mean_packet_bytes": mean_len,
               "step": "one packet of every stream"}
        if self.spec.get("in_order"):
            cfg["l2"] = "flushed between timed steps (256 MiB write, outside the event pair)"
            cfg["device_resident_calls"] = "queued in order on the bench's stream (opus_decode_group_device with a stream)"
        else:
            cfg["l2"] = ("inputs larger than L2: every step reads its own packet arena from HBM (%d arenas, %.0f MB); the "
                         "decoders' state is the streams' own working set.  isolated_steps gives the per-step figure with L2 "
                         "flushed between steps" % (self.nbuf, self.nbuf * self.n * self.maxlen / 1e6))
            cfg["device_resident_calls"] = ("queued back to back on the library's streams (opus_decode_group_device, cuda_stream "
                                            "NULL): the parse side of a call runs beside the rebuild / synthesis of the call before")
        if len(self.names) > 1:
            cfg["mix"] = {nm: int((self.fix == k).sum()) for k, nm in enumerate(self.names)}
        return dict(metric=self.spec["metric"], data=self.spec["data"], config=cfg,
                    alg_bytes=int(self.n * (mean_len + pcm_bytes)),
                    h2d=int(self.n * self.maxlen + 3 * 4 * self.n), d2h=int(self.n * self.fsz * self.ch * 2 + 4 * self.n),
                    api="opus_decode_group_submit/_wait (pinned host buffers, three calls in flight)",
                    sync_api="opus_decode_group_flat",
                    kernels={"plan_silk": "plan_silk_kernel", "celt_parse": "celt_parse_kernel",
                             "celt_rebuild": "celt_rebuild_kernel", "celt_synth": "celt_synth_kernel"},
                    ncu=self.spec.get("ncu"))

    def close(self):
        self.bd.close()


class EncodeWorkload(object):
    """Config 3: n encoders, real audio (tests/golden/audio_48k_stereo_5s.pcm) rotated per stream, 128 kbit/s CBR."""

    kind = "encode"
    FS, CH, SPF, BITRATE, CX, MAXB = 48000, 2, 960, 128000, 10, 1275

    def __init__(self, key, n, total_steps, use_ref, gpu=True):
        self.n, self.use_ref = n, use_ref
        if not gpu:
            return
        import torch
        import concentus_b200 as cb
        import enc_cases
        self.torch, self.cb, self.lib, self.ec = torch, cb, cb.load(), enc_cases
        self.pcm = enc_cases.real_pcm(self.CH)
        self.rot = enc_cases.c3_rotations(n)
        self.rot[0] = 0                                        # stream 0 plays the excerpt from its start (golden gate)
        self.units_per_step = n * self.SPF / self.FS
        self.be = cb.OpusBatchEncoder(n, self.FS, self.CH, enc_cases.LOWDELAY)
        self.ctls = [(enc_cases.SET_BITRATE, self.BITRATE), (enc_cases.SET_VBR, 0), (enc_cases.SET_COMPLEXITY, self.CX)]
        for req, val in self.ctls:
            self.be.ctl(req, val)
        nbuf = min(total_steps, enc_cases.REAL_FRAMES)
        self.h_rows = [torch.from_numpy(enc_cases.c3_rows(self.pcm, self.CH, self.rot, t)).pin_memory() for t in range(nbuf)]
        self.d_rows = [r.cuda() for r in self.h_rows]
        self.d_data = torch.zeros((n, self.MAXB), dtype=torch.uint8, device="cuda")
        self.d_len = torch.zeros(n, dtype=torch.int32, device="cuda")
        self.t = 0
        torch.cuda.synchronize()

    def step_device(self, t):
        self.be.encode_device(self.d_rows[t % len(self.d_rows)].data_ptr(), self.SPF, self.d_data.data_ptr(), self.MAXB,
                              self.d_len.data_ptr())

    def warmup(self, W):
        for _ in range(W):
            self.step_device(self.t)
            self.t += 1
        self.lib.opus_b200_synchronize()
        if int((self.d_len <= 0).sum().item()):
            raise ParityError("encode errors in warm-up")

    def gate(self):
        t = self.t - 1
        lens = self.d_len.cpu().numpy()
        data = self.d_data.cpu().numpy()
        fr = self.be.final_ranges()
        res = {"step": t, "golden_streams": 0, "reference_streams": 0, "twins_equal": True}
        gold = read_bit(os.path.join(GOLDEN, "encr_st_128k_cbr.bit"))
        if t < len(gold):
            if data[0, :lens[0]].tobytes() != gold[t][0] or int(fr[0]) != gold[t][1]:
                raise ParityError("stream 0: packet %d differs from the reference encoder's golden packet" % t)
            res["golden_streams"] = 1
        order = np.argsort(self.rot, kind="stable")
        ks = self.rot[order]
        starts = np.nonzero(np.concatenate([[True], ks[1:] != ks[:-1]]))[0]
        reps = order[starts]
        rep_of = np.repeat(reps, np.diff(np.concatenate([starts, [self.n]])))
        same = (lens[order] == lens[rep_of]) & (fr[order] == fr[rep_of]) & (data[order] == data[rep_of]).all(axis=1)
        if not bool(same.all()):
            raise ParityError("streams with the same PCM produced different packets: %s ..." % order[~same][:8].tolist())
        res["distinct_streams"] = int(len(reps))
        ref = _ref_lib() if self.use_ref else None
        if ref is not None:
            ref.opus_encoder_create.restype = ctypes.c_void_p
            ref.opus_encoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
            ref.opus_encode.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int32]
            ref.opus_encoder_destroy.argtypes = [ctypes.c_void_p]
            buf = (ctypes.c_ubyte * self.MAXB)()
            v = ctypes.c_uint32(0)
            frames = self.pcm.reshape(self.ec.REAL_FRAMES, self.SPF * self.CH)
            for i in reps:
                err = ctypes.c_int(0)
                h = ref.opus_encoder_create(self.FS, self.CH, self.ec.LOWDELAY, ctypes.byref(err))
                for req, val in self.ctls:
                    ref.opus_encoder_ctl(ctypes.c_void_p(h), ctypes.c_int(req), ctypes.c_int(val))
                for s in range(t + 1):
                    row = np.ascontiguousarray(frames[(self.rot[i] + s) % self.ec.REAL_FRAMES])
                    nb = ref.opus_encode(h, row.ctypes.data, self.SPF, buf, self.MAXB)
                ref.opus_encoder_ctl(ctypes.c_void_p(h), 4031, ctypes.byref(v))
                ref.opus_encoder_destroy(h)
                if nb != lens[i] or bytes(buf[:nb]) != data[i, :nb].tobytes() or v.value != int(fr[i]):
                    raise ParityError("stream %d (start frame %d): packet differs from the compiled reference at step %d"
                                      % (i, self.rot[i], t))
                res["reference_streams"] += 1
        return res

    def timed_device(self, K, flush):
        # the encoder's device entry runs on the library's own stream: time it with the library's events
        self.lib.opus_b200_encode_profile_begin()
        for k in range(K):
            flush.fill_(k & 0xFF)
            self.torch.cuda.synchronize()
            self.step_device(self.t)
            self.t += 1
        self.lib.opus_b200_synchronize()
        ms = ctypes.c_double(0)
        cnt = ctypes.c_longlong(0)
        self.lib.opus_b200_encode_profile_end(ctypes.byref(ms), ctypes.byref(cnt))
        if int((self.d_len <= 0).sum().item()):
            raise ParityError("encode errors in the timed region")
        return ms.value, int(cnt.value)

    def _host_bufs(self):
        if not hasattr(self, "h_data"):
            torch = self.torch
            self.h_data = [torch.zeros((self.n, self.MAXB), dtype=torch.uint8).pin_memory() for _ in range(3)]
            self.h_len = [torch.zeros(self.n, dtype=torch.int32).pin_memory() for _ in range(3)]

    def _call(self, fn, t, slot):
        return fn(self.be._g, self.h_rows[t % len(self.h_rows)].data_ptr(), self.SPF, self.h_data[slot].data_ptr(), self.MAXB,
                  self.h_len[slot].data_ptr())

    def e2e_sync(self, K):
        self._host_bufs()
        for _ in range(3):
            assert self._call(self.lib.opus_encode_group_flat, self.t, 0) == 0
            self.t += 1
        s0 = time.perf_counter()
        for _ in range(K):
            rc = self._call(self.lib.opus_encode_group_flat, self.t, 0)
            assert rc == 0 and bool((self.h_len[0] > 0).all())
            self.t += 1
        return time.perf_counter() - s0

    def e2e_pipelined(self, K, barrier=None):
        self._host_bufs()
        for k in range(3):
            assert self._call(self.lib.opus_encode_group_submit, self.t, k % 3) == 0
            self.t += 1
        assert self.lib.opus_encode_group_wait(self.be._g) == 0
        if barrier:
            barrier()
        e0 = time.perf_counter()
        for k in range(K):
            assert self._call(self.lib.opus_encode_group_submit, self.t, k % 3) == 0
            if k >= 2:
                assert bool((self.h_len[(k - 2) % 3] > 0).all())
            self.t += 1
        assert self.lib.opus_encode_group_wait(self.be._g) == 0
        e1 = time.perf_counter()
        assert all(bool((l > 0).all()) for l in self.h_len)
        return e1 - e0

    def describe(self):
        mean_len = float(self.BITRATE / 8 / 50)
        return dict(metric=ENC_METRIC,
                    data="synthetic batch of real audio: 250 x 20 ms of the reference's AudioData/48Khz Stereo.raw "
                         "(tests/golden/audio_48k_stereo_5s.pcm), rotated per stream (seed-1234 LCG)",
                    config={"workload": ENC_WORKLOAD, "streams_per_gpu": self.n, "frame_ms": 20, "packet_bytes": mean_len,
                            "l2": "flushed between timed steps (256 MiB write, outside the event pair)",
                            "step": "one 20 ms frame of every stream"},
                    alg_bytes=int(self.n * (self.SPF * self.CH * 2 + mean_len)),
                    h2d=int(self.n * self.SPF * self.CH * 2), d2h=int(self.n * self.MAXB + 4 * self.n),
                    api="opus_encode_group_submit/_wait (pinned host buffers, two calls in flight)",
                    sync_api="opus_encode_group_flat", kernels={"encode": "opus_encode_kernel"}, ncu=True)

    def close(self):
        self.be.close()


class MSWorkload(object):
    """Config 5: n six-channel multistream decoders (4 elementary streams each); host-buffer entry only."""

    kind = "ms"
    FS, CH, SPF = 48000, 6, 960

    def __init__(self, key, n, total_steps, use_ref, gpu=True):
        self.n, self.use_ref = n, use_ref
        self.meta = json.load(open(os.path.join(GOLDEN, "index_ms.json")))["ms51_256k"]
        pk = [p for p, _ in read_bit(os.path.join(GOLDEN, "ms51_256k.bit"))]
        self.npk = len(pk)
        self.maxlen = max(len(p) for p in pk)
        self.table = np.zeros((self.npk, self.maxlen), dtype=np.uint8)
        self.plen = np.zeros(self.npk, dtype=np.int32)
        for i, p in enumerate(pk):
            self.table[i, :len(p)] = np.frombuffer(p, dtype=np.uint8)
            self.plen[i] = len(p)
        self.rot = (37 * np.arange(n)) % self.npk
        self.base_off = (np.arange(n, dtype=np.int32) * self.maxlen).astype(np.int32)
        self.units_per_step = n * self.SPF / self.FS
        self.lost_fraction = float((self.plen == 0).mean())
        if not gpu:
            return
        import torch
        import concentus_b200 as cb
        self.torch, self.cb, self.lib = torch, cb, cb.load()
        nbuf = min(total_steps, self.npk)
        self.inputs = []
        for t in range(nbuf):
            idx = (self.rot + t) % self.npk
            lens = self.plen[idx].copy()
            offs = np.where(lens > 0, self.base_off, -1).astype(np.int32)
            self.inputs.append((torch.from_numpy(np.ascontiguousarray(self.table[idx]).reshape(-1)).pin_memory(),
                                torch.from_numpy(offs).pin_memory(), torch.from_numpy(lens).pin_memory()))
        m = self.meta
        self.bd = cb.OpusMSBatchDecoder(self.FS, self.CH, m["streams"], m["coupled"], m["mapping"], n)
        self.h_pcm = [torch.zeros(n * self.SPF * self.CH, dtype=torch.int16).pin_memory() for _ in range(4)]
        self.h_ret = [torch.zeros(n, dtype=torch.int32).pin_memory() for _ in range(4)]
        self.t = 0

    def _call(self, fn, t, slot):
        a, o, l = self.inputs[t % len(self.inputs)]
        return fn(self.bd._g, a.data_ptr(), a.numel(), o.data_ptr(), l.data_ptr(), self.h_pcm[slot].data_ptr(),
                  self.SPF * self.CH, self.SPF, None, self.h_ret[slot].data_ptr())

    def warmup(self, W):
        for _ in range(W):
            assert self._call(self.lib.opus_multistream_decode_group_flat, self.t, 0) == 0
            if not bool((self.h_ret[0] == self.SPF).all()):
                raise ParityError("multistream decode errors in warm-up")
            self.t += 1

    def gate(self):
        t = self.t - 1
        pcm = self.h_pcm[0].numpy().reshape(self.n, self.SPF, self.CH)
        res = {"step": t, "golden_streams": 0, "reference_streams": 0, "twins_equal": True}
        if t < self.npk:                                     # decoder 0 plays the fixture from its start
            if crc32(pcm[0]) != self.meta["pcm_crc32"][t]:
                raise ParityError("decoder 0: PCM CRC differs from the golden one at packet %d" % t)
            res["golden_streams"] = 1
        order = np.argsort(self.rot, kind="stable")
        ks = self.rot[order]
        starts = np.nonzero(np.concatenate([[True], ks[1:] != ks[:-1]]))[0]
        reps = order[starts]
        rep_of = np.repeat(reps, np.diff(np.concatenate([starts, [self.n]])))
        same = (pcm[order] == pcm[rep_of]).reshape(self.n, -1).all(axis=1)
        if not bool(same.all()):
            raise ParityError("decoders with the same packets disagree: %s ..." % order[~same][:8].tolist())
        res["distinct_streams"] = int(len(reps))
        ref = _ref_lib() if self.use_ref else None
        if ref is not None:
            m = self.meta
            ref.opus_multistream_decoder_create.restype = ctypes.c_void_p
            ref.opus_multistream_decoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                            ctypes.c_char_p, ctypes.POINTER(ctypes.c_int)]
            ref.opus_multistream_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int32, ctypes.c_void_p,
                                                    ctypes.c_int, ctypes.c_int]
            ref.opus_multistream_decoder_destroy.argtypes = [ctypes.c_void_p]
            out = np.zeros((self.SPF, self.CH), dtype=np.int16)
            for i in reps:
                err = ctypes.c_int(0)
                h = ref.opus_multistream_decoder_create(self.FS, self.CH, m["streams"], m["coupled"], bytes(m["mapping"]),
                                                        ctypes.byref(err))
                for s in range(t + 1):
                    r = int((self.rot[i] + s) % self.npk)
                    ln = int(self.plen[r])
                    ret = ref.opus_multistream_decode(h, self.table[r, :ln].tobytes() if ln else None, ln, out.ctypes.data,
                                                      self.SPF, 0)
                ref.opus_multistream_decoder_destroy(h)
                if ret != self.SPF or not np.array_equal(out, pcm[i]):
                    raise ParityError("decoder %d (start %d): output differs from the compiled reference at step %d"
                                      % (i, self.rot[i], t))
                res["reference_streams"] += 1
        return res

    def timed_device(self, K, flush):
        return None

    def profile(self, K, flush):
        self.lib.opus_b200_profile_begin()
        for _ in range(K):
            assert self._call(self.lib.opus_multistream_decode_group_flat, self.t, 0) == 0
            self.t += 1
        ms4 = (ctypes.c_double * 4)()
        cnt4 = (ctypes.c_longlong * 4)()
        self.lib.opus_b200_profile_end(ms4, cnt4)
        return {"celt_parse": ms4[1] / K, "celt_rebuild": ms4[2] / K, "celt_synth": ms4[3] / K}

    def e2e_sync(self, K):
        s0 = time.perf_counter()
        for _ in range(K):
            rc = self._call(self.lib.opus_multistream_decode_group_flat, self.t, 0)
            assert rc == 0 and bool((self.h_ret[0] == self.SPF).all())
            self.t += 1
        return time.perf_counter() - s0

    def e2e_pipelined(self, K, barrier=None):
        fn = self.lib.opus_multistream_decode_group_submit
        for k in range(4):
            assert self._call(fn, self.t, k % 4) == 0
            self.t += 1
        assert self.lib.opus_multistream_decode_group_wait(self.bd._g) == 0
        if barrier:
            barrier()
        e0 = time.perf_counter()
        for k in range(K):
            assert self._call(fn, self.t, k % 4) == 0
            if k >= 3:
                assert bool((self.h_ret[(k - 3) % 4] == self.SPF).all())
            self.t += 1
        assert self.lib.opus_multistream_decode_group_wait(self.bd._g) == 0
        e1 = time.perf_counter()
        assert all(bool((r == self.SPF).all()) for r in self.h_ret)
        return e1 - e0

    def describe(self):
        mean_len = float(self.plen[self.plen > 0].mean())
        return dict(metric=MS_METRIC,
                    data="synthetic: rotations of one reference-encoded 5.1 stream (tests/golden/ms51_256k.bit)",
                    config={"workload": MS_WORKLOAD, "streams_per_gpu": self.n, "elementary_streams": self.n * self.meta["streams"],
                            "frame_ms": 20, "mean_packet_bytes": mean_len, "lost_packet_fraction": self.lost_fraction,
                            "l2": "inputs and outputs larger than L2 across steps (12 MB PCM per step, host round trip each step)",
                            "step": "one 20 ms frame of every decoder"},
                    alg_bytes=int(self.n * (mean_len + self.SPF * self.CH * 2)),
                    h2d=int(self.n * self.maxlen + 12 * self.n), d2h=int(self.n * self.SPF * self.CH * 2 + 4 * self.n),
                    api="opus_multistream_decode_group_submit/_wait (pinned host buffers, three calls in flight)",
                    sync_api="opus_multistream_decode_group_flat",
                    kernels={"celt_parse": "celt_parse_kernel", "celt_rebuild": "celt_rebuild_kernel",
                             "celt_synth": "celt_synth_kernel"}, ncu=None)

    def close(self):
        self.bd.close()


def make_workload(key, n, total_steps, use_ref, gpu=True):
    if key in DECODE_SPECS:
        return DecodeWorkload(key, n or DECODE_SPECS[key]["n"], total_steps, use_ref, gpu)
    if key == "3":
        return EncodeWorkload(key, n or 4096, total_steps, use_ref, gpu)
    if key == "5":
        return MSWorkload(key, n or 1024, total_steps, use_ref, gpu)
    raise SystemExit("unknown --config " + key)


# ---------------------------------------------------------------------------------------------------------------
# Reference arm / CPU baseline: the compiled reference on the host cores
# ---------------------------------------------------------------------------------------------------------------
def _run_ref(exe, argv):
    path = os.path.join(REF_DIR, exe)
    if not os.path.exists(path):
        return None
    try:
        out = subprocess.run([path] + [str(a) for a in argv], stdout=subprocess.PIPE, text=True, check=True).stdout.split()
    except (subprocess.CalledProcessError, OSError):       # e.g. the -O3 x86-64-v3 build on a host without AVX2
        return None
    return float(out[0]), float(out[1])


def reference_pass(key, frames, cores, build=""):
    """One bounded pass of the reference over a sample of the configuration's workload: (audio seconds, wall
    seconds, description).  `frames` = packets per stream.  build "" = -O2 (the reference's own flags), "_o3"."""
    if key in DECODE_SPECS:
        spec = DECODE_SPECS[key]
        per_core = {"2": 16, "2het": 16, "4": 96, "4h": 48}[key]
        audio = wall = 0.0
        parts = []
        for nm, wgt in spec["fixtures"]:
            streams = max(cores, int(round(per_core * cores * wgt)))
            r = _run_ref("ref_bench" + build, [os.path.join(GOLDEN, nm + ".bit"), spec["fs"], spec["ch"], streams, frames, cores])
            if r is None:
                return None
            audio += r[0]
            wall += r[1]
            parts.append("%d x %s" % (streams, nm))
        return audio, wall, "%s of %d streams, %d packets each" % (" + ".join(parts), spec["n"], frames)
    if key == "3":
        streams = 8 * cores
        r = _run_ref("ref_enc_bench" + build, [os.path.join(GOLDEN, "audio_48k_stereo_5s.pcm"), 48000, 2, 128000, 10, 960,
                                               streams, frames, cores, 0])
        return None if r is None else (r[0], r[1], "%d of 4096 streams x %d frames" % (streams, frames))
    if key == "5":
        meta = json.load(open(os.path.join(GOLDEN, "index_ms.json")))["ms51_256k"]
        streams = 8 * cores
        r = _run_ref("ref_bench" + build, [os.path.join(GOLDEN, "ms51_256k.bit"), 48000, 6, streams, frames, cores,
                                           meta["streams"], meta["coupled"], ",".join(str(m) for m in meta["mapping"])])
        return None if r is None else (r[0], r[1], "%d of 1024 decoders x %d frames" % (streams, frames))
    return None


def run_reference(args, rank):
    """Reference arm: the reference's own decoder / encoder on the host cores, all threads; every step is one
    bounded pass (1 s of audio per sampled stream).  Rank 0 alone works."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    d = make_workload(args.config, args.streams, 0, False, gpu=False).describe()   # same metric / data / config as the GPU arm
    frames = 50
    audio = wall = 0.0
    sample = ""
    for it in range(args.warmup + args.steps):
        r = reference_pass(args.config, frames, cores)
        if r is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built (needs /root/reference at build time)"}))
            return
        if it >= args.warmup:
            audio += r[0]
            wall += r[1]
        sample = r[2]
    value = audio / wall
    sample = "%s per step, %d steps" % (sample, args.steps)
    o3 = reference_pass(args.config, frames, cores, "_o3")
    line = {
        "impl": "reference", "metric": d["metric"], "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": d["data"],
        "config": d["config"],
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample,
                         "build": "-O2 (the reference's Makefile.unix flags)",
                         "O3_x86-64-v3": (o3[0] / o3[1]) if o3 else None},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def cpu_baseline(key):
    cores = os.cpu_count() or 1
    frames = 200
    o2 = reference_pass(key, frames, cores)
    if o2 is None:
        return {"value": None, "unit": UNIT, "cores": cores, "kind": "reference", "sample": "oracle/_ref missing"}
    o3 = reference_pass(key, frames, cores, "_o3")
    return {"value": o2[0] / o2[1], "unit": UNIT, "cores": cores, "kind": "reference", "sample": o2[2] + ", one pass",
            "O2": o2[0] / o2[1], "O3_x86-64-v3": (o3[0] / o3[1]) if o3 else None,
            "note": "value = the -O2 build (the reference's own Makefile.unix flags); the -O3 build is reported beside it"}


# ---------------------------------------------------------------------------------------------------------------
def run_threads(args):
    """One process, one host thread per GPU, host-buffer pipelined entry (what a single-process P/Invoke host does)."""
    import torch
    ndev = args.threads_per_process
    K, W = args.steps, args.warmup
    bar = threading.Barrier(ndev)
    res, errs = [None] * ndev, []

    def work(d):
        try:
            torch.cuda.set_device(d)
            wl = make_workload(args.config, args.streams, W + K + 16, False)
            wl.warmup(W)
            wl.gate()
            dt = wl.e2e_pipelined(K, barrier=bar.wait)
            res[d] = (dt, wl.units_per_step)
            bar.wait()
            wl.close()
        except Exception as e:
            errs.append((d, repr(e)))
            bar.abort()

    th = [threading.Thread(target=work, args=(d,)) for d in range(ndev)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    if errs:
        raise SystemExit("thread mode failed: %r" % errs)
    dt = max(r[0] for r in res)
    units = sum(r[1] for r in res)
    d = make_workload(args.config, args.streams, 0, False, gpu=False).describe()
    print(json.dumps({"mode": "threads", "metric": d["metric"], "value": units * K / dt, "unit": UNIT, "n_gpus": ndev, "steps": K,
                      "warmup": W, "ms_per_step": 1e3 * dt / K, "higher_is_better": True, "scaling": "weak", "dtype": DTYPE,
                      "config": dict(d["config"], host="one process, %d threads, thread d drives GPU d (submit/wait entry)" % ndev),
                      "e2e": {"value": units * K / dt, "unit": UNIT}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="2", choices=["2", "2het", "3", "4", "4h", "5"])
    ap.add_argument("--streams", type=int, default=0, help="streams per GPU (default: the configuration's size)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--threads-per-process", type=int, default=0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank)
        return
    if args.threads_per_process > 0:
        run_threads(args)
        return

    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))
    K, W = args.steps, args.warmup
    use_ref = (rank == 0 and not args.no_cpu_baseline)
    wl = make_workload(args.config, args.streams, W + 3 * K + 32, use_ref)
    lib = wl.lib
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up, then the parity gate on its last step ----
    wl.warmup(W)
    gate = wl.gate()
    barrier()

    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = lib.opus_b200_launch_count()
    # ---- timed: K steps, device-resident inputs, CUDA events on the launching stream ----
    barrier()
    wall0 = time.perf_counter()
    timed = wl.timed_device(K, flush)
    barrier()
    wall1 = time.perf_counter()
    launches = lib.opus_b200_launch_count() - launches0
    kern_ms = {}
    isolated = None
    if wl.kind == "decode":
        iso_ms = float(sum(a.elapsed_time(b) for a, b in timed))
        wl.check_ret("timed region")
        isolated = {"ms_per_step": iso_ms / K,
                    "what": "each step alone on one stream, L2 flushed between steps (256 MiB write outside the event pair)"}
        kern_ms = wl.profile(min(K, 50), flush)
        barrier()
        if wl.spec.get("in_order"):
            # SILK-only packets have no parse work for the run-ahead to hide: the device-resident calls go on the
            # bench's own stream, in order, and `value` is the per-step figure (L2 flushed between steps)
            dev_ms = iso_ms
        else:
            dev_ms = wl.timed_pipelined(K)
    elif wl.kind == "encode":
        dev_ms, cnt = timed
        kern_ms = {"encode": dev_ms / max(1, cnt)}
    else:
        dev_ms = None
        kern_ms = wl.profile(min(K, 50), flush)

    # ---- e2e: host pinned inputs -> C-ABI -> host pinned outputs, wall clock around the calls ----
    barrier()
    e2e_sync_s = wl.e2e_sync(K)
    barrier()
    e2e_s = wl.e2e_pipelined(K)
    if dev_ms is None:
        launches = lib.opus_b200_launch_count() - launches0
        dev_ms = 1e3 * e2e_s                                    # no device-resident entry: value = end to end
    sampler.stop_flag = True
    sampler.join(timeout=2)

    # ---- reduce over ranks: max time ----
    if world > 1:
        tt = torch.tensor([dev_ms, e2e_s, e2e_sync_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dev_ms, e2e_s, e2e_sync_s = float(tt[0].item()), float(tt[1].item()), float(tt[2].item())

    if world > 1 and isolated is not None:
        tt = torch.tensor([isolated["ms_per_step"]], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        isolated["ms_per_step"] = float(tt[0].item())
    d = wl.describe()
    units = wl.units_per_step
    if isolated is not None:
        isolated["value"] = world * units / (isolated["ms_per_step"] * 1e-3)
    value = world * units * K / (dev_ms * 1e-3)
    e2e_value = world * units * K / e2e_s
    peak, peak_kind = peaks()
    dom = max(kern_ms, key=kern_ms.get)
    achieved = d["alg_bytes"] / (kern_ms[dom] * 1e-3) / 1e9 if kern_ms[dom] > 0 else 0.0
    line = {
        "metric": d["metric"], "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": DTYPE, "data": d["data"], "config": d["config"],
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": d["h2d"], "d2h_bytes_per_step": d["d2h"],
                "ms_per_step": 1e3 * e2e_s / K, "api": d["api"],
                "synchronous_call_value": world * units * K / e2e_sync_s, "synchronous_call_api": d["sync_api"]},
        "gpu_launches": int(launches),
        "clocks": sampler.summary(),
        "parity_gate": gate,
        "isolated_steps": isolated,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(NCU_FILES[dom]) if (dom in NCU_FILES and d["ncu"]) else None, "peak_kind": peak_kind,
                     "kernel": d["kernels"][dom], "kernel_ms": kern_ms, "algorithmic_bytes_per_launch": d["alg_bytes"],
                     "note": "integer/latency-bound path: the HBM fraction is small by construction, profiles/ holds the "
                             "issue-slot utilisation per kernel"},
        "wall_ms_per_step": 1e3 * (wall1 - wall0) / K,
    }
    if rank == 0 and not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_baseline(args.config)
    if rank == 0:
        print(json.dumps(line))
    wl.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
