#!/usr/bin/env python3
"""Timeline of the decode kernels of a few pipelined device-resident calls (config 2): start / end of every launch
relative to the first one, from event pairs recorded on the launching streams (opus_b200_trace_begin/_end).
Shows whether the parse kernel of call k+1 runs beside the rebuild / synthesis kernels of call k."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    cfg = sys.argv[2] if len(sys.argv) > 2 else "2"
    host = len(sys.argv) > 3 and sys.argv[3] == "host"          # the pipelined host-buffer entry instead
    wl = bench.make_workload(cfg, 0, 40, False)
    wl.warmup(5)
    lib = wl.lib
    if host:
        wl.e2e_pipelined(4)
    for _ in range(3):
        wl.step_device(wl.t, own_stream=False)
        wl.t += 1
    lib.opus_b200_synchronize()
    lib.opus_b200_trace_begin()
    if host:
        import time
        t0 = time.perf_counter()
        wl.e2e_pipelined(steps)
        print("host wall per step (incl. warm-up calls): %.3f ms" % (1e3 * (time.perf_counter() - t0) / (steps + 4)))
    for _ in range(0 if host else steps):
        wl.step_device(wl.t, own_stream=False)
        wl.t += 1
    cap = 256
    kinds = (ctypes.c_int * cap)()
    t0 = (ctypes.c_double * cap)()
    t1 = (ctypes.c_double * cap)()
    n = lib.opus_b200_trace_end(kinds, t0, t1, cap)
    names = ["plan", "parse", "rebuild", "synth"]
    for i in range(n):
        print("%-8s %8.3f .. %8.3f ms  (%.3f)" % (names[kinds[i]], t0[i], t1[i], t1[i] - t0[i]))
    wl.close()


if __name__ == "__main__":
    main()
