"""Threads of ONE process driving the library at the same time (what a Concentus / P-Invoke host does): the
library's lock is per GPU and is not held while a call waits for its results, so (a) two threads with their own
decoder groups on one GPU and (b) one thread per GPU must both produce exactly the single-threaded output.
Checked against the golden PCM CRCs of the reference decoder."""
import ctypes
import threading

import numpy as np
import pytest

import concentus_b200 as cb
from conftest import crc32, golden_case

pytestmark = pytest.mark.gpu


def _worker(device, name, n, frames, rot, out, errs):
    try:
        import torch
        torch.cuda.set_device(device)
        meta, pk = golden_case(name)
        ch, fs = meta["channels"], meta["fs"]
        bd = cb.OpusBatchDecoder(fs, ch, n)
        try:
            crcs = []
            for t in range(frames):
                ret, pcm = bd.decode([pk[t][0]] * n, 960)
                assert (ret == ret[0]).all()
                assert all(np.array_equal(pcm[0], pcm[i]) for i in range(1, n, 7))
                crcs.append(crc32(pcm[0][:ret[0]]))
                assert (bd.final_ranges() == pk[t][1]).all()
            out[(device, name, rot)] = (crcs, meta["pcm_crc32"][:frames])
        finally:
            bd.close()
    except Exception as e:                     # surfaced by the main thread
        errs.append((device, name, repr(e)))


def _run(jobs):
    out, errs = {}, []
    th = [threading.Thread(target=_worker, args=j + (out, errs)) for j in jobs]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=300)
    assert not errs, errs
    assert len(out) == len(jobs)
    for key, (got, want) in out.items():
        assert got == want, "PCM differs from the reference for %r" % (key,)


def test_two_threads_one_gpu():
    _run([(0, "c2_96k_cbr_st", 64, 24, 0), (0, "vbr_128k_st", 48, 24, 1), (0, "mono_64k", 40, 24, 2)])


def test_one_thread_per_gpu():
    import torch
    ndev = torch.cuda.device_count()
    if ndev < 2:
        pytest.skip("needs two or more GPUs")
    names = ["c2_96k_cbr_st", "vbr_128k_st"]
    _run([(d, names[d % 2], 96, 24, d) for d in range(min(ndev, 8))])


def test_encoders_follow_their_gpu():
    """An encoder group created on GPU 1 and called while GPU 0 is current must still run on GPU 1."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two or more GPUs")
    import enc_cases
    from conftest import GOLDEN, read_bit
    import os
    case = {c.name: c for c in enc_cases.golden_cases()}["cbr_2ch_960_mix"]
    gold = read_bit(os.path.join(GOLDEN, "enc_%s.bit" % case.name))
    torch.cuda.set_device(1)
    be = cb.OpusBatchEncoder(8, case.fs, case.ch, case.app)
    torch.cuda.set_device(0)
    try:
        for req, val in case.ctls:
            be.ctl(req, val)
        for t, frame in enumerate(case.frames()[:4]):
            data, lens = be.encode(np.tile(frame, (8, 1)), case.frame, case.max_bytes)
            for i in range(8):
                assert data[i, :lens[i]].tobytes() == gold[t][0]
    finally:
        be.close()
