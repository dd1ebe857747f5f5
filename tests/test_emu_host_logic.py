"""CPU-side check of the kernel SOURCES' logic: csrc/*.h compiled as plain C++ (lane loops become
loops, tests/emu) must reproduce the reference decoder's PCM and final range on every golden
fixture.  This is a host-logic test, not the product path (the product is the CUDA build)."""
import ctypes

import numpy as np
import pytest

from conftest import crc32, golden_case, golden_index

CASES = sorted(golden_index().keys())


@pytest.mark.parametrize("name", CASES)
def test_emu_matches_golden(emu_lib, name):
    meta, pk = golden_case(name)
    ch, fs = meta["channels"], meta["fs"]
    st = emu_lib.emu_celt_create(ch, 48000 // fs)
    spf = meta["samples_per_frame"]
    try:
        for n, (p, rng) in enumerate(pk):
            toc = p[0]
            assert toc & 0x80 and (toc & 3) == 0
            cfg = toc >> 3
            LM = cfg & 3
            end = [13, 17, 19, 21][(cfg - 16) >> 2]
            C = 2 if toc & 4 else 1
            pcm = np.zeros((spf, ch), dtype=np.int16)
            fr = ctypes.c_uint(0)
            r = emu_lib.emu_celt_decode(st, p[1:], len(p) - 1, LM, C, 0, end, pcm.ctypes.data, ctypes.byref(fr))
            assert r == spf
            assert fr.value == rng, "final range, packet %d" % n
            assert crc32(pcm) == meta["pcm_crc32"][n], "pcm, packet %d" % n
    finally:
        emu_lib.emu_celt_destroy(st)
