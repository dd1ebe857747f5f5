"""Do the golden fixtures really take every branch of the band rebuild?

The rebuild half of the CELT decoder (csrc/cb_celt_bands3.h) works from the records the parse half leaves on the
tape: leaves with and without pulses, masked / zero / noise / folded fills, the stereo modes, the block-layout
changes.  A parity test only proves what its inputs exercise, so the host-emulation build counts the branches
(CB_COV) and this test decodes the CELT fixtures -- each packet checked against the reference's PCM and final range,
as test_emu_host_logic does -- and then requires every counter to be non-zero.
"""
import ctypes

import numpy as np

from conftest import crc32, golden_case, golden_index

NAMES = ["EMPTY_IN_TREE", "EMPTY_WHOLE_CALL", "MASKED", "ZERO", "NOISE", "FOLD", "SEGMENTS3", "SWITCH", "JOINT2",
         "N1_STEREO", "N1_MONO", "SIDE_FIRST", "DUAL", "RECOMBINE", "TIME_DIVIDE", "SCRATCH", "INV"]


def _read(lib):
    buf = (ctypes.c_long * 64)()
    lib.emu_cov_read.argtypes = [ctypes.c_void_p, ctypes.c_int]
    n = lib.emu_cov_read(buf, 64)
    assert n == len(NAMES), "tests/test_emu_coverage.py NAMES is out of step with cb_celt_bands3.h"
    return dict(zip(NAMES, list(buf)[:n]))


def test_fixtures_cover_the_rebuild_paths(emu_lib):
    _read(emu_lib)                                   # clear
    total = dict.fromkeys(NAMES, 0)
    per_case = {}
    for name in sorted(golden_index().keys()):
        meta, pk = golden_case(name)
        ch, fs = meta["channels"], meta["fs"]
        st = emu_lib.emu_celt_create(ch, 48000 // fs)
        spf = meta["samples_per_frame"]
        try:
            for n, (p, rng) in enumerate(pk):
                toc = p[0]
                cfg = toc >> 3
                LM = cfg & 3
                end = [13, 17, 19, 21][(cfg - 16) >> 2]
                C = 2 if toc & 4 else 1
                pcm = np.zeros((spf, ch), dtype=np.int16)
                fr = ctypes.c_uint(0)
                r = emu_lib.emu_celt_decode(st, p[1:], len(p) - 1, LM, C, 0, end, pcm.ctypes.data, ctypes.byref(fr))
                assert r == spf
                assert fr.value == rng and crc32(pcm) == meta["pcm_crc32"][n], (name, n)
        finally:
            emu_lib.emu_celt_destroy(st)
        per_case[name] = _read(emu_lib)
        for k, v in per_case[name].items():
            total[k] += v
    missing = [k for k, v in total.items() if v == 0]
    assert not missing, "no golden fixture takes these rebuild branches: %s\n%s" % (missing, total)
