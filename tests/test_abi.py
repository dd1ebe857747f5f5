"""The C-ABI library loads, exports every symbol include/opus_b200.h declares, and its host-only
helpers (packet inspection) agree with the reference.  No GPU compute here."""
import ctypes
import os
import re

import numpy as np
import pytest

import concentus_b200 as cb
from conftest import ROOT, RefDecoder, cuda_available, golden_case, needs_ref


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "opus_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(opus_[a-z0-9_]+)\s*\(", hdr))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    lib = cb.load()
    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), name
    assert set(declared) == set(cb.EXPORTS)


def test_version_string_says_fixed():
    assert b"-fixed" in cb.load().opus_get_version_string()      # celt/celt.c:289-294


def test_strerror():
    lib = cb.load()
    assert lib.opus_strerror(0) == b"success"
    assert lib.opus_strerror(-4) == b"corrupted stream"
    assert lib.opus_strerror(1) == b"unknown error"


def test_create_rejects_bad_arguments():
    lib = cb.load()
    err = ctypes.c_int(0)
    for fs, ch in [(44100, 2), (48000, 3), (48000, 0), (0, 1)]:
        h = lib.opus_decoder_create(fs, ch, ctypes.byref(err))
        assert not h and err.value == cb.OPUS_BAD_ARG


@pytest.mark.skipif(cuda_available(), reason="only meaningful on a machine without a GPU")
def test_no_cpu_fallback_without_gpu():
    lib = cb.load()
    err = ctypes.c_int(0)
    h = lib.opus_decoder_create(48000, 2, ctypes.byref(err))
    assert not h and err.value < 0       # fails loudly, never decodes on the CPU


@needs_ref
def test_packet_helpers_match_reference():
    lib = cb.load()
    ref = RefDecoder.lib()
    rng = np.random.RandomState(1234)
    _, pk = golden_case("vbr_128k_st")
    packets = [p for p, _ in pk[:40]]
    for _ in range(3000):                      # random heads in the style of test_opus_decode.c:214-359
        ln = int(rng.randint(1, 64))
        packets.append(bytes(rng.randint(0, 256, size=ln, dtype=np.uint8).tolist()))
    for p in packets:
        buf = ctypes.create_string_buffer(p, len(p))
        ptr = ctypes.cast(buf, ctypes.c_void_p)
        assert lib.opus_packet_get_bandwidth(ptr) == ref.opus_packet_get_bandwidth(buf)
        assert lib.opus_packet_get_nb_channels(ptr) == ref.opus_packet_get_nb_channels(buf)
        assert lib.opus_packet_get_nb_frames(ptr, len(p)) == ref.opus_packet_get_nb_frames(buf, len(p))
        for fs in (48000, 16000, 8000):
            assert lib.opus_packet_get_samples_per_frame(ptr, fs) == ref.opus_packet_get_samples_per_frame(buf, fs)
            assert lib.opus_packet_get_nb_samples(ptr, len(p), fs) == ref.opus_packet_get_nb_samples(buf, len(p), fs)
        sz_a = (ctypes.c_int16 * 48)()
        sz_b = (ctypes.c_int16 * 48)()
        toc_a, toc_b = ctypes.c_ubyte(0), ctypes.c_ubyte(0)
        off_a, off_b = ctypes.c_int(0), ctypes.c_int(0)
        ca = lib.opus_packet_parse(ptr, len(p), ctypes.byref(toc_a), None, sz_a, ctypes.byref(off_a))
        cr = ref.opus_packet_parse(buf, len(p), ctypes.byref(toc_b), None, sz_b, ctypes.byref(off_b))
        assert ca == cr
        if ca > 0:
            assert toc_a.value == toc_b.value and off_a.value == off_b.value
            assert list(sz_a[:ca]) == list(sz_b[:ca])
