import ctypes
import json
import os
import struct
import subprocess
import sys
import zlib

import numpy as np
import pytest

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libopus_ref.so")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA GPU (run on the B200 box)")


def read_bit(path):
    """opus_demo .bit container: BE32 length, BE32 encoder final range, payload."""
    d = open(path, "rb").read()
    o, out = 0, []
    while o < len(d):
        ln, rng = struct.unpack(">II", d[o:o + 8])
        o += 8
        out.append((d[o:o + ln], rng))
        o += ln
    return out


def golden_index():
    return json.load(open(os.path.join(GOLDEN, "index.json")))


def golden_case(name):
    meta = golden_index()[name]
    return meta, read_bit(os.path.join(GOLDEN, name + ".bit"))


def crc32(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


class RefDecoder(object):
    """The UNMODIFIED reference decoder (oracle/_ref/libopus_ref.so, built from /root/reference)."""

    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            lib = ctypes.CDLL(REF_SO)
            lib.opus_decoder_create.restype = ctypes.c_void_p
            lib.opus_decoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
            lib.opus_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int32, ctypes.c_void_p,
                                        ctypes.c_int, ctypes.c_int]
            lib.opus_decoder_destroy.argtypes = [ctypes.c_void_p]
            lib.opus_decoder_destroy.restype = None
            cls._lib = lib
        return cls._lib

    def __init__(self, fs, ch):
        err = ctypes.c_int(0)
        self.h = self.lib().opus_decoder_create(fs, ch, ctypes.byref(err))
        assert self.h and err.value == 0
        self.ch = ch

    def decode_raw(self, data, frame_size, fec=0):
        pcm = np.zeros((max(frame_size, 1), self.ch), dtype=np.int16)
        ret = self.lib().opus_decode(self.h, data, len(data) if data else 0, pcm.ctypes.data, frame_size, fec)
        return ret, pcm

    @property
    def final_range(self):
        v = ctypes.c_uint32(0)
        self.lib().opus_decoder_ctl(ctypes.c_void_p(self.h), 4031, ctypes.byref(v))
        return v.value

    def close(self):
        if self.h:
            self.lib().opus_decoder_destroy(self.h)
            self.h = None

    __del__ = close


def have_ref():
    return os.path.exists(REF_SO)


needs_ref = pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (needs /root/reference; run make -C oracle ref)")


@pytest.fixture(scope="session")
def emu_lib():
    """Lane-loop emulation of the kernel sources (tests/emu), compiled with g++ on demand."""
    src = os.path.join(ROOT, "tests", "emu", "emu_celt.cpp")
    so = os.path.join(ROOT, "tests", "emu", "libemu_celt.so")
    csrc = os.path.join(ROOT, "concentus_b200", "csrc")
    newest = max([os.path.getmtime(src)] + [os.path.getmtime(os.path.join(csrc, f)) for f in os.listdir(csrc)])
    if not os.path.exists(so) or os.path.getmtime(so) < newest:
        subprocess.run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-Wall", "-Wno-unused-function", "-o", so, src],
                       check=True)
    lib = ctypes.CDLL(so)
    lib.emu_celt_create.restype = ctypes.c_void_p
    lib.emu_celt_create.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.emu_celt_destroy.argtypes = [ctypes.c_void_p]
    lib.emu_celt_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    return lib


class RefEncoder(object):
    """The UNMODIFIED reference encoder (oracle/_ref/libopus_ref.so)."""

    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            lib = ctypes.CDLL(REF_SO)
            lib.opus_encoder_create.restype = ctypes.c_void_p
            lib.opus_encoder_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
            lib.opus_encode.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int32]
            lib.opus_encoder_destroy.argtypes = [ctypes.c_void_p]
            lib.opus_encoder_destroy.restype = None
            cls._lib = lib
        return cls._lib

    def __init__(self, fs, ch, app):
        err = ctypes.c_int(0)
        self.h = self.lib().opus_encoder_create(fs, ch, app, ctypes.byref(err))
        assert self.h and err.value == 0

    def ctl(self, req, val):
        return self.lib().opus_encoder_ctl(ctypes.c_void_p(self.h), ctypes.c_int(req), ctypes.c_int(val))

    def encode(self, pcm, frame_size, max_bytes):
        """Returns (return code, packet bytes, final range)."""
        pcm = np.ascontiguousarray(pcm, dtype=np.int16)
        buf = (ctypes.c_ubyte * max(max_bytes, 1))()
        n = self.lib().opus_encode(self.h, pcm.ctypes.data, frame_size, buf, max_bytes)
        v = ctypes.c_uint32(0)
        self.lib().opus_encoder_ctl(ctypes.c_void_p(self.h), 4031, ctypes.byref(v))
        return n, bytes(buf[:max(n, 0)]), v.value

    def close(self):
        if self.h:
            self.lib().opus_encoder_destroy(self.h)
            self.h = None

    __del__ = close


class EmuEncoder(object):
    """Host-emulation build of the encoder kernel source (tests/emu/emu_enc.cpp)."""

    def __init__(self, lib, fs, ch, app):
        self.lib = lib
        self.h = lib.emu_enc_create(fs, ch, app)
        assert self.h

    def ctl(self, req, val):
        return self.lib.emu_enc_ctl(self.h, req, val, None)

    def encode(self, pcm, frame_size, max_bytes):
        pcm = np.ascontiguousarray(pcm, dtype=np.int16)
        buf = (ctypes.c_ubyte * max(max_bytes, 1))()
        n = self.lib.emu_enc_encode(self.h, pcm.ctypes.data, frame_size, buf, max_bytes)
        return n, bytes(buf[:max(n, 0)]), self.lib.emu_enc_final_range(self.h)

    def close(self):
        if self.h:
            self.lib.emu_enc_destroy(self.h)
            self.h = None

    __del__ = close


@pytest.fixture(scope="session")
def emu_enc_lib():
    """Host build of the encoder headers (tests/emu/emu_enc.cpp), compiled with g++ on demand."""
    src = os.path.join(ROOT, "tests", "emu", "emu_enc.cpp")
    so = os.path.join(ROOT, "tests", "emu", "libemu_enc.so")
    csrc = os.path.join(ROOT, "concentus_b200", "csrc")
    newest = max([os.path.getmtime(src)] + [os.path.getmtime(os.path.join(csrc, f)) for f in os.listdir(csrc)])
    if not os.path.exists(so) or os.path.getmtime(so) < newest:
        subprocess.run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-Wall", "-Wno-unused-function", "-o", so, src],
                       check=True)
    lib = ctypes.CDLL(so)
    lib.emu_enc_create.restype = ctypes.c_void_p
    lib.emu_enc_create.argtypes = [ctypes.c_int] * 3
    lib.emu_enc_destroy.argtypes = [ctypes.c_void_p]
    lib.emu_enc_encode.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    lib.emu_enc_ctl.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    lib.emu_enc_final_range.argtypes = [ctypes.c_void_p]
    lib.emu_enc_final_range.restype = ctypes.c_uint
    return lib


def silk_golden_index():
    """SILK-only fixtures plus the hybrid / mode-transition fixtures (same replay format)."""
    idx = json.load(open(os.path.join(GOLDEN, "index_silk.json")))
    idx.update(json.load(open(os.path.join(GOLDEN, "index_hybrid.json"))))
    return idx


class EmuOpusDecoder(object):
    """Host-emulation build of the whole decode path (plan + CELT passes + SILK pass), one stream."""

    def __init__(self, lib, fs, ch):
        lib.emu_opus_create.restype = ctypes.c_void_p
        lib.emu_opus_create.argtypes = [ctypes.c_int, ctypes.c_int]
        lib.emu_opus_destroy.argtypes = [ctypes.c_void_p]
        lib.emu_opus_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                        ctypes.c_int]
        lib.emu_opus_final_range.argtypes = [ctypes.c_void_p]
        lib.emu_opus_final_range.restype = ctypes.c_uint
        self.lib, self.ch = lib, ch
        self.h = lib.emu_opus_create(fs, ch)

    def decode_raw(self, data, frame_size, fec=0):
        pcm = np.zeros((max(frame_size, 1), self.ch), dtype=np.int16)
        ret = self.lib.emu_opus_decode(self.h, data, len(data) if data else 0, pcm.ctypes.data, frame_size, int(fec))
        return ret, pcm

    @property
    def final_range(self):
        return self.lib.emu_opus_final_range(self.h)

    def close(self):
        if self.h:
            self.lib.emu_opus_destroy(self.h)
            self.h = None

    __del__ = close


def enc_golden_index():
    return json.load(open(os.path.join(GOLDEN, "enc_index.json")))


def cuda_available():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False
