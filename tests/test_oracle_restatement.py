"""Pins the plain-C oracle restatement (oracle/oracle_celt.c, oracle_opus.c): it must reproduce
the golden fixtures (reference encoder + decoder output) bit for bit, and -- where the compiled
reference is available -- follow it through losses, DTX, FEC requests, multi-frame packets,
corrupted packets, decoder gain and 5.1 multistream decode."""
import ctypes
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, RefDecoder, crc32, golden_case, golden_index, needs_ref, read_bit
from scenarios import fuzz_scenario, loss_scenario

CASES = sorted(golden_index().keys())


@pytest.fixture(scope="module")
def oracle():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True)
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_build", "liboracle.so"))
    lib.oracle_opus_create.restype = ctypes.c_void_p
    lib.oracle_opus_create.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.oracle_opus_destroy.argtypes = [ctypes.c_void_p]
    lib.oracle_opus_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    lib.oracle_opus_final_range.argtypes = [ctypes.c_void_p]
    lib.oracle_opus_final_range.restype = ctypes.c_uint32
    lib.oracle_opus_set_gain.argtypes = [ctypes.c_void_p, ctypes.c_int]
    lib.oracle_ms_create.restype = ctypes.c_void_p
    lib.oracle_ms_create.argtypes = [ctypes.c_int] * 4 + [ctypes.c_char_p]
    lib.oracle_ms_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    lib.oracle_ms_final_range.argtypes = [ctypes.c_void_p]
    lib.oracle_ms_final_range.restype = ctypes.c_uint32
    lib.oracle_ms_destroy.argtypes = [ctypes.c_void_p]
    return lib


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_golden(oracle, name):
    meta, pk = golden_case(name)
    st = oracle.oracle_opus_create(meta["fs"], meta["channels"])
    spf = meta["samples_per_frame"]
    try:
        for n, (p, rng) in enumerate(pk):
            pcm = np.zeros((spf, meta["channels"]), dtype=np.int16)
            r = oracle.oracle_opus_decode(st, p, len(p), pcm.ctypes.data, spf, 0)
            assert r == spf
            assert oracle.oracle_opus_final_range(st) == rng, "final range, packet %d" % n
            assert crc32(pcm) == meta["pcm_crc32"][n], "pcm, packet %d" % n
    finally:
        oracle.oracle_opus_destroy(st)


def test_oracle_multistream_matches_golden(oracle):
    meta = json.load(open(os.path.join(GOLDEN, "index_ms.json")))["ms51_256k"]
    pk = read_bit(os.path.join(GOLDEN, "ms51_256k.bit"))
    ms = oracle.oracle_ms_create(48000, 6, meta["streams"], meta["coupled"], bytes(meta["mapping"]))
    try:
        for n, (p, rng) in enumerate(pk):
            pcm = np.zeros((960, 6), dtype=np.int16)
            r = oracle.oracle_ms_decode(ms, p if p else None, len(p), pcm.ctypes.data, 960, 0)
            assert r == 960
            assert crc32(pcm) == meta["pcm_crc32"][n], "pcm, packet %d" % n
            assert oracle.oracle_ms_final_range(ms) == rng
    finally:
        oracle.oracle_ms_destroy(ms)


def _follow(oracle, meta, calls, gain=0):
    fs, ch = meta["fs"], meta["channels"]
    ref = RefDecoder(fs, ch)
    st = oracle.oracle_opus_create(fs, ch)
    if gain:
        assert RefDecoder.lib().opus_decoder_ctl(ctypes.c_void_p(ref.h), 4034, ctypes.c_int(gain)) == 0
        oracle.oracle_opus_set_gain(st, gain)
    try:
        for n, (data, fsz, fec) in enumerate(calls):
            ra, pa = ref.decode_raw(data, fsz, fec)
            pb = np.zeros((fsz, ch), dtype=np.int16)
            rb = oracle.oracle_opus_decode(st, data, len(data) if data else 0, pb.ctypes.data, fsz, fec)
            assert ra == rb, "return value, call %d" % n
            if ra > 0:
                assert np.array_equal(pa[:ra], pb[:ra]), "pcm, call %d" % n
            assert ref.final_range == oracle.oracle_opus_final_range(st), "final range, call %d" % n
    finally:
        oracle.oracle_opus_destroy(st)


@needs_ref
@pytest.mark.parametrize("name", ["c2_96k_cbr_st", "f2_5_96k_st", "mono_12k", "st_to_mono_out", "mono_to_16k_out", "low_24k_st"])
def test_oracle_follows_reference_through_losses(oracle, name):
    meta, calls = loss_scenario(name, seed=len(name) * 7 + 1)
    _follow(oracle, meta, calls)
    _follow(oracle, meta, calls[:60], gain=-2000)


@needs_ref
@pytest.mark.parametrize("name", ["c2_96k_cbr_st", "vbr_128k_st", "f5_64k_st", "mono_to_st_out"])
def test_oracle_follows_reference_through_damage(oracle, name):
    meta, calls = fuzz_scenario(name, seed=len(name) + 3)
    _follow(oracle, meta, calls)
