"""Host-logic check (lane-loop emulation of the kernel sources, CPU) of the Opus layer and the
concealment code against the compiled reference: losses, DTX, FEC requests, multi-frame packets,
corrupted and truncated packets.  Return value, PCM and final range must all be identical."""
import ctypes

import numpy as np
import pytest

from conftest import RefDecoder, needs_ref
from scenarios import fuzz_scenario, loss_scenario, random_payload_scenario

LOSS = ["c2_96k_cbr_st", "vbr_128k_st", "f10_64k_st", "f2_5_96k_st", "mono_12k", "low_24k_st",
        "st_to_mono_out", "mono_to_st_out", "st_to_24k_out", "mono_to_16k_out", "nb_20k_mono"]
FUZZ = ["c2_96k_cbr_st", "vbr_128k_st", "f5_64k_st", "f2_5_96k_st", "mono_64k", "low_24k_st",
        "st_to_24k_out", "mono_to_st_out"]


def _run(emu_lib, meta, calls):
    emu_lib.emu_opus_create.restype = ctypes.c_void_p
    emu_lib.emu_opus_create.argtypes = [ctypes.c_int, ctypes.c_int]
    emu_lib.emu_opus_destroy.argtypes = [ctypes.c_void_p]
    emu_lib.emu_opus_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p,
                                        ctypes.c_int, ctypes.c_int]
    emu_lib.emu_opus_final_range.argtypes = [ctypes.c_void_p]
    emu_lib.emu_opus_final_range.restype = ctypes.c_uint
    fs, ch = meta["fs"], meta["channels"]
    ref = RefDecoder(fs, ch)
    st = emu_lib.emu_opus_create(fs, ch)
    try:
        for n, (data, fsz, fec) in enumerate(calls):
            ra, pa = ref.decode_raw(data, fsz, fec)
            pb = np.zeros((fsz, ch), dtype=np.int16)
            rb = emu_lib.emu_opus_decode(st, data, len(data) if data else 0, pb.ctypes.data, fsz, fec)
            assert ra == rb, "return value, call %d" % n
            if ra > 0:
                assert np.array_equal(pa[:ra], pb[:ra]), "pcm, call %d" % n
            assert ref.final_range == emu_lib.emu_opus_final_range(st), "final range, call %d" % n
    finally:
        emu_lib.emu_opus_destroy(st)


@needs_ref
@pytest.mark.parametrize("name", LOSS)
def test_emu_loss_concealment(emu_lib, name):
    meta, calls = loss_scenario(name, seed=len(name) * 7 + 1)
    assert sum(1 for d, _, f in calls if d is None or len(d) <= 1 or f) > 20
    _run(emu_lib, meta, calls)


@needs_ref
@pytest.mark.parametrize("name", FUZZ)
def test_emu_multiframe_and_corruption(emu_lib, name):
    meta, calls = fuzz_scenario(name, seed=len(name) + 3)
    _run(emu_lib, meta, calls)


@needs_ref
@pytest.mark.parametrize("gain", [-3000, 256, 1537, 5000])
def test_emu_decode_gain(emu_lib, gain):
    """OPUS_SET_GAIN (opus_decoder.c:567-577, 867-876): Q8 dB gain applied to the int16 output."""
    from conftest import golden_case
    emu_lib.emu_opus_create.restype = ctypes.c_void_p
    emu_lib.emu_opus_create.argtypes = [ctypes.c_int, ctypes.c_int]
    emu_lib.emu_opus_set_gain.argtypes = [ctypes.c_void_p, ctypes.c_int]
    emu_lib.emu_opus_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p,
                                        ctypes.c_int, ctypes.c_int]
    meta, pk = golden_case("vbr_128k_st")
    ref = RefDecoder(48000, 2)
    assert RefDecoder.lib().opus_decoder_ctl(ctypes.c_void_p(ref.h), 4034, ctypes.c_int(gain)) == 0
    st = emu_lib.emu_opus_create(48000, 2)
    emu_lib.emu_opus_set_gain(st, gain)
    for n, (p, _) in enumerate(pk[:40]):
        data = None if n in (7, 8, 20) else p
        ra, pa = ref.decode_raw(data, 960)
        pb = np.zeros((960, 2), dtype=np.int16)
        rb = emu_lib.emu_opus_decode(st, data, len(data) if data else 0, pb.ctypes.data, 960, 0)
        assert ra == rb == 960 and np.array_equal(pa, pb), n


@needs_ref
@pytest.mark.parametrize("name,fsz", [("c2_96k_cbr_st", 40320), ("c2_96k_cbr_st", 46080), ("f10_64k_st", 23040)])
def test_emu_long_concealment(emu_lib, name, fsz):
    """One lost call that asks for most of a second: up to 48 concealment pieces, PCM offsets beyond 32767 samples
    (the plan's offsets are 32-bit).  libopus conceals any frame_size in pieces (opus_decoder.c:613-627)."""
    from conftest import golden_case
    meta, pk = golden_case(name)
    spf = meta["frame_size"] if "frame_size" in meta else None
    calls = []
    for t in range(6):
        d = pk[t][0]
        calls.append((d, 5760, 0))
    calls.append((None, fsz, 0))
    calls.append((pk[6][0], 5760, 0))
    _run(emu_lib, meta, calls)


def test_emu_max_passes_cap(emu_lib):
    """A device-resident call launches exactly max_frames passes: a stream that needs more (a two-frame packet, a
    loss that takes several concealment pieces) must fail with OPUS_BAD_ARG and leave its state alone."""
    from conftest import golden_case
    emu_lib.emu_opus_create.restype = ctypes.c_void_p
    emu_lib.emu_opus_create.argtypes = [ctypes.c_int, ctypes.c_int]
    emu_lib.emu_opus_destroy.argtypes = [ctypes.c_void_p]
    emu_lib.emu_opus_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p,
                                        ctypes.c_int, ctypes.c_int]
    emu_lib.emu_opus_decode_capped.argtypes = emu_lib.emu_opus_decode.argtypes + [ctypes.c_int]
    emu_lib.emu_opus_final_range.argtypes = [ctypes.c_void_p]
    emu_lib.emu_opus_final_range.restype = ctypes.c_uint
    emu_lib.emu_opus_last_packet_duration.argtypes = [ctypes.c_void_p]
    meta, pk = golden_case("f10_64k_st")          # 10 ms frames
    a, b = emu_lib.emu_opus_create(48000, 2), emu_lib.emu_opus_create(48000, 2)
    pa, pb = np.zeros((960, 2), np.int16), np.zeros((960, 2), np.int16)
    try:
        for t in range(4):
            for st, p in ((a, pa), (b, pb)):
                assert emu_lib.emu_opus_decode(st, pk[t][0], len(pk[t][0]), p.ctypes.data, 960, 0) == 480
        # a 20 ms loss after 10 ms frames needs two pieces: one pass is not enough
        assert emu_lib.emu_opus_decode_capped(a, None, 0, pa.ctypes.data, 960, 0, 1) == -1
        assert emu_lib.emu_opus_last_packet_duration(a) == 480
        # two frames in one packet (code 1), one pass allowed
        two = bytes([pk[4][0][0] | 1]) + pk[4][0][1:] + pk[4][0][1:]
        assert emu_lib.emu_opus_decode_capped(a, two, len(two), pa.ctypes.data, 960, 0, 1) == -1
        assert emu_lib.emu_opus_final_range(a) == emu_lib.emu_opus_final_range(b)
        # the failed calls left no trace: both decoders continue identically
        for t in range(4, 8):
            ra = emu_lib.emu_opus_decode_capped(a, pk[t][0], len(pk[t][0]), pa.ctypes.data, 960, 0, 1)
            rb = emu_lib.emu_opus_decode(b, pk[t][0], len(pk[t][0]), pb.ctypes.data, 960, 0)
            assert ra == rb == 480 and np.array_equal(pa, pb)
            assert emu_lib.emu_opus_final_range(a) == emu_lib.emu_opus_final_range(b)
    finally:
        emu_lib.emu_opus_destroy(a)
        emu_lib.emu_opus_destroy(b)


@needs_ref
@pytest.mark.parametrize("fs,ch,seed", [(48000, 2, 1), (48000, 1, 2), (24000, 2, 3), (16000, 1, 4)])
def test_emu_random_celt_payloads(emu_lib, fs, ch, seed):
    """Random payloads behind a CELT-only TOC (scenarios.random_payload_scenario): the tape the parse half writes and
    the rebuild half's reading of it must give the reference's PCM and final range for any bit pattern."""
    meta, calls = random_payload_scenario(fs, ch, seed)
    _run(emu_lib, meta, calls)
