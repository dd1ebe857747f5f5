"""5.1 multistream decode (BASELINE config 5): 6 channels = 4 elementary streams (2 coupled),
mapping {0,4,1,2,3,5}, packets from the reference's surround encoder, 5 % of the packets lost.
CPU: lane-loop emulation vs the golden fixture; GPU: C-ABI (single, batch, 1024-stream group)."""
import ctypes
import json
import os

import numpy as np
import pytest

import concentus_b200 as cb
from conftest import GOLDEN, crc32, have_ref, needs_ref, read_bit, REF_SO


# ms51_256k: config 5 (CELT on every stream); ms51_80k_voip: hybrid + CELT elementary streams side by side;
# ms51_40k_voip: SILK-only + CELT elementary streams (the surround encoder's own mode decisions at low rates)
MS_CASES = ["ms51_256k", "ms51_80k_voip", "ms51_40k_voip"]


def ms_case(name="ms51_256k"):
    meta = json.load(open(os.path.join(GOLDEN, "index_ms.json")))[name]
    return meta, read_bit(os.path.join(GOLDEN, name + ".bit"))


@pytest.mark.parametrize("case", MS_CASES)
def test_emu_multistream_matches_golden(emu_lib, case):
    meta, pk = ms_case(case)
    emu_lib.emu_ms_create.restype = ctypes.c_void_p
    emu_lib.emu_ms_create.argtypes = [ctypes.c_int] * 4 + [ctypes.c_char_p]
    emu_lib.emu_ms_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    emu_lib.emu_ms_final_range.argtypes = [ctypes.c_void_p]
    emu_lib.emu_ms_final_range.restype = ctypes.c_uint
    emu_lib.emu_ms_destroy.argtypes = [ctypes.c_void_p]
    ms = emu_lib.emu_ms_create(48000, 6, meta["streams"], meta["coupled"], bytes(meta["mapping"]))
    assert len(meta["lost"]) >= 5
    try:
        for n, (p, rng) in enumerate(pk):
            pcm = np.zeros((960, 6), dtype=np.int16)
            r = emu_lib.emu_ms_decode(ms, p if p else None, len(p), pcm.ctypes.data, 960, 0)
            assert r == 960, (n, r)
            assert crc32(pcm) == meta["pcm_crc32"][n], "pcm of packet %d" % n
            assert emu_lib.emu_ms_final_range(ms) == rng, "final range of packet %d" % n
    finally:
        emu_lib.emu_ms_destroy(ms)


@pytest.mark.gpu
@pytest.mark.parametrize("case", MS_CASES)
def test_gpu_multistream_single_matches_golden(case):
    meta, pk = ms_case(case)
    dec = cb.OpusMSDecoder(48000, 6, meta["streams"], meta["coupled"], meta["mapping"])
    try:
        for n, (p, rng) in enumerate(pk[:100]):
            r, pcm = dec.decode_raw(p, 960)
            assert r == 960, (n, r)
            assert crc32(pcm) == meta["pcm_crc32"][n], "pcm of packet %d" % n
            assert dec.final_range == rng
    finally:
        dec.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", MS_CASES)
def test_gpu_multistream_1024_streams(case):
    """Config 5 at full size: 1024 six-channel decoders in one group; all play the fixture (same
    loss pattern), every decoder must reproduce the golden PCM of every packet."""
    meta, pk = ms_case(case)
    n = 1024
    bd = cb.OpusMSBatchDecoder(48000, 6, meta["streams"], meta["coupled"], meta["mapping"], n)
    try:
        for t, (p, rng) in enumerate(pk[:24]):
            ret, pcm = bd.decode([p] * n, 960)
            assert (ret == 960).all()
            assert crc32(pcm[0]) == meta["pcm_crc32"][t], "packet %d" % t
            assert (pcm == pcm[0]).all(), "streams differ at packet %d" % t
    finally:
        bd.close()


@pytest.mark.gpu
def test_gpu_multistream_errors():
    meta, pk = ms_case()
    dec = cb.OpusMSDecoder(48000, 6, meta["streams"], meta["coupled"], meta["mapping"])
    try:
        good = pk[0][0]
        r, _ = dec.decode_raw(good[:5], 960)           # shorter than 2*streams-1 bytes
        assert r == cb.OPUS_INVALID_PACKET
        r, _ = dec.decode_raw(good, 480)               # packet holds 960 samples
        assert r == cb.OPUS_BUFFER_TOO_SMALL
        r, _ = dec.decode_raw(good[:40], 960)          # truncated: a sub-packet runs past the end
        assert r < 0
        r, pcm = dec.decode_raw(good, 960)             # failed calls left the state untouched
        assert r == 960 and crc32(pcm) == meta["pcm_crc32"][0]
    finally:
        dec.close()
    lib = cb.load()
    err = ctypes.c_int(0)
    bad = (ctypes.c_ubyte * 6)(0, 4, 1, 2, 3, 9)       # mapping refers to a channel that does not exist
    assert not lib.opus_multistream_decoder_create(48000, 6, 4, 2, bad, ctypes.byref(err)) and err.value == cb.OPUS_BAD_ARG


@pytest.mark.gpu
def test_gpu_multistream_pipelined_submit_matches_golden():
    """opus_multistream_decode_group_submit/_wait: 32 decoders replay the 5.1 fixture (with its lost packets)
    from staggered offsets, ten calls queued back to back; decoder 0 must reproduce the golden CRCs."""
    import torch
    meta, pk = ms_case("ms51_256k")
    lib = cb.load()
    n, steps = 32, 10
    bd = cb.OpusMSBatchDecoder(48000, 6, meta["streams"], meta["coupled"], meta["mapping"], n)
    maxlen = max(len(p) for p, _ in pk)
    try:
        bufs = []
        for t in range(steps):
            arena = torch.zeros(n * maxlen, dtype=torch.uint8).pin_memory()
            offs = torch.zeros(n, dtype=torch.int32).pin_memory()
            lens = torch.zeros(n, dtype=torch.int32).pin_memory()
            for i in range(n):
                p = pk[(t + 5 * i) % len(pk)][0] if i else pk[t][0]
                lens[i] = len(p)
                offs[i] = i * maxlen if p else -1
                if p:
                    arena[i * maxlen:i * maxlen + len(p)] = torch.frombuffer(bytearray(p), dtype=torch.uint8)
            pcm = torch.zeros(n * 960 * 6, dtype=torch.int16).pin_memory()
            ret = torch.zeros(n, dtype=torch.int32).pin_memory()
            bufs.append((arena, offs, lens, pcm, ret))
            assert lib.opus_multistream_decode_group_submit(bd._g, arena.data_ptr(), arena.numel(), offs.data_ptr(),
                                                            lens.data_ptr(), pcm.data_ptr(), 960 * 6, 960, None,
                                                            ret.data_ptr()) == 0
        assert lib.opus_multistream_decode_group_wait(bd._g) == 0
        for t, (_, _, _, pcm, ret) in enumerate(bufs):
            assert bool((ret == 960).all()), t
            assert crc32(pcm.numpy().reshape(n, 960, 6)[0]) == meta["pcm_crc32"][t], t
    finally:
        bd.close()
