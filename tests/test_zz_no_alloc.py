"""opus_decode_batch_device_static(): the batch decode for callers that may not allocate GPU
memory (a real-time decode thread that cannot afford cudaMalloc/cudaFree stalling the device, or a
host that pre-carved all GPU memory).  Its scratch is a fixed array inside the library, so this test
holds it to the ordinary group entry point (opus_decode_group_device, whose scratch is allocated
when the group is created): same packets, two calls in a row -- the second call finds the first
call's records still in the shared array -- and the PCM and return codes must be identical
(bit-exact, the tolerance every decode test here uses).  Collected last on purpose (file name), and
it also passes when run on its own in a fresh process.
"""
import ctypes

import numpy as np
import pytest

import concentus_b200 as cb
from conftest import crc32, golden_case

pytestmark = pytest.mark.gpu


def test_static_scratch_decode_matches_group_decode():
    import torch
    lib = cb.load()
    meta, pk = golden_case("c2_96k_cbr_st")
    n, spf, ch = 96, 960, 2
    stream = torch.cuda.Stream()
    grp = cb.OpusBatchDecoder(48000, ch, n)                       # reference path: group scratch
    solo = [cb.OpusDecoder(48000, ch) for _ in range(n)]          # new path: decoders only, no group
    handles = (ctypes.c_void_p * n)(*[d._h for d in solo])
    try:
        for t in range(2):                                        # two calls in a row
            packets = [pk[(11 * i + t) % len(pk)][0] for i in range(n)]
            plen = len(packets[0])
            arena = torch.from_numpy(np.frombuffer(b"".join(packets), dtype=np.uint8).copy()).cuda()
            offs = torch.arange(n, dtype=torch.int32, device="cuda") * plen
            lens = torch.full((n,), plen, dtype=torch.int32, device="cuda")
            pcm_a = torch.zeros(n * spf * ch, dtype=torch.int16, device="cuda")
            pcm_b = torch.zeros(n * spf * ch, dtype=torch.int16, device="cuda")
            ret_a = torch.zeros(n, dtype=torch.int32, device="cuda")
            ret_b = torch.zeros(n, dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            sp = ctypes.c_void_p(stream.cuda_stream)
            rc = lib.opus_decode_group_device(grp._g, arena.data_ptr(), offs.data_ptr(), lens.data_ptr(),
                                              pcm_a.data_ptr(), spf * ch, spf, None, ret_a.data_ptr(), 1, plen, sp)
            assert rc == 0
            rc = lib.opus_decode_batch_device_static(handles, n, arena.data_ptr(), offs.data_ptr(), lens.data_ptr(),
                                                     pcm_b.data_ptr(), spf * ch, spf, ret_b.data_ptr(), 1, plen, sp)
            assert rc == 0
            stream.synchronize()
            assert (ret_a.cpu().numpy() == spf).all()
            assert torch.equal(ret_a, ret_b)
            assert torch.equal(pcm_a, pcm_b), "static-scratch decode differs from group decode (call %d)" % t
            # and both equal the reference's output for stream 0
            first = pcm_b[:spf * ch].cpu().numpy().reshape(spf, ch)
            if t == 0:
                assert crc32(first) == meta["pcm_crc32"][0]
    finally:
        grp.close()
        for d in solo:
            d.close()


def test_static_scratch_rejects_oversized_batches():
    import torch
    lib = cb.load()
    n = 257                                                       # OPUS_B200_STATIC_MAX_STREAMS + 1
    decs = [cb.OpusDecoder(48000, 1) for _ in range(n)]
    handles = (ctypes.c_void_p * n)(*[d._h for d in decs])
    dummy = torch.zeros(16, dtype=torch.int32, device="cuda")
    try:
        rc = lib.opus_decode_batch_device_static(handles, n, dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(),
                                                 dummy.data_ptr(), 960, 960, dummy.data_ptr(), 1, 240, None)
        assert rc == cb.OPUS_BAD_ARG
    finally:
        for d in decs:
            d.close()
