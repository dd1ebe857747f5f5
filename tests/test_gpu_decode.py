"""Parity tests proper: the CUDA library, called through its C ABI, against the golden fixtures
(reference output) and against the compiled reference itself (oracle/_ref) on the same packets.
Bar: bit-exact PCM and equal OPUS_GET_FINAL_RANGE per packet (SURVEY 8c "Parity definition")."""
import ctypes

import numpy as np
import pytest

import concentus_b200 as cb
from conftest import RefDecoder, crc32, golden_case, golden_index, have_ref, needs_ref

pytestmark = pytest.mark.gpu

CASES = sorted(golden_index().keys())


@pytest.mark.parametrize("name", CASES)
def test_single_stream_matches_golden(name):
    """opus_decoder_create / opus_decode / OPUS_GET_FINAL_RANGE, packet by packet (config 1 style)."""
    meta, pk = golden_case(name)
    dec = cb.OpusDecoder(meta["fs"], meta["channels"])
    spf = meta["samples_per_frame"]
    try:
        for n, (p, rng) in enumerate(pk[:120]):
            ret, pcm = dec.decode_raw(p, spf)
            assert ret == spf, (n, ret)
            assert crc32(pcm[:ret]) == meta["pcm_crc32"][n], "pcm of packet %d" % n
            if n % 8 == 0:
                assert dec.final_range == rng, "final range of packet %d" % n
        assert dec.last_packet_duration == spf
    finally:
        dec.close()


def test_heterogeneous_batch_matches_golden():
    """One opus_decode_batch call carries 17 streams of different frame size, bandwidth, channel
    count and output rate; every stream must equal its own golden, for all 250 packets."""
    lib = cb.load()
    metas, pks, decs = [], [], []
    for name in CASES:
        meta, pk = golden_case(name)
        metas.append(meta)
        pks.append(pk)
        decs.append(cb.OpusDecoder(meta["fs"], meta["channels"]))
    n = len(decs)
    handles = (ctypes.c_void_p * n)(*[d._h for d in decs])
    frame_size = 960
    try:
        for t in range(250):
            bufs = [ctypes.create_string_buffer(pks[i][t][0], len(pks[i][t][0])) for i in range(n)]
            data = (ctypes.c_void_p * n)(*[ctypes.cast(b, ctypes.c_void_p) for b in bufs])
            lens = np.array([len(pks[i][t][0]) for i in range(n)], dtype=np.int32)
            outs = [np.zeros((frame_size, metas[i]["channels"]), dtype=np.int16) for i in range(n)]
            pcm = (ctypes.c_void_p * n)(*[o.ctypes.data for o in outs])
            ret = np.zeros(n, dtype=np.int32)
            rc = lib.opus_decode_batch(handles, data, lens.ctypes.data, pcm, frame_size, None, n, ret.ctypes.data)
            assert rc == 0
            for i in range(n):
                spf = metas[i]["samples_per_frame"]
                assert ret[i] == spf, (CASES[i], t, ret[i])
                assert crc32(outs[i][:spf]) == metas[i]["pcm_crc32"][t], (CASES[i], t)
        fr = np.zeros(n, dtype=np.uint32)
        assert lib.opus_decoder_get_final_range_batch(handles, n, fr.ctypes.data) == 0
        for i in range(n):
            assert int(fr[i]) == pks[i][249][1], CASES[i]
    finally:
        for d in decs:
            d.close()


@needs_ref
@pytest.mark.parametrize("n", [4096, 700, 2500, 4200])
def test_4096_streams_full_size(n):
    """BASELINE config 2 at full size: 4096 independent 48 kHz stereo 96 kbps streams on one GPU (and three other
    batch sizes: the entropy kernel picks its CTA shape from the batch -- 4..28 warps per CTA, one or two waves).
    Stream i plays the fixture's packets starting at packet (37*i mod 250), so streams differ.
    Checks: (1) every stream's final range equals the range the reference ENCODER stored;
    (2) streams with the same rotation are identical (checksum of checksums);
    (3) one stream per rotation is compared sample by sample with the compiled reference."""
    meta, pk = golden_case("c2_96k_cbr_st")
    frames, spf = 6 if n == 4096 else 3, 960
    bd = cb.OpusBatchDecoder(48000, 2, n)
    rot = [(37 * i) % 250 for i in range(n)]
    refs = {}
    try:
        for t in range(frames):
            packets = [pk[(rot[i] + t) % 250][0] for i in range(n)]
            ret, pcm = bd.decode(packets, spf)
            assert (ret == spf).all()
            fr = bd.final_ranges()
            want = np.array([pk[(rot[i] + t) % 250][1] for i in range(n)], dtype=np.uint32)
            assert (fr == want).all(), "final ranges, frame %d" % t
            first = {}
            for i in range(n):
                r = rot[i]
                if r not in first:
                    first[r] = i
                    if r not in refs:
                        refs[r] = RefDecoder(48000, 2)
                    rr, rp = refs[r].decode_raw(packets[i], spf)
                    assert rr == spf
                    assert np.array_equal(rp, pcm[i]), "stream %d frame %d differs from the reference" % (i, t)
                else:
                    assert np.array_equal(pcm[i], pcm[first[r]]), "stream %d != stream %d" % (i, first[r])
    finally:
        bd.close()


def test_errors_do_not_poison_neighbours():
    """Per-stream return codes: bad packets / bad arguments fail alone (opus_decoder.c:608-690)."""
    meta, pk = golden_case("c2_96k_cbr_st")
    n = 8
    bd = cb.OpusBatchDecoder(48000, 2, n)
    try:
        good = pk[0][0]
        packets = [good] * n
        packets[1] = bytes([good[0] | 0x03, 0x00]) + good[2:]      # code 3 with zero frames -> INVALID_PACKET
        packets[2] = bytes([good[0] | 0x01]) + good[1:240]         # code 1 with odd payload -> INVALID_PACKET
        packets[3] = bytes([(good[0] & 0x07) | (3 << 3)]) + good[1:]  # SILK 60 ms TOC: 2880 > frame_size
        ret, pcm = bd.decode(packets, 960)
        assert ret[0] == 960 and ret[4] == 960 and ret[7] == 960
        assert ret[1] == cb.OPUS_INVALID_PACKET
        assert ret[2] == cb.OPUS_INVALID_PACKET
        assert ret[3] == cb.OPUS_BUFFER_TOO_SMALL
        assert crc32(pcm[0]) == meta["pcm_crc32"][0]
        assert np.array_equal(pcm[0], pcm[7])
        # second packet: streams that failed kept their state untouched (still "cold")
        ret2, pcm2 = bd.decode([pk[1][0]] * n, 960)
        assert (ret2 == 960).all()
        assert crc32(pcm2[0]) == meta["pcm_crc32"][1]
        # frame_size too small for the packet -> BUFFER_TOO_SMALL for everyone, state untouched
        ret3, _ = bd.decode([pk[2][0]] * n, 480)
        assert (ret3 == cb.OPUS_BUFFER_TOO_SMALL).all()
        ret4, pcm4 = bd.decode([pk[2][0]] * n, 960)
        assert crc32(pcm4[0]) == meta["pcm_crc32"][2]
    finally:
        bd.close()


def test_decode_float_and_ctl():
    meta, pk = golden_case("mono_64k")
    lib = cb.load()
    dec = cb.OpusDecoder(48000, 1)
    dec2 = cb.OpusDecoder(48000, 1)
    try:
        p = pk[0][0]
        out = np.zeros(960, dtype=np.float32)
        buf = ctypes.create_string_buffer(p, len(p))
        r = lib.opus_decode_float(dec._h, ctypes.cast(buf, ctypes.c_void_p), len(p), out.ctypes.data, 960, 0)
        assert r == 960
        ret, pcm = dec2.decode_raw(p, 960)
        assert np.array_equal(out, pcm[:, 0].astype(np.float32) / 32768.0)     # opus_decoder.c:745-748
        assert dec.bandwidth == 1105                                           # OPUS_BANDWIDTH_FULLBAND
        v = ctypes.c_int32(0)
        assert lib.opus_decoder_ctl(dec._h, cb.OPUS_GET_SAMPLE_RATE_REQUEST, ctypes.byref(v)) == 0 and v.value == 48000
        assert lib.opus_decoder_ctl(dec._h, 12345, ctypes.byref(v)) == cb.OPUS_UNIMPLEMENTED
        # reset: decoding packet 0 again from a reset decoder reproduces the cold-start output
        dec2.decode_raw(pk[1][0], 960)
        dec2.reset_state()
        ret, pcm_again = dec2.decode_raw(p, 960)
        assert np.array_equal(pcm_again, pcm)
    finally:
        dec.close()
        dec2.close()


@needs_ref
def test_decode_gain_matches_reference():
    """OPUS_SET_GAIN through opus_decoder_ctl, applied on the device (opus_decoder.c:567-577)."""
    meta, pk = golden_case("vbr_128k_st")
    for gain in (-3000, 1537):
        ref = RefDecoder(48000, 2)
        assert RefDecoder.lib().opus_decoder_ctl(ctypes.c_void_p(ref.h), 4034, ctypes.c_int(gain)) == 0
        dec = cb.OpusDecoder(48000, 2)
        dec.gain = gain
        assert dec.gain == gain
        try:
            for n, (p, _) in enumerate(pk[:30]):
                data = None if n in (7, 8, 20) else p
                ra, pa = ref.decode_raw(data, 960)
                rb, pb = dec.decode_raw(data, 960)
                assert ra == rb == 960 and np.array_equal(pa, pb), (gain, n)
        finally:
            dec.close()


def test_pipelined_submit_matches_golden():
    """opus_decode_group_submit/_wait: 64 streams replay the config-2 golden from different offsets, six calls
    in flight back to back (two at a time in the pipeline) with page-locked buffers; every call's PCM equals
    the synchronous path's and stream 0 equals the golden CRCs."""
    import torch
    meta, pk = golden_case("c2_96k_cbr_st")
    n, spf, ch, steps = 64, 960, 2, 12
    plen = len(pk[0][0])
    assert all(len(p) == plen for p, _ in pk)
    bd = cb.OpusBatchDecoder(48000, ch, n)
    sync = cb.OpusBatchDecoder(48000, ch, n)
    offs = (np.arange(n, dtype=np.int32) * plen).astype(np.int32)
    lens = np.full(n, plen, dtype=np.int32)
    arenas = [torch.zeros(n * plen, dtype=torch.uint8).pin_memory() for _ in range(steps)]
    pcms = [torch.zeros(n * spf * ch, dtype=torch.int16).pin_memory() for _ in range(steps)]
    rets = [torch.zeros(n, dtype=torch.int32).pin_memory() for _ in range(steps)]
    try:
        for t in range(steps):
            a = arenas[t].numpy()
            for i in range(n):
                a[i * plen:(i + 1) * plen] = np.frombuffer(pk[(3 * i + t) % len(pk)][0], dtype=np.uint8)
            bd.submit_flat(a, n * plen, offs, lens, spf, pcms[t].numpy(), rets[t].numpy())
        bd.wait()
        for t in range(steps):
            r2, p2 = sync.decode_flat(arenas[t].numpy(), n * plen, offs, lens, spf)
            assert (rets[t].numpy() == spf).all() and (r2 == spf).all()
            assert np.array_equal(pcms[t].numpy().reshape(n, spf, ch), p2), t
            assert crc32(pcms[t].numpy().reshape(n, spf, ch)[0]) == meta["pcm_crc32"][t], t
        assert np.array_equal(bd.final_ranges(), sync.final_ranges())
    finally:
        bd.close()
        sync.close()
