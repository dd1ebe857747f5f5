"""SILK / hybrid decode parity on the GPU: the CUDA library through its C ABI against the committed SILK and
hybrid / mode-transition goldens (reference encoder packets from the reference's audio, with loss / FEC / DTX
call sequences) and against the compiled reference on batches that mix SILK, hybrid and CELT streams whose
modes change under way.  Bit-exact PCM + final range."""
import os

import numpy as np
import pytest

import concentus_b200 as cb
from conftest import GOLDEN, RefDecoder, crc32, golden_case, have_ref, needs_ref, read_bit, silk_golden_index
from scenarios import silk_calls, silk_fuzz_calls
from test_emu_silk import TRANSITION_SCRIPTS, scripted_packets, silk_packets

pytestmark = pytest.mark.gpu

CASES = sorted(silk_golden_index().keys())


@pytest.mark.parametrize("name", CASES)
def test_single_stream_silk_matches_golden(name):
    meta = silk_golden_index()[name]
    pk = read_bit(os.path.join(GOLDEN, name + ".bit"))
    dec = cb.OpusDecoder(meta["fs"], meta["channels"])
    try:
        for k, (idx, use_fec) in enumerate(silk_calls(len(pk), meta["loss"], meta["fec"], seed=11)):
            data = pk[idx][0] if idx is not None else None
            ret, pcm = dec.decode_raw(data, meta["samples_per_frame"], use_fec)
            assert ret == meta["ret"][k], (name, k, ret)
            assert crc32(pcm[:ret]) == meta["pcm_crc32"][k], (name, "pcm of call", k)
            assert dec.final_range == meta["final_range"][k], (name, k)
    finally:
        dec.close()


@needs_ref
def test_batch_mixing_silk_and_celt_streams():
    """One batch call per 20 ms carries WB/MB/NB SILK streams (mono and stereo packets, some lost) next to
    CELT streams; every stream must equal the reference decoder run on its own packets."""
    fs, ch, nfr = 48000, 2, 30
    streams = []
    for i, (efs, ech) in enumerate([(16000, 1), (16000, 2), (12000, 1), (8000, 1), (8000, 2), (16000, 1)] * 4):
        streams.append(silk_packets(efs, ech, 20, 12000 * ech + efs // 2 + 500 * i, nfr, seed=20 + i))
    _, celt = golden_case("c2_96k_cbr_st")
    for i in range(8):
        streams.append([celt[(37 * i + t) % len(celt)][0] for t in range(nfr)])
    n = len(streams)
    bd = cb.OpusBatchDecoder(fs, ch, n)
    refs = [RefDecoder(fs, ch) for _ in range(n)]
    rng = np.random.default_rng(4)
    try:
        for t in range(nfr):
            pkts = [None if (t > 2 and rng.random() < 0.1) else streams[i][t] for i in range(n)]
            ret, pcm = bd.decode(pkts, 960)
            fr = bd.final_ranges()
            for i in range(n):
                r1, p1 = refs[i].decode_raw(pkts[i], 960)
                assert ret[i] == r1, (i, t, ret[i], r1)
                assert np.array_equal(pcm[i][:r1], p1[:r1]), (i, t)
                assert int(fr[i]) == refs[i].final_range, (i, t)
    finally:
        bd.close()


@needs_ref
@pytest.mark.parametrize("fs,ch,ms", [(48000, 2, 20), (16000, 1, 10), (24000, 2, 40)])
def test_batch_of_streams_changing_mode(fs, ch, ms):
    """Every stream of the batch follows its own script of mode changes (SILK-only / hybrid / CELT-only at
    varying bandwidths, mono and stereo packets, redundant frames at the joints), shifted in time against
    the others, with some packets lost: each batch call holds a different mix of routes."""
    names = sorted(TRANSITION_SCRIPTS)
    streams = []
    for i in range(20):
        pk = scripted_packets(1 + i % 2, ms, TRANSITION_SCRIPTS[names[i % len(names)]], 26000 + 1500 * i, seed=40 + i)
        streams.append(pk[i % 5:])
    nfr = min(len(p) for p in streams)
    n = len(streams)
    fsz = fs * ms // 1000
    bd = cb.OpusBatchDecoder(fs, ch, n)
    refs = [RefDecoder(fs, ch) for _ in range(n)]
    rng = np.random.default_rng(8)
    try:
        for t in range(nfr):
            pkts = [None if (t > 1 and rng.random() < 0.12) else streams[i][t] for i in range(n)]
            ret, pcm = bd.decode(pkts, fsz)
            fr = bd.final_ranges()
            for i in range(n):
                r1, p1 = refs[i].decode_raw(pkts[i], fsz)
                assert ret[i] == r1, (i, t, ret[i], r1)
                assert np.array_equal(pcm[i][:r1], p1[:r1]), (i, t)
                assert int(fr[i]) == refs[i].final_range, (i, t)
    finally:
        bd.close()


def _full_size_rotations(name, fs, fsz, n, steps):
    """16384 streams replay a golden fixture from 64 different start offsets.  Every one of the 64 distinct
    histories is held to a reference decoder of its own (PCM, return value, final range), every copy of it must
    agree, and the stream that starts at packet 0 must also reproduce the golden CRCs."""
    meta = silk_golden_index()[name]
    pk = [p for p, _ in read_bit(os.path.join(GOLDEN, name + ".bit"))]
    bd = cb.OpusBatchDecoder(fs, 1, n)
    refs = [RefDecoder(fs, 1) for _ in range(64)] if have_ref() else None
    try:
        rot = (3 * (np.arange(n) % 64))
        for t in range(steps):
            pkts = [pk[(rot[i] + t) % len(pk)] for i in range(n)]
            ret, pcm = bd.decode(pkts, fsz)
            assert (ret == fsz).all()
            fr = bd.final_ranges()
            for i in range(64):
                assert (pcm[i::64] == pcm[i]).all(), (t, i)
                assert (fr[i::64] == fr[i]).all(), (t, i)
                if refs is not None:
                    r1, p1 = refs[i].decode_raw(pkts[i], fsz)
                    assert r1 == fsz
                    assert np.array_equal(pcm[i], p1[:fsz]), (t, i)
                    assert int(fr[i]) == refs[i].final_range, (t, i)
            assert crc32(pcm[0]) == meta["pcm_crc32"][t], t       # stream 0 plays the fixture from its start
    finally:
        bd.close()


def test_full_size_silk_batch_properties():
    """Config-4 size: 16384 mono 16 kHz SILK wideband streams, 20 ms (thread-per-stream kernel at full size)."""
    _full_size_rotations("silk_wb_mono_32k", 16000, 320, 16384, 6)


@needs_ref
@pytest.mark.parametrize("name", ["silk_wb_mono_32k", "hyb_fb_mono_32k", "trans_all_st_to24"])
def test_decoder_queries_follow_the_reference(name):
    """OPUS_GET_PITCH / GET_BANDWIDTH / GET_LAST_PACKET_DURATION after every packet of SILK, hybrid and
    mode-switching streams (opus_decoder.c:836-849: the pitch comes from the SILK decoder unless the last
    frame was CELT-only)."""
    import ctypes
    meta = silk_golden_index()[name]
    pk = read_bit(os.path.join(GOLDEN, name + ".bit"))
    dec = cb.OpusDecoder(meta["fs"], meta["channels"])
    ref = RefDecoder(meta["fs"], meta["channels"])

    def rget(req):
        v = ctypes.c_int32(0)
        assert ref.lib().opus_decoder_ctl(ctypes.c_void_p(ref.h), req, ctypes.byref(v)) == 0
        return v.value
    try:
        for k, (p, _) in enumerate(pk[:60]):
            data = None if (k % 11 == 7) else p
            r1, _ = dec.decode_raw(data, meta["samples_per_frame"])
            r2, _ = ref.decode_raw(data, meta["samples_per_frame"])
            assert r1 == r2
            assert dec.pitch == rget(4033), (name, k)
            assert dec.bandwidth == rget(4009), (name, k)
            assert dec.last_packet_duration == rget(4039), (name, k)
    finally:
        dec.close()
        ref.close()


@needs_ref
@pytest.mark.parametrize("fs,ch", [(48000, 2), (16000, 1)])
def test_batch_of_corrupted_streams_follows_reference(fs, ch):
    """Sixteen streams of damaged / truncated / random packets with random TOC bytes in one batch per call:
    every route (CELT, SILK, hybrid, transitions, concealment, errors) mixes in each launch, and each stream
    must still do exactly what the reference does with the same garbage."""
    names = ["silk_wb_mono_32k", "silk_wb_st_40k", "silk_nb_st_40ms", "silk_wb_mono_60ms", "hyb_fb_mono_32k",
             "hyb_swb_st_48k", "trans_all_st_to24", "trans_all_mono_10ms_loss"]
    streams = []
    for i in range(16):
        pk = [p for p, _ in read_bit(os.path.join(GOLDEN, names[i % len(names)] + ".bit"))]
        streams.append(silk_fuzz_calls(pk, seed=100 + i, random_toc=(i % 2 == 1)))
    steps = min(len(c) for c in streams)
    n = len(streams)
    fsz = fs * 60 // 1000
    bd = cb.OpusBatchDecoder(fs, ch, n)
    refs = [RefDecoder(fs, ch) for _ in range(n)]
    try:
        for t in range(steps):
            pkts = [streams[i][t] for i in range(n)]
            ret, pcm = bd.decode(pkts, fsz)
            fr = bd.final_ranges()
            for i in range(n):
                r1, p1 = refs[i].decode_raw(pkts[i], fsz)
                assert ret[i] == r1, (i, t, ret[i], r1)
                if r1 > 0:
                    assert np.array_equal(pcm[i][:r1], p1[:r1]), (i, t)
                assert int(fr[i]) == refs[i].final_range, (i, t)
    finally:
        bd.close()


def test_full_size_hybrid_batch_properties():
    """The hybrid set of config 4 at full size: 16384 mono 48 kHz streams of hybrid fullband packets (SILK kernel +
    CELT layer kernels at full occupancy)."""
    _full_size_rotations("hyb_fb_mono_32k", 48000, 960, 16384, 5)
