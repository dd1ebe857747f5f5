"""Encoder parity on the GPU: the CUDA library, called through its C ABI, against the committed golden
packets (made by the reference encoder) and against the compiled reference itself (oracle/_ref) on the
same seeded PCM.  Bar: identical packet bytes and OPUS_GET_FINAL_RANGE for every frame."""
import ctypes
import os

import numpy as np
import pytest

import concentus_b200 as cb
import enc_cases
from conftest import GOLDEN, RefEncoder, have_ref, needs_ref, read_bit

pytestmark = pytest.mark.gpu

CASES = {c.name: c for c in enc_cases.golden_cases()}


@pytest.mark.parametrize("name", sorted(CASES))
def test_single_stream_matches_golden(name):
    """opus_encoder_create / opus_encoder_ctl / opus_encode / OPUS_GET_FINAL_RANGE, frame by frame."""
    c = CASES[name]
    gold = read_bit(os.path.join(GOLDEN, "enc_%s.bit" % name))
    enc = cb.OpusEncoder(c.fs, c.ch, c.app)
    try:
        for req, val in c.ctls:
            enc.ctl(req, val)
        for i, frame in enumerate(c.frames()):
            pkt = enc.encode(frame, c.frame, c.max_bytes)
            assert pkt == gold[i][0], (name, i, len(pkt), len(gold[i][0]))
            assert enc.final_range == gold[i][1], (name, i)
    finally:
        enc.close()


@needs_ref
def test_batch_of_different_streams_matches_reference():
    """One launch per frame carries 96 streams with their own signal, bit rate, VBR mode and complexity."""
    n, frame, ch, nframes = 96, 960, 2, 10
    rng = np.random.default_rng(9)
    be = cb.OpusBatchEncoder(n, 48000, ch, enc_cases.LOWDELAY)
    refs = []
    pcms = []
    try:
        for s in range(n):
            r = RefEncoder(48000, ch, enc_cases.LOWDELAY)
            ctls = [(enc_cases.SET_BITRATE, int(rng.choice([24000, 48000, 64000, 96000, 160000, 320000])))]
            if s % 3 == 1:
                ctls.append((enc_cases.SET_VBR, 0))
            if s % 3 == 2:
                ctls.append((enc_cases.SET_VBR_CONSTRAINT, 0))
            if s % 4 == 0:
                ctls.append((enc_cases.SET_COMPLEXITY, int(rng.integers(0, 11))))
            for req, val in ctls:
                assert r.ctl(req, val) == 0
                be.encoders[s].ctl(req, val)
            refs.append(r)
            pcms.append(enc_cases.signal(enc_cases.KINDS[s % len(enc_cases.KINDS)], frame * nframes, ch, 1000 + s))
        for t in range(nframes):
            rows = np.stack([p[t * frame * ch:(t + 1) * frame * ch] for p in pcms])
            data, lens = be.encode(rows, frame, 1275)
            fr = be.final_ranges()
            for s in range(n):
                rn, rpkt, rrng = refs[s].encode(rows[s], frame, 1275)
                assert lens[s] == rn, (s, t, lens[s], rn)
                assert data[s, :rn].tobytes() == rpkt, (s, t)
                assert int(fr[s]) == rrng, (s, t)
    finally:
        be.close()


@needs_ref
def test_sweep_matches_reference():
    """Randomized sweep over sample rates, frame sizes, rates, buffers and applications (single-stream API)."""
    for c in enc_cases.sweep_cases(count=40):
        r = RefEncoder(c.fs, c.ch, c.app)
        e = cb.OpusEncoder(c.fs, c.ch, c.app)
        try:
            for req, val in c.ctls:
                assert r.ctl(req, val) == 0
                e.ctl(req, val)
            for i, frame in enumerate(c.frames()):
                rn, rpkt, rrng = r.encode(frame, c.frame, c.max_bytes)
                pkt = e.encode(frame, c.frame, c.max_bytes)
                assert pkt == rpkt, (c.name, c.fs, c.ch, c.frame, c.kind, c.ctls, c.max_bytes, i)
                assert e.final_range == rrng, (c.name, i)
        finally:
            e.close()


def test_full_size_batch_properties():
    """Config-3 size (4096 stereo streams, 20 ms): size-independent properties.  Streams are 64 distinct
    signals repeated 64 times: copies must produce identical packets; the first 64 must equal the
    golden-checked single-stream path; every packet must decode (own decoder) to the right length."""
    import torch
    n, frame, ch, nframes, distinct = 4096, 960, 2, 4, 64
    be = cb.OpusBatchEncoder(n, 48000, ch, enc_cases.LOWDELAY)
    singles = [cb.OpusEncoder(48000, ch, enc_cases.LOWDELAY) for _ in range(4)]
    try:
        be.ctl(enc_cases.SET_BITRATE, 96000)
        for e in singles:
            e.ctl(enc_cases.SET_BITRATE, 96000)
        base = np.stack([enc_cases.signal(enc_cases.KINDS[s % len(enc_cases.KINDS)], frame * nframes, ch, 50 + s)
                         for s in range(distinct)])
        d_pcm = torch.empty((n, frame * ch), dtype=torch.int16, device="cuda")
        d_data = torch.zeros((n, 1275), dtype=torch.uint8, device="cuda")
        d_len = torch.zeros(n, dtype=torch.int32, device="cuda")
        dec = cb.OpusDecoder(48000, ch)
        for t in range(nframes):
            rows = base[:, t * frame * ch:(t + 1) * frame * ch]
            d_pcm.copy_(torch.from_numpy(np.tile(rows, (n // distinct, 1))))
            torch.cuda.synchronize()
            be.encode_device(d_pcm.data_ptr(), frame, d_data.data_ptr(), 1275, d_len.data_ptr())
            cb.load().opus_b200_synchronize()
            lens = d_len.cpu().numpy()
            data = d_data.cpu().numpy()
            assert (lens > 2).all() and (lens <= 1275).all()
            for s in range(distinct):
                same = data[s::distinct]
                assert (lens[s::distinct] == lens[s]).all()
                assert (same[:, :lens[s]] == same[0, :lens[s]]).all(), (t, s)
            for k, e in enumerate(singles):
                assert e.encode(rows[k], frame, 1275) == data[k, :lens[k]].tobytes(), (t, k)
            ret, _ = dec.decode_raw(data[0, :lens[0]].tobytes(), frame)
            assert ret == frame
    finally:
        be.close()
        for e in singles:
            e.close()


def test_float_entry_point_and_errors():
    lib = cb.load()
    enc = cb.OpusEncoder(48000, 1, enc_cases.LOWDELAY)
    enc2 = cb.OpusEncoder(48000, 1, enc_cases.LOWDELAY)
    try:
        x = enc_cases.signal("mix", 960, 1, 4)
        xf = (x.astype(np.float32) / 32768.0)
        buf = np.zeros(1275, dtype=np.uint8)
        n = lib.opus_encode_float(enc._h, xf.ctypes.data, 960, buf.ctypes.data, 1275)
        assert n > 0 and buf[:n].tobytes() == enc2.encode(x, 960, 1275)
        assert lib.opus_encode(enc._h, x.ctypes.data, 1000, buf.ctypes.data, 1275) == -1     # bad frame size
        assert lib.opus_encode(enc._h, x.ctypes.data, 960, buf.ctypes.data, 0) == -1
        assert lib.opus_encoder_ctl(ctypes.c_void_p(enc._h), 4010, ctypes.c_int(11)) == -1
        assert lib.opus_encoder_ctl(ctypes.c_void_p(enc._h), 4999, ctypes.c_int(0)) == -5
        err = ctypes.c_int(0)
        assert not lib.opus_encoder_create(44100, 1, 2049, ctypes.byref(err)) and err.value == -1
        # frames the reference would code with SILK are refused, not mis-coded
        voip = cb.OpusEncoder(16000, 1, enc_cases.AUDIO)
        voip.ctl(enc_cases.SET_BITRATE, 12000)
        assert lib.opus_encode(voip._h, x.ctypes.data, 320, buf.ctypes.data, 1275) == -5
        voip.close()
    finally:
        enc.close()
        enc2.close()


def test_pipelined_submit_matches_synchronous_calls():
    """opus_encode_group_submit/_wait against opus_encode_group_flat on a twin group: same packets, lengths and
    final ranges for ten frames queued back to back (pageable and page-locked buffers both work)."""
    import torch
    n, frame, ch, nframes = 48, 960, 2, 10
    a = cb.OpusBatchEncoder(n, 48000, ch, enc_cases.LOWDELAY)
    b = cb.OpusBatchEncoder(n, 48000, ch, enc_cases.LOWDELAY)
    try:
        for g in (a, b):
            g.ctl(enc_cases.SET_BITRATE, 96000)
        sig = np.stack([enc_cases.signal(enc_cases.KINDS[s % len(enc_cases.KINDS)], frame * nframes, ch, 300 + s)
                        for s in range(n)])
        rows = [torch.from_numpy(np.ascontiguousarray(sig[:, t * frame * ch:(t + 1) * frame * ch])) for t in range(nframes)]
        rows = [r.pin_memory() if t % 2 == 0 else r for t, r in enumerate(rows)]
        datas = [torch.zeros((n, 1275), dtype=torch.uint8).pin_memory() for _ in range(nframes)]
        lens = [np.zeros(n, dtype=np.int32) for _ in range(nframes)]
        for t in range(nframes):
            a.submit(rows[t].numpy(), frame, datas[t].numpy(), lens[t])
        a.wait()
        for t in range(nframes):
            d2, l2 = b.encode(rows[t].numpy(), frame, 1275)
            assert np.array_equal(lens[t], l2), t
            for s in range(n):
                assert np.array_equal(datas[t].numpy()[s, :l2[s]], d2[s, :l2[s]]), (t, s)
        assert np.array_equal(a.final_ranges(), b.final_ranges())
    finally:
        a.close()
        b.close()


@needs_ref
@pytest.mark.parametrize("ms,vbr,maxb", [(40, 1, 1275), (60, 1, 1275), (60, 0, 1275), (40, 0, 250), (60, 1, 3000)])
def test_long_frames_match_reference(ms, vbr, maxb):
    """40 / 60 ms frames (two or three 20 ms CELT frames in one packet, opus_encoder.c:1362-1438): a batch of
    24 stereo 48 kHz streams with their own signals and bit rates, eight frames, against the reference."""
    n, ch, fs, nframes = 24, 2, 48000, 8
    frame = fs * ms // 1000
    be = cb.OpusBatchEncoder(n, fs, ch, enc_cases.LOWDELAY)
    refs, pcms = [], []
    try:
        for s in range(n):
            r = RefEncoder(fs, ch, enc_cases.LOWDELAY)
            for req, val in ((enc_cases.SET_BITRATE, 32000 + 9000 * s), (enc_cases.SET_VBR, vbr)):
                assert r.ctl(req, val) == 0
                be.encoders[s].ctl(req, val)
            refs.append(r)
            pcms.append(enc_cases.signal(enc_cases.KINDS[s % len(enc_cases.KINDS)], frame * nframes, ch, 2000 + s))
        for t in range(nframes):
            rows = np.stack([p[t * frame * ch:(t + 1) * frame * ch] for p in pcms])
            data, lens = be.encode(rows, frame, maxb)
            fr = be.final_ranges()
            for s in range(n):
                rn, rpkt, rrng = refs[s].encode(rows[s], frame, maxb)
                assert lens[s] == rn, (s, t, lens[s], rn)
                assert data[s, :rn].tobytes() == rpkt, (s, t)
                assert int(fr[s]) == rrng, (s, t)
    finally:
        be.close()


@needs_ref
def test_long_frames_single_stream_and_mixed_rates():
    """opus_encode with 40 / 60 ms frames on single encoders at 16 / 24 / 48 kHz (the kernel's stage rendezvous
    count follows the longest sub-frame count of the launch)."""
    for fs, ch in ((16000, 1), (24000, 2), (48000, 1)):
        for ms in (40, 60):
            enc, ref = cb.OpusEncoder(fs, ch, enc_cases.LOWDELAY), RefEncoder(fs, ch, enc_cases.LOWDELAY)
            try:
                frame = fs * ms // 1000
                pcm = enc_cases.signal("chirp", frame * 4, ch, 5)
                for t in range(4):
                    x = pcm[t * frame * ch:(t + 1) * frame * ch]
                    n1, p1, r1 = ref.encode(x, frame, 1500)
                    p2 = enc.encode(x, frame, 1500)
                    assert n1 == len(p2) and p1 == p2, (fs, ch, ms, t, n1, len(p2))
                    assert enc.final_range == r1
            finally:
                enc.close()


@needs_ref
def test_voip_application_and_surround_stream_settings_match_reference():
    """Sixteen streams with the settings that used to be refused: OPUS_APPLICATION_VOIP coded by CELT (forced mode or a
    rate above the speech threshold: VOIP high-pass, opus_encoder.c:325-360), an LFE stream (OPUS_SET_LFE), and
    streams with a surround mask (OPUS_SET_ENERGY_MASK, a new one before every frame).  Against the reference."""
    n, fs, frame, nframes = 16, 48000, 960, 8
    rng = np.random.RandomState(5)
    rlib = RefEncoder.lib()
    lib = cb.load()
    for ch in (1, 2):
        encs = [cb.OpusEncoder(fs, ch, enc_cases.VOIP if s % 4 < 2 else enc_cases.AUDIO) for s in range(n)]
        refs = [RefEncoder(fs, ch, enc_cases.VOIP if s % 4 < 2 else enc_cases.AUDIO) for s in range(n)]
        try:
            for s in range(n):
                kind = s % 4                       # 0 VOIP forced CELT, 1 VOIP high rate, 2 masked, 3 LFE (mono) / masked
                settings = [(enc_cases.SET_BITRATE, (150000 if kind == 1 else 40000 + 3000 * s) * ch)]
                if kind != 1:
                    settings.append((11002, 1002))
                if kind == 3 and ch == 1:
                    settings += [(10024, 1), (enc_cases.SET_BITRATE, 4000 + 500 * s)]
                for req, val in settings:
                    assert refs[s].ctl(req, val) == 0
                    encs[s].ctl(req, val)
            pcms = [enc_cases.signal(enc_cases.KINDS[s % len(enc_cases.KINDS)], frame * nframes, ch, 3100 + s) for s in range(n)]
            keep = []
            for t in range(nframes):
                for s in range(n):
                    if s % 4 >= 2 and not (s % 4 == 3 and ch == 1):
                        m = np.ascontiguousarray(rng.randint(-2500, 400, size=21 * ch).astype(np.int16))
                        keep.append(m)
                        assert rlib.opus_encoder_ctl(ctypes.c_void_p(refs[s].h), ctypes.c_int(10026), ctypes.c_void_p(m.ctypes.data)) == 0
                        assert lib.opus_encoder_ctl(ctypes.c_void_p(encs[s]._h), ctypes.c_int(10026), ctypes.c_void_p(m.ctypes.data)) == 0
                for s in range(n):
                    x = pcms[s][t * frame * ch:(t + 1) * frame * ch]
                    rn, rpkt, rrng = refs[s].encode(x, frame, 1275)
                    pkt = encs[s].encode(x, frame, 1275)
                    assert len(pkt) == rn and pkt == rpkt, (ch, s, t, len(pkt), rn)
                    assert encs[s].get(4031) & 0xFFFFFFFF == rrng, (ch, s, t)
        finally:
            for e in encs:
                e.close()


# ---- real audio (excerpts of the reference's AudioData; tests/golden/make_golden_enc_real.py) -----------------
REAL = {c.name: c for c in enc_cases.real_cases()}


@pytest.mark.parametrize("name", sorted(REAL))
def test_real_audio_matches_golden(name):
    """Every frame of the real-audio excerpt through the C ABI: packet bytes and final range equal the reference
    encoder's (procedure of CSharp/ParityTest/TestDriver.cs:253-275 with its random bitrate changes :219-224).
    32 copies of the stream per launch so that the batch kernel shape is the one under test."""
    c = REAL[name]
    gold = read_bit(os.path.join(GOLDEN, "encr_%s.bit" % name))
    pcm = enc_cases.real_pcm(c.ch)
    n = 32
    be = cb.OpusBatchEncoder(n, 48000, c.ch, c.app)
    try:
        for req, val in c.ctls:
            be.ctl(req, val)
        sched = c.bitrate_schedule()
        step = c.frame * c.ch
        for t in range(c.nframes):
            if t in sched:
                be.ctl(enc_cases.SET_BITRATE, sched[t])
            data, lens = be.encode(np.tile(pcm[t * step:(t + 1) * step], (n, 1)), c.frame, c.max_bytes)
            assert (lens == len(gold[t][0])).all(), (name, t, lens[:4], len(gold[t][0]))
            for i in (0, 13, 31):
                assert data[i, :lens[i]].tobytes() == gold[t][0], (name, t, i)
            assert (be.final_ranges() == gold[t][1]).all(), (name, t)
    finally:
        be.close()


@needs_ref
@pytest.mark.parametrize("vbr", [0, 1])
def test_4096_real_audio_streams_every_rotation(vbr):
    """Config 3 as SURVEY 8d C3 states it: 4096 streams, each the real-audio excerpt rotated by its own number of
    frames, 128 kbit/s, complexity 10, 20 ms, CBR (and VBR).  EVERY distinct rotation is held to a reference
    encoder of its own, packet bytes + final range, on every frame."""
    n, ch, frames = 4096, 2, 10
    pcm = enc_cases.real_pcm(ch)
    rot = enc_cases.c3_rotations(n)
    distinct = sorted(set(int(r) for r in rot))
    assert len(distinct) > 200
    first = {r: int(np.nonzero(rot == r)[0][0]) for r in distinct}
    refs = {}
    for r in distinct:
        e = RefEncoder(48000, ch, enc_cases.LOWDELAY)
        for req, val in ((enc_cases.SET_BITRATE, 128000), (enc_cases.SET_VBR, vbr), (enc_cases.SET_COMPLEXITY, 10)):
            assert e.ctl(req, val) == 0
        refs[r] = e
    be = cb.OpusBatchEncoder(n, 48000, ch, enc_cases.LOWDELAY)
    try:
        be.ctl(enc_cases.SET_BITRATE, 128000)
        be.ctl(enc_cases.SET_VBR, vbr)
        be.ctl(enc_cases.SET_COMPLEXITY, 10)
        for t in range(frames):
            rows = enc_cases.c3_rows(pcm, ch, rot, t)
            data, lens = be.encode(rows, 960, 1275)
            fr = be.final_ranges()
            for r in distinct:
                i = first[r]
                rn, rpkt, rrng = refs[r].encode(rows[i], 960, 1275)
                same = np.nonzero(rot == r)[0]
                assert (lens[same] == rn).all(), (t, r, lens[same][:4], rn)
                for j in same:
                    assert data[j, :rn].tobytes() == rpkt, (t, r, int(j))
                assert (fr[same] == rrng).all(), (t, r)
    finally:
        be.close()
