"""CPU-side parity of the encoder kernel source (host-emulation build) against the committed golden
packets (made by the reference encoder, tests/golden/make_golden_enc.py) and, where oracle/_ref is
built, against the reference encoder itself on a wider sweep.  Bit-exact: packet bytes + final range."""
import os

import numpy as np
import pytest

import enc_cases
from conftest import GOLDEN, EmuEncoder, RefEncoder, enc_golden_index, needs_ref, read_bit

CASES = {c.name: c for c in enc_cases.golden_cases()}


def test_golden_index_matches_case_list():
    idx = enc_golden_index()
    assert [m["name"] for m in idx] == [c.name for c in enc_cases.golden_cases()]
    for m in idx:
        c = CASES[m["name"]]
        assert enc_cases.pcm_md5(c.pcm()) == m["pcm_md5"], "synthetic PCM drifted: regenerate goldens"


@pytest.mark.parametrize("name", sorted(CASES))
def test_emu_encoder_matches_golden(emu_enc_lib, name):
    c = CASES[name]
    gold = read_bit(os.path.join(GOLDEN, "enc_%s.bit" % name))
    e = EmuEncoder(emu_enc_lib, c.fs, c.ch, c.app)
    for req, val in c.ctls:
        assert e.ctl(req, val) == 0
    for i, frame in enumerate(c.frames()):
        n, pkt, rng = e.encode(frame, c.frame, c.max_bytes)
        assert n == len(gold[i][0]), (name, i, n)
        assert pkt == gold[i][0], (name, i)
        assert rng == gold[i][1], (name, i)


@needs_ref
def test_emu_encoder_matches_reference_sweep(emu_enc_lib):
    for c in enc_cases.sweep_cases():
        r = RefEncoder(c.fs, c.ch, c.app)
        e = EmuEncoder(emu_enc_lib, c.fs, c.ch, c.app)
        for req, val in c.ctls:
            assert r.ctl(req, val) == e.ctl(req, val)
        for i, frame in enumerate(c.frames()):
            a, b = r.encode(frame, c.frame, c.max_bytes), e.encode(frame, c.frame, c.max_bytes)
            assert a == b, (c.name, c.fs, c.ch, c.frame, c.kind, c.ctls, c.max_bytes, i, a[0], b[0])


@needs_ref
def test_emu_encoder_midstream_changes(emu_enc_lib):
    """Bit-rate, frame size, bandwidth and complexity changed between frames of one stream."""
    c = enc_cases.EncCase("mid", 48000, 2, enc_cases.LOWDELAY, 120, "mix", 400, [(enc_cases.SET_BITRATE, 64000)])
    pcm = c.pcm()
    r = RefEncoder(c.fs, c.ch, c.app)
    e = EmuEncoder(emu_enc_lib, c.fs, c.ch, c.app)
    pos = 0
    rng = np.random.default_rng(3)
    step = 0
    while True:
        frame = int(rng.choice([120, 240, 480, 960]))
        if pos + frame > 120 * 400:
            break
        if step % 7 == 3:
            br = int(rng.choice([16000, 32000, 64000, 128000]))
            assert r.ctl(enc_cases.SET_BITRATE, br) == e.ctl(enc_cases.SET_BITRATE, br)
        if step % 11 == 5:
            cx = int(rng.integers(0, 11))
            assert r.ctl(enc_cases.SET_COMPLEXITY, cx) == e.ctl(enc_cases.SET_COMPLEXITY, cx)
        if step % 13 == 6:
            bw = int(rng.choice([-1000, 1101, 1103, 1104, 1105]))
            assert r.ctl(enc_cases.SET_BANDWIDTH, bw) == e.ctl(enc_cases.SET_BANDWIDTH, bw)
        chunk = pcm[pos * 2:(pos + frame) * 2]
        a, b = r.encode(chunk, frame, 1275), e.encode(chunk, frame, 1275)
        assert a == b, (step, frame, a[0], b[0])
        pos += frame
        step += 1
    assert step > 50


@needs_ref
def test_emu_encoder_error_codes(emu_enc_lib):
    """Argument errors agree with the reference (opus_encoder.c:983-991)."""
    r = RefEncoder(48000, 1, enc_cases.LOWDELAY)
    e = EmuEncoder(emu_enc_lib, 48000, 1, enc_cases.LOWDELAY)
    pcm = np.zeros(4000, dtype=np.int16)
    for frame in (100, 119, 121, 500, 1000, 1920, 2880):
        a, b = r.encode(pcm, frame, 1275), e.encode(pcm, frame, 1275)
        assert a[0] == b[0], frame
    # out-of-range ctl values
    for req, val in [(4002, 0), (4002, -5), (4010, 11), (4010, -1), (4006, 2), (4022, 2), (4008, 1100), (4036, 7),
                     (4040, 4999), (4014, 101), (11002, 999), (4000, 2050)]:
        assert r.ctl(req, val) == e.ctl(req, val), (req, val)


@needs_ref
@pytest.mark.parametrize("ms", [40, 60])
def test_emu_long_frames_match_reference(emu_enc_lib, ms):
    """40 / 60 ms frames in CELT-only mode: two or three 20 ms frames in one packet (opus_encoder.c:1362-1438),
    decisions taken once for the long frame.  Packets and final range against the reference for both
    applications that reach CELT, mono / stereo, three rates, VBR / CBR, generous and tight buffers."""
    for app in (enc_cases.LOWDELAY, 2049):
        for ch in (1, 2):
            for fs in (48000, 24000, 16000):
                for vbr, maxb, rate in ((1, 1275, 64000), (0, 1275, 64000), (1, 300, 64000), (0, 200, 64000), (1, 3800, 250000)):
                    ref, emu = RefEncoder(fs, ch, app), EmuEncoder(emu_enc_lib, fs, ch, app)
                    for e in (ref, emu):
                        assert e.ctl(enc_cases.SET_BITRATE, rate * ch) == 0
                        assert e.ctl(enc_cases.SET_VBR, vbr) == 0
                        if app == 2049:
                            assert e.ctl(11002, 1002) == 0            # OPUS_SET_FORCE_MODE(MODE_CELT_ONLY)
                    frame = fs * ms // 1000
                    pcm = enc_cases.signal(enc_cases.KINDS[(ch + ms) % len(enc_cases.KINDS)], frame * 5, ch, 77)
                    for t in range(5):
                        x = pcm[t * frame * ch:(t + 1) * frame * ch]
                        a, b = ref.encode(x, frame, maxb), emu.encode(x, frame, maxb)
                        assert a == b, (app, ch, fs, ms, vbr, maxb, t, a[0], b[0])


@needs_ref
def test_emu_lfe_and_surround_mask_match_reference(emu_enc_lib):
    """The two settings only a surround multistream encoder uses on its elementary encoders: OPUS_SET_LFE
    (celt_encoder.c:1649-2117 `st->lfe` branches, quant_bands.c:232,266,302) and OPUS_SET_ENERGY_MASK
    (celt_encoder.c:1726-1801 masking terms; opus_encoder.c:1792 no stereo width reduction).  Masks here are
    random Q10 values in the range surround_analysis produces (and beyond it, to hit the clamps); a new one is
    set before every frame, as the multistream encoder does."""
    import ctypes
    rng = np.random.RandomState(11)
    emu_enc_lib.emu_enc_set_mask.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    rlib = RefEncoder.lib()
    for ch, lfe, rate, vbr, cvbr, ms, fs in [(1, 1, 6000, 1, 1, 20, 48000), (1, 1, 3500, 1, 0, 20, 48000),
                                            (1, 1, 8000, 0, 0, 10, 48000), (1, 1, 5000, 1, 1, 40, 48000),
                                            (2, 0, 96000, 1, 1, 20, 48000), (2, 0, 64000, 1, 0, 20, 48000),
                                            (1, 0, 48000, 1, 1, 20, 48000), (2, 0, 80000, 0, 0, 20, 48000),
                                            (2, 0, 40000, 1, 1, 10, 48000), (1, 0, 32000, 1, 0, 5, 48000),
                                            (2, 0, 120000, 1, 1, 60, 48000), (2, 0, 70000, 1, 1, 20, 24000)]:
        ref, emu = RefEncoder(fs, ch, 2049), EmuEncoder(emu_enc_lib, fs, ch, 2049)
        for e in (ref, emu):
            assert e.ctl(enc_cases.SET_BITRATE, rate) == 0
            assert e.ctl(enc_cases.SET_VBR, vbr) == 0
            assert e.ctl(4020, cvbr) == 0
            assert e.ctl(11002, 1002) == 0
            if lfe:
                assert e.ctl(10024, 1) == 0
        frame = fs * ms // 1000
        nfr = 12
        pcm = enc_cases.signal(enc_cases.KINDS[(ch + ms + rate // 1000) % len(enc_cases.KINDS)], frame * nfr, ch, 5)
        masks = []                                   # keep the arrays alive: the reference holds the pointer
        for t in range(nfr):
            if not lfe:
                scale = [300, 1500, 4000][t % 3]
                m = np.ascontiguousarray(rng.randint(-scale, scale // 4 + 1, size=21 * ch).astype(np.int16))
                if t % 5 == 4:
                    m[:] = -3000                      # heavy masking everywhere
                masks.append(m)
                assert rlib.opus_encoder_ctl(ctypes.c_void_p(ref.h), ctypes.c_int(10026), ctypes.c_void_p(m.ctypes.data)) == 0
                emu_enc_lib.emu_enc_set_mask(emu.h, m.ctypes.data)
            x = pcm[t * frame * ch:(t + 1) * frame * ch]
            a, b = ref.encode(x, frame, 1275), emu.encode(x, frame, 1275)
            assert a == b, (ch, lfe, rate, vbr, ms, fs, t, a[0], b[0])


@needs_ref
def test_emu_surround_analysis_matches_reference(emu_enc_lib):
    """surround_analysis (opus_multistream_encoder.c:224-360) -- the per-channel masks a surround multistream
    encoder hands to its elementary encoders -- against the reference's own function on 3..8 channel frames of
    2.5..60 ms at three rates, the window / pre-emphasis memories carried over several frames."""
    import ctypes
    rlib = RefEncoder.lib()
    rlib.opus_custom_mode_create.restype = ctypes.c_void_p
    rlib.opus_custom_mode_create.argtypes = [ctypes.c_int32, ctypes.c_int, ctypes.c_void_p]
    mode = rlib.opus_custom_mode_create(48000, 960, None)
    assert mode
    COPY = ctypes.CFUNCTYPE(None, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int)

    def copy_in(dst, dst_stride, src, src_stride, src_channel, frame_size):
        s = np.ctypeslib.as_array(ctypes.cast(src, ctypes.POINTER(ctypes.c_int16)), shape=(frame_size * src_stride,))
        d = np.ctypeslib.as_array(ctypes.cast(dst, ctypes.POINTER(ctypes.c_int16)), shape=(frame_size * dst_stride,))
        d[::dst_stride] = s[src_channel::src_stride]
    cb = COPY(copy_in)
    rlib.surround_analysis.restype = None
    rlib.surround_analysis.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int] * 4 + [COPY, ctypes.c_int]
    emu_enc_lib.emu_surround_analysis.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int] + [ctypes.c_void_p] * 3
    rng = np.random.RandomState(3)
    for channels in (3, 4, 5, 6, 7, 8):
        for fs, ms10 in ((48000, 200), (48000, 25), (48000, 100), (48000, 400), (48000, 600), (24000, 200), (16000, 50),
                         (24000, 400), (8000, 200)):
            len_ = fs * ms10 // 10000
            mem_r = np.zeros(channels * 120, dtype=np.int32); pre_r = np.zeros(channels, dtype=np.int32)
            mem_e = mem_r.copy(); pre_e = pre_r.copy()
            for t in range(4):
                pcm = np.zeros((len_, channels), dtype=np.int16)
                for c in range(channels):
                    kind = enc_cases.KINDS[(c + t + channels) % len(enc_cases.KINDS)]
                    gain = [1.0, 0.3, 0.05, 0.0][(c + t) % 4] if t else 1.0
                    pcm[:, c] = (enc_cases.signal(kind, len_, 1, 100 * t + c).astype(np.float64) * gain).astype(np.int16)
                pcm = np.ascontiguousarray(pcm.reshape(-1))
                out_r = np.zeros(21 * channels, dtype=np.int16); out_e = out_r.copy()
                rlib.surround_analysis(mode, pcm.ctypes.data, out_r.ctypes.data, mem_r.ctypes.data, pre_r.ctypes.data,
                                       len_, 120, channels, fs, cb, 0)
                emu_enc_lib.emu_surround_analysis(pcm.ctypes.data, channels, len_, fs, mem_e.ctypes.data,
                                                  pre_e.ctypes.data, out_e.ctypes.data)
                assert np.array_equal(out_r, out_e), (channels, fs, ms10, t, out_r[:8], out_e[:8])
                assert np.array_equal(mem_r, mem_e) and np.array_equal(pre_r, pre_e), (channels, fs, ms10, t)


@needs_ref
def test_emu_voip_application_in_celt_mode_matches_reference(emu_enc_lib):
    """OPUS_APPLICATION_VOIP when the frame is coded by CELT (forced, or chosen because the rate is high, or
    because the frame is shorter than 10 ms): the input goes through the VOIP high-pass (hp_cutoff + silk_biquad_alt,
    opus_encoder.c:325-360) instead of the DC rejection, and the mode thresholds move (voice_est 115, +8000)."""
    for ch, fs, rate, force, ms10 in [(1, 48000, 64000, 1, 200), (2, 48000, 96000, 1, 200), (1, 16000, 32000, 1, 100),
                                      (2, 24000, 64000, 1, 200), (1, 48000, 128000, 0, 200), (2, 48000, 200000, 0, 200),
                                      (1, 48000, 24000, 0, 50), (2, 8000, 48000, 1, 400), (1, 48000, 40000, 0, 25)]:
        ref, emu = RefEncoder(fs, ch, enc_cases.VOIP), EmuEncoder(emu_enc_lib, fs, ch, enc_cases.VOIP)
        for e in (ref, emu):
            assert e.ctl(enc_cases.SET_BITRATE, rate) == 0
            if force:
                assert e.ctl(11002, 1002) == 0
        frame = fs * ms10 // 10000
        pcm = enc_cases.signal(enc_cases.KINDS[(ch + ms10 + rate // 1000) % len(enc_cases.KINDS)], frame * 10, ch, 21)
        pcm = (pcm.astype(np.int32) + 900).clip(-32768, 32767).astype(np.int16)       # a DC offset for the high-pass
        for t in range(10):
            if t == 5:
                assert ref.ctl(4028, 0) == emu.ctl(4028, 0) == 0                    # OPUS_RESET_STATE
            x = pcm[t * frame * ch:(t + 1) * frame * ch]
            a, b = ref.encode(x, frame, 1275), emu.encode(x, frame, 1275)
            assert a == b, (ch, fs, rate, force, ms10, t, a[0], b[0])
            assert a[0] > 0 and (a[1][0] & 0x80), "expected a CELT packet"


# ---- real audio: excerpts of the reference's AudioData, packets of the reference encoder -----------------------
REAL = {c.name: c for c in enc_cases.real_cases()}


def real_index():
    import json
    return json.load(open(os.path.join(GOLDEN, "enc_real_index.json")))


def test_real_audio_index_matches_case_list():
    import hashlib
    idx = real_index()
    assert [m["name"] for m in idx] == [c.name for c in enc_cases.real_cases()]
    for m in idx:
        assert hashlib.md5(enc_cases.real_pcm(m["ch"]).tobytes()).hexdigest() == m["pcm_md5"]


@pytest.mark.parametrize("name", sorted(REAL))
def test_emu_encoder_real_audio_matches_golden(emu_enc_lib, name):
    """Every frame of the real-audio excerpt: packet bytes and final range equal the reference encoder's
    (the procedure of CSharp/ParityTest/TestDriver.cs:253-275, incl. its random bitrate changes :219-224)."""
    c = REAL[name]
    gold = read_bit(os.path.join(GOLDEN, "encr_%s.bit" % name))
    pcm = enc_cases.real_pcm(c.ch)
    e = EmuEncoder(emu_enc_lib, 48000, c.ch, c.app)
    for req, val in c.ctls:
        assert e.ctl(req, val) == 0
    sched = c.bitrate_schedule()
    step = c.frame * c.ch
    for t in range(c.nframes):
        if t in sched:
            assert e.ctl(enc_cases.SET_BITRATE, sched[t]) == 0
        n, pkt, rng = e.encode(pcm[t * step:(t + 1) * step], c.frame, c.max_bytes)
        assert n == len(gold[t][0]), (name, t, n, len(gold[t][0]))
        assert pkt == gold[t][0], (name, t)
        assert rng == gold[t][1], (name, t)
