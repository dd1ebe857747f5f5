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
