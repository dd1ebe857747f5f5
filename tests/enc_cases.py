"""Seeded synthetic PCM and the encoder parity cases shared by the CPU (host-emulation) and GPU tests and
by tests/golden/make_golden_enc.py.  Nothing here reads /root/reference."""
import hashlib

import numpy as np

LOWDELAY, AUDIO, VOIP = 2051, 2049, 2048
SET_BITRATE, SET_VBR, SET_COMPLEXITY, SET_VBR_CONSTRAINT = 4002, 4006, 4010, 4020
SET_FORCE_CHANNELS, SET_BANDWIDTH, SET_MAX_BANDWIDTH, SET_LOSS, SET_LSB_DEPTH = 4022, 4008, 4004, 4014, 4036
SET_PREDICTION_DISABLED, SET_FORCE_MODE, MODE_CELT_ONLY = 4042, 11002, 1002

KINDS = ["tone", "noise", "clicks", "chirp", "silence", "mix", "loud"]


def signal(kind, n, ch, seed, fs=48000):
    """n samples per channel, interleaved int16."""
    rng = np.random.default_rng(seed)
    t = np.arange(n) / fs
    if kind == "tone":
        x = 8000 * np.sin(2 * np.pi * 440 * t) + 3000 * np.sin(2 * np.pi * 1234.5 * t)
    elif kind == "noise":
        x = rng.normal(0, 3000, n)
    elif kind == "clicks":
        x = rng.normal(0, 100, n)
        for k in range(0, n, 3777):
            x[k:k + 40] += 20000 * rng.uniform(-1, 1, min(40, n - k))
    elif kind == "chirp":
        x = 10000 * np.sin(2 * np.pi * (100 * t + 2000 * t * t))
    elif kind == "silence":
        x = np.zeros(n)
    elif kind == "mix":
        x = 6000 * np.sin(2 * np.pi * 220 * t) * (np.sin(2 * np.pi * 3 * t) > 0) + rng.normal(0, 800, n)
        x[n // 3: n // 3 + 2000] = 0
    elif kind == "loud":
        x = rng.uniform(-32768, 32767, n)
    else:
        raise ValueError(kind)
    if ch == 2:
        y = 0.6 * x + 0.4 * np.roll(x, 37) + rng.normal(0, 200, n)
        x = np.stack([x, y], axis=1).reshape(-1)
    return np.clip(np.round(x), -32768, 32767).astype(np.int16)


def pcm_md5(pcm):
    return hashlib.md5(np.ascontiguousarray(pcm).tobytes()).hexdigest()


class EncCase(object):
    def __init__(self, name, fs, ch, app, frame, kind, nframes, ctls, max_bytes=1275, seed=1):
        self.name, self.fs, self.ch, self.app, self.frame = name, fs, ch, app, frame
        self.kind, self.nframes, self.ctls, self.max_bytes, self.seed = kind, nframes, ctls, max_bytes, seed

    def pcm(self):
        return signal(self.kind, self.frame * self.nframes, self.ch, self.seed, self.fs)

    def frames(self):
        p = self.pcm()
        step = self.frame * self.ch
        return [np.ascontiguousarray(p[i * step:(i + 1) * step]) for i in range(self.nframes)]


def golden_cases():
    """The committed golden set: small, but touching every frame size, both channel counts, CBR / VBR /
    CVBR, several complexities and sample rates, forced mono, narrow bandwidths and tiny buffers."""
    c = []
    k = 0
    for ch in (1, 2):
        for frame in (960, 480, 240, 120):
            kind = KINDS[k % len(KINDS)]; k += 1
            c.append(EncCase("ld_%dch_%d_%s" % (ch, frame, kind), 48000, ch, LOWDELAY, frame, kind, 12,
                             [(SET_BITRATE, 64000 * ch)]))
    c.append(EncCase("cbr_2ch_960_mix", 48000, 2, LOWDELAY, 960, "mix", 12, [(SET_BITRATE, 96000), (SET_VBR, 0)]))
    c.append(EncCase("cbr_1ch_240_clicks", 48000, 1, LOWDELAY, 240, "clicks", 16, [(SET_BITRATE, 32000), (SET_VBR, 0)]))
    c.append(EncCase("vbr_2ch_960_clicks", 48000, 2, LOWDELAY, 960, "clicks", 12,
                     [(SET_BITRATE, 96000), (SET_VBR, 1), (SET_VBR_CONSTRAINT, 0)]))
    c.append(EncCase("audio_forced_2ch_960_mix", 48000, 2, AUDIO, 960, "mix", 12,
                     [(SET_BITRATE, 96000), (SET_FORCE_MODE, MODE_CELT_ONLY)]))
    c.append(EncCase("audio_auto_2ch_960_tone", 48000, 2, AUDIO, 960, "tone", 12, [(SET_BITRATE, 128000)]))
    c.append(EncCase("audio_forced_1ch_480_noise", 48000, 1, AUDIO, 480, "noise", 12,
                     [(SET_BITRATE, 48000), (SET_FORCE_MODE, MODE_CELT_ONLY)]))
    for cx in (0, 2, 4, 7, 10):
        c.append(EncCase("cx%d_2ch_960_mix" % cx, 48000, 2, LOWDELAY, 960, "mix", 10,
                         [(SET_BITRATE, 80000), (SET_COMPLEXITY, cx)]))
    for fs in (8000, 16000, 24000):
        c.append(EncCase("fs%d_1ch_20ms_chirp" % fs, fs, 1, LOWDELAY, fs // 50, "chirp", 12, [(SET_BITRATE, 24000)]))
    c.append(EncCase("fs12000_2ch_10ms_noise", 12000, 2, LOWDELAY, 120, "noise", 12, [(SET_BITRATE, 40000)]))
    c.append(EncCase("forcemono_2ch_960_mix", 48000, 2, LOWDELAY, 960, "mix", 12,
                     [(SET_BITRATE, 64000), (SET_FORCE_CHANNELS, 1)]))
    c.append(EncCase("lowrate_2ch_960_mix", 48000, 2, LOWDELAY, 960, "mix", 20, [(SET_BITRATE, 24000)]))
    c.append(EncCase("bw_nb_1ch_960_tone", 48000, 1, LOWDELAY, 960, "tone", 10, [(SET_BITRATE, 32000), (SET_BANDWIDTH, 1101)]))
    c.append(EncCase("bw_swb_2ch_480_loud", 48000, 2, LOWDELAY, 480, "loud", 10, [(SET_BITRATE, 64000), (SET_MAX_BANDWIDTH, 1104)]))
    c.append(EncCase("loss10_2ch_960_chirp", 48000, 2, LOWDELAY, 960, "chirp", 12, [(SET_BITRATE, 64000), (SET_LOSS, 10)]))
    c.append(EncCase("nopred_2ch_960_mix", 48000, 2, LOWDELAY, 960, "mix", 10, [(SET_BITRATE, 64000), (SET_PREDICTION_DISABLED, 1)]))
    c.append(EncCase("tiny8_1ch_960_noise", 48000, 1, LOWDELAY, 960, "noise", 8, [(SET_BITRATE, 64000)], max_bytes=8))
    c.append(EncCase("tiny2_cbr_2ch_120_noise", 48000, 2, LOWDELAY, 120, "noise", 8, [(SET_BITRATE, 64000), (SET_VBR, 0)], max_bytes=2))
    c.append(EncCase("max_2ch_960_loud", 48000, 2, LOWDELAY, 960, "loud", 8, [(SET_BITRATE, 510000)]))
    c.append(EncCase("silence_2ch_960", 48000, 2, LOWDELAY, 960, "silence", 6, [(SET_BITRATE, 64000)]))
    c.append(EncCase("silence_1ch_480", 48000, 1, LOWDELAY, 480, "silence", 6, [(SET_BITRATE, 64000)]))
    c.append(EncCase("silence_cbr_1ch_960", 48000, 1, LOWDELAY, 960, "silence", 6, [(SET_BITRATE, 32000), (SET_VBR, 0)]))
    return c


def sweep_cases(seed=5, count=60):
    """A wider randomized sweep, checked against the compiled reference when it is present."""
    import random
    r = random.Random(seed)
    out = []
    for i in range(count):
        fs = r.choice([48000, 48000, 48000, 24000, 16000, 12000, 8000])
        ch = r.choice([1, 2])
        frame = fs // r.choice([50, 100, 200, 400])
        ctls = [(SET_BITRATE, r.choice([6000, 12000, 24000, 32000, 48000, 64000, 96000, 128000, 256000, 510000]))]
        if r.random() < 0.3:
            ctls.append((SET_VBR, 0))
        elif r.random() < 0.3:
            ctls.append((SET_VBR_CONSTRAINT, 0))
        if r.random() < 0.4:
            ctls.append((SET_COMPLEXITY, r.randrange(0, 11)))
        app = r.choice([LOWDELAY, LOWDELAY, AUDIO])
        if app == AUDIO:
            ctls.append((SET_FORCE_MODE, MODE_CELT_ONLY))
        out.append(EncCase("sweep%d" % i, fs, ch, app, frame, r.choice(KINDS), 12, ctls,
                           max_bytes=r.choice([1275, 1275, 400, 60, 20]), seed=100 + i))
    return out


# ---- real audio (tests/golden/audio_48k_*.pcm, excerpts of the reference's AudioData; make_golden_enc_real.py) ----
REAL_FRAMES, REAL_SPF = 250, 960


class RealCase(object):
    """20 ms frames of the committed real-audio excerpt through one encoder configuration.  varrate_seed != 0:
    the bitrate is re-drawn on about every tenth frame, like the reference's ParityTest driver does
    (CSharp/ParityTest/TestDriver.cs:219-224: random.Next(6, 510) * 1024), from a fixed seed."""

    def __init__(self, name, ch, app, ctls, nframes=REAL_FRAMES, varrate_seed=0, max_bytes=1275):
        self.name, self.ch, self.app, self.ctls, self.nframes = name, ch, app, ctls, nframes
        self.varrate_seed, self.max_bytes = varrate_seed, max_bytes
        self.fs, self.frame = 48000, REAL_SPF

    def bitrate_schedule(self):
        """{frame index: new bitrate}"""
        if not self.varrate_seed:
            return {}
        rng = np.random.default_rng(self.varrate_seed)
        out = {}
        for t in range(self.nframes):
            if rng.random() < 0.1:
                out[t] = int(rng.integers(6, 510)) * 1024
        return out


def real_pcm(ch):
    import os
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    return np.fromfile(os.path.join(here, "audio_48k_%s_5s.pcm" % ("stereo" if ch == 2 else "mono")), dtype="<i2")


def real_cases():
    cx10 = (SET_COMPLEXITY, 10)
    return [
        RealCase("st_128k_cbr", 2, LOWDELAY, [(SET_BITRATE, 128000), (SET_VBR, 0), cx10]),
        RealCase("st_128k_vbr", 2, LOWDELAY, [(SET_BITRATE, 128000), (SET_VBR, 1), (SET_VBR_CONSTRAINT, 0), cx10]),
        RealCase("st_96k_cvbr", 2, LOWDELAY, [(SET_BITRATE, 96000), cx10]),
        RealCase("st_parity", 2, AUDIO, [(SET_BITRATE, 128 * 1024), (SET_FORCE_MODE, MODE_CELT_ONLY), cx10]),  # ParityTest flavour
        RealCase("st_varrate", 2, AUDIO, [(SET_BITRATE, 64 * 1024), (SET_FORCE_MODE, MODE_CELT_ONLY), cx10], varrate_seed=11),
        RealCase("st_cx5_64k", 2, LOWDELAY, [(SET_BITRATE, 64000), (SET_COMPLEXITY, 5)], nframes=120),
        RealCase("mono_64k_cbr", 1, LOWDELAY, [(SET_BITRATE, 64000), (SET_VBR, 0), cx10]),
        RealCase("mono_parity", 1, AUDIO, [(SET_BITRATE, 64 * 1024), (SET_FORCE_MODE, MODE_CELT_ONLY), cx10]),
        RealCase("mono_varrate", 1, LOWDELAY, [(SET_BITRATE, 48 * 1024), cx10], varrate_seed=5),
    ]


def c3_rotations(n, seed=1234):
    """stream i plays the excerpt rotated by a whole number of 20 ms frames drawn from a fixed-seed
    LCG (the reference's celt_lcg_rand constants); returns the start frame of every stream."""
    rot = np.zeros(n, dtype=np.int64)
    s = seed & 0xFFFFFFFF
    for i in range(n):
        s = (1664525 * s + 1013904223) & 0xFFFFFFFF
        rot[i] = (s >> 8) % REAL_FRAMES
    return rot


def c3_rows(pcm, ch, rot, t):
    """int16 [n, 960*ch]: frame t of every stream (frame (rot[i] + t) mod 250 of the excerpt)."""
    frames = pcm.reshape(REAL_FRAMES, REAL_SPF * ch)
    return np.ascontiguousarray(frames[(rot + t) % REAL_FRAMES])
