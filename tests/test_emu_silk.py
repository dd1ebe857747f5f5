"""CPU-side parity of the SILK / hybrid decode path (host build of the kernel sources): SILK-only and hybrid
packets, mode transitions and redundant CELT frames made by the reference encoder from the reference's test
audio, replayed with losses / in-band FEC / DTX, against
the committed goldens (PCM CRC32, return value and final range per decode call) and, where oracle/_ref is
built, against the reference decoder on synthetic voice at every bandwidth / frame size / output rate."""
import os

import numpy as np
import pytest

from conftest import (GOLDEN, EmuOpusDecoder, RefDecoder, RefEncoder, crc32, needs_ref, read_bit, silk_golden_index)
from scenarios import silk_calls, silk_fuzz_calls

CASES = sorted(silk_golden_index().keys())


def replay(dec, meta, pk):
    """Run the fixture's decode-call sequence on `dec`; yields (call number, ret, pcm, final range)."""
    for k, (idx, use_fec) in enumerate(silk_calls(len(pk), meta["loss"], meta["fec"], seed=11)):
        data = pk[idx][0] if idx is not None else None
        ret, pcm = dec.decode_raw(data, meta["samples_per_frame"], use_fec)
        yield k, ret, pcm, dec.final_range


@pytest.mark.parametrize("name", CASES)
def test_emu_silk_matches_golden(emu_lib, name):
    meta = silk_golden_index()[name]
    pk = read_bit(os.path.join(GOLDEN, name + ".bit"))
    dec = EmuOpusDecoder(emu_lib, meta["fs"], meta["channels"])
    n = 0
    for k, ret, pcm, rng in replay(dec, meta, pk):
        assert ret == meta["ret"][k], (name, k, ret)
        assert crc32(pcm[:ret]) == meta["pcm_crc32"][k], (name, "pcm of call", k)
        assert rng == meta["final_range"][k], (name, k)
        n += 1
    assert n == len(meta["ret"])


def voice(n, ch, seed, fs):
    rng = np.random.default_rng(seed)
    t = np.arange(n) / fs
    f0 = 120 + 40 * np.sin(2 * np.pi * 0.7 * t)
    ph = 2 * np.pi * np.cumsum(f0) / fs
    voiced = sum(np.sin(k * ph) / k for k in range(1, 12)) * 4000
    env = (np.sin(2 * np.pi * 1.3 * t) > -0.2) * (0.3 + 0.7 * np.abs(np.sin(2 * np.pi * 2.1 * t)))
    x = voiced * env + rng.normal(0, 300, n) * (1 - env) * 3
    x[n // 2: n // 2 + fs // 3] = 0
    if ch == 2:
        y = 0.7 * x + 0.3 * np.roll(x, 11) + rng.normal(0, 50, n)
        x = np.stack([x, y], axis=1).reshape(-1)
    return np.clip(np.round(x), -32768, 32767).astype(np.int16)


def silk_packets(fs_enc, ch, ms, bitrate, nframes, ctls=(), seed=1):
    enc = RefEncoder(fs_enc, ch, 2048)
    assert enc.ctl(4002, bitrate) == 0
    for r, v in ctls:
        assert enc.ctl(r, v) == 0
    fr = fs_enc * ms // 1000
    pcm = voice(fr * nframes, ch, seed, fs_enc)
    return [enc.encode(pcm[i * fr * ch:(i + 1) * fr * ch], fr, 1275)[1] for i in range(nframes)]


@needs_ref
@pytest.mark.parametrize("fs_enc", [16000, 12000, 8000])
def test_emu_silk_matches_reference_all_shapes(emu_lib, fs_enc):
    for ch in (1, 2):
        for ms in (10, 20, 40, 60):
            pk = silk_packets(fs_enc, ch, ms, 14000 * ch + fs_enc // 2, 25, ctls=[(4012, 1), (4014, 25)])
            for fs_dec, chd in ((fs_enc, ch), (48000, 3 - ch), (8000, ch), (24000, 2), (12000, 1), (16000, 2)):
                for loss, fec in ((0.0, False), (0.25, False), (0.25, True)):
                    ref, emu = RefDecoder(fs_dec, chd), EmuOpusDecoder(emu_lib, fs_dec, chd)
                    fsz = fs_dec * ms // 1000
                    for idx, use_fec in silk_calls(len(pk), loss, fec, seed=5):
                        data = pk[idx] if idx is not None else None
                        r1, p1 = ref.decode_raw(data, fsz, use_fec)
                        r2, p2 = emu.decode_raw(data, fsz, use_fec)
                        assert r1 == r2, (fs_enc, ch, ms, fs_dec, chd, loss, fec, idx)
                        assert np.array_equal(p1[:r1], p2[:r1]), (fs_enc, ch, ms, fs_dec, chd, loss, fec, idx)
                        assert ref.final_range == emu.final_range


@needs_ref
def test_emu_silk_bandwidth_and_channel_switches(emu_lib):
    """One decoder fed NB, then WB, then MB, then stereo, then mono SILK packets (resampler re-init, stereo
    state transitions), with a few concealed frames at the joints."""
    segs = [silk_packets(8000, 1, 20, 12000, 12, seed=2), silk_packets(16000, 1, 20, 24000, 12, seed=3),
            silk_packets(12000, 1, 20, 18000, 12, seed=4), silk_packets(16000, 2, 20, 40000, 12, seed=5),
            silk_packets(16000, 1, 20, 24000, 12, seed=6), silk_packets(8000, 2, 20, 20000, 12, seed=7)]
    for fs_dec, chd in ((48000, 2), (16000, 1), (24000, 2)):
        ref, emu = RefDecoder(fs_dec, chd), EmuOpusDecoder(emu_lib, fs_dec, chd)
        fsz = fs_dec // 50
        k = 0
        for seg in segs:
            for p in seg + [None, None]:
                r1, p1 = ref.decode_raw(p, fsz)
                r2, p2 = emu.decode_raw(p, fsz)
                assert r1 == r2 and np.array_equal(p1[:r1], p2[:r1]), (fs_dec, chd, k)
                assert ref.final_range == emu.final_range
                k += 1


FORCE_MODE, SET_BANDWIDTH = 11002, 4008
MODE = {"silk": 1000, "hybrid": 1001, "celt": 1002, "auto": -1000}
TRANSITION_SCRIPTS = {
    "hybrid": [(20, "hybrid", None)],
    "silk>celt>silk": [(7, "silk", 1103), (7, "celt", None), (7, "silk", 1101), (5, "celt", 1103)],
    "hybrid>celt>hybrid": [(6, "hybrid", None), (6, "celt", None), (6, "hybrid", 1104), (5, "celt", 1104)],
    "silk>hybrid>silk": [(6, "silk", 1103), (6, "hybrid", None), (6, "silk", 1102), (6, "hybrid", 1104)],
    "mixed": [(4, "celt", None), (5, "silk", 1101), (3, "hybrid", 1105), (4, "celt", 1102), (5, "hybrid", 1104),
              (3, "silk", 1103), (3, "celt", 1105)],
}


def scripted_packets(ch, ms, script, bitrate, seed=1, app=2049, ctls=()):
    """48 kHz reference encoder with its mode (and bandwidth) forced per segment: hybrid packets, mode
    transitions and the redundant CELT frames that go with them."""
    enc = RefEncoder(48000, ch, app)
    assert enc.ctl(4002, bitrate) == 0
    for r, v in ctls:
        assert enc.ctl(r, v) == 0
    fr = 48 * ms
    pcm = voice(fr * sum(n for n, _, _ in script), ch, seed, 48000)
    out, k = [], 0
    for n, mode, bw in script:
        assert enc.ctl(FORCE_MODE, MODE[mode]) == 0
        assert enc.ctl(SET_BANDWIDTH, bw if bw else -1000) == 0
        for _ in range(n):
            out.append(enc.encode(pcm[k * fr * ch:(k + 1) * fr * ch], fr, 1275)[1])
            k += 1
    return out


@needs_ref
@pytest.mark.parametrize("script", sorted(TRANSITION_SCRIPTS))
def test_emu_hybrid_and_transitions_match_reference(emu_lib, script):
    for ch in (1, 2):
        for ms in (20, 10, 40):
            pk = scripted_packets(ch, ms, TRANSITION_SCRIPTS[script], 28000 * ch, ctls=[(4012, 1), (4014, 20)])
            for fs_dec, chd in ((48000, ch), (16000, 3 - ch), (24000, 2), (8000, 1)):
                for loss, fec in ((0.0, False), (0.2, False), (0.2, True)):
                    ref, emu = RefDecoder(fs_dec, chd), EmuOpusDecoder(emu_lib, fs_dec, chd)
                    fsz = fs_dec * ms // 1000
                    for idx, use_fec in silk_calls(len(pk), loss, fec, seed=9):
                        data = pk[idx] if idx is not None else None
                        r1, p1 = ref.decode_raw(data, fsz, use_fec)
                        r2, p2 = emu.decode_raw(data, fsz, use_fec)
                        assert r1 == r2, (script, ch, ms, fs_dec, chd, loss, fec, idx)
                        assert np.array_equal(p1[:r1], p2[:r1]), (script, ch, ms, fs_dec, chd, loss, fec, idx)
                        assert ref.final_range == emu.final_range, (script, ch, ms, fs_dec, chd, loss, fec, idx)


@needs_ref
@pytest.mark.parametrize("name", ["silk_wb_mono_32k", "silk_wb_st_40k", "silk_nb_st_40ms", "silk_wb_mono_60ms",
                                  "hyb_fb_mono_32k", "hyb_swb_st_48k", "trans_all_st_to24", "trans_all_mono_10ms_loss"])
@pytest.mark.parametrize("random_toc", [False, True])
def test_emu_silk_corrupted_packets_follow_reference(emu_lib, name, random_toc):
    """Damaged, truncated and random payloads (optionally with random TOC bytes, which throw the decoder
    between SILK, hybrid and CELT on garbage): return value, PCM and final range must still equal the
    reference's for every call -- the reference decodes garbage deterministically."""
    meta = silk_golden_index()[name]
    pk = [p for p, _ in read_bit(os.path.join(GOLDEN, name + ".bit"))]
    calls = silk_fuzz_calls(pk, seed=len(name) * 3 + random_toc, random_toc=random_toc)
    ref, emu = RefDecoder(meta["fs"], meta["channels"]), EmuOpusDecoder(emu_lib, meta["fs"], meta["channels"])
    # a frame size that holds any one-frame packet of any configuration (60 ms)
    fsz = meta["fs"] * 60 // 1000
    for k, data in enumerate(calls):
        want = fsz if data is not None else meta["samples_per_frame"]
        r1, p1 = ref.decode_raw(data, want)
        r2, p2 = emu.decode_raw(data, want)
        assert r1 == r2, (name, k, r1, r2)
        if r1 > 0:
            assert np.array_equal(p1[:r1], p2[:r1]), (name, k)
        assert ref.final_range == emu.final_range, (name, k)
