"""Developer tool: SILK-only packets made by the compiled reference encoder, decoded by the compiled
reference decoder and by the host-emulation build of the decoder headers; reports the first mismatch."""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import enc_cases
from conftest import RefEncoder, RefDecoder

emu = ctypes.CDLL(os.path.join(ROOT, "tests/emu/libemu_celt.so"))
emu.emu_opus_create.restype = ctypes.c_void_p
emu.emu_opus_create.argtypes = [ctypes.c_int, ctypes.c_int]
emu.emu_opus_decode.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
emu.emu_opus_final_range.argtypes = [ctypes.c_void_p]
emu.emu_opus_final_range.restype = ctypes.c_uint


def speechlike(n, ch, seed, fs):
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


def make_packets(fs_enc, ch, frame_ms, bitrate, nframes, ctls=(), seed=1, app=2048):
    enc = RefEncoder(fs_enc, ch, app)
    enc.ctl(4002, bitrate)
    for r, v in ctls:
        assert enc.ctl(r, v) == 0
    frame = fs_enc * frame_ms // 1000
    pcm = speechlike(frame * nframes, ch, seed, fs_enc)
    out = []
    for i in range(nframes):
        n, pkt, rng = enc.encode(pcm[i * frame * ch:(i + 1) * frame * ch], frame, 1275)
        assert n > 0
        out.append((pkt, rng))
    return out


def compare(packets, fs_dec, ch_dec, frame_ms, loss=None, fec=False, verbose=True):
    ref = RefDecoder(fs_dec, ch_dec)
    e = emu.emu_opus_create(fs_dec, ch_dec)
    fsz = fs_dec * frame_ms // 1000
    rng = np.random.default_rng(7)
    prev_lost = False
    for i, (pkt, _) in enumerate(packets):
        lost = loss is not None and rng.random() < loss
        if lost:
            calls = [(None, 0)]
        elif fec and prev_lost:
            calls = [(pkt, 1), (pkt, 0)]
        else:
            calls = [(pkt, 0)]
        if fec and lost:
            calls = []          # decided at the next packet
            prev_lost = True
            continue
        if loss is not None and not fec:
            pass
        for data, f in calls:
            r1, p1 = ref.decode_raw(data, fsz, f)
            p2 = np.zeros((fsz, ch_dec), dtype=np.int16)
            r2 = emu.emu_opus_decode(e, data, len(data) if data else 0, p2.ctypes.data, fsz, f)
            fr1, fr2 = ref.final_range, emu.emu_opus_final_range(e)
            if r1 != r2 or (r1 > 0 and not np.array_equal(p1[:r1], p2[:r1])) or fr1 != fr2:
                if verbose:
                    bad = np.nonzero((p1[:max(r1, 0)] != p2[:max(r1, 0)]).any(axis=1))[0] if r1 > 0 and r1 == r2 else []
                    print("MISMATCH packet", i, "fec", f, "ret", r1, r2, "rng %08x %08x" % (fr1, fr2),
                          "first bad sample", bad[0] if len(bad) else None, "toc %02x" % (pkt[0] if pkt else 0))
                return i
        prev_lost = lost
    return -1


if __name__ == "__main__":
    bad = 0
    for fs_enc, bw in ((16000, "WB"), (12000, "MB"), (8000, "NB")):
        for ch in (1, 2):
            for ms in (20, 10, 40, 60):
                pk = make_packets(fs_enc, ch, ms, 16000 * ch + fs_enc, 40)
                modes = set(p[0][0] >> 3 for p in pk)
                for fs_dec in (fs_enc, 48000, 8000, 24000):
                    for chd in (1, 2):
                        r = compare(pk, fs_dec, chd, ms)
                        print("ok  " if r < 0 else "FAIL", bw, "ch", ch, ms, "ms -> dec", fs_dec, chd, "tocs", sorted(modes))
                        bad += r >= 0
                        if bad > 4:
                            sys.exit(1)
    print("failures", bad)
