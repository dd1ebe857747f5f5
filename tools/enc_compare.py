"""Developer tool: encode the same synthetic PCM with the compiled reference (oracle/_ref) and the
host-emulation build of the encoder headers, and report the first differing packet."""
import ctypes, sys, os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ref = ctypes.CDLL(os.path.join(ROOT, "oracle/_ref/libopus_ref.so"))
emu = ctypes.CDLL(os.path.join(ROOT, "oracle/_build/libemu_enc.so"))
ref.opus_encoder_create.restype = ctypes.c_void_p
ref.opus_encoder_create.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
ref.opus_encode.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
ref.opus_encoder_ctl.argtypes = [ctypes.c_void_p, ctypes.c_int]
emu.emu_enc_create.restype = ctypes.c_void_p
emu.emu_enc_create.argtypes = [ctypes.c_int] * 3
emu.emu_enc_encode.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
emu.emu_enc_ctl.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
emu.emu_enc_final_range.argtypes = [ctypes.c_void_p]
emu.emu_enc_final_range.restype = ctypes.c_uint


def signal(kind, n, ch, seed, fs=48000):
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


def run(fs, ch, app, frame, kind, nframes, ctls, max_bytes=1500, seed=1, verbose=True):
    err = ctypes.c_int(0)
    r = ref.opus_encoder_create(fs, ch, app, ctypes.byref(err))
    e = emu.emu_enc_create(fs, ch, app)
    for req, val in ctls:
        a = ref.opus_encoder_ctl(ctypes.c_void_p(r), req, ctypes.c_int(val))
        b = emu.emu_enc_ctl(e, req, val, None)
        assert a == b, ("ctl", req, val, a, b)
    pcm = signal(kind, frame * nframes, ch, seed, fs)
    bufr = (ctypes.c_ubyte * max_bytes)()
    bufe = (ctypes.c_ubyte * max_bytes)()
    fr = ctypes.c_uint(0)
    for i in range(nframes):
        chunk = np.ascontiguousarray(pcm[i * frame * ch:(i + 1) * frame * ch])
        lr = ref.opus_encode(r, chunk.ctypes.data, frame, bufr, max_bytes)
        le = emu.emu_enc_encode(e, chunk.ctypes.data, frame, bufe, max_bytes)
        ref.opus_encoder_ctl(ctypes.c_void_p(r), 4031, ctypes.byref(fr))
        fe = emu.emu_enc_final_range(e)
        if lr != le or bytes(bufr[:max(lr, 0)]) != bytes(bufe[:max(le, 0)]) or fr.value != fe:
            if verbose:
                print(f"MISMATCH frame {i}: len ref {lr} emu {le} rng {fr.value:08x} {fe:08x}")
                if lr > 0 and le > 0:
                    a, b = bytes(bufr[:lr]), bytes(bufe[:le])
                    k = next((j for j in range(min(lr, le)) if a[j] != b[j]), min(lr, le))
                    print("  first differing byte", k, a[:8].hex(), b[:8].hex())
            return i
    return -1


if __name__ == "__main__":
    LOWDELAY, AUDIO = 2051, 2049
    bad = 0
    cases = []
    for kind in ["tone", "noise", "clicks", "chirp", "silence", "mix", "loud"]:
        for ch in (1, 2):
            for frame in (960, 480, 240, 120):
                cases.append((48000, ch, LOWDELAY, frame, kind, 30, [(4002, 64000 * ch)]))
    for c in cases:
        res = run(*c)
        print(("ok  " if res < 0 else "FAIL"), c[:5], c[6])
        bad += res >= 0
        if bad > 3:
            break
    print("failures", bad)
