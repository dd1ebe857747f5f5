"""Real-audio encoder fixtures: excerpts of the reference's own test audio and the packets the reference
encoder (oracle/_ref/libopus_ref.so) makes from them.

Run in the build container (needs /root/reference/AudioData and `make -C oracle ref`):
    python tests/golden/make_golden_enc_real.py
Writes
    tests/golden/audio_48k_stereo_5s.pcm   250 x 20 ms of AudioData/48Khz Stereo.raw (from 2.0 s), int16 LE interleaved
    tests/golden/audio_48k_mono_5s.pcm     the same span of AudioData/48Khz Mono.raw
    tests/golden/encr_<case>.bit           opus_demo framing (be32 length, be32 final range, payload) per case
    tests/golden/enc_real_index.json       case parameters, md5 of the PCM, total bytes
The cases follow SURVEY 8d C3 (128 kbit/s CBR, complexity 10, 20 ms) and the reference's own ParityTest flavour
(CSharp/ParityTest/TestDriver.cs: OPUS_APPLICATION_AUDIO + forced CELT mode, bitrate 128*1024, VBR), plus the
schedule of random bitrate changes that driver applies (:219-224), here from a fixed seed."""
import ctypes
import hashlib
import json
import os
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import enc_cases  # noqa: E402

AUDIO_DIR = "/root/reference/AudioData"
START_S, FRAMES, SPF = 2.0, 250, 960


def excerpt(name, ch):
    raw = np.fromfile(os.path.join(AUDIO_DIR, name), dtype="<i2")
    a = int(START_S * 48000) * ch
    return np.ascontiguousarray(raw[a:a + FRAMES * SPF * ch])


def main():
    ref = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libopus_ref.so"))
    ref.opus_encoder_create.restype = ctypes.c_void_p
    ref.opus_encoder_create.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    ref.opus_encode.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    ref.opus_encoder_ctl.argtypes = [ctypes.c_void_p, ctypes.c_int]
    ref.opus_encoder_destroy.argtypes = [ctypes.c_void_p]
    pcm = {2: excerpt("48Khz Stereo.raw", 2), 1: excerpt("48Khz Mono.raw", 1)}
    pcm[2].tofile(os.path.join(HERE, "audio_48k_stereo_5s.pcm"))
    pcm[1].tofile(os.path.join(HERE, "audio_48k_mono_5s.pcm"))
    index = []
    for c in enc_cases.real_cases():
        err = ctypes.c_int(0)
        st = ref.opus_encoder_create(48000, c.ch, c.app, ctypes.byref(err))
        assert st and err.value == 0
        for req, val in c.ctls:
            assert ref.opus_encoder_ctl(st, req, ctypes.c_int(val)) == 0
        buf = (ctypes.c_ubyte * c.max_bytes)()
        fr = ctypes.c_uint(0)
        lens = []
        sched = c.bitrate_schedule()
        with open(os.path.join(HERE, "encr_%s.bit" % c.name), "wb") as f:
            for t in range(c.nframes):
                if t in sched:
                    assert ref.opus_encoder_ctl(st, enc_cases.SET_BITRATE, ctypes.c_int(sched[t])) == 0
                frame = np.ascontiguousarray(pcm[c.ch][t * SPF * c.ch:(t + 1) * SPF * c.ch])
                n = ref.opus_encode(st, frame.ctypes.data, SPF, buf, c.max_bytes)
                assert n > 0, (c.name, n)
                assert buf[0] & 0x80, "the reference picked SILK / hybrid for %s frame %d" % (c.name, t)
                ref.opus_encoder_ctl(st, 4031, ctypes.byref(fr))
                f.write(struct.pack(">II", n, fr.value))
                f.write(bytes(buf[:n]))
                lens.append(n)
        ref.opus_encoder_destroy(st)
        index.append(dict(name=c.name, ch=c.ch, app=c.app, nframes=c.nframes, ctls=c.ctls, max_bytes=c.max_bytes,
                          varrate_seed=c.varrate_seed,
                          pcm_md5=hashlib.md5(pcm[c.ch].tobytes()).hexdigest(), bytes=sum(lens)))
        print(c.name, lens[:8], "total", sum(lens))
    with open(os.path.join(HERE, "enc_real_index.json"), "w") as f:
        json.dump(index, f, indent=1)


if __name__ == "__main__":
    main()
