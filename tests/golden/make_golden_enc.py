"""Generate the encoder golden packets with the reference encoder compiled under oracle/_ref.

Run in the build container (needs oracle/_ref/libopus_ref.so, i.e. `make -C oracle ref`):
    python tests/golden/make_golden_enc.py
Writes tests/golden/enc_<case>.bit (opus_demo framing: be32 length, be32 final range, payload) and
tests/golden/enc_index.json (case parameters + md5 of the synthetic PCM that was encoded)."""
import ctypes
import json
import os
import struct
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import enc_cases  # noqa: E402


def main():
    ref = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libopus_ref.so"))
    ref.opus_encoder_create.restype = ctypes.c_void_p
    ref.opus_encoder_create.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    ref.opus_encode.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    ref.opus_encoder_ctl.argtypes = [ctypes.c_void_p, ctypes.c_int]
    ref.opus_encoder_destroy.argtypes = [ctypes.c_void_p]
    index = []
    for c in enc_cases.golden_cases():
        err = ctypes.c_int(0)
        st = ref.opus_encoder_create(c.fs, c.ch, c.app, ctypes.byref(err))
        assert st and err.value == 0
        for req, val in c.ctls:
            assert ref.opus_encoder_ctl(st, req, ctypes.c_int(val)) == 0
        buf = (ctypes.c_ubyte * c.max_bytes)()
        fr = ctypes.c_uint(0)
        lens = []
        with open(os.path.join(HERE, "enc_%s.bit" % c.name), "wb") as f:
            for frame in c.frames():
                n = ref.opus_encode(st, frame.ctypes.data, c.frame, buf, c.max_bytes)
                assert n > 0, (c.name, n)
                ref.opus_encoder_ctl(st, 4031, ctypes.byref(fr))
                f.write(struct.pack(">II", n, fr.value))
                f.write(bytes(buf[:n]))
                lens.append(n)
        ref.opus_encoder_destroy(st)
        index.append(dict(name=c.name, fs=c.fs, ch=c.ch, app=c.app, frame=c.frame, kind=c.kind, nframes=c.nframes,
                          ctls=c.ctls, max_bytes=c.max_bytes, seed=c.seed, pcm_md5=enc_cases.pcm_md5(c.pcm()),
                          bytes=sum(lens)))
        print(c.name, lens[:6])
    with open(os.path.join(HERE, "enc_index.json"), "w") as f:
        json.dump(index, f, indent=1)


if __name__ == "__main__":
    main()
