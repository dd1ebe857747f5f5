#!/usr/bin/env python3
"""Golden fixture for 5.1 multistream decode (BASELINE config 5), from the UNMODIFIED reference.

Uses oracle/_ref/libopus_ref.so through ctypes: opus_multistream_surround_encoder_create(48000, 6,
mapping family 1) -> packets from AudioData/48Khz 5.1.raw (960-sample frames, 256 kbps, complexity
10, every elementary encoder forced to CELT with OPUS_SET_FORCE_MODE as the reference's own parity
harness does, CSharp/ParityTest/TestDriver.cs:425-470), then opus_multistream_decode with a seeded
5 % loss mask.  Keeps packets (.bit container, final range = decoder's) and per-packet PCM CRC32.
Build container only (needs /root/reference)."""
import ctypes, json, os, struct, zlib
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.normpath(os.path.join(HERE, "..", ".."))
lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libopus_ref.so"))
AUDIO = "/root/reference/AudioData/48Khz 5.1.raw"
FRAMES = 200


def gen(name, bitrate, force_celt, application, index):
    pcm = np.fromfile(AUDIO, dtype=np.int16).reshape(-1, 6)
    streams, coupled = ctypes.c_int(0), ctypes.c_int(0)
    mapping = (ctypes.c_ubyte * 6)()
    err = ctypes.c_int(0)
    lib.opus_multistream_surround_encoder_create.restype = ctypes.c_void_p
    enc = lib.opus_multistream_surround_encoder_create(48000, 6, 1, ctypes.byref(streams), ctypes.byref(coupled),
                                                       mapping, application, ctypes.byref(err))
    assert enc and err.value == 0
    enc = ctypes.c_void_p(enc)
    lib.opus_multistream_encoder_ctl(enc, 4002, ctypes.c_int(bitrate))       # OPUS_SET_BITRATE
    lib.opus_multistream_encoder_ctl(enc, 4010, ctypes.c_int(10))           # OPUS_SET_COMPLEXITY
    for s in range(streams.value if force_celt else 0):                     # force CELT on every stream
        sub = ctypes.c_void_p(0)
        assert lib.opus_multistream_encoder_ctl(enc, 5120, ctypes.c_int(s), ctypes.byref(sub)) == 0
        assert lib.opus_encoder_ctl(sub, 11002, ctypes.c_int(1002)) == 0    # OPUS_SET_FORCE_MODE(MODE_CELT_ONLY)
    lib.opus_multistream_decoder_create.restype = ctypes.c_void_p
    dec = ctypes.c_void_p(lib.opus_multistream_decoder_create(48000, 6, streams.value, coupled.value, mapping,
                                                              ctypes.byref(err)))
    assert dec and err.value == 0
    out = (ctypes.c_ubyte * 4000)()
    rng = np.random.RandomState(1234)
    packets, crcs, rets = [], [], []
    for t in range(FRAMES):
        frame = np.ascontiguousarray(pcm[t * 960:(t + 1) * 960])
        n = lib.opus_multistream_encode(enc, frame.ctypes.data_as(ctypes.c_void_p), 960, out, 4000)
        assert n > 0
        pkt = bytes(out[:n])
        lost = rng.rand() < 0.05 and t > 3
        dpcm = np.zeros((960, 6), dtype=np.int16)
        if lost:
            r = lib.opus_multistream_decode(dec, None, 0, dpcm.ctypes.data_as(ctypes.c_void_p), 960, 0)
            pkt = b""
        else:
            r = lib.opus_multistream_decode(dec, pkt, n, dpcm.ctypes.data_as(ctypes.c_void_p), 960, 0)
        assert r == 960, r
        fr = ctypes.c_uint32(0)
        lib.opus_multistream_decoder_ctl(dec, 4031, ctypes.byref(fr))
        packets.append((pkt, fr.value))
        crcs.append(zlib.crc32(dpcm.tobytes()) & 0xFFFFFFFF)
    tocs = set()
    for p, _ in packets:                                                    # TOC config of every elementary stream
        q = p
        for s in range(streams.value):
            if not q:
                break
            tocs.add(q[0] >> 3)
            if s == streams.value - 1:
                break
            # self-delimited framing (opus.c:190-343): skip this sub-packet via the reference's own parser
            off = ctypes.c_int(0)
            sizes = (ctypes.c_short * 48)()
            toc = ctypes.c_ubyte(0)
            cnt = lib.opus_packet_parse_impl(q, len(q), 1, ctypes.byref(toc), None, sizes, None, ctypes.byref(off))
            assert cnt > 0
            q = q[off.value:]
    with open(os.path.join(HERE, name + ".bit"), "wb") as f:
        for p, r in packets:
            f.write(struct.pack(">II", len(p), r) + p)
    index.update({name: {"fs": 48000, "channels": 6, "streams": streams.value, "coupled": coupled.value,
                             "mapping": list(mapping), "frames": FRAMES, "samples_per_frame": 960,
                             "pcm_crc32": crcs, "lost": [i for i, (p, _) in enumerate(packets) if not p],
                         "toc_configs": sorted(tocs)}})
    print(name, "toc configs", sorted(tocs), "streams", streams.value, "coupled", coupled.value, "mapping", list(mapping),
          "avg bytes", sum(len(p) for p, _ in packets) / FRAMES, "lost", sum(1 for p, _ in packets if not p))


def main():
    index = {}
    gen("ms51_256k", 256000, True, 2049, index)      # config 5: CELT on every stream
    gen("ms51_80k_voip", 80000, False, 2048, index)  # low rate, automatic modes: SILK / hybrid / CELT (LFE) streams side by side
    gen("ms51_40k_voip", 40000, False, 2048, index)  # lower still: SILK-only streams
    json.dump(index, open(os.path.join(HERE, "index_ms.json"), "w"))


if __name__ == "__main__":
    main()
