#!/usr/bin/env python3
"""Generate the SILK decode golden fixtures from the UNMODIFIED reference (oracle/_ref/libopus_ref.so).

Runs only in the build container (needs /root/reference/AudioData).  For every case the reference
encoder (OPUS_APPLICATION_VOIP, so the packets are SILK-only) produces packets from the reference's own
test audio; the reference decoder decodes them -- with the listed loss / FEC pattern -- and we keep the
packets in opus_demo's .bit container plus, per decode call, the CRC32 of the PCM it returned.
The tests replay the same call sequence (tests/scenarios.py:silk_calls) on our decoders.
"""
import json, os, struct, sys, zlib
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.normpath(os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import RefDecoder, RefEncoder  # noqa: E402
from scenarios import silk_calls  # noqa: E402

AUDIO = "/root/reference/AudioData"
VOIP = 2048
# name, encoder Fs, encoder channels, frame ms, bitrate, extra encoder ctls, decoder Fs, decoder channels, frames, skip, loss
CASES = [
    ("silk_wb_mono_32k", 16000, 1, 20, 32000, [], 16000, 1, 200, 50, 0.0),          # config 4
    ("silk_wb_mono_16k_to48", 16000, 1, 20, 16000, [], 48000, 1, 120, 100, 0.0),
    ("silk_wb_st_40k", 16000, 2, 20, 40000, [], 16000, 2, 120, 50, 0.0),
    ("silk_mb_mono_20k", 12000, 1, 20, 20000, [], 12000, 1, 100, 50, 0.0),
    ("silk_nb_mono_12k_to48", 8000, 1, 20, 12000, [], 48000, 2, 100, 50, 0.0),
    ("silk_wb_mono_10ms", 16000, 1, 10, 28000, [], 24000, 1, 160, 100, 0.0),
    ("silk_wb_mono_60ms", 16000, 1, 60, 24000, [], 16000, 1, 40, 10, 0.0),
    ("silk_nb_st_40ms", 8000, 2, 40, 24000, [], 8000, 2, 50, 10, 0.0),
    ("silk_wb_mono_loss", 16000, 1, 20, 24000, [], 16000, 1, 150, 50, 0.15),
    ("silk_wb_mono_fec", 16000, 1, 20, 28000, [(4012, 1), (4014, 20)], 16000, 1, 150, 50, 0.15),
    ("silk_wb_st_fec_to48", 16000, 2, 20, 40000, [(4012, 1), (4014, 20)], 48000, 2, 100, 50, 0.2),
    ("silk_wb_mono_dtx", 16000, 1, 20, 16000, [(4016, 1)], 16000, 1, 250, 0, 0.0),
]


def main():
    index = {}
    for name, efs, ech, ms, br, ctls, dfs, dch, frames, skip, loss in CASES:
        raw = np.fromfile(os.path.join(AUDIO, "%dKhz %s.raw" % (efs // 1000, "Stereo" if ech == 2 else "Mono")), dtype=np.int16)
        enc = RefEncoder(efs, ech, VOIP)
        assert enc.ctl(4002, br) == 0
        for r, v in ctls:
            assert enc.ctl(r, v) == 0
        fr = efs * ms // 1000
        pk = []
        for i in range(skip + frames):
            n, p, rng = enc.encode(raw[i * fr * ech:(i + 1) * fr * ech], fr, 1275)
            assert n > 0
            if i >= skip:
                assert not (p[0] & 0x80) and (p[0] & 0x60) != 0x60, "not a SILK-only packet"
                pk.append((p, rng))
        with open(os.path.join(HERE, name + ".bit"), "wb") as f:
            for p, rng in pk:
                f.write(struct.pack(">II", len(p), rng) + p)
        fec = any(r == 4012 for r, _ in ctls)
        dec = RefDecoder(dfs, dch)
        fsz = dfs * ms // 1000
        crcs, rets, rngs = [], [], []
        for idx, use_fec in silk_calls(len(pk), loss, fec, seed=11):
            data = pk[idx][0] if idx is not None and idx >= 0 else None
            ret, pcm = dec.decode_raw(data, fsz, use_fec)
            rets.append(ret)
            crcs.append(zlib.crc32(np.ascontiguousarray(pcm[:max(ret, 0)]).tobytes()) & 0xFFFFFFFF)
            rngs.append(dec.final_range)
        index[name] = {"fs": dfs, "channels": dch, "frame_ms": ms, "frames": len(pk), "loss": loss, "fec": fec,
                       "samples_per_frame": fsz, "ret": rets, "pcm_crc32": crcs, "final_range": rngs,
                       "encoder": "opus_encode VOIP %d Hz %d ch %d ms %d bps ctls %s on AudioData (frames %d..%d)" %
                                  (efs, ech, ms, br, ctls, skip, skip + frames)}
        print(name, len(pk), "packets", sum(len(p) for p, _ in pk), "bytes,", len(rets), "decode calls")
    json.dump(index, open(os.path.join(HERE, "index_silk.json"), "w"), indent=0)


if __name__ == "__main__":
    main()
