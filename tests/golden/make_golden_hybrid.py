#!/usr/bin/env python3
"""Generate the hybrid / mode-transition decode golden fixtures from the UNMODIFIED reference
(oracle/_ref/libopus_ref.so).

Runs only in the build container (needs /root/reference/AudioData).  The reference encoder runs at 48 kHz
on the reference's own test audio with its mode forced per segment (OPUS_SET_FORCE_MODE: SILK-only,
hybrid, CELT-only) and the bandwidth set per segment, so the packet sequence contains hybrid packets,
every kind of mode transition and the redundant CELT frames the encoder attaches to them.  The reference
decoder decodes the packets -- with the listed loss / FEC pattern -- and we keep the packets in opus_demo's
.bit container plus, per decode call, return value, CRC32 of the PCM and the final range.
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
FORCE_MODE, SET_BANDWIDTH, AUTO = 11002, 4008, -1000
MODE = {"silk": 1000, "hybrid": 1001, "celt": 1002, "auto": AUTO}
NB, MB, WB, SWB, FB = 1101, 1102, 1103, 1104, 1105
# name, channels, frame ms, bitrate, application, extra ctls, segments [(frames, mode, bandwidth)], decoder Fs, decoder channels, skip, loss
CASES = [
    ("hyb_fb_mono_32k", 1, 20, 32000, 2048, [], [(120, "hybrid", FB)], 48000, 1, 50, 0.0),
    ("hyb_swb_st_48k", 2, 20, 48000, 2048, [], [(100, "hybrid", SWB)], 48000, 2, 50, 0.0),
    ("hyb_fb_mono_10ms_to16", 1, 10, 36000, 2048, [], [(120, "hybrid", FB)], 16000, 1, 100, 0.0),
    ("hyb_fb_st_loss", 2, 20, 56000, 2048, [], [(100, "hybrid", FB)], 48000, 2, 50, 0.15),
    ("hyb_fb_mono_fec", 1, 20, 40000, 2048, [(4012, 1), (4014, 20)], [(100, "hybrid", FB)], 48000, 1, 50, 0.15),
    ("hyb_fb_mono_40ms", 1, 40, 32000, 2048, [], [(40, "hybrid", FB)], 48000, 2, 20, 0.0),
    ("hyb_auto_voip_mono", 1, 20, 28000, 2048, [], [(150, "auto", AUTO)], 48000, 1, 0, 0.0),
    ("trans_silk_celt_mono", 1, 20, 32000, 2049, [],
     [(10, "silk", WB), (10, "celt", FB), (10, "silk", WB), (10, "celt", WB), (10, "silk", NB)], 48000, 1, 20, 0.0),
    ("trans_hyb_celt_st", 2, 20, 56000, 2049, [],
     [(8, "hybrid", FB), (8, "celt", FB), (8, "hybrid", SWB), (8, "celt", SWB), (8, "hybrid", FB)], 48000, 2, 20, 0.0),
    ("trans_silk_hyb_mono", 1, 20, 30000, 2048, [],
     [(8, "silk", WB), (8, "hybrid", FB), (8, "silk", MB), (8, "hybrid", SWB), (8, "silk", WB)], 48000, 1, 20, 0.0),
    ("trans_all_st_to24", 2, 20, 48000, 2049, [],
     [(5, "celt", FB), (6, "silk", NB), (4, "hybrid", FB), (5, "celt", MB), (6, "hybrid", SWB), (4, "silk", WB),
      (4, "celt", FB), (5, "hybrid", FB), (5, "silk", MB)], 24000, 2, 20, 0.0),
    ("trans_all_mono_10ms_loss", 1, 10, 30000, 2049, [],
     [(8, "celt", FB), (9, "silk", WB), (7, "hybrid", FB), (8, "celt", WB), (9, "hybrid", SWB), (7, "silk", NB),
      (8, "celt", SWB)], 48000, 2, 40, 0.12),
]


def main():
    index = {}
    for name, ch, ms, br, app, ctls, segs, dfs, dch, skip, loss in CASES:
        raw = np.fromfile(os.path.join(AUDIO, "48Khz %s.raw" % ("Stereo" if ch == 2 else "Mono")), dtype=np.int16)
        enc = RefEncoder(48000, ch, app)
        assert enc.ctl(4002, br) == 0
        for r, v in ctls:
            assert enc.ctl(r, v) == 0
        fr = 48 * ms
        k = skip
        pk = []
        for frames, mode, bw in segs:
            assert enc.ctl(FORCE_MODE, MODE[mode]) == 0
            assert enc.ctl(SET_BANDWIDTH, bw) == 0
            for _ in range(frames):
                n, p, rng = enc.encode(raw[k * fr * ch:(k + 1) * fr * ch], fr, 1275)
                assert n > 0
                pk.append((p, rng))
                k += 1
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
        tocs = sorted(set(p[0] >> 3 for p, _ in pk))
        index[name] = {"fs": dfs, "channels": dch, "frame_ms": ms, "frames": len(pk), "loss": loss, "fec": fec,
                       "samples_per_frame": fsz, "ret": rets, "pcm_crc32": crcs, "final_range": rngs, "toc_configs": tocs,
                       "encoder": "opus_encode app %d 48000 Hz %d ch %d ms %d bps ctls %s segments %s on AudioData from frame %d" %
                                  (app, ch, ms, br, ctls, segs, skip)}
        print(name, len(pk), "packets", sum(len(p) for p, _ in pk), "bytes,", len(rets), "decode calls, toc configs", tocs)
    json.dump(index, open(os.path.join(HERE, "index_hybrid.json"), "w"), indent=0)


if __name__ == "__main__":
    main()
