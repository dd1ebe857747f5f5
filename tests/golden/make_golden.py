#!/usr/bin/env python3
"""Generate the golden fixtures in this directory from the UNMODIFIED reference.

Runs only in the build container: needs /root/reference/AudioData and oracle/_ref/opus_demo
(`make -C oracle ref`).  For every case the reference encoder (opus_demo -e, libopus 1.1.2
FIXED_POINT) produces packets, the reference decoder (opus_demo -d) produces PCM; we keep the
first FRAMES packets in opus_demo's own .bit container (BE32 length, BE32 encoder final range,
payload -- opus-fix/src/opus_demo.c:68-80,750-760) and, per packet, the CRC32 of the decoded
PCM.  The tests hold our decoders (oracle restatement, lane-loop emulation, CUDA library) to
these bytes.
"""
import json, os, struct, subprocess, sys, tempfile, zlib

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.normpath(os.path.join(HERE, "..", ".."))
DEMO = os.path.join(ROOT, "oracle", "_ref", "opus_demo")
AUDIO = "/root/reference/AudioData"
FRAMES = 250

# name, decode Fs, decode channels, encoder args (after 'restricted-lowdelay 48000 <ch>'), input, skip packets
CASES = [
    ("c1_128k_cbr_st", 48000, 2, 2, ["128000", "-cbr", "-complexity", "10", "-framesize", "20"], "48Khz Stereo.raw", 100),
    ("c2_96k_cbr_st", 48000, 2, 2, ["96000", "-cbr", "-complexity", "10", "-framesize", "20"], "48Khz Stereo.raw", 0),
    ("vbr_128k_st", 48000, 2, 2, ["128000", "-complexity", "10", "-framesize", "20"], "48Khz Stereo.raw", 300),
    ("low_24k_st", 48000, 2, 2, ["24000", "-complexity", "10", "-framesize", "20"], "48Khz Stereo.raw", 200),
    ("f10_64k_st", 48000, 2, 2, ["64000", "-complexity", "10", "-framesize", "10"], "48Khz Stereo.raw", 400),
    ("f5_64k_st", 48000, 2, 2, ["64000", "-complexity", "10", "-framesize", "5"], "48Khz Stereo.raw", 800),
    ("f2_5_96k_st", 48000, 2, 2, ["96000", "-complexity", "10", "-framesize", "2.5"], "48Khz Stereo.raw", 1600),
    ("mono_64k", 48000, 1, 1, ["64000", "-complexity", "10", "-framesize", "20"], "48Khz Mono.raw", 150),
    ("mono_12k", 48000, 1, 1, ["12000", "-complexity", "10", "-framesize", "20"], "48Khz Mono.raw", 50),
    ("hi_510k_st", 48000, 2, 2, ["510000", "-complexity", "10", "-framesize", "20"], "48Khz Stereo.raw", 500),
    ("swb_48k_st", 48000, 2, 2, ["48000", "-bandwidth", "SWB", "-complexity", "10", "-framesize", "20"], "48Khz Stereo.raw", 250),
    ("wb_40k_f10_st", 48000, 2, 2, ["40000", "-bandwidth", "WB", "-complexity", "10", "-framesize", "10"], "48Khz Stereo.raw", 0),
    ("nb_20k_mono", 48000, 1, 1, ["20000", "-bandwidth", "NB", "-complexity", "10", "-framesize", "20"], "48Khz Mono.raw", 300),
    # two more rates for the mixed-bitrate batch of bench.py --config 2het
    ("het_64k_cbr_st", 48000, 2, 2, ["64000", "-cbr", "-complexity", "10", "-framesize", "20"], "48Khz Stereo.raw", 450),
    ("het_96k_vbr_st", 48000, 2, 2, ["96000", "-complexity", "10", "-framesize", "20"], "48Khz Stereo.raw", 550),
    # decoder channel count / rate different from the stream's
    ("st_to_mono_out", 48000, 1, 2, ["64000", "-complexity", "10", "-framesize", "20"], "48Khz Stereo.raw", 100),
    ("mono_to_st_out", 48000, 2, 1, ["48000", "-complexity", "10", "-framesize", "20"], "48Khz Mono.raw", 100),
    ("st_to_24k_out", 24000, 2, 2, ["64000", "-complexity", "10", "-framesize", "20"], "48Khz Stereo.raw", 100),
    ("mono_to_16k_out", 16000, 1, 1, ["32000", "-complexity", "10", "-framesize", "10"], "48Khz Mono.raw", 100),
]


def read_bit(path):
    d = open(path, "rb").read()
    o, out = 0, []
    while o < len(d):
        ln, rng = struct.unpack(">II", d[o:o + 8])
        o += 8
        out.append((d[o:o + ln], rng))
        o += ln
    return out


def main():
    index = {}
    with tempfile.TemporaryDirectory() as tmp:
        for name, dfs, dch, ech, eargs, wav, skip in CASES:
            bit = os.path.join(tmp, name + ".bit")
            subprocess.run([DEMO, "-e", "restricted-lowdelay", "48000", str(ech)] + eargs +
                           [os.path.join(AUDIO, wav), bit], check=True, stdout=subprocess.DEVNULL,
                           stderr=subprocess.DEVNULL)
            pk = read_bit(bit)[skip:skip + FRAMES]
            # re-pack the slice and decode it from a cold decoder with the reference
            sl = os.path.join(HERE, name + ".bit")
            with open(sl, "wb") as f:
                for p, rng in pk:
                    f.write(struct.pack(">II", len(p), rng) + p)
            dec = os.path.join(tmp, name + ".dec")
            subprocess.run([DEMO, "-d", str(dfs), str(dch), sl, dec], check=True, stdout=subprocess.DEVNULL,
                           stderr=subprocess.DEVNULL)
            pcm = open(dec, "rb").read()
            toc = pk[0][0][0]
            spf = (dfs << ((toc >> 3) & 3)) // 400          # CELT-only TOC (opus.c:173-177)
            bpf = spf * dch * 2
            assert len(pcm) == bpf * len(pk), (name, len(pcm), bpf, len(pk))
            crcs = [zlib.crc32(pcm[i * bpf:(i + 1) * bpf]) & 0xFFFFFFFF for i in range(len(pk))]
            index[name] = {"fs": dfs, "channels": dch, "stream_channels": ech, "frames": len(pk),
                           "samples_per_frame": spf, "pcm_crc32": crcs,
                           "pcm_crc32_all": zlib.crc32(pcm) & 0xFFFFFFFF,
                           "encoder": "opus_demo -e restricted-lowdelay 48000 %d %s '%s' (packets %d..%d)" %
                                      (ech, " ".join(eargs), wav, skip, skip + len(pk))}
            print(name, len(pk), "packets", os.path.getsize(sl), "bytes")
    json.dump(index, open(os.path.join(HERE, "index.json"), "w"), indent=0)


if __name__ == "__main__":
    main()
