"""Developer tool: hybrid packets, mode transitions and redundant frames made by the compiled reference
encoder; decoded by the compiled reference decoder and the host-emulation build of the decoder headers."""
import sys
import numpy as np
from silk_compare import RefEncoder, speechlike, compare

FORCE_MODE = 11002
MODES = {"silk": 1000, "hybrid": 1001, "celt": 1002, "auto": -1000}


def scripted_packets(ch, frame_ms, script, bitrate=32000, seed=1, app=2048, ctls=()):
    """script: list of (nframes, mode name, bandwidth or None)"""
    fs = 48000
    enc = RefEncoder(fs, ch, app)
    enc.ctl(4002, bitrate)
    for r, v in ctls:
        assert enc.ctl(r, v) == 0
    frame = fs * frame_ms // 1000
    total = sum(s[0] for s in script)
    pcm = speechlike(frame * total, ch, seed, fs)
    out = []
    k = 0
    for n, mode, bw in script:
        assert enc.ctl(FORCE_MODE, MODES[mode]) == 0
        assert enc.ctl(4008, bw if bw else -1000) == 0
        for _ in range(n):
            nb, pkt, rng = enc.encode(pcm[k * frame * ch:(k + 1) * frame * ch], frame, 1275)
            assert nb > 0
            out.append((pkt, rng))
            k += 1
    return out


SCRIPTS = {
    "hybrid": [(30, "hybrid", None)],
    "hybrid_swb": [(30, "hybrid", 1104)],
    "silk>celt": [(10, "silk", 1103), (10, "celt", None), (10, "silk", 1103)],
    "hybrid>celt>hybrid": [(8, "hybrid", None), (8, "celt", None), (8, "hybrid", None)],
    "silk>hybrid>silk": [(8, "silk", 1103), (8, "hybrid", None), (8, "silk", 1102), (6, "hybrid", 1104)],
    "mixed": [(4, "celt", None), (5, "silk", 1101), (3, "hybrid", 1105), (4, "celt", 1102), (5, "hybrid", 1104),
              (3, "silk", 1103), (3, "celt", 1105)],
    "auto": [(40, "auto", None)],
}

if __name__ == "__main__":
    bad = 0
    for name, script in SCRIPTS.items():
        for ch in (1, 2):
            for ms in (20, 10):
                pk = scripted_packets(ch, ms, script, bitrate=28000 * ch)
                tocs = sorted(set(p[0][0] >> 3 for p in pk))
                for fs_dec, chd in ((48000, ch), (48000, 3 - ch), (16000, ch), (24000, 2), (8000, 1)):
                    for loss in (None, 0.15):
                        r = compare(pk, fs_dec, chd, ms, loss=loss)
                        print("ok  " if r < 0 else "FAIL", name, "ch", ch, ms, "ms -> dec", fs_dec, chd, "loss", loss, "tocs", tocs)
                        bad += r >= 0
                        if bad > 4:
                            sys.exit(1)
    print("failures", bad)
