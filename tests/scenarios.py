"""Deterministic per-stream call sequences (packet, frame_size, fec) built from the golden fixtures:
loss bursts, DTX-style 1-byte packets, FEC requests, multi-frame packets (code 1/2/3 with padding),
bit corruption, truncation and garbage -- the cases the reference's own robustness tests exercise
(opus-fix/tests/test_opus_decode.c:214-359; CSharp/ParityTest/TestDriver.cs:326-380)."""
import numpy as np

from conftest import golden_case


def _enc_size(n):
    if n < 252:
        return bytes([n])
    b0 = 252 + (n & 3)
    return bytes([b0, (n - b0) >> 2])


def pack_frames(frames, toc, rng):
    """Repacketise CELT frames (payloads without TOC) into one Opus packet (src/opus.c:190-343)."""
    m = len(frames)
    toc &= 0xFC
    if m == 1:
        return bytes([toc]) + frames[0]
    if m == 2 and len(frames[0]) == len(frames[1]) and rng.rand() < 0.5:
        return bytes([toc | 1]) + frames[0] + frames[1]
    if m == 2 and rng.rand() < 0.5:
        return bytes([toc | 2]) + _enc_size(len(frames[0])) + frames[0] + frames[1]
    cbr = all(len(f) == len(frames[0]) for f in frames) and rng.rand() < 0.5
    pad = int(rng.randint(0, 300)) if rng.rand() < 0.4 else 0
    out = bytes([toc | 3, m | (0 if cbr else 0x80) | (0x40 if pad else 0)])
    if pad:
        p = pad
        while p > 254:
            out += bytes([255])
            p -= 254
        out += bytes([p])
    if not cbr:
        for f in frames[:-1]:
            out += _enc_size(len(f))
    return out + b"".join(frames) + bytes(pad)


def loss_scenario(name, seed):
    """One call per golden packet; ~20 % of the calls are losses of various kinds."""
    meta, pk = golden_case(name)
    fs, spf = meta["fs"], meta["samples_per_frame"]
    rng = np.random.RandomState(seed)
    calls, burst = [], 0
    for n, (p, _) in enumerate(pk):
        u = rng.rand()
        if burst > 0:
            lost, burst = True, burst - 1
        elif u < 0.08:
            lost = True
        elif u < 0.10:
            lost, burst = True, int(rng.randint(2, 9))
        else:
            lost = False
        if n == 0 and (seed & 1):
            lost = True                                   # loss before any packet: zeros
        fsz, fec, data = spf, 0, p
        if lost:
            v = rng.rand()
            if v < 0.3:
                data = p[:1]                              # TOC only: DTX / "conceal this frame"
            elif v < 0.4:
                fec = 1                                   # FEC request on a CELT packet -> PLC
            else:
                data = None
            if rng.rand() < 0.15 and spf * 2 <= fs // 50 * 3:
                fsz = spf * 2                             # conceal two frames' worth in one call
        calls.append((data, fsz, fec))
    return meta, calls


def fuzz_scenario(name, seed):
    """Multi-frame packets, corrupted / truncated / random payloads, too-small frame_size."""
    meta, pk = golden_case(name)
    fs, spf = meta["fs"], meta["samples_per_frame"]
    rng = np.random.RandomState(seed)
    calls, i = [], 0
    maxm = min(6, (fs * 3 // 25) // spf)
    while i < len(pk) - 8:
        m = int(rng.randint(1, maxm + 1))
        frames = [pk[i + k][0][1:] for k in range(m)]
        toc = pk[i][0][0]
        i += m
        u = rng.rand()
        if u < 0.15:
            frames = [bytearray(f) for f in frames]
            for f in frames:
                for _ in range(rng.randint(1, 6)):
                    if len(f):
                        f[rng.randint(0, len(f))] = rng.randint(0, 256)
            frames = [bytes(f) for f in frames]
        elif u < 0.22:
            k = rng.randint(0, m)
            frames[k] = frames[k][:rng.randint(0, len(frames[k]) + 1)]
        elif u < 0.27:
            frames = [bytes(rng.randint(0, 256, size=rng.randint(2, 200), dtype=np.uint8).tolist()) for _ in range(m)]
        data = pack_frames(frames, toc, rng)
        if rng.rand() < 0.05:
            data = data[:rng.randint(1, len(data) + 1)]
        fsz = m * spf if rng.rand() < 0.9 else max(1, m * spf - spf // 2)
        calls.append((data, fsz, 0))
    return meta, calls


def silk_calls(npackets, loss, fec, seed=11):
    """The decode-call sequence of the SILK loss/FEC fixtures: yields (packet index or None, decode_fec).
    A lost packet is concealed (None, 0); with FEC enabled the concealment is deferred to the arrival of
    the next packet, whose in-band redundancy is decoded first: (next, 1) then (next, 0)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    lost = [bool(loss > 0 and rng.random() < loss) for _ in range(npackets)]
    lost[0] = False
    out = []
    pending = False
    for i in range(npackets):
        if lost[i]:
            if fec:
                if pending:
                    out.append((None, 0))     # two losses in a row: the older one is plainly concealed
                pending = True
            else:
                out.append((None, 0))
            continue
        if pending:
            out.append((i, 1))
            pending = False
        out.append((i, 0))
    return out


def silk_fuzz_calls(packets, seed, random_toc=False):
    """Corrupted, truncated and random payloads for the SILK / hybrid fixtures (the TOC is kept unless
    random_toc, which also throws the decoder between modes on garbage): list of packet bytes or None."""
    import numpy as np
    rng = np.random.RandomState(seed)
    out = []
    for p in packets:
        u = rng.rand()
        if u < 0.35:
            b = bytearray(p)
            for _ in range(rng.randint(1, 8)):
                if len(b) > 1:
                    b[rng.randint(1, len(b))] = rng.randint(0, 256)
            p = bytes(b)
        elif u < 0.5 and len(p) > 2:
            p = p[:rng.randint(1, len(p))]
        elif u < 0.6:
            p = bytes([p[0]]) + bytes(rng.randint(0, 256, size=rng.randint(1, 120), dtype=np.uint8).tolist())
        elif u < 0.65:
            p = None
        if p is not None and random_toc and rng.rand() < 0.3:
            p = bytes([int(rng.randint(0, 256)) & 0xFC]) + p[1:]       # any configuration, one frame
        out.append(p)
    return out


def random_payload_scenario(fs, ch, seed, n=160):
    """Packets with a CELT-only TOC and RANDOM payload bytes: whatever band layout, stereo mode, split tree, tf /
    anti-collapse combination the bits happen to spell (things no encoder would emit included).  Golden and
    corrupted-golden packets stay close to what the encoder produces; these do not.  -> (meta, calls)."""
    rng = np.random.default_rng(seed)
    calls = []
    for k in range(n):
        stereo = int(rng.integers(0, 2))
        cfg = int(rng.integers(16, 32))                       # CELT-only NB .. FB, 2.5 .. 20 ms
        toc = (cfg << 3) | (stereo << 2)
        size = int(rng.choice([2, 3, 6, 12, 25, 50, 100, 160, 240, 400, 700, 1275]))
        if int(rng.integers(0, 5)) == 0:
            body = bytes([int(rng.choice([0, 255]))]) * (size - 1)
        else:
            body = rng.integers(0, 256, size - 1, dtype=np.uint8).tobytes()
        spf = fs // 400 << (cfg & 3)
        calls.append((bytes([toc]) + body, spf, 0))
        if k % 23 == 5:
            calls.append((None, spf, 0))                      # and a loss now and then
    return {"fs": fs, "channels": ch}, calls
