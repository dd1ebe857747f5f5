#!/usr/bin/env python3
"""Generate concentus_b200/csrc/celt_tables.h (and oracle/celt_tables_c.h).

The CELT static mode (48 kHz / 960) is normative *data* of RFC 6716 / libopus 1.1.2.  This
script reads the numeric tables out of the reference tree (read-only) where they cannot be
derived cheaply, re-derives the ones that can (PVQ U(n,k) recurrence) and cross-checks them against the reference values, then emits our own header in
our own layout.  Run in the build container only (needs /root/reference); the emitted header
is committed.

Reference locations: celt/modes.c:42-63, celt/static_modes_fixed.h:14-506,
celt/quant_bands.c:46-140, celt/cwrs.c:213-428, celt/rate.c:42-48.
"""
import math, re, sys, os

REF = "/root/reference/opus-fix/celt"

def strip_custom_modes(src):
    # drop '#if defined(CUSTOM_MODES) ... #endif' blocks (single level) used inside tables
    out, skip = [], 0
    for line in src.splitlines():
        s = line.strip()
        if skip:
            if s.startswith("#if"):
                skip += 1
            elif s.startswith("#endif"):
                skip -= 1
            elif s.startswith("#else") and skip == 1:
                skip = 0
            continue
        if re.match(r"#\s*if\s+defined\(CUSTOM_MODES\)", s):
            skip = 1
            continue
        out.append(line)
    return "\n".join(out)

def grab(path, name, strip=False):
    src = open(os.path.join(REF, path)).read()
    if strip:
        src = strip_custom_modes(src)
    m = re.search(re.escape(name) + r"\s*(\[[^\]]*\])*\s*=\s*\{", src)
    assert m, name
    i = m.end()
    depth, j = 1, i
    while depth:
        c = src[j]
        depth += (c == "{") - (c == "}")
        j += 1
    body = src[i:j - 1]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    body = re.sub(r"#.*", "", body)
    vals = [int(t.rstrip("Uu")) for t in re.findall(r"-?\d+[Uu]?", body)]
    return vals

def pvq_u_table():
    # U(n,k)=U(n-1,k)+U(n,k-1)+U(n-1,k-1); U(0,0)=1, U(n,0)=U(0,k)=0 otherwise (cwrs.c:105-118)
    maxk = [176, 176, 176, 176, 176, 176, 96, 54, 37, 28, 24, 19, 18, 16, 14]
    K = 178
    U = [[0] * K for _ in range(15)]
    U[0][0] = 1
    for n in range(1, 15):
        for k in range(1, K):
            U[n][k] = U[n - 1][k] + U[n][k - 1] + U[n - 1][k - 1]
    data, rows = [], []
    for n in range(15):
        rows.append(len(data) - n)          # so that row[n][k] = data[rows[n]+k] for k>=n
        for k in range(n, maxk[n] + 1):
            v = U[n][k]
            assert v < 2 ** 32, (n, k)
            data.append(v)
    return data, rows

def fmt(vals, per=16):
    lines = []
    for i in range(0, len(vals), per):
        lines.append("  " + ",".join(str(v) + ("u" if v > 2147483647 else "") for v in vals[i:i + per]) + ",")
    return "\n".join(lines)

def main():
    T = {}
    T["eband5ms"] = ("int16_t", grab("modes.c", "eband5ms"))
    T["band_alloc"] = ("uint8_t", grab("modes.c", "band_allocation"))
    T["window120"] = ("int16_t", grab("static_modes_fixed.h", "window120"))
    T["logN400"] = ("int16_t", grab("static_modes_fixed.h", "logN400"))
    T["cache_index50"] = ("int16_t", grab("static_modes_fixed.h", "cache_index50"))
    T["cache_bits50"] = ("uint8_t", grab("static_modes_fixed.h", "cache_bits50"))
    T["cache_caps50"] = ("uint8_t", grab("static_modes_fixed.h", "cache_caps50"))
    tw = grab("static_modes_fixed.h", "fft_twiddles48000_960")
    assert len(tw) == 960
    # (generated upstream with the fixed-point celt_cos_norm, _kiss_fft_guts.h:136-140; taken as data)
    assert tw[0] == 32767 and tw[1] == 0 and tw[2 * 120] == 0 and tw[2 * 120 + 1] == -32767
    T["fft_tw480"] = ("int16_t", tw)
    for n in (480, 240, 120, 60):
        T["fft_bitrev%d" % n] = ("int16_t", grab("static_modes_fixed.h", "fft_bitrev%d" % n))
        assert sorted(T["fft_bitrev%d" % n][1]) == list(range(n))
    T["mdct_tw960"] = ("int16_t", grab("static_modes_fixed.h", "mdct_twiddles960"))
    assert len(T["mdct_tw960"][1]) == 1800
    em = grab("quant_bands.c", "eMeans")          # first definition = FIXED_POINT Q4 ints
    T["eMeans"] = ("int8_t", em[:25])
    T["e_prob_model"] = ("uint8_t", grab("quant_bands.c", "e_prob_model"))
    assert len(T["e_prob_model"][1]) == 4 * 2 * 42
    T["log2_frac_table"] = ("uint8_t", grab("rate.c", "LOG2_FRAC_TABLE"))
    ref_u = grab("cwrs.c", "CELT_PVQ_U_DATA", strip=True)
    data, rows = pvq_u_table()
    assert data == ref_u, "PVQ U table mismatch"
    T["pvq_u_data"] = ("uint32_t", data)
    T["pvq_u_row"] = ("int16_t", rows)
    # lane-indexed copies for the encoder's parallel index computation (kernel header only)
    T["eband5ms_g"] = T["eband5ms"]
    T["eMeans_g"] = T["eMeans"]
    eb = T["eband5ms"][1]
    T["band_of_bin_g"] = ("uint8_t", [max(b for b in range(21) if eb[b] <= j) for j in range(eb[21])])   # band of a 2.5 ms bin
    T["cache_bits50_g"] = (T["cache_bits50"][0], T["cache_bits50"][1] + [0] * 64)   # padded: lanes read 64 entries of a row
    T["band_alloc_g"] = T["band_alloc"]
    # derived: the budget above which quant_partition splits a band (bands.c:893: b > cache[cache[0]] + 12), per
    # (LM + 1, band) like cache_index50 -- one read instead of two dependent ones in the parse walk
    ci, cbits = T["cache_index50"][1], T["cache_bits50"][1]
    T["split_thresh"] = ("int16_t", [cbits[i + cbits[i]] + 12 if i >= 0 else 32767 for i in ci])
    # derived: bits2pulses (rate.h:48-70) as a table.  The 6-step bisection over a row of cache_bits50 depends on
    # (row, bits) only and is constant from bits = 256 on (no entry exceeds 255), so every answer fits in
    # 23 rows x 257 budgets; pulse_row maps (LM + 1, band) to the row.
    def bits2pulses(row, bits):
        cache = cbits[row:]
        lo, hi = 0, cache[0]
        bits -= 1
        for _ in range(6):
            mid = (lo + hi + 1) >> 1
            if cache[mid] >= bits:
                hi = mid
            else:
                lo = mid
        return lo if bits - (-1 if lo == 0 else cache[lo]) <= cache[hi] - bits else hi
    rows_d = sorted(set(i for i in ci if i >= 0))
    T["pulse_row"] = ("uint8_t", [rows_d.index(i) if i >= 0 else 255 for i in ci])
    inv = []
    for r in rows_d:
        inv += [bits2pulses(r, b) for b in range(257)]
        assert all(bits2pulses(r, b) == bits2pulses(r, 256) for b in range(256, 4096))
    T["bits2pulses"] = ("uint8_t", inv)
    T["pvq_u_data_g"] = ("uint32_t", data)
    T["pvq_u_row_g"] = ("int16_t", rows)
    assert len(T["eband5ms"][1]) == 22 and len(T["band_alloc"][1]) == 231

    out = []
    out.append("/* GENERATED by tools/gen_tables.py -- do not edit.\n"
               " * Static tables of the single CELT mode (48000 Hz, 960-sample frames, overlap 120),\n"
               " * i.e. normative data of RFC 6716 as shipped in libopus 1.1.2 (see generator for the\n"
               " * reference file:line of each table).  CB_UTABLE = tables indexed uniformly by a\n"
               " * warp (constant cache); CB_GTABLE = tables indexed per lane (global, L1-cached). */\n")
    out.append("#pragma once\n#include <stdint.h>\n")
    lane_indexed = {"window120", "fft_tw480", "fft_bitrev480", "fft_bitrev240", "fft_bitrev120",
                    "fft_bitrev60", "mdct_tw960", "eband5ms_g", "eMeans_g", "band_of_bin_g", "cache_bits50_g", "band_alloc_g", "pvq_u_data_g", "pvq_u_row_g"}
    for name, (ty, vals) in T.items():
        q = "CB_GTABLE" if name in lane_indexed else "CB_UTABLE"
        out.append("%s %s ct_%s[%d] = {\n%s\n};\n" % (q, ty, name, len(vals), fmt(vals)))
    hdr = "\n".join(out)
    here = os.path.dirname(os.path.abspath(__file__))
    p = os.path.join(here, "..", "concentus_b200", "csrc", "celt_tables.h")
    open(p, "w").write(hdr)
    print("wrote", os.path.normpath(p), len(hdr), "bytes")
    # plain-C copy for the oracle (test infrastructure): same data, static const arrays
    oc = ["/* GENERATED by tools/gen_tables.py -- do not edit.  Static tables of the CELT 48 kHz/960 mode\n"
          " * (normative RFC 6716 data, libopus 1.1.2) for the plain-C oracle. */\n#ifndef ORACLE_TABLES_H\n"
          "#define ORACLE_TABLES_H\n#include <stdint.h>\n"]
    for name, (ty, vals) in T.items():
        if name.endswith("_g") or name in ("split_thresh", "pulse_row", "bits2pulses"):
            continue
        oc.append("static const %s ot_%s[%d] = {\n%s\n};\n" % (ty, name, len(vals), fmt(vals)))
    oc.append("#endif\n")
    p2 = os.path.join(here, "..", "oracle", "oracle_tables.h")
    open(p2, "w").write("\n".join(oc))
    print("wrote", os.path.normpath(p2))

if __name__ == "__main__":
    main()
