#!/usr/bin/env python3
"""Join an ncu source-page CSV (per-SASS executed counts + stall samples) with nvdisasm line info
of the same .so, and aggregate by source function.
usage: ncu_by_function.py <source.csv> <lib.so> <kernel-substring>"""
import collections, csv, os, re, subprocess, sys, tempfile

HELPER_FILES = ("cb_common.h", "sm_30_intrinsics.hpp", "sm_32_intrinsics.hpp", "sm_80_rt.hpp", "device_functions.hpp")

def sass_lines(lib, pat, outer=True):
    """One (file, line) per SASS instruction of the sections matching pat.  With outer=True an
    instruction inlined from a helper header (fixed-point primitives, shuffle intrinsics) is
    attributed to the nearest enclosing frame that is not a helper."""
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
        text = ""
        for f in sorted(os.listdir(tmp)):
            if f.endswith(".cubin") and "-" not in f:
                text += subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, f)], stdout=subprocess.PIPE, text=True).stdout
    sec, chain, out = None, [], []
    fresh = True
    for l in text.splitlines():
        m = re.match(r'\s*\.section\s+\.text\.(\S+),', l)
        if m:
            sec = m.group(1)
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            if fresh:
                chain = []
                fresh = False
            chain.append((m.group(1).split('/')[-1], int(m.group(2))))
            continue
        if sec and pat in sec and re.match(r'\s+/\*[0-9a-f]{4,}\*/', l):
            cur = chain[0] if chain else None
            if outer:
                for fr in chain:
                    if fr[0] not in HELPER_FILES:
                        cur = fr
                        break
            out.append(cur)
            fresh = True
    return out

def main():
    src_csv, lib, pat = sys.argv[1:4]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    rows = list(csv.reader(open(src_csv)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
    ix = {h: i for i, h in enumerate(rows[hi])}
    data = rows[hi + 1:]
    lines = sass_lines(lib, pat)
    assert len(lines) == len(data), (len(lines), len(data))
    srcs = {}
    def func_of(fn, line):
        if fn not in srcs:
            try:
                srcs[fn] = open(os.path.join(root, "concentus_b200", "csrc", fn)).read().splitlines()
            except Exception:
                srcs[fn] = []
        L = srcs[fn]
        for i in range(min(line, len(L)) - 1, -1, -1):
            m = re.match(r'(CB_FN_NOINLINE|CB_FN|CB_HD|static|__global__|template).*?\b(\w+)\s*\(', L[i])
            if m and not L[i].startswith(' '):
                return m.group(2)
        return '?'
    agg = collections.defaultdict(lambda: [0, 0, 0, 0])   # static, executed, samples, no_inst
    for loc, r in zip(lines, data):
        key = (loc[0], func_of(*loc)) if loc else ('?', '?')
        a = agg[key]
        a[0] += 1
        a[1] += int(r[ix['Instructions Executed']] or 0)
        a[2] += int(r[ix['# Samples']] or 0)
        a[3] += int(r[ix['stall_no_inst']] or 0)
    if len(sys.argv) > 4:                                  # per-line view of one function
        want = sys.argv[4]
        per = collections.defaultdict(lambda: [0, 0, 0])
        for loc, r in zip(lines, data):
            if loc and func_of(*loc) == want:
                a = per[loc]
                a[0] += 1
                a[1] += int(r[ix['Instructions Executed']] or 0)
                a[2] += int(r[ix['# Samples']] or 0)
        for loc, v in sorted(per.items(), key=lambda kv: kv[0]):
            src = srcs.get(loc[0], [])
            text = src[loc[1] - 1].strip() if 0 < loc[1] <= len(src) else ''
            print("%s:%-4d %5d %11d %7d  %s" % (loc[0][:16], loc[1], v[0], v[1], v[2], text[:90]))
        return
    tot = [sum(v[k] for v in agg.values()) for k in range(4)]
    print("%-40s %7s %12s %6s %8s %6s %8s" % ("function", "static", "executed", "%exec", "samples", "%smp", "no_inst"))
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][2])[:40]:
        print("%-40s %7d %12d %6.1f %8d %6.1f %8d" % (k[0][:14] + ":" + k[1], v[0], v[1], 100.0 * v[1] / tot[1], v[2], 100.0 * v[2] / tot[2], v[3]))
    print("TOTAL static %d executed %d samples %d no_inst %d" % tuple(tot))

if __name__ == "__main__":
    main()
