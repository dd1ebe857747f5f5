#!/usr/bin/env python3
"""Attribute SASS instruction counts of a kernel to source functions (needs -lineinfo).
usage: sass_by_function.py <lib.so> <kernel-substring>"""
import collections, os, re, subprocess, sys, tempfile

def main():
    lib, pat = os.path.abspath(sys.argv[1]), sys.argv[2]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
        text = ""
        for f in os.listdir(tmp):
            if f.endswith(".cubin"):
                r = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], stdout=subprocess.PIPE, text=True)
                text += r.stdout
    sec, cur = None, None
    cnt = collections.Counter()
    for l in text.splitlines():
        m = re.match(r'\s*\.section\s+\.text\.(\S+),', l)
        if m:
            sec = m.group(1)
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split('/')[-1], int(m.group(2)))
            continue
        if sec and pat in sec and re.match(r'\s+/\*[0-9a-f]{4,}\*/', l):
            cnt[cur] += 1
    src = {}
    def func_of(fn, line):
        if fn not in src:
            try:
                src[fn] = open(os.path.join(root, "concentus_b200", "csrc", fn)).read().splitlines()
            except Exception:
                src[fn] = []
        L = src[fn]
        for i in range(min(line, len(L)) - 1, -1, -1):
            m = re.match(r'(CB_FN|CB_HD|static|__global__|template).*?\b(\w+)\s*\(', L[i])
            if m and not L[i].startswith(' '):
                return m.group(2)
        return '?'
    byf = collections.Counter()
    for (fn, ln), c in cnt.items():
        byf[(fn, func_of(fn, ln))] += c
    print("total", sum(byf.values()))
    for k, c in byf.most_common(45):
        print("%6d  %s:%s" % (c, k[0], k[1]))

if __name__ == "__main__":
    main()
