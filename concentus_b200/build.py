"""Build libopus_b200.so (CUDA, sm_100a) in-tree with nvcc.  No JIT cache, no torch extension:
the product is a plain C-ABI shared library (see include/opus_b200.h)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libopus_b200.so")
SOURCES = ["cb_kernels.cu", "cb_parse_kernels.cu", "opus_api.cu", "cb_noalloc.cu", "cb_enc_kernels.cu", "opus_enc_api.cu", "cb_silk_kernels.cu", "cb_general_kernels.cu", "opus_repacketizer.cu", "opus_ms_enc_api.cu", "cb_surround_kernels.cu", "cb_selftest.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--compiler-options", "-fPIC", "-shared",
]


OBJ = os.path.join(HERE, "_obj")


def _newest(paths):
    return max(os.path.getmtime(p) for p in paths)


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    for src in SOURCES:
        obj = os.path.join(OBJ, src[:-3] + ".o")
        if _object_stale(obj, os.path.join(OBJ, src[:-3] + ".d"), os.path.join(CSRC, src)) or os.path.getmtime(obj) > t:
            return True
    return False


def _deps(dfile):
    """Prerequisites listed in a make-style dependency file written by nvcc -MD."""
    txt = open(dfile).read().replace("\\\n", " ")
    deps = [os.path.join(CSRC, t) for t in txt.split(":", 1)[1].split()]      # nvcc ran with cwd = csrc
    local = [d for d in deps if os.path.dirname(os.path.abspath(d)) == CSRC]
    if not local:
        raise IndexError("no local prerequisites in " + dfile)
    return [d for d in deps if os.path.exists(d)]


def _object_stale(obj, dfile, src):
    if not (os.path.exists(obj) and os.path.exists(dfile)):
        return True
    t = os.path.getmtime(obj)
    if os.path.getmtime(src) > t or os.path.getmtime(__file__) > t:
        return True
    try:
        return any(os.path.getmtime(d) > t for d in _deps(dfile))
    except (OSError, IndexError):
        return True


def build(force=False, verbose=False):
    """Compile the library if sources are newer than the .so.  Returns the path.
    One object per translation unit, compiled concurrently; an object is reused while none of the files
    it included has changed."""
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "nvcc")
    os.makedirs(OBJ, exist_ok=True)
    jobs = []
    objs = []
    for src in SOURCES:
        obj = os.path.join(OBJ, src[:-3] + ".o")
        dfile = os.path.join(OBJ, src[:-3] + ".d")
        objs.append(obj)
        if not force and not _object_stale(obj, dfile, os.path.join(CSRC, src)):
            continue
        cmd = [nvcc] + NVCC_FLAGS[:-1] + (["-Xptxas", "-v"] if verbose else []) + ["-MD", "-MF", dfile, "-c", "-o", obj, src]
        jobs.append((src, subprocess.Popen(cmd, cwd=CSRC, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, pr in jobs:
        out = pr.communicate()[0]
        if pr.returncode != 0:
            sys.stderr.write(out)
            failed = True
        elif verbose:
            print(out)
    if failed:
        raise RuntimeError("nvcc failed building libopus_b200.so")
    res = subprocess.run([nvcc] + NVCC_FLAGS + ["-o", LIB] + objs, cwd=CSRC, stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError("nvcc failed linking libopus_b200.so")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
