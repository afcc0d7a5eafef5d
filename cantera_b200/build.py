"""Builds cantera_b200/libctb.so (the C-ABI CUDA library, include/ctb_clib.h) in-tree with nvcc for
sm_100a. No JIT cache: the .so sits next to this file so it travels with the repository snapshot."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libctb.so")
# (source, extra nvcc flags, object name); reactor_advance.cu is compiled once per group width
SOURCES = [(s, [], s + ".o") for s in ("mech_pack.cpp", "kin_kernels.cu", "engine.cu", "reactor.cu",
                                       "reactor_kernels.cu", "reactor_batch.cu", "ctb_clib.cu", "derivs_kernels.cu", "transport.cu")]
SOURCES += [("reactor_advance.cu", ["-DCTB_ADV_NT=%d" % nt], "reactor_advance_%d.o" % nt) for nt in (128, 256, 64, 32)]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-O2", "--fmad=true",
              "-diag-suppress", "177"]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def _deps():
    out = []
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for d, _, files in os.walk(root):
            out += [os.path.join(d, f) for f in files if f.endswith((".h", ".cuh", ".cu", ".cpp"))]
    return out


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in _deps())


def build(force=False, verbose=False):
    """Compile every translation unit and link libctb.so. Returns the library path."""
    if not force and not needs_build():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdr_time = max(os.path.getmtime(p) for p in _deps() if p.endswith((".h", ".cuh")))

    def compile_one(item):
        src, extra, oname = item
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, oname)
        if not force and os.path.exists(o) and os.path.getmtime(o) > max(os.path.getmtime(s), hdr_time):
            return o
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + extra + ["-x", "cu", "-c", s, "-o", o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        if verbose:
            sys.stderr.write(r.stderr)
        return o

    with ThreadPoolExecutor(max_workers=min(os.cpu_count() or 4, len(SOURCES))) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs + ["-cudart", "static"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
