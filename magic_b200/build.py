"""In-tree nvcc build of the sm_100a C-ABI library (libmagic_b200.so) and the native self-tests.

    python -m magic_b200.build            # build everything that is out of date
    python -m magic_b200.build --force

nvcc cross-compiles without a GPU.  Outputs stay inside magic_b200/ so they travel to the GPU box
with the repo snapshot (they are git-ignored, not gpurun-ignored).
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "build")
LIB = os.path.join(HERE, "libmagic_b200.so")
SELFTEST = os.path.join(CSRC, "build", "selftest_gemm")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
CFLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-DMAGIC_WAIT_TRAP", "-I", INCLUDE, "-I", CSRC]

LIB_SOURCES = ["gemm.cu", "kernels.cu", "engine.cu", "capi.cu"]
SELFTEST_SOURCES = ["selftest_gemm.cu", "gemm.cu", "kernels.cu"]


def _deps():
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))] + \
           [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE)] if os.path.isdir(INCLUDE) else []


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources if os.path.exists(s))


def _compile(src, force):
    obj = os.path.join(OBJ, os.path.basename(src).replace(".cu", ".o"))
    if force or _stale(obj, [src] + _deps()):
        cmd = [NVCC] + ARCH + CFLAGS + ["-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    return obj


def build(force: bool = False, verbose: bool = True) -> str:
    os.makedirs(OBJ, exist_ok=True)
    srcs = sorted(set(LIB_SOURCES + SELFTEST_SOURCES))
    srcs = [os.path.join(CSRC, s) for s in srcs if os.path.exists(os.path.join(CSRC, s))]
    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = dict(zip([os.path.basename(s) for s in srcs], ex.map(lambda s: _compile(s, force), srcs)))
    lib_objs = [objs[s] for s in LIB_SOURCES if s in objs]
    if len(lib_objs) == len(LIB_SOURCES) and (force or _stale(LIB, lib_objs)):
        cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + lib_objs + ["-cudart", "static", "-ldl"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    st_objs = [objs[s] for s in SELFTEST_SOURCES if s in objs]
    if len(st_objs) == len(SELFTEST_SOURCES) and (force or _stale(SELFTEST, st_objs)):
        cmd = [NVCC] + ARCH + ["-o", SELFTEST] + st_objs + ["-cudart", "static"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("selftest link failed:\n%s\n%s" % (r.stdout, r.stderr))
    if verbose:
        print("built:", LIB if os.path.exists(LIB) else "(library sources incomplete)", SELFTEST)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv)
