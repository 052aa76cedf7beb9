"""In-tree build of libpathtrace_b200.so (nvcc, sm_100a only).

The .so lands in pathtrace_b200/lib/ so it travels with the repo snapshot to the GPU box; it is
git-ignored.  nvcc cross-compiles without a GPU, so this runs on the CPU-only build container too.
"""
from __future__ import annotations

import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIBDIR = os.path.join(_HERE, "lib")
LIB = os.environ.get("PT_LIB") or os.path.join(LIBDIR, "libpathtrace_b200.so")   # PT_LIB: development override (kernel variants)
SOURCES = ["capi.cu", "lbvh.cu", "wavefront.cu", "smem_wavefront.cu", "group.cu"]
HEADERS = ["pt_internal.h", "pt_math.cuh", "pt_shade.cuh", "pt_wave.cuh", os.path.join("..", "..", "include", "pathtrace_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
              "-ccbin", "g++"]
# The render kernels compile plain `/` and sqrtf() to the approximate MUFU sequences (2 ulp): that is the free-form shading
# math (statistical parity only).  Everything under the bit-exact contract (camera rays, intersection, LBVH) is written with
# explicit __f*_rn intrinsics, which these flags do not touch; denormals and FMA contraction keep their defaults.
FAST_SHADING_FLAGS = ["-prec-div=false", "-prec-sqrt=false"]
FAST_SHADING_SOURCES = {"wavefront.cu", "smem_wavefront.cu"}


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    raise RuntimeError("nvcc not found")


def is_stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(LIBDIR, "obj")
    os.makedirs(objdir, exist_ok=True)
    nvcc = _nvcc()

    def compile_one(src):
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        fast = FAST_SHADING_FLAGS if src in FAST_SHADING_SOURCES and not os.environ.get("PT_PRECISE_SHADING") else []
        cmd = [nvcc, *NVCC_FLAGS, *fast, *os.environ.get("PT_NVCC_EXTRA", "").split(), "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(r.stderr)
        return obj

    with ThreadPoolExecutor(len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    r = subprocess.run([nvcc, "-shared", "-ccbin", "g++", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-ldl"],
                       capture_output=True, text=True)
    if r.returncode:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


HOST_DIR = os.path.join(os.path.dirname(_HERE), "host")
HOST_BIN = os.path.join(LIBDIR, "pathtrace_host")


def build_host(force: bool = False) -> str:
    """The compiled-language host side above the C ABI (host/pathtrace_host.hpp: the reference's World / Object / Camera API in
    C++17) and its headless `main` (host/main.cpp, the stand-in for src/main.rs), linked against the in-tree library."""
    lib = build_library()
    srcs = [os.path.join(HOST_DIR, "main.cpp"), os.path.join(HOST_DIR, "pathtrace_host.hpp"), os.path.join(_HERE, "..", "include", "pathtrace_b200.h")]
    if not force and os.path.exists(HOST_BIN) and all(os.path.getmtime(HOST_BIN) >= os.path.getmtime(d) for d in srcs + [lib]):
        return HOST_BIN
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", srcs[0], "-o", HOST_BIN, "-L" + LIBDIR, "-lpathtrace_b200", "-Wl,-rpath,$ORIGIN"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode:
        raise RuntimeError(f"host build failed:\n{r.stdout}\n{r.stderr}")
    return HOST_BIN


if __name__ == "__main__":
    import sys
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_host(force="--force" in sys.argv))
