"""Build librtce.so (the C-ABI CUDA library) in-tree for sm_100a with nvcc.

The library is plain CUDA C++ (no torch types, no pybind): `include/rtce.h` is its whole
interface.  nvcc cross-compiles without a GPU, so this runs in the CPU-only build container.
"""
from __future__ import annotations

import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "librtce.so")
SOURCES = ["rtce_api.cu", "rtce_kernels.cu", "rtce_kernels_large.cu", "rtce_kernels_large_reg.cu", "rtce_kernels_fused.cu", "rtce_kernels_strip.cu", "rtce_kernels_coarse.cu", "rtce_kernels_shade.cu",
           "rtce_measure.cu"]
DEPS = ["rtce_api_trace.inc", "rtce_device.cuh", "rtce_march.cuh", "rtce_shade_math.cuh", "rtce_kernels.cuh", "rtce_host.hpp", "../../include/rtce.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    # canonical arithmetic: never contract mul+add into FMA; IEEE sqrt/div; no flush-to-zero
    "--fmad=false", "--prec-sqrt=true", "--prec-div=true", "--ftz=false",
    "-Xcompiler", "-fPIC,-fvisibility=default,-O2",
    "-Xptxas", "-v",
    "-shared",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    files = [os.path.join(CSRC, f) for f in SOURCES + DEPS if os.path.exists(os.path.join(CSRC, f))]
    files.append(os.path.abspath(__file__))
    return any(os.path.getmtime(f) > t for f in files)


def build_library(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return LIB_PATH
    srcs = [os.path.join(CSRC, f) for f in SOURCES if os.path.exists(os.path.join(CSRC, f))]
    env = dict(os.environ)
    env.pop("CC", None)   # the image exports CC/CXX wrappers that lack libgomp specs; use system g++
    env.pop("CXX", None)
    # kernel-tuning experiments: extra -D flags / another output file (the shipped library uses neither)
    extra = os.environ.get("RTCE_NVCC_EXTRA", "").split()
    lib_out = os.environ.get("RTCE_LIB_OUT", LIB_PATH)
    cmd = [_nvcc(), *NVCC_FLAGS, *extra, "-ccbin", "/usr/bin/g++", "-o", lib_out + ".tmp", *srcs]
    out = subprocess.run(cmd, capture_output=True, text=True, env=env)
    log = out.stdout + out.stderr
    with open(os.path.join(PKG_DIR, "build_ptxas.log"), "w") as fh:
        fh.write(" ".join(cmd) + "\n" + log)
    if out.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + log[-6000:])
    os.replace(lib_out + ".tmp", lib_out)
    if verbose:
        print(log)
    return lib_out


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
