"""Build librtce.so (the C-ABI CUDA library) in-tree for sm_100a with nvcc.

The library is plain CUDA C++ (no torch types, no pybind): `include/rtce.h` is its whole
interface.  nvcc cross-compiles without a GPU, so this runs in the CPU-only build container.
Every translation unit is compiled on its own (`nvcc -c`, in parallel, only when stale) and the
objects are linked into the shared library; NCCL is NOT linked (rtce_comm.cu loads it on first use).
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "librtce.so")
OBJ_DIR = os.path.join(PKG_DIR, "build")
SOURCES = ["rtce_api.cu", "rtce_api_trace.cu", "rtce_comm.cu", "rtce_camera.cu", "rtce_kernels.cu", "rtce_kernels_accum.cu",
           "rtce_kernels_large_reg.cu", "rtce_kernels_large_ev.cu", "rtce_kernels_strip.cu", "rtce_kernels_shade.cu",
           "rtce_measure.cu", "rtce_noalloc.cu"]
DEPS = ["rtce_ctx.hpp", "rtce_device.cuh", "rtce_march.cuh", "rtce_shade_math.cuh", "rtce_kernels.cuh", "rtce_host.hpp",
        "../../include/rtce.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    # canonical arithmetic: never contract mul+add into FMA; IEEE sqrt/div; no flush-to-zero
    "--fmad=false", "--prec-sqrt=true", "--prec-div=true", "--ftz=false",
    "-Xcompiler", "-fPIC,-fvisibility=default,-O2,-ffp-contract=off",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _sources():
    return [f for f in SOURCES if os.path.exists(os.path.join(CSRC, f))]


def _digest(files, extra=()) -> str:
    """Content hash of the inputs of a build step (mtimes do not survive the copy to the GPU box)."""
    h = hashlib.sha1()
    for f in files:
        h.update(os.path.basename(f).encode())
        if os.path.exists(f):
            with open(f, "rb") as fh:
                h.update(fh.read())
    h.update(" ".join([*NVCC_FLAGS, *extra]).encode())
    return h.hexdigest()


def _stale(target: str, digest: str) -> bool:
    stamp = target + ".sha1"
    if not os.path.exists(target) or not os.path.exists(stamp):
        return True
    with open(stamp) as fh:
        return fh.read().strip() != digest


def _stamp(target: str, digest: str) -> None:
    with open(target + ".sha1", "w") as fh:
        fh.write(digest)


def build_library(force: bool = False, verbose: bool = False) -> str:
    # kernel-tuning experiments: extra -D flags / another output file (the shipped library uses neither)
    extra = os.environ.get("RTCE_NVCC_EXTRA", "").split()
    lib_out = os.environ.get("RTCE_LIB_OUT", LIB_PATH)
    tag = hashlib.sha1(" ".join(extra).encode()).hexdigest()[:8] if extra else "default"
    obj_dir = os.path.join(OBJ_DIR, tag)
    os.makedirs(obj_dir, exist_ok=True)
    deps = [os.path.join(CSRC, d) for d in DEPS] + [os.path.abspath(__file__)]
    env = dict(os.environ)
    env.pop("CC", None)   # the image exports CC/CXX wrappers that lack libgomp specs; use system g++
    env.pop("CXX", None)
    nvcc = _nvcc()
    all_inputs = [os.path.join(CSRC, f) for f in _sources()] + deps
    lib_digest = _digest(all_inputs, extra)
    if not force and not _stale(lib_out, lib_digest):
        return lib_out          # e.g. on the GPU box: the shipped library matches the shipped sources
    jobs = []
    digests = {}
    for f in _sources():
        src = os.path.join(CSRC, f)
        obj = os.path.join(obj_dir, f[:-3] + ".o")
        digests[obj] = _digest([src] + deps, extra)
        if force or _stale(obj, digests[obj]):
            jobs.append((f, [nvcc, *NVCC_FLAGS, *extra, "-ccbin", "/usr/bin/g++", "-c", src, "-o", obj]))
    logs = {}

    def run(job):
        name, cmd = job
        out = subprocess.run(cmd, capture_output=True, text=True, env=env)
        return name, cmd, out

    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            for name, cmd, out in ex.map(run, jobs):
                logs[name] = " ".join(cmd) + "\n" + out.stdout + out.stderr
                with open(os.path.join(obj_dir, name[:-3] + ".log"), "w") as fh:
                    fh.write(logs[name])
                if out.returncode != 0:
                    raise RuntimeError(f"nvcc failed on {name}:\n" + logs[name][-6000:])
                _stamp(os.path.join(obj_dir, name[:-3] + ".o"), digests[os.path.join(obj_dir, name[:-3] + ".o")])
    objs = [os.path.join(obj_dir, f[:-3] + ".o") for f in _sources()]
    if True:
        cmd = [nvcc, "-shared", "-ccbin", "/usr/bin/g++", "-o", lib_out + ".tmp", *objs, "-ldl"]
        out = subprocess.run(cmd, capture_output=True, text=True, env=env)
        if out.returncode != 0:
            raise RuntimeError("link failed:\n" + (out.stdout + out.stderr)[-6000:])
        os.replace(lib_out + ".tmp", lib_out)
        _stamp(lib_out, lib_digest)
        # one ptxas log of the whole library (registers / spills per kernel), regenerated from the per-file logs
        with open(os.path.join(PKG_DIR, "build_ptxas.log"), "w") as fh:
            for f in _sources():
                p = os.path.join(obj_dir, f[:-3] + ".log")
                if os.path.exists(p):
                    fh.write(open(p).read())
    if verbose:
        for name in logs:
            print(logs[name])
    return lib_out


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
