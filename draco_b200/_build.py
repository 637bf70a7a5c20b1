"""In-tree build of ``libdgibbs.so`` (hand-written sm_100a CUDA + the C ABI) with nvcc.

The shared object is written next to this file so it travels with the source tree; it is
never installed into site-packages.  ``build()`` is incremental (per-object timestamps).
"""

from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdgibbs.so")
# NRML likelihood object: separate library because it links cuBLAS / cuSOLVER
NRML_LIB = os.path.join(HERE, "libdgnrml.so")
NRML_SOURCES = ["nrml.cu", "wiener_op.cu", "wiener_own.cu", "nrml_own.cu"]
NRML_HEADERS = [os.path.join("..", "..", "include", "dgnrml.h"), os.path.join("..", "..", "include", "dgwiener.h"), "wiener_own.h", "nrml_own.h", "ozaki_gemm.cuh",
                "dense_core.cuh", "tma_pipe.cuh", "tma_host.h", "common.cuh", "internal.h"]
SOURCES = ["dgibbs_api.cu", "setup_kernels.cu", "sampler_dense.cu", "sampler_diag.cu", "sampler_cross.cu", "sampler_freq.cu", "project_fft.cu", "cut_kernels.cu", "finish_kernels.cu", "delay_fft.cu"]
HEADERS = ["common.cuh", "internal.h", "dense_core.cuh", "tma_pipe.cuh", "tma_host.h", "window.cuh", "fft_fused.cuh", os.path.join("..", "..", "include", "dgibbs.h")]
NVCC_FLAGS = [
    "-gencode",
    "arch=compute_100a,code=sm_100a",
    "-O3",
    "-lineinfo",
    "-std=c++17",
    "-Xcompiler",
    "-fPIC",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libdgibbs.so")


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a and link ``libdgibbs.so``.  Returns its path."""
    nvcc = _nvcc()
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in HEADERS]
    objs = []
    procs = []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s, *hdrs]):
            cmd = [nvcc, *NVCC_FLAGS, "-c", s, "-o", o]
            if verbose:
                print(" ".join(cmd))
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    nobjs = []
    for src in NRML_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src.replace(".cu", ".o"))
        nobjs.append(o)
        if force or _stale(o, [s, *[os.path.join(CSRC, h) for h in NRML_HEADERS]]):
            cmd = [nvcc, *NVCC_FLAGS, "-c", s, "-o", o]
            if verbose:
                print(" ".join(cmd))
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for cmd, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd) + "\n" + out.decode())
    if force or procs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-o", LIB, *objs]
        subprocess.run(cmd, check=True)
    if force or procs or _stale(NRML_LIB, nobjs):
        cmd = [nvcc, "-shared", "-o", NRML_LIB, *nobjs, "-lcublas", "-lcusolver"]
        subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(verbose=True))
