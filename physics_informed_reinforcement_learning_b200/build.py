"""Build the C-ABI CUDA library in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo).

Three translation units (compiled to objects under csrc/build/, linked into libpirl_rcwa.so):
  rcwa_api.cu      C ABI + the single-SM kernels (rcwa_kernels / rcwa_regtile / rcwa_tail2 / rcwa_grad)
  rcwa_cluster.cu  thread-block-cluster kernels for the large orders (rcwa_cluster.cuh)
  rcwa_smm.cu      scattering-matrix (algo = SMM) layer solve (rcwa_smm.cuh)
"""
from __future__ import annotations

import os
import shutil
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
OBJ_DIR = os.path.join(CSRC, "build")
LIB_PATH = os.path.join(PKG_DIR, "libpirl_rcwa.so")
HEADER = os.path.join("..", "..", "include", "pirl_rcwa.h")
# source -> files whose change makes its object stale
UNITS = {
    "rcwa_api.cu": ["rcwa_api.cu", "rcwa_common.cuh", "rcwa_kernels.cuh", "rcwa_regtile.cuh", "rcwa_tail2.cuh",
                    "rcwa_grad.cuh", "rcwa_cluster.h", "rcwa_smm.h", HEADER],
    "rcwa_cluster.cu": ["rcwa_cluster.cu", "rcwa_cluster.cuh", "rcwa_cluster_tail.cuh", "rcwa_cluster.h", "rcwa_common.cuh"],
    "rcwa_smm.cu": ["rcwa_smm.cu", "rcwa_smm.cuh", "rcwa_smm.h", "rcwa_common.cuh"],
}
SOURCES = list(UNITS)

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the CUDA library cannot be built (there is no CPU fallback)")


def _obj(src: str) -> str:
    return os.path.join(OBJ_DIR, os.path.splitext(src)[0] + ".o")


def _mtime(rel: str) -> float:
    p = os.path.join(CSRC, rel)
    return os.path.getmtime(p) if os.path.exists(p) else 0.0


def _unit_stale(src: str) -> bool:
    o = _obj(src)
    if not os.path.exists(o):
        return True
    t = os.path.getmtime(o)
    return any(_mtime(d) > t for d in UNITS[src])


def is_stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(_mtime(d) > t for deps in UNITS.values() for d in deps)


def _run(cmd, verbose):
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)


def build_library(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB_PATH
    nvcc = find_nvcc()
    os.makedirs(OBJ_DIR, exist_ok=True)
    procs = []
    for src in SOURCES:
        if force or _unit_stale(src):
            extra = os.environ.get("PIRL_NVCC_EXTRA", "").split()      # e.g. -DPIRL_CK_PRODUCT_MMA (experiments)
            cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas=-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", _obj(src)]
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)))
    for src, p in procs:       # the units compile in parallel
        out, err = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n" + out + err)
        if verbose:
            print(err)
    _run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC"] + [_obj(s) for s in SOURCES] + ["-o", LIB_PATH], verbose)
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
