"""Builds the CUDA shared library in-tree: bioinformatics_b200/libhsq_b200.so (sm_100a only)."""
import os
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(PKG, "csrc", "hsq_capi.cu")
OUT = os.environ.get("HSQ_BUILD_OUT") or os.path.join(PKG, "libhsq_b200.so")  # HSQ_BUILD_OUT: experiment builds
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-shared"]


def sources():
    d = os.path.join(PKG, "csrc")
    inc = os.path.join(os.path.dirname(PKG), "include")
    return [os.path.join(d, f) for f in sorted(os.listdir(d))] + [os.path.join(inc, "hsq_b200.h")]


def stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(s) > t for s in sources())


def build_native(force=False, verbose=False):
    if not force and not stale():
        return OUT
    extra = os.environ.get("HSQ_NVCC_EXTRA", "").split()  # development experiments only
    cmd = [NVCC] + FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT, SRC]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return OUT


if __name__ == "__main__":
    build_native(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(OUT)
