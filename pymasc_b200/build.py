"""Compile libpymasc_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU), and
libpymasc_io.so (host-side BAM / BigWig decoding, g++ + zlib).

    python -m pymasc_b200.build [--force]
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libpymasc_b200.so")
SOURCES = ["pmb_api.cu"]
HEADERS = ["pmb_build.cuh", "pmb_window.cuh", "pmb_probe.cuh", os.path.join("..", "..", "include", "pymasc_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--shared", "-Xcompiler", "-fPIC"]


IO_OUT = os.path.join(HERE, "libpymasc_io.so")
IO_SOURCES = ["pmb_io.cpp"]
IO_DEPS = IO_SOURCES + ["pmb_inflate.h", "pmb_crc32.h", os.path.join("..", "..", "include", "pymasc_io.h")]
IO_FLAGS = ["-O3", "-std=c++17", "-shared", "-fPIC", "-Wall"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    # every header next to the sources counts (a stale library behind a changed .cuh is a silent lie)
    deps = set(SOURCES + HEADERS) | {f for f in os.listdir(CSRC) if f.endswith((".cuh", ".cu", ".h"))}
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in deps)


def build_io(force=False):
    """libpymasc_io.so: plain C++ (no CUDA), links zlib and pthreads."""
    if not force and os.path.exists(IO_OUT) and all(
            os.path.getmtime(os.path.join(CSRC, f)) <= os.path.getmtime(IO_OUT) for f in IO_DEPS):
        return IO_OUT
    cxx = shutil.which("g++") or "g++"
    env = dict(os.environ)
    env.pop("CC", None)
    subprocess.check_call([cxx] + IO_FLAGS + ["-o", IO_OUT] + IO_SOURCES + ["-lz", "-lpthread"], cwd=CSRC, env=env)
    return IO_OUT


def build(force=False, verbose=False):
    build_io(force)
    if not force and not stale():
        return OUT
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT] + SOURCES
    env = dict(os.environ)
    env.pop("CC", None)   # the image exports a gcc wrapper; let nvcc pick the system host compiler
    subprocess.check_call(cmd, cwd=CSRC, env=env)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
