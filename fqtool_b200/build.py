"""Build fqtool_b200/csrc/libfqk.so in-tree with nvcc for sm_100a (no JIT cache).

    python -m fqtool_b200.build [--force]
"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "libfqk.so")
SOURCES = ["fqk_api.cu", "fqk_kernels.cu", "fqk_fused.cu", "fqk_synth.cu", "fqk_prepass.cu", "fqk_text.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(HERE, "..", "include", "*.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    cmd = [nvcc] + NVCC_FLAGS + os.environ.get("FQK_NVCC_EXTRA", "").split() + ["-o", LIB] + srcs + ["-ldl"]
    if verbose:
        print(" ".join(cmd))
    subprocess.run(cmd, check=True)
    return LIB


HOST = os.path.join(HERE, "host")
HOST_EXE = os.path.join(HOST, "fqk_run")


def build_host(force=False, verbose=False):
    """C++ host adapter + command-line harness (g++), linked against libfqk.so."""
    build(force=False)
    srcs = [os.path.join(HOST, f) for f in ("fqk_host.cpp", "fqk_run.cpp")]
    deps = srcs + [os.path.join(HOST, "fqk_host.hpp"), LIB, os.path.join(HERE, "..", "include", "fqk.h")]
    if not force and os.path.exists(HOST_EXE) and all(os.path.getmtime(d) <= os.path.getmtime(HOST_EXE) for d in deps):
        return HOST_EXE
    cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(HERE, "..", "include")] + srcs + \
          ["-o", HOST_EXE, "-L" + CSRC, "-lfqk", "-lz", "-Wl,-rpath,$ORIGIN/../csrc"]
    if verbose:
        print(" ".join(cmd))
    subprocess.run(cmd, check=True)
    return HOST_EXE


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
    print(build_host(force="--force" in sys.argv, verbose=True))
