"""TEST INFRASTRUCTURE: build and call the CPU emulation of the fused paired-end kernel's trim role
(tests/emu/emu_pe.cpp compiles fqtool_b200/csrc/fqk_pe4.cuh with g++, 32 lanes = 32 fibers in lock step)."""
import ctypes as C
import os
import subprocess

import numpy as np

from fqtool_b200 import abi, batch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libemu_pe.so")
ROOT = os.path.dirname(os.path.dirname(HERE))


def build():
    src = os.path.join(HERE, "emu_pe.cpp")
    csrc = os.path.join(ROOT, "fqtool_b200", "csrc")
    deps = [src, os.path.join(HERE, "cuda_emu.h")] + [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith((".cuh", ".h"))]
    if os.path.exists(LIB) and all(os.path.getmtime(d) <= os.path.getmtime(LIB) for d in deps):
        return LIB
    subprocess.run(["g++", "-std=c++17", "-O2", "-g", "-w", "-fPIC", "-shared", "-fvisibility=hidden", "-Wl,-Bsymbolic", "-I" + os.path.join(ROOT, "include"),
                    src, "-o", LIB], check=True)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.emu_pe_run.argtypes = [C.c_void_p] * 7 + [C.c_int, C.c_int]
        _lib.emu_pe_run.restype = C.c_int
    return _lib


def run_pe(opt, r1, r2, chunks=1, tile=64):
    """Same result triple as tests.util.run_cuda for a paired-end option set, computed by the emulated kernel."""
    params = opt.to_params()
    L = abi.FqStatsLayout()
    abi.load().fqk_stats_layout(C.byref(params), C.byref(L))
    flat = np.zeros(L.total_u64, dtype=np.uint64)
    n = len(r1[2]) - 1
    bounds = np.linspace(0, n, chunks + 1).astype(int)
    outs = []
    for lo, hi in zip(bounds[:-1], bounds[1:]):
        b1 = batch.pack(batch.slice_reads(r1, lo, hi), int(lo)); b2 = batch.pack(batch.slice_reads(r2, lo, hi), int(lo))
        m = b1.n
        o1 = np.zeros(m, dtype=abi.READ_RESULT_DTYPE); o2 = np.zeros(m, dtype=abi.READ_RESULT_DTYPE)
        op = np.zeros(m, dtype=abi.PAIR_RESULT_DTYPE)
        if m:
            maxlen = int(max(b1.lengths.max(), b2.lengths.max()))
            s1, s2 = b1.c_struct(), b2.c_struct()
            rc = lib().emu_pe_run(C.addressof(params), C.addressof(s1), C.addressof(s2), abi.ptr(o1), abi.ptr(o2), abi.ptr(op),
                                  abi.ptr(flat), tile, maxlen)
            if rc:
                raise RuntimeError(f"emulated kernel raised the device error flag {rc}")
        outs.append((o1, o2, op))
    recs = tuple(np.concatenate([o[i] for o in outs]) for i in range(3))
    return recs, flat, None
