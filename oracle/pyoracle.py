"""TEST INFRASTRUCTURE ONLY: ctypes access to the CPU oracle.

``Oracle("port")`` drives oracle/libfqoracle.so (the plain-C restatement),
``Oracle("reference")`` drives oracle/_ref/libfqref.so (the reference's own
unmodified C++ behind oracle/ref_driver.cpp).  Only tests/, smoke() and
bench.py's CPU-baseline legs import this module; fqtool_b200/ never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from fqtool_b200 import abi

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_LIB = os.path.join(HERE, "libfqoracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libfqref.so")


def build(quiet=True):
    """make the port and, when /root/reference is present, the reference build."""
    subprocess.run(["make", "-C", HERE, "-j8"], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


def have_reference():
    return os.path.exists(REF_LIB)


_libs = {}


def _load(kind):
    if kind in _libs:
        return _libs[kind]
    path = PORT_LIB if kind == "port" else REF_LIB
    if not os.path.exists(path):
        raise RuntimeError(f"{path} missing: run `make -C oracle`")
    lib = C.CDLL(path)
    pre = "fqo_" if kind == "port" else "fqr_"
    vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
    f = getattr(lib, pre + "create"); f.argtypes = [C.POINTER(abi.FqParams)]; f.restype = vp
    f = getattr(lib, pre + "destroy"); f.argtypes = [vp]; f.restype = None
    f = getattr(lib, pre + "process_se"); f.argtypes = [vp, u32, u64, vp, vp, vp, vp]; f.restype = C.c_int
    f = getattr(lib, pre + "process_pe"); f.argtypes = [vp, u32, u64, vp, vp, vp, vp, vp, vp, vp, vp, vp]; f.restype = C.c_int
    f = getattr(lib, pre + "finalize"); f.argtypes = [vp, vp]; f.restype = C.c_int
    f = getattr(lib, pre + "dup_table"); f.argtypes = [vp, vp, vp, vp]; f.restype = C.c_int
    f = getattr(lib, pre + "get_corrections"); f.argtypes = [vp, vp, u32, C.POINTER(u32)]; f.restype = C.c_int
    lib.fqo_layout.argtypes = [C.POINTER(abi.FqParams), C.POINTER(abi.FqStatsLayout)]
    lib.fqo_layout.restype = C.c_int
    if kind == "reference":
        lib.fqr_bench.argtypes = [C.POINTER(abi.FqParams), C.c_int, u32, vp, vp, vp, vp, vp, vp,
                                  C.POINTER(C.c_double), C.POINTER(u64)]
        lib.fqr_bench.restype = C.c_int
    _libs[kind] = (lib, pre)
    return _libs[kind]


def layout(params):
    lib, _ = _load("port")
    L = abi.FqStatsLayout()
    lib.fqo_layout(C.byref(params), C.byref(L))
    return L


class Oracle:
    """One run of the CPU path: process batches in input order, then finalize()."""

    def __init__(self, params, kind="port"):
        self.lib, self.pre = _load(kind)
        self.kind = kind
        self.params = params
        self.layout = layout(params)
        self.h = getattr(self.lib, self.pre + "create")(C.byref(params))
        if not self.h:
            raise RuntimeError("oracle create failed")

    def close(self):
        if self.h:
            getattr(self.lib, self.pre + "destroy")(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def process_se(self, reads, first_index=0):
        seq, qual, off = reads
        n = len(off) - 1
        out = np.zeros(n, dtype=abi.READ_RESULT_DTYPE)
        rc = getattr(self.lib, self.pre + "process_se")(self.h, n, first_index, abi.ptr(seq), abi.ptr(qual),
                                                        abi.ptr(off), abi.ptr(out))
        if rc:
            raise RuntimeError(f"oracle process_se rc={rc}")
        return out

    def process_pe(self, reads1, reads2, first_index=0):
        s1, q1, o1 = reads1
        s2, q2, o2 = reads2
        n = len(o1) - 1
        assert len(o2) - 1 == n
        out1 = np.zeros(n, dtype=abi.READ_RESULT_DTYPE)
        out2 = np.zeros(n, dtype=abi.READ_RESULT_DTYPE)
        outp = np.zeros(n, dtype=abi.PAIR_RESULT_DTYPE)
        rc = getattr(self.lib, self.pre + "process_pe")(
            self.h, n, first_index, abi.ptr(s1), abi.ptr(q1), abi.ptr(o1), abi.ptr(s2), abi.ptr(q2), abi.ptr(o2),
            abi.ptr(out1), abi.ptr(out2), abi.ptr(outp))
        if rc:
            raise RuntimeError(f"oracle process_pe rc={rc}")
        return out1, out2, outp

    def corrections(self):
        """Edits of the last process_pe call, sorted (pair, mate, pos)."""
        n = C.c_uint32(0)
        f = getattr(self.lib, self.pre + "get_corrections")
        f(self.h, None, 0, C.byref(n))
        out = np.zeros(n.value, dtype=abi.CORRECTION_DTYPE)
        f(self.h, abi.ptr(out) if n.value else None, n.value, C.byref(n))
        return np.sort(out, order=["pair", "mate_base", "pos"])

    def finalize(self):
        flat = np.zeros(self.layout.total_u64, dtype=np.uint64)
        getattr(self.lib, self.pre + "finalize")(self.h, abi.ptr(flat))
        return flat

    def dup_table(self):
        n = self.layout.dup_bins
        dups = np.zeros(n, dtype=np.uint64); counts = np.zeros(n, dtype=np.uint32); gc = np.zeros(n, dtype=np.uint8)
        rc = getattr(self.lib, self.pre + "dup_table")(self.h, abi.ptr(dups), abi.ptr(counts), abi.ptr(gc))
        if rc:
            raise RuntimeError("dup table unavailable")
        return dups, counts, gc


def bench_reference(params, threads, reads1, reads2=None):
    """Timed multi-threaded run of the reference's own hot path (fqr_bench).
    Returns (seconds, passed)."""
    lib, _ = _load("reference")
    s1, q1, o1 = reads1
    n = len(o1) - 1
    sec = C.c_double(0); passed = C.c_uint64(0)
    if reads2 is None:
        a = (None, None, None)
    else:
        a = tuple(abi.ptr(x) for x in reads2)
    rc = lib.fqr_bench(C.byref(params), threads, n, abi.ptr(s1), abi.ptr(q1), abi.ptr(o1), a[0], a[1], a[2],
                       C.byref(sec), C.byref(passed))
    if rc:
        raise RuntimeError("fqr_bench failed")
    return sec.value, passed.value


# ---------------------------------------------------------------- adapter detection pre-pass
def adapter_kmer_counts(reads, keylen=10, first_pos=20, shift_tail=1, kind="port"):
    """counts[4^keylen]: oracle port (kind="port") or the loop run with the reference's own Evaluator::seq2int."""
    import numpy as np
    lib = C.CDLL(PORT_LIB if kind == "port" else REF_LIB)
    f = lib.fqo_adapter_kmer_counts if kind == "port" else lib.fqr_adapter_kmer_counts
    seq = np.ascontiguousarray(reads[0]); off = np.ascontiguousarray(reads[2], dtype=np.uint64)
    counts = np.zeros(1 << (2 * keylen), dtype=np.uint64)
    f.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    rc = f(len(off) - 1, abi.ptr(seq), abi.ptr(off), keylen, first_pos, shift_tail, abi.ptr(counts))
    assert rc == 0
    return counts


def adapter_seed_hits(reads, seed, keylen=10, first_pos=20, shift_tail=1):
    import numpy as np
    lib = C.CDLL(PORT_LIB)
    seq = np.ascontiguousarray(reads[0]); off = np.ascontiguousarray(reads[2], dtype=np.uint64)
    n = C.c_uint32(0)
    f = lib.fqo_adapter_seed_hits
    f.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
    f(len(off) - 1, abi.ptr(seq), abi.ptr(off), keylen, first_pos, shift_tail, seed, None, None, 0, C.byref(n))
    ri = np.zeros(max(n.value, 1), dtype=np.uint32); po = np.zeros(max(n.value, 1), dtype=np.uint32)
    f(len(off) - 1, abi.ptr(seq), abi.ptr(off), keylen, first_pos, shift_tail, seed, abi.ptr(ri), abi.ptr(po), n.value, C.byref(n))
    return ri[:n.value], po[:n.value]


def reference_adapter_with_seed(reads, seed, keylen=10, trim_tail1=0):
    """Evaluator::getAdapterWithSeed of the reference build on these reads."""
    import numpy as np
    lib = C.CDLL(REF_LIB)
    seq = np.ascontiguousarray(reads[0]); off = np.ascontiguousarray(reads[2], dtype=np.uint64)
    out = C.create_string_buffer(256)
    f = lib.fqr_adapter_with_seed
    f.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_uint32]
    assert f(len(off) - 1, abi.ptr(seq), abi.ptr(off), seed, keylen, trim_tail1, out, 256) == 0
    return out.value.decode()
