"""Adapter auto-detection pre-pass: host mirror of Evaluator::evaluateAdapterSeq
(/root/reference/src/evaluator.cpp:229-390) and Evaluator::getAdapterWithSeed (src/evaluator.cpp:392-425).

The two per-base loops (10-mer counting over the first 256 K reads; the scan for windows equal to a seed) run on the
GPU through fqk_prepass_kmer_counts / fqk_prepass_seed_hits (include/fqk.h).  What stays here is what the reference
does on a handful of values: ranking 4^10 counts, and the NucleotideTree walk (src/nucleotidetree.cpp:40-90) over the
reads that carry the seed.  The table of known Illumina adapters (src/knownadapters.h) is the reference's data and is
not reproduced: `known` may be passed by the caller; the drop-in build uses the reference's own matchKnownAdapter.
"""
import ctypes as C

import numpy as np

from . import abi

READ_LIMIT = 256 * 1024            # src/evaluator.cpp:235-236
BASE_LIMIT = 151 * READ_LIMIT
KEYLEN, FIRST_POS, TOPNUM, FOLD_THRESHOLD = 10, 20, 10, 20


def _ascii_batch(reads):
    seq = np.ascontiguousarray(reads[0])
    off = np.ascontiguousarray(reads[2], dtype=np.uint64)
    b = abi.FqAsciiBatch()
    b.n_reads = len(off) - 1; b.first_index = 0
    b.off, b.seq, b.qual = abi.ptr(off), abi.ptr(seq), None
    return b, (seq, off)


def kmer_counts(reads, keylen=KEYLEN, first_pos=FIRST_POS, shift_tail=1, device=0):
    """counts[4^keylen] of all-ATCG windows (device)."""
    lib = abi.load()
    b, keep = _ascii_batch(reads)
    counts = np.zeros(1 << (2 * keylen), dtype=np.uint64)
    rc = lib.fqk_prepass_kmer_counts(device, C.byref(b), keylen, first_pos, shift_tail, abi.ptr(counts))
    if rc:
        raise RuntimeError(f"fqk_prepass_kmer_counts rc={rc}")
    return counts


def seed_hits(reads, seed, keylen=KEYLEN, first_pos=FIRST_POS, shift_tail=1, device=0):
    """(read, pos) of the windows whose key equals seed, sorted (device)."""
    lib = abi.load()
    b, keep = _ascii_batch(reads)
    cap = 1 << 16
    while True:
        ri = np.zeros(cap, dtype=np.uint32); po = np.zeros(cap, dtype=np.uint32); n = C.c_uint32(0)
        rc = lib.fqk_prepass_seed_hits(device, C.byref(b), keylen, first_pos, shift_tail, int(seed), abi.ptr(ri), abi.ptr(po),
                                       cap, C.byref(n))
        if rc:
            raise RuntimeError(f"fqk_prepass_seed_hits rc={rc}")
        if n.value <= cap:
            break
        cap = n.value
    order = np.lexsort((po[:n.value], ri[:n.value]))
    return ri[:n.value][order], po[:n.value][order]


def int2seq(val, n):                         # Evaluator::int2seq, src/evaluator.cpp:51-61
    return "".join("ATCG"[(val >> (2 * (n - 1 - i))) & 3] for i in range(n))


def rank_keys(counts, keylen=KEYLEN):
    """The top-10 selection of src/evaluator.cpp:285-337: keys that are not low-complexity (a base six or more times), not
    GC-rich (C+G >= 8) and do not start with GGGG, ordered by count; among equal counts the LATER key comes first (the
    insertion loop compares with `<`).  Returns (top keys, total of the eligible counts)."""
    size = 1 << (2 * keylen)
    k = np.arange(size, dtype=np.int64)
    per = np.zeros((4, size), dtype=np.int32)
    for i in range(keylen):
        d = (k >> (2 * i)) & 3
        for b in range(4):
            per[b] += (d == b)
    ok = (per.max(axis=0) < keylen - 4) & (per[2] + per[3] < keylen - 2) & ((k >> 12) != 0xFF)
    c = counts.astype(np.int64).copy()
    c[0] = 0
    total = int(c[ok].sum())
    ks = k[ok]
    order = np.lexsort((-ks, -c[ok]))[:TOPNUM]
    return [int(x) for x in ks[order]], total, c


class _Tree:                                  # NucleotideTree, src/nucleotidetree.cpp:40-90
    def __init__(self):
        self.root = {}

    def add(self, s):
        node = self.root
        for ch in s:
            if ch == "N":
                break
            slot = ord(ch) & 7
            nxt = node.get(slot)
            if nxt is None:
                nxt = node[slot] = [0, ch, {}]
            nxt[0] += 1
            node = nxt[2]

    def dominant(self):
        out, node, reached = [], self.root, True
        while True:
            total = sum(v[0] for v in node.values())
            if total < 50:
                break
            nxt = None
            for slot in range(8):
                v = node.get(slot)
                if v is not None and v[0] / total >= 0.95:
                    nxt = v; break
            if nxt is None:
                reached = False
                break
            out.append(nxt[1]); node = nxt[2]
        return "".join(out), reached


def adapter_with_seed(reads, seed, trim_tail1=0, keylen=KEYLEN, device=0, known=(), hits_fn=None):
    """Evaluator::getAdapterWithSeed: adapter grown from the reads that carry the seed ('' when no dominant path).
    hits_fn replaces the device scan (tests hand the CPU oracle in to check the host logic without a GPU)."""
    shift = max(1, trim_tail1)
    seq, _, off = reads
    ri, po = (hits_fn(reads, seed, keylen, FIRST_POS, shift) if hits_fn else seed_hits(reads, seed, keylen, FIRST_POS, shift, device))
    fwd, bwd = _Tree(), _Tree()
    for r, p in zip(ri.tolist(), po.tolist()):
        s = bytes(seq[int(off[r]):int(off[r + 1])]).decode()
        fwd.add(s[p + keylen:p + keylen + max(0, len(s) - keylen - shift - p)])
        bwd.add(s[:p][::-1])
    f, r1 = fwd.dominant()
    b, r2 = bwd.dominant()
    adapter = (b[::-1] + int2seq(seed, keylen) + f)[:60]
    for ka in sorted(known):                  # matchKnownAdapter, src/evaluator.cpp:427-445 (std::map order)
        if len(adapter) >= len(ka) and adapter[:len(ka)] == ka:
            return ka
    return adapter if (r1 and r2) else ""


def detect_adapter(reads, trim_tail1=0, device=0, known=(), counts_fn=None, hits_fn=None):
    """Evaluator::evaluateAdapterSeq on reads already in memory (the caller applies READ_LIMIT / BASE_LIMIT).
    counts_fn / hits_fn replace the two device calls (CPU tests of the host logic)."""
    n = len(reads[2]) - 1
    if n < 10000:
        return ""
    shift = max(1, trim_tail1)
    counts = counts_fn(reads, KEYLEN, FIRST_POS, shift) if counts_fn else kmer_counts(reads, KEYLEN, FIRST_POS, shift, device)
    top, total, c = rank_keys(counts)
    size = 1 << (2 * KEYLEN)
    for key in top:
        if key == 0:
            continue
        cnt = int(c[key])
        if cnt < 10 or cnt * size < total * FOLD_THRESHOLD:
            break
        s = int2seq(key, KEYLEN)
        if sum(1 for i in range(len(s) - 1) if s[i] != s[i + 1]) < 3:
            continue
        a = adapter_with_seed(reads, key, trim_tail1, KEYLEN, device, known, hits_fn)
        if a:
            return a
    return ""
