"""Packed read batches (the fq_batch layout of include/fqk.h) and ASCII helpers.

ASCII reads are a triple ``(seq, qual, off)``: flat uint8 arrays plus uint64
offsets of n+1 entries, exactly what a FASTQ parser holds in memory.
"""
import ctypes as C
import gzip

import numpy as np

from . import abi


class Batch:
    """Host-side fq_batch: numpy arrays in the packed layout."""

    def __init__(self, offsets, lengths, bases2, qual, first_index=0):
        self.offsets, self.lengths, self.bases2, self.qual = offsets, lengths, bases2, qual
        self.first_index = int(first_index)
        self.n = len(lengths)
        self.total_bases = len(qual)

    def c_struct(self):
        b = abi.FqBatch()
        b.n_reads = self.n
        b.first_index = self.first_index
        b.total_bases = self.total_bases
        b.offsets, b.lengths = abi.ptr(self.offsets), abi.ptr(self.lengths)
        b.bases2, b.qual = abi.ptr(self.bases2), abi.ptr(self.qual)
        return b

    @property
    def n_bases(self):
        return int(self.lengths.astype(np.int64).sum())

    @property
    def nbytes(self):
        return self.offsets.nbytes + self.lengths.nbytes + self.bases2.nbytes + self.qual.nbytes


def pack(reads, first_index=0):
    """ASCII (seq, qual, off) -> Batch through the library's host packer (fqk_pack_ascii)."""
    lib = abi.load()
    seq, qual, off = reads
    seq = np.ascontiguousarray(seq, dtype=np.uint8)
    qual = np.ascontiguousarray(qual, dtype=np.uint8)
    off = np.ascontiguousarray(off, dtype=np.uint64)
    n = len(off) - 1
    total = int(lib.fqk_packed_bases(n, abi.ptr(off)))
    offsets = np.zeros(max(n, 1), dtype=np.uint32)[:n]
    lengths = np.zeros(max(n, 1), dtype=np.uint16)[:n]
    bases2 = np.zeros(total // 4, dtype=np.uint8)
    q = np.zeros(total, dtype=np.uint8)
    rc = lib.fqk_pack_ascii(n, abi.ptr(seq), abi.ptr(qual), abi.ptr(off), abi.ptr(offsets), abi.ptr(lengths),
                            abi.ptr(bases2), abi.ptr(q))
    if rc:
        raise ValueError({abi_code: msg for abi_code, msg in (
            (-4, "base outside {A,C,G,T,N} (or quality >= 128): input contract of the packer"),
            (-6, "read longer than FQK_MAX_READ_LEN or batch larger than 2^32 bases"))}.get(rc, f"pack failed rc={rc}"))
    return Batch(offsets, lengths, bases2, q, first_index)


_CODE2CHAR = np.frombuffer(b"ATCG", dtype=np.uint8)


def unpack(batch):
    """Batch -> ASCII (seq, qual, off); inverse of pack (numpy, used by tests and bench sampling)."""
    n = batch.n
    lengths = batch.lengths.astype(np.int64)
    off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lengths, out=off[1:].view(np.int64))
    total = int(off[-1])
    # position of every base in the padded buffers
    idx = np.repeat(batch.offsets.astype(np.int64) - off[:-1].astype(np.int64), lengths) + np.arange(total, dtype=np.int64)
    q = batch.qual[idx]
    codes = (batch.bases2[idx >> 2] >> ((idx & 3) * 2).astype(np.uint8)) & 3
    seq = _CODE2CHAR[codes]
    isn = (q & 0x80) != 0
    seq = np.where(isn, np.uint8(ord("N")), seq).astype(np.uint8)
    return seq, (q & 0x7F).astype(np.uint8), off


def read_fastq(path, with_names=False):
    """Minimal FASTQ(.gz) reader for tests: returns (seq, qual, off)[, names]."""
    op = gzip.open if str(path).endswith(".gz") else open
    with op(path, "rb") as f:
        lines = f.read().split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    names, seqs, quals = lines[0::4], lines[1::4], lines[3::4]
    off = np.zeros(len(seqs) + 1, dtype=np.uint64)
    np.cumsum([len(s) for s in seqs], out=off[1:].view(np.int64))
    seq = np.frombuffer(b"".join(seqs), dtype=np.uint8).copy()
    qual = np.frombuffer(b"".join(quals), dtype=np.uint8).copy()
    if with_names:
        return (seq, qual, off), names
    return seq, qual, off


def from_strings(seqs, quals):
    """Lists of str/bytes -> ASCII triple."""
    seqs = [s.encode() if isinstance(s, str) else s for s in seqs]
    quals = [q.encode() if isinstance(q, str) else q for q in quals]
    assert all(len(s) == len(q) for s, q in zip(seqs, quals))
    off = np.zeros(len(seqs) + 1, dtype=np.uint64)
    if seqs:
        np.cumsum([len(s) for s in seqs], out=off[1:].view(np.int64))
    return (np.frombuffer(b"".join(seqs), dtype=np.uint8).copy(),
            np.frombuffer(b"".join(quals), dtype=np.uint8).copy(), off)


def slice_reads(reads, lo, hi):
    seq, qual, off = reads
    a, b = int(off[lo]), int(off[hi])
    return seq[a:b], qual[a:b], (off[lo:hi + 1] - off[lo]).astype(np.uint64)
