"""Host-side mirror of the cross-rank merge that fqk_finalize performs with NCCL
(SURVEY.md section 8e), written against torch.distributed so it runs on any backend.

* the flat accumulator buffer is a plain sum (Stats::merge src/stats.cpp:815-863,
  FilterResult::merge src/filterresult.cpp:52-102);
* the duplication table is the monoid (min kmer32, count of reads carrying it, smallest
  global index among them) plus the key's first read (smallest global index, its GC byte).
"""
import numpy as np
import torch
import torch.distributed as dist

_SIGN = np.uint64(1 << 63)


def shard_range(n, rank, world):
    """Contiguous range of global read indices owned by a rank."""
    return n * rank // world, n * (rank + 1) // world


def allreduce_flat(flat):
    t = torch.from_numpy(flat.view(np.int64).copy())          # two's complement sum == uint64 sum
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.numpy().view(np.uint64)


def _min_u64(a):
    """unsigned 64-bit min over ranks with a signed reduction (flip the sign bit)."""
    t = torch.from_numpy((a ^ _SIGN).view(np.int64).copy())
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return t.numpy().view(np.uint64) ^ _SIGN


def allreduce_dup(kmer, count, idx, first):
    """kmer u64 (valid where count > 0), count u32, idx u32 (smallest global index carrying kmer),
    first u64 = (first global index of the key << 8 | gc byte), all ones where the key is empty.
    Returns the reference's (mDups, mCounts, mGC)."""
    kmin = _min_u64(np.where(count > 0, kmer, np.uint64(0xFFFFFFFFFFFFFFFF)))
    keep = (count > 0) & (kmer == kmin)
    c = torch.from_numpy(np.where(keep, count, 0).astype(np.int64)); dist.all_reduce(c, op=dist.ReduceOp.SUM)
    i = torch.from_numpy(np.where(keep, idx, 0xFFFFFFFF).astype(np.int64)); dist.all_reduce(i, op=dist.ReduceOp.MIN)
    f = _min_u64(first)
    counts = c.numpy().astype(np.uint32)
    gc = np.where((counts > 0) & ((f >> np.uint64(8)).astype(np.int64) == i.numpy()), f & np.uint64(0xFF), 0).astype(np.uint8)
    return np.where(counts > 0, kmin, 0).astype(np.uint64), counts, gc
