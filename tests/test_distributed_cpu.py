"""world_size-2 gloo test of the N>1 host logic: contiguous sharding by global read index and
the cross-rank merge (plain sum for the accumulators; min/sum protocol for the duplication
table).  Each rank runs the CPU oracle on its shard; the merged result must equal the oracle
on the whole input (= the reference's -w 1 run)."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fqtool_b200 import batch, distributed, synth
from tests import configs


def _worker(rank, world, port, cfg, paired, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import pyoracle
    opt = configs.options(cfg)
    if paired:
        r1, r2 = synth.pe_novaseq(600, 100, seed=5, frag_mean=150, frag_sd=60)
    else:
        r1, r2 = synth.se_adapter_polyx(1500, 100, seed=6, dup_p=0.3), None
    n = len(r1[2]) - 1
    lo, hi = distributed.shard_range(n, rank, world)
    o = pyoracle.Oracle(opt.to_params(), "port")
    if paired:
        o.process_pe(batch.slice_reads(r1, lo, hi), batch.slice_reads(r2, lo, hi), lo)
    else:
        o.process_se(batch.slice_reads(r1, lo, hi), lo)
    flat = distributed.allreduce_flat(o.finalize())
    out = {"flat": flat}
    if opt.duplicate.enabled:
        # the oracle's sequential table gives (kmer, count, gc-of-first-if-min); rebuild the monoid fields
        dups, counts, gc = o.dup_table()
        keys, idx, first = _scan_keys(r1, r2, lo, hi, opt.duplicate.keylen, len(counts))
        # gc byte of the first read of each key (src/duplicate.cpp:83-91,113-127: uint8 counter, std::round)
        fgc = np.zeros(len(counts), dtype=np.uint64)
        for k in np.nonzero(first != np.uint64(0xFFFFFFFFFFFFFFFF))[0]:
            g = int(first[k])
            s1 = r1[0][int(r1[2][g]):int(r1[2][g + 1])]
            ngc = int(((s1 == ord("C")) | (s1 == ord("G"))).sum()); tot = len(s1)
            if paired:
                s2 = r2[0][int(r2[2][g]):int(r2[2][g + 1])]
                ngc += int(((s2 == ord("C")) | (s2 == ord("G"))).sum()); tot += len(s2)
            fgc[k] = int(np.floor(255.0 * (ngc & 0xFF) / tot + 0.5)) & 0xFF
        firstw = np.where(first != np.uint64(0xFFFFFFFFFFFFFFFF), (first << np.uint64(8)) | fgc, first)
        out["dup"] = distributed.allreduce_dup(dups, counts, idx, firstw)
    if rank == 0:
        q.put(out)
    dist.destroy_process_group()


_CODE = {ord("A"): 0, ord("T"): 1, ord("C"): 2, ord("G"): 3}


def _code(seq, n):
    v = 0
    for ch in seq[:n]:
        if int(ch) not in _CODE:
            return None
        v = (v << 2) | _CODE[int(ch)]
    return v


def _scan_keys(r1, r2, lo, hi, keylen, bins):
    """For every key: smallest global index carrying the minimal kmer32 (idx) and the key's first index."""
    first = np.full(bins, 0xFFFFFFFFFFFFFFFF, dtype=np.uint64)
    best = {}
    for g in range(lo, hi):
        s1 = r1[0][int(r1[2][g]):int(r1[2][g + 1])]
        if r2 is None:
            if len(s1) < 32:
                continue
            key = _code(s1, keylen); km = _code(s1[max(0, len(s1) - 37):], 32)
        else:
            s2 = r2[0][int(r2[2][g]):int(r2[2][g + 1])]
            if len(s1) < 32 or len(s2) < 32:
                continue
            key = _code(s1, keylen); km = _code(s2, 32)
        if key is None or km is None:
            continue
        key &= 0xFFFFFFFF
        if first[key] == np.uint64(0xFFFFFFFFFFFFFFFF):
            first[key] = g
        if key not in best or km < best[key][0]:
            best[key] = (km, g)
    idx = np.full(bins, 0xFFFFFFFF, dtype=np.uint32)
    for k, (_, g) in best.items():
        idx[k] = g
    return None, idx, first


@pytest.mark.parametrize("cfg,paired", [("cfg4", False), ("cfg1", True), ("pe_dup_kmer", True)])
def test_two_rank_merge_equals_single_run(cfg, paired):
    from oracle import pyoracle
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, cfg, paired, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    opt = configs.options(cfg)
    if paired:
        r1, r2 = synth.pe_novaseq(600, 100, seed=5, frag_mean=150, frag_sd=60)
    else:
        r1, r2 = synth.se_adapter_polyx(1500, 100, seed=6, dup_p=0.3), None
    o = pyoracle.Oracle(opt.to_params(), "port")
    o.process_se(r1) if r2 is None else o.process_pe(r1, r2)
    assert np.array_equal(out["flat"], o.finalize())
    if opt.duplicate.enabled:
        for a, b in zip(out["dup"], o.dup_table()):
            assert np.array_equal(a, b)
