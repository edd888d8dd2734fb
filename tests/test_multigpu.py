"""GPU (>= 2 devices): one process per GPU, contiguous shards, the NCCL all-reduce inside
fqk_finalize (sum for the accumulators, min/sum protocol for the duplication table) against
the CPU oracle run on the whole input."""
import os

import numpy as np
import pytest

from fqtool_b200 import batch, distributed, synth
from tests import configs

pytestmark = pytest.mark.gpu


def _data(paired):
    if paired:
        return synth.pe_novaseq(3000, 100, seed=5, frag_mean=150, frag_sd=60)
    return synth.se_adapter_polyx(20000, 100, seed=6, dup_p=0.3), None


def _worker(rank, world, port, cfg, paired, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fqtool_b200.engine import Engine
    uid = [Engine.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, 0)
    opt = configs.options(cfg)
    r1, r2 = _data(paired)
    n = len(r1[2]) - 1
    lo, hi = distributed.shard_range(n, rank, world)
    e = Engine(opt.to_params(), device=rank, rank=rank, nranks=world, nccl_unique_id=uid[0])
    if paired:
        e.process_pe(batch.pack(batch.slice_reads(r1, lo, hi), lo), batch.pack(batch.slice_reads(r2, lo, hi), lo))
    else:
        e.process_se(batch.pack(batch.slice_reads(r1, lo, hi), lo))
    flat = e.finalize()
    dup = e.dup_table() if opt.duplicate.enabled else None
    q.put((rank, flat, dup))
    dist.barrier()
    e.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("cfg,paired", [("cfg2", False), ("cfg4", False), ("pe_dup_kmer", True), ("cfg3", True)])
def test_nccl_merge(cfg, paired):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from oracle import pyoracle
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, world, port, cfg, paired, q)) for r in range(world)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    opt = configs.options(cfg)
    r1, r2 = _data(paired)
    o = pyoracle.Oracle(opt.to_params(), "port")
    o.process_se(r1) if r2 is None else o.process_pe(r1, r2)
    want = o.finalize()
    for rank, flat, dup in outs:                      # every rank holds the merged result
        assert np.array_equal(flat, want), f"rank {rank}"
        if dup is not None:
            for a, b in zip(dup, o.dup_table()):
                assert np.array_equal(a, b), f"rank {rank}"


def _worker_err(rank, world, port, q):
    """rank 1's shard holds a read longer than max_cycles: every rank must come back from fqk_finalize with an error."""
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fqtool_b200.engine import Engine, FqkError
    uid = [Engine.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, 0)
    opt = configs.options("cfg2", max_cycles=128)
    reads = synth.se_hiseq(2000, 150 if rank == 1 else 100, seed=11 + rank)
    e = Engine(opt.to_params(), device=rank, rank=rank, nranks=world, nccl_unique_id=uid[0])
    msgs = []
    try:
        e.process_se(batch.pack(reads, rank * 2000))
    except FqkError as ex:
        msgs.append("process: " + str(ex))
    try:
        e.finalize()
        msgs.append("finalize: ok")
    except FqkError as ex:
        msgs.append("finalize: " + str(ex))
    q.put((rank, msgs))
    dist.barrier()
    e.close()
    dist.destroy_process_group()


def test_error_on_one_shard_does_not_hang():
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31600 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker_err, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    outs = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert outs[1][0].startswith("process: fqk_process_se rc="), outs          # the shard with the over-length read fails ...
    assert outs[1][-1].startswith("finalize: fqk_finalize rc="), outs          # ... keeps failing in finalize ...
    assert "another rank reported an error" in outs[0][-1], outs               # ... and the healthy rank is told so instead of hanging
