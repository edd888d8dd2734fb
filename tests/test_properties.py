"""GPU, bench-sized batches: size-independent properties of the hot path on device-generated
workloads (millions of reads, far beyond what the CPU oracle checks in seconds):
  * two independent CUDA implementations (SWAR warp-specialised kernel vs generic kernels) agree
    bit-exactly on records, accumulators and duplication tables;
  * linearity: accumulators of one big batch == sum over its halves (Stats::merge is a plain sum);
  * conservation: filter counters sum to the reads, per-cycle totals are consistent with lengths;
  * a random 50k-read slice of the same batch matches the CPU oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [("cfg2", 2_000_000), ("cfg4", 1_000_000), ("cfg3", 600_000), ("cfg5", 300_000)]


def _run(cfg, n, impl, halves=False):
    import torch
    from fqtool_b200 import abi, workload
    from fqtool_b200.engine import Engine
    dev = torch.device("cuda", 0)
    opt = workload.options_for(cfg)
    e = Engine(opt.to_params(), device=0)
    e.set_impl(impl)
    parts = [(0, n)] if not halves else [(0, n // 2), (n // 2, n - n // 2)]
    recs = []
    keep = None
    for first, cnt in parts:
        b1, b2 = workload.generate(cfg, cnt, first, dev, e.stream)
        o1 = torch.zeros(cnt * 16, dtype=torch.uint8, device=dev)
        o2 = torch.zeros(cnt * 16, dtype=torch.uint8, device=dev) if b2 else None
        op = torch.zeros(cnt * 24, dtype=torch.uint8, device=dev) if b2 else None
        if b2 is None:
            e.process_se_device(b1.c_struct(), o1.data_ptr())
        else:
            e.process_pe_device(b1.c_struct(), b2.c_struct(), o1.data_ptr(), o2.data_ptr(), op.data_ptr())
        e.sync()
        r = [o1.cpu().numpy().view(abi.READ_RESULT_DTYPE)]
        if b2 is not None:
            r += [o2.cpu().numpy().view(abi.READ_RESULT_DTYPE), op.cpu().numpy().view(abi.PAIR_RESULT_DTYPE)]
        recs.append(r)
        keep = (b1, b2)
    flat = e.finalize()
    dup = e.dup_table() if opt.duplicate.enabled else None
    layout = e.layout
    e.close()
    out = [np.concatenate([p[i] for p in recs]) for i in range(len(recs[0]))]
    return out, flat, dup, layout, opt, keep


@pytest.mark.parametrize("cfg,n", CASES)
def test_two_implementations_linearity_conservation(cfg, n):
    from fqtool_b200 import batch, workload
    from fqtool_b200.report import RunStats
    from tests import util
    swar = _run(cfg, n, 0)
    generic = _run(cfg, n, 1)
    util.assert_same((tuple(swar[0]), swar[1], swar[2]), (tuple(generic[0]), generic[1], generic[2]))
    halves = _run(cfg, n, 0, halves=True)
    util.assert_same((tuple(swar[0]), swar[1], swar[2]), (tuple(halves[0]), halves[1], halves[2]))
    # conservation
    recs, flat, _, L, opt, (b1, b2) = swar
    rs = RunStats(flat, L)
    paired = b2 is not None
    assert int(rs.filter.sum()) == n * (2 if paired else 1)
    Lr = workload.CONFIGS[cfg]["L"]
    pre1 = rs.stats["pre1"]
    assert pre1.reads == n and pre1.length_sum == n * Lr and int(pre1.total_base[:Lr].sum()) == n * Lr
    assert np.array_equal(pre1.total_base, pre1.content.sum(axis=0)) and np.array_equal(pre1.total_qual, pre1.qualsum.sum(axis=0))
    post1 = rs.stats["post1"]
    assert int(post1.total_base.sum()) == post1.length_sum and np.all(np.diff(post1.total_base.astype(np.int64)) <= 0)
    if not paired:
        passed = (recs[0]["result"] == 0) & ((recs[0]["flags"] & 1) == 0)
        assert post1.reads == int(passed.sum()) == int(rs.filter[0])
        assert post1.length_sum == int(recs[0]["len"][passed].astype(np.int64).sum())
    # oracle on a slice of the same device-generated stream
    from oracle import pyoracle
    lo, cnt = n // 3, 50_000 if not paired else 20_000
    h1 = b1.to_host(); r1 = batch.unpack(h1)
    r1s = batch.slice_reads(r1, lo - (n - b1.n), lo - (n - b1.n) + cnt) if b1.n != n else batch.slice_reads(r1, lo, lo + cnt)
    o = pyoracle.Oracle(opt.to_params(), "port")
    if paired:
        r2 = batch.unpack(b2.to_host())
        o1, o2, op = o.process_pe(r1s, batch.slice_reads(r2, lo, lo + cnt), lo)
        util.assert_records(recs[0][lo:lo + cnt], o1, "rec1 slice")
        util.assert_records(recs[1][lo:lo + cnt], o2, "rec2 slice")
        util.assert_records(recs[2][lo:lo + cnt], op, "pair slice")
    else:
        util.assert_records(recs[0][lo:lo + cnt], o.process_se(r1s, lo), "rec slice")
