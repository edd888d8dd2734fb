"""Shared helpers: run a pack of reads through the oracle or the CUDA path and compare."""
import glob
import os

import numpy as np

from fqtool_b200 import batch
from tests import configs

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_files():
    return sorted(glob.glob(os.path.join(GOLDEN, "*.npz")))


def load_golden(path):
    d = np.load(path)
    g = {k: d[k] for k in d.files}
    g["cfg"] = str(g["cfg"])
    g["r1"] = (g["seq1"], g["qual1"], g["off1"])
    g["r2"] = (g["seq2"], g["qual2"], g["off2"]) if "seq2" in g else None
    return g


def run_oracle(opt, r1, r2=None, kind="port", chunks=1):
    from oracle import pyoracle
    o = pyoracle.Oracle(opt.to_params(), kind)
    o._corr = []

    def pe(a, b, i):
        r = o.process_pe(a, b, i)
        if opt.correction.enabled:
            o._corr.append((i, o.corrections()))
        return r
    out = _run_chunks(lambda a, i: (o.process_se(a, i),), pe, r1, r2, chunks)
    flat = o.finalize()
    dup = o.dup_table() if opt.duplicate.enabled else None
    o.close()
    if opt.correction.enabled:
        return out, flat, dup, _merge_corr(o._corr)
    return out, flat, dup


def run_cuda(opt, r1, r2=None, chunks=1, impl=0, device=0):
    """Through the C ABI with host buffers (fqk_process_se / fqk_process_pe)."""
    from fqtool_b200.engine import Engine
    e = Engine(opt.to_params(), device=device)
    e.set_impl(impl)
    corr = []

    def pe(a, b, i):
        r = e.process_pe(batch.pack(a, i), batch.pack(b, i))
        if opt.correction.enabled:
            corr.append((i, e.corrections()))
        return r
    out = _run_chunks(lambda a, i: (e.process_se(batch.pack(a, i)),), pe, r1, r2, chunks)
    flat = e.finalize()
    dup = e.dup_table() if opt.duplicate.enabled else None
    e.close()
    if opt.correction.enabled:
        return out, flat, dup, _merge_corr(corr)
    return out, flat, dup


def _merge_corr(parts):
    """Per-chunk edit lists -> one list with global pair indices, sorted."""
    from fqtool_b200 import abi
    if not parts:
        return np.zeros(0, dtype=abi.CORRECTION_DTYPE)
    out = []
    for first, c in parts:
        c = c.copy(); c["pair"] += np.uint32(first); out.append(c)
    return np.sort(np.concatenate(out), order=["pair", "mate_base", "pos"])


def _run_chunks(se, pe, r1, r2, chunks):
    n = len(r1[2]) - 1
    bounds = np.linspace(0, n, chunks + 1).astype(int)
    recs = []
    for lo, hi in zip(bounds[:-1], bounds[1:]):
        if r2 is None:
            recs.append(se(batch.slice_reads(r1, lo, hi), int(lo)))
        else:
            recs.append(pe(batch.slice_reads(r1, lo, hi), batch.slice_reads(r2, lo, hi), int(lo)))
    return tuple(np.concatenate([r[i] for r in recs]) for i in range(len(recs[0])))


def assert_records(x, y, what=""):
    assert len(x) == len(y)
    for name in x.dtype.names:
        if name == "reserved":
            continue
        bad = np.nonzero(x[name] != y[name])[0]
        assert bad.size == 0, f"{what} field {name}: {bad.size} mismatches, first at read {bad[0]}: {x[bad[0]]} vs {y[bad[0]]}"


def assert_same(a, b):
    """(records, flat, dup) bit-exact."""
    (ra, fa, da), (rb, fb, db) = a[:3], b[:3]
    if len(a) > 3 or len(b) > 3:
        assert len(a[3]) == len(b[3]) and np.array_equal(a[3], b[3]), f"base corrections differ: {len(a[3])} vs {len(b[3])}"
    for i, (x, y) in enumerate(zip(ra, rb)):
        assert_records(x, y, f"records[{i}]")
    bad = np.nonzero(fa != fb)[0]
    assert bad.size == 0, f"flat accumulators differ at {bad[:8]}: {fa[bad[:8]]} vs {fb[bad[:8]]}"
    if da is not None:
        for x, y, nm in zip(da, db, ("dups", "counts", "gc")):
            bad = np.nonzero(x != y)[0]
            assert bad.size == 0, f"dup {nm} differ at keys {bad[:8]}: {x[bad[:8]]} vs {y[bad[:8]]}"


def assert_matches_golden(res, g):
    recs, flat, dup = res[:3]
    if "corr" in g:
        assert np.array_equal(res[3], g["corr"]), "base corrections differ from the reference"
    assert_records(recs[0], g["rec1"], "rec1")
    if g["r2"] is not None:
        assert_records(recs[1], g["rec2"], "rec2")
        assert_records(recs[2], g["recp"], "recp")
    bad = np.nonzero(flat != g["flat"])[0]
    assert bad.size == 0, f"flat accumulators differ at {bad[:8]}: {flat[bad[:8]]} vs {g['flat'][bad[:8]]}"
    if "dup_keys" in g:
        dups, counts, gc = dup
        nz = np.nonzero(counts)[0]
        assert np.array_equal(nz.astype(np.uint32), g["dup_keys"])
        assert np.array_equal(dups[nz], g["dup_kmer"]) and np.array_equal(counts[nz], g["dup_counts"])
        assert np.array_equal(gc[nz], g["dup_gc"])
