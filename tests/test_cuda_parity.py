"""GPU parity: the CUDA path through the C ABI (host buffers) against the CPU oracle on the
same seeded inputs.  Bit-exact records, accumulators and duplication tables."""
import numpy as np
import pytest

from fqtool_b200 import synth
from tests import configs, util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("impl", [0, 1, 3], ids=["auto", "generic", "fused_ws"])
@pytest.mark.parametrize("name", list(configs.SE))
def test_se(name, impl):
    opt = configs.options(name)
    for reads in (synth.se_hiseq(6000, 150, seed=31), synth.se_adapter_polyx(6000, 100, seed=32),
                  synth.ragged(6000, seed=33)):
        util.assert_same(util.run_cuda(opt, reads, chunks=3, impl=impl), util.run_oracle(opt, reads))


@pytest.mark.parametrize("impl", [0, 1, 3], ids=["auto", "generic", "two_phase"])
@pytest.mark.parametrize("name", list(configs.PE))
def test_pe(name, impl):
    opt = configs.options(name)
    r1, r2 = synth.pe_novaseq(4000, 150, seed=41, frag_mean=220, frag_sd=90)
    util.assert_same(util.run_cuda(opt, r1, r2, chunks=2, impl=impl), util.run_oracle(opt, r1, r2))
    a, b = synth.ragged(4000, seed=42), synth.ragged(4000, seed=43)
    util.assert_same(util.run_cuda(opt, a, b, impl=impl), util.run_oracle(opt, a, b))
    a = synth.ragged(3000, seed=44, lmin=20, lmax=200, n_rate=0.01, lowq_rate=0.05)
    b = synth.revcomp_reads(a)
    util.assert_same(util.run_cuda(opt, a, b, impl=impl), util.run_oracle(opt, a, b))


@pytest.mark.parametrize("impl", [0, 3], ids=["auto", "two_phase"])
def test_long_reads_250(impl):
    opt = configs.options("cfg5")
    r1, r2 = synth.pe_novaseq(2000, 250, seed=45, frag_mean=420, frag_sd=120)
    util.assert_same(util.run_cuda(opt, r1, r2, impl=impl), util.run_oracle(opt, r1, r2))
    opt = configs.options("cfg3", max_cycles=1024)
    util.assert_same(util.run_cuda(opt, r1, r2, impl=impl), util.run_oracle(opt, r1, r2))


def test_pe_many_tiles_and_blocks():
    """Enough pairs for several tiles per block of the fused paired-end kernel on every SM, ragged tails included."""
    for name in ("cfg3", "cfg1", "pe_merge_cut"):
        opt = configs.options(name)
        r1, r2 = synth.pe_novaseq(60000, 150, seed=51, frag_mean=260, frag_sd=100)
        util.assert_same(util.run_cuda(opt, r1, r2), util.run_oracle(opt, r1, r2))
    # 250 bp: two chunks per tile, so a trim warp runs up to four tiles ahead of the slowest one (tile-tag handshake)
    opt = configs.options("cfg5", max_cycles=256)
    r1, r2 = synth.pe_novaseq(40000, 250, seed=52, frag_mean=420, frag_sd=120)
    util.assert_same(util.run_cuda(opt, r1, r2), util.run_oracle(opt, r1, r2))


def test_empty_and_tiny_batches():
    opt = configs.options("cfg2")
    from fqtool_b200 import batch
    empty = batch.from_strings([], [])
    util.assert_same(util.run_cuda(opt, empty), util.run_oracle(opt, empty))
    one = batch.from_strings(["ACGTN"], ["IIII#"])
    util.assert_same(util.run_cuda(opt, one), util.run_oracle(opt, one))


def test_read_longer_than_max_cycles_is_an_error():
    from fqtool_b200.engine import FqkError
    opt = configs.options("cfg2", max_cycles=64)
    with pytest.raises(FqkError):
        util.run_cuda(opt, synth.se_hiseq(10, 100, seed=1))


def test_pipelined_submit_wait_matches_sync():
    """fqk_submit_se / fqk_wait (two packs in flight) give the same records and accumulators as fqk_process_se."""
    from fqtool_b200 import abi, batch
    from fqtool_b200.engine import Engine
    opt = configs.options("cfg2", max_cycles=256)
    reads = synth.ragged(9000, seed=77)
    want = util.run_cuda(opt, reads, chunks=6)
    e = Engine(opt.to_params())
    bounds = np.linspace(0, 9000, 7).astype(int)
    packs = [batch.pack(batch.slice_reads(reads, lo, hi), int(lo)) for lo, hi in zip(bounds[:-1], bounds[1:])]
    outs = [np.zeros(p.n, dtype=abi.READ_RESULT_DTYPE) for p in packs]
    pending = None
    for p, o in zip(packs, outs):
        t = e.submit_se(p, o)
        if pending is not None:
            e.wait(pending)
        pending = t
    e.wait(pending)
    got = (np.concatenate(outs),), e.finalize(), None
    util.assert_same(got, want)
    e.close()


@pytest.mark.parametrize("name", ["cfg3", "pe_dup_kmer", "cfg5"])
def test_pipelined_submit_pe_matches_sync(name):
    """fqk_submit_pe / fqk_wait (two pair packs in flight) give the same records, accumulators and duplication table as
    fqk_process_pe, i.e. as the oracle; with -c the pipelined form is refused."""
    from fqtool_b200 import abi, batch
    from fqtool_b200.engine import Engine, FqkError
    opt = configs.options(name, max_cycles=320)
    r1, r2 = synth.pe_novaseq(9001, 150, seed=31, frag_mean=230, frag_sd=80)
    want = util.run_oracle(opt, r1, r2)
    e = Engine(opt.to_params())
    bounds = np.linspace(0, 9001, 8).astype(int)
    packs = [(batch.pack(batch.slice_reads(r1, lo, hi), int(lo)), batch.pack(batch.slice_reads(r2, lo, hi), int(lo)))
             for lo, hi in zip(bounds[:-1], bounds[1:])]
    outs = [(np.zeros(p[0].n, dtype=abi.READ_RESULT_DTYPE), np.zeros(p[0].n, dtype=abi.READ_RESULT_DTYPE),
             np.zeros(p[0].n, dtype=abi.PAIR_RESULT_DTYPE)) for p in packs]
    pending = None
    for (p1, p2), o in zip(packs, outs):
        t = e.submit_pe(p1, p2, o)
        if pending is not None:
            e.wait(pending)
        pending = t
    e.wait(pending)
    recs = tuple(np.concatenate([o[k] for o in outs]) for k in range(3))
    got = (recs, e.finalize(), e.dup_table() if opt.duplicate.enabled else None)
    util.assert_same(got, want[:3])
    e.close()
    e = Engine(configs.options("pe_correct_merge").to_params())
    with pytest.raises(FqkError):
        e.submit_pe(packs[0][0], packs[0][1], outs[0])
    e.close()


@pytest.mark.parametrize("name", ["cfg2", "cfg4", "cfg3", "pe_correct_merge"])
def test_ascii_entry_points_match_packed_path(name):
    """fqk_process_*_ascii (device-side packing) against the ORACLE (and hence the host-packed path); foreign alphabet is
    an error."""
    from fqtool_b200 import batch
    from fqtool_b200.engine import Engine, FqkError
    opt = configs.options(name)
    paired = name in configs.PE
    r1 = synth.ragged(5000, seed=91); r2 = synth.ragged(5000, seed=92) if paired else None
    want = util.run_oracle(opt, r1, r2)
    e = Engine(opt.to_params())
    recs = e.process_pe_ascii(r1, r2) if paired else (e.process_se_ascii(r1),)
    got = (tuple(recs), e.finalize(), e.dup_table() if opt.duplicate.enabled else None)
    util.assert_same(got, want[:3])
    if not paired:
        bad = batch.from_strings(["ACGTNACGTRACGT"], ["IIIIIIIIIIIIII"])
        with pytest.raises(FqkError):
            e.process_se_ascii(bad)
    e.close()


@pytest.mark.parametrize("name", ["cfg2", "cfg4", "se_all_cuts"])
def test_pipelined_ascii_submit_matches_oracle(name):
    """fqk_submit_se_ascii / fqk_wait: two packs in flight, per-read offsets / lengths / packed layout derived on the device
    (k_ascii_index + prefix sum); records, accumulators and the duplication table against the oracle."""
    from fqtool_b200 import abi, batch
    from fqtool_b200.engine import Engine, FqkError
    opt = configs.options(name)
    reads = synth.ragged(9000, seed=78)
    want = util.run_oracle(opt, reads)
    e = Engine(opt.to_params())
    bounds = np.linspace(0, 9000, 8).astype(int)
    packs = [batch.slice_reads(reads, lo, hi) for lo, hi in zip(bounds[:-1], bounds[1:])]
    outs = [np.zeros(len(p[2]) - 1, dtype=abi.READ_RESULT_DTYPE) for p in packs]
    pending, keep = None, []
    for p, o, lo in zip(packs, outs, bounds[:-1]):
        t, k = e.submit_se_ascii(p, o, int(lo))
        keep.append(k)
        if pending is not None:
            e.wait(pending)
        pending = t
    e.wait(pending)
    got = ((np.concatenate(outs),), e.finalize(), e.dup_table() if opt.duplicate.enabled else None)
    util.assert_same(got, want[:3])
    e.close()
    # an over-long read is reported by the wait (the index is built on the device)
    e = Engine(configs.options("cfg2", max_cycles=64).to_params())
    o = np.zeros(10, dtype=abi.READ_RESULT_DTYPE)
    t, k = e.submit_se_ascii(synth.se_hiseq(10, 100, seed=1), o)
    with pytest.raises(FqkError):
        e.wait(t)
    e.close()
