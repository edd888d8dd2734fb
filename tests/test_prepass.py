"""Adapter auto-detection pre-pass (Evaluator::evaluateAdapterSeq src/evaluator.cpp:229-390,
getAdapterWithSeed :392-425).  CPU: the oracle port against the loop run with the reference's own Evaluator::seq2int,
and the host logic (key ranking, NucleotideTree walk) against the reference's getAdapterWithSeed.  GPU: the two device
calls of include/fqk.h against the oracle, and the whole detection."""
import numpy as np
import pytest

from fqtool_b200 import prepass, synth
from oracle import pyoracle

ADAPTER1 = "AGATCGGAAGAGCACACGTCTGAACTCCAGTCA"


def _reads(n, L, seed):
    return synth.se_adapter_polyx(n, L, seed=seed)          # 12 % adapter read-through, polyX tails, N bases


def _o_counts(reads, k, p, s):
    return pyoracle.adapter_kmer_counts(reads, k, p, s, "port")


def _o_hits(reads, seed, k, p, s):
    return pyoracle.adapter_seed_hits(reads, seed, k, p, s)


@pytest.mark.parametrize("k,first,shift", [(10, 20, 1), (10, 20, 3), (6, 0, 0), (12, 5, 2)])
def test_oracle_counts_match_reference_seq2int(k, first, shift):
    if not pyoracle.have_reference():
        pytest.skip("reference build absent")
    reads = _reads(3000, 100, 5)
    a = pyoracle.adapter_kmer_counts(reads, k, first, shift, "port")
    b = pyoracle.adapter_kmer_counts(reads, k, first, shift, "reference")
    assert a.sum() > 0 and np.array_equal(a, b)
    rag = synth.ragged(2000, seed=9)                          # lengths 0..: windows that do not exist
    assert np.array_equal(pyoracle.adapter_kmer_counts(rag, k, first, shift, "port"),
                          pyoracle.adapter_kmer_counts(rag, k, first, shift, "reference"))


def test_seed_hits_agree_with_counts():
    reads = _reads(4000, 100, 6)
    counts = _o_counts(reads, 10, 20, 1)
    for seed in np.argsort(counts)[-3:]:
        ri, po = _o_hits(reads, int(seed), 10, 20, 1)
        assert len(ri) == counts[seed]
        seq, _, off = reads
        for r, p in list(zip(ri, po))[:20]:
            assert bytes(seq[int(off[r]) + p:int(off[r]) + p + 10]).decode() == prepass.int2seq(int(seed), 10)


def test_host_logic_matches_reference_getAdapterWithSeed():
    if not pyoracle.have_reference():
        pytest.skip("reference build absent")
    reads = _reads(12000, 100, 7)
    counts = _o_counts(reads, 10, 20, 1)
    top, total, c = prepass.rank_keys(counts)
    assert top[0] != 0 and c[top[0]] >= c[top[1]] >= c[top[2]]
    for seed in top[:4]:
        want = pyoracle.reference_adapter_with_seed(reads, seed, 10, 0)
        # the reference maps a grown adapter onto its table of known adapters (src/knownadapters.h, the reference's data, not
        # reproduced here); the synthetic reads carry TruSeq read 1, which is in that table
        got = prepass.adapter_with_seed(reads, seed, 0, hits_fn=_o_hits, known=(ADAPTER1,))
        assert got == want, (prepass.int2seq(seed, 10), got, want)
    assert prepass.detect_adapter(reads, counts_fn=_o_counts, hits_fn=_o_hits, known=(ADAPTER1,)) == ADAPTER1
    # random reads: no key stands out, the reference finds nothing either
    rnd = synth.se_hiseq(10000, 100, seed=3)
    assert prepass.detect_adapter(rnd, counts_fn=_o_counts, hits_fn=_o_hits) == ""


def test_rank_keys_tie_break_and_filters():
    counts = np.zeros(1 << 20, dtype=np.uint64)
    ka = int("".join(str("ATCG".index(ch)) for ch in "ATCGATCGAT"), 4)
    kb = int("".join(str("ATCG".index(ch)) for ch in "TACGTACGTA"), 4)
    kg = int("".join(str("ATCG".index(ch)) for ch in "GGGGATCATC"), 4)      # starts with GGGG: skipped
    kl = int("".join(str("ATCG".index(ch)) for ch in "AAAAAATCGT"), 4)      # six A: low complexity
    for k_ in (ka, kb, kg, kl):
        counts[k_] = 100
    top, total, _ = prepass.rank_keys(counts)
    assert top[:2] == sorted([ka, kb], reverse=True) and total == 200       # equal counts: the later key ranks first


@pytest.mark.gpu
@pytest.mark.parametrize("k,first,shift", [(10, 20, 1), (6, 0, 0), (12, 5, 2)])
def test_device_counts_and_hits_match_oracle(k, first, shift):
    for reads in (_reads(20000, 100, 8), synth.ragged(5000, seed=10)):
        want = _o_counts(reads, k, first, shift)
        got = prepass.kmer_counts(reads, k, first, shift)
        assert np.array_equal(got, want)
        seed = int(np.argmax(want))
        ri, po = prepass.seed_hits(reads, seed, k, first, shift)
        wr, wp = _o_hits(reads, seed, k, first, shift)
        assert np.array_equal(ri, wr) and np.array_equal(po, wp)


@pytest.mark.gpu
def test_device_detection_finds_the_adapter():
    reads = _reads(30000, 100, 11)
    got = prepass.detect_adapter(reads, known=(ADAPTER1,))
    assert got == prepass.detect_adapter(reads, counts_fn=_o_counts, hits_fn=_o_hits, known=(ADAPTER1,)) == ADAPTER1
    assert prepass.detect_adapter(_reads(9999, 100, 12)) == ""               # fewer than 10000 records: no detection
