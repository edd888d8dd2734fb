"""Pins the plain-C oracle (oracle/fq_oracle.c) against the reference's own unmodified
C++ (oracle/_ref/libfqref.so, built from /root/reference/src by oracle/Makefile):
identical per-read records, identical raw Stats / FilterResult / insert-size arrays,
identical duplication tables.  CPU only.  Skipped where the reference build is absent."""
import os

import numpy as np
import pytest

from fqtool_b200 import batch, synth
from oracle import pyoracle
from tests import configs, util

REF_TESTDATA = "/root/reference/testdata"
needs_ref = pytest.mark.skipif(not pyoracle.have_reference(), reason="oracle/_ref not built (no /root/reference)")


def run(kind, opt, r1, r2=None, chunks=1):
    return util.run_oracle(opt, r1, r2, kind=kind, chunks=chunks)


assert_same = util.assert_same


@needs_ref
@pytest.mark.parametrize("name", list(configs.SE))
def test_se_synthetic(name):
    opt = configs.options(name)
    for reads in (synth.se_hiseq(3000, 150, seed=11), synth.se_adapter_polyx(3000, 100, seed=12),
                  synth.ragged(3000, seed=13)):
        assert_same(run("port", opt, reads, chunks=3), run("reference", opt, reads))


@needs_ref
@pytest.mark.parametrize("name", list(configs.PE))
def test_pe_synthetic(name):
    opt = configs.options(name)
    r1, r2 = synth.pe_novaseq(2000, 150, seed=21, frag_mean=220, frag_sd=90)
    assert_same(run("port", opt, r1, r2, chunks=2), run("reference", opt, r1, r2))
    a, b = synth.ragged(2000, seed=22), synth.ragged(2000, seed=23)
    assert_same(run("port", opt, a, b), run("reference", opt, a, b))
    # pairs that really overlap, ragged lengths: read2 = revcomp of a window of read1's fragment
    a = synth.ragged(1500, seed=24, lmin=20, lmax=200, n_rate=0.01, lowq_rate=0.05)
    assert_same(run("port", opt, a, synth.revcomp_reads(a)), run("reference", opt, a, synth.revcomp_reads(a)))


@needs_ref
@pytest.mark.skipif(not os.path.exists(REF_TESTDATA), reason="reference testdata absent")
@pytest.mark.parametrize("name", ["cfg1", "cfg3", "cfg5", "pe_adapters", "pe_dup_kmer"])
def test_pe_testdata(name):
    opt = configs.options(name)
    r1 = batch.read_fastq(f"{REF_TESTDATA}/r1.fq.gz"); r2 = batch.read_fastq(f"{REF_TESTDATA}/r2.fq.gz")
    assert_same(run("port", opt, r1, r2, chunks=4), run("reference", opt, r1, r2))


@needs_ref
@pytest.mark.skipif(not os.path.exists(REF_TESTDATA), reason="reference testdata absent")
@pytest.mark.parametrize("name", ["cfg2", "cfg4", "se_all_cuts", "se_polyg"])
def test_se_testdata(name):
    opt = configs.options(name)
    r1 = batch.read_fastq(f"{REF_TESTDATA}/r1.fq.gz")
    assert_same(run("port", opt, r1, chunks=4), run("reference", opt, r1))
