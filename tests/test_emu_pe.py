"""CPU check of the fused paired-end kernel's device logic: the trim role of k_pe_v4 (the source nvcc compiles,
fqtool_b200/csrc/fqk_pe4.cuh) runs one emulated warp at a time on the CPU and its records, exception lists and tail
corrections, consumed through the statistics role's contract, must reproduce the oracle bit for bit."""
import pytest

from fqtool_b200 import synth
from tests import configs, util
from tests.emu import emu

FUSED_PE = [n for n in configs.PE if not n.startswith("pe_correct") and n != "pe_dup_kmer"]


@pytest.mark.parametrize("name", FUSED_PE)
def test_emulated_pe_kernel_matches_oracle(name):
    opt = configs.options(name)
    r1, r2 = synth.pe_novaseq(700, 150, seed=41, frag_mean=220, frag_sd=90)
    util.assert_same(emu.run_pe(opt, r1, r2, chunks=2), util.run_oracle(opt, r1, r2))
    a, b = synth.ragged(600, seed=42), synth.ragged(600, seed=43)
    util.assert_same(emu.run_pe(opt, a, b), util.run_oracle(opt, a, b))
    a = synth.ragged(500, seed=44, lmin=20, lmax=200, n_rate=0.01, lowq_rate=0.05)
    b = synth.revcomp_reads(a)
    util.assert_same(emu.run_pe(opt, a, b, tile=32), util.run_oracle(opt, a, b))


def test_emulated_pe_kernel_250():
    r1, r2 = synth.pe_novaseq(300, 250, seed=45, frag_mean=420, frag_sd=120)
    for name, mc in (("cfg5", 512), ("cfg3", 1024)):
        opt = configs.options(name, max_cycles=mc)
        util.assert_same(emu.run_pe(opt, r1, r2), util.run_oracle(opt, r1, r2))
