"""Generate tests/golden/*.npz from the REFERENCE build (oracle/_ref/libfqref.so, i.e. the
reference's own unmodified C++ compiled from /root/reference/src by oracle/Makefile).

    python tests/golden/make_golden.py

Run in the build container (where /root/reference exists); the .npz files are committed
so that the GPU box, which has no /root/reference, can check both the oracle port and the
CUDA path against the reference's outputs.  Each file holds the ASCII inputs, the option
CLI string, the per-read records and the run-level accumulators.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from fqtool_b200 import batch, synth          # noqa: E402
from oracle import pyoracle                   # noqa: E402
from tests import configs                     # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
TD = "/root/reference/testdata"


def micro_cases():
    """Hand-made edge cases of SURVEY.md section 8a/8c ([probed] items)."""
    I, H = "I", "#"
    seqs, quals = [], []
    def add(s, q):
        seqs.append(s); quals.append(q)
    add("ACGT" * 10, H * 4 + I * 36)                       # ####I... : cut_front starts at base 5
    add("ACGT" * 10, I * 36 + H * 4)                       # ...I#### : cut_tail keeps 35 of 40
    add("ACGT" * 10, H + I * 39)                           # bad first base for cut_right
    add("A" * 50, I * 27 + H * 23)                         # 23 low-quality bases in 50: fails -q
    add("A" * 50, I * 28 + H * 22)                         # 22: passes
    add("ACGTT" * 60, I * 277 + H * 23)                    # 23 of 300: fails (limit is 22 whatever the length)
    add("ACGTT" * 60, I * 278 + H * 22)
    add("ACGTACGTACGTACGTACGT" + "G" * 12, I * 32)          # 12 G tail
    add("ACGTACGTACGTACGTACGT" + "G" * 8 + "GT", I * 30)    # 8 G + GT: scanned length counts
    add("ACGTACGTACGTACGTACGT" + "GGGGTGGGGGG", I * 31)     # G tail with one mismatch
    add("NNNNACGTACGTNNNN" * 3, I * 48)                    # N skipping at both ends
    add("N" * 40, I * 40)                                  # all N
    add("", "")                                            # empty read
    add("ACG", "III")                                      # shorter than any window
    add("AGATCGGAAGAGCACACGTCTGAACTCCAGTCA"[4:] + "ACGT" * 5, I * 49)   # adapter at pos -4
    add("ACGTTGCA" * 4 + "AGATCGGAAGAGCACACGTCTGAACTCCAGTCA", I * 65)   # adapter inside
    add("ACGTTGCA" * 4 + "AGATCGGTAGAGCACACGTCTGAACTCCAGTCA"[:20], I * 52)   # adapter with a mismatch, truncated
    add("ACGTTGCATT" * 3 + "A" * 25, I * 55)               # polyA tail
    add("ACGTTGCATT" * 3 + "N" * 12, I * 30 + H * 12)      # polyN tail
    return batch.from_strings(seqs, quals)


def save(name, cfg, r1, r2=None, kind="reference"):
    opt = configs.options(cfg)
    o = pyoracle.Oracle(opt.to_params(), kind)
    d = {"cfg": np.array(cfg), "seq1": r1[0], "qual1": r1[1], "off1": r1[2]}
    if r2 is None:
        d["rec1"] = o.process_se(r1)
    else:
        d["seq2"], d["qual2"], d["off2"] = r2
        d["rec1"], d["rec2"], d["recp"] = o.process_pe(r1, r2)
        if opt.correction.enabled:
            d["corr"] = o.corrections()
    d["flat"] = o.finalize()
    if opt.duplicate.enabled:
        dups, counts, gc = o.dup_table()
        nz = np.nonzero(counts)[0]
        d["dup_keys"], d["dup_kmer"], d["dup_counts"], d["dup_gc"] = nz.astype(np.uint32), dups[nz], counts[nz], gc[nz]
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
    print(name, {k: (v.shape if hasattr(v, "shape") else v) for k, v in d.items() if k.startswith("rec")})


def main():
    if not pyoracle.have_reference():
        raise SystemExit("oracle/_ref/libfqref.so missing: run `make -C oracle` where /root/reference exists")
    micro = micro_cases()
    for cfg in ("cfg2", "cfg4", "se_cut_right", "se_all_cuts", "se_polyg", "se_polyx1", "se_force"):
        save(f"micro_{cfg}", cfg, micro)
    rag = synth.ragged(600, seed=101)
    for cfg in ("cfg2", "cfg4", "se_all_cuts", "se_polyg", "se_kmer9_dup13"):
        save(f"ragged_{cfg}", cfg, rag)
    save("hiseq_cfg2", "cfg2", synth.se_hiseq(800, 150, seed=102))
    save("adapter_cfg4", "cfg4", synth.se_adapter_polyx(800, 100, seed=103))
    p1, p2 = synth.pe_novaseq(500, 150, seed=104, frag_mean=200, frag_sd=80)
    for cfg in ("cfg1", "cfg3", "cfg5", "pe_adapters", "pe_dup_kmer", "pe_overlap_limit0", "pe_merge_cut", "pe_correct", "pe_correct_merge"):
        save(f"novaseq_{cfg}", cfg, p1, p2)
    a, b = synth.ragged(500, seed=105), synth.ragged(500, seed=106)
    for cfg in ("cfg1", "cfg3", "pe_merge_discard", "pe_merge_cut"):
        save(f"raggedpe_{cfg}", cfg, a, b)
    a = synth.ragged(500, seed=107, lmin=20, lmax=200, n_rate=0.01, lowq_rate=0.05)
    for cfg in ("cfg3", "pe_overlap_params"):
        save(f"revcomp_{cfg}", cfg, a, synth.revcomp_reads(a))
    # the reference's own testdata (first 1500 pairs) and its NovaSeq polyG pair
    if os.path.exists(TD):
        r1 = batch.slice_reads(batch.read_fastq(f"{TD}/r1.fq.gz"), 0, 1500)
        r2 = batch.slice_reads(batch.read_fastq(f"{TD}/r2.fq.gz"), 0, 1500)
        save("testdata_cfg1", "cfg1", r1, r2)
        save("testdata_cfg3", "cfg3", r1, r2)
        save("testdata_pe_correct_merge", "pe_correct_merge", r1, r2)
        save("testdata_cfg2", "cfg2", r1)
        save("testdata_cfg4", "cfg4", r1)
        g1, g2 = batch.read_fastq(f"{TD}/polygr1.fq"), batch.read_fastq(f"{TD}/polygr2.fq")
        save("polyg_pair_cfg3", "cfg3", g1, g2)


if __name__ == "__main__":
    main()
