"""Tiny end-to-end runs of every kernel family for compute-sanitizer (memcheck).
    compute-sanitizer --tool memcheck python tools/sanitize_run.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fqtool_b200 import synth
from tests import configs, util

for name, impl in (("cfg2", 0), ("cfg2", 2), ("cfg4", 0), ("se_all_cuts", 0), ("se_polyg", 1)):
    util.run_cuda(configs.options(name, max_cycles=256), synth.ragged(3000, seed=1), chunks=2, impl=impl)
    util.run_cuda(configs.options(name, max_cycles=160), synth.se_hiseq(3000, 150, seed=2), impl=impl)
r1, r2 = synth.pe_novaseq(1500, 150, seed=3, frag_mean=200, frag_sd=80)
for name in ("cfg1", "cfg3", "pe_dup_kmer", "pe_adapters", "pe_merge_cut"):
    util.run_cuda(configs.options(name, max_cycles=320), r1, r2)
    util.run_cuda(configs.options(name, max_cycles=512), synth.ragged(1500, seed=4), synth.ragged(1500, seed=5), impl=0)
print("sanitize_run done")
