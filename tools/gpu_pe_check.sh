#!/bin/bash
# NEEDS a tuning build: FQK_NVCC_EXTRA=-DFQK_TUNING_SWITCHES python -m fqtool_b200.build --force (the shipped libfqk.so ignores FQK_V4_* / FQK_V3_*)
# GPU check of the fused paired-end kernel: parity tests, then kernel timings with role switches.
tag=${1:-pe}
o=gpurun_out; mkdir -p $o
timeout 600 python -m pytest tests/test_cuda_parity.py -m gpu -x -q -k "test_pe or 250" > $o/${tag}_tests.log 2>&1; echo "tests rc=$?" | tee -a $o/${tag}_tests.log
tail -n 5 $o/${tag}_tests.log
for c in cfg3 cfg5; do
  for d in 0 1 2; do
    FQK_V4_DEBUG=$d timeout 60 python tools/prof_run.py $c 4000000 10 > $o/${tag}_prof_${c}_dbg$d.log 2>&1; echo "$c dbg=$d: $(grep 'kernel ms' $o/${tag}_prof_${c}_dbg$d.log)"
  done
done
