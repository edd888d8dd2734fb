#!/bin/bash
# NEEDS a tuning build: FQK_NVCC_EXTRA=-DFQK_TUNING_SWITCHES python -m fqtool_b200.build --force (the shipped libfqk.so ignores FQK_V4_* / FQK_V3_*)
# kernel timings of the fused paired-end kernel with role/part switches (FQK_V4_DEBUG bits: 1 no trim work, 2 no statistics
# work, 4 no first analyze, 8 no second analyze, 16 no tail corrections, 32 no polyG)
tag=${1:-pe}; cfgs=${2:-cfg3}; dbgs=${3:-0 2}
o=gpurun_out; mkdir -p $o
for c in $cfgs; do
  for d in $dbgs; do
    FQK_V4_DEBUG=$d timeout 60 python tools/prof_run.py $c 4000000 10 > $o/${tag}_prof_${c}_dbg$d.log 2>&1; echo "$c dbg=$d: $(grep 'kernel ms' $o/${tag}_prof_${c}_dbg$d.log)"
  done
done
