#!/bin/bash
# NEEDS a tuning build: FQK_NVCC_EXTRA=-DFQK_TUNING_SWITCHES python -m fqtool_b200.build --force (the shipped libfqk.so ignores FQK_V4_* / FQK_V3_*)
# like gpu_pe_time.sh with tile size / stage count overrides: gpu_pe_time2.sh <tag> <cfg> "<T:NS:dbg> ..."
tag=${1:-pe}; c=${2:-cfg3}; shift 2
o=gpurun_out; mkdir -p $o
for spec in $*; do
  IFS=: read T NS d <<< "$spec"
  FQK_V4_T=$T FQK_V4_NS=$NS FQK_V4_DEBUG=$d timeout 60 python tools/prof_run.py $c 4000000 10 > $o/${tag}_prof_${c}_T${T}_NS${NS}_dbg$d.log 2>&1
  echo "$c T=$T NS=$NS dbg=$d: $(grep 'kernel ms' $o/${tag}_prof_${c}_T${T}_NS${NS}_dbg$d.log) $(grep v4prof $o/${tag}_prof_${c}_T${T}_NS${NS}_dbg$d.log | tail -1)"
done
