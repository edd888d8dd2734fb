#!/bin/bash
# NEEDS a tuning build: FQK_NVCC_EXTRA=-DFQK_TUNING_SWITCHES python -m fqtool_b200.build --force (the shipped libfqk.so ignores FQK_V4_* / FQK_V3_*)
# per-kernel times of the split paired-end path (ncu launch list) and tile-shape experiments for the trim kernel
o=gpurun_out; mkdir -p $o
for c in ${1:-cfg3 cfg5}; do
  timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_pe_v4 --csv --log-file $o/split_$c.csv python tools/prof_run.py $c 4000000 3 > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("$o/split_$c.csv")) if len(r)>5 and r[0].isdigit()]
for r in rows[-2:]: print("$c", r[4][60:110], r[-1], r[-2])
PY
  for cfgs in "32 12" "64 8" "64 6" "64 4" "96 5" "128 4" "128 3"; do set -- $cfgs
    echo "$c T=$1 NS=$2: $(FQK_V4_TMODE=1 FQK_V4_T=$1 FQK_V4_NS=$2 timeout 60 python tools/prof_run.py $c 4000000 5 2>&1 | grep 'kernel ms')"
  done
  echo "$c counter hand-out: $(FQK_V4_DEBUG=512 timeout 60 python tools/prof_run.py $c 4000000 5 2>&1 | grep 'kernel ms')"
  echo "$c no stats work:   $(FQK_V4_DEBUG=2 timeout 60 python tools/prof_run.py $c 4000000 5 2>&1 | grep 'kernel ms')"
  echo "$c no trim work:    $(FQK_V4_DEBUG=1 timeout 60 python tools/prof_run.py $c 4000000 5 2>&1 | grep 'kernel ms')"
done
