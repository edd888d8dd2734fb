#!/bin/bash
# Evidence run for the roofline fields of bench.py:
#   1. ncu launch list of the bench command (kernel SHARES of the step),
#   2. one --set full capture per configuration (4 M units, the hot kernels only) -> dram bytes per launch for
#      profiles/traffic.json (python tools/ncu_summary.py full, NCU_CFG=<cfg>, run where ncu can read the reports).
# Usage (repo root): gpurun --timeout 1500 -- 'bash tools/gpu_traffic.sh TAG'
tag=${1:-traffic}
o=gpurun_out
mkdir -p $o
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extra > $o/${tag}_ncu_bench.log 2>&1; echo "launch list rc=$?"
for c in cfg2 cfg3 cfg4 cfg5; do
    timeout 400 ncu --set full --clock-control none -k regex:'k_se_v3|k_pe_v4|k_kmer_se|k_dup$' -c 3 -f -o $o/${tag}_$c \
        python tools/prof_run.py $c 4000000 1 > $o/${tag}_ncu_$c.log 2>&1; echo "$c full capture rc=$?"
    # gpurun brings back at most 64 MiB: keep the raw metric table, drop the report
    ncu -i $o/${tag}_$c.ncu-rep --page raw --csv > $o/${tag}_${c}_raw.csv 2>> $o/${tag}_ncu_$c.log && rm -f $o/${tag}_$c.ncu-rep
done
ls -la $o/${tag}_*
