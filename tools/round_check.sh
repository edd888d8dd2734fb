#!/bin/bash
# Round-end check on a GPU box: parity suite, smoke, bench (both arms), ncu launch list of the bench command.
# Usage (from the repo root): gpurun --timeout 1500 -- 'bash tools/round_check.sh TAG'
tag=${1:-rc}
o=gpurun_out
mkdir -p $o
timeout 900 python -m pytest tests -m gpu -x -q > $o/${tag}_tests.log 2>&1; echo "tests rc=$?" | tee -a $o/${tag}_tests.log
timeout 300 python __graft_entry__.py --smoke > $o/${tag}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $o/${tag}_smoke.log
for c in cfg2 cfg3 cfg4 cfg5; do timeout 200 python tools/prof_run.py $c 4000000 10 > $o/${tag}_prof_$c.log 2>&1; tail -2 $o/${tag}_prof_$c.log | head -1; done
timeout 600 python bench.py > $o/${tag}_bench.json 2> $o/${tag}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $o/${tag}_bench_ref.json 2> $o/${tag}_bench_ref.err; echo "bench ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $o/${tag}_ncu_bench.log 2>&1; echo "ncu rc=$?"
tail -3 $o/${tag}_tests.log; cat $o/${tag}_smoke.log | tail -2; cat $o/${tag}_bench.json $o/${tag}_bench_ref.json
