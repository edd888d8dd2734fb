#!/bin/bash
# host-side string path (fqk::HotPath: pack + fqk_process_* + slice + concat) timed on a larger FASTQ: the test data x200
o=gpurun_out; mkdir -p $o /tmp/hp
(for i in $(seq 200); do zcat tests/golden/fq/r1.fq.gz; done) > /tmp/hp/r1.fq
(for i in $(seq 200); do zcat tests/golden/fq/r2.fq.gz; done) > /tmp/hp/r2.fq
for pack in 100000 400000; do
echo "SE cfg2 pack=$pack: $(fqtool_b200/host/fqk_run -i /tmp/hp/r1.fq -o /tmp/hp/o1.fq -q --enable_cut_front --enable_cut_tail --max_item_in_pack $pack --max_cycles 160)"
echo "PE cfg3 pack=$pack: $(fqtool_b200/host/fqk_run -i /tmp/hp/r1.fq -I /tmp/hp/r2.fq -o /tmp/hp/o1.fq -O /tmp/hp/o2.fq --merge_output /tmp/hp/m.fq -g -a -m --max_item_in_pack $pack --max_cycles 320)"
done
