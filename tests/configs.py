"""Named option sets used across the tests (CLI strings of the reference, main.cpp:14-122)."""
from fqtool_b200.options import from_cli

ADAPTER1 = "AGATCGGAAGAGCACACGTCTGAACTCCAGTCA"
ADAPTER2 = "AGATCGGAAGAGCGTCGTGTAGGGAAAGAGTGT"

SE = {
    # BASELINE.json configs[1] and configs[3]
    "cfg2": "-q --enable_cut_front --enable_cut_tail",
    "cfg4": f"-a --adapter_of_read1 {ADAPTER1} -x -d --dup_ana_key_len 12 --kmer --kmer_length 6",
    "se_plain": "",
    "se_q_only": "-q",
    "se_cut_right": "-q --enable_cut_front --enable_cut_right --cut_right_window 5 --cut_right_mean_qual 25",
    "se_force": "-f 3 -t 7 -b 90 -q -l --min_length 30 --max_length 85",
    "se_force_tail_only": "-t 5",
    "se_all_cuts": "-f 2 -t 1 --enable_cut_front --enable_cut_tail --cut_front_window 6 --cut_front_mean_qual 28 "
                   "--cut_tail_window 3 --cut_tail_mean_qual 15 -q -Q 15 -N 2 -e 25 -y -Y 0.4 -l",
    "se_polyg": "-g --min_len_detect_polyG 6 -x --base_to_trim AGN --min_len_detect_polyX 5 "
                "--max_mismatches_polyX 2 --one_mismatch_each_polyX 6 -q",
    "se_polyx1": "-x --min_len_detect_polyX 1 -g --min_len_detect_polyG 1",
    "se_short_adapter": "-a --adapter_of_read1 AGATCGGAAG -q --kmer --kmer_length 4",
    "se_kmer9_dup13": "-d --dup_ana_key_len 13 --kmer --kmer_length 9",
}
PE = {
    # BASELINE.json configs[0], [2], [4]
    "cfg1": "-q -a -g --enable_cut_tail",
    "cfg3": "-g -a -m",
    "cfg5": "-q -g -a",
    "pe_plain": "",
    "pe_merge_discard": "-m --discard_unmerged -q -l",
    "pe_adapters": f"-a --adapter_of_read1 {ADAPTER1} --adapter_of_read2 {ADAPTER2} -x -q --enable_cut_front "
                   "--enable_cut_tail -B 120 -b 130",
    "pe_dup_kmer": "-d --kmer --kmer_length 5 -q -F 2 -T 3 -f 1",
    "pe_overlap_params": "-a -m --min_overlap_len 20 --max_diff_for_overlap 2 -q",
    "pe_overlap_limit0": "-a -m --max_diff_for_overlap 0",
    "pe_merge_cut": "-m -a -g -x --enable_cut_front --enable_cut_right -q -l --min_length 40",
    # PE base correction (BaseCorrector, src/basecorrector.cpp)
    "pe_correct": "-c -a -q",
    "pe_correct_only": "-c --enable_cut_front",
    "pe_correct_merge": f"-c -m -a -g -q --adapter_of_read1 {ADAPTER1} -x",
}


def options(name, paired=None, max_cycles=512):
    cli = SE[name] if name in SE else PE[name]
    if paired is None:
        paired = name in PE
    argv = cli.split() + (["-I", "r2.fq"] if paired else [])
    o = from_cli(argv)
    o.maxCycles = max_cycles
    return o
