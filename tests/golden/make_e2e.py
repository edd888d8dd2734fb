"""End-to-end goldens from the WHOLE reference program (oracle/_ref/fqtool_ref -w 1).

    python tests/golden/make_e2e.py

Writes tests/golden/fq/{r1,r2}.fq.gz (the first 2000 pairs of the reference's testdata, names
included) and tests/golden/e2e_expected.json with the md5 of every output file the reference
binary produces for a set of command lines.  The GPU test (tests/test_e2e_host.py) runs the
same command lines through fqtool_b200/host/fqk_run (C++ host adapter + CUDA hot path) and
requires byte-identical files.
"""
import gzip
import hashlib
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
OUT = os.path.dirname(os.path.abspath(__file__))
TD = "/root/reference/testdata"
REF = os.path.join(ROOT, "oracle", "_ref", "fqtool_ref")
NPAIRS = 2000
A1 = "AGATCGGAAGAGCACACGTCTGAACTCCAGTCA"
A2 = "AGATCGGAAGAGCGTCGTGTAGGGAAAGAGTGT"

CASES = {
    # name: (paired, flags)   -- flags of BASELINE.json configs 1-4 plus routing-heavy variants
    "cfg1": (True, "-q -a -g --enable_cut_tail"),
    "cfg2_r1": (False, "-q --enable_cut_front --enable_cut_tail"),
    "cfg3": (True, "-g -a -m"),
    "cfg4_r1": (False, f"-a --adapter_of_read1 {A1} -x -d --dup_ana_key_len 12 --kmer --kmer_length 6"),
    "pe_failed_unpaired": (True, f"-q -a --adapter_of_read1 {A1} --adapter_of_read2 {A2} -x --enable_cut_front --enable_cut_tail -l --min_length 60"),
    "pe_merge_keep": (True, "-m -a -q -l --min_length 50 --enable_cut_right"),
    "pe_merge_discard": (True, "-m --discard_unmerged -q"),
    "se_failed": (False, "-q -l --min_length 100 -f 2 -t 3 -g -y -Y 0.35"),
    "pe_correct": (True, "-c -a -q"),
    "pe_correct_merge": (True, "-c -m -a -g -q"),
}


def md5(path):
    return hashlib.md5(open(path, "rb").read()).hexdigest() if os.path.exists(path) else None


def outputs(case, paired, d):
    o = {"-o": f"{d}/out1.fq"}
    if paired:
        o["-O"] = f"{d}/out2.fq"
    if "-m" in CASES[case][1].split():
        o["--merge_output"] = f"{d}/merged.fq"
    if "failed" in case:
        o["--failed_out"] = f"{d}/failed.fq"
    if case == "pe_failed_unpaired":
        o["--unpaired_read1"] = f"{d}/un1.fq"; o["--unpaired_read2"] = f"{d}/un2.fq"
    return o


def command(exe, case, fq1, fq2, d):
    paired, flags = CASES[case]
    cmd = [exe, "-i", fq1] + (["-I", fq2] if paired else [])
    for k, v in outputs(case, paired, d).items():
        cmd += [k, v]
    return cmd + flags.split()


def main():
    if not os.path.exists(REF):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "-j8", "_ref/fqtool_ref"], check=True)
    os.makedirs(os.path.join(OUT, "fq"), exist_ok=True)
    fqs = []
    for name in ("r1", "r2"):
        lines = gzip.open(f"{TD}/{name}.fq.gz", "rb").read().split(b"\n")[:4 * NPAIRS]
        path = os.path.join(OUT, "fq", f"{name}.fq.gz")
        with gzip.GzipFile(path, "wb", mtime=0) as f:
            f.write(b"\n".join(lines) + b"\n")
        fqs.append(path)
    expected = {}
    for case in CASES:
        with tempfile.TemporaryDirectory() as d:
            cmd = command(REF, case, fqs[0], fqs[1], d) + ["-w", "1", "-J", f"{d}/report.json", "-H", f"{d}/report.html"]
            subprocess.run(cmd, check=True, cwd=d, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            files = {os.path.basename(v): md5(v) for v in outputs(case, CASES[case][0], d).values()}
            rep = json.load(open(f"{d}/report.json"))
            fr = rep.get("FilterResult", {})
            expected[case] = {"files": files, "filter_result": {k: fr[k] for k in fr if isinstance(fr[k], (int, float))}}
            print(case, files)
    json.dump(expected, open(os.path.join(OUT, "e2e_expected.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    sys.path.insert(0, ROOT)
    main()
