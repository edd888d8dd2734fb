"""Whole-program timing of the drop-in: the reference program (oracle/_ref/fqtool_ref) at
-w 1 and -w <cores> beside the same program with the worker bodies replaced by libfqk.so calls (oracle/_ref/fqtool_fqk)
at -w 1 and -w 4, on the same uncompressed synthetic FASTQ in /dev/shm, plain .fq output, wall clock of the whole run
(reader thread, workers, writer threads, reports).  Run on the GPU box:  python tools/dropin_time.py [reads] > out.json
The numbers say what a USER of the drop-in sees today; the kernels are not the limiter here."""
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fqtool_b200 import synth                                                    # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "fqtool_ref")
FQK = os.path.join(ROOT, "oracle", "_ref", "fqtool_fqk")
FLAGS = {"cfg2": "-q --enable_cut_front --enable_cut_tail", "cfg5": "-q -g -a"}


def write_fastq(path, reads, mate):
    seq, qual, off = reads
    n = len(off) - 1
    L = int(off[1] - off[0])
    assert (np.diff(off) == L).all()
    digits = (np.arange(n, dtype=np.int64)[:, None] // (10 ** np.arange(9, -1, -1, dtype=np.int64))[None, :] % 10 + 48).astype(np.uint8)
    head = np.frombuffer(b"@SIM:1:FCX:1:", dtype=np.uint8); tail = np.frombuffer(b" %d:N:0:ATCACG" % mate, dtype=np.uint8)
    nlen = len(head) + 10 + len(tail)
    rec = nlen + 1 + L + 1 + 2 + L + 1
    t = np.empty((n, rec), dtype=np.uint8)
    t[:, :len(head)] = head; t[:, len(head):len(head) + 10] = digits; t[:, len(head) + 10:nlen] = tail; t[:, nlen] = 10
    t[:, nlen + 1:nlen + 1 + L] = seq.reshape(n, L); t[:, nlen + 1 + L] = 10
    t[:, nlen + 2 + L] = ord("+"); t[:, nlen + 3 + L] = 10
    t[:, nlen + 4 + L:nlen + 4 + 2 * L] = qual.reshape(n, L); t[:, rec - 1] = 10
    t.tofile(path)


def run(exe, flags, inputs, workers, d):
    os.makedirs(d, exist_ok=True)
    cmd = [exe, "-i", inputs[0], "-o", os.path.join(d, "o1.fq")]
    if len(inputs) > 1:
        cmd += ["-I", inputs[1], "-O", os.path.join(d, "o2.fq")]
    cmd += flags.split() + ["-w", str(workers), "-J", os.path.join(d, "r.json"), "-H", os.path.join(d, "r.html")]
    t0 = time.perf_counter()
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=d)
    sec = time.perf_counter() - t0
    assert r.returncode == 0, (cmd, r.stderr[-500:])
    summ = json.load(open(os.path.join(d, "r.json")))["Summary"]
    for f in os.listdir(d):
        os.remove(os.path.join(d, f))
    return sec, summ


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
    tmp = "/dev/shm/fqk_dropin_time"
    os.makedirs(tmp, exist_ok=True)
    cores = os.cpu_count() or 1
    out = {"reads": n, "read_length": 150, "host_cores": cores, "input": "uncompressed FASTQ in /dev/shm, .fq output in /dev/shm", "runs": []}
    try:
        for cfg in ("cfg2", "cfg5"):
            paired = cfg == "cfg5"
            if paired:
                r1, r2 = synth.pe_novaseq(n, 150, seed=3, frag_mean=250, frag_sd=90)
                inputs = [os.path.join(tmp, "r1.fq"), os.path.join(tmp, "r2.fq")]
                write_fastq(inputs[0], r1, 1); write_fastq(inputs[1], r2, 2)
            else:
                inputs = [os.path.join(tmp, "r1.fq")]
                write_fastq(inputs[0], synth.se_hiseq(n, 150, seed=3), 1)
            bases = n * 150 * (2 if paired else 1)
            want = None
            for name, exe, w in (("reference -w 1", REF, 1), (f"reference -w {cores}", REF, cores), ("drop-in -w 1", FQK, 1), ("drop-in -w 4", FQK, 4)):
                run(exe, FLAGS[cfg], inputs, w, os.path.join(tmp, "out"))            # warm-up (page cache, CUDA context creation cost stays in)
                sec, summ = run(exe, FLAGS[cfg], inputs, w, os.path.join(tmp, "out"))
                want = want or summ
                assert summ == want, f"{cfg} {name}: report.json Summary differs from the reference's"
                out["runs"].append({"config": cfg, "flags": FLAGS[cfg], "program": name, "seconds": round(sec, 3),
                                    "gbases_per_s": round(bases / sec / 1e9, 4)})
                print(out["runs"][-1], file=sys.stderr)
    finally:
        subprocess.run(["rm", "-rf", tmp])
    print(json.dumps(out))


if __name__ == "__main__":
    main()
