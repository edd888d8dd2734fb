"""Stall-reason totals of an ncu capture, overall and for SASS whose text matches a regex.

    python tools/ncu_stalls.py rep.ncu-rep [regex]
"""
import csv, subprocess, sys, re
from collections import defaultdict
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = None
tot = defaultdict(int); inst = 0; samples = 0
opc = defaultdict(lambda: [0, 0])
for r in rows:
    if r and r[0] == "Address":
        hdr = r; continue
    if hdr is None or len(r) != len(hdr):
        continue
    d = dict(zip(hdr, r))
    n = int(d["Instructions Executed"] or 0); s = int(d["# Samples"] or 0)
    inst += n; samples += s
    op = d["Source"].split()[0] if d["Source"].split() else "?"
    if op.startswith("@"):
        op = d["Source"].split()[1]
    op = op.split(".")[0]
    opc[op][0] += n; opc[op][1] += s
    for k in hdr:
        if k.startswith("stall_") and "Not Issued" not in k:
            tot[k] += int(d[k] or 0)
print("instructions", inst, "samples", samples)
for k, v in sorted(tot.items(), key=lambda x: -x[1])[:12]:
    print(f"  {k:28s} {100.0 * v / max(samples, 1):6.2f}%")
print("opcodes by instructions:")
for k, (n, s) in sorted(opc.items(), key=lambda x: -x[1][0])[:28]:
    print(f"  {k:10s} instr {100.0 * n / inst:6.2f}%  samples {100.0 * s / max(samples, 1):6.2f}%")
