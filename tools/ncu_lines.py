"""Per-source-line instruction counts and stall samples from an ncu capture (needs -lineinfo).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top]
"""
import csv
import subprocess
import sys
from collections import defaultdict

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
fname, hdr = "?", None
agg = defaultdict(lambda: [0, 0, 0, ""])   # (file, line) -> [warp instr, samples, sass count, text]
cur = None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r; continue
    if hdr is None:
        continue
    d = dict(zip(range(len(r)), r))
    if r[0] != "":
        cur = (fname, int(r[0])); agg[cur][3] = r[1].strip()[:90]
    if len(r) > 7 and r[2]:
        try:
            agg[cur][0] += int(r[7]); agg[cur][1] += int(r[6] or 0); agg[cur][2] += 1
        except ValueError:
            pass
tot_i = sum(v[0] for v in agg.values()); tot_s = sum(v[1] for v in agg.values())
print(f"total warp instr {tot_i}  samples {tot_s}")
print("by file:")
byf = defaultdict(lambda: [0, 0])
for (f, l), v in agg.items():
    byf[f][0] += v[0]; byf[f][1] += v[1]
for f, v in sorted(byf.items(), key=lambda kv: -kv[1][0]):
    print(f"  {f:20s} instr {100*v[0]/max(tot_i,1):5.1f}%  samples {100*v[1]/max(tot_s,1):5.1f}%")
print("top lines by instructions:")
for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"  {f}:{l:<5d} instr {100*v[0]/max(tot_i,1):5.2f}%  samples {100*v[1]/max(tot_s,1):5.2f}%  sass {v[2]:4d}  {v[3]}")
print("top lines by stall samples:")
for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"  {f}:{l:<5d} instr {100*v[0]/max(tot_i,1):5.2f}%  samples {100*v[1]/max(tot_s,1):5.2f}%  sass {v[2]:4d}  {v[3]}")
