"""Per-source-line instruction counts and stall samples from an ncu capture (needs -lineinfo).
Every SASS instruction is attributed once, to the most specific source line that lists it (inlined code
is listed under the callee line and under every call site).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top] [file:lo-hi=name ...]
"""
import csv
import subprocess
import sys
from collections import defaultdict

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
buckets = []
for a in sys.argv[3:]:
    rng, name = a.split("=")
    f, r = rng.split(":")
    lo, hi = r.split("-")
    buckets.append((f, int(lo), int(hi), name))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))


def I(x):
    try:
        return int(x)
    except ValueError:
        return 0


fname, cur = "?", None
text = {}
nsass = defaultdict(int)
addr = {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]; continue
    if r[0] in ("Function Name", "Line No"):
        continue
    if r[0] != "":
        cur = (fname, I(r[0])); text[cur] = r[1].strip()[:80]
    if len(r) > 7 and r[2]:
        nsass[cur] += 1
        a = addr.setdefault(r[2], {"n": I(r[7]), "s": I(r[6]), "lines": []})
        a["lines"].append(cur)
agg = defaultdict(lambda: [0, 0])
bagg = defaultdict(lambda: [0, 0])
for a in addr.values():
    best = min(a["lines"], key=lambda l: nsass[l])
    agg[best][0] += a["n"]; agg[best][1] += a["s"]
    name = "other"
    for f, lo, hi, nm in buckets:
        if any(l[0] == f and lo <= l[1] <= hi for l in a["lines"]):
            name = nm; break
    bagg[name][0] += a["n"]; bagg[name][1] += a["s"]
ti = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
print(f"total warp instr {ti}  samples {ts}")
if buckets:
    for k, v in sorted(bagg.items(), key=lambda kv: -kv[1][0]):
        print(f"  {k:24s} instr {100*v[0]/max(ti,1):5.1f}%  samples {100*v[1]/max(ts,1):5.1f}%")
print("top lines by instructions:")
for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"  {f}:{l:<5d} instr {100*v[0]/max(ti,1):5.2f}%  samples {100*v[1]/max(ts,1):5.2f}%  {text.get((f,l),'')}")
print("top lines by stall samples:")
for (f, l), v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top // 2]:
    print(f"  {f}:{l:<5d} instr {100*v[0]/max(ti,1):5.2f}%  samples {100*v[1]/max(ts,1):5.2f}%  {text.get((f,l),'')}")
