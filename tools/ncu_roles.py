"""Stall reasons per source-line bucket of an ncu capture (source page exported to CSV).

    ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass > src.csv
    python tools/ncu_roles.py src.csv file:lo-hi=name ...
"""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
buckets = []
for a in sys.argv[2:]:
    rng, name = a.split("=")
    f, r = rng.split(":")
    lo, hi = r.split("-")
    buckets.append((f, int(lo), int(hi), name))
hdr = next(r for r in rows if r and r[0] == "Line No")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]


def I(x):
    try:
        return int(x)
    except ValueError:
        return 0


fname, cur = "?", None
addr = {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]; continue
    if r[0] in ("Function Name", "Line No"):
        continue
    if r[0] != "":
        cur = (fname, I(r[0]))
    if len(r) > 7 and r[2]:
        a = addr.setdefault(r[2], {"n": I(r[7]), "st": {h: I(r[i]) for i, h in stall_cols}, "lines": []})
        a["lines"].append(cur)
agg = defaultdict(lambda: defaultdict(int))
for a in addr.values():
    name = "other"
    for f, lo, hi, nm in buckets:
        if any(l[0] == f and lo <= l[1] <= hi for l in a["lines"]):
            name = nm; break
    agg[name]["instr"] += a["n"]
    for h, v in a["st"].items():
        agg[name][h] += v
keys = ["stall_selected", "stall_no_inst", "stall_wait", "stall_short_sb", "stall_branch_resolving", "stall_barrier", "stall_math",
        "stall_not_selected", "stall_long_sb", "stall_mio", "stall_dispatch", "stall_sleep"]
print(f"{'bucket':22s} {'instr':>12s} {'samples':>8s} " + " ".join(f"{k[6:14]:>8s}" for k in keys))
for name, d in sorted(agg.items(), key=lambda kv: -kv[1]["instr"]):
    tot = sum(v for k, v in d.items() if k.startswith("stall_"))
    print(f"{name:22s} {d['instr']:12d} {tot:8d} " + " ".join(f"{d[k]:8d}" for k in keys))
