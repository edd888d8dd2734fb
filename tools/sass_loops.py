"""List the loops (backward branches) of a kernel's SASS with their instruction mix.

    cuobjdump -sass lib.so | python tools/sass_loops.py <function-substring> [min-instr]
"""
import re
import sys
from collections import Counter

pat = sys.argv[1]
mini = int(sys.argv[2]) if len(sys.argv) > 2 else 8
ins = []          # (addr, text)
on = False
for line in sys.stdin:
    if "Function :" in line:
        on = pat in line
        if on:
            ins = []
        continue
    if not on:
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
addr_idx = {a: i for i, (a, _) in enumerate(ins)}
print(f"{len(ins)} instructions")
for i, (a, t) in enumerate(ins):
    m = re.search(r"\bBRA(?:\.\w+)*\s+(?:\w+,\s*)?(0x[0-9a-f]+)", t)
    if not m:
        continue
    tgt = int(m.group(1), 16)
    if tgt <= a and tgt in addr_idx:
        j = addr_idx[tgt]
        body = [x[1] for x in ins[j:i + 1]]
        if len(body) < mini:
            continue
        ops = Counter(re.sub(r"^@!?U?P\d+\s+", "", b).split()[0].split(".")[0] for b in body)
        print(f"loop {tgt:#x}..{a:#x}: {len(body)} instr  " + " ".join(f"{k}:{v}" for k, v in ops.most_common(14)))
