"""Small driver for ncu: a few launches of the hot path on one device-resident batch.

    python tools/prof_run.py [cfg] [reads] [launches] [impl]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from fqtool_b200 import workload
from fqtool_b200.engine import Engine

cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4_000_000
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
impl = int(sys.argv[4]) if len(sys.argv) > 4 else 0
dev = torch.device("cuda", 0)
eng = Engine(workload.options_for(cfg).to_params(), device=0)
eng.set_impl(impl)
b1, b2 = workload.generate(cfg, n, 0, dev, eng.stream)
o1 = torch.empty(n * 16, dtype=torch.uint8, device=dev)
o2 = torch.empty(n * 16, dtype=torch.uint8, device=dev) if b2 else None
op = torch.empty(n * 24, dtype=torch.uint8, device=dev) if b2 else None
s1, s2 = b1.c_struct(), (b2.c_struct() if b2 else None)
eng.sync()
eng.set_kernel_timing(True)
for _ in range(reps):
    if b2 is None:
        eng.process_se_device(s1, o1.data_ptr())
    else:
        eng.process_pe_device(s1, s2, o1.data_ptr(), o2.data_ptr(), op.data_ptr())
eng.sync()
ms, cnt = eng.kernel_times()
print("kernel ms per launch [all, trim, stats, dup, fused]:", [round(m / max(c, 1), 4) for m, c in zip(ms, cnt)])
import ctypes
if hasattr(eng.lib, "fqk_debug_prof"):
    buf = (ctypes.c_ulonglong * 16)()
    if eng.lib.fqk_debug_prof(buf, 0):
        names = ["trim wait full", "trim chunk", "trim n_bases", "stats wait ready", "stats P", "stats exceptions", "producer wait free",
                 "producer issue", "block total", "trim sweep", "trim cut", "trim tail", "trim result", "-", "-", "-"]
        tot = buf[8] or 1
        print("cycle accounting (sum over warps / block total):")
        for nm, v in zip(names, buf):
            if v:
                print(f"  {nm:20s} {v / tot:8.3f}")
flat = eng.finalize()
print("reads counted:", int(flat[eng.layout.stats_off[0] + eng.layout.s_reads]))
