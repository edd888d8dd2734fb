"""cfg4 with parts of its option set switched off (which kernel runs, what each part costs)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fqtool_b200 import workload
from fqtool_b200.engine import Engine
from fqtool_b200.options import from_cli
A = workload.ADAPTER1
variants = {"full": f"-a --adapter_of_read1 {A} -x -d --dup_ana_key_len 12 --kmer --kmer_length 6",
            "no kmer": f"-a --adapter_of_read1 {A} -x -d --dup_ana_key_len 12",
            "no kmer, no dup": f"-a --adapter_of_read1 {A} -x",
            "adapter only": f"-a --adapter_of_read1 {A}",
            "polyX only": "-x",
            "kmer only": "--kmer --kmer_length 6",
            "plain": ""}
n = int(os.environ.get("PARTS_N", "4000000"))
only = os.environ.get("PARTS_ONLY")
reps = int(os.environ.get("PARTS_REPS", "6"))
dev = torch.device("cuda", 0)
for name, cli in variants.items():
    if only and name != only:
        continue
    o = from_cli(cli.split()); o.maxCycles = 128
    eng = Engine(o.to_params(), device=0)
    b1, _ = workload.generate("cfg4", n, 0, dev, eng.stream)
    o1 = torch.empty(n * 16, dtype=torch.uint8, device=dev)
    s1 = b1.c_struct(); eng.sync(); eng.set_kernel_timing(True)
    for _ in range(reps):
        eng.process_se_device(s1, o1.data_ptr())
    eng.sync(); ms, cnt = eng.kernel_times()
    print(f"{name:18s} ms per launch [all, trim, stats, dup, fused]:", [round(m / max(c, 1), 3) for m, c in zip(ms, cnt)])
    eng.close(); del b1, o1
