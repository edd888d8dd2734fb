#!/usr/bin/env python
"""bench.py -- the hot-path benchmark (contract in the task statement, section 4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config cfg2] [--impl reference]

A step is one pass of the per-read QC+trim hot path over the whole HBM-resident
synthetic workload of this rank (cfg2 at N=1: 100 M SE reads of 150 bp in 25 batches
of 4 M reads).  `value` = Gbases/s over all ranks with inputs resident in HBM; `e2e` =
the same metric through the host-buffer C-ABI call (pinned host batches, H2D + kernels
+ D2H of the records inside the timed region).  One rank per GPU; accumulators are
merged by one NCCL all-reduce per run inside fqk_finalize (timed, after the K steps).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries the ONE JSON line and nothing else: the original stdout is kept aside for it and file descriptor 1 is
# pointed at stderr, so that whatever a library prints (NCCL's "NCCL version ..." banner still reached stdout with
# NCCL_DEBUG_FILE=/dev/stderr under torchrun) lands on stderr
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
sys.stdout.flush()
_JSON_FD = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_JSON_FD, (json.dumps(line) + "\n").encode())


BATCH_READS = 4_000_000


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="cfg2")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "generic"])
    ap.add_argument("--reads", type=int, default=0, help="reads (pairs) per GPU; 0 = the config's size capped by HBM")
    ap.add_argument("--cpu-sample", type=int, default=4_000_000, help="reads of the CPU-baseline sample")
    ap.add_argument("--e2e-steps", type=int, default=1)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


# ------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.rows, self.proc = gpu, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm = sorted(int(r[1]) for r in self.rows if len(r) >= 9 and r[1].isdigit())
        mx = [int(r[2]) for r in self.rows if len(r) >= 9 and r[2].isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v == "Active":
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------ CPU baseline
def cpu_reference_run(cfg, sample_reads, steps, warmup, torch_dev=None):
    """The reference's own hot path on all host cores (oracle/_ref, else the oracle port
    single-threaded), on a bounded sample of the same synthetic workload."""
    import numpy as np
    from fqtool_b200 import batch, synth, workload
    from oracle import pyoracle
    c = workload.CONFIGS[cfg]
    opt = workload.options_for(cfg)
    if torch_dev is not None:
        import torch
        b1, b2 = workload.generate(cfg, sample_reads, 0, torch_dev, None)
        torch.cuda.synchronize()
        r1 = batch.unpack(b1.to_host()); r2 = batch.unpack(b2.to_host()) if b2 else None
        sample = f"first {sample_reads} {'pairs' if c['paired'] else 'reads'} of the device-generated {cfg} stream"
        del b1, b2
    else:
        if c["paired"]:
            r1, r2 = synth.pe_novaseq(sample_reads, c["L"], seed=c["seed"], frag_mean=c["frag"][0], frag_sd=c["frag"][1])
        elif cfg == "cfg4":
            r1, r2 = synth.se_adapter_polyx(sample_reads, c["L"], seed=c["seed"]), None
        else:
            r1, r2 = synth.se_hiseq(sample_reads, c["L"], seed=c["seed"]), None
        sample = f"{sample_reads} host-generated (numpy) {cfg} reads: no GPU visible"
    bases = sample_reads * c["L"] * (2 if c["paired"] else 1)
    cores = os.cpu_count() or 1
    times = []
    if pyoracle.have_reference():
        kind = "reference"
        for i in range(warmup + steps):
            sec, _ = pyoracle.bench_reference(opt.to_params(), cores, r1, r2)
            if i >= warmup:
                times.append(sec)
    else:
        kind, cores = "port", 1
        for i in range(warmup + steps):
            o = pyoracle.Oracle(opt.to_params(), "port")
            t0 = time.perf_counter()
            o.process_se(r1) if r2 is None else o.process_pe(r1, r2)
            if i >= warmup:
                times.append(time.perf_counter() - t0)
            o.close()
    sec = sum(times) / len(times)
    return {"value": bases / sec / 1e9, "unit": "Gbases/s", "cores": cores, "kind": kind, "sample": sample,
            "ms_per_step": sec * 1e3}


def run_reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from fqtool_b200 import workload
    dev = None
    try:
        import torch
        if torch.cuda.is_available():
            dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    except Exception:
        dev = None
    c = workload.CONFIGS[a.config]
    cb = cpu_reference_run(a.config, a.cpu_sample, a.steps, min(a.warmup, 1), dev)
    line = {"impl": "reference", "metric": "Gbases/s per-read QC+trim", "value": cb["value"], "unit": "Gbases/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": min(a.warmup, 1), "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": c["desc"], "flags": c["cli"], "config": a.config},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": "Gbases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ----------------------------------------------------------------- B200 arm
def main():
    a = parse()
    if a.impl == "reference":
        return run_reference_arm(a)
    import numpy as np
    import torch
    import torch.distributed as dist
    from fqtool_b200 import abi, workload
    from fqtool_b200.engine import Engine

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    cfg = a.config
    c = workload.CONFIGS[cfg]
    opt = workload.options_for(cfg)
    params = opt.to_params()

    # one NCCL communicator inside the library (ncclUniqueId travels through torch.distributed)
    uid = None
    if world > 1:
        t = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(Engine.nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(t, 0)
        uid = bytes(t.cpu().numpy().tobytes())
    eng = Engine(params, device=local, rank=rank, nranks=world, nccl_unique_id=uid)
    if a.impl == "generic":
        eng.set_impl(1)
    stream = torch.cuda.ExternalStream(eng.stream, device=dev)

    # ---- workload: per-GPU share, resident in HBM (weak scaling: fixed per GPU) ----
    per_read_bytes = ((c["L"] + 15) // 16 * 16) * 1.25 + 6 + 16
    per_unit = per_read_bytes * (2 if c["paired"] else 1) + (24 if c["paired"] else 0)
    free_b, _ = torch.cuda.mem_get_info(dev)
    cap = int(free_b * 0.80 / per_unit)
    reads = a.reads or min(c["reads"], cap)
    reads = max(1, reads)
    nb = (reads + BATCH_READS - 1) // BATCH_READS
    first = rank * reads
    batches, outs = [], []
    for i in range(nb):
        n = min(BATCH_READS, reads - i * BATCH_READS)
        b1, b2 = workload.generate(cfg, n, first + i * BATCH_READS, dev, eng.stream)
        o1 = torch.empty(n * 16, dtype=torch.uint8, device=dev)
        o2 = torch.empty(n * 16, dtype=torch.uint8, device=dev) if b2 else None
        op = torch.empty(n * 24, dtype=torch.uint8, device=dev) if b2 else None
        batches.append((b1, b2, b1.c_struct(), b2.c_struct() if b2 else None)); outs.append((o1, o2, op))
    eng.sync()
    units = reads
    bases_per_step = units * c["L"] * (2 if c["paired"] else 1)

    def step():
        for (b1, b2, s1, s2), (o1, o2, op) in zip(batches, outs):
            if b2 is None:
                eng.process_se_device(s1, o1.data_ptr())
            else:
                eng.process_pe_device(s1, s2, o1.data_ptr(), o2.data_ptr(), op.data_ptr())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(a.warmup):
        step()
    eng.finalize()                               # warm-up of the run-level merge too (NCCL connects lazily)
    eng.reset()
    eng.set_kernel_timing(True)
    eng.kernel_times()
    sampler = ClockSampler(local); sampler.start()
    launches0 = eng.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(a.steps):
        step()
    flat = eng.finalize()                       # the once-per-run NCCL all-reduce + D2H of the accumulators
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = eng.launch_count - launches0
    clocks = sampler.stop()
    kms, kn = eng.kernel_times()
    eng.set_kernel_timing(False)
    tmax = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms_total = float(tmax.item())
    value = bases_per_step * world * a.steps / (ms_total / 1e3) / 1e9

    # sanity of the timed work: every read of every step was counted (pre1 reads over all ranks)
    L = eng.layout
    reads_counted = int(flat[L.stats_off[0] + L.s_reads])
    assert reads_counted == units * world * a.steps, (reads_counted, units, world, a.steps)

    # ---- roofline of the dominant kernel (live CUDA-event durations on the launching stream) ----
    peaks = {}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    peak = peaks.get("hbm_gbs", 6650.0)
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    fused_name = "k_se_v3" if (not c["paired"] and not opt.to_params().kmer_len and a.impl == "b200") else "k_fused_se_ws"
    names = {abi.T_TRIM: "k_trim", abi.T_STATS: "k_stats", abi.T_DUP: "k_dup", abi.T_FUSED: fused_name}
    dom = max(names, key=lambda s: kms[s])
    per_launch_ms = kms[dom] / max(kn[dom], 1)
    alg = workload.algorithmic_bytes(cfg, min(BATCH_READS, reads))
    achieved = alg / (per_launch_ms / 1e3) / 1e9 if per_launch_ms > 0 else 0.0
    roofline = {"bound": "hbm", "kernel": names[dom], "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg, "avg_launch_ms": per_launch_ms,
                "kernel_ms_share": {names[s]: kms[s] / max(kms[abi.T_ALL], 1e-9) for s in names if kn[s]},
                "whole_path_achieved": workload.algorithmic_bytes(cfg, units) * a.steps / (kms[abi.T_ALL] / 1e3) / 1e9
                if kms[abi.T_ALL] > 0 else None}
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        roofline["traffic"] = json.load(open(tr)).get(names[dom])

    # ---- end to end through the host-buffer C-ABI call ----
    n_host = min(4, nb)
    hbs = [(batches[i][0].to_host(pinned=True), batches[i][1].to_host(pinned=True) if batches[i][1] else None)
           for i in range(n_host)]
    hout = [(torch.empty(b[0].n * 16, dtype=torch.uint8, pin_memory=True),
             torch.empty(b[0].n * 16, dtype=torch.uint8, pin_memory=True) if b[1] else None,
             torch.empty(b[0].n * 24, dtype=torch.uint8, pin_memory=True) if b[1] else None) for b in hbs]
    h2d = d2h = 0; e2e_units = 0

    def e2e_step():
        nonlocal h2d, d2h, e2e_units
        h2d = d2h = 0; e2e_units = 0
        pending = None
        for i in range(nb):
            (h1, h2), (r1, r2, rp) = hbs[i % n_host], hout[i % n_host]
            if h2 is None:
                # two packs in flight: the upload of pack i overlaps kernels + download of pack i-1
                t = eng.submit_se(h1, r1.numpy().view(abi.READ_RESULT_DTYPE))
                if pending is not None:
                    eng.wait(pending)
                pending = t
                h2d += h1.nbytes; d2h += h1.n * 16
            else:
                eng.process_pe(h1, h2, (r1.numpy().view(abi.READ_RESULT_DTYPE), r2.numpy().view(abi.READ_RESULT_DTYPE),
                                        rp.numpy().view(abi.PAIR_RESULT_DTYPE)))
                h2d += h1.nbytes + h2.nbytes; d2h += h1.n * (16 + 16 + 24)
            e2e_units += h1.n
        if pending is not None:
            eng.wait(pending)
    e2e_step()                                   # warm-up (allocates the staging buffers)
    eng.reset()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.e2e_steps):
        e2e_step()
    flat2 = eng.finalize()
    d2h_final = flat2.nbytes
    barrier()
    dt = time.perf_counter() - t0
    tt = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    e2e_bases = e2e_units * c["L"] * (2 if c["paired"] else 1) * a.e2e_steps * world
    e2e = {"value": e2e_bases / float(tt.item()) / 1e9, "unit": "Gbases/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h + d2h_final // max(a.e2e_steps, 1), "steps": a.e2e_steps,
           "ms_per_step": float(tt.item()) * 1e3 / a.e2e_steps,
           "note": f"{nb} fqk_submit_se/fqk_wait (SE, two packs in flight) or fqk_process_pe calls per step over {n_host} distinct pinned host batches of {BATCH_READS} reads"}

    # ---- the same through the ASCII entry point (FASTQ characters in, packing on the GPU) ----
    e2e_ascii = None
    if not c["paired"]:
        from fqtool_b200 import batch as fbatch
        seq, qual, off = fbatch.unpack(hbs[0][0])
        pseq = torch.empty(len(seq), dtype=torch.uint8, pin_memory=True); pseq.numpy()[:] = seq
        pqual = torch.empty(len(qual), dtype=torch.uint8, pin_memory=True); pqual.numpy()[:] = qual
        areads = (pseq.numpy(), pqual.numpy(), off)
        eng.process_se_ascii(areads)
        eng.reset(); barrier()
        t0 = time.perf_counter()
        reps = min(nb, 8)
        for _ in range(reps):
            eng.process_se_ascii(areads)
        eng.finalize(); barrier()
        dta = time.perf_counter() - t0
        e2e_ascii = {"value": reps * hbs[0][0].n * c["L"] / dta / 1e9, "unit": "Gbases/s",
                     "h2d_bytes_per_call": int(2 * len(seq) + 10 * hbs[0][0].n), "calls": reps,
                     "note": "fqk_process_se_ascii: FASTQ characters over PCIe (2 B/base), 2-bit packing by k_pack_ascii on the GPU"}
        del pseq, pqual

    cpu_baseline = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        del batches, outs
        torch.cuda.empty_cache()
        cb = cpu_reference_run(cfg, a.cpu_sample, 1, 0, dev)
        cpu_baseline = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}

    if rank == 0:
        line = {"metric": "Gbases/s per-read QC+trim", "value": value, "unit": "Gbases/s", "n_gpus": world,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_total / a.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": c["desc"], "config": cfg, "flags": c["cli"], "reads_per_gpu": reads,
                           "read_length": c["L"], "batches_per_step": nb, "batch_reads": BATCH_READS,
                           "l2": "inputs larger than L2 (%.1f GB resident per GPU)" % (reads * per_unit / 1e9),
                           "impl": a.impl},
                "clocks": clocks, "e2e": e2e, "e2e_ascii": e2e_ascii, "gpu_launches": launches, "roofline": roofline,
                "cpu_baseline": cpu_baseline}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
