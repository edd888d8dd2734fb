#!/usr/bin/env python
"""bench.py -- the hot-path benchmark (contract in the task statement, section 4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config cfg2] [--impl reference]

A step is one pass of the per-read QC+trim hot path over the whole HBM-resident
synthetic workload of this rank (cfg2 at N=1: 100 M SE reads of 150 bp in 25 batches
of 4 M reads).  `value` = Gbases/s over all ranks with inputs resident in HBM; `e2e` =
the same metric through the host-buffer C-ABI call (pinned host batches, H2D + kernels
+ D2H of the records inside the timed region; median of --e2e-steps steps).  One rank
per GPU; accumulators are merged by one NCCL all-reduce per run inside fqk_finalize
(timed, after the K steps).

The same JSON line also carries, under `configs`, the other synthetic BASELINE.json
configurations measured with the same protocol on smaller resident workloads: cfg3 and
cfg4 (weak scaling, --extra-units per GPU; cfg4's fqk_finalize runs the four
duplication-table all-reduces and is timed as `finalize_ms`) and cfg5 as a STRONG
scaling record (`scaling_cfg5`: a fixed total of --cfg5-total pairs sharded contiguously
by global pair index over the ranks).  At N > 1 `merge_check` compares the NCCL-merged
accumulators and duplication table of two small sharded streams bit-exactly with the
same streams run on one rank.
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
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "generic", "sample"])
    ap.add_argument("--sample-out", default="", help="--impl sample: write the first --cpu-sample units of --config's stream here (.npz)")
    ap.add_argument("--check-units", type=int, default=4_000_000,
                    help="units of every timed stream whose records and accumulators are compared with the oracle (0 = skip)")
    ap.add_argument("--reads", type=int, default=0, help="reads (pairs) per GPU; 0 = the config's size capped by HBM")
    ap.add_argument("--cpu-sample", type=int, default=4_000_000, help="reads of the CPU-baseline sample")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the cfg3/cfg4/cfg5 records and the merge check")
    ap.add_argument("--extra-units", type=int, default=16_000_000, help="reads (pairs) per GPU of the cfg3/cfg4 records")
    ap.add_argument("--cfg5-total", type=int, default=128_000_000, help="pairs of the cfg5 strong-scaling record (whole job)")
    ap.add_argument("--extra-steps", type=int, default=5)
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
def device_sample(cfg, n, torch_dev):
    """First n units of the device-generated stream of cfg as host reads (seq, qual, off) per mate."""
    import torch
    from fqtool_b200 import batch, workload
    b1, b2 = workload.generate(cfg, n, 0, torch_dev, None)
    torch.cuda.synchronize()
    r1 = batch.unpack(b1.to_host()); r2 = batch.unpack(b2.to_host()) if b2 else None
    return r1, r2


def write_sample(a):
    """--impl sample (a helper process of the reference arm): the sample comes from the same counter-based generator the B200
    arm times, which lives in libfqk.so; it is produced HERE and handed over as a file so that the process that times the
    reference never loads the library or touches the GPU."""
    import numpy as np
    import torch
    r1, r2 = device_sample(a.config, a.cpu_sample, torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))
    arrs = {"s1": r1[0], "q1": r1[1], "o1": r1[2]}
    if r2 is not None:
        arrs.update({"s2": r2[0], "q2": r2[1], "o2": r2[2]})
    np.savez(a.sample_out, **arrs)


def oracle_flat_threaded(params, r1, r2, threads):
    """Records and flat accumulators of the CPU oracle (the reference's own objects when oracle/_ref is built, else the
    port) over the reads, split into `threads` contiguous chunks run side by side: every entry of the flat buffer is a sum
    over reads, so the chunk results add up (the duplication table is not part of this check: it is order-dependent and
    covered by the parity tests and merge_check)."""
    import numpy as np
    from concurrent.futures import ThreadPoolExecutor
    from oracle import pyoracle
    kind = "reference" if pyoracle.have_reference() else "port"
    n = len(r1[2]) - 1
    cuts = [n * t // threads for t in range(threads + 1)]

    def sub(r, lo, hi):
        seq, qual, off = r
        a, b = int(off[lo]), int(off[hi])
        return seq[a:b], qual[a:b], (off[lo:hi + 1] - off[lo]).astype(np.uint64)

    def work(t):
        lo, hi = cuts[t], cuts[t + 1]
        o = pyoracle.Oracle(params, kind)
        rec = (o.process_se(sub(r1, lo, hi), lo),) if r2 is None else o.process_pe(sub(r1, lo, hi), sub(r2, lo, hi), lo)
        flat = o.finalize()
        o.close()
        return rec, flat
    with ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(work, range(threads)))
    recs = tuple(np.concatenate([p[0][k] for p in parts]) for k in range(len(parts[0][0])))
    flat = np.sum([p[1] for p in parts], axis=0, dtype=np.uint64)
    return kind, recs, flat


def oracle_check(ctx, cfg, units, reads=None):
    """The first `units` units of cfg's timed stream through the product path (fqk_process_*_device on a fresh context) against
    the CPU oracle: every field of every record and every entry of the flat accumulator buffer, bit for bit."""
    import numpy as np
    import torch
    from fqtool_b200 import abi, batch, workload
    params = workload.options_for(cfg).to_params()
    eng = ctx.engine(params, single=True)
    b1, b2 = workload.generate(cfg, units, 0, ctx.dev, eng.stream)
    o1 = torch.zeros(units * 16, dtype=torch.uint8, device=ctx.dev)
    if b2 is None:
        eng.process_se_device(b1.c_struct(), o1.data_ptr())
    else:
        o2 = torch.zeros(units * 16, dtype=torch.uint8, device=ctx.dev); op = torch.zeros(units * 24, dtype=torch.uint8, device=ctx.dev)
        eng.process_pe_device(b1.c_struct(), b2.c_struct(), o1.data_ptr(), o2.data_ptr(), op.data_ptr())
    eng.sync()
    flat = eng.finalize()
    L = eng.layout
    recs = [o1.cpu().numpy().view(abi.READ_RESULT_DTYPE)]
    if b2 is not None:
        recs += [o2.cpu().numpy().view(abi.READ_RESULT_DTYPE), op.cpu().numpy().view(abi.PAIR_RESULT_DTYPE)]
    if reads is None:
        reads = (batch.unpack(b1.to_host()), batch.unpack(b2.to_host()) if b2 else None)
    eng.close()
    t0 = time.perf_counter()
    kind, orecs, oflat = oracle_flat_threaded(params, reads[0], reads[1], os.cpu_count() or 1)
    sec = time.perf_counter() - t0
    for k, (x, y) in enumerate(zip(recs, orecs)):
        for name in x.dtype.names:
            if name == "reserved":
                continue
            bad = np.nonzero(x[name] != y[name])[0]
            assert bad.size == 0, f"oracle_check {cfg}: records[{k}].{name} differs for {bad.size} units, first {bad[0]}"
    bad = np.nonzero(flat != oflat)[0]
    assert bad.size == 0, f"oracle_check {cfg}: flat accumulators differ at {bad[:8]}"
    return {"status": "ok", "units": units, "oracle": kind, "oracle_seconds": round(sec, 2),
            "reads_passed": int(flat[L.filter_off]), "flat_u64": int(len(flat)),
            "what": "first units of the timed stream: every record field and every flat accumulator entry == CPU oracle, bit-exact"}


def cpu_reference_run(cfg, sample_reads, steps, warmup, torch_dev=None, sample_file=None):
    """The reference's own hot path on all host cores (oracle/_ref, else the oracle port
    single-threaded), on a bounded sample of the same synthetic workload."""
    import numpy as np
    from fqtool_b200 import batch, synth, workload
    from oracle import pyoracle
    c = workload.CONFIGS[cfg]
    opt = workload.options_for(cfg)
    if sample_file is not None:
        z = np.load(sample_file)
        r1 = (z["s1"], z["q1"], z["o1"]); r2 = (z["s2"], z["q2"], z["o2"]) if "s2" in z.files else None
        sample = f"first {sample_reads} {'pairs' if c['paired'] else 'reads'} of the device-generated {cfg} stream"
    elif torch_dev is not None:
        r1, r2 = device_sample(cfg, sample_reads, torch_dev)
        sample = f"first {sample_reads} {'pairs' if c['paired'] else 'reads'} of the device-generated {cfg} stream"
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
            "ms_per_step": sec * 1e3, "reads": (r1, r2)}


def per_unit_bytes(c):
    per_read_bytes = ((c["L"] + 15) // 16 * 16) * 1.25 + 6 + 16
    return per_read_bytes * (2 if c["paired"] else 1) + (24 if c["paired"] else 0)


def config_dict(a, c, reads):
    """`config` of the JSON line: the same keys and values from both arms (the reference arm times a bounded sample of this
    workload and says so in cpu_baseline.sample)."""
    nb = (reads + BATCH_READS - 1) // BATCH_READS
    return {"workload": c["desc"], "config": a.config, "flags": c["cli"], "reads_per_gpu": reads, "read_length": c["L"],
            "batches_per_step": nb, "batch_reads": BATCH_READS,
            "l2": "inputs larger than L2 (%.1f GB resident per GPU)" % (reads * per_unit_bytes(c) / 1e9)}


def run_reference_arm(a):
    """The reference's own hot path on the host cores.  This process loads neither libfqk.so nor CUDA: the sample (first
    --cpu-sample units of the stream the B200 arm times) is produced by a helper process (--impl sample) and read from a
    file; without a GPU the numpy generator stands in."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import tempfile
    from fqtool_b200 import workload
    c = workload.CONFIGS[a.config]
    tmpdir = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else tempfile.gettempdir()
    path = os.path.join(tmpdir, f"fqk_sample_{os.getpid()}.npz")
    sample_file = None
    try:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "sample", "--config", a.config, "--cpu-sample",
                            str(a.cpu_sample), "--sample-out", path], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True)
        if r.returncode == 0 and os.path.exists(path):
            sample_file = path
        else:
            sys.stderr.write("bench.py --impl reference: no device sample (%s); using the numpy generator\n" % r.stderr.strip()[-300:])
        cb = cpu_reference_run(a.config, a.cpu_sample, a.steps, a.warmup, None, sample_file)
    finally:
        if os.path.exists(path):
            os.remove(path)
    line = {"impl": "reference", "metric": "Gbases/s per-read QC+trim", "value": cb["value"], "unit": "Gbases/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": config_dict(a, c, a.reads or c["reads"]),
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": "Gbases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ----------------------------------------------------------------- B200 arm
def kernel_name(cfg, params, impl):
    from fqtool_b200 import workload
    c = workload.CONFIGS[cfg]
    if c["paired"]:
        return "k_pe_v4" if impl == "b200" else "k_trim_pe+k_stats_pe"
    return "k_se_v3" if (impl == "b200" and params.kmer_len <= 13) else "k_fused_se_ws"   # k-mers run as their own pass (k_kmer_se)


class Job:
    """One configuration resident in HBM on this rank: engine, batches, record buffers."""

    def __init__(self, ctx, cfg, units, first):
        import torch
        from fqtool_b200 import workload
        self.ctx, self.cfg, self.units, self.c = ctx, cfg, units, workload.CONFIGS[cfg]
        self.opt = workload.options_for(cfg)
        self.params = self.opt.to_params()
        self.eng = ctx.engine(self.params)
        if ctx.a.impl == "generic":
            self.eng.set_impl(1)
        self.stream = torch.cuda.ExternalStream(self.eng.stream, device=ctx.dev)
        self.batches, self.outs = [], []
        dev = ctx.dev
        for i in range((units + BATCH_READS - 1) // BATCH_READS):
            n = min(BATCH_READS, units - i * BATCH_READS)
            b1, b2 = workload.generate(cfg, n, first + i * BATCH_READS, dev, self.eng.stream)
            o1 = torch.empty(n * 16, dtype=torch.uint8, device=dev)
            o2 = torch.empty(n * 16, dtype=torch.uint8, device=dev) if b2 else None
            op = torch.empty(n * 24, dtype=torch.uint8, device=dev) if b2 else None
            self.batches.append((b1, b2, b1.c_struct(), b2.c_struct() if b2 else None)); self.outs.append((o1, o2, op))
        self.eng.sync()
        self.nb = len(self.batches)
        self.bases_per_step = units * self.c["L"] * (2 if self.c["paired"] else 1)

    def step(self):
        e = self.eng
        for (b1, b2, s1, s2), (o1, o2, op) in zip(self.batches, self.outs):
            if b2 is None:
                e.process_se_device(s1, o1.data_ptr())
            else:
                e.process_pe_device(s1, s2, o1.data_ptr(), o2.data_ptr(), op.data_ptr())

    def timed(self, steps, warmup, sampler=None):
        """warmup untimed steps, then `steps` steps + fqk_finalize between two events; max over ranks."""
        import torch
        from fqtool_b200 import abi, workload
        ctx, eng = self.ctx, self.eng
        for _ in range(warmup):
            self.step()
        eng.finalize()                               # warm-up of the run-level merge too (NCCL connects lazily)
        eng.reset()
        eng.set_kernel_timing(True)
        eng.kernel_times()
        if sampler:
            sampler.start()
        launches0 = eng.launch_count
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        ctx.barrier()
        e0.record(self.stream)
        for _ in range(steps):
            self.step()
        e1.record(self.stream)
        flat = eng.finalize()                       # the once-per-run NCCL all-reduce(s) + D2H of the accumulators
        e2.record(self.stream)
        ctx.barrier()
        ms_total = ctx.max_over_ranks(e0.elapsed_time(e2))
        fin_ms = ctx.max_over_ranks(e1.elapsed_time(e2))
        launches = eng.launch_count - launches0
        kms, kn = eng.kernel_times()
        eng.set_kernel_timing(False)
        L = eng.layout
        total_units = ctx.sum_over_ranks(self.units)
        counted = int(flat[L.stats_off[0] + L.s_reads])
        assert counted == total_units * steps, (self.cfg, counted, total_units, steps)   # every read of every step was counted
        total_bases = total_units * self.c["L"] * (2 if self.c["paired"] else 1)
        # roofline of the dominant kernel: live CUDA-event durations on the launching stream
        fused = kernel_name(self.cfg, self.params, ctx.a.impl)
        stats_name = "k_kmer_se" if (not self.c["paired"] and self.params.kmer_len and ctx.a.impl == "b200") else "k_stats"
        names = {abi.T_TRIM: "k_trim", abi.T_STATS: stats_name, abi.T_DUP: "k_dup", abi.T_FUSED: fused}
        dom = max(names, key=lambda s: kms[s])
        per_launch_ms = kms[dom] / max(kn[dom], 1)
        per_launch_units = min(BATCH_READS, self.units)
        alg = workload.algorithmic_bytes(self.cfg, per_launch_units)
        achieved = alg / (per_launch_ms / 1e3) / 1e9 if per_launch_ms > 0 else 0.0
        roofline = {"bound": "hbm", "kernel": names[dom], "achieved": achieved, "peak": ctx.peak, "unit": "GB/s",
                    "frac": achieved / ctx.peak, "traffic": ctx.traffic.get(self.cfg, {}).get(names[dom]),
                    "traffic_source": "profiles/traffic.json: dram bytes read + written per launch of 4 M units, ncu --set full capture of this "
                                      "kernel on this configuration (null = not captured for this configuration)",
                    "peak_source": ctx.peak_src,
                    "algorithmic_bytes_per_launch": alg, "avg_launch_ms": per_launch_ms,
                    "kernel_ms_share": {names[s]: kms[s] / max(kms[abi.T_ALL], 1e-9) for s in names if kn[s]},
                    "ms_per_launch": {names[s]: kms[s] / kn[s] for s in names if kn[s]},
                    "whole_path_achieved": workload.algorithmic_bytes(self.cfg, self.units) * steps / (kms[abi.T_ALL] / 1e3) / 1e9
                    if kms[abi.T_ALL] > 0 else None}
        if roofline["whole_path_achieved"]:
            # all kernels of the path together (cfg4: k_se_v3 + k_kmer_se + k_dup), against the same peak
            roofline["whole_path_frac"] = roofline["whole_path_achieved"] / ctx.peak
        return {"value": total_bases * steps / (ms_total / 1e3) / 1e9, "unit": "Gbases/s", "ms_per_step": ms_total / steps,
                "finalize_ms": fin_ms, "steps": steps, "warmup": warmup, "units_per_gpu": self.units,
                "units_total": total_units, "batches_per_step": self.nb, "gpu_launches": launches, "roofline": roofline,
                "flags": self.c["cli"]}

    def e2e(self, steps):
        """End to end through the host-buffer C-ABI call for this configuration: every batch of the step goes H2D from pinned
        host memory, through the kernels, records D2H, then fqk_finalize.
        Two packs in flight (fqk_submit_pe / fqk_submit_se + fqk_wait); up to two distinct pinned host batches are cycled.  Median over steps of the max over ranks (wall clock)."""
        import torch
        from fqtool_b200 import abi
        ctx, eng = self.ctx, self.eng
        paired = self.c["paired"]
        n_host = min(2, self.nb)
        hbs = [(self.batches[i][0].to_host(pinned=True), self.batches[i][1].to_host(pinned=True) if paired else None) for i in range(n_host)]
        pin = lambda nbytes: torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        hout = [(pin(h[0].n * 16), pin(h[0].n * 16) if paired else None, pin(h[0].n * 24) if paired else None) for h in hbs]
        h2d = d2h = units = 0

        def step():
            nonlocal h2d, d2h, units
            h2d = d2h = units = 0
            pending = None
            for i in range(self.nb):
                (h1, h2), (r1, r2, rp) = hbs[i % n_host], hout[i % n_host]
                if paired:
                    t = eng.submit_pe(h1, h2, (r1.numpy().view(abi.READ_RESULT_DTYPE), r2.numpy().view(abi.READ_RESULT_DTYPE),
                                               rp.numpy().view(abi.PAIR_RESULT_DTYPE)))   # two pair packs in flight
                    h2d += h1.nbytes + h2.nbytes; d2h += h1.n * (16 + 16 + 24)
                else:
                    t = eng.submit_se(h1, r1.numpy().view(abi.READ_RESULT_DTYPE))     # two packs in flight
                    h2d += h1.nbytes; d2h += h1.n * 16
                if pending is not None:
                    eng.wait(pending)
                pending = t
                units += h1.n
            if pending is not None:
                eng.wait(pending)
            return eng.finalize()
        eng.reset()
        step()
        eng.reset()
        times = []
        for _ in range(max(steps, 1)):
            ctx.barrier()
            t0 = time.perf_counter()
            flat = step()
            ctx.barrier()
            times.append(ctx.max_over_ranks(time.perf_counter() - t0))
        med = sorted(times)[len(times) // 2]
        bases = units * self.c["L"] * (2 if paired else 1) * ctx.world
        return {"value": bases / med / 1e9, "unit": "Gbases/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h + flat.nbytes,
                "steps": len(times), "ms_per_step": med * 1e3, "h2d_gbs_per_gpu": h2d / med / 1e9,
                "note": f"{self.nb} {'fqk_submit_pe' if paired else 'fqk_submit_se'} / fqk_wait calls (two packs in flight) + one "
                        f"fqk_finalize per step, {n_host} distinct pinned host batches (fq_batch layout)"}

    def close(self):
        self.batches = self.outs = None
        self.eng.close()


class Ctx:
    """Rank plumbing: torch.distributed for barriers and the ncclUniqueId hand-over; one NCCL communicator per engine
    inside the library."""

    def __init__(self, a):
        import torch
        import torch.distributed as dist
        self.a, self.torch, self.dist = a, torch, dist
        self.rank = int(os.environ.get("RANK", "0")); self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.dev)
        peaks = {}
        pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(pk):
            peaks = json.load(open(pk))
        self.peak = peaks.get("hbm_gbs", 6650.0)
        self.peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        tr = os.path.join(ROOT, "profiles", "traffic.json")
        self.traffic = json.load(open(tr)) if os.path.exists(tr) else {}

    def engine(self, params, single=False):
        from fqtool_b200.engine import Engine
        torch, dist = self.torch, self.dist
        if self.world == 1 or single:
            return Engine(params, device=self.local)
        t = torch.zeros(128, dtype=torch.uint8, device=self.dev)
        if self.rank == 0:
            t.copy_(torch.frombuffer(bytearray(Engine.nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(t, 0)
        return Engine(params, device=self.local, rank=self.rank, nranks=self.world, nccl_unique_id=bytes(t.cpu().numpy().tobytes()))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def _red(self, x, op):
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max_over_ranks(self, x):
        return self._red(x, self.dist.ReduceOp.MAX)

    def sum_over_ranks(self, x):
        return int(round(self._red(float(x), self.dist.ReduceOp.SUM)))

    def free(self):
        import gc
        gc.collect()
        self.torch.cuda.empty_cache()


def merge_check(ctx):
    """N > 1: the NCCL merge inside fqk_finalize (one sum all-reduce of the flat buffer; min/sum/min/min protocol for the
    duplication table, Duplicate::addRecord src/duplicate.cpp:46-62, Stats::merge src/stats.cpp:815-863) against the same
    stream run on ONE rank: flat accumulators and mDups/mCounts/mGC bit-exact on every rank."""
    import numpy as np
    from fqtool_b200 import workload
    torch, dist = ctx.torch, ctx.dist
    out = {}
    for cfg, total in (("cfg4", 300_000), ("pe_dup", 150_001)):
        c = workload.CONFIGS[cfg]
        params = workload.options_for(cfg).to_params()
        lo = total * ctx.rank // ctx.world; hi = total * (ctx.rank + 1) // ctx.world

        def run(eng, a, b):
            b1, b2 = workload.generate(cfg, b - a, a, ctx.dev, eng.stream)
            o1 = torch.empty((b - a) * 16, dtype=torch.uint8, device=ctx.dev)
            if b2 is None:
                eng.process_se_device(b1.c_struct(), o1.data_ptr())
            else:
                o2 = torch.empty((b - a) * 16, dtype=torch.uint8, device=ctx.dev)
                op = torch.empty((b - a) * 24, dtype=torch.uint8, device=ctx.dev)
                eng.process_pe_device(b1.c_struct(), b2.c_struct(), o1.data_ptr(), o2.data_ptr(), op.data_ptr())
            eng.sync()
            flat = eng.finalize()
            return flat, eng.dup_table()
        eng = ctx.engine(params)
        flat, dup = run(eng, lo, hi)
        flat_again = eng.finalize()                 # the merge leaves the local accumulators and the local table intact
        dup_again = eng.dup_table()
        eng.close()
        ok = np.array_equal(flat, flat_again) and all(np.array_equal(x, y) for x, y in zip(dup, dup_again))
        if ctx.rank == 0:
            one = ctx.engine(params, single=True)
            want, wdup = run(one, 0, total)
            one.close()
            ref = [torch.from_numpy(want.view(np.int64)).to(ctx.dev)] + [torch.from_numpy(x.copy()).to(ctx.dev) for x in
                   (wdup[0].view(np.int64), wdup[1].view(np.int32), wdup[2])]
        else:
            ref = [torch.empty(len(flat), dtype=torch.int64, device=ctx.dev), torch.empty(len(dup[0]), dtype=torch.int64, device=ctx.dev),
                   torch.empty(len(dup[1]), dtype=torch.int32, device=ctx.dev), torch.empty(len(dup[2]), dtype=torch.uint8, device=ctx.dev)]
        for t in ref:
            dist.broadcast(t, 0)
        mine = [flat.view(np.int64), dup[0].view(np.int64), dup[1].view(np.int32), dup[2]]
        ok = ok and all(np.array_equal(t.cpu().numpy(), m) for t, m in zip(ref, mine))
        bad = ctx.sum_over_ranks(0 if ok else 1)
        assert bad == 0, f"merge_check {cfg}: {bad} rank(s) differ from the single-rank run"
        out[cfg] = {"units": total, "dup_keys_nonzero": int((dup[1] != 0).sum()), "flat_u64": int(len(flat))}
    return {"status": "ok", "streams": out,
            "what": "merged flat accumulators + duplication table of N contiguous shards == one-rank run, bit-exact on every rank; second fqk_finalize idempotent"}


def main():
    a = parse()
    if a.impl == "reference":
        return run_reference_arm(a)
    if a.impl == "sample":
        return write_sample(a)
    import numpy as np
    from fqtool_b200 import abi, workload
    ctx = Ctx(a)
    torch, dist, dev, rank, world = ctx.torch, ctx.dist, ctx.dev, ctx.rank, ctx.world
    cfg = a.config
    c = workload.CONFIGS[cfg]

    # ---- workload: per-GPU share, resident in HBM (weak scaling: fixed per GPU) ----
    per_unit = per_unit_bytes(c)
    free_b, _ = torch.cuda.mem_get_info(dev)
    cap = int(free_b * 0.80 / per_unit)
    reads = max(1, a.reads or min(c["reads"], cap))
    job = Job(ctx, cfg, reads, rank * reads)
    eng, nb, batches = job.eng, job.nb, job.batches
    sampler = ClockSampler(ctx.local)
    main_rec = job.timed(a.steps, a.warmup, sampler)
    clocks = sampler.stop()

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
                t = eng.submit_pe(h1, h2, (r1.numpy().view(abi.READ_RESULT_DTYPE), r2.numpy().view(abi.READ_RESULT_DTYPE),
                                           rp.numpy().view(abi.PAIR_RESULT_DTYPE)))
                if pending is not None:
                    eng.wait(pending)
                pending = t
                h2d += h1.nbytes + h2.nbytes; d2h += h1.n * (16 + 16 + 24)
            e2e_units += h1.n
        if pending is not None:
            eng.wait(pending)
        return eng.finalize()                        # the step's metric: accumulators to the host (NCCL-merged at N > 1)
    e2e_step()                                       # warm-up (allocates the staging buffers)
    eng.reset()
    times = []
    for _ in range(max(a.e2e_steps, 1)):
        ctx.barrier()
        t0 = time.perf_counter()
        flat2 = e2e_step()
        ctx.barrier()
        times.append(ctx.max_over_ranks(time.perf_counter() - t0))
    d2h_final = flat2.nbytes
    ts = sorted(times)
    med = ts[len(ts) // 2]
    e2e_bases = e2e_units * c["L"] * (2 if c["paired"] else 1) * world
    e2e = {"value": e2e_bases / med / 1e9, "unit": "Gbases/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h + d2h_final, "steps": len(ts), "ms_per_step": med * 1e3,
           "statistic": "median over steps of the max over ranks", "best": e2e_bases / ts[0] / 1e9, "worst": e2e_bases / ts[-1] / 1e9,
           "ms_per_step_all": [round(t * 1e3, 2) for t in times],
           "note": f"{nb} fqk_submit_se / fqk_submit_pe + fqk_wait calls (two packs in flight) + one fqk_finalize per step over "
                   f"{n_host} distinct pinned host batches of {BATCH_READS} reads (already in the fq_batch layout; host packing is outside)"}

    # ---- what the box allows: the same bytes moved by bare copies, no library and no kernel ----
    # (all ranks at once, like the step: at N > 1 this is the ceiling of N GPUs pulling pinned host memory over PCIe together)
    up = [torch.empty(h[0].nbytes + (h[1].nbytes if h[1] else 0), dtype=torch.uint8, pin_memory=True) for h in hbs]
    d_in = [torch.empty_like(up[0], device=dev) for _ in range(2)]
    d_out = torch.empty(hbs[0][0].n * (56 if c["paired"] else 16), dtype=torch.uint8, device=dev)
    down = [torch.empty(d_out.numel(), dtype=torch.uint8, pin_memory=True) for _ in range(2)]
    s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def copy_step():
        for i in range(nb):
            with torch.cuda.stream(s_up):
                d_in[i & 1].copy_(up[i % n_host], non_blocking=True)
            with torch.cuda.stream(s_dn):
                down[i & 1].copy_(d_out, non_blocking=True)
        s_up.synchronize(); s_dn.synchronize()
    copy_step()
    ctimes = []
    for _ in range(3):
        ctx.barrier()
        t0 = time.perf_counter()
        copy_step()
        ctx.barrier()
        ctimes.append(ctx.max_over_ranks(time.perf_counter() - t0))
    cmed = sorted(ctimes)[1]
    copy_bytes = nb * up[0].numel()
    e2e["copy_ceiling"] = {"value": e2e_bases / cmed / 1e9, "unit": "Gbases/s", "ms_per_step": cmed * 1e3,
                           "h2d_gbs_per_gpu": copy_bytes / cmed / 1e9, "h2d_gbs_e2e_per_gpu": h2d / med / 1e9,
                           "e2e_over_ceiling": (e2e_bases / med) / (e2e_bases / cmed),
                           "note": "the step's H2D and D2H bytes as bare cudaMemcpyAsync between pinned host and device buffers on two "
                                   "streams (upload and download overlapped), all ranks at once, no library call and no kernel: "
                                   "the end-to-end number cannot exceed this on this box"}
    del up, d_in, d_out, down

    # ---- the same through the ASCII entry point (FASTQ characters in, packing on the GPU) ----
    e2e_ascii = e2e_text = e2e_fastq = None
    if not c["paired"]:
        from fqtool_b200 import batch as fbatch
        seq, qual, off = fbatch.unpack(hbs[0][0])
        pseq = torch.empty(len(seq), dtype=torch.uint8, pin_memory=True); pseq.numpy()[:] = seq
        pqual = torch.empty(len(qual), dtype=torch.uint8, pin_memory=True); pqual.numpy()[:] = qual
        areads = (pseq.numpy(), pqual.numpy(), off)
        eng.process_se_ascii(areads)
        aout = [torch.empty(hbs[0][0].n * 16, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
        poff = torch.empty(len(off), dtype=torch.int64, pin_memory=True); poff.numpy()[:] = off.astype(np.int64)
        areads_p = (pseq.numpy(), pqual.numpy(), poff.numpy().view(np.uint64))

        def ascii_step(reps):
            pend = None; keep = []
            for i in range(reps):
                t, k = eng.submit_se_ascii(areads_p, aout[i & 1].numpy().view(abi.READ_RESULT_DTYPE))
                keep.append(k)
                if pend is not None:
                    eng.wait(pend)
                pend = t
            eng.wait(pend)
        ascii_step(2)
        eng.reset(); ctx.barrier()
        t0 = time.perf_counter()
        reps = min(nb, 8)
        ascii_step(reps)
        eng.finalize(); ctx.barrier()
        dta = ctx.max_over_ranks(time.perf_counter() - t0)
        e2e_ascii = {"value": world * reps * hbs[0][0].n * c["L"] / dta / 1e9, "unit": "Gbases/s",
                     "h2d_bytes_per_call": int(2 * len(seq) + 8 * (hbs[0][0].n + 1)), "d2h_bytes_per_call": int(16 * hbs[0][0].n), "calls": reps,
                     "note": "fqk_submit_se_ascii / fqk_wait, two packs in flight: FASTQ characters over PCIe (2 B/base); per-read index, "
                             "prefix sum and 2-bit packing on the GPU, nothing per read on the host"}
        # ---- and with the FASTQ records assembled on the device (text in, text out) ----
        import ctypes as C
        n1 = hbs[0][0].n
        digits = (np.arange(n1, dtype=np.int64)[:, None] // (10 ** np.arange(9, -1, -1, dtype=np.int64))[None, :] % 10 + 48).astype(np.uint8)
        head, tail = np.frombuffer(b"@SIM:1:FCX:1:", dtype=np.uint8), np.frombuffer(b" 1:N:0:ATCACG", dtype=np.uint8)
        names = np.concatenate([np.broadcast_to(head, (n1, len(head))), digits, np.broadcast_to(tail, (n1, len(tail)))], axis=1).reshape(-1)
        nlen = len(head) + 10 + len(tail)
        pin = lambda a: (lambda t: (t.numpy().__setitem__(slice(None), a), t)[1])(torch.empty(a.shape, dtype=torch.uint8, pin_memory=True))
        pnames = pin(names); pstr = pin(np.full(n1, ord("+"), dtype=np.uint8))
        noff = np.arange(n1 + 1, dtype=np.uint64) * np.uint64(nlen); soff = np.arange(n1 + 1, dtype=np.uint64)
        # every array of the pack pinned: a pageable source makes cudaMemcpyAsync wait for the stream, which serialises the pipeline
        pin64 = lambda a: (lambda t: (t.numpy().__setitem__(slice(None), a.astype(np.int64)), t)[1])(torch.empty(len(a), dtype=torch.int64, pin_memory=True))
        pnoff, psoff = pin64(noff), pin64(soff)
        tb, keep = eng._text((pnames.numpy(), pnoff.numpy().view(np.uint64), pseq.numpy(), pqual.numpy(), poff.numpy().view(np.uint64),
                              pstr.numpy(), psoff.numpy().view(np.uint64)), 0)
        cap = int(len(names) + 2 * len(seq) + n1 * 8)
        tout = torch.empty(cap, dtype=torch.uint8, pin_memory=True)
        touts = (abi.FqTextOut * abi.CH_COUNT)()
        touts[abi.CH_OUT1].data = tout.data_ptr(); touts[abi.CH_OUT1].cap = cap
        passed = C.c_uint64(0)

        def text_call():
            rc = eng.lib.fqk_process_se_text(eng.h, C.byref(tb), 0, None, touts, C.byref(passed))
            assert rc == 0, eng.lib.fqk_last_error(eng.h)
        text_call()
        # two packs in flight: fqk_submit_se_text / fqk_wait_text, the upload of pack k+1 over the assembly + download of pack k
        tout2 = torch.empty(cap, dtype=torch.uint8, pin_memory=True)
        touts2 = (abi.FqTextOut * abi.CH_COUNT)()
        touts2[abi.CH_OUT1].data = tout2.data_ptr(); touts2[abi.CH_OUT1].cap = cap
        slots = (touts, touts2)

        def text_pipe(reps):
            pend = None
            for i in range(reps + 1):
                tk = None
                if i < reps:
                    tk = C.c_uint64(0)
                    rc = eng.lib.fqk_submit_se_text(eng.h, C.byref(tb), 0, None, slots[i & 1], C.byref(tk))
                    assert rc == 0, eng.lib.fqk_last_error(eng.h)
                if pend is not None:
                    rc = eng.lib.fqk_wait_text(eng.h, pend.value, C.byref(passed))
                    assert rc == 0, eng.lib.fqk_last_error(eng.h)
                pend = tk
        text_pipe(2)
        eng.reset(); ctx.barrier()
        t0 = time.perf_counter()
        text_pipe(reps)
        eng.finalize(); ctx.barrier()
        dtt = ctx.max_over_ranks(time.perf_counter() - t0)
        e2e_text = {"value": world * reps * n1 * c["L"] / dtt / 1e9, "unit": "Gbases/s", "calls": reps,
                    "h2d_bytes_per_call": int(2 * len(seq) + len(names) + n1 * (1 + 8 + 8 + 8)),
                    "d2h_bytes_per_call": int(touts[abi.CH_OUT1].len), "records_out": int(touts[abi.CH_OUT1].records),
                    "note": "fqk_submit_se_text / fqk_wait_text, two packs in flight: FASTQ text in (names, bases, qualities), trimmed FASTQ "
                            "text of the passing reads out; packing, trimming, statistics and record assembly on the GPU, no per-read "
                            "work on the host"}
        # ---- and from the decompressed FASTQ TEXT itself (newline scan on the device too) ----
        L1 = c["L"]
        rec_bytes = nlen + 1 + L1 + 1 + 2 + L1 + 1
        traw = torch.empty(n1 * rec_bytes, dtype=torch.uint8, pin_memory=True)
        tv = traw.numpy().reshape(n1, rec_bytes)
        tv[:, :nlen] = names.reshape(n1, nlen); tv[:, nlen] = 10
        tv[:, nlen + 1:nlen + 1 + L1] = seq.reshape(n1, L1); tv[:, nlen + 1 + L1] = 10
        tv[:, nlen + 2 + L1] = ord("+"); tv[:, nlen + 3 + L1] = 10
        tv[:, nlen + 4 + L1:nlen + 4 + 2 * L1] = qual.reshape(n1, L1); tv[:, rec_bytes - 1] = 10
        rb = abi.FqRawBlock(); rb.text = traw.data_ptr(); rb.bytes = traw.numel(); rb.first_index = 0
        nr, used = C.c_uint32(0), C.c_uint64(0)

        def fastq_call():
            rc = eng.lib.fqk_process_se_fastq(eng.h, C.byref(rb), 0, None, 0, touts, C.byref(passed), C.byref(nr), C.byref(used))
            assert rc == 0 and nr.value == n1 and used.value == traw.numel(), eng.lib.fqk_last_error(eng.h)
        fastq_call()

        def fastq_pipe(reps):                    # two blocks in flight: fqk_submit_se_fastq / fqk_wait_fastq
            pend = None
            for i in range(reps + 1):
                tk = None
                if i < reps:
                    tk = C.c_uint64(0)
                    rc = eng.lib.fqk_submit_se_fastq(eng.h, C.byref(rb), C.byref(tk))
                    assert rc == 0, eng.lib.fqk_last_error(eng.h)
                if pend is not None:
                    rc = eng.lib.fqk_wait_fastq(eng.h, pend.value, 0, None, 0, slots[i & 1], C.byref(passed), C.byref(nr), C.byref(used))
                    assert rc == 0 and nr.value == n1 and used.value == traw.numel(), eng.lib.fqk_last_error(eng.h)
                pend = tk
        fastq_pipe(2)
        eng.reset(); ctx.barrier()
        t0 = time.perf_counter()
        fastq_pipe(reps)
        eng.finalize(); ctx.barrier()
        dtf = ctx.max_over_ranks(time.perf_counter() - t0)
        e2e_fastq = {"value": world * reps * n1 * L1 / dtf / 1e9, "unit": "Gbases/s", "calls": reps,
                     "h2d_bytes_per_call": int(traw.numel()), "d2h_bytes_per_call": int(touts[abi.CH_OUT1].len),
                     "note": "fqk_submit_se_fastq / fqk_wait_fastq, two blocks in flight: decompressed FASTQ text in, trimmed FASTQ text out; newline scan, packing, hot path "
                             "and record assembly on the GPU (what FqReader, the worker body and Read::toString do in the reference)"}
        del pseq, pqual, areads, seq, qual, off, pnames, pstr, tout, tout2, traw, tv
    del hbs, hout, batches
    job.close(); del job, eng
    ctx.free()

    # ---- the other BASELINE.json configurations, same protocol, smaller resident workloads ----
    extra, scaling_cfg5, mcheck = {}, None, None
    if not a.no_extra and cfg == "cfg2":
        for xc in ("cfg3", "cfg4"):
            j = Job(ctx, xc, a.extra_units, rank * a.extra_units)
            r = j.timed(a.extra_steps, 3)
            r["scaling"] = "weak"; r["workload"] = workload.CONFIGS[xc]["desc"]
            r["e2e"] = j.e2e(3)
            extra[xc] = r
            j.close(); del j; ctx.free()
        # cfg5: strong scaling of a fixed total, contiguous shards by global pair index
        tot = a.cfg5_total
        free_b, _ = torch.cuda.mem_get_info(dev)
        tot = min(tot, int(free_b * 0.85 / 708) * world // BATCH_READS * BATCH_READS)
        lo, hi = tot * rank // world, tot * (rank + 1) // world
        j = Job(ctx, "cfg5", hi - lo, lo)
        r = j.timed(max(3, a.extra_steps // 2 + 1), 3)
        r["scaling"] = "strong"; r["workload"] = workload.CONFIGS["cfg5"]["desc"]; r["pairs_total"] = tot
        r["e2e"] = j.e2e(3)
        extra["cfg5"] = r
        scaling_cfg5 = {"n_gpus": world, "pairs_total": tot, "pairs_per_gpu": hi - lo, "value": r["value"], "unit": "Gbases/s",
                        "ms_per_step": r["ms_per_step"], "finalize_ms": r["finalize_ms"], "scaling": "strong",
                        "roofline_frac": r["roofline"]["frac"],
                        "note": "fixed total sharded contiguously by global pair index; one NCCL all-reduce of the flat buffer in fqk_finalize"}
        j.close(); del j; ctx.free()
        if world > 1:
            mcheck = merge_check(ctx)

    cpu_baseline = oracle_checks = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        ctx.free()
        cb = cpu_reference_run(cfg, a.cpu_sample, 3, 1, dev)
        cpu_baseline = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        cpu_baseline["protocol"] = "1 warm-up run, mean of 3 timed runs"
        # the checker: what the timed kernels compute on the first units of every timed stream == the CPU oracle
        if a.check_units > 0:
            oracle_checks = {}
            nmain = min(a.check_units, reads)
            oracle_checks[cfg] = oracle_check(ctx, cfg, nmain, cb["reads"] if nmain == a.cpu_sample else None)
            del cb
            for xc in extra:
                oracle_checks[xc] = oracle_check(ctx, xc, min(a.check_units, extra[xc]["units_per_gpu"]))
                ctx.free()

    if rank == 0:
        line = {"metric": "Gbases/s per-read QC+trim", "value": main_rec["value"], "unit": "Gbases/s", "n_gpus": world,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": main_rec["ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": config_dict(a, c, reads), "impl": a.impl,
                "clocks": clocks, "e2e": e2e, "e2e_ascii": e2e_ascii, "e2e_text": e2e_text, "e2e_fastq": e2e_fastq, "gpu_launches": main_rec["gpu_launches"],
                "roofline": main_rec["roofline"], "finalize_ms": main_rec["finalize_ms"], "cpu_baseline": cpu_baseline,
                "configs": extra or None, "scaling_cfg5": scaling_cfg5, "merge_check": mcheck, "oracle_check": oracle_checks}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
