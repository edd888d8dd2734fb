"""Device-resident synthetic workloads (BASELINE.json configs) for bench.py and the GPU tests.

torch is used for device memory only; reads are generated straight into HBM in the
fq_batch layout by the generators of csrc/fqk_synth.cu and processed in place through
the device-pointer entry points of the C ABI.
"""
import ctypes as C

import numpy as np
import torch

from . import abi
from .batch import Batch

SYNTH_SE_HISEQ, SYNTH_PE_NOVASEQ, SYNTH_SE_ADAPTER = 2, 3, 4

# name -> (paired, read length, generator model, seed, fragment mean/sd, reference CLI flags)
ADAPTER1 = "AGATCGGAAGAGCACACGTCTGAACTCCAGTCA"
CONFIGS = {
    "cfg2": dict(paired=False, L=150, model=SYNTH_SE_HISEQ, seed=20261101, frag=(0.0, 0.0), reads=100_000_000,
                 cli="-q --enable_cut_front --enable_cut_tail",
                 desc="synthetic SE 150bp Illumina-like 100M reads: -q + cut_front/cut_tail window 4"),
    "cfg3": dict(paired=True, L=150, model=SYNTH_PE_NOVASEQ, seed=20261102, frag=(300.0, 90.0), reads=200_000_000,
                 cli="-g -a -m", desc="synthetic PE 2x150bp NovaSeq-like 200M pairs: polyG + overlap adapter trim + merge"),
    "cfg4": dict(paired=False, L=100, model=SYNTH_SE_ADAPTER, seed=20261103, frag=(0.0, 0.0), reads=100_000_000,
                 cli=f"-a --adapter_of_read1 {ADAPTER1} -x -d --dup_ana_key_len 12 --kmer --kmer_length 6",
                 desc="synthetic SE 100bp, explicit adapter, polyX, duplication key 12, k-mer 6"),
    "cfg5": dict(paired=True, L=250, model=SYNTH_PE_NOVASEQ, seed=20261104, frag=(420.0, 120.0), reads=1_000_000_000,
                 cli="-q -g -a", desc="synthetic PE 2x250bp 1B pairs sharded across GPUs"),
    # not a BASELINE config: small paired stream with the duplication table and k-mers on, for bench.py's merge check
    "pe_dup": dict(paired=True, L=150, model=SYNTH_PE_NOVASEQ, seed=20261105, frag=(260.0, 80.0), reads=1_000_000,
                   cli="-d --kmer --kmer_length 5 -q -F 2 -T 3 -f 1", desc="synthetic PE 2x150bp, duplication + k-mer 5 (merge check)"),
}


def _declare(lib):
    if getattr(lib, "_synth_declared", False):
        return
    vp, u32, u64 = C.c_void_p, C.c_uint32, C.c_uint64
    lib.fqk_synth_total_bases.argtypes = [u32, C.c_int]; lib.fqk_synth_total_bases.restype = u64
    lib.fqk_synth_generate.argtypes = [vp, C.c_int, u64, u64, u32, C.c_int, C.c_double, C.c_double] + [vp] * 8
    lib.fqk_synth_generate.restype = C.c_int
    lib._synth_declared = True


class DeviceBatch:
    """fq_batch whose four arrays live in HBM (torch tensors own the memory)."""

    def __init__(self, n, L, first_index, device):
        lib = abi.load(); _declare(lib)
        self.n, self.L, self.first_index = n, L, first_index
        self.total = int(lib.fqk_synth_total_bases(n, L))
        self.offsets = torch.empty(max(n, 1), dtype=torch.int32, device=device)
        self.lengths = torch.empty(max(n, 1), dtype=torch.int16, device=device)
        # zero-filled: the generator leaves the (<64 bases) tail padding of the batch untouched
        self.bases2 = torch.zeros(max(self.total // 4, 1), dtype=torch.uint8, device=device)
        self.qual = torch.zeros(max(self.total, 1), dtype=torch.uint8, device=device)

    def c_struct(self):
        b = abi.FqBatch()
        b.n_reads, b.first_index, b.total_bases = self.n, self.first_index, self.total
        b.max_len = self.L
        b.offsets, b.lengths = self.offsets.data_ptr(), self.lengths.data_ptr()
        b.bases2, b.qual = self.bases2.data_ptr(), self.qual.data_ptr()
        return b

    def to_host(self, n=None, pinned=False):
        """Copy the first n reads back as a host Batch (numpy views of torch CPU tensors)."""
        n = self.n if n is None else min(n, self.n)
        Lp = (self.L + 15) // 16 * 16
        tot = (n * Lp + 63) // 64 * 64
        def cp(t, k):
            h = torch.empty(k, dtype=t.dtype, pin_memory=pinned)
            h.copy_(t[:k])
            return h
        off, ln, b2, q = cp(self.offsets, n), cp(self.lengths, n), cp(self.bases2, tot // 4), cp(self.qual, tot)
        hb = Batch(off.numpy().view(np.uint32), ln.numpy().view(np.uint16), b2.numpy(), q.numpy(), self.first_index)
        hb._keep = (off, ln, b2, q)
        return hb

    @property
    def n_bases(self):
        return self.n * self.L

    @property
    def nbytes(self):
        return self.n * 6 + self.total // 4 + self.total


def generate(cfg, n, first_index, device, stream_ptr):
    """Generate reads [first_index, first_index+n) of a named config; returns (batch1, batch2|None)."""
    lib = abi.load(); _declare(lib)
    c = CONFIGS[cfg]
    b1 = DeviceBatch(n, c["L"], first_index, device)
    b2 = DeviceBatch(n, c["L"], first_index, device) if c["paired"] else None
    # the buffers were zero-filled on torch's stream; the generator runs on the caller's (non-blocking) stream
    torch.cuda.current_stream(device).synchronize()
    a2 = [b2.offsets.data_ptr(), b2.lengths.data_ptr(), b2.bases2.data_ptr(), b2.qual.data_ptr()] if b2 else [None] * 4
    rc = lib.fqk_synth_generate(stream_ptr, c["model"], c["seed"], first_index, n, c["L"], c["frag"][0], c["frag"][1],
                                b1.offsets.data_ptr(), b1.lengths.data_ptr(), b1.bases2.data_ptr(), b1.qual.data_ptr(), *a2)
    if rc:
        raise RuntimeError(f"fqk_synth_generate rc={rc}")
    return b1, b2


def options_for(cfg, max_cycles=None):
    from .options import from_cli
    c = CONFIGS[cfg]
    o = from_cli(c["cli"].split() + (["-I", "read2"] if c["paired"] else []))
    if max_cycles is None:
        # merged reads can reach L1 + L2 - overlapRequire cycles (peprocessor.cpp:361)
        need = 2 * c["L"] if (c["paired"] and o.mergePE.enabled) else c["L"]
        max_cycles = (need + 31) // 32 * 32
    o.maxCycles = max_cycles
    return o


def algorithmic_bytes(cfg, units):
    """1.25 B/base + 12 B/read (+ 12 B/pair)."""
    c = CONFIGS[cfg]
    per_read = 1.25 * c["L"] + 12
    return units * (2 * per_read + 12 if c["paired"] else per_read)
