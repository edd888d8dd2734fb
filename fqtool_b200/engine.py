"""One GPU context behind the C ABI (include/fqk.h): the object a host thread owns.

``Engine`` is the Python face of ``fqk_ctx``: process packs of reads (host
buffers, H2D/D2H inside the call), accumulate run-level statistics on the
device, ``finalize()`` once per run (NCCL all-reduce across ranks).
"""
import ctypes as C

import numpy as np

from . import abi


class FqkError(RuntimeError):
    pass


class Engine:
    def __init__(self, params, device=0, rank=0, nranks=1, nccl_unique_id=None):
        self.lib = abi.load()
        self.params = params
        self.h = C.c_void_p()
        uid = None
        if nccl_unique_id is not None:
            self._uid = (C.c_char * 128).from_buffer_copy(bytes(nccl_unique_id))
            uid = C.cast(self._uid, C.c_void_p)
        rc = self.lib.fqk_create(C.byref(self.h), C.byref(params), device, rank, nranks, uid)
        if rc:
            self.h = None
            raise FqkError(f"fqk_create rc={rc}: {self.lib.fqk_last_error(None).decode()}")
        self.layout = abi.FqStatsLayout()
        self.lib.fqk_stats_layout(C.byref(params), C.byref(self.layout))

    @staticmethod
    def nccl_unique_id():
        lib = abi.load()
        buf = (C.c_char * 128)()
        rc = lib.fqk_nccl_unique_id(C.cast(buf, C.c_void_p))
        if rc:
            raise FqkError(f"fqk_nccl_unique_id rc={rc}")
        return bytes(buf)

    def _check(self, rc, what):
        if rc:
            raise FqkError(f"{what} rc={rc}: {self.lib.fqk_last_error(self.h).decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.lib.fqk_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    # ---- host-buffer path (the drop-in call) ----
    def process_se(self, batch, out=None):
        if out is None:
            out = np.zeros(batch.n, dtype=abi.READ_RESULT_DTYPE)
        b = batch.c_struct()
        self._check(self.lib.fqk_process_se(self.h, C.byref(b), abi.ptr(out)), "fqk_process_se")
        return out

    @staticmethod
    def _ascii(reads, first_index):
        seq, qual = np.ascontiguousarray(reads[0]), np.ascontiguousarray(reads[1])
        off = np.ascontiguousarray(reads[2], dtype=np.uint64)      # converted ONCE; the caller keeps this array alive
        b = abi.FqAsciiBatch()
        b.n_reads = len(off) - 1; b.first_index = first_index
        b.off, b.seq, b.qual = abi.ptr(off), abi.ptr(seq), abi.ptr(qual)
        return b, (seq, qual, off)

    def process_se_ascii(self, reads, first_index=0):
        """ASCII (seq, qual, off) straight to the GPU: packing runs on the device (fqk_process_se_ascii)."""
        b, keep = self._ascii(reads, first_index)
        out = np.zeros(b.n_reads, dtype=abi.READ_RESULT_DTYPE)
        self._check(self.lib.fqk_process_se_ascii(self.h, C.byref(b), abi.ptr(out)), "fqk_process_se_ascii")
        return out

    def process_pe_ascii(self, reads1, reads2, first_index=0):
        b1, k1 = self._ascii(reads1, first_index); b2, k2 = self._ascii(reads2, first_index)
        n = b1.n_reads
        out = (np.zeros(n, dtype=abi.READ_RESULT_DTYPE), np.zeros(n, dtype=abi.READ_RESULT_DTYPE),
               np.zeros(n, dtype=abi.PAIR_RESULT_DTYPE))
        self._check(self.lib.fqk_process_pe_ascii(self.h, C.byref(b1), C.byref(b2), abi.ptr(out[0]), abi.ptr(out[1]),
                                                  abi.ptr(out[2])), "fqk_process_pe_ascii")
        return out

    # ---- text entry points: FASTQ records assembled on the device (fqk_process_*_text) ----
    @staticmethod
    def _text(rec, first_index):
        """rec = (names, name_off, seq, qual, off, strands, strand_off) as uint8 / uint64 numpy arrays."""
        names, noff, seq, qual, off, strands, soff = rec
        keep = [np.ascontiguousarray(names), np.ascontiguousarray(noff, dtype=np.uint64), np.ascontiguousarray(seq),
                np.ascontiguousarray(qual), np.ascontiguousarray(off, dtype=np.uint64), np.ascontiguousarray(strands),
                np.ascontiguousarray(soff, dtype=np.uint64)]
        b = abi.FqTextBatch()
        b.reads.n_reads = len(keep[4]) - 1; b.reads.first_index = first_index
        b.reads.off, b.reads.seq, b.reads.qual = abi.ptr(keep[4]), abi.ptr(keep[2]), abi.ptr(keep[3])
        b.name_off, b.names, b.strand_off, b.strands = abi.ptr(keep[1]), abi.ptr(keep[0]), abi.ptr(keep[6]), abi.ptr(keep[5])
        return b, keep

    def _text_call(self, fn, args_before, nbytes_hint, writers, recs):
        bufs = [np.zeros(max(nbytes_hint, 16), dtype=np.uint8) for _ in range(abi.CH_COUNT)]
        outs = (abi.FqTextOut * abi.CH_COUNT)()
        for ch in range(abi.CH_COUNT):
            outs[ch].data = abi.ptr(bufs[ch]); outs[ch].cap = bufs[ch].nbytes
        passed = C.c_uint64(0)
        rc = fn(self.h, *args_before, writers, *recs, outs, C.byref(passed))
        self._check(rc, "fqk_process_*_text")
        return ([bytes(bufs[ch][:outs[ch].len]) for ch in range(abi.CH_COUNT)], [int(outs[ch].records) for ch in range(abi.CH_COUNT)],
                int(passed.value))

    def process_se_text(self, rec, writers=0, first_index=0):
        """-> (text block per channel, records per channel, readPassed, records array)."""
        b, keep = self._text(rec, first_index)
        out = np.zeros(b.reads.n_reads, dtype=abi.READ_RESULT_DTYPE)
        hint = int(len(keep[0]) + 2 * len(keep[2]) + len(keep[5]) + 40 * b.reads.n_reads + 64)
        blocks, nrec, passed = self._text_call(self.lib.fqk_process_se_text, (C.byref(b),), hint, writers, (abi.ptr(out),))
        return blocks, nrec, passed, out

    def process_pe_text(self, rec1, rec2, writers=0, first_index=0):
        b1, k1 = self._text(rec1, first_index); b2, k2 = self._text(rec2, first_index)
        n = b1.reads.n_reads
        out = (np.zeros(n, dtype=abi.READ_RESULT_DTYPE), np.zeros(n, dtype=abi.READ_RESULT_DTYPE), np.zeros(n, dtype=abi.PAIR_RESULT_DTYPE))
        hint = int(sum(len(k[0]) + 2 * len(k[2]) + len(k[5]) for k in (k1, k2)) + 80 * n + 64)
        blocks, nrec, passed = self._text_call(self.lib.fqk_process_pe_text, (C.byref(b1), C.byref(b2)), hint, writers,
                                               (abi.ptr(out[0]), abi.ptr(out[1]), abi.ptr(out[2])))
        return blocks, nrec, passed, out

    def submit_se_text(self, rec, writers=0, first_index=0):
        """Pipelined fqk_process_se_text; returns a handle for wait_text (it keeps every buffer of the pack alive)."""
        b, keep = self._text(rec, first_index)
        hint = int(len(keep[0]) + 2 * len(keep[2]) + len(keep[5]) + 40 * b.reads.n_reads + 64)
        bufs, outs = self._fastq_outs(hint)
        out = np.zeros(b.reads.n_reads, dtype=abi.READ_RESULT_DTYPE)
        t = C.c_uint64(0)
        self._check(self.lib.fqk_submit_se_text(self.h, C.byref(b), writers, abi.ptr(out), outs, C.byref(t)), "fqk_submit_se_text")
        return {"ticket": t.value, "keep": (b, keep), "bufs": bufs, "outs": outs, "records": out}

    def wait_text(self, handle):
        passed = C.c_uint64(0)
        self._check(self.lib.fqk_wait_text(self.h, handle["ticket"], C.byref(passed)), "fqk_wait_text")
        outs, bufs = handle["outs"], handle["bufs"]
        return ([bytes(bufs[ch][:outs[ch].len]) for ch in range(abi.CH_COUNT)], [int(outs[ch].records) for ch in range(abi.CH_COUNT)],
                int(passed.value), handle["records"])

    # ---- raw FASTQ blocks: decompressed text in, FASTQ text out (fqk_process_*_fastq) ----
    def _fastq_outs(self, nbytes):
        bufs = [np.zeros(max(nbytes, 16), dtype=np.uint8) for _ in range(abi.CH_COUNT)]
        outs = (abi.FqTextOut * abi.CH_COUNT)()
        for ch in range(abi.CH_COUNT):
            outs[ch].data = abi.ptr(bufs[ch]); outs[ch].cap = bufs[ch].nbytes
        return bufs, outs

    def process_se_fastq(self, text, writers=0, first_index=0):
        """text: bytes / uint8 array of a decompressed FASTQ block -> (blocks per channel, reads, bytes consumed, readPassed)."""
        t = np.frombuffer(bytes(text), dtype=np.uint8) if not isinstance(text, np.ndarray) else text
        b = abi.FqRawBlock(); b.text = abi.ptr(t) if len(t) else None; b.bytes = len(t); b.first_index = first_index
        bufs, outs = self._fastq_outs(len(t) + 64 * (len(t) // 8 + 1))
        passed, n, used = C.c_uint64(0), C.c_uint32(0), C.c_uint64(0)
        self._check(self.lib.fqk_process_se_fastq(self.h, C.byref(b), writers, None, 0, outs, C.byref(passed), C.byref(n), C.byref(used)),
                    "fqk_process_se_fastq")
        return [bytes(bufs[ch][:outs[ch].len]) for ch in range(abi.CH_COUNT)], n.value, used.value, int(passed.value)

    def process_pe_fastq(self, text1, text2, writers=0, first_index=0):
        t1 = np.frombuffer(bytes(text1), dtype=np.uint8); t2 = np.frombuffer(bytes(text2), dtype=np.uint8)
        b1 = abi.FqRawBlock(); b1.text = abi.ptr(t1) if len(t1) else None; b1.bytes = len(t1); b1.first_index = first_index
        b2 = abi.FqRawBlock(); b2.text = abi.ptr(t2) if len(t2) else None; b2.bytes = len(t2); b2.first_index = first_index
        bufs, outs = self._fastq_outs(len(t1) + len(t2) + 64 * ((len(t1) + len(t2)) // 8 + 1))
        passed, n, u1, u2 = C.c_uint64(0), C.c_uint32(0), C.c_uint64(0), C.c_uint64(0)
        self._check(self.lib.fqk_process_pe_fastq(self.h, C.byref(b1), C.byref(b2), writers, None, None, None, 0, outs, C.byref(passed),
                                                  C.byref(n), C.byref(u1), C.byref(u2)), "fqk_process_pe_fastq")
        return [bytes(bufs[ch][:outs[ch].len]) for ch in range(abi.CH_COUNT)], n.value, (u1.value, u2.value), int(passed.value)

    def submit_se_fastq(self, text, first_index=0):
        """Pipelined fqk_process_se_fastq, first half (upload + newline count); returns a handle for wait_fastq."""
        t = np.frombuffer(bytes(text), dtype=np.uint8) if not isinstance(text, np.ndarray) else text
        b = abi.FqRawBlock(); b.text = abi.ptr(t) if len(t) else None; b.bytes = len(t); b.first_index = first_index
        tk = C.c_uint64(0)
        self._check(self.lib.fqk_submit_se_fastq(self.h, C.byref(b), C.byref(tk)), "fqk_submit_se_fastq")
        return {"ticket": tk.value, "keep": (t, b)}

    def wait_fastq(self, handle, writers=0):
        nbytes = len(handle["keep"][0])
        bufs, outs = self._fastq_outs(nbytes + 64 * (nbytes // 8 + 1))
        passed, n, used = C.c_uint64(0), C.c_uint32(0), C.c_uint64(0)
        self._check(self.lib.fqk_wait_fastq(self.h, handle["ticket"], writers, None, 0, outs, C.byref(passed), C.byref(n), C.byref(used)),
                    "fqk_wait_fastq")
        return [bytes(bufs[ch][:outs[ch].len]) for ch in range(abi.CH_COUNT)], n.value, used.value, int(passed.value)

    def submit_se(self, batch, out):
        """Pipelined fqk_process_se: returns a ticket; `batch` and `out` must stay alive until wait(ticket)."""
        b = batch.c_struct()
        t = C.c_uint64(0)
        self._check(self.lib.fqk_submit_se(self.h, C.byref(b), abi.ptr(out), C.byref(t)), "fqk_submit_se")
        return t.value

    def submit_pe(self, batch1, batch2, out):
        """Pipelined fqk_process_pe: returns a ticket; the batches and out = (records1, records2, pairs) must stay alive
        until wait(ticket)."""
        b1, b2 = batch1.c_struct(), batch2.c_struct()
        t = C.c_uint64(0)
        self._check(self.lib.fqk_submit_pe(self.h, C.byref(b1), C.byref(b2), abi.ptr(out[0]), abi.ptr(out[1]), abi.ptr(out[2]),
                                           C.byref(t)), "fqk_submit_pe")
        return t.value

    def submit_se_ascii(self, reads, out, first_index=0):
        """Pipelined fqk_process_se_ascii; returns (ticket, keep): `keep` and `out` must stay alive until wait(ticket)."""
        b, keep = self._ascii(reads, first_index)
        t = C.c_uint64(0)
        self._check(self.lib.fqk_submit_se_ascii(self.h, C.byref(b), abi.ptr(out), C.byref(t)), "fqk_submit_se_ascii")
        return t.value, keep

    def wait(self, ticket):
        self._check(self.lib.fqk_wait(self.h, ticket), "fqk_wait")

    def process_pe(self, batch1, batch2, out=None):
        n = batch1.n
        if out is None:
            out = (np.zeros(n, dtype=abi.READ_RESULT_DTYPE), np.zeros(n, dtype=abi.READ_RESULT_DTYPE),
                   np.zeros(n, dtype=abi.PAIR_RESULT_DTYPE))
        b1, b2 = batch1.c_struct(), batch2.c_struct()
        self._check(self.lib.fqk_process_pe(self.h, C.byref(b1), C.byref(b2), abi.ptr(out[0]), abi.ptr(out[1]),
                                            abi.ptr(out[2])), "fqk_process_pe")
        return out

    # ---- device-resident path (pointers are raw device addresses) ----
    def process_se_device(self, dev_batch, out_ptr):
        self._check(self.lib.fqk_process_se_device(self.h, C.byref(dev_batch), out_ptr), "fqk_process_se_device")

    def process_pe_device(self, dev_batch1, dev_batch2, out1_ptr, out2_ptr, outp_ptr):
        self._check(self.lib.fqk_process_pe_device(self.h, C.byref(dev_batch1), C.byref(dev_batch2), out1_ptr, out2_ptr,
                                                   outp_ptr), "fqk_process_pe_device")

    def corrections(self):
        """Bases rewritten by -c in the last process_pe call, sorted (pair, mate, pos)."""
        n = C.c_uint32(0)
        self._check(self.lib.fqk_get_corrections(self.h, None, 0, C.byref(n)), "fqk_get_corrections")
        out = np.zeros(n.value, dtype=abi.CORRECTION_DTYPE)
        if n.value:
            self._check(self.lib.fqk_get_corrections(self.h, abi.ptr(out), n.value, C.byref(n)), "fqk_get_corrections")
        return np.sort(out, order=["pair", "mate_base", "pos"])

    def sync(self):
        self._check(self.lib.fqk_sync(self.h), "fqk_sync")

    @property
    def stream(self):
        return self.lib.fqk_stream(self.h)

    @property
    def launch_count(self):
        return int(self.lib.fqk_launch_count(self.h))

    def set_impl(self, impl):
        self._check(self.lib.fqk_set_impl(self.h, impl), "fqk_set_impl")

    def set_kernel_timing(self, on):
        self._check(self.lib.fqk_set_kernel_timing(self.h, int(on)), "fqk_set_kernel_timing")

    def kernel_times(self):
        ms = (C.c_double * abi.T_SLOTS)()
        n = (C.c_uint64 * abi.T_SLOTS)()
        self._check(self.lib.fqk_kernel_times(self.h, ms, n), "fqk_kernel_times")
        return list(ms), list(n)

    def reset(self):
        self._check(self.lib.fqk_reset(self.h), "fqk_reset")

    def finalize(self):
        flat = np.zeros(self.layout.total_u64, dtype=np.uint64)
        self._check(self.lib.fqk_finalize(self.h, abi.ptr(flat)), "fqk_finalize")
        return flat

    def dup_table(self):
        n = self.layout.dup_bins
        dups = np.zeros(n, dtype=np.uint64); counts = np.zeros(n, dtype=np.uint32); gc = np.zeros(n, dtype=np.uint8)
        self._check(self.lib.fqk_dup_table(self.h, abi.ptr(dups), abi.ptr(counts), abi.ptr(gc)), "fqk_dup_table")
        return dups, counts, gc
