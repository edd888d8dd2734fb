"""Host-side view of the run-level accumulators returned by ``fqk_finalize``.

The reference's post-run reductions stay on the host:
``Stats::summarize`` (/root/reference/src/stats.cpp:147-228),
``Duplicate::statAll`` (/root/reference/src/duplicate.cpp:131-166) and the
insert-size peak (/root/reference/src/peprocessor.cpp:248-259) consume the raw
uint64 arrays unchanged.
"""
import numpy as np

STATS_NAMES = ("pre1", "pre2", "post1", "post2")
BIN = {"A": ord("A") & 7, "T": ord("T") & 7, "C": ord("C") & 7, "G": ord("G") & 7, "N": ord("N") & 7}


class StatsView:
    """One Stats object: raw arrays in the reference's indexing [8][C]."""

    def __init__(self, flat, L, which):
        C = int(L.max_cycles)
        s = flat[int(L.stats_off[which]):int(L.stats_off[which]) + int(L.stats_stride)]
        self.C = C
        self.q20 = s[L.s_q20:L.s_q20 + 8 * C].reshape(8, C)
        self.q30 = s[L.s_q30:L.s_q30 + 8 * C].reshape(8, C)
        self.content = s[L.s_content:L.s_content + 8 * C].reshape(8, C)
        self.qualsum = s[L.s_qualsum:L.s_qualsum + 8 * C].reshape(8, C)
        self.total_base = s[L.s_total_base:L.s_total_base + C]
        self.total_qual = s[L.s_total_qual:L.s_total_qual + C]
        self.reads = int(s[L.s_reads])
        self.length_sum = int(s[L.s_length_sum])
        self.kmer = s[L.s_kmer:L.s_kmer + int(L.kmer_bins)]

    def summarize(self):
        """Stats::summarize: cycles = first cycle with no base; totals over those cycles."""
        tb = self.total_base
        zero = np.nonzero(tb == 0)[0]
        cycles = int(zero[0]) if zero.size else self.C
        bases = int(tb[:cycles].sum())
        q20 = int(self.q20[:, :cycles].sum()); q30 = int(self.q30[:, :cycles].sum())
        gc = int(self.content[BIN["G"], :cycles].sum() + self.content[BIN["C"], :cycles].sum())
        with np.errstate(divide="ignore", invalid="ignore"):
            mean_qual = self.total_qual[:cycles] / tb[:cycles]
        out = {"reads": self.reads, "bases": bases, "q20": q20, "q30": q30, "gc": gc, "cycles": cycles,
               "mean_length": self.length_sum // self.reads if self.reads else 0, "mean_qual_curve": mean_qual}
        if self.kmer.size:
            out["kmer_min"], out["kmer_max"] = int(self.kmer.min()), int(self.kmer.max())
        return out


class RunStats:
    def __init__(self, flat, layout):
        self.flat, self.L = flat, layout
        L = layout
        self.stats = {n: StatsView(flat, L, i) for i, n in enumerate(STATS_NAMES)}
        self.filter = flat[L.filter_off:L.filter_off + 32]
        self.adapter_reads = int(flat[L.adapter_reads_off]); self.adapter_bases = int(flat[L.adapter_bases_off])
        self.polyx_reads = flat[L.polyx_reads_off:L.polyx_reads_off + 5]
        self.polyx_bases = flat[L.polyx_bases_off:L.polyx_bases_off + 5]
        self.merged_pairs = int(flat[L.merged_pairs_off])
        self.insert_hist = flat[L.insert_hist_off:L.insert_hist_off + L.insert_hist_len]

    def insert_peak(self):
        """PairEndProcessor::getPeakInsertSize: first arg-max below insertSizeMax."""
        h = self.insert_hist[:-1]
        return int(np.argmax(h)) if h.size and h.max() > 0 else 0


def dup_stat_all(counts, gc, hist_size=32):
    """Duplicate::statAll, src/duplicate.cpp:131-166: (hist, meanGC, rate)."""
    counts = counts.astype(np.int64)
    nz = counts > 0
    total = int(counts[nz].sum()); dupn = int((counts[nz] - 1).sum())
    slot = np.where(counts > hist_size, hist_size - 1, counts)[nz]
    hist = np.bincount(slot, minlength=hist_size).astype(np.uint64)[:hist_size]
    gsum = np.bincount(slot, weights=gc[nz].astype(np.float64), minlength=hist_size)[:hist_size]
    mean_gc = np.where(hist > 0, gsum / 255.0 / np.maximum(hist, 1), gsum)
    return hist, mean_gc, (dupn / total if total else 0.0)
