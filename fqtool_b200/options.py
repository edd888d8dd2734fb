"""Host-side mirror of the hot-path part of the reference's ``Options`` tree.

Same field names, defaults and derived values as
/root/reference/src/options.h:14-386 and ``Options::update``
(/root/reference/src/options.cpp:24-58), including the quirks SURVEY.md
section 5 lists: CLI11's ``add_flag(bool&)`` first zeroes the bool
(src/CLI.hpp:2760), so ``qualFilter.enabled``, ``duplicate.enabled`` and
``adapter.enableTriming`` are *false* unless their flag is given, and
``lowQualityBaseLimit`` is ``int(lowQualityRatio * 151)`` whatever the read
length (update() runs before evaluateReadLen(), src/main.cpp:124-129).
"""
from types import SimpleNamespace as NS

from . import abi


class Options:
    def __init__(self):
        self.in2 = ""
        self.interleavedInput = False
        self.insertSizeMax = 512          # options.cpp:17
        self.overlapDiffLimit = 5         # options.cpp:18
        self.overlapRequire = 30          # options.cpp:19
        self.trim = NS(front1=0, tail1=0, maxLen1=0, front2=0, tail2=0, maxLen2=0)
        self.qualitycut = NS(enableFront=False, enableTail=False, enableRright=False,
                             qualityShared=20, windowSizeShared=4,
                             qualityFront=20, qualityTail=20, qualityRight=20,
                             windowSizeFront=4, windowSizeTail=4, windowSizeRight=4)
        self.polyGTrim = NS(enabled=False, minLen=10, maxMismatch=1, allowedOneMismatchForEach=10)
        self.polyXTrim = NS(enabled=False, trimChr="ATCGN", minLen=10, maxMismatch=1,
                            allowedOneMismatchForEach=10)
        # the three defaults below are `true` in options.h but CLI11 resets them (see module doc)
        self.qualFilter = NS(enabled=False, lowQualityLimit=20, lowQualityBaseLimit=40, nBaseLimit=5,
                             lowQualityRatio=0.15, averageQualityLimit=0.0)
        self.duplicate = NS(enabled=False, keylen=12, histSize=32)
        self.adapter = NS(enableTriming=False, inputAdapterSeqR1="", inputAdapterSeqR2="",
                          adapterSeqR1Provided=False, adapterSeqR2Provided=False)
        self.lengthFilter = NS(enabled=False, minReadLength=15, maxReadLength=0)
        self.complexityFilter = NS(enabled=False, threshold=0.3)
        self.correction = NS(enabled=False)
        self.mergePE = NS(enabled=False, discardUnmerged=False)
        self.kmer = NS(enabled=False, kmerLen=0)
        self.est = NS(seqLen1=151, seqLen2=151)
        self.maxCycles = 512              # capacity of per-cycle arrays on the device
        self._updated = False

    def isPaired(self):                   # options.cpp:73-75
        return len(self.in2) > 0 or self.interleavedInput

    def update(self):
        """Options::update, options.cpp:24-58 (hot-path lines only)."""
        if self._updated:
            return self
        self.qualFilter.lowQualityLimit += 33
        self.adapter.adapterSeqR1Provided = bool(self.adapter.inputAdapterSeqR1)
        self.adapter.adapterSeqR2Provided = bool(self.adapter.inputAdapterSeqR2)
        self.qualFilter.lowQualityBaseLimit = int(self.qualFilter.lowQualityRatio * self.est.seqLen1)
        self.polyXTrim.trimChr = self.polyXTrim.trimChr.upper()
        if set(self.polyXTrim.trimChr) - set("ATCGN"):      # Options::validate, options.cpp:67-70
            raise ValueError("Can only trim nucleotides ATCGN")
        self._updated = True
        return self

    def to_params(self):
        """Flat POD for the C ABI (fq_params, include/fqk.h)."""
        self.update()
        p = abi.FqParams()
        p.abi_version = abi.FQK_ABI_VERSION
        p.is_paired = int(self.isPaired())
        t = self.trim
        p.trim_front1, p.trim_tail1, p.trim_max_len1 = t.front1, t.tail1, t.maxLen1
        p.trim_front2, p.trim_tail2, p.trim_max_len2 = t.front2, t.tail2, t.maxLen2
        q = self.qualitycut
        p.cut_front_enabled, p.cut_tail_enabled, p.cut_right_enabled = int(q.enableFront), int(q.enableTail), int(q.enableRright)
        p.cut_front_window, p.cut_tail_window, p.cut_right_window = q.windowSizeFront, q.windowSizeTail, q.windowSizeRight
        p.cut_front_qual, p.cut_tail_qual, p.cut_right_qual = q.qualityFront, q.qualityTail, q.qualityRight
        g = self.polyGTrim
        p.polyg_enabled, p.polyg_min_len = int(g.enabled), g.minLen
        p.polyg_max_mismatch, p.polyg_one_mismatch_each = g.maxMismatch, g.allowedOneMismatchForEach
        x = self.polyXTrim
        p.polyx_enabled = int(x.enabled)
        p.polyx_base_mask = sum(1 << i for i, b in enumerate("ATCGN") if b in x.trimChr)
        p.polyx_min_len, p.polyx_max_mismatch, p.polyx_one_mismatch_each = x.minLen, x.maxMismatch, x.allowedOneMismatchForEach
        f = self.qualFilter
        p.qual_filter_enabled, p.low_quality_limit = int(f.enabled), f.lowQualityLimit
        p.low_quality_base_limit, p.n_base_limit = f.lowQualityBaseLimit, f.nBaseLimit
        p.average_quality_limit = float(f.averageQualityLimit)
        l = self.lengthFilter
        p.length_filter_enabled, p.min_read_length, p.max_read_length = int(l.enabled), l.minReadLength, l.maxReadLength
        p.complexity_enabled, p.complexity_threshold = int(self.complexityFilter.enabled), float(self.complexityFilter.threshold)
        a = self.adapter
        p.adapter_trim_enabled = int(a.enableTriming)
        a1, a2 = a.inputAdapterSeqR1.encode(), a.inputAdapterSeqR2.encode()
        if len(a1) > abi.FQK_MAX_ADAPTER or len(a2) > abi.FQK_MAX_ADAPTER:
            raise ValueError("adapter longer than FQK_MAX_ADAPTER")
        p.adapter_r1_len, p.adapter_r2_len = len(a1), len(a2)
        p.adapter_r1, p.adapter_r2 = a1, a2
        p.overlap_require, p.overlap_diff_limit, p.insert_size_max = self.overlapRequire, self.overlapDiffLimit, self.insertSizeMax
        p.correction_enabled = int(self.correction.enabled)
        p.merge_enabled, p.merge_discard_unmerged = int(self.mergePE.enabled), int(self.mergePE.discardUnmerged)
        p.kmer_len = self.kmer.kmerLen if self.kmer.enabled else 0
        p.dup_enabled, p.dup_keylen = int(self.duplicate.enabled), self.duplicate.keylen
        p.max_cycles = self.maxCycles
        return p


def from_cli(argv):
    """Parse the subset of the reference's CLI flags that reach the hot path
    (flag names of /root/reference/src/main.cpp:14-122)."""
    o = Options()
    it = iter(argv)
    flag = {"-q": ("qualFilter", "enabled"), "-a": ("adapter", "enableTriming"), "-g": ("polyGTrim", "enabled"),
            "-x": ("polyXTrim", "enabled"), "-d": ("duplicate", "enabled"), "-m": ("mergePE", "enabled"),
            "--discard_unmerged": ("mergePE", "discardUnmerged"), "-l": ("lengthFilter", "enabled"),
            "-y": ("complexityFilter", "enabled"), "--kmer": ("kmer", "enabled"), "-c": ("correction", "enabled"),
            "--enable_cut_front": ("qualitycut", "enableFront"), "--enable_cut_tail": ("qualitycut", "enableTail"),
            "--enable_cut_right": ("qualitycut", "enableRright")}
    val = {"-f": ("trim", "front1", int), "-t": ("trim", "tail1", int), "-b": ("trim", "maxLen1", int),
           "-F": ("trim", "front2", int), "-T": ("trim", "tail2", int), "-B": ("trim", "maxLen2", int),
           "--adapter_of_read1": ("adapter", "inputAdapterSeqR1", str), "--adapter_of_read2": ("adapter", "inputAdapterSeqR2", str),
           "--min_len_detect_polyG": ("polyGTrim", "minLen", int), "--max_mismatches_polyG": ("polyGTrim", "maxMismatch", int),
           "--one_mismatch_each_polyG": ("polyGTrim", "allowedOneMismatchForEach", int),
           "--base_to_trim": ("polyXTrim", "trimChr", str), "--min_len_detect_polyX": ("polyXTrim", "minLen", int),
           "--max_mismatches_polyX": ("polyXTrim", "maxMismatch", int), "--one_mismatch_each_polyX": ("polyXTrim", "allowedOneMismatchForEach", int),
           # -W / -M only write the *Shared fields, which nothing reads (dead flags)
           "-W": ("qualitycut", "windowSizeShared", int), "-M": ("qualitycut", "qualityShared", int),
           "--cut_front_window": ("qualitycut", "windowSizeFront", int), "--cut_tail_window": ("qualitycut", "windowSizeTail", int),
           "--cut_right_window": ("qualitycut", "windowSizeRight", int), "--cut_front_mean_qual": ("qualitycut", "qualityFront", int),
           "--cut_tail_mean_qual": ("qualitycut", "qualityTail", int), "--cut_right_mean_qual": ("qualitycut", "qualityRight", int),
           "-Q": ("qualFilter", "lowQualityLimit", int), "-U": ("qualFilter", "lowQualityRatio", float),
           "-N": ("qualFilter", "nBaseLimit", int), "-e": ("qualFilter", "averageQualityLimit", float),
           "--min_length": ("lengthFilter", "minReadLength", int), "--max_length": ("lengthFilter", "maxReadLength", int),
           "-Y": ("complexityFilter", "threshold", float), "--dup_ana_key_len": ("duplicate", "keylen", int),
           "--kmer_length": ("kmer", "kmerLen", int)}
    top = {"--min_overlap_len": ("overlapRequire", int), "--max_diff_for_overlap": ("overlapDiffLimit", int)}
    for a in it:
        if a in flag:
            setattr(getattr(o, flag[a][0]), flag[a][1], True)
        elif a in val:
            g, n, ty = val[a]
            setattr(getattr(o, g), n, ty(next(it)))
        elif a in top:
            setattr(o, top[a][0], top[a][1](next(it)))
        elif a == "-I":
            o.in2 = next(it)
        elif a in ("-i", "-o", "-O", "-w", "-J", "-H", "--merge_output", "--failed_out",
                   "--unpaired_read1", "--unpaired_read2", "-z"):
            next(it)            # host-side I/O flags: not part of the hot path
        else:
            raise ValueError(f"flag {a!r} is outside the hot-path scope of this package")
    return o
