"""ctypes mirror of include/fqk.h and loader of the CUDA library.

The product path has no CPU fallback: if ``libfqk.so`` (built by
``__graft_entry__.build()`` / ``fqtool_b200.build``) is missing, ``load()``
raises.  Nothing in this package imports ``oracle/``.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FQK_LIB") or os.path.join(HERE, "csrc", "libfqk.so")   # FQK_LIB: A/B builds of the same library (development aid)

FQK_ABI_VERSION = 1
FQK_MAX_ADAPTER = 128
FQK_MAX_READ_LEN = 4096
FQK_PAD = 16

# result codes, /root/reference/src/common.h:9-16
PASS_FILTER, FAIL_POLY_X, FAIL_OVERLAP, FAIL_N_BASE = 0, 4, 8, 12
FAIL_LENGTH, FAIL_TOO_LONG, FAIL_QUALITY, FAIL_COMPLEXITY = 16, 17, 20, 24
FILTER_RESULT_TYPES = 32

F_DROPPED, F_POLYG, F_POLYX, F_POLYX_SHIFT = 0x01, 0x02, 0x04, 3
F_ADAPTER_SEQ, F_ADAPTER_OV = 0x40, 0x80
P_OV1_DONE, P_OV1_OVERLAP, P_OV2_DONE, P_OV2_OVERLAP = 0x01, 0x02, 0x04, 0x08
P_MERGED, P_MERGE_PROC = 0x10, 0x20
T_ALL, T_TRIM, T_STATS, T_DUP, T_FUSED, T_SLOTS = 0, 1, 2, 3, 4, 5


class FqParams(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("is_paired", C.c_int32),
        ("trim_front1", C.c_int32), ("trim_tail1", C.c_int32), ("trim_max_len1", C.c_int32),
        ("trim_front2", C.c_int32), ("trim_tail2", C.c_int32), ("trim_max_len2", C.c_int32),
        ("cut_front_enabled", C.c_int32), ("cut_tail_enabled", C.c_int32), ("cut_right_enabled", C.c_int32),
        ("cut_front_window", C.c_int32), ("cut_tail_window", C.c_int32), ("cut_right_window", C.c_int32),
        ("cut_front_qual", C.c_int32), ("cut_tail_qual", C.c_int32), ("cut_right_qual", C.c_int32),
        ("polyg_enabled", C.c_int32), ("polyg_min_len", C.c_int32), ("polyg_max_mismatch", C.c_int32),
        ("polyg_one_mismatch_each", C.c_int32),
        ("polyx_enabled", C.c_int32), ("polyx_base_mask", C.c_int32), ("polyx_min_len", C.c_int32),
        ("polyx_max_mismatch", C.c_int32), ("polyx_one_mismatch_each", C.c_int32),
        ("qual_filter_enabled", C.c_int32), ("low_quality_limit", C.c_int32),
        ("low_quality_base_limit", C.c_int32), ("n_base_limit", C.c_int32),
        ("average_quality_limit", C.c_double),
        ("length_filter_enabled", C.c_int32), ("min_read_length", C.c_int32), ("max_read_length", C.c_int32),
        ("complexity_enabled", C.c_int32), ("reserved0", C.c_int32),
        ("complexity_threshold", C.c_double),
        ("adapter_trim_enabled", C.c_int32), ("adapter_r1_len", C.c_int32), ("adapter_r2_len", C.c_int32),
        ("reserved1", C.c_int32),
        ("adapter_r1", C.c_char * FQK_MAX_ADAPTER), ("adapter_r2", C.c_char * FQK_MAX_ADAPTER),
        ("overlap_require", C.c_int32), ("overlap_diff_limit", C.c_int32), ("insert_size_max", C.c_int32),
        ("correction_enabled", C.c_int32),
        ("merge_enabled", C.c_int32), ("merge_discard_unmerged", C.c_int32),
        ("kmer_len", C.c_int32),
        ("dup_enabled", C.c_int32), ("dup_keylen", C.c_int32),
        ("max_cycles", C.c_int32),
    ]


class FqBatch(C.Structure):
    _fields_ = [
        ("n_reads", C.c_uint32), ("max_len", C.c_uint32),
        ("first_index", C.c_uint64), ("total_bases", C.c_uint64),
        ("offsets", C.c_void_p), ("lengths", C.c_void_p),
        ("bases2", C.c_void_p), ("qual", C.c_void_p),
    ]


class FqAsciiBatch(C.Structure):
    _fields_ = [("n_reads", C.c_uint32), ("reserved", C.c_uint32), ("first_index", C.c_uint64),
                ("off", C.c_void_p), ("seq", C.c_void_p), ("qual", C.c_void_p)]


class FqTextBatch(C.Structure):
    _fields_ = [("reads", FqAsciiBatch), ("name_off", C.c_void_p), ("names", C.c_void_p),
                ("strand_off", C.c_void_p), ("strands", C.c_void_p)]


class FqRawBlock(C.Structure):
    _fields_ = [("text", C.c_void_p), ("bytes", C.c_uint64), ("first_index", C.c_uint64)]


class FqTextOut(C.Structure):
    _fields_ = [("data", C.c_void_p), ("cap", C.c_uint64), ("len", C.c_uint64), ("records", C.c_uint64)]


CH_OUT1, CH_OUT2, CH_MERGED, CH_FAILED, CH_UNPAIRED1, CH_UNPAIRED2, CH_SINGLE, CH_COUNT = 0, 1, 2, 3, 4, 5, 6, 7
W_FAILED, W_UNPAIRED1, W_STDOUT = 1, 2, 4


class FqStatsLayout(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "total_u64", "max_cycles", "kmer_bins")] + [("stats_off", C.c_uint64 * 4)] + [
        (n, C.c_uint64) for n in (
            "stats_stride", "s_q20", "s_q30", "s_content", "s_qualsum", "s_total_base",
            "s_total_qual", "s_reads", "s_length_sum", "s_kmer", "filter_off",
            "adapter_reads_off", "adapter_bases_off", "polyx_reads_off", "polyx_bases_off",
            "merged_pairs_off", "insert_hist_off", "insert_hist_len", "dup_bins", "correction_off",
            "corrected_reads_off")]


# numpy views of the result records (same memory layout as the C structs)
READ_RESULT_DTYPE = np.dtype([
    ("start", "<u2"), ("len", "<u2"), ("result", "u1"), ("flags", "u1"),
    ("adapter_pos", "<i2"), ("adapter_len", "<u2"), ("polyg_len", "<u2"),
    ("polyx_len", "<u2"), ("reserved", "<u2")])
PAIR_RESULT_DTYPE = np.dtype([
    ("ov1_offset", "<i2"), ("ov1_len", "<u2"), ("ov1_diff", "<u2"),
    ("ov2_offset", "<i2"), ("ov2_len", "<u2"), ("ov2_diff", "<u2"),
    ("merge_len1", "<u2"), ("merge_len2", "<u2"),
    ("flags", "u1"), ("merged_result", "u1"), ("insert_size", "<u2"), ("reserved", "<u4")])
CORRECTION_DTYPE = np.dtype([("pair", "<u4"), ("pos", "<u2"), ("mate_base", "u1"), ("qual", "u1")])
assert READ_RESULT_DTYPE.itemsize == 16 and PAIR_RESULT_DTYPE.itemsize == 24 and CORRECTION_DTYPE.itemsize == 8

EXPORTS = [
    "fqk_default_params", "fqk_stats_layout", "fqk_create", "fqk_destroy", "fqk_last_error",
    "fqk_nccl_unique_id", "fqk_process_se", "fqk_process_pe", "fqk_submit_se", "fqk_wait", "fqk_process_se_ascii", "fqk_process_pe_ascii", "fqk_process_se_device",
    "fqk_process_pe_device", "fqk_sync", "fqk_stream", "fqk_launch_count",
    "fqk_set_kernel_timing", "fqk_kernel_times", "fqk_set_impl", "fqk_get_corrections", "fqk_reset", "fqk_finalize",
    "fqk_dup_table", "fqk_packed_bases", "fqk_pack_ascii", "fqk_host_alloc", "fqk_host_free",
    "fqk_prepass_kmer_counts", "fqk_prepass_seed_hits", "fqk_process_se_text", "fqk_process_pe_text",
    "fqk_process_se_fastq", "fqk_process_pe_fastq", "fqk_submit_se_ascii",
    "fqk_submit_se_text", "fqk_wait_text", "fqk_submit_se_fastq", "fqk_wait_fastq",
    "fqk_prepass_kmer_counts_static", "fqk_submit_pe",
]

_lib = None


def load():
    """dlopen libfqk.so and declare prototypes; raises if the library is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m fqtool_b200.build` "
            "(there is no CPU fallback for the hot path)")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, i32, u32, u64 = C.c_void_p, C.c_int, C.c_uint32, C.c_uint64
    P = C.POINTER
    lib.fqk_default_params.argtypes = [P(FqParams)]; lib.fqk_default_params.restype = None
    lib.fqk_stats_layout.argtypes = [P(FqParams), P(FqStatsLayout)]; lib.fqk_stats_layout.restype = i32
    lib.fqk_create.argtypes = [P(vp), P(FqParams), i32, i32, i32, vp]; lib.fqk_create.restype = i32
    lib.fqk_destroy.argtypes = [vp]; lib.fqk_destroy.restype = None
    lib.fqk_last_error.argtypes = [vp]; lib.fqk_last_error.restype = C.c_char_p
    lib.fqk_nccl_unique_id.argtypes = [vp]; lib.fqk_nccl_unique_id.restype = i32
    lib.fqk_process_se.argtypes = [vp, P(FqBatch), vp]; lib.fqk_process_se.restype = i32
    lib.fqk_process_pe.argtypes = [vp, P(FqBatch), P(FqBatch), vp, vp, vp]; lib.fqk_process_pe.restype = i32
    lib.fqk_process_se_ascii.argtypes = [vp, P(FqAsciiBatch), vp]; lib.fqk_process_se_ascii.restype = i32
    lib.fqk_process_pe_ascii.argtypes = [vp, P(FqAsciiBatch), P(FqAsciiBatch), vp, vp, vp]; lib.fqk_process_pe_ascii.restype = i32
    lib.fqk_submit_se.argtypes = [vp, P(FqBatch), vp, P(u64)]; lib.fqk_submit_se.restype = i32
    lib.fqk_wait.argtypes = [vp, u64]; lib.fqk_wait.restype = i32
    lib.fqk_submit_pe.argtypes = [vp, P(FqBatch), P(FqBatch), vp, vp, vp, P(u64)]; lib.fqk_submit_pe.restype = i32
    lib.fqk_submit_se_ascii.argtypes = [vp, P(FqAsciiBatch), vp, P(u64)]; lib.fqk_submit_se_ascii.restype = i32
    lib.fqk_process_se_device.argtypes = [vp, P(FqBatch), vp]; lib.fqk_process_se_device.restype = i32
    lib.fqk_process_pe_device.argtypes = [vp, P(FqBatch), P(FqBatch), vp, vp, vp]
    lib.fqk_process_pe_device.restype = i32
    lib.fqk_sync.argtypes = [vp]; lib.fqk_sync.restype = i32
    lib.fqk_stream.argtypes = [vp]; lib.fqk_stream.restype = vp
    lib.fqk_launch_count.argtypes = [vp]; lib.fqk_launch_count.restype = u64
    lib.fqk_set_kernel_timing.argtypes = [vp, i32]; lib.fqk_set_kernel_timing.restype = i32
    lib.fqk_kernel_times.argtypes = [vp, P(C.c_double), P(u64)]; lib.fqk_kernel_times.restype = i32
    lib.fqk_set_impl.argtypes = [vp, i32]; lib.fqk_set_impl.restype = i32
    lib.fqk_get_corrections.argtypes = [vp, vp, u32, P(u32)]; lib.fqk_get_corrections.restype = i32
    lib.fqk_reset.argtypes = [vp]; lib.fqk_reset.restype = i32
    lib.fqk_finalize.argtypes = [vp, vp]; lib.fqk_finalize.restype = i32
    lib.fqk_dup_table.argtypes = [vp, vp, vp, vp]; lib.fqk_dup_table.restype = i32
    lib.fqk_packed_bases.argtypes = [u32, vp]; lib.fqk_packed_bases.restype = u64
    lib.fqk_pack_ascii.argtypes = [u32, vp, vp, vp, vp, vp, vp, vp]; lib.fqk_pack_ascii.restype = i32
    lib.fqk_host_alloc.argtypes = [C.c_size_t]; lib.fqk_host_alloc.restype = vp
    lib.fqk_process_se_text.argtypes = [vp, P(FqTextBatch), u32, vp, P(FqTextOut), P(u64)]; lib.fqk_process_se_text.restype = i32
    lib.fqk_process_pe_text.argtypes = [vp, P(FqTextBatch), P(FqTextBatch), u32, vp, vp, vp, P(FqTextOut), P(u64)]
    lib.fqk_process_pe_text.restype = i32
    lib.fqk_process_se_fastq.argtypes = [vp, P(FqRawBlock), u32, vp, u32, P(FqTextOut), P(u64), P(u32), P(u64)]
    lib.fqk_process_se_fastq.restype = i32
    lib.fqk_process_pe_fastq.argtypes = [vp, P(FqRawBlock), P(FqRawBlock), u32, vp, vp, vp, u32, P(FqTextOut), P(u64), P(u32), P(u64), P(u64)]
    lib.fqk_process_pe_fastq.restype = i32
    lib.fqk_submit_se_text.argtypes = [vp, P(FqTextBatch), u32, vp, P(FqTextOut), P(u64)]; lib.fqk_submit_se_text.restype = i32
    lib.fqk_wait_text.argtypes = [vp, u64, P(u64)]; lib.fqk_wait_text.restype = i32
    lib.fqk_submit_se_fastq.argtypes = [vp, P(FqRawBlock), P(u64)]; lib.fqk_submit_se_fastq.restype = i32
    lib.fqk_wait_fastq.argtypes = [vp, u64, u32, vp, u32, P(FqTextOut), P(u64), P(u32), P(u64)]; lib.fqk_wait_fastq.restype = i32
    lib.fqk_prepass_kmer_counts.argtypes = [i32, vp, i32, i32, i32, vp]; lib.fqk_prepass_kmer_counts.restype = i32
    lib.fqk_prepass_kmer_counts_static.argtypes = [i32, u32, vp, vp, i32, i32, i32, vp, vp]
    lib.fqk_prepass_kmer_counts_static.restype = i32
    lib.fqk_prepass_seed_hits.argtypes = [i32, vp, i32, i32, i32, u32, vp, vp, u32, vp]; lib.fqk_prepass_seed_hits.restype = i32
    lib.fqk_host_free.argtypes = [vp]; lib.fqk_host_free.restype = None
    _lib = lib
    return lib


def ptr(a):
    """void* of a numpy array (must be C-contiguous)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)
