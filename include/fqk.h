/*
 * fqk.h -- C ABI of the B200-native per-read QC + trimming hot path of fqtool.
 *
 * This is the drop-in boundary.  The reference has no
 * plugin/FFI API; the seam these entry points replace is the per-pack worker
 * body of
 *     SingleEndProcessor::processSingleEnd(ReadPack*, ThreadConfig*)
 *         /root/reference/src/seprocessor.cpp:290-388
 *     PairEndProcessor::processPairEnd(ReadPairPack*, ThreadConfig*)
 *         /root/reference/src/peprocessor.cpp:261-508
 * plus the post-join merges Stats::merge (src/stats.cpp:815-863),
 * FilterResult::merge (src/filterresult.cpp:52-102) and the Duplicate table
 * (src/duplicate.cpp:46-166).
 *
 * Plain pointers and sizes only; no C++/torch types.  Every function returns
 * 0 on success and a negative FQK_E_* code on error (never throws); the text
 * of the last error of a context is available from fqk_last_error().
 */
#ifndef FQK_H
#define FQK_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FQK_ABI_VERSION 1

/* error codes */
#define FQK_OK            0
#define FQK_E_INVALID    -1   /* bad argument / unsupported parameter value   */
#define FQK_E_CUDA       -2   /* CUDA runtime error                           */
#define FQK_E_NOMEM      -3   /* host or device allocation failed             */
#define FQK_E_ALPHABET   -4   /* base outside {A,C,G,T,N} given to the packer */
#define FQK_E_NCCL       -5   /* NCCL unavailable or failed                   */
#define FQK_E_LIMIT      -6   /* read longer than FQK_MAX_READ_LEN etc.       */

/* result codes of Filter::passFilter, /root/reference/src/common.h:9-16 */
#define FQK_PASS_FILTER      0
#define FQK_FAIL_POLY_X      4
#define FQK_FAIL_OVERLAP     8
#define FQK_FAIL_N_BASE     12
#define FQK_FAIL_LENGTH     16
#define FQK_FAIL_TOO_LONG   17
#define FQK_FAIL_QUALITY    20
#define FQK_FAIL_COMPLEXITY 24
#define FQK_FILTER_RESULT_TYPES 32

#define FQK_MAX_ADAPTER   128      /* bases of an explicit adapter            */
#define FQK_MAX_READ_LEN  4096     /* bases per read accepted by the kernels  */
#define FQK_PAD           16       /* every read starts at a multiple of this */
#define FQK_BATCH_PAD     64       /* total_bases of a batch is a multiple of this */

/*
 * fq_params: flat POD mirror of the hot-path fields of the reference's
 * Options tree (/root/reference/src/options.h:311-386), already passed through
 * Options::update() (src/options.cpp:24-58): low_quality_limit is Phred+33 and
 * low_quality_base_limit is the int the reference derives (22 by default).
 */
typedef struct fq_params {
    int32_t abi_version;            /* FQK_ABI_VERSION                                  */
    int32_t is_paired;              /* Options::isPaired()            options.cpp:73    */
    /* ForceTrimOptions options.h:239-255 */
    int32_t trim_front1, trim_tail1, trim_max_len1;
    int32_t trim_front2, trim_tail2, trim_max_len2;
    /* QualityCutOptions options.h:104-131 (window/quality Front, Tail, Right) */
    int32_t cut_front_enabled, cut_tail_enabled, cut_right_enabled;
    int32_t cut_front_window,  cut_tail_window,  cut_right_window;
    int32_t cut_front_qual,    cut_tail_qual,    cut_right_qual;
    /* PolyGTrimmerOptions options.h:39-51 */
    int32_t polyg_enabled, polyg_min_len, polyg_max_mismatch, polyg_one_mismatch_each;
    /* PolyXTrimmerOptions options.h:54-68; base mask bit0 A, bit1 T, bit2 C, bit3 G, bit4 N */
    int32_t polyx_enabled, polyx_base_mask, polyx_min_len, polyx_max_mismatch, polyx_one_mismatch_each;
    /* QualityFilterOptions options.h:197-213 */
    int32_t qual_filter_enabled, low_quality_limit, low_quality_base_limit, n_base_limit;
    double  average_quality_limit;
    /* ReadLengthFilterOptions options.h:183-194 */
    int32_t length_filter_enabled, min_read_length, max_read_length;
    /* LowComplexityFilterOptions options.h:172-180 */
    int32_t complexity_enabled, reserved0;
    double  complexity_threshold;
    /* AdapterOptions options.h:216-236: enableTriming + explicit adapters (ASCII, A/C/G/T/N) */
    int32_t adapter_trim_enabled, adapter_r1_len, adapter_r2_len, reserved1;
    char    adapter_r1[FQK_MAX_ADAPTER];
    char    adapter_r2[FQK_MAX_ADAPTER];
    /* Options::overlapRequire / overlapDiffLimit / insertSizeMax options.cpp:17-19 */
    int32_t overlap_require, overlap_diff_limit, insert_size_max;
    /* CorrectionOptions options.h:163-169: PE base correction by overlap analysis (-c) */
    int32_t correction_enabled;
    /* MergePEReadsOptions options.h:27-36 */
    int32_t merge_enabled, merge_discard_unmerged;
    /* KmerOptions options.h:279-287; 0 = off, else 4..15 */
    int32_t kmer_len;
    /* DuplicationAnalysisOptions options.h:91-101; key length 12..14 on device */
    int32_t dup_enabled, dup_keylen;
    /* capacity of every per-cycle array (the reference grows Stats buffers on
     * demand, src/stats.cpp:240-242; here the caller states the maximum read
     * length, merged reads included).  0 = derive 512. */
    int32_t max_cycles;
} fq_params;

/*
 * fq_batch: one pack of reads in the B200 layout.  Bases are 2 bits each
 * (A=0 T=1 C=2 G=3, the reference's own k-mer code, src/evaluator.cpp:9-23;
 * N is stored as 0), four per byte, base i of the buffer in bits 2*(i%4) of
 * byte i/4.  Qualities are one byte per base: Phred+33 in bits 0-6 and the
 * N flag in bit 7.  Read r occupies base positions
 * [offsets[r], offsets[r]+lengths[r]) of both buffers and offsets[r] is a
 * multiple of FQK_PAD, so every read starts on a 16-byte quality boundary and
 * a 4-byte base boundary.  Batches are dense and ordered: offsets[r+1] ==
 * offsets[r] + lengths[r] rounded up to FQK_PAD; total_bases is the end of the
 * last read rounded up to FQK_BATCH_PAD.  All padding bytes MUST be zero (the
 * packer writes them so) and bases2/qual must be 16-byte aligned.
 * Buffers are caller-owned and must stay valid until the call returns.
 */
typedef struct fq_batch {
    uint32_t        n_reads;
    uint32_t        max_len;       /* longest read of the batch, 0 = unknown (the host entry points then
                                      scan lengths[]; the device entry points assume max_cycles, which keeps
                                      paired-end batches off the fused kernel when max_cycles > 496)      */
    uint64_t        first_index;   /* global index of read 0 (input order)    */
    uint64_t        total_bases;   /* padded extent: bases2 has total/4 bytes */
    const uint32_t* offsets;       /* [n_reads]                               */
    const uint16_t* lengths;       /* [n_reads]                               */
    const uint8_t*  bases2;        /* [total_bases/4]                         */
    const uint8_t*  qual;          /* [total_bases]                           */
} fq_batch;

/* flags of fq_read_result */
#define FQK_F_DROPPED      0x01   /* trimAndCut returned NULL (filter.cpp:69-189)      */
#define FQK_F_POLYG        0x02   /* trimPolyG tallied this read (polyx.cpp:31-36)     */
#define FQK_F_POLYX        0x04   /* trimPolyX tallied this read (polyx.cpp:82-100)    */
#define FQK_F_POLYX_SHIFT  3      /* bits 3-5: poly base index 0 A,1 T,2 C,3 G,4 N     */
#define FQK_F_ADAPTER_SEQ  0x40   /* trimBySequence hit (adaptertrimmer.cpp:71-88)     */
#define FQK_F_ADAPTER_OV   0x80   /* trimByOverlapAnalysis hit (adaptertrimmer.cpp:16) */

/*
 * fq_read_result: what the host needs to rebuild the trimmed read and the
 * per-read tallies.  The kept read is original[start, start+len).
 */
typedef struct fq_read_result {
    uint16_t start;        /* first kept base in the original read                 */
    uint16_t len;          /* kept length after every trim (0 when dropped)        */
    uint8_t  result;       /* passFilter code of this read (FQK_PASS_FILTER ...)   */
    uint8_t  flags;        /* FQK_F_*                                              */
    int16_t  adapter_pos;  /* trimBySequence pos (may be <0) / overlap trim length */
    uint16_t adapter_len;  /* length of the tallied adapter text                   */
    uint16_t polyg_len;    /* bases tallied by addPolyXTrimmed(3, .)               */
    uint16_t polyx_len;    /* bases tallied by addPolyXTrimmed(poly, .)            */
    uint16_t reserved;
} fq_read_result;

/* flags of fq_pair_result */
#define FQK_P_OV1_DONE     0x01  /* first analyze() ran (peprocessor.cpp:303 or :330)   */
#define FQK_P_OV1_OVERLAP  0x02
#define FQK_P_OV2_DONE     0x04  /* merge-time analyze() ran (peprocessor.cpp:354)      */
#define FQK_P_OV2_OVERLAP  0x08
#define FQK_P_MERGED       0x10  /* merged read passed the filter and was emitted       */
#define FQK_P_MERGE_PROC   0x20  /* mergeProcessed == true (peprocessor.cpp:366,384)    */

/*
 * fq_pair_result: OverlapResult (src/overlapanalysis.h:12-20) of both analyze
 * calls, the merge geometry of OverlapAnalysis::merge (overlapanalysis.cpp:74-93)
 * and the result code recorded by addFilterResult for the pair.
 */
typedef struct fq_pair_result {
    int16_t  ov1_offset;  uint16_t ov1_len;  uint16_t ov1_diff;
    int16_t  ov2_offset;  uint16_t ov2_len;  uint16_t ov2_diff;
    uint16_t merge_len1;  uint16_t merge_len2;
    uint8_t  flags;        /* FQK_P_*                                               */
    uint8_t  merged_result;/* passFilter(merged) when FQK_P_OV2_OVERLAP             */
    uint16_t insert_size;  /* value counted in the insert-size histogram, 0xFFFF none */
    uint32_t reserved;
} fq_pair_result;

/*
 * fq_stats_layout: where each run-level accumulator lives inside the flat
 * uint64 buffer returned by fqk_finalize (and all-reduced across ranks).  All
 * offsets are in uint64 units.  A Stats block s (0 pre1, 1 pre2, 2 post1,
 * 3 post2) starts at stats_off[s] and holds, in the reference's own indexing
 * (bin = base & 7, src/stats.cpp:249; C = max_cycles):
 *   q20[8][C], q30[8][C], content[8][C], qualsum[8][C]   (stats.h mCycle*)
 *   total_base[C], total_qual[C], reads, length_sum, kmer[4^k]
 */
typedef struct fq_stats_layout {
    uint64_t total_u64;
    uint64_t max_cycles;
    uint64_t kmer_bins;            /* 4^k or 0                                  */
    uint64_t stats_off[4];
    uint64_t stats_stride;         /* u64 per Stats block                       */
    uint64_t s_q20, s_q30, s_content, s_qualsum;   /* offsets inside a block    */
    uint64_t s_total_base, s_total_qual, s_reads, s_length_sum, s_kmer;
    uint64_t filter_off;           /* mFilterReadStats[32] filterresult.h:26    */
    uint64_t adapter_reads_off;    /* mTrimmedAdapterReads                      */
    uint64_t adapter_bases_off;    /* mTrimmedAdapterBases                      */
    uint64_t polyx_reads_off;      /* mTrimmedPolyXReads[5]                     */
    uint64_t polyx_bases_off;      /* mTrimmedPolyXBases[5]                     */
    uint64_t merged_pairs_off;     /* mMergedPairs                              */
    uint64_t insert_hist_off;      /* mInsertSizeHist[insert_size_max+1]        */
    uint64_t insert_hist_len;
    uint64_t dup_bins;             /* 4^keylen or 0 (table is returned apart)   */
    uint64_t correction_off;       /* mCorrectionMatrix[64] filterresult.h      */
    uint64_t corrected_reads_off;  /* mCorrectedReads                           */
} fq_stats_layout;

/*
 * fq_correction: one base rewritten by BaseCorrector::correctByOverlapAnalysis
 * (src/basecorrector.cpp:14-70).  The host applies it to its own strings before slicing.
 */
typedef struct fq_correction {
    uint32_t pair;        /* index of the pair inside the batch                         */
    uint16_t pos;         /* position in the ORIGINAL (untrimmed) read                  */
    uint8_t  mate_base;   /* bit 7: 0 = read1, 1 = read2; bits 0-6: new base (ASCII)    */
    uint8_t  qual;        /* new quality character                                      */
} fq_correction;

typedef struct fqk_ctx fqk_ctx;

/* fill p with the reference's defaults (options.h constructors + update()) */
void fqk_default_params(fq_params* p);

int  fqk_stats_layout(const fq_params* p, fq_stats_layout* out);

/*
 * One context per GPU, owned by one host thread.  nccl_unique_id is the 128
 * byte ncclUniqueId shared by all ranks (from fqk_nccl_unique_id on rank 0),
 * or NULL when nranks == 1.
 */
int  fqk_create(fqk_ctx** out, const fq_params* p, int device_ordinal,
                int rank, int nranks, const void* nccl_unique_id);
void fqk_destroy(fqk_ctx* ctx);
const char* fqk_last_error(const fqk_ctx* ctx);   /* ctx may be NULL: create errors */

int  fqk_nccl_unique_id(void* out128);

/*
 * Host-buffer entry points: replace the loop bodies of processSingleEnd /
 * processPairEnd.  Batches live in (preferably pinned) host memory; results
 * are written to host arrays of n_reads entries.  H2D, kernels and D2H are
 * all inside the call.
 */
int  fqk_process_se(fqk_ctx* ctx, const fq_batch* in, fq_read_result* out);
int  fqk_process_pe(fqk_ctx* ctx, const fq_batch* r1, const fq_batch* r2,
                    fq_read_result* out1, fq_read_result* out2, fq_pair_result* outp);

/*
 * ASCII entry points: the host hands over what a FASTQ parser holds (flat sequence and quality
 * characters, read r at [off[r], off[r+1])) and the 2-bit packing runs on the GPU (k_pack_ascii),
 * so the host does no per-base work at all.  PCIe carries 2 B/base instead of 1.25.  A base
 * outside {A,C,G,T,N} or a quality >= 128 makes the call fail with FQK_E_ALPHABET.
 */
typedef struct fq_ascii_batch {
    uint32_t        n_reads;
    uint32_t        reserved;
    uint64_t        first_index;
    const uint64_t* off;           /* [n_reads + 1]                           */
    const char*     seq;           /* [off[n_reads]]                          */
    const char*     qual;          /* [off[n_reads]]                          */
} fq_ascii_batch;
int  fqk_process_se_ascii(fqk_ctx* ctx, const fq_ascii_batch* in, fq_read_result* out);
int  fqk_process_pe_ascii(fqk_ctx* ctx, const fq_ascii_batch* r1, const fq_ascii_batch* r2,
                          fq_read_result* out1, fq_read_result* out2, fq_pair_result* outp);

/*
 * Text entry points: device-side output compaction and FASTQ record assembly.  The host hands
 * over what FqReader parsed (names, sequences, third lines, qualities as flat character buffers) and gets back, per
 * output channel, ONE contiguous block of FASTQ text in input order: exactly the strings the reference's worker
 * bodies build with Read::toString / toStringWithTag (src/read.h:166-176), the routing ladders of
 * src/seprocessor.cpp:340-349 and src/peprocessor.cpp:353-429 and OverlapAnalysis::merge (src/overlapanalysis.cpp:
 * 74-104, name quirk included).  The writer threads only pass the blocks on; no per-read string is built on the host.
 * `writers` says which optional writers exist (the reference tests its WriterThread pointers).  outs[ch].data is a
 * caller-owned (preferably pinned) buffer of outs[ch].cap bytes, or NULL for a channel the caller does not want;
 * on return outs[ch].len / .records hold the bytes and records of the channel.  A channel that does not fit makes the
 * call fail with FQK_E_LIMIT (outs[ch].len then holds the size needed).  The records (out / out1 / out2 / outp) are
 * still returned when the pointers are not NULL: the adapter text tallies are keyed by strings on the host.
 */
#define FQK_CH_OUT1      0
#define FQK_CH_OUT2      1
#define FQK_CH_MERGED    2
#define FQK_CH_FAILED    3
#define FQK_CH_UNPAIRED1 4
#define FQK_CH_UNPAIRED2 5
#define FQK_CH_SINGLE    6      /* interleaved pairs for stdout (Options::outputToSTDOUT) */
#define FQK_CH_COUNT     7
#define FQK_W_FAILED     1u     /* mFailedWriter != NULL        */
#define FQK_W_UNPAIRED1  2u     /* mUnPairedLeftWriter != NULL  */
#define FQK_W_STDOUT     4u     /* Options::outputToSTDOUT      */
typedef struct fq_text_batch {
    fq_ascii_batch  reads;         /* sequence and quality characters            */
    const uint64_t* name_off;      /* [n_reads + 1]                              */
    const char*     names;         /* first lines (with the '@'), no newline     */
    const uint64_t* strand_off;    /* [n_reads + 1]                              */
    const char*     strands;       /* third lines ("+..."), no newline           */
} fq_text_batch;
typedef struct fq_text_out {
    char*    data;                 /* caller-owned buffer, NULL = channel unused */
    uint64_t cap;
    uint64_t len;                  /* out: bytes written (or needed)             */
    uint64_t records;              /* out: FASTQ records in the block            */
} fq_text_out;
int  fqk_process_se_text(fqk_ctx* ctx, const fq_text_batch* in, uint32_t writers, fq_read_result* out,
                         fq_text_out outs[FQK_CH_COUNT], uint64_t* read_passed);
int  fqk_process_pe_text(fqk_ctx* ctx, const fq_text_batch* r1, const fq_text_batch* r2, uint32_t writers,
                         fq_read_result* out1, fq_read_result* out2, fq_pair_result* outp,
                         fq_text_out outs[FQK_CH_COUNT], uint64_t* read_passed);

/* Pipelined form of fqk_process_se_text: fqk_submit_se_text enqueues the upload, the packing, the hot path and the sizes
 * of the text blocks; fqk_wait_text assembles and downloads them and fills outs[].len / .records (the array given to
 * submit, which must stay valid, like every buffer of the pack, until the wait returns).  Keep two packs in flight
 * (submit k+1, then wait k): the upload of pack k+1 overlaps the assembly and download of pack k.  Every submitted
 * pack must be waited for, in order, before the pack after the next one is submitted. */
int  fqk_submit_se_text(fqk_ctx* ctx, const fq_text_batch* in, uint32_t writers, fq_read_result* out,
                        fq_text_out outs[FQK_CH_COUNT], uint64_t* ticket);
int  fqk_wait_text(fqk_ctx* ctx, uint64_t ticket, uint64_t* read_passed);

/*
 * Raw FASTQ entry points: the host hands over the DECOMPRESSED TEXT of a FASTQ block as it
 * comes out of gzread and gets FASTQ text back; the newline scan of FqReader::getLine / FqReader::read
 * (src/fqreader.cpp:90-195), the packing, the hot path and the record assembly all run on the device.  Record r is
 * lines 4r .. 4r+3 of the block ('\r' before the newline is dropped); a trailing incomplete record is left alone and
 * *consumed says how many bytes the processed records took, so the caller carries the rest into its next block.
 * For pairs the two blocks are matched by position and min(records1, records2) pairs are processed.  Blocks that are
 * not well-formed four-line FASTQ ('@' name, '+' third line, equal lengths) are refused with FQK_E_INVALID: the
 * reference's skipping of stray lines (src/fqreader.cpp:167-170) is not reproduced.  out / out1 / out2 / outp may be
 * NULL; when given they must hold out_cap >= records entries.
 */
typedef struct fq_raw_block {
    const char* text;              /* decompressed FASTQ text (preferably pinned)  */
    uint64_t    bytes;
    uint64_t    first_index;       /* global index of the block's first read       */
} fq_raw_block;
int  fqk_process_se_fastq(fqk_ctx* ctx, const fq_raw_block* in, uint32_t writers, fq_read_result* out, uint32_t out_cap,
                          fq_text_out outs[FQK_CH_COUNT], uint64_t* read_passed, uint32_t* n_reads, uint64_t* consumed);
int  fqk_process_pe_fastq(fqk_ctx* ctx, const fq_raw_block* in1, const fq_raw_block* in2, uint32_t writers,
                          fq_read_result* out1, fq_read_result* out2, fq_pair_result* outp, uint32_t out_cap,
                          fq_text_out outs[FQK_CH_COUNT], uint64_t* read_passed, uint32_t* n_pairs,
                          uint64_t* consumed1, uint64_t* consumed2);

/* Pipelined form of fqk_process_se_fastq: fqk_submit_se_fastq uploads the block and counts its newlines (asynchronous);
 * fqk_wait_fastq does the rest and returns what fqk_process_se_fastq returns.  With two blocks in flight (submit k+1,
 * then wait k) the upload of block k+1 overlaps the kernels and the download of block k.  The block's text must stay
 * valid until its wait returns; blocks are independent (the caller cuts them at record boundaries, e.g. with the
 * *consumed of a previous call on a probe, or by scanning back to a line that starts with '@' after a quality line). */
int  fqk_submit_se_fastq(fqk_ctx* ctx, const fq_raw_block* in, uint64_t* ticket);
int  fqk_wait_fastq(fqk_ctx* ctx, uint64_t ticket, uint32_t writers, fq_read_result* out, uint32_t out_cap,
                    fq_text_out outs[FQK_CH_COUNT], uint64_t* read_passed, uint32_t* n_reads, uint64_t* consumed);

/*
 * Pipelined form of fqk_process_se for hosts that keep two packs in flight: fqk_submit_se enqueues
 * upload, kernels and download of a pack and returns a ticket; fqk_wait blocks until that pack's
 * records are in `out`.  The upload of pack k+1 overlaps the kernels and the download of pack k.
 * Buffers of a pack (pinned host memory) must stay valid until its fqk_wait returns.  Do not mix
 * with fqk_process_se on the same context without waiting for every ticket first.  A ticket is handed out only when the
 * pack was enqueued (a failed fqk_submit_se consumes none).  The first error of any pack sticks to the context: every
 * later fqk_wait (also for tickets already recycled) and fqk_finalize return it until fqk_reset.
 */
int  fqk_submit_se(fqk_ctx* ctx, const fq_batch* in, fq_read_result* out, uint64_t* ticket);
/* the same for FASTQ characters (fq_ascii_batch): per-read offsets, lengths and the packed layout are all derived on the
 * device, the call only enqueues copies and kernels; in->off / seq / qual must stay valid until fqk_wait */
int  fqk_submit_se_ascii(fqk_ctx* ctx, const fq_ascii_batch* in, fq_read_result* out, uint64_t* ticket);
/* the same for pairs (PairEndProcessor's workers run side by side, /root/reference/src/peprocessor.cpp:546-566; here one
 * caller thread keeps two pair packs in flight): both batches and the three record arrays must stay valid until
 * fqk_wait(ticket).  Not available with correction_enabled (the edit list of fqk_get_corrections belongs to ONE pack):
 * FQK_E_INVALID, use fqk_process_pe. */
int  fqk_submit_pe(fqk_ctx* ctx, const fq_batch* r1, const fq_batch* r2, fq_read_result* out1, fq_read_result* out2,
                   fq_pair_result* pairs, uint64_t* ticket);
int  fqk_wait(fqk_ctx* ctx, uint64_t ticket);

/*
 * Device-resident entry points (all pointers are device pointers of the
 * context's GPU, fq_batch itself is a host struct): kernels only, launched on
 * the context's stream without synchronising.  fqk_sync waits for them.
 */
/* With correction_enabled (-c) fqk_process_pe_device REWRITES the bases and qualities of the caller's device
 * buffers in place (BaseCorrector::correctByOverlapAnalysis mutates the Read objects too, src/basecorrector.cpp:14-70);
 * the fq_batch pointers are const only for the other option sets.  A resident batch must therefore not be processed
 * twice with -c: the second pass would see corrected data.  The host entry points work on their own staged copy. */
int  fqk_process_se_device(fqk_ctx* ctx, const fq_batch* in_dev, fq_read_result* out_dev);
int  fqk_process_pe_device(fqk_ctx* ctx, const fq_batch* r1_dev, const fq_batch* r2_dev,
                           fq_read_result* out1_dev, fq_read_result* out2_dev,
                           fq_pair_result* outp_dev);
int  fqk_sync(fqk_ctx* ctx);
void* fqk_stream(fqk_ctx* ctx);                   /* cudaStream_t of the context       */
uint64_t fqk_launch_count(const fqk_ctx* ctx);    /* kernels launched so far           */
/* Device time of the kernels of the process_* calls, measured with CUDA events
 * on the context's stream while timing is enabled.  Slots: FQK_T_ALL whole
 * kernel sequence of a call, FQK_T_TRIM, FQK_T_STATS, FQK_T_DUP, FQK_T_FUSED.
 * fqk_kernel_times synchronises, returns the totals since the last call and
 * resets them. */
#define FQK_T_ALL   0
#define FQK_T_TRIM  1
#define FQK_T_STATS 2
#define FQK_T_DUP   3
#define FQK_T_FUSED 4
#define FQK_T_SLOTS 5
int  fqk_set_kernel_timing(fqk_ctx* ctx, int enabled);
int  fqk_kernel_times(fqk_ctx* ctx, double ms[FQK_T_SLOTS], uint64_t launches[FQK_T_SLOTS]);
/* implementation selector for A/B tests: 0 = automatic (fused kernels when the
 * configuration allows), 1 = generic kernels only, 3 = the first warp-specialised
 * fused kernel (k_fused_se_ws) for single-end batches and the two-phase path
 * (k_trim_pe + statistics passes) for paired-end batches; 2 is retired */
int  fqk_set_impl(fqk_ctx* ctx, int impl);

/* corrections made by the last fqk_process_pe / fqk_process_pe_device call (synchronises);
 * *n receives the total, at most cap entries are written (unordered) */
int  fqk_get_corrections(fqk_ctx* ctx, fq_correction* out, uint32_t cap, uint32_t* n);

/* zero every accumulator (Stats, FilterResult counters, insert histogram, dup table) */
int  fqk_reset(fqk_ctx* ctx);

/*
 * End of run: merge accumulators over ranks (one NCCL all-reduce of the flat
 * buffer; min/sum passes for the duplication table) and copy them to the host.
 * flat has fq_stats_layout.total_u64 entries.  With nranks > 1 the call is COLLECTIVE: every rank must call it the
 * same number of times.  The ranks first agree (one all-reduce(max) of a status word) on whether every shard is
 * error-free; if one is not, every rank skips the merge and returns an error (its own, or FQK_E_INVALID "another rank
 * reported an error"), so a bad read on one shard cannot leave the other ranks blocked inside NCCL.  The local
 * accumulators and the local duplication table are not modified by the merge: processing may continue and a later
 * fqk_finalize merges again.  fqk_dup_table returns the table of the last fqk_finalize.
 */
int  fqk_finalize(fqk_ctx* ctx, uint64_t* flat);
/* Duplicate::mDups / mCounts / mGC (src/duplicate.h:64-66) after fqk_finalize;
 * any pointer may be NULL.  Arrays have fq_stats_layout.dup_bins entries. */
int  fqk_dup_table(fqk_ctx* ctx, uint64_t* dups, uint32_t* counts, uint8_t* gc);

/*
 * Host-side packer: ASCII reads -> fq_batch layout.  seq/qual are flat ASCII
 * buffers, read r at [src_off[r], src_off[r+1]).  Output arrays are caller
 * allocated: offsets[n], lengths[n], bases2[fqk_packed_bases(...)/4],
 * qual[fqk_packed_bases(...)].  Returns FQK_E_ALPHABET for a base outside
 * {A,C,G,T,N} and FQK_E_LIMIT for a read longer than FQK_MAX_READ_LEN.
 */
uint64_t fqk_packed_bases(uint32_t n, const uint64_t* src_off);
int  fqk_pack_ascii(uint32_t n, const char* seq, const char* qual, const uint64_t* src_off,
                    uint32_t* offsets, uint16_t* lengths, uint8_t* bases2, uint8_t* qual_out);

/*
 * Adapter auto-detection pre-pass, device part.  Replaces the two per-base loops of
 * Evaluator::evaluateAdapterSeq (/root/reference/src/evaluator.cpp:229-390: k-mer counting over the first 256 K reads,
 * k = 10, positions 20 .. len - k - shiftTail) and Evaluator::getAdapterWithSeed (src/evaluator.cpp:392-407: every
 * window whose key equals the seed).  Keys are Evaluator::seq2int's (A0 T1 C2 G3, first base most significant); a
 * window holding any other character is not counted.  Context-free (the pre-pass runs before the processors exist):
 * `device` is a CUDA ordinal; there is no CPU fallback.  The host keeps ranking the keys and the NucleotideTree walk.
 *   fqk_prepass_kmer_counts  counts[4^keylen] (caller allocated), keylen 1..15
 *   fqk_prepass_seed_hits    *n = number of windows equal to seed; the first min(*n, cap) found are written to
 *                            read_idx[] / pos[] in no particular order
 */
int  fqk_prepass_kmer_counts(int device, const fq_ascii_batch* reads, int keylen, int first_pos, int shift_tail,
                             uint64_t* counts);
int  fqk_prepass_seed_hits(int device, const fq_ascii_batch* reads, int keylen, int first_pos, int shift_tail, uint32_t seed,
                           uint32_t* read_idx, uint32_t* pos, uint32_t cap, uint32_t* n);

/*
 * fqk_prepass_kmer_counts_static -- no counterpart in the reference (the reference counts on the CPU and has nothing to
 * allocate on a GPU); it exists for callers of THIS library that must not allocate device memory.
 *
 * Same counts as fqk_prepass_kmer_counts, but the call allocates nothing: no cudaMalloc / cudaFree and no staging
 * buffer.  fqk_prepass_kmer_counts takes its working histogram from cudaMalloc on every call; that is a problem for
 *   - a thread that runs the pre-pass while other threads of the process have packs in flight on the same GPU
 *     (fqk_submit_*): cudaMalloc / cudaFree hold up every stream of the device, so each call would stall all workers;
 *   - a process whose HBM is already fully handed out (its own pool, long-lived contexts), where cudaMalloc fails.
 * Here the histogram is an array built into libfqk.so: 4 MiB of device memory on EVERY GPU that loads the library,
 * held until the process exits.  It has room for 4^10 bins, so keylen is limited to FQK_PREPASS_STATIC_MAX_KEYLEN
 * (the reference's Evaluator uses 10); a longer key returns FQK_E_LIMIT and fqk_last_error(NULL) says why.
 *
 * Everything the caller passes is DEVICE memory it owns: seq_dev = the reads' characters, off_dev[n_reads + 1] = start
 * of each read in seq_dev (off_dev[n_reads] = end of the last), counts_dev[4^keylen] = output (overwritten).
 * `stream` is a cudaStream_t (NULL = the default stream); the work is queued on it and the call returns without
 * waiting: counts_dev is complete once that stream has been synchronised.  `device` becomes the current device.
 * The first call on each device also looks up the array's device address (one-time set-up): make that first call
 * before any stream capture on that device.
 * All calls on a GPU share the one array and it keeps the previous call's contents; every call clears the bins it uses
 * first.  Calls that overlap on the GPU (two streams, two threads) race on it and both get wrong counts: keep them on
 * one stream or synchronise between them.
 */
#define FQK_PREPASS_STATIC_MAX_KEYLEN 10
int  fqk_prepass_kmer_counts_static(int device, uint32_t n_reads, const char* seq_dev, const uint64_t* off_dev, int keylen,
                                    int first_pos, int shift_tail, uint64_t* counts_dev, void* stream);

/* pinned host memory helpers for callers without a CUDA toolchain */
void* fqk_host_alloc(size_t bytes);
void  fqk_host_free(void* p);

#ifdef __cplusplus
}
#endif
#endif /* FQK_H */
