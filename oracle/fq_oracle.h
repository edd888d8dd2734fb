/*
 * fq_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C) of the reference's per-read QC + trimming hot
 * path, used as the parity checker for the CUDA path.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library; the product (fqtool_b200/) never does.
 *
 * Parity pin: this restatement is checked against the reference's own
 * unmodified C++ (oracle/_ref, built from /root/reference/src by
 * oracle/Makefile) on testdata and synthetic reads (tests/test_oracle_vs_ref.py)
 * and against golden vectors generated from that build (tests/golden/).
 *
 * Inputs are ASCII, exactly what the reference sees: flat seq/qual buffers,
 * read r at [off[r], off[r+1]).  Records and the flat accumulator buffer use
 * the same structs/layout as the C ABI (include/fqk.h).
 */
#ifndef FQ_ORACLE_H
#define FQ_ORACLE_H
#include "../include/fqk.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fqo_ctx fqo_ctx;

int  fqo_layout(const fq_params* p, fq_stats_layout* out);
fqo_ctx* fqo_create(const fq_params* p);
void fqo_destroy(fqo_ctx* c);

int  fqo_process_se(fqo_ctx* c, uint32_t n, uint64_t first_index,
                    const char* seq, const char* qual, const uint64_t* off,
                    fq_read_result* out);
int  fqo_process_pe(fqo_ctx* c, uint32_t n, uint64_t first_index,
                    const char* seq1, const char* qual1, const uint64_t* off1,
                    const char* seq2, const char* qual2, const uint64_t* off2,
                    fq_read_result* out1, fq_read_result* out2, fq_pair_result* outp);
/* bases rewritten by the last fqo_process_pe call (BaseCorrector, -c) */
int  fqo_get_corrections(fqo_ctx* c, fq_correction* out, uint32_t cap, uint32_t* n);
/* copy accumulators (fq_stats_layout.total_u64 entries) */
int  fqo_finalize(fqo_ctx* c, uint64_t* flat);
int  fqo_dup_table(fqo_ctx* c, uint64_t* dups, uint32_t* counts, uint8_t* gc);
/* merge another context's accumulators into c (Stats::merge, FilterResult::merge;
 * dup tables by the (min kmer32, count) rule); used by the multi-thread baseline */
int  fqo_merge(fqo_ctx* c, const fqo_ctx* other);


/* Adapter auto-detection pre-pass (Evaluator::evaluateAdapterSeq, src/evaluator.cpp:229-390): counts[4^keylen] of the
 * windows [pos, pos+keylen), pos = first_pos .. len - keylen - shift_tail, whose characters are all in ATCG (key =
 * Evaluator::seq2int, src/evaluator.cpp:3-49); and the (read, pos) list of the windows equal to `seed`
 * (Evaluator::getAdapterWithSeed, src/evaluator.cpp:392-407), in input order. */
int  fqo_adapter_kmer_counts(uint32_t n, const char* seq, const uint64_t* off, int keylen, int first_pos, int shift_tail,
                             uint64_t* counts);
int  fqo_adapter_seed_hits(uint32_t n, const char* seq, const uint64_t* off, int keylen, int first_pos, int shift_tail,
                           uint32_t seed, uint32_t* read_idx, uint32_t* pos, uint32_t cap, uint32_t* nhits);

#ifdef __cplusplus
}
#endif
#endif
