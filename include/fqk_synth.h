/*
 * fqk_synth.h -- synthetic Illumina-shaped workloads generated directly in HBM
 * in the fq_batch layout.  Benchmark tooling: not part
 * of the reference-facing boundary of fqk.h.  Counter-based RNG keyed by
 * (seed, global read index, mate, position): any chunk of the stream can be
 * generated independently on any rank.
 */
#ifndef FQK_SYNTH_H
#define FQK_SYNTH_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define FQK_SYNTH_SE_HISEQ    2   /* cfg2: SE, HiSeq-like quality decay, 8 % Q2 tail collapse, N p=0.002   */
#define FQK_SYNTH_PE_NOVASEQ  3   /* cfg3/cfg5: PE fragments, adapter read-through + polyG, binned quality */
#define FQK_SYNTH_SE_ADAPTER  4   /* cfg4: SE, 12 % adapter+polyA, 3 % polyX tails, 10 % duplicated reads  */

/* padded bases of n fixed-length reads: n * ceil(L/16)*16 */
uint64_t fqk_synth_total_bases(uint32_t n, int L);

/*
 * Fill device arrays (caller-allocated: offsets[n] u32, lengths[n] u16,
 * bases2[total/4], qual[total]) with n reads of length L, global indices
 * [first_index, first_index+n).  For FQK_SYNTH_PE_NOVASEQ pass both mates
 * (read2 arrays non-NULL) and the fragment-length model; otherwise read2
 * arrays are NULL.  Runs on `stream` (cudaStream_t) without synchronising.
 */
int fqk_synth_generate(void* stream, int model, uint64_t seed, uint64_t first_index, uint32_t n, int L,
                       double frag_mean, double frag_sd,
                       uint32_t* off1, uint16_t* len1, uint8_t* bases1, uint8_t* qual1,
                       uint32_t* off2, uint16_t* len2, uint8_t* bases2, uint8_t* qual2);
#ifdef __cplusplus
}
#endif
#endif
