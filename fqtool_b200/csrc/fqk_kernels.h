// fqk_kernels.h -- kernel declarations shared by fqk_kernels.cu / fqk_fused.cu / fqk_api.cu
#pragma once
#include "fqk_trim.cuh"

__global__ void k_trim_se(const __grid_constant__ DevParams D, DevBatch B,
                          fq_read_result* __restrict__ out, uint64_t* __restrict__ flat);
__global__ void k_trim_pe(const __grid_constant__ DevParams D, DevBatch B1, DevBatch B2,
                          fq_read_result* __restrict__ out1, fq_read_result* __restrict__ out2,
                          fq_pair_result* __restrict__ outp, uint64_t* __restrict__ flat, uint32_t* __restrict__ maxlen2,
                          fq_correction* __restrict__ corr, uint32_t* __restrict__ ncorr, uint32_t corr_cap);
__global__ void k_stats_se(const __grid_constant__ DevParams D, DevBatch B,
                           const fq_read_result* __restrict__ res, uint64_t* __restrict__ flat, int Cs);
__global__ void k_stats_pe(const __grid_constant__ DevParams D, DevBatch B1, DevBatch B2,
                           const fq_read_result* __restrict__ res1, const fq_read_result* __restrict__ res2,
                           const fq_pair_result* __restrict__ resp, uint64_t* __restrict__ flat, int Cs, int phase);
__global__ void k_dup(const __grid_constant__ DevParams D, DevBatch B1, DevBatch B2, int paired,
                      unsigned long long* tab128, unsigned long long* first64);
__global__ void k_dup_finish(uint64_t bins, const unsigned long long* tab128, const unsigned long long* first64,
                             uint64_t* dups, uint32_t* counts, uint8_t* gcs);
__global__ void k_pack_ascii(uint32_t n, const char* __restrict__ seq, const char* __restrict__ qual,
                             const uint32_t* __restrict__ src_off, const uint32_t* __restrict__ dst_off,
                             const uint16_t* __restrict__ lengths, uint8_t* __restrict__ bases2,
                             uint8_t* __restrict__ qout, uint32_t* __restrict__ err, const uint32_t* __restrict__ qsrc_off = nullptr);
__global__ void k_kmer_se(const __grid_constant__ DevParams D, DevBatch B, const fq_read_result* __restrict__ recs,
                          uint64_t* __restrict__ flat, int use_smem);
