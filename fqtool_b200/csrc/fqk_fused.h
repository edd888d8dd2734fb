// fqk_fused.h -- SWAR per-cycle statistics kernel for single-end batches (see fqk_fused.cu)
#pragma once
#include "fqk_trim.cuh"
void fused_configure(const cudaDeviceProp& prop);
bool swar_stats_supported_se(const DevParams& D);
// returns the number of kernels launched, or < 0 on error
int  swar_stats_launch_se(const DevParams& D, const DevBatch& B, const fq_read_result* res, uint64_t* flat, int sms,
                          cudaStream_t st);

// fused trim + filter + stats kernel (one pass over HBM, bulk-async staged tiles)
bool fused_supported_se(const DevParams& D, const DevBatch& B);
int  fused_launch_se(const DevParams& D, const DevBatch& B, fq_read_result* out, uint64_t* flat, int sms, cudaStream_t st);
