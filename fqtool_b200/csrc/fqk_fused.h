// fqk_fused.h -- the fused single-end kernel (trim + filter + pre/post stats in one pass)
#pragma once
#include "fqk_trim.cuh"
void fused_configure(const cudaDeviceProp& prop);
bool fused_supported_se(const DevParams& D);
// returns the number of kernels launched, or < 0 on error
int  fused_launch_se(const DevParams& D, const DevBatch& B, fq_read_result* out, uint64_t* flat, int sms, cudaStream_t st);
