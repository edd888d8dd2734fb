// fqk_fused.h -- fused single-end kernel (trim + filter + pre/post stats in one HBM pass); see fqk_fused.cu
#pragma once
#include "fqk_trim.cuh"
void fused_configure(const cudaDeviceProp& prop);
bool fused_supported_se(const DevParams& D, const DevBatch& B);
// returns the number of kernels launched, or < 0 on error
// variant 0: warp-specialised kernel (default), 2: single-role kernel
int  fused_launch_se(const DevParams& D, const DevBatch& B, fq_read_result* out, uint64_t* flat, int sms, cudaStream_t st,
                     int variant);
