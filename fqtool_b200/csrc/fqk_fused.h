// fqk_fused.h -- fused single-end kernel (trim + filter + pre/post stats in one HBM pass); see fqk_fused.cu
#pragma once
#include <cstdlib>
#include "fqk_trim.cuh"

// A/B and timing switches (FQK_V3_DEBUG, FQK_V4_DEBUG, FQK_V4_T ... : some of them SKIP work) exist only in a library built
// with -DFQK_TUNING_SWITCHES (FQK_NVCC_EXTRA=-DFQK_TUNING_SWITCHES python -m fqtool_b200.build --force, as tools/gpu_pe_*.sh
// expect).  The shipped libfqk.so reads no environment variable on its kernel path.
#ifdef FQK_TUNING_SWITCHES
inline const char* fqk_tuning_env(const char* name){ return getenv(name); }
#else
inline const char* fqk_tuning_env(const char*){ return nullptr; }
#endif
void fused_configure(const cudaDeviceProp& prop);
bool fused_supported_se(const DevParams& D, const DevBatch& B);
// returns the number of kernels launched, or < 0 on error
// variant 0: decoupled warp-specialised kernel k_se_v3 (default; k-mers fall to k_fused_se_ws), 2: single-role kernel,
// 3: k_fused_se_ws
int  fused_launch_se(const DevParams& D, const DevBatch& B, fq_read_result* out, uint64_t* flat, int sms, cudaStream_t st,
                     int variant);

// paired-end per-cycle statistics (SWAR kernel in statistics-only mode) from the records of k_trim_pe
bool ws_stats_supported_pe(const DevParams& D, const DevBatch& B1, const DevBatch& B2);
int  ws_stats_launch_pe(const DevParams& D, const DevBatch& B1, const DevBatch& B2, const fq_read_result* r1,
                        const fq_read_result* r2, const fq_pair_result* rp, uint64_t* flat, int sms, cudaStream_t st,
                        const uint32_t* maxlen2 /* device: longest raw read of each mate, or nullptr */);

// fused paired-end kernel k_pe_v4 (fqk_pe4.cuh): trim + filter + all four Stats objects in one pass
bool pe4_supported(const DevParams& D, const DevBatch& B1, const DevBatch& B2);
int  pe4_launch(const DevParams& D, const DevBatch& B1, const DevBatch& B2, fq_read_result* o1, fq_read_result* o2,
                fq_pair_result* op, uint64_t* flat, int sms, cudaStream_t st);
