// fqk_fused.cu -- placeholder until the fused SE kernel lands (generic kernels are used)
#include "fqk_fused.h"
void fused_configure(const cudaDeviceProp&){}
bool fused_supported_se(const DevParams&){ return false; }
int  fused_launch_se(const DevParams&, const DevBatch&, fq_read_result*, uint64_t*, int, cudaStream_t){ return -1; }
