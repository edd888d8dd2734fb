// Integer pipe micro-benchmark for sm_100a: warp-instructions per cycle per SM for dp4a, imad, lop3, prmt
// and pairwise mixes (to see which share a pipe).  nvcc -arch=sm_100a -o /tmp/mb tools/microbench_int.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define N_ACC 8
template <int OP>
__global__ void k(uint32_t* out, int iters, uint32_t seed){
    uint32_t a[N_ACC], b[N_ACC];
    for(int i = 0; i < N_ACC; ++i){ a[i] = seed * (i + 1) + threadIdx.x; b[i] = seed ^ (i * 77u); }
    const uint32_t m = seed | 0x01010101u;
    for(int it = 0; it < iters; ++it){
#pragma unroll
        for(int i = 0; i < N_ACC; ++i){
            if(OP == 0) asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b[i]), "r"(m));
            if(OP == 1) asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b[i]), "r"(m));
            if(OP == 2) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b[i]), "r"(m));
            if(OP == 3) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b[i]), "r"(m));
            if(OP == 4){ // dp4a + lop3
                asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b[i]), "r"(m));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(m), "r"(m));
            }
            if(OP == 5){ // dp4a + imad
                asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b[i]), "r"(m));
                asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(b[i]) : "r"(m), "r"(m));
            }
            if(OP == 6){ // imad + lop3
                asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b[i]), "r"(m));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(m), "r"(m));
            }
            if(OP == 7) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b[i]));
            if(OP == 8) asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b[i]), "r"(m));
        }
    }
    uint32_t s = 0;
    for(int i = 0; i < N_ACC; ++i) s += a[i] + b[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP>
void run(const char* name, int per_iter){
    uint32_t* out; cudaMalloc(&out, 148 * 1024 * 4);
    const int iters = 20000;
    for(int threads : {128, 256, 512, 1024}){
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        k<OP><<<148, threads>>>(out, 100, 12345u);
        cudaEventRecord(e0);
        k<OP><<<148, threads>>>(out, iters, 12345u);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double winstr = (double)iters * N_ACC * per_iter * (threads / 32);   // per SM
        int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
        printf("%-12s threads/SM %4d: %.3f warp-instr/ns/SM  (%.2f per cycle at %d MHz nominal)\n", name, threads,
               winstr / (ms * 1e6), winstr / (ms * 1e6) / (clk / 1e6), clk / 1000);
    }
    cudaFree(out);
}
int main(){
    run<0>("dp4a.u32", 1); run<8>("dp4a.mixed", 1); run<1>("imad", 1); run<2>("lop3", 1); run<3>("prmt", 1); run<7>("iadd", 1);
    run<4>("dp4a+lop3", 2); run<5>("dp4a+imad", 2); run<6>("imad+lop3", 2);
    return 0;
}
