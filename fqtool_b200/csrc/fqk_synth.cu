// fqk_synth.cu -- device generators of synthetic reads (see include/fqk_synth.h).
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/fqk.h"
#include "../../include/fqk_synth.h"

__device__ __forceinline__ uint64_t mix64(uint64_t z){
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
// counter-based stream: (seed, read, lane) -> sequence indexed by ctr
struct Rng {
    uint64_t key;
    __device__ Rng(uint64_t seed, uint64_t read, uint32_t lane){ key = mix64(mix64(seed ^ (read * 0xD1342543DE82EF95ull)) + lane); }
    __device__ __forceinline__ uint64_t u64(uint32_t ctr) const { return mix64(key + (uint64_t)ctr * 0x2545F4914F6CDD1Dull); }
    __device__ __forceinline__ float uni(uint32_t ctr) const { return (float)((u64(ctr) >> 40) + 1) * (1.0f / 16777217.0f); }
};
__device__ __forceinline__ float gauss(const Rng& r, uint32_t ctr){
    const uint64_t x = r.u64(ctr);
    const float u1 = (float)((x >> 40) + 1) * (1.0f / 16777217.0f);
    const float u2 = (float)((x >> 8) & 0xFFFFFF) * (1.0f / 16777216.0f);
    return sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
}

__constant__ char c_ad1[34] = "AGATCGGAAGAGCACACGTCTGAACTCCAGTCA";   // TruSeq R1, src/knownadapters.h:14
__constant__ char c_ad2[34] = "AGATCGGAAGAGCGTCGTGTAGGGAAAGAGTGT";   // TruSeq R2, src/knownadapters.h:15
__device__ __forceinline__ int code_of(char c){ return c == 'A' ? 0 : c == 'T' ? 1 : c == 'C' ? 2 : 3; }

// write 16 bases (codes 0..3, or 4 for N) + qualities of one read chunk
__device__ __forceinline__ void store16(uint8_t* qual, uint8_t* bases2, uint64_t pos, const int* cls, const int* q, int valid){
    uint32_t w[4] = {0, 0, 0, 0}, b = 0;
#pragma unroll
    for(int i = 0; i < 16; ++i){
        uint32_t qq = 0;
        if(i < valid){
            qq = (uint32_t)(q[i] + 33);
            if(cls[i] == 4) qq |= 0x80; else b |= (uint32_t)cls[i] << (2 * i);
        }
        w[i >> 2] |= qq << (8 * (i & 3));
    }
    *reinterpret_cast<uint4*>(qual + pos) = make_uint4(w[0], w[1], w[2], w[3]);
    *reinterpret_cast<uint32_t*>(bases2 + (pos >> 2)) = b;
}

__global__ void k_synth_se(int model, uint64_t seed, uint64_t first, uint32_t n, int L,
                           uint32_t* off, uint16_t* len, uint8_t* bases2, uint8_t* qual){
    const int Lp = (L + 15) / 16 * 16;
    for(uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x){
        const uint64_t g = first + r;
        const uint64_t base = (uint64_t)r * Lp;
        off[r] = (uint32_t)base; len[r] = (uint16_t)L;
        Rng rq(seed, g, 1), rm(seed, g, 2);
        uint64_t src = g;                                   // index whose bases this read carries
        int coll_start = L, ad_pos = L, px_len = 0, px_base = 0;
        float q0 = 38.0f, qslope = 8.0f;
        if(model == FQK_SYNTH_SE_HISEQ){
            if(rm.uni(0) < 0.08f){
                int k = 1 + (int)(logf(rm.uni(1)) / logf(0.8f));   // geometric(0.2)
                coll_start = L - min(k, 15);
            }
        }else{
            q0 = 37.0f; qslope = 6.0f;
            if(g > 0 && rm.uni(0) < 0.10f) src = rm.u64(1) % g;     // duplicate of an earlier read
            Rng rs(seed, src, 2);
            if(rs.uni(2) < 0.12f) ad_pos = 20 + (int)(rs.u64(3) % (uint64_t)max(L - 20, 1));
            else if(rs.uni(4) < 0.03f / 0.88f){ px_len = 10 + (int)(rs.u64(5) % 31); px_base = (int)(rs.u64(6) & 3); }
        }
        Rng rb(seed, src, 0), rx(seed, src, 3);
        for(int c0 = 0; c0 < Lp; c0 += 16){
            int cls[16], q[16];
#pragma unroll
            for(int j = 0; j < 16; ++j){
                const int i = c0 + j;
                int b = (int)(rb.u64(i) >> 62);
                if(i >= ad_pos) b = (i - ad_pos < 33) ? code_of(c_ad1[i - ad_pos]) : 0;          // adapter then polyA
                else if(px_len && i >= L - px_len){ b = (rx.uni(i) < 0.1f) ? (int)(rx.u64(i + 4096) & 3) : px_base; }
                const float mean = q0 - qslope * (float)i / (float)max(L - 1, 1);
                int qq = (int)rintf(mean + 4.0f * gauss(rq, i));
                qq = min(41, max(2, qq));
                if(i >= coll_start) qq = 2;
                if(rq.uni(i + 8192) < 0.002f){ b = 4; qq = 2; }
                cls[j] = b; q[j] = qq;
            }
            store16(qual, bases2, base + c0, cls, q, L - c0);
        }
    }
}

// NovaSeq-binned quality {2,12,23,37}, degrading toward the 3' end
__device__ __forceinline__ int nova_q(float u, float t){
    const float p3 = 0.12f * t;
    const float c0 = 0.01f + 0.2f * p3, c1 = c0 + 0.03f + 0.3f * p3, c2 = c1 + 0.08f + 0.5f * p3;
    return u < c0 ? 2 : u < c1 ? 12 : u < c2 ? 23 : 37;
}

__global__ void k_synth_pe(uint64_t seed, uint64_t first, uint32_t n, int L, float fmean, float fsd,
                           uint32_t* off1, uint16_t* len1, uint8_t* b1, uint8_t* q1,
                           uint32_t* off2, uint16_t* len2, uint8_t* b2, uint8_t* q2){
    const int Lp = (L + 15) / 16 * 16;
    for(uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x){
        const uint64_t g = first + r;
        const uint64_t base = (uint64_t)r * Lp;
        off1[r] = off2[r] = (uint32_t)base; len1[r] = len2[r] = (uint16_t)L;
        Rng rf(seed, g, 0), rm(seed, g, 2);
        int F = (int)rintf(fmean + fsd * gauss(rm, 0));
        F = min(700, max(35, F));
        for(int mate = 0; mate < 2; ++mate){
            Rng rq(seed, g, 4 + mate), re(seed, g, 6 + mate);
            uint8_t* qo = mate ? q2 : q1; uint8_t* bo = mate ? b2 : b1;
            const char* ad = mate ? c_ad2 : c_ad1;
            for(int c0 = 0; c0 < Lp; c0 += 16){
                int cls[16], q[16];
#pragma unroll
                for(int j = 0; j < 16; ++j){
                    const int i = c0 + j;
                    int b;
                    if(i < F){
                        const int p = mate ? F - 1 - i : i;                  // read2 = revcomp of the fragment end
                        b = (int)(rf.u64(p) >> 62);
                        if(mate) b ^= 1;                                     // A<->T, C<->G in the 2-bit code
                    }else if(i - F < 33) b = code_of(ad[i - F]);
                    else b = 3;                                              // polyG fill (two-colour chemistry)
                    if(re.uni(i) < 0.003f) b = (int)(re.u64(i + 4096) & 3);  // substitution error
                    int qq = nova_q(rq.uni(i), (float)i / (float)max(L - 1, 1));
                    if(re.uni(i + 8192) < 0.001f){ b = 4; qq = 2; }
                    cls[j] = b; q[j] = qq;
                }
                store16(qo, bo, base + c0, cls, q, L - c0);
            }
        }
    }
}

extern "C" uint64_t fqk_synth_total_bases(uint32_t n, int L){
    const uint64_t t = (uint64_t)n * (uint64_t)((L + 15) / 16 * 16);
    return (t + FQK_BATCH_PAD - 1) / FQK_BATCH_PAD * FQK_BATCH_PAD;
}

extern "C" int fqk_synth_generate(void* stream, int model, uint64_t seed, uint64_t first_index, uint32_t n, int L,
                                  double frag_mean, double frag_sd,
                                  uint32_t* off1, uint16_t* len1, uint8_t* bases1, uint8_t* qual1,
                                  uint32_t* off2, uint16_t* len2, uint8_t* bases2, uint8_t* qual2){
    if(L < 1 || L > FQK_MAX_READ_LEN || fqk_synth_total_bases(n, L) > 0xFFFFFFFFull) return FQK_E_LIMIT;
    if(n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = (int)((n + 127) / 128 < 148 * 16 ? (n + 127) / 128 : 148 * 16);
    if(model == FQK_SYNTH_PE_NOVASEQ){
        if(!off2 || !len2 || !bases2 || !qual2) return FQK_E_INVALID;
        k_synth_pe<<<grid, 128, 0, st>>>(seed, first_index, n, L, (float)frag_mean, (float)frag_sd,
                                         off1, len1, bases1, qual1, off2, len2, bases2, qual2);
    }else if(model == FQK_SYNTH_SE_HISEQ || model == FQK_SYNTH_SE_ADAPTER){
        k_synth_se<<<grid, 128, 0, st>>>(model, seed, first_index, n, L, off1, len1, bases1, qual1);
    }else return FQK_E_INVALID;
    return cudaGetLastError() == cudaSuccess ? 0 : FQK_E_CUDA;
}
