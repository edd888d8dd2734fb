// fqk_prepass.cu -- device part of the adapter auto-detection pre-pass:
//   Evaluator::evaluateAdapterSeq   /root/reference/src/evaluator.cpp:229-390  (10-mer counting over the first 256 K reads)
//   Evaluator::getAdapterWithSeed   /root/reference/src/evaluator.cpp:392-425  (every window equal to the seed key)
// The reference walks every read with a rolling 2-bit key (Evaluator::seq2int, src/evaluator.cpp:3-49: A0 T1 C2 G3,
// -1 as soon as the window holds another character, recomputed from scratch after that), i.e. window [pos, pos+k)
// is counted iff all its k characters are in {A,T,C,G}.  Positions run from first_pos to len - k - shift_tail.
// Here: one warp per read, a lane per window, the key built from the ASCII characters of the window (k <= 15, the
// reads are the FASTQ characters a parser holds: fq_ascii_batch).  What the host keeps: ranking the keys and the
// NucleotideTree walk over the few reads that carry the seed (string work on a handful of reads).
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include "../../include/fqk.h"

namespace {

__device__ __forceinline__ int code_of_char(unsigned char c){            // Evaluator::seq2int's switch
    return c == 'A' ? 0 : c == 'T' ? 1 : c == 'C' ? 2 : c == 'G' ? 3 : -1;
}
// key of window [pos, pos+k) of a read, -1 when a character is outside ATCG
__device__ __forceinline__ int window_key(const char* s, int pos, int k){
    int key = 0;
    for(int i = 0; i < k; ++i){
        const int c = code_of_char((unsigned char)s[pos + i]);
        if(c < 0) return -1;
        key = (key << 2) | c;
    }
    return key;
}

// seed < 0: count every window into counts[4^k]; seed >= 0: append (read, pos) of the windows equal to seed
__global__ void k_prepass_kmers(uint32_t n, const char* __restrict__ seq, const uint64_t* __restrict__ off, int k, int first_pos,
                                int shift_tail, int seed, uint32_t* __restrict__ counts, uint32_t* __restrict__ hit_read,
                                uint32_t* __restrict__ hit_pos, uint32_t cap, uint32_t* __restrict__ nhits){
    const int lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for(uint32_t r = warp; r < n; r += nwarps){
        const uint64_t o = off[r];
        const int len = (int)(off[r + 1] - o);
        const int last = len - k - shift_tail;                        // for(pos = first_pos; pos <= len - keylen - shiftTail; ++pos)
        const char* s = seq + o;
        for(int pos = first_pos + lane; pos <= last; pos += 32){
            const int key = window_key(s, pos, k);
            if(key < 0) continue;
            if(seed < 0) atomicAdd(&counts[key], 1u);
            else if(key == seed){
                const uint32_t i = atomicAdd(nhits, 1u);
                if(i < cap){ hit_read[i] = r; hit_pos[i] = (uint32_t)pos; }
            }
        }
    }
}

struct Tmp {
    void* p = nullptr;
    ~Tmp(){ if(p) cudaFree(p); }
    bool alloc(size_t b){ return cudaMalloc(&p, b ? b : 1) == cudaSuccess; }
};

int check_batch(const fq_ascii_batch* b, int k, int first_pos, int shift_tail){
    if(!b || !b->off || (!b->seq && b->n_reads) || k < 1 || k > 15 || first_pos < 0 || shift_tail < 0) return FQK_E_INVALID;
    return 0;
}
// characters and offsets of the batch to the device (offsets rebased to the first read)
int upload(const fq_ascii_batch* b, Tmp& dseq, Tmp& doff, cudaStream_t st){
    const uint32_t n = b->n_reads;
    const uint64_t chars = b->off[n] - b->off[0];
    if(!dseq.alloc(chars + 16) || !doff.alloc((size_t)(n + 1) * 8)) return FQK_E_NOMEM;
    if(cudaMemcpyAsync(dseq.p, b->seq + b->off[0], chars, cudaMemcpyHostToDevice, st) != cudaSuccess) return FQK_E_CUDA;
    if(b->off[0] == 0){
        if(cudaMemcpyAsync(doff.p, b->off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, st) != cudaSuccess) return FQK_E_CUDA;
    }else{
        uint64_t* h = new uint64_t[n + 1];
        for(uint32_t i = 0; i <= n; ++i) h[i] = b->off[i] - b->off[0];
        const cudaError_t e = cudaMemcpyAsync(doff.p, h, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, st);
        cudaStreamSynchronize(st);
        delete[] h;
        if(e != cudaSuccess) return FQK_E_CUDA;
    }
    return 0;
}
int grid_for(uint32_t n){ return (int)((n + 7) / 8 < 148 * 8 ? (n + 7) / 8 : 148 * 8); }

} // namespace

extern "C" int fqk_prepass_kmer_counts(int device, const fq_ascii_batch* reads, int keylen, int first_pos, int shift_tail,
                                       uint64_t* counts){
    int rc = check_batch(reads, keylen, first_pos, shift_tail);
    if(rc || !counts) return FQK_E_INVALID;
    int ndev = 0;
    if(cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return FQK_E_CUDA;   // no CPU fallback
    if(cudaSetDevice(device) != cudaSuccess) return FQK_E_CUDA;
    const size_t bins = (size_t)1 << (2 * keylen);
    for(size_t i = 0; i < bins; ++i) counts[i] = 0;
    if(reads->n_reads == 0) return 0;
    Tmp dseq, doff, dcnt;
    if(!dcnt.alloc(bins * 4)) return FQK_E_NOMEM;
    rc = upload(reads, dseq, doff, 0); if(rc) return rc;
    if(cudaMemsetAsync(dcnt.p, 0, bins * 4, 0) != cudaSuccess) return FQK_E_CUDA;
    k_prepass_kmers<<<grid_for(reads->n_reads), 256>>>(reads->n_reads, (const char*)dseq.p, (const uint64_t*)doff.p, keylen, first_pos,
                                                      shift_tail, -1, (uint32_t*)dcnt.p, nullptr, nullptr, 0, nullptr);
    uint32_t* h = new uint32_t[bins];
    const cudaError_t e = cudaMemcpy(h, dcnt.p, bins * 4, cudaMemcpyDeviceToHost);
    if(e == cudaSuccess) for(size_t i = 0; i < bins; ++i) counts[i] = h[i];
    delete[] h;
    return e == cudaSuccess ? 0 : FQK_E_CUDA;
}

extern "C" int fqk_prepass_seed_hits(int device, const fq_ascii_batch* reads, int keylen, int first_pos, int shift_tail, uint32_t seed,
                                     uint32_t* read_idx, uint32_t* pos, uint32_t cap, uint32_t* n){
    int rc = check_batch(reads, keylen, first_pos, shift_tail);
    if(rc || !n || (cap && (!read_idx || !pos)) || seed >= (1u << (2 * keylen))) return FQK_E_INVALID;
    *n = 0;
    int ndev = 0;
    if(cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return FQK_E_CUDA;
    if(cudaSetDevice(device) != cudaSuccess) return FQK_E_CUDA;
    if(reads->n_reads == 0) return 0;
    Tmp dseq, doff, dr, dp, dn;
    if(!dr.alloc((size_t)cap * 4) || !dp.alloc((size_t)cap * 4) || !dn.alloc(4)) return FQK_E_NOMEM;
    rc = upload(reads, dseq, doff, 0); if(rc) return rc;
    if(cudaMemsetAsync(dn.p, 0, 4, 0) != cudaSuccess) return FQK_E_CUDA;
    k_prepass_kmers<<<grid_for(reads->n_reads), 256>>>(reads->n_reads, (const char*)dseq.p, (const uint64_t*)doff.p, keylen, first_pos,
                                                      shift_tail, (int)seed, nullptr, (uint32_t*)dr.p, (uint32_t*)dp.p, cap, (uint32_t*)dn.p);
    uint32_t cnt = 0;
    if(cudaMemcpy(&cnt, dn.p, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return FQK_E_CUDA;
    *n = cnt;                                                             // the total; at most cap entries are written (unordered)
    const uint32_t m = cnt < cap ? cnt : cap;
    if(m){
        if(cudaMemcpy(read_idx, dr.p, (size_t)m * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return FQK_E_CUDA;
        if(cudaMemcpy(pos, dp.p, (size_t)m * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return FQK_E_CUDA;
    }
    return 0;
}
