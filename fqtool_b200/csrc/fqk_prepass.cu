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
#include <mutex>
#include <string>
#include <cuda_runtime.h>
#include "../../include/fqk.h"

void fqk_set_thread_error(const char* msg);                               // fqk_api.cu: text behind fqk_last_error(NULL)

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

// ---------------------------------------------------------------------------------------------------------------
// fqk_prepass_kmer_counts_static: the k-mer count of the pre-pass for callers that must not allocate.
//
// fqk_prepass_kmer_counts above takes its 32-bit histogram (the table the lanes add into) from cudaMalloc and gives it
// back with cudaFree on every call.  Two kinds of caller cannot afford that:
//   * a thread that runs the pre-pass while other threads of the same process have packs in flight on this GPU through
//     fqk_submit_* : cudaMalloc and cudaFree wait for and block every stream of the device, so each call would stall the
//     upload / kernel / download overlap of all the workers;
//   * a process that has handed the whole of HBM to its own pool (or to long-lived fqk contexts) and has nothing left
//     for the driver to hand out, so cudaMalloc fails there.
// For them the histogram lives in the library image itself: a file-scope device array that every GPU which loads
// libfqk.so holds from load until exit (4 MiB per GPU, whether or not this entry point is ever called).  4 MiB is
// 4^10 32-bit bins, i.e. key lengths up to FQK_PREPASS_STATIC_MAX_KEYLEN = 10, the length the reference's
// Evaluator uses; longer keys are refused (FQK_E_LIMIT), use fqk_prepass_kmer_counts for them.
// Inputs and the output are the caller's own DEVICE buffers, so nothing is allocated here: no cudaMalloc / cudaFree,
// no host staging buffer, no library routine that allocates.
//
// The array's name is only usable as an address inside device code; on the host it is a stand-in symbol, and a kernel
// handed `g_prepass_static_hist` from host code would fault (or scribble on host memory where the GPU can reach it).
// So the first call on each device asks the runtime for the real address, with that device current
// (cudaGetSymbolAddress), and remembers it per device.  That first call does this one-time set-up; make it before any
// stream capture begins on that device.
//
// All calls on one GPU share the one array and it keeps whatever the previous call left: each call clears the bins it
// is going to use (stream-ordered, before its kernel).  Two calls that overlap on the GPU (different streams, or two
// threads on the same device) race on the bins and both results are garbage: serialise them (same stream, or a
// synchronise between them).
namespace {

constexpr int      kStaticMaxKeylen = FQK_PREPASS_STATIC_MAX_KEYLEN;
constexpr uint32_t kStaticBins      = 1u << (2 * kStaticMaxKeylen);       // 4^10 = 1 Mi bins = 4 MiB
constexpr int      kStaticMaxDevices = 64;

static __device__ __align__(256) uint32_t g_prepass_static_hist[kStaticBins];

uint32_t*  g_static_hist_addr[kStaticMaxDevices] = {nullptr};             // device address of the array, per device
std::mutex g_static_hist_mtx;

// 32-bit bins of the scratch histogram -> the caller's 64-bit counts (the layout fqk_prepass_kmer_counts returns)
__global__ void k_prepass_widen(const uint32_t* __restrict__ hist, uint64_t* __restrict__ counts, uint32_t bins){
    for(uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < bins; i += gridDim.x * blockDim.x) counts[i] = hist[i];
}

int static_fail(int code, const std::string& msg){ fqk_set_thread_error(msg.c_str()); return code; }

// address of the scratch array on `device` (which must be the current device); looked up once per device
int static_hist_address(int device, uint32_t** out){
    std::lock_guard<std::mutex> lk(g_static_hist_mtx);
    if(!g_static_hist_addr[device]){
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_prepass_static_hist);
        if(e != cudaSuccess || !p)
            return static_fail(FQK_E_CUDA, std::string("fqk_prepass_kmer_counts_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        g_static_hist_addr[device] = (uint32_t*)p;
    }
    *out = g_static_hist_addr[device];
    return 0;
}

} // namespace

extern "C" int fqk_prepass_kmer_counts_static(int device, uint32_t n_reads, const char* seq_dev, const uint64_t* off_dev, int keylen,
                                              int first_pos, int shift_tail, uint64_t* counts_dev, void* stream){
    if(keylen > kStaticMaxKeylen)
        return static_fail(FQK_E_LIMIT, "fqk_prepass_kmer_counts_static: keylen " + std::to_string(keylen) + " needs 4^keylen bins, the built-in scratch "
                           "array holds 4^" + std::to_string(kStaticMaxKeylen) + " (keylen <= " + std::to_string(kStaticMaxKeylen) +
                           "); use fqk_prepass_kmer_counts for longer keys");
    if(keylen < 1 || first_pos < 0 || shift_tail < 0 || !counts_dev || !off_dev || (n_reads && !seq_dev))
        return static_fail(FQK_E_INVALID, "fqk_prepass_kmer_counts_static: bad argument (keylen 1..10, first_pos/shift_tail >= 0, device pointers set)");
    int ndev = 0;
    if(cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
        return static_fail(FQK_E_CUDA, "fqk_prepass_kmer_counts_static: no such CUDA device (there is no CPU fallback)");
    if(device >= kStaticMaxDevices)
        return static_fail(FQK_E_LIMIT, "fqk_prepass_kmer_counts_static: device ordinal beyond " + std::to_string(kStaticMaxDevices - 1));
    cudaError_t e = cudaSetDevice(device);
    if(e != cudaSuccess) return static_fail(FQK_E_CUDA, std::string("fqk_prepass_kmer_counts_static: cudaSetDevice: ") + cudaGetErrorString(e));
    uint32_t* hist = nullptr;
    int rc = static_hist_address(device, &hist); if(rc) return rc;
    const cudaStream_t st = (cudaStream_t)stream;
    const uint32_t bins = 1u << (2 * keylen);
    // the array still holds the previous call's bins: clear the ones this call counts into
    e = cudaMemsetAsync(hist, 0, (size_t)bins * 4, st);
    if(e != cudaSuccess) return static_fail(FQK_E_CUDA, std::string("fqk_prepass_kmer_counts_static: cudaMemsetAsync: ") + cudaGetErrorString(e));
    if(n_reads)
        k_prepass_kmers<<<grid_for(n_reads), 256, 0, st>>>(n_reads, seq_dev, off_dev, keylen, first_pos, shift_tail, -1, hist, nullptr, nullptr, 0,
                                                          nullptr);
    k_prepass_widen<<<(int)((bins + 255) / 256 < 148 * 4 ? (bins + 255) / 256 : 148 * 4), 256, 0, st>>>(hist, counts_dev, bins);
    e = cudaGetLastError();
    if(e != cudaSuccess) return static_fail(FQK_E_CUDA, std::string("fqk_prepass_kmer_counts_static: launch: ") + cudaGetErrorString(e));
    return 0;
}
