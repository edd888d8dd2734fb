// fqk_dev.cuh -- device-side data model shared by the kernels (sm_100a only).
//
// Layout in HBM (DESIGN.md section "Data layout"): a batch is four flat arrays
//   offsets[n] u32, lengths[n] u16, bases2[total/4] u8 (2-bit, A0 T1 C2 G3, N as 0),
//   qual[total] u8 (Phred+33 in bits 0-6, N flag in bit 7)
// every read starts at a multiple of FQK_PAD = 16 bases, i.e. on a 16-byte
// quality boundary and a 4-byte base boundary.
#pragma once
#include <cstdint>
#ifndef FQK_EMU
#include <cuda_runtime.h>
#endif
#include "../../include/fqk.h"

struct DevBatch {
    uint32_t n;
    uint64_t first_index;
    uint64_t total;        // padded extent in bases (multiple of FQK_BATCH_PAD)
    const uint32_t* offsets;
    const uint16_t* lengths;
    const uint8_t* bases2;
    const uint8_t* qual;
    uint32_t max_len;      // longest read of the batch, 0 = unknown (fq_batch::max_len)
};

// base classes used on the device: 0 A, 1 T, 2 C, 3 G, 4 N
// (the order of PolyX's atcgBases, /root/reference/src/polyx.cpp:49, and of
// the 2-bit k-mer code, /root/reference/src/evaluator.cpp:9-23)
#define CLS_N 4

// Stats bin = base & 7 (/root/reference/src/stats.cpp:249): A 1, T 4, C 3, G 7, N 6
__device__ __forceinline__ int cls_to_bin(int cls){
    return (0x67341 >> (cls * 4)) & 0xF;
}

// A view of one read inside a batch (positions relative to the view start).
struct ReadView {
    const uint8_t* q;      // quality bytes of base 0 of the ORIGINAL read (16-byte aligned)
    const uint8_t* b2;     // packed bases of the ORIGINAL read (4-byte aligned)
    int start;             // first base of the current view in the original read
    int len;               // current length
    __device__ __forceinline__ int qual(int i) const { return q[start + i] & 0x7F; }
    __device__ __forceinline__ bool isN(int i) const { return (q[start + i] & 0x80) != 0; }
    __device__ __forceinline__ int code(int i) const {
        int p = start + i;
        return (b2[p >> 2] >> ((p & 3) * 2)) & 3;
    }
    __device__ __forceinline__ int cls(int i) const { return isN(i) ? CLS_N : code(i); }
};

__device__ __forceinline__ ReadView make_view(const DevBatch& b, uint32_t r){
    ReadView v;
    uint32_t off = b.offsets[r];
    v.q = b.qual + off;
    v.b2 = b.bases2 + (off >> 2);
    v.start = 0;
    v.len = b.lengths[r];
    return v;
}

// Read::resize, /root/reference/src/read.h:181-187
__device__ __forceinline__ void view_resize(ReadView& v, int len){
    if(len > v.len || len < 0) return;
    v.len = len;
}

// Accumulator layout on the device = fq_stats_layout (include/fqk.h)
struct DevLayout {
    fq_stats_layout L;
};

__device__ __forceinline__ void add_u64(uint64_t* p, uint64_t v){
    atomicAdd(reinterpret_cast<unsigned long long*>(p), static_cast<unsigned long long>(v));
}
