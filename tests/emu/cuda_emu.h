// cuda_emu.h -- TEST INFRASTRUCTURE.  Just enough of the CUDA device vocabulary to compile the per-warp device
// logic of fqtool_b200/csrc (fqk_dev.cuh, fqk_trim.cuh, fqk_pe.cuh, fqk_smem.cuh, the trim role of fqk_pe4.cuh)
// with g++ and run ONE warp of it on the CPU: 32 lanes are 32 fibers (ucontext) that run in lock step; every warp
// primitive (__shfl_sync, __ballot_sync, __reduce_*_sync, ...) is a rendezvous of all 32 fibers.  "Shared
// memory" is a byte array addressed by 32-bit offsets.  Nothing here is on the product path: the product is the
// CUDA build of the same headers; this exists so that the CPU test suite (no GPU in the build container) can
// check the device logic against the oracle.
#pragma once
#define FQK_EMU 1
#include <ucontext.h>
#include <algorithm>
#include <cassert>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

// keep <cuda_runtime.h> out: fqk_dev.cuh includes it, we provide the few types it would have given us
#define __CUDA_RUNTIME_H__
#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __noinline__
#define __restrict__
#define __grid_constant__
#define __launch_bounds__(...)
struct uint2 { uint32_t x, y; };
struct uint4 { uint32_t x, y, z, w; };
static inline uint2 make_uint2(uint32_t x, uint32_t y){ return uint2{x, y}; }
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w){ return uint4{x, y, z, w}; }
using std::min; using std::max;
static inline int min(int a, unsigned b){ return (int)std::min<long long>(a, b); }
static inline uint32_t min(uint32_t a, int b){ return (uint32_t)std::min<long long>(a, b); }

// ---- scalar intrinsics
static inline int __popc(uint32_t x){ return __builtin_popcount(x); }
static inline int __ffs(int x){ return __builtin_ffs(x); }
static inline int __clz(int x){ return x ? __builtin_clz((unsigned)x) : 32; }
static inline uint32_t __brev(uint32_t x){
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1); x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0F0F0F0Fu) | ((x & 0x0F0F0F0Fu) << 4); return __builtin_bswap32(x);
}
static inline uint32_t __funnelshift_r(uint32_t lo, uint32_t hi, uint32_t sh){
    sh &= 31; return sh ? (uint32_t)((((uint64_t)hi << 32) | lo) >> sh) : lo;
}
static inline uint32_t __byte_perm(uint32_t a, uint32_t b, uint32_t s){
    const uint64_t v = ((uint64_t)b << 32) | a; uint32_t r = 0;
    for(int i = 0; i < 4; ++i){ const uint32_t sel = (s >> (4 * i)) & 0xF; uint32_t by = (uint32_t)(v >> (8 * (sel & 7))) & 0xFF; if(sel & 8) by = (by & 0x80) ? 0xFF : 0; r |= by << (8 * i); }
    return r;
}
static inline uint32_t __dp4a(uint32_t a, uint32_t b, uint32_t c){
    for(int i = 0; i < 4; ++i) c += ((a >> (8 * i)) & 0xFF) * ((b >> (8 * i)) & 0xFF);
    return c;
}
static inline uint32_t dp4(uint32_t a, uint32_t b, uint32_t c){ return __dp4a(a, b, c); }
template <class T> static inline T atomicAdd(T* p, T v){ T o = *p; *p = o + v; return o; }
static inline uint32_t atomicOr(uint32_t* p, uint32_t v){ uint32_t o = *p; *p = o | v; return o; }
static inline uint32_t atomicMax(uint32_t* p, uint32_t v){ uint32_t o = *p; *p = std::max(o, v); return o; }

// ---- "shared memory"
static uint8_t* emu_smem = nullptr;                  // set by the harness (size EMU_SMEM_BYTES)
static inline uint32_t lds32(uint32_t a){ uint32_t v; memcpy(&v, emu_smem + a, 4); return v; }
static inline uint32_t lds16(uint32_t a){ uint16_t v; memcpy(&v, emu_smem + a, 2); return v; }
static inline uint32_t lds8(uint32_t a){ return emu_smem[a]; }
static inline uint2 lds64(uint32_t a){ uint2 v; memcpy(&v, emu_smem + a, 8); return v; }
static inline uint4 lds128(uint32_t a){ assert((a & 15) == 0); uint4 v; memcpy(&v, emu_smem + a, 16); return v; }
static inline void sts32(uint32_t a, uint32_t v){ memcpy(emu_smem + a, &v, 4); }
static inline void sts64(uint32_t a, uint32_t x, uint32_t y){ memcpy(emu_smem + a, &x, 4); memcpy(emu_smem + a + 4, &y, 4); }
static inline void sts16(uint32_t a, uint32_t v){ uint16_t h = (uint16_t)v; memcpy(emu_smem + a, &h, 2); }
static inline void reds_add(uint32_t a, uint32_t v){ sts32(a, lds32(a) + v); }
static inline uint32_t atoms_add(uint32_t a, uint32_t v){ const uint32_t o = lds32(a); sts32(a, o + v); return o; }
static inline uint32_t smem_u32(const void* p){ return (uint32_t)((const uint8_t*)p - emu_smem); }
static inline uint8_t* smem_ptr(uint32_t a){ return emu_smem + a; }

// ---- one warp = 32 fibers in lock step
namespace emu {
constexpr int LANES = 32;
constexpr size_t STACK = 512 * 1024;
static ucontext_t sched_ctx, lane_ctx[LANES];
static uint8_t* stacks = nullptr;
static int cur = -1;
static bool done[LANES];
static uint32_t xch[2][LANES]; static int xsite[2][LANES]; static int seq[LANES];
static void (*body)(int lane, void* arg); static void* body_arg;

static void trampoline(int lane){ body(lane, body_arg); done[lane] = true; swapcontext(&lane_ctx[lane], &sched_ctx); }
// rendezvous: returns the buffer that holds every lane's value of this primitive
static inline int exchange(uint32_t v, int site){
    const int l = cur, b = seq[l] & 1; seq[l]++;
    xch[b][l] = v; xsite[b][l] = site;
    swapcontext(&lane_ctx[l], &sched_ctx);
    for(int k = 0; k < LANES; ++k) if(xsite[b][k] != site){ fprintf(stderr, "emu: divergent warp primitive (lane %d at %d, lane %d at %d)\n", l, site, k, xsite[b][k]); abort(); }
    return b;
}
static void run_warp(void (*fn)(int, void*), void* arg){
    if(!stacks) stacks = (uint8_t*)malloc(STACK * LANES);
    body = fn; body_arg = arg;
    for(int l = 0; l < LANES; ++l){
        done[l] = false; seq[l] = 0;
        getcontext(&lane_ctx[l]);
        lane_ctx[l].uc_stack.ss_sp = stacks + STACK * l; lane_ctx[l].uc_stack.ss_size = STACK; lane_ctx[l].uc_link = &sched_ctx;
        makecontext(&lane_ctx[l], (void (*)())trampoline, 1, l);
    }
    for(;;){
        int alive = 0;
        for(int l = 0; l < LANES; ++l){ if(done[l]) continue; ++alive; cur = l; swapcontext(&sched_ctx, &lane_ctx[l]); }
        if(!alive) break;
        int nd = 0; for(int l = 0; l < LANES; ++l) nd += done[l];
        if(nd != 0 && nd != LANES){ fprintf(stderr, "emu: some lanes left the warp function while others wait in a primitive\n"); abort(); }
    }
    cur = -1;
}
}  // namespace emu

static inline uint32_t __shfl_sync(uint32_t, uint32_t v, int src){ const int b = emu::exchange(v, 1); return emu::xch[b][src & 31]; }
static inline int __shfl_sync(uint32_t m, int v, int src){ return (int)__shfl_sync(m, (uint32_t)v, src); }
static inline uint32_t __shfl_down_sync(uint32_t, uint32_t v, int d){ const int l = emu::cur, b = emu::exchange(v, 2); return l + d < 32 ? emu::xch[b][l + d] : v; }
static inline uint32_t __shfl_xor_sync(uint32_t, uint32_t v, int x){ const int l = emu::cur, b = emu::exchange(v, 3); return emu::xch[b][(l ^ x) & 31]; }
static inline uint32_t __ballot_sync(uint32_t, bool p){ const int b = emu::exchange(p ? 1u : 0u, 4); uint32_t r = 0; for(int k = 0; k < 32; ++k) r |= emu::xch[b][k] << k; return r; }
static inline bool __all_sync(uint32_t m, bool p){ return __ballot_sync(m, p) == 0xFFFFFFFFu; }
static inline bool __any_sync(uint32_t m, bool p){ return __ballot_sync(m, p) != 0u; }
static inline uint32_t __reduce_add_sync(uint32_t, uint32_t v){ const int b = emu::exchange(v, 5); uint32_t r = 0; for(int k = 0; k < 32; ++k) r += emu::xch[b][k]; return r; }
static inline uint32_t __reduce_min_sync(uint32_t, uint32_t v){ const int b = emu::exchange(v, 6); uint32_t r = ~0u; for(int k = 0; k < 32; ++k) r = std::min(r, emu::xch[b][k]); return r; }
static inline uint32_t __reduce_max_sync(uint32_t, uint32_t v){ const int b = emu::exchange(v, 7); uint32_t r = 0; for(int k = 0; k < 32; ++k) r = std::max(r, emu::xch[b][k]); return r; }
static inline void __syncwarp(){ emu::exchange(0, 8); }
static inline void __threadfence_block(){}
