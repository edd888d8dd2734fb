// fqk_smem.cuh -- shared-memory / barrier / bulk-copy primitives and the per-read helpers that work on reads
// staged in shared memory (addresses are 32-bit shared-window addresses).  Used by the fused kernels of
// fqk_fused.cu (k_fused_se_ws, k_se_v3, k_pe_v4).
//
// With -DFQK_EMU the same helpers compile with g++ for tests/emu (a lock-step CPU emulation of one warp of the
// trim role, used by the CPU test suite to check the device logic against the oracle without a GPU): "shared
// memory" is then a byte array and the barrier / bulk-copy primitives are not available.
#pragma once
#include "fqk_trim.cuh"

// statistics vocabulary of the fused kernels: kinds of per-cycle counters and sets of bases
constexpr int KN = 0, KN20 = 1, KN30 = 2, KQS = 3;                 // kinds
constexpr int S_ALL = 0, S_B0 = 1, S_B1 = 2, S_B01 = 3, S_N = 4;   // sets
constexpr int NSETS = 5;
// per-read descriptor of a tile, 8 bytes: x = (offset in tile)/16 | len << 17, y = start | plen << 13 | kind << 26
// kind: 0 common (passing, untrimmed), 1 raw only, 2 raw + shifted kept part for post1, 3 tail of a merged read
struct Desc { uint32_t x, y; };
__device__ __forceinline__ Desc make_desc(uint32_t off, uint32_t len, uint32_t start, uint32_t plen, uint32_t kind){
    Desc d; d.x = (off >> 4) | (len << 17); d.y = start | (plen << 13) | (kind << 26); return d;
}

#ifndef FQK_EMU
__device__ __forceinline__ uint32_t lds32(uint32_t a){ uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds32_volatile(uint32_t a){ uint32_t v; asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }
// A trim warp may run several tiles ahead of the producer.  mbarrier.try_wait.parity cannot tell "phase n has not
// started" from "phase n-1 is complete" when the waiter is two phases ahead of the barrier (it would return at once
// and the warp would read a stale stage), so the producer tags the stage with the tile index before it arms the
// barrier and a consumer waits for the tag first: once the tag is there the barrier is in the right phase.
__device__ __forceinline__ void wait_tile_tag(uint32_t tag_addr, uint32_t tile){
    while(lds32_volatile(tag_addr) != tile) __nanosleep(64);
}
__device__ __forceinline__ uint32_t lds16(uint32_t a){ uint32_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds8(uint32_t a){ uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint2 lds64(uint32_t a){ uint2 v; asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a)); return v; }
__device__ __forceinline__ uint4 lds128(uint32_t a){
    uint4 v; asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v;
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v){ asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void sts64(uint32_t a, uint32_t x, uint32_t y){ asm volatile("st.shared.v2.u32 [%0], {%1, %2};" :: "r"(a), "r"(x), "r"(y) : "memory"); }
__device__ __forceinline__ void sts16(uint32_t a, uint32_t v){ asm volatile("st.shared.u16 [%0], %1;" :: "r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void reds_add(uint32_t a, uint32_t v){ asm volatile("red.shared.add.u32 [%0], %1;" :: "r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t atoms_add(uint32_t a, uint32_t v){ uint32_t o; asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(o) : "r"(a), "r"(v) : "memory"); return o; }
__device__ __forceinline__ uint32_t smem_u32(const void* p){ return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint8_t* smem_ptr(uint32_t a){ return reinterpret_cast<uint8_t*>(__cvta_shared_to_generic((size_t)a)); }
__device__ __forceinline__ uint32_t dp4(uint32_t a, uint32_t b, uint32_t c){ return __dp4a(a, b, c); }

// ------------------------------------------------------------- async staging
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count){
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes){
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar){
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity){
    uint32_t ok = 0;
    const uint32_t a = smem_u32(bar);
    while(!ok){
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    }
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar){
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
// mbarrier wait for the long waits between roles: try_wait with a suspend-time hint parks the thread in hardware
// until the phase completes (or the hint expires), so a waiting warp issues almost nothing; a spinning warp would
// take issue slots from the working warps (12 % of all instructions in the first version)
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity, uint32_t ns){
    uint32_t ok = 0;
    const uint32_t a = smem_u32(bar);
    for(;;){
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(a), "r"(parity), "r"(ns) : "memory");
        if(ok) break;
    }
}
// the same wait issued by ONE lane of a converged warp; the other lanes are parked at the warp barrier
__device__ __forceinline__ void mbar_wait_lane0(uint64_t* bar, uint32_t parity, uint32_t ns){
    if((threadIdx.x & 31) == 0) mbar_wait_sleep(bar, parity, ns);
    __syncwarp();
}
// Named barriers (bar.sync / bar.arrive) for the two handshakes that have no transaction count: a warp blocked in
// bar.sync is parked by the hardware and issues nothing, whereas mbarrier.try_wait returns after a few tens of
// nanoseconds whatever the hint, and the retry loops of the waiting warps took 16 % of all issued instructions
// (profiles/r01d_k_se_v3_lines.txt).
__device__ __forceinline__ void nb_sync(int id, int nthreads){ asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(nthreads) : "memory"); }
__device__ __forceinline__ void nb_arrive(int id, int nthreads){ asm volatile("bar.arrive %0, %1;" :: "r"(id), "r"(nthreads) : "memory"); }
// mbarrier wait with a short back-off between polls (the `full` barriers: completed by the bulk copies)
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t parity){
    uint32_t ok = 0, ns = 32;
    const uint32_t a = smem_u32(bar);
    for(;;){
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
        if(ok) break;
        __nanosleep(ns);
        ns = min(ns * 2u, 256u);
    }
}
#endif  // !FQK_EMU (the emulation header provides lds*/sts*/reds_add/atoms_add/smem_ptr/dp4 on a byte array)

// Filter::passFilter (src/filter.cpp:15-25) counts over the kept part [start, start+len) of a read staged
// in shared memory at address qa (16-byte aligned), 16 quality bytes per load.  Low-quality bases: byte + k
// carries into bit 7 iff byte >= limit; the four carry words of a chunk are merged into one word before the
// popcount.  N bases: bit 7 of the stored byte (chunks without N skip the masking).
__device__ __forceinline__ uint32_t byte_range_mask(int p, int start, int end){      // bytes of word at position p inside [start, end)
    const int nlo = min(4, max(0, start - p)), nhi = min(4, max(0, p + 4 - end));
    return (nlo + nhi >= 4) ? 0u : ((0xFFFFFFFFu << (8 * nlo)) & (0xFFFFFFFFu >> (8 * nhi)));
}
__device__ __forceinline__ void swar_counts(uint32_t qa, int start, int len, int lim, bool want_total,
                                            int& lowq, int& nn, int& total){
    lowq = nn = total = 0;
    if(len <= 0) return;
    const int end = start + len;
    const uint32_t k = (uint32_t)(0x80 - lim) * 0x01010101u;
    int ge_bytes = 0;                                                  // bytes with q >= limit
    for(int c = start >> 4; c <= (end - 1) >> 4; ++c){
        uint4 v = lds128(qa + 16 * c);
        const int p = 16 * c;
        if(p < start || p + 16 > end){                                 // edge chunk: zero the bytes outside the read part
            v.x &= byte_range_mask(p, start, end); v.y &= byte_range_mask(p + 4, start, end);
            v.z &= byte_range_mask(p + 8, start, end); v.w &= byte_range_mask(p + 12, start, end);
        }
        const uint32_t nf = (v.x | v.y | v.z | v.w) & 0x80808080u;
        if(nf){
            nn += __popc(v.x & 0x80808080u) + __popc(v.y & 0x80808080u) + __popc(v.z & 0x80808080u) + __popc(v.w & 0x80808080u);
            v.x &= 0x7F7F7F7Fu; v.y &= 0x7F7F7F7Fu; v.z &= 0x7F7F7F7Fu; v.w &= 0x7F7F7F7Fu;
        }
        const uint32_t g = ((v.x + k) & 0x80808080u) | (((v.y + k) & 0x80808080u) >> 1) |
                           (((v.z + k) & 0x80808080u) >> 2) | (((v.w + k) & 0x80808080u) >> 3);
        ge_bytes += __popc(g);
        if(want_total){
            total = (int)dp4(v.x, 0x01010101u, (uint32_t)total); total = (int)dp4(v.y, 0x01010101u, (uint32_t)total);
            total = (int)dp4(v.z, 0x01010101u, (uint32_t)total); total = (int)dp4(v.w, 0x01010101u, (uint32_t)total);
        }
    }
    lowq = len - ge_bytes;                                             // zeroed bytes never reach the limit (limit >= 33)
    total -= 33 * len;
}

// Filter::trimAndCut (src/filter.cpp:69-189) on a read staged in shared memory (qa = shared address of quality
// byte 0).  Same statements as dev_trim_and_cut (fqk_trim.cuh), explicit shared loads instead of generic ones.
__device__ __forceinline__ int sq7(uint32_t qa, int i){ return (int)(lds8(qa + (uint32_t)i) & 0x7Fu); }
__device__ __forceinline__ bool sisn(uint32_t qa, int i){ return (lds8(qa + (uint32_t)i) & 0x80u) != 0; }
// quality values (flag masked) of positions p .. p+7 of a staged read as two words; p may be unaligned or negative
__device__ __forceinline__ void ldq8(uint32_t qa, int p, uint32_t& lo, uint32_t& hi){
    const uint32_t addr = qa + (uint32_t)p, a = addr & ~3u, sh = (addr & 3u) * 8u;
    const uint32_t w0 = lds32(a), w1 = lds32(a + 4), w2 = lds32(a + 8);
    lo = __funnelshift_r(w0, w1, sh) & 0x7F7F7F7Fu; hi = __funnelshift_r(w1, w2, sh) & 0x7F7F7F7Fu;
}
__device__ __forceinline__ bool s_trim_and_cut(const fq_params& P, uint32_t qa, int l, int F, int T, int& ostart, int& olen){
    const bool ef = P.cut_front_enabled, er = P.cut_right_enabled, et = P.cut_tail_enabled;
    ostart = 0; olen = l;
    if(F == 0 && T == 0 && !ef && !er && !et) return true;
    int rlen = l - F - T;
    if(rlen < 0) return false;
    if(!ef && !er && !et){ ostart = F; olen = rlen; return true; }
    if(ef){
        const int w = P.cut_front_window;
        if(l - F - T - w <= 0) return false;
        const int thr = w * (33 + P.cut_front_qual);
        int s = F, sum = 0;
        if(w == 4){                              // four window positions per step from seven bytes
            const int lim = l - T;
            s = lim - 4;                         // where the scalar loop ends when no window passes
            for(int base = F; base + 4 < lim; base += 4){
                uint32_t lo, hi; ldq8(qa, base, lo, hi);
                uint32_t m = 0;
#pragma unroll
                for(int j = 0; j < 4; ++j){
                    const int sj = (int)dp4(j ? __funnelshift_r(lo, hi, 8 * j) : lo, 0x01010101u, 0u);
                    m |= (sj >= thr && base + j + 4 < lim) ? (1u << j) : 0u;
                }
                if(m){ s = base + __ffs((int)m) - 1; break; }
            }
        }else{
            for(int i = 0; i < w - 1; ++i) sum += sq7(qa, s + i);
            for(; s + w < l - T; ++s){
                sum += sq7(qa, s + w - 1);
                if(s > F) sum -= sq7(qa, s - 1);
                if(sum >= thr) break;
            }
        }
        if(s > 0) s += w - 1;
        while(s < l && sisn(qa, s)) ++s;
        F = s;
        rlen = l - F - T;
    }
    if(er){
        const int w = P.cut_right_window;
        if(l - F - T - w <= 0) return false;
        const int thr = w * (33 + P.cut_right_qual);
        int s = F, sum = 0;
        bool found = false;
        for(int i = 0; i < w - 1; ++i) sum += sq7(qa, s + i);
        for(; s + w < l - T; ++s){
            sum += sq7(qa, s + w - 1);
            if(s > F) sum -= sq7(qa, s - 1);
            if(sum < thr){ found = true; break; }
        }
        if(found){
            const int qr = 33 + P.cut_right_qual;
            while(s < l - 1 && sq7(qa, s) >= qr) ++s;
            rlen = s - F;
        }
    }
    if(!er && et){
        const int w = P.cut_tail_window;
        if(l - F - T - w <= 0) return false;
        const int thr = w * (33 + P.cut_tail_qual);
        int t = l - T - 1, sum = 0;
        if(w == 4){
            int top = t;
            t = F + 3;                           // where the scalar loop ends when no window passes
            for(; top - 4 >= F; top -= 4){
                uint32_t lo, hi; ldq8(qa, top - 6, lo, hi);
                uint32_t m = 0;
#pragma unroll
                for(int j = 0; j < 4; ++j){      // candidate t = top - j: window [t-3, t] starts at byte 3-j of (lo, hi)
                    const int sj = (int)dp4(j < 3 ? __funnelshift_r(lo, hi, 8 * (3 - j)) : lo, 0x01010101u, 0u);
                    m |= (sj >= thr && top - j - 4 >= F) ? (1u << j) : 0u;
                }
                if(m){ t = top - (__ffs((int)m) - 1); break; }
            }
        }else{
            for(int i = 0; i < w - 1; ++i) sum += sq7(qa, t - i);
            for(; t - w >= F; --t){
                sum += sq7(qa, t - w + 1);
                if(t < l - T - 1) sum -= sq7(qa, t + 1);
                if(sum >= thr) break;
            }
        }
        if(t < l - 1) t -= w - 1;
        while(t >= 0 && sisn(qa, t)) --t;
        rlen = t - F + 1;
    }
    if(rlen <= 0 || F >= l - 1) return false;
    ostart = F; olen = rlen;
    return true;
}

// passFilter's counts over a sub-range of a staged read (the parts a trim removed)
__device__ __noinline__ void range_counts(uint32_t qa, int from, int n, int lim, int& lowq, int& nn){
    int total; swar_counts(qa, from, n, lim, false, lowq, nn, total);
}
// number of N bases of a read; nmask flags the 32-byte granules that hold one
__device__ __forceinline__ int n_count(uint32_t qa, int len, uint32_t nmask){
    int nn = 0;
    while(nmask){
        const int gi = __ffs((int)nmask) - 1;
        nmask &= nmask - 1;
        const int nw = ((len + 3) >> 2) - 8 * gi;              // the granule may reach into the next read
        const uint4 v0 = lds128(qa + 32u * gi), v1 = lds128(qa + 32u * gi + 16u);
        const uint32_t wv8[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for(int h = 0; h < 8; ++h) nn += (h < nw) ? __popc(wv8[h] & 0x80808080u) : 0;
    }
    return nn;
}
