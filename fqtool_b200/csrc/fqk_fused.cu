// fqk_fused.cu -- SWAR per-cycle statistics kernel for single-end batches.
//
// Stats::statRead (/root/reference/src/stats.cpp:237-295) as a register-resident
// SIMD-within-a-register reduction.  A thread owns one 4-cycle word of the read
// (4 quality bytes + one byte of 2-bit bases); consecutive threads own consecutive
// words, so a group of W = max_cycles/4 threads streams one read with fully
// coalesced loads, and G = 256/W groups per block stream different reads.  Each
// thread keeps, for four "sets" of bases (all, code bit0, code bit1, both bits),
// packed per-cycle counters in registers: n, n(q>20), n(q>30) as 4x8-bit lanes and
// the quality sum as 2x(2x16-bit) lanes.  Registers are drained into per-block
// shared-memory u32 arrays every <= 255 reads, and the block converts sets into
// the reference's bins (A = all - bit0 - bit1 + both - N, T = bit0 - both,
// C = bit1 - both, G = both) when it retires.  N bases (flag in quality bit 7,
// stored with code 0) take a rare per-byte path.  Per-cycle totals come from a
// per-block read-length histogram, which relies on the zero padding of the batch
// layout.  No atomics on the per-base path.
//
// Pre-filter stats (pre1) use the aligned fast path.  Post-filter stats (post1)
// reuse the same registers when a passing read is untouched by trimming (the
// common case); otherwise the kept part is re-read at its shifted position.
#include "fqk_fused.h"

namespace {

constexpr int KN = 0, KN20 = 1, KN30 = 2, KQS = 3;        // kinds
constexpr int S_ALL = 0, S_B0 = 1, S_B1 = 2, S_B01 = 3, S_N = 4;   // sets
constexpr int THREADS = 256;

struct Acc {           // one accumulator bank: 4 sets x (n, n20, n30, qsE, qsO); S_ALL has no n
    uint32_t n[4], n20[4], n30[4], qe[4], qo[4];
    __device__ __forceinline__ void zero(){
#pragma unroll
        for(int s = 0; s < 4; ++s){ n[s] = n20[s] = n30[s] = qe[s] = qo[s] = 0; }
    }
};

// shared-memory layout of one Stats object: u32 [kind 4][set 5][C] then lenhist[C+1]
__device__ __forceinline__ uint32_t* sm_cell(uint32_t* sm, int C, int kind, int set){ return sm + (kind * 5 + set) * C; }

// rare path: a word holding N bases (quality bit 7); every flagged byte goes to the N set
__device__ __noinline__ void acc_n_bytes(uint32_t* sm, int C, uint32_t qraw, int cyc0){
#pragma unroll
    for(int j = 0; j < 4; ++j){
        const uint32_t b = (qraw >> (8 * j)) & 0xFF;
        if(b & 0x80){
            const uint32_t q = b & 0x7F;
            atomicAdd(sm_cell(sm, C, KN, S_N) + cyc0 + j, 1u);
            if(q > '5') atomicAdd(sm_cell(sm, C, KN20, S_N) + cyc0 + j, 1u);
            if(q > '?') atomicAdd(sm_cell(sm, C, KN30, S_N) + cyc0 + j, 1u);
            atomicAdd(sm_cell(sm, C, KQS, S_N) + cyc0 + j, q);
        }
    }
}

// one word: qraw = 4 quality bytes (N flag in bit 7; bytes outside the read are 0),
// b = the 4 2-bit codes (0 outside the read and for N)
__device__ __forceinline__ void acc_word(Acc& a, uint32_t qraw, uint32_t b){
    const uint32_t q7 = qraw & 0x7F7F7F7Fu;
    const uint32_t t7 = (q7 + 0x4A4A4A4Au) >> 7;          // bit0 of byte j: q >= 54  (q > '5')
    const uint32_t s6 = q7 >> 6;                          // bit0 of byte j: q >= 64  (q > '?')
    // 2-bit codes -> one byte per cycle
    uint32_t x = (b | (b << 12)) & 0x000F000Fu;
    x = (x | (x << 6)) & 0x03030303u;
    uint32_t m[4];
    m[S_ALL] = 0x01010101u;
    m[S_B0] = x & 0x01010101u;
    m[S_B1] = (x >> 1) & 0x01010101u;
    m[S_B01] = m[S_B0] & m[S_B1];
    // set "all": counts come from the length histogram
    a.n20[0] += t7 & 0x01010101u;
    a.n30[0] += s6 & 0x01010101u;
    a.qe[0] += __byte_perm(q7, 0, 0x4240);
    a.qo[0] += __byte_perm(q7, 0, 0x4341);
#pragma unroll
    for(int s = 1; s < 4; ++s){
        a.n[s] += m[s];
        a.n20[s] += t7 & m[s];
        a.n30[s] += s6 & m[s];
        const uint32_t qx = q7 & (m[s] * 0xFFu);
        a.qe[s] += __byte_perm(qx, 0, 0x4240);
        a.qo[s] += __byte_perm(qx, 0, 0x4341);
    }
}

// drain a register bank into one or two shared-memory Stats objects (cycles 4w .. 4w+3)
__device__ __forceinline__ void drain(Acc& a, uint32_t* smA, uint32_t* smB, int C, int w){
    const int c0 = 4 * w;
#pragma unroll
    for(int s = 0; s < 4; ++s){
#pragma unroll
        for(int j = 0; j < 4; ++j){
            const uint32_t vn = (a.n[s] >> (8 * j)) & 0xFF, v20 = (a.n20[s] >> (8 * j)) & 0xFF, v30 = (a.n30[s] >> (8 * j)) & 0xFF;
            const uint32_t vq = (j & 1) ? ((a.qo[s] >> (8 * (j - 1))) & 0xFFFF) : ((a.qe[s] >> (8 * j)) & 0xFFFF);
            if(s && vn){ atomicAdd(sm_cell(smA, C, KN, s) + c0 + j, vn); if(smB) atomicAdd(sm_cell(smB, C, KN, s) + c0 + j, vn); }
            if(v20){ atomicAdd(sm_cell(smA, C, KN20, s) + c0 + j, v20); if(smB) atomicAdd(sm_cell(smB, C, KN20, s) + c0 + j, v20); }
            if(v30){ atomicAdd(sm_cell(smA, C, KN30, s) + c0 + j, v30); if(smB) atomicAdd(sm_cell(smB, C, KN30, s) + c0 + j, v30); }
            if(vq){ atomicAdd(sm_cell(smA, C, KQS, s) + c0 + j, vq); if(smB) atomicAdd(sm_cell(smB, C, KQS, s) + c0 + j, vq); }
        }
    }
    a.zero();
}

// block retirement: sets -> bins, raw quality chars -> Phred (minus 33 per base), add to the global u64 arrays
__device__ void retire(const uint32_t* sm, int C, const fq_stats_layout& L, uint64_t* S){
    const uint32_t* lenhist = sm + 20 * C;
    const uint64_t GC = L.max_cycles;
    // n_all[i] = number of reads longer than i (suffix sum of the length histogram), done by one warp-free loop per thread
    for(int i = threadIdx.x; i < C; i += blockDim.x){
        uint32_t nall = 0;
        for(int l = i + 1; l <= C; ++l) nall += lenhist[l];
        if(!nall) continue;
        uint32_t v[4][5];                     // [kind][set]
#pragma unroll
        for(int k = 0; k < 4; ++k)
#pragma unroll
            for(int s = 0; s < 5; ++s) v[k][s] = sm[(k * 5 + s) * C + i];
        v[KN][S_ALL] = nall;
        uint64_t tb = 0, tq = 0;
#pragma unroll
        for(int cls = 0; cls < 5; ++cls){
            int64_t r[4];
#pragma unroll
            for(int k = 0; k < 4; ++k){
                const int64_t all = v[k][S_ALL], b0 = v[k][S_B0], b1 = v[k][S_B1], b01 = v[k][S_B01], nn = v[k][S_N];
                r[k] = cls == 0 ? all - b0 - b1 + b01 - nn : cls == 1 ? b0 - b01 : cls == 2 ? b1 - b01 : cls == 3 ? b01 : nn;
            }
            if(!r[KN]) continue;
            const int64_t qs = r[KQS] - 33 * r[KN];
            const int bin = cls_to_bin(cls);
            add_u64(S + L.s_content + bin * GC + i, (uint64_t)r[KN]);
            if(r[KN20]) add_u64(S + L.s_q20 + bin * GC + i, (uint64_t)r[KN20]);
            if(r[KN30]) add_u64(S + L.s_q30 + bin * GC + i, (uint64_t)r[KN30]);
            if(qs) add_u64(S + L.s_qualsum + bin * GC + i, (uint64_t)qs);
            tb += (uint64_t)r[KN]; tq += (uint64_t)qs;
        }
        add_u64(S + L.s_total_base + i, tb);
        add_u64(S + L.s_total_qual + i, tq);
    }
    // reads / length_sum from the histogram
    uint64_t nr = 0, ls = 0;
    for(int l = threadIdx.x; l <= C; l += blockDim.x){ nr += lenhist[l]; ls += (uint64_t)lenhist[l] * l; }
#pragma unroll
    for(int o = 16; o; o >>= 1){ nr += __shfl_xor_sync(0xffffffffu, nr, o); ls += __shfl_xor_sync(0xffffffffu, ls, o); }
    if((threadIdx.x & 31) == 0 && nr){ add_u64(S + L.s_reads, nr); add_u64(S + L.s_length_sum, ls); }
}

struct Desc { uint32_t off; uint16_t len; uint16_t start; uint16_t plen; uint16_t kind; };   // kind: 0 common, 1 other (no post), 2 other + post

// the kernel: reads [blk_lo, blk_hi) of the batch per block, tiles of TILE reads
template <int TILE>
__global__ void __launch_bounds__(THREADS, 2)
k_stats_swar_se(const __grid_constant__ DevParams D, DevBatch B, const fq_read_result* __restrict__ res,
                uint64_t* __restrict__ flat, int W, int reads_per_block){
    extern __shared__ uint32_t sm[];
    const int C = 4 * W;                              // cycles covered (>= max read length of the batch)
    uint32_t* smPre = sm;
    uint32_t* smPost = sm + (20 * C + C + 1);
    Desc* desc = reinterpret_cast<Desc*>(smPost + (20 * C + C + 1));
    for(int i = threadIdx.x; i < 2 * (20 * C + C + 1); i += THREADS) sm[i] = 0;
    const int G = THREADS / W;
    const int g = threadIdx.x / W, w = threadIdx.x - g * W;
    const bool active = g < G;
    const uint32_t lo = (uint32_t)min((uint64_t)B.n, (uint64_t)blockIdx.x * reads_per_block);
    const uint32_t hi = (uint32_t)min((uint64_t)B.n, (uint64_t)lo + reads_per_block);
    const uint32_t* __restrict__ q32 = reinterpret_cast<const uint32_t*>(B.qual);
    Acc common, preo, posto;
    common.zero(); preo.zero(); posto.zero();
    int left = 255;                                   // reads this thread may still add before its 8-bit lanes must drain
    __syncthreads();
    for(uint32_t t0 = lo; t0 < hi; t0 += TILE){
        const int nt = (int)min((uint32_t)TILE, hi - t0);
        // ---- tile descriptors (coalesced) ----
        for(int i = threadIdx.x; i < nt; i += THREADS){
            const uint32_t r = t0 + i;
            Desc d; d.off = B.offsets[r]; d.len = B.lengths[r];
            const fq_read_result rec = res[r];
            const bool pass = !(rec.flags & FQK_F_DROPPED) && rec.result == FQK_PASS_FILTER;
            d.start = rec.start; d.plen = rec.len;
            d.kind = pass ? ((rec.start == 0 && rec.len == d.len) ? 0 : 2) : 1;
            if(d.len > C || (d.off & 15)){ atomicOr(D.err, d.len > C ? 1u : 2u); d.len = 0; d.kind = 1; }
            desc[i] = d;
            atomicAdd(smPre + 20 * C + d.len, 1u);
            if(pass) atomicAdd(smPost + 20 * C + d.plen, 1u);
        }
        __syncthreads();
        if(active){
            for(int i = g; i < nt; i += G){
                const Desc d = desc[i];
                if(--left == 0){
                    drain(common, smPre, smPost, C, w); drain(preo, smPre, nullptr, C, w); drain(posto, smPost, nullptr, C, w);
                    left = 255;
                }
                const int nw = (d.len + 3) >> 2;
                if(w < nw){
                    const uint32_t wi = (d.off >> 2) + w;
                    const uint32_t qraw = __ldg(q32 + wi);
                    const uint32_t b = __ldg(B.bases2 + wi);
                    if(qraw & 0x80808080u) acc_n_bytes(smPre, C, qraw, 4 * w);
                    if(d.kind == 0){
                        acc_word(common, qraw, b);
                        if(qraw & 0x80808080u) acc_n_bytes(smPost, C, qraw, 4 * w);
                    }else acc_word(preo, qraw, b);
                }
                if(d.kind == 2){
                    const int nv = (int)d.plen - 4 * w;               // kept bases covered by this thread
                    if(nv > 0){
                        const uint32_t p = d.off + d.start + 4 * w;   // first original position
                        const uint32_t a0 = p >> 2, sh = (p & 3);
                        const uint32_t w0 = __ldg(q32 + a0), w1 = sh ? __ldg(q32 + a0 + 1) : 0u;
                        uint32_t qraw = __funnelshift_r(w0, w1, sh * 8);
                        const uint32_t c0 = __ldg(B.bases2 + a0), c1 = sh ? __ldg(B.bases2 + a0 + 1) : 0u;
                        uint32_t b = ((c0 | (c1 << 8)) >> (2 * sh)) & 0xFF;
                        if(nv < 4){ qraw &= (1u << (8 * nv)) - 1u; b &= (1u << (2 * nv)) - 1u; }
                        if(qraw & 0x80808080u) acc_n_bytes(smPost, C, qraw, 4 * w);
                        acc_word(posto, qraw, b);
                    }
                }
            }
        }
        __syncthreads();
    }
    if(active){ drain(common, smPre, smPost, C, w); drain(preo, smPre, nullptr, C, w); drain(posto, smPost, nullptr, C, w); }
    __syncthreads();
    retire(smPre, C, D.L, flat + D.L.stats_off[0]);
    retire(smPost, C, D.L, flat + D.L.stats_off[2]);
}


// ===========================================================================
// Fused single-end kernel: trim + filter + pre/post statistics in ONE pass over HBM.
//
// Persistent blocks walk their share of the batch in tiles of T consecutive reads.
// The quality and base bytes of a tile are one contiguous range of the batch, so a
// single elected thread stages them into shared memory with two bulk async copies
// (cp.async.bulk global->shared, completion on an mbarrier) while the block is still
// working on the previous tile (two stages).  Phase 1: one thread per read runs the
// reference's per-read sequence (processSingleEnd, src/seprocessor.cpp:313-340) on the
// staged bytes and writes the 16-byte record.  Phase 2: the SWAR statistics of
// acc_word() over the same staged bytes.  HBM is read exactly once per base.
// ===========================================================================
__device__ __forceinline__ uint32_t smem_u32(const void* p){ return (uint32_t)__cvta_generic_to_shared(p); }
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

// Filter::passFilter (src/filter.cpp:3-52) counts over the kept part [start, start+len) of a staged
// read, four quality bytes per step: low-quality bases via a per-byte carry test, N bases via bit 7.
__device__ __forceinline__ void swar_counts(const uint32_t* q32, int start, int len, int lim,
                                            int& lowq, int& nn, int& total){
    lowq = nn = total = 0;
    if(len <= 0) return;
    const int first = start >> 2, last = (start + len - 1) >> 2;
    const uint32_t k = (uint32_t)(0x80 - lim) * 0x01010101u;          // byte + k carries into bit 7 iff byte >= lim
    for(int wi = first; wi <= last; ++wi){
        uint32_t qraw = q32[wi];
        uint32_t mask = 0xFFFFFFFFu;
        if(wi == first) mask &= 0xFFFFFFFFu << (8 * (start & 3));
        if(wi == last) mask &= 0xFFFFFFFFu >> (8 * (3 - ((start + len - 1) & 3)));
        const uint32_t q7 = qraw & 0x7F7F7F7Fu;
        const uint32_t ge = (q7 + k) & 0x80808080u;
        lowq += __popc(~ge & 0x80808080u & mask);
        nn += __popc(qraw & 0x80808080u & mask);
        total = __dp4a(q7 & mask, 0x01010101u, (uint32_t)total);
    }
    total -= 33 * len;
}

struct Stage {              // shared-memory staging area of one tile
    uint32_t* off; uint16_t* len; uint8_t* q; uint8_t* b;
};

__global__ void __launch_bounds__(THREADS, 2)
k_fused_se(const __grid_constant__ DevParams D, DevBatch B, fq_read_result* __restrict__ out,
           uint64_t* __restrict__ flat, int W, int T, int reads_per_block){
    extern __shared__ __align__(128) uint8_t smraw[];
    const fq_params& P = D.P;
    const int C = 4 * W, Cp = (C + 15) & ~15;
    const int statw = 20 * C + C + 1;
    uint32_t* smPre = reinterpret_cast<uint32_t*>(smraw);
    uint32_t* smPost = smPre + statw;
    BlockCounters* bc = reinterpret_cast<BlockCounters*>(smPost + statw);
    uint64_t* mbar = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(bc + 1) + 7) & ~uintptr_t(7));
    Desc* desc = reinterpret_cast<Desc*>(mbar + 2);
    uint8_t* sp = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(desc + T) + 15) & ~uintptr_t(15));
    const uint32_t qcap = (uint32_t)T * Cp, bcap = qcap / 4 + 16;
    Stage st[2];
#pragma unroll
    for(int s = 0; s < 2; ++s){
        st[s].q = sp; sp += qcap;
        st[s].b = sp; sp += bcap;
        st[s].off = reinterpret_cast<uint32_t*>(sp); sp += ((T * 4 + 15) & ~15);
        st[s].len = reinterpret_cast<uint16_t*>(sp); sp += ((T * 2 + 15) & ~15);
    }
    for(int i = threadIdx.x; i < 2 * statw; i += THREADS) smPre[i] = 0;
    counters_zero(bc);
    if(threadIdx.x == 0){ mbar_init(&mbar[0], 1); mbar_init(&mbar[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

    const int G = THREADS / W;
    const int g = threadIdx.x / W, w = threadIdx.x - g * W;
    const bool active = g < G;
    const uint32_t lo = (uint32_t)min((uint64_t)B.n, (uint64_t)blockIdx.x * reads_per_block);
    const uint32_t hi = (uint32_t)min((uint64_t)B.n, (uint64_t)lo + reads_per_block);
    const int ntiles = (int)((hi - lo + T - 1) / T);

    // issue the two bulk copies of a tile whose offsets/lengths are already in st.off/st.len
    auto issue = [&](const Stage& s, int nt, uint64_t* bar){
        const uint32_t a = s.off[0];
        const uint32_t lastoff = s.off[nt - 1], lastlen = s.len[nt - 1];
        uint64_t e = (uint64_t)lastoff + ((lastlen + 15u) & ~15u);
        bool bad = (a & 15u) || e < a || e - a > qcap || e > B.total;
        if(bad){ atomicOr(D.err, 2u); e = a; }
        const uint32_t qbytes = (uint32_t)(e - a);
        const uint32_t b0 = (a >> 2) & ~15u;                               // base bytes start, aligned down to 16
        const uint32_t bbytes = qbytes ? ((uint32_t)((e >> 2) + 15u) & ~15u) - b0 : 0u;
        mbar_expect_tx(bar, qbytes + bbytes);
        if(qbytes){
            bulk_g2s(s.q, B.qual + a, qbytes, bar);
            bulk_g2s(s.b, B.bases2 + b0, bbytes, bar);
        }
    };
    auto load_meta = [&](const Stage& s, uint32_t t0, int nt){
        for(int i = threadIdx.x; i < nt; i += THREADS){ s.off[i] = B.offsets[t0 + i]; s.len[i] = B.lengths[t0 + i]; }
    };

    if(ntiles > 0) load_meta(st[0], lo, (int)min((uint32_t)T, hi - lo));
    __syncthreads();
    if(threadIdx.x == 0 && ntiles > 0) issue(st[0], (int)min((uint32_t)T, hi - lo), &mbar[0]);

    Acc common, preo, posto;
    common.zero(); preo.zero(); posto.zero();
    int left = 255;
    Tally tl = {};
    const bool fast_filter = !P.complexity_enabled;

    for(int k = 0; k < ntiles; ++k){
        const Stage& S = st[k & 1];
        const Stage& Snext = st[(k + 1) & 1];
        const uint32_t t0 = lo + (uint32_t)k * T;
        const int nt = (int)min((uint32_t)T, hi - t0);
        const uint32_t t1 = t0 + T;
        const int nt1 = (k + 1 < ntiles) ? (int)min((uint32_t)T, hi - t1) : 0;
        // metadata of the next tile: loads in flight during phase 1
        uint32_t moff = 0, mlen = 0;
        if((int)threadIdx.x < nt1){ moff = B.offsets[t1 + threadIdx.x]; mlen = B.lengths[t1 + threadIdx.x]; }
        mbar_wait(&mbar[k & 1], (uint32_t)(k >> 1) & 1u);
        const uint32_t a0 = S.off[0];
        const uint32_t bshift = (a0 >> 2) & 15u;
        // ---------------- phase 1: one thread per read ----------------
        for(int i = threadIdx.x; i < nt; i += THREADS){
            Desc d; d.off = S.off[i] - a0; d.len = S.len[i];
            const uint32_t padded = (d.len + 15u) & ~15u;
            bool ok = !(S.off[i] & 15u) && S.off[i] >= a0 && d.off + padded <= qcap && d.len <= C;
            if(ok && i + 1 < nt) ok = (S.off[i + 1] == S.off[i] + padded);            // dense, ordered batch
            if(!ok){ atomicOr(D.err, d.len > C ? 1u : 2u); d.off = 0; d.len = 0; }
            ReadView v; v.q = S.q + d.off; v.b2 = S.b + bshift + (d.off >> 2); v.start = 0; v.len = d.len;
            fq_read_result rec = {};
            if(fast_filter){
                const bool alive = dev_trim_and_cut(P, v, P.trim_front1, P.trim_tail1);
                if(alive){
                    if(P.polyg_enabled) dev_trim_polyg(v, P.polyg_min_len, P.polyg_max_mismatch, P.polyg_one_mismatch_each, rec, tl);
                    if(P.adapter_trim_enabled && P.adapter_r1_len > 0) dev_trim_by_sequence(v, D.ad1, P.adapter_r1_len, rec, tl);
                    if(P.polyx_enabled) dev_trim_polyx(v, P.polyx_base_mask, P.polyx_min_len, P.polyx_max_mismatch, P.polyx_one_mismatch_each, rec, tl);
                    if(P.trim_max_len1 > 0 && P.trim_max_len1 < v.len) view_resize(v, P.trim_max_len1);
                }
                int result = FQK_PASS_FILTER;
                if(!alive || v.len == 0) result = FQK_FAIL_LENGTH;
                else{
                    if(P.qual_filter_enabled){
                        int lowq, nn, total;
                        swar_counts(reinterpret_cast<const uint32_t*>(S.q + d.off), v.start, v.len, P.low_quality_limit, lowq, nn, total);
                        if(lowq > P.low_quality_base_limit) result = FQK_FAIL_QUALITY;
                        else if(P.average_quality_limit > 0 && P.average_quality_limit > (double)total / v.len) result = FQK_FAIL_QUALITY;
                        else if(nn > P.n_base_limit) result = FQK_FAIL_N_BASE;
                    }
                    if(result == FQK_PASS_FILTER && P.length_filter_enabled){
                        if(v.len < P.min_read_length) result = FQK_FAIL_LENGTH;
                        else if(P.max_read_length > 0 && v.len > P.max_read_length) result = FQK_FAIL_TOO_LONG;
                    }
                }
                rec.result = (uint8_t)result;
                if(!alive){ rec.flags |= FQK_F_DROPPED; rec.start = 0; rec.len = 0; }
                else { rec.start = (uint16_t)v.start; rec.len = (uint16_t)v.len; }
            }else se_read(D, v, rec, tl);
            out[t0 + i] = rec;
            atomicAdd(&bc->filter[rec.result], 1u);
            const bool pass = !(rec.flags & FQK_F_DROPPED) && rec.result == FQK_PASS_FILTER;
            d.start = rec.start; d.plen = rec.len;
            d.kind = pass ? ((rec.start == 0 && rec.len == d.len) ? 0 : 2) : 1;
            desc[i] = d;
            atomicAdd(smPre + 20 * C + d.len, 1u);
            if(pass) atomicAdd(smPost + 20 * C + d.plen, 1u);
        }
        if((int)threadIdx.x < nt1){ Snext.off[threadIdx.x] = moff; Snext.len[threadIdx.x] = (uint16_t)mlen; }
        __syncthreads();
        if(threadIdx.x == 0 && nt1 > 0) issue(Snext, nt1, &mbar[(k + 1) & 1]);
        // ---------------- phase 2: SWAR statistics ----------------
        if(active){
            const uint8_t* sb = S.b + bshift;
            for(int i = g; i < nt; i += G){
                const Desc d = desc[i];
                if(--left == 0){
                    drain(common, smPre, smPost, C, w); drain(preo, smPre, nullptr, C, w); drain(posto, smPost, nullptr, C, w);
                    left = 255;
                }
                const int nw = (d.len + 3) >> 2;
                if(w < nw){
                    const uint32_t qraw = *reinterpret_cast<const uint32_t*>(S.q + d.off + 4 * w);
                    const uint32_t b = sb[(d.off >> 2) + w];
                    if(qraw & 0x80808080u) acc_n_bytes(smPre, C, qraw, 4 * w);
                    if(d.kind == 0){
                        acc_word(common, qraw, b);
                        if(qraw & 0x80808080u) acc_n_bytes(smPost, C, qraw, 4 * w);
                    }else acc_word(preo, qraw, b);
                }
                if(d.kind == 2){
                    const int nv = (int)d.plen - 4 * w;
                    if(nv > 0){
                        const uint32_t p = d.off + d.start + 4 * w;
                        const uint32_t a1 = p & ~3u, sh = p & 3u;
                        const uint32_t w0 = *reinterpret_cast<const uint32_t*>(S.q + a1);
                        const uint32_t w1 = sh ? *reinterpret_cast<const uint32_t*>(S.q + a1 + 4) : 0u;
                        uint32_t qraw = __funnelshift_r(w0, w1, sh * 8);
                        const uint32_t c0 = sb[p >> 2], c1 = sh ? sb[(p >> 2) + 1] : 0u;
                        uint32_t b = ((c0 | (c1 << 8)) >> (2 * sh)) & 0xFF;
                        if(nv < 4){ qraw &= (1u << (8 * nv)) - 1u; b &= (1u << (2 * nv)) - 1u; }
                        if(qraw & 0x80808080u) acc_n_bytes(smPost, C, qraw, 4 * w);
                        acc_word(posto, qraw, b);
                    }
                }
            }
        }
        __syncthreads();
    }
    if(active){ drain(common, smPre, smPost, C, w); drain(preo, smPre, nullptr, C, w); drain(posto, smPost, nullptr, C, w); }
    counters_add_tally(bc, tl);
    __syncthreads();
    retire(smPre, C, D.L, flat + D.L.stats_off[0]);
    retire(smPost, C, D.L, flat + D.L.stats_off[2]);
    counters_flush(bc, D.L, flat);
}

constexpr int TILE_READS = 504;
size_t g_smem_optin = 0;
constexpr size_t SMEM_BUDGET = 100 * 1024;        // two blocks per SM

size_t smem_bytes(int W){ return (size_t)2 * (20 * 4 * W + 4 * W + 1) * 4 + TILE_READS * sizeof(Desc) + 16; }

size_t fused_smem_bytes(int W, int T){
    const int C = 4 * W, Cp = (C + 15) & ~15;
    size_t s = (size_t)2 * (20 * C + C + 1) * 4 + sizeof(BlockCounters) + 8 + 16 + (size_t)T * sizeof(Desc) + 16;
    s += 2 * ((size_t)T * Cp + (size_t)T * Cp / 4 + 16 + ((T * 4 + 15) & ~15) + ((T * 2 + 15) & ~15));
    return s + 128;
}
// largest tile (multiple of the group count, at most THREADS reads) that keeps two blocks per SM
int fused_tile(int W){
    const int G = THREADS / W;
    for(int T = THREADS / G * G; T >= G; T -= G)
        if(fused_smem_bytes(W, T) <= SMEM_BUDGET) return T;
    return 0;
}

} // namespace

void fused_configure(const cudaDeviceProp& prop){
    g_smem_optin = prop.sharedMemPerBlockOptin;
    cudaFuncSetAttribute(k_stats_swar_se<TILE_READS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::min<size_t>(g_smem_optin, 200 * 1024));
    cudaFuncSetAttribute(k_fused_se, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::min<size_t>(g_smem_optin, SMEM_BUDGET + 1024));
}

bool swar_stats_supported_se(const DevParams& D){
    if(D.P.kmer_len != 0) return false;                    // k-mer histogram: generic kernel
    const int W = (int)(D.L.max_cycles + 3) / 4;
    if(W < 8 || W > THREADS) return false;
    return smem_bytes(W) + 2048 <= std::min<size_t>(g_smem_optin, 200 * 1024);
}

int swar_stats_launch_se(const DevParams& D, const DevBatch& B, const fq_read_result* res, uint64_t* flat, int sms,
                         cudaStream_t st){
    const int W = (int)(D.L.max_cycles + 3) / 4;
    const int blocks_target = sms * 2;
    int rpb = (int)((B.n + blocks_target - 1) / blocks_target);
    rpb = (rpb + TILE_READS - 1) / TILE_READS * TILE_READS;
    const int grid = (int)((B.n + rpb - 1) / rpb);
    k_stats_swar_se<TILE_READS><<<grid, THREADS, smem_bytes(W), st>>>(D, B, res, flat, W, rpb);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}

bool fused_supported_se(const DevParams& D, const DevBatch& B){
    if(!swar_stats_supported_se(D)) return false;
    if((reinterpret_cast<uintptr_t>(B.qual) | reinterpret_cast<uintptr_t>(B.bases2)) & 15) return false;
    if(B.total % FQK_BATCH_PAD) return false;
    const int W = (int)(D.L.max_cycles + 3) / 4;
    return fused_tile(W) >= THREADS / W && fused_smem_bytes(W, fused_tile(W)) <= g_smem_optin;
}

int fused_launch_se(const DevParams& D, const DevBatch& B, fq_read_result* out, uint64_t* flat, int sms, cudaStream_t st){
    const int W = (int)(D.L.max_cycles + 3) / 4;
    const int T = fused_tile(W);
    const int blocks_target = sms * 2;
    int rpb = (int)((B.n + blocks_target - 1) / blocks_target);
    rpb = (rpb + T - 1) / T * T;
    const int grid = (int)((B.n + rpb - 1) / rpb);
    k_fused_se<<<grid, THREADS, fused_smem_bytes(W, T), st>>>(D, B, out, flat, W, T, rpb);
    return cudaGetLastError() == cudaSuccess ? 1 : -1;
}
