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

constexpr int TILE_READS = 504;
size_t g_smem_optin = 0;

size_t smem_bytes(int W){ return (size_t)2 * (20 * 4 * W + 4 * W + 1) * 4 + TILE_READS * sizeof(Desc) + 16; }

} // namespace

void fused_configure(const cudaDeviceProp& prop){
    g_smem_optin = prop.sharedMemPerBlockOptin;
    cudaFuncSetAttribute(k_stats_swar_se<TILE_READS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::min<size_t>(g_smem_optin, 200 * 1024));
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
