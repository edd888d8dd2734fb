// fqk_fused.cu -- the fused kernels: trim + filter + pre/post per-cycle statistics in ONE pass over HBM (sm_100a).
//   k_se_v3       single-end batches without k-mers (fqk_se3.cuh): three warp roles, P - D + Z statistics
//   k_pe_v4       paired-end batches (fqk_pe4.cuh): both mates staged per tile, warp-cooperative overlap analysis
//   k_fused_se_ws single-end with k-mers, and the statistics-only passes of the older paired-end path
//
// Data movement.  Persistent blocks (one per SM) walk their share of the batch in tiles of T consecutive reads.
// The quality and base bytes of a tile are one contiguous range of the dense batch, so they are staged into
// shared memory with bulk async copies (cp.async.bulk global->shared::cluster, completion counted on an
// mbarrier) while the block is still computing on earlier tiles.  HBM is read once per base; records are
// written once.
//
// Statistics (Stats::statRead, /root/reference/src/stats.cpp:237-295) are register-resident reductions.  A
// thread owns one 4-cycle word of the read (4 quality bytes + one byte of 2-bit bases); consecutive threads own
// consecutive words, so a group of W threads covers a read and several groups cover different reads.  For five
// "sets" of bases (all, code bit0, code bit1, both bits, N) the per-cycle counters n, n(q>20), n(q>30) and the
// quality sums are accumulated either four reads at a time with dp4a (QuadAcc) or one read at a time as byte
// lanes (Acc1).  When the block retires, sets are converted to the reference's bins: A = all - bit0 - bit1 +
// both - N, T = bit0 - both, C = bit1 - both, G = both.  Per-cycle base totals come from a per-block read-length
// histogram, which relies on the zero padding of the batch layout.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include "fqk_fused.h"
#include "fqk_pe.cuh"
#include "fqk_smem.cuh"

// Optional cycle accounting per warp role (nvcc -DFQK_V3_PROF; read back with fqk_debug_prof): development aid only.
#ifdef FQK_V3_PROF
__device__ unsigned long long g_v3prof[16];
// per block accumulators in dynamic shared memory, flushed once at the end (global atomics per probe distort the timing)
__device__ __forceinline__ unsigned long long* v3prof_base(){ extern __shared__ __align__(128) uint8_t smraw[]; return reinterpret_cast<unsigned long long*>(smraw); }
#define s_v3prof v3prof_base()
#define PROF_T(var) const long long var = clock64()
#define PROF_ADD(slot, t0) do{ if((threadIdx.x & 31) == 0) atomicAdd(&s_v3prof[slot], (unsigned long long)(clock64() - (t0))); }while(0)
#else
#define PROF_T(var)
#define PROF_ADD(slot, t0)
#endif
extern "C" int fqk_debug_prof(unsigned long long* out16, int reset){
#ifdef FQK_V3_PROF
    cudaDeviceSynchronize();
    if(out16) cudaMemcpyFromSymbol(out16, g_v3prof, sizeof(unsigned long long) * 16);
    if(reset){ unsigned long long z[16] = {}; cudaMemcpyToSymbol(g_v3prof, z, sizeof(z)); }
    return 1;
#else
    (void)out16; (void)reset; return 0;
#endif
}

namespace {



// level 1: 8-bit count lanes, 16-bit sum lanes (n[S_ALL] is unused: totals come from lengths)
struct Acc1 {
    uint32_t n[NSETS], n20[NSETS], n30[NSETS], qe[NSETS], qo[NSETS];
    __device__ __forceinline__ void zero(){
#pragma unroll
        for(int s = 0; s < NSETS; ++s){ n[s] = n20[s] = n30[s] = qe[s] = qo[s] = 0; }
    }
};
// shared-memory layout of one Stats object: u32 [kind 4][set 5][C] then lenhist[C+1]
__device__ __forceinline__ uint32_t* sm_cell(uint32_t* sm, int C, int kind, int set){ return sm + (kind * NSETS + set) * C; }
__device__ __forceinline__ void sm_add(uint32_t* smA, uint32_t* smB, int C, int kind, int set, int cyc, uint32_t v){
    if(v){
        atomicAdd(sm_cell(smA, C, kind, set) + cyc, v);
        if(smB) atomicAdd(sm_cell(smB, C, kind, set) + cyc, v);
    }
}

// one word: qraw = 4 quality bytes (N flag in bit 7; bytes outside the read are 0),
// b = the four 2-bit codes (0 outside the read and for N)
__device__ __forceinline__ void acc_word(Acc1& a, uint32_t qraw, uint32_t b){
    const uint32_t q7 = qraw & 0x7F7F7F7Fu;
    const uint32_t t7 = (q7 + 0x4A4A4A4Au) >> 7;          // bit0 of byte j: q >= 54  (q > '5')
    const uint32_t s6 = q7 >> 6;                          // bit0 of byte j: q >= 64  (q > '?')
    uint32_t x = (b | (b << 12)) & 0x000F000Fu;           // 2-bit codes -> one byte per cycle
    x = (x | (x << 6)) & 0x03030303u;
    uint32_t m[4];
    m[S_B0] = x & 0x01010101u;
    m[S_B1] = (x >> 1) & 0x01010101u;
    m[S_B01] = m[S_B0] & m[S_B1];
    a.n20[S_ALL] += t7 & 0x01010101u;
    a.n30[S_ALL] += s6 & 0x01010101u;
    a.qe[S_ALL] += __byte_perm(q7, 0, 0x4240);
    a.qo[S_ALL] += __byte_perm(q7, 0, 0x4341);
#pragma unroll
    for(int s = 1; s < 4; ++s){
        a.n[s] += m[s];
        a.n20[s] += t7 & m[s];
        a.n30[s] += s6 & m[s];
        const uint32_t qx = q7 & (m[s] * 0xFFu);
        a.qe[s] += __byte_perm(qx, 0, 0x4240);
        a.qo[s] += __byte_perm(qx, 0, 0x4341);
    }
    if(qraw & 0x80808080u){                               // rare: N bases in this word
        const uint32_t mn = (qraw >> 7) & 0x01010101u;
        a.n[S_N] += mn;
        a.n20[S_N] += t7 & mn;
        a.n30[S_N] += s6 & mn;
        const uint32_t qx = q7 & (mn * 0xFFu);
        a.qe[S_N] += __byte_perm(qx, 0, 0x4240);
        a.qo[S_N] += __byte_perm(qx, 0, 0x4341);
    }
}

// level 1 -> shared arrays of one or two Stats objects (cycles 4w .. 4w+3); rare banks only
__device__ __noinline__ void drain1(uint32_t* vals /*[NSETS*5]*/, uint32_t* smA, uint32_t* smB, int C, int w){
    const int c0 = 4 * w;
    for(int s = 0; s < NSETS; ++s){
        const uint32_t n = vals[s * 5 + 0], n20 = vals[s * 5 + 1], n30 = vals[s * 5 + 2], qe = vals[s * 5 + 3], qo = vals[s * 5 + 4];
        for(int j = 0; j < 4; ++j){
            if(s) sm_add(smA, smB, C, KN, s, c0 + j, (n >> (8 * j)) & 0xFF);
            sm_add(smA, smB, C, KN20, s, c0 + j, (n20 >> (8 * j)) & 0xFF);
            sm_add(smA, smB, C, KN30, s, c0 + j, (n30 >> (8 * j)) & 0xFF);
            const uint32_t qv = (j & 1) ? ((qo >> (8 * (j - 1))) & 0xFFFF) : ((qe >> (8 * j)) & 0xFFFF);
            sm_add(smA, smB, C, KQS, s, c0 + j, qv);
        }
    }
}
__device__ __forceinline__ void drain_l1(Acc1& a, uint32_t* smA, uint32_t* smB, int C, int w){
    uint32_t vals[NSETS * 5];
#pragma unroll
    for(int s = 0; s < NSETS; ++s){ vals[s * 5] = a.n[s]; vals[s * 5 + 1] = a.n20[s]; vals[s * 5 + 2] = a.n30[s]; vals[s * 5 + 3] = a.qe[s]; vals[s * 5 + 4] = a.qo[s]; }
    drain1(vals, smA, smB, C, w);
    a.zero();
}
// block retirement: sets -> bins, raw quality chars -> Phred (minus 33 per base), add to the global u64 arrays.
// Everything is linear and evaluated in wrapping signed arithmetic, so a block may hold a partial view
// (e.g. only the tail of merged reads: length histogram entries +1 at the end, -1 at the start of the range).
__device__ void retire(const uint32_t* sm, int C, const fq_stats_layout& L, uint64_t* S){
    const uint32_t* lenhist = sm + 4 * NSETS * C;
    const uint64_t GC = L.max_cycles;
    for(int i = threadIdx.x; i < C; i += blockDim.x){
        int32_t nall = 0;                                 // reads covering cycle i
        for(int l = i + 1; l <= C; ++l) nall += (int32_t)lenhist[l];
        uint32_t v[4][NSETS]; uint32_t any = (uint32_t)nall;
#pragma unroll
        for(int k = 0; k < 4; ++k)
#pragma unroll
            for(int s = 0; s < NSETS; ++s){ v[k][s] = sm[(k * NSETS + s) * C + i]; any |= v[k][s]; }
        if(!any) continue;
        int64_t tb = 0, tq = 0;
#pragma unroll
        for(int cls = 0; cls < 5; ++cls){
            int64_t r[4];
#pragma unroll
            for(int k = 0; k < 4; ++k){
                const int64_t all = k == KN ? (int64_t)nall : (int64_t)v[k][S_ALL];
                const int64_t b0 = v[k][S_B0], b1 = v[k][S_B1], b01 = v[k][S_B01], nn = v[k][S_N];
                r[k] = cls == 0 ? all - b0 - b1 + b01 - nn : cls == 1 ? b0 - b01 : cls == 2 ? b1 - b01 : cls == 3 ? b01 : nn;
            }
            const int64_t qs = r[KQS] - 33 * r[KN];
            const int bin = cls_to_bin(cls);
            if(r[KN]) add_u64(S + L.s_content + bin * GC + i, (uint64_t)r[KN]);
            if(r[KN20]) add_u64(S + L.s_q20 + bin * GC + i, (uint64_t)r[KN20]);
            if(r[KN30]) add_u64(S + L.s_q30 + bin * GC + i, (uint64_t)r[KN30]);
            if(qs) add_u64(S + L.s_qualsum + bin * GC + i, (uint64_t)qs);
            tb += r[KN]; tq += qs;
        }
        if(tb) add_u64(S + L.s_total_base + i, (uint64_t)tb);
        if(tq) add_u64(S + L.s_total_qual + i, (uint64_t)tq);
    }
    int64_t nr = 0, ls = 0;                               // reads / length_sum from the histogram
    for(int l = threadIdx.x; l <= C; l += blockDim.x){ const int64_t h = (int32_t)lenhist[l]; nr += h; ls += h * l; }
#pragma unroll
    for(int o = 16; o; o >>= 1){ nr += __shfl_xor_sync(0xffffffffu, nr, o); ls += __shfl_xor_sync(0xffffffffu, ls, o); }
    if((threadIdx.x & 31) == 0){ if(nr) add_u64(S + L.s_reads, (uint64_t)nr); if(ls) add_u64(S + L.s_length_sum, (uint64_t)ls); }
}


// ===========================================================================
// Warp-specialised variant (the default): 8 statistics warps + 8 trim warps per block.
//   trim warps (threads 256..511): bulk-copy producer, phase 1 of tile k+1 (records, per-tile lists of
//       "common" and "other" reads) while the statistics warps reduce tile k;
//   statistics warps (threads 0..255): phase 2 over the lists (common reads through the dp4a quad path),
//       no block-wide barrier in steady state.
// Three staging buffers; mbarriers: full[s] (bulk copies landed), ready[s] (lists written), free[s]
// (statistics done with the stage).  Register budget split with setmaxnreg (168 / 80).
// ===========================================================================
// Two splits: NSTAT = 128 statistics threads + 512 trim threads (single-end: phase 1 is the long pole, and it is
// latency-bound, so it gets the warps) or 256 + 256 (statistics-only passes of paired-end batches).
constexpr int NS = 3, MS = 4;
constexpr int ws_trim_threads(int nstat){ return nstat == 128 ? 512 : 256; }
constexpr int KMER_SMEM_MAX = 6;


// phase 1 of one read staged in shared memory: the per-read sequence of processSingleEnd
// (src/seprocessor.cpp:313-340); returns the record
__device__ __forceinline__ fq_read_result phase1_read(const DevParams& D, const uint8_t* q, const uint8_t* b2, int len,
                                                      bool fast_filter, Tally& tl){
    const fq_params& P = D.P;
    ReadView v; v.q = q; v.b2 = b2; v.start = 0; v.len = len;
    fq_read_result rec = {};
    if(!fast_filter){ se_read(D, v, rec, tl); return rec; }
    const bool alive = dev_trim_and_cut(P, v, P.trim_front1, P.trim_tail1);
    if(alive){
        if(P.polyg_enabled) dev_trim_polyg(v, P.polyg_min_len, P.polyg_max_mismatch, P.polyg_one_mismatch_each, rec, tl);
        if(P.adapter_trim_enabled && P.adapter_r1_len > 0) trim_by_sequence_any(D, v, 0, rec, tl);
        if(P.polyx_enabled) dev_trim_polyx(v, P.polyx_base_mask, P.polyx_min_len, P.polyx_max_mismatch, P.polyx_one_mismatch_each, rec, tl);
        if(P.trim_max_len1 > 0 && P.trim_max_len1 < v.len) view_resize(v, P.trim_max_len1);
    }
    int result = FQK_PASS_FILTER;
    if(!alive || v.len == 0) result = FQK_FAIL_LENGTH;
    else{
        if(P.qual_filter_enabled){
            int lowq, nn, total;
            swar_counts(smem_u32(q), v.start, v.len, P.low_quality_limit,
                        P.average_quality_limit > 0, lowq, nn, total);
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
    return rec;
}

// k-mers ending at view indices 4w .. 4w+3 of the view [vstart, vstart+vlen) of a staged read (tile offset off).
// The 16 bases ending at the thread's last position are one unaligned 32-bit window of the packed bases;
// reversing its 2-bit groups puts earlier bases at higher bits = Evaluator::seq2int's key order
// (src/evaluator.cpp:3-49).  A k-mer counts iff all k bases are A/T/C/G: reads without N skip the check.
__device__ __forceinline__ void kmers_word(uint32_t sq0, uint32_t sb0, uint32_t off, uint32_t vstart, uint32_t vlen, uint32_t hasN,
                                           int w, int K, uint64_t* histA, uint64_t* histB, uint32_t smhist = 0){
    if((uint32_t)(4 * w) >= vlen) return;
    const int e = (int)(off + vstart) + 4 * w + 3;                   // tile position of view index 4w+3
    const int s0 = e - 15, bi = s0 >> 2;
    const int a = bi & ~3, sh = 8 * (bi & 3) + 2 * (s0 & 3);
    const uint32_t win = __funnelshift_r(lds32(sb0 + a), lds32(sb0 + a + 4), sh);
    uint32_t rw = __brev(win);
    rw = ((rw >> 1) & 0x55555555u) | ((rw & 0x55555555u) << 1);
    const uint32_t mask = (K >= 16) ? 0xFFFFFFFFu : ((1u << (2 * K)) - 1u);
#pragma unroll
    for(int j = 0; j < 4; ++j){
        const int i = 4 * w + j;
        if(i < K - 1 || (uint32_t)i >= vlen) continue;
        bool ok = true;
        if(hasN){
            const uint32_t qa = sq0 + off + vstart + (uint32_t)i;
            for(int t = 0; t < K; ++t) ok &= !(lds8(qa - t) & 0x80u);
        }
        if(ok){
            const uint32_t key = (rw >> (2 * (3 - j))) & mask;
            if(smhist) asm volatile("red.shared.add.u32 [%0], 1;" :: "r"(smhist + 4 * key) : "memory");
            else { add_u64(histA + key, 1); if(histB) add_u64(histB + key, 1); }
        }
    }
}

// What the kernel does with a batch.  mode 0: single-end, records are computed here (phase 1).
// mode 1: statistics of one mate of a paired-end batch from the records k_trim_pe wrote (pre = raw read,
// post = the mate's own contribution following src/peprocessor.cpp:353-429).  mode 2: read2's batch once
// more, only the reverse-complemented tails that OverlapAnalysis::merge appends to merged reads
// (src/overlapanalysis.cpp:80-93), which belong to post1 at cycles [len1, len1+len2).
struct StatsMode {
    int mode, mate, merge_enabled, pre_obj, post_obj;
    const fq_read_result* recA;      // records of this mate
    const fq_read_result* recB;      // records of the other mate
    const fq_pair_result* recP;
    const uint32_t* maxlen;          // device scalar: longest raw read of this mate's batch (or nullptr)
};

// ---------------------------------------------------------------------------
// Quad path of the statistics warps: four "common" reads at a time.  The thread's 4-cycle word of each of the
// four reads is byte-transposed (8 PRMT), so that a register holds ONE cycle of FOUR reads; every per-cycle sum
// is then a single dp4a into a 32-bit accumulator (no byte lanes, no folding):  sum over reads of a*b with
// a = quality / threshold flag, b = class mask.  Flags and masks stay where they are in the byte (weights
// 128 / 64 for q>20 / q>30, 4^j or 2*4^j for the code bits of cycle j); the weights are divided out when the
// accumulators are drained.  Largest step: 4 bases * 128 * 128 = 2^16 per quad, so 2^15 quads fit in 32 bits.
constexpr int QUAD_LIMIT = 30000;
struct QuadAcc {
    uint32_t a20[4], a30[4], aqs[4];                        // all bases of cycle j
    uint32_t n[3][4], n20[3][4], n30[3][4], qs[3][4];       // code bit0, bit1, both
    uint32_t nn[4], nn20[4], nn30[4], nqs[4];               // N bases
    __device__ __forceinline__ void zero(){
#pragma unroll
        for(int j = 0; j < 4; ++j){
            a20[j] = a30[j] = aqs[j] = nn[j] = nn20[j] = nn30[j] = nqs[j] = 0;
#pragma unroll
            for(int s = 0; s < 3; ++s) n[s][j] = n20[s][j] = n30[s][j] = qs[s][j] = 0;
        }
    }
};

// q0..q3: the thread's quality word of four reads (0 outside a read), b0..b3: their code bytes
__device__ __forceinline__ void quad_accumulate(QuadAcc& A, uint32_t q0, uint32_t q1, uint32_t q2, uint32_t q3,
                                                uint32_t b0, uint32_t b1, uint32_t b2, uint32_t b3){
    const uint32_t t0 = __byte_perm(q0, q1, 0x5140), t1 = __byte_perm(q0, q1, 0x7362);
    const uint32_t t2 = __byte_perm(q2, q3, 0x5140), t3 = __byte_perm(q2, q3, 0x7362);
    uint32_t qT[4];                                           // qT[j]: byte r = cycle j of read r
    qT[0] = __byte_perm(t0, t2, 0x5410); qT[1] = __byte_perm(t0, t2, 0x7632);
    qT[2] = __byte_perm(t1, t3, 0x5410); qT[3] = __byte_perm(t1, t3, 0x7632);
    const uint32_t wc = __byte_perm(__byte_perm(b0, b1, 0x3340), __byte_perm(b2, b3, 0x3340), 0x5410);   // byte r = codes of read r
    const uint32_t wc1 = wc >> 1;
    const uint32_t ones = 0x01010101u;
    uint32_t q7[4], t7[4], s6[4];
#pragma unroll
    for(int j = 0; j < 4; ++j){
        const uint32_t one = ones << (2 * j);
        q7[j] = qT[j] & 0x7F7F7F7Fu;
        t7[j] = (q7[j] + 0x4A4A4A4Au) & 0x80808080u;          // q >= 54 (q > '5')
        s6[j] = q7[j] & 0x40404040u;                          // q >= 64 (q > '?')
        uint32_t m[3];
        m[0] = wc & one; m[1] = wc & (one << 1); m[2] = wc & wc1 & one;
        A.a20[j] = dp4(t7[j], ones, A.a20[j]);
        A.a30[j] = dp4(s6[j], ones, A.a30[j]);
        A.aqs[j] = dp4(q7[j], ones, A.aqs[j]);
#pragma unroll
        for(int s = 0; s < 3; ++s){
            A.n[s][j] = dp4(m[s], ones, A.n[s][j]);
            A.n20[s][j] = dp4(t7[j], m[s], A.n20[s][j]);
            A.n30[s][j] = dp4(s6[j], m[s], A.n30[s][j]);
            A.qs[s][j] = dp4(q7[j], m[s], A.qs[s][j]);
        }
    }
    if((qT[0] | qT[1] | qT[2] | qT[3]) & 0x80808080u){        // rare: N bases among the 16
#pragma unroll
        for(int j = 0; j < 4; ++j){
            const uint32_t mn = (qT[j] >> 7) & ones;
            A.nn[j] = dp4(mn, ones, A.nn[j]);
            A.nn20[j] = dp4(t7[j], mn, A.nn20[j]);
            A.nn30[j] = dp4(s6[j], mn, A.nn30[j]);
            A.nqs[j] = dp4(q7[j], mn, A.nqs[j]);
        }
    }
}
// accumulators -> shared arrays of both Stats objects (cycles 4w .. 4w+3); v = the 76 accumulators, laid out by quad_drain
__device__ __noinline__ void quad_drain1(const uint32_t* v, uint32_t* smA, uint32_t* smB, int C, int w){
    for(int j = 0; j < 4; ++j){
        const int c = 4 * w + j;
        sm_add(smA, smB, C, KN20, S_ALL, c, v[0 + j] >> 7);
        sm_add(smA, smB, C, KN30, S_ALL, c, v[4 + j] >> 6);
        sm_add(smA, smB, C, KQS, S_ALL, c, v[8 + j]);
        for(int s = 0; s < 3; ++s){
            const int sh = 2 * j + (s == 1 ? 1 : 0);
            sm_add(smA, smB, C, KN, S_B0 + s, c, v[12 + s * 4 + j] >> sh);
            sm_add(smA, smB, C, KN20, S_B0 + s, c, v[24 + s * 4 + j] >> (7 + sh));
            sm_add(smA, smB, C, KN30, S_B0 + s, c, v[36 + s * 4 + j] >> (6 + sh));
            sm_add(smA, smB, C, KQS, S_B0 + s, c, v[48 + s * 4 + j] >> sh);
        }
        sm_add(smA, smB, C, KN, S_N, c, v[60 + j]);
        sm_add(smA, smB, C, KN20, S_N, c, v[64 + j] >> 7);
        sm_add(smA, smB, C, KN30, S_N, c, v[68 + j] >> 6);
        sm_add(smA, smB, C, KQS, S_N, c, v[72 + j]);
    }
}
__device__ __forceinline__ void quad_drain(QuadAcc& A, uint32_t* smA, uint32_t* smB, int C, int w){
    uint32_t v[76];
#pragma unroll
    for(int j = 0; j < 4; ++j){
        v[j] = A.a20[j]; v[4 + j] = A.a30[j]; v[8 + j] = A.aqs[j];
#pragma unroll
        for(int s = 0; s < 3; ++s){
            v[12 + s * 4 + j] = A.n[s][j]; v[24 + s * 4 + j] = A.n20[s][j]; v[36 + s * 4 + j] = A.n30[s][j]; v[48 + s * 4 + j] = A.qs[s][j];
        }
        v[60 + j] = A.nn[j]; v[64 + j] = A.nn20[j]; v[68 + j] = A.nn30[j]; v[72 + j] = A.nqs[j];
    }
    quad_drain1(v, smA, smB, C, w);
    A.zero();
}

template <int MODE, bool KMER, int NSTAT>
__global__ void __launch_bounds__(NSTAT + ws_trim_threads(NSTAT), 1)
k_fused_se_ws(const __grid_constant__ DevParams D, DevBatch B, fq_read_result* __restrict__ out,
              uint64_t* __restrict__ flat, int W, int T, int reads_per_block, const StatsMode M){
    constexpr int WS_STATS = NSTAT, WS_TRIM = (NSTAT == 128 ? 512 : 256), WS_THREADS = WS_STATS + WS_TRIM;
    extern __shared__ __align__(128) uint8_t smraw[];
    const fq_params& P = D.P;
    const int C = 4 * W, Cp = (C + 15) & ~15;
    const int statw = 4 * NSETS * C + C + 1;
    uint32_t* smPre = reinterpret_cast<uint32_t*>(smraw);
    uint32_t* smPost = smPre + statw;
    BlockCounters* bc = reinterpret_cast<BlockCounters*>(smPost + statw);
    uint64_t* mbar = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(bc + 1) + 7) & ~uintptr_t(7));
    uint64_t* full = mbar; uint64_t* ready = mbar + NS; uint64_t* freeb = mbar + 2 * NS;
    uint64_t* metab = mbar + 3 * NS;                                     // offsets/lengths of a tile landed
    uint32_t* zero_word = reinterpret_cast<uint32_t*>(mbar + 3 * NS + MS);
    uint32_t* blkmax = zero_word + 2;                                    // longest raw read of this block's share
    // k-mer histograms privatised per block when 3 x 4^k counters fit (k <= 6): common / pre-only / post-only
    const uint32_t kbins = (KMER && P.kmer_len <= KMER_SMEM_MAX) ? (1u << (2 * P.kmer_len)) : 0u;
    uint32_t* khist = reinterpret_cast<uint32_t*>((reinterpret_cast<uintptr_t>(zero_word + 4) + 15) & ~uintptr_t(15));
    uint8_t* sp = reinterpret_cast<uint8_t*>(khist + 3 * kbins);
    // stage layout: q[qcap] b[bcap] clist[T] olist[2T] cnt[4]; after the NS stages a ring of MS metadata slots off[T] len[T]
    // clist: full-width common reads from the front, shorter common reads from the back
    const uint32_t qcap = (uint32_t)T * Cp, bcap = ((qcap / 4 + 15u) & ~15u) + 16;
    const uint32_t mo = (T * 4 + 15) & ~15, ml = (T * 2 + 15) & ~15;
    const uint32_t o_b = qcap, o_cl = o_b + bcap, o_ol = o_cl + mo, o_cnt = o_ol + 2 * mo;
    const uint32_t stage_bytes = o_cnt + 16, meta_bytes = mo + ml;
    uint8_t* metap = sp + (size_t)NS * stage_bytes;
    const uint32_t sp32 = smem_u32(sp), zero_addr = smem_u32(zero_word);

    for(int i = threadIdx.x; i < 2 * statw; i += WS_THREADS) smPre[i] = 0;
    for(int i = threadIdx.x; i < (int)(3 * kbins); i += WS_THREADS) khist[i] = 0;
    counters_zero(bc);
    if(threadIdx.x == 0){
        for(int s = 0; s < NS; ++s){ mbar_init(&full[s], 1); mbar_init(&ready[s], WS_TRIM); mbar_init(&freeb[s], WS_STATS); }
        for(int s = 0; s < MS; ++s) mbar_init(&metab[s], 1);
        zero_word[0] = 0; zero_word[1] = 0; blkmax[0] = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const uint32_t lo = (uint32_t)min((uint64_t)B.n, (uint64_t)blockIdx.x * reads_per_block);
    const uint32_t hi = (uint32_t)min((uint64_t)B.n, (uint64_t)lo + reads_per_block);
    const int ntiles = (int)((hi - lo + T - 1) / T);

    // threads per read (Wr): enough 4-cycle words for the longest raw read this block will see.  Paired-end:
    // the batch maximum from k_trim_pe (max_cycles also has to hold merged reads, twice the raw length).
    int Wr = W;
    if(MODE == 0){
        uint32_t mx = 0;
        for(uint32_t i = lo + threadIdx.x; i < hi; i += WS_THREADS) mx = max(mx, (uint32_t)B.lengths[i]);
#pragma unroll
        for(int o = 16; o; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if((threadIdx.x & 31) == 0) atomicMax(blkmax, mx);
        __syncthreads();
        Wr = min(W, max(8, (int)((blkmax[0] + 3u) >> 2)));
    }else if(MODE == 1 && M.maxlen) Wr = min(W, max(8, (int)((*M.maxlen + 3u) >> 2)));
    const uint32_t fullmin = 4u * (uint32_t)(Wr - 1);                    // a read longer than this has a byte in every word

    if(threadIdx.x >= WS_STATS){
        // ============================== trim warps ==============================
        // registers released here are what the statistics warps can claim: 512 x (96 - 72) = 128 x (192 - 96), or
        // 256 x (128 - 80) >= 256 x (168 - 128); asking for more than was released blocks forever
        if(NSTAT == 128) asm volatile("setmaxnreg.dec.sync.aligned.u32 72;\n");
        else asm volatile("setmaxnreg.dec.sync.aligned.u32 80;\n");
        const int tt = threadIdx.x - WS_STATS;
        Tally tl = {};
        const bool fast_filter = !P.complexity_enabled;
        // One producer thread.  Metadata (offsets, lengths) of tile k is bulk-copied into ring slot k % MS one
        // tile ahead of the data, whose two bulk copies need the first and last offset of the tile.
        auto issue_meta = [&](int k){
            const int m = k % MS;
            uint32_t* so = reinterpret_cast<uint32_t*>(metap + (size_t)m * meta_bytes);
            uint16_t* sl = reinterpret_cast<uint16_t*>(metap + (size_t)m * meta_bytes + mo);
            const uint32_t t0 = lo + (uint32_t)k * T;
            const uint32_t nt = min((uint32_t)T, hi - t0), nb = nt & ~7u;            // bulk copies move multiples of 16 bytes
            for(uint32_t i = nb; i < nt; ++i){ so[i] = B.offsets[t0 + i]; sl[i] = B.lengths[t0 + i]; }   // last tile of the batch
            mbar_expect_tx(&metab[m], nb * 6u);
            if(nb){
                bulk_g2s(so, B.offsets + t0, nb * 4u, &metab[m]);
                bulk_g2s(sl, B.lengths + t0, nb * 2u, &metab[m]);
            }
        };
        auto issue_data = [&](int k){
            const int s = k % NS, m = k % MS;
            uint8_t* S = sp + (size_t)s * stage_bytes;
            const uint32_t* so = reinterpret_cast<const uint32_t*>(metap + (size_t)m * meta_bytes);
            const uint16_t* sl = reinterpret_cast<const uint16_t*>(metap + (size_t)m * meta_bytes + mo);
            const uint32_t t0 = lo + (uint32_t)k * T;
            const int nt = (int)min((uint32_t)T, hi - t0);
            mbar_wait_sleep(&metab[m], (uint32_t)(k / MS) & 1u, 20000);
            uint32_t* c = reinterpret_cast<uint32_t*>(S + o_cnt); c[0] = 0; c[1] = 0; c[2] = 0;
            const uint32_t a = so[0];
            uint64_t e = (uint64_t)so[nt - 1] + ((sl[nt - 1] + 15u) & ~15u);
            const bool bad = (a & 15u) || e < a || e - a > qcap || e > B.total;
            if(bad){ atomicOr(D.err, 2u); e = a; }
            const uint32_t qbytes = (uint32_t)(e - a);
            const uint32_t b0 = (a >> 2) & ~15u;
            const uint32_t bbytes = qbytes ? (((uint32_t)(e >> 2) + 15u) & ~15u) - b0 : 0u;
            mbar_expect_tx(&full[s], qbytes + bbytes);
            if(qbytes){
                bulk_g2s(S, B.qual + a, qbytes, &full[s]);
                bulk_g2s(S + o_b, B.bases2 + b0, bbytes, &full[s]);
            }
        };
        if(tt == 0){
            for(int k = 0; k < min(3, ntiles); ++k) issue_meta(k);
            for(int k = 0; k < min(2, ntiles); ++k) issue_data(k);
        }
        for(int k = 0; k < ntiles; ++k){
            const int s = k % NS;
            uint8_t* S = sp + (size_t)s * stage_bytes;
            const uint32_t* So = reinterpret_cast<const uint32_t*>(metap + (size_t)(k % MS) * meta_bytes);
            const uint16_t* Sl = reinterpret_cast<const uint16_t*>(metap + (size_t)(k % MS) * meta_bytes + mo);
            uint32_t* clist = reinterpret_cast<uint32_t*>(S + o_cl);
            uint2* olist = reinterpret_cast<uint2*>(S + o_ol);
            uint32_t* cnt = reinterpret_cast<uint32_t*>(S + o_cnt);
            const uint32_t t0 = lo + (uint32_t)k * T;
            const int nt = (int)min((uint32_t)T, hi - t0);
            mbar_wait_sleep(&metab[k % MS], (uint32_t)(k / MS) & 1u, 20000);
            mbar_wait_sleep(&full[s], (uint32_t)(k / NS) & 1u, 20000);
            const uint32_t a0 = So[0];
            const uint32_t bshift = (a0 >> 2) & 15u;
            // read i of tile k goes to thread (i + k*T) mod WS_TRIM: when T is not a multiple of the trim threads the
            // remainder lands on different warps tile after tile (the warps run up to two tiles apart)
            const int rot = (int)(((uint32_t)k * (uint32_t)T) % (uint32_t)WS_TRIM);
            for(int i = (tt - rot + WS_TRIM) % WS_TRIM; i < nt; i += WS_TRIM){
                uint32_t doff = So[i] - a0, dlen = Sl[i];
                const uint32_t padded = (dlen + 15u) & ~15u;
                bool ok = !(So[i] & 15u) && So[i] >= a0 && doff + padded <= qcap && dlen <= (uint32_t)C;
                if(ok && i + 1 < nt) ok = (So[i + 1] == So[i] + padded);            // dense, ordered batch
                if(!ok){ atomicOr(D.err, dlen > (uint32_t)C ? 1u : 2u); doff = 0; dlen = 0; }
                uint32_t hasN = 0;                                    // k-mer validity needs to know whether the raw read holds any N
                if(KMER){
                    uint32_t any = 0;
                    const uint32_t qa = smem_u32(S + doff);
                    for(uint32_t c = 0; c < padded; c += 16){ const uint4 v = lds128(qa + c); any |= v.x | v.y | v.z | v.w; }
                    hasN = (any & 0x80808080u) ? 1u : 0u;
                }
                if(MODE == 0){
                    const fq_read_result rec = phase1_read(D, S + doff, S + o_b + bshift + (doff >> 2), (int)dlen, fast_filter, tl);
                    out[t0 + i] = rec;
                    atomicAdd(&bc->filter[rec.result], 1u);
                    const bool pass = !(rec.flags & FQK_F_DROPPED) && rec.result == FQK_PASS_FILTER;
                    atomicAdd(smPre + 4 * NSETS * C + dlen, 1u);
                    if(pass) atomicAdd(smPost + 4 * NSETS * C + rec.len, 1u);
                    if(pass && rec.start == 0 && rec.len == dlen){
                        const uint32_t e = doff | hasN | (dlen << 17);
                        if(dlen > fullmin) clist[atomicAdd(&cnt[0], 1u)] = e;
                        else clist[T - 1 - atomicAdd(&cnt[2], 1u)] = e;
                    }else{
                        const Desc d = make_desc(doff, dlen, rec.start, rec.len, pass ? 2u : 1u);
                        olist[atomicAdd(&cnt[1], 1u)] = make_uint2(d.x, d.y | hasN << 28);
                    }
                }else{
                    const fq_read_result ra = M.recA[t0 + i], rb = M.recB[t0 + i];
                    const fq_pair_result pr = M.recP[t0 + i];
                    const bool alive = !(ra.flags & FQK_F_DROPPED) && !(rb.flags & FQK_F_DROPPED);
                    if(MODE == 1){
                        bool own = false; uint32_t start = ra.start, plen = ra.len;
                        if(alive){
                            if(!M.merge_enabled) own = ra.result == FQK_PASS_FILTER && rb.result == FQK_PASS_FILTER;   // :392-401
                            else if(pr.flags & FQK_P_OV2_OVERLAP){                                                      // :355-366
                                if((pr.flags & FQK_P_MERGED) && M.mate == 0){ own = true; plen = pr.merge_len1; }       // read1 prefix of the merged read
                            }else if(pr.flags & FQK_P_MERGE_PROC) own = ra.result == FQK_PASS_FILTER;                  // :367-380
                        }
                        atomicAdd(smPre + 4 * NSETS * C + dlen, 1u);
                        if(own) atomicAdd(smPost + 4 * NSETS * C + plen, 1u);
                        if(own && start == 0 && plen == dlen){
                            const uint32_t e = doff | hasN | (dlen << 17);
                            if(dlen > fullmin) clist[atomicAdd(&cnt[0], 1u)] = e;
                            else clist[T - 1 - atomicAdd(&cnt[2], 1u)] = e;
                        }else{
                            const Desc d = make_desc(doff, dlen, start, plen, own ? 2u : 1u);
                            olist[atomicAdd(&cnt[1], 1u)] = make_uint2(d.x, d.y | hasN << 28);
                        }
                    }else if(alive && (pr.flags & FQK_P_MERGED) && pr.merge_len2 > 0){
                        // tail of the merged read: revcomp(read2 kept part)[ol, ol+len2) lands on cycles [len1, len1+len2);
                        // cycle c reads original position pmax - c of read2
                        const uint32_t len1 = pr.merge_len1, len2 = pr.merge_len2;
                        const uint32_t pmax = (uint32_t)ra.start + ra.len - 1u - pr.ov2_len + len1;
                        atomicAdd(smPost + 4 * NSETS * C + len1 + len2, 1u);
                        atomicAdd(smPost + 4 * NSETS * C + len1, 0xFFFFFFFFu);          // -1: the tail does not cover [0, len1)
                        const Desc d = make_desc(doff, pmax, len1, len2, 3u);
                        olist[atomicAdd(&cnt[1], 1u)] = make_uint2(d.x, d.y);
                    }
                }
            }
            mbar_arrive(&ready[s]);
            if(tt == 0 && k + 2 < ntiles){
                // stage of tile k+2 = stage of tile k-1: once the statistics warps released it, every thread of the block is
                // done with tile k-1, whose metadata slot is the one tile k+3 lands in
                const int k2 = k + 2, n2 = k2 / NS;
                if(n2 >= 1) mbar_wait_sleep(&freeb[k2 % NS], (uint32_t)(n2 - 1) & 1u, 20000);
                if(k + 3 < ntiles) issue_meta(k + 3);
                issue_data(k2);
            }
        }
        counters_add_tally(bc, tl);
    }else{
        // ============================ statistics warps ============================
        if(NSTAT == 128) asm volatile("setmaxnreg.inc.sync.aligned.u32 192;\n");
        else asm volatile("setmaxnreg.inc.sync.aligned.u32 168;\n");
        const int G = WS_STATS / Wr;
        const int g = threadIdx.x / Wr, w = threadIdx.x - g * Wr;
        const bool active = g < G;
        QuadAcc qa;
        Acc1 preo, posto;
        qa.zero(); preo.zero(); posto.zero();
        int usedQ = 0, leftO = 255;
        for(int k = 0; k < ntiles; ++k){
            const int s = k % NS;
            const uint32_t S32 = sp32 + (uint32_t)s * stage_bytes;
            mbar_wait_sleep(&ready[s], (uint32_t)(k / NS) & 1u, 20000);
            if(active){
                const uint32_t a0 = lds32(smem_u32(metap) + (uint32_t)(k % MS) * meta_bytes);
                const uint32_t bshift = (a0 >> 2) & 15u;
                const uint32_t sq0 = S32, sb0 = S32 + o_b + bshift;
                const uint32_t sqw = sq0 + 4 * w, sbw = sb0 + w;
                const int nc = (int)lds32(S32 + o_cnt), no = (int)lds32(S32 + o_cnt + 4), nr = (int)lds32(S32 + o_cnt + 8);
                const uint32_t cl = S32 + o_cl;
                // ---- common reads: contribute identically to pre1 and post1, four at a time ----
                {
                    const int nq = nc >> 2, rem = nc & 3, items = rem + nr;
                    const int mine = (nq > g ? (nq - g + G - 1) / G : 0) + (items + 3) / 4;
                    if(usedQ + mine > QUAD_LIMIT){ quad_drain(qa, smPre, smPost, C, w); usedQ = 0; }
                    usedQ += mine;
                    // full-width reads: every thread of the group has a word inside each read, no range checks
                    uint32_t q[4] = {0, 0, 0, 0}, b[4] = {0, 0, 0, 0};
                    auto fetch = [&](int qd){
                        const uint4 e = lds128(cl + 16 * qd);
                        const uint32_t o0 = e.x & 0x1FFF0u, o1 = e.y & 0x1FFF0u, o2 = e.z & 0x1FFF0u, o3 = e.w & 0x1FFF0u;
                        q[0] = lds32(sqw + o0); q[1] = lds32(sqw + o1); q[2] = lds32(sqw + o2); q[3] = lds32(sqw + o3);
                        b[0] = lds8(sbw + (o0 >> 2)); b[1] = lds8(sbw + (o1 >> 2)); b[2] = lds8(sbw + (o2 >> 2)); b[3] = lds8(sbw + (o3 >> 2));
                    };
                    if(g < nq) fetch(g);
                    for(int qd = g; qd < nq; qd += G){
                        const uint32_t c0 = q[0], c1 = q[1], c2 = q[2], c3 = q[3], d0 = b[0], d1 = b[1], d2 = b[2], d3 = b[3];
                        if(qd + G < nq) fetch(qd + G);
                        quad_accumulate(qa, c0, c1, c2, c3, d0, d1, d2, d3);
                    }
                    // the last full-width reads and the shorter ones: same reduction with range checks
                    for(int qd = g; 4 * qd < items; qd += G){
#pragma unroll
                        for(int r = 0; r < 4; ++r){
                            const int it = 4 * qd + r;
                            uint32_t e = 0;
                            if(it < rem) e = lds32(cl + 4 * (4 * nq + it));
                            else if(it < items) e = lds32(cl + 4 * (T - 1 - (it - rem)));
                            const uint32_t off = e & 0x1FFF0u, len = e >> 17;
                            const bool in = (uint32_t)(4 * w) < len;
                            q[r] = in ? lds32(sqw + off) : 0u;
                            b[r] = in ? lds8(sbw + (off >> 2)) : 0u;
                        }
                        quad_accumulate(qa, q[0], q[1], q[2], q[3], b[0], b[1], b[2], b[3]);
                    }
                }
                // ---- other reads: raw part -> pre1, shifted kept part of passing reads -> post1 ----
                const uint32_t ol = S32 + o_ol;
                for(int i = g; i < no; i += G){
                    const uint2 dc = lds64(ol + 8 * i);
                    const uint32_t off = (dc.x & 0x1FFFFu) << 4, len = dc.x >> 17, kind = (dc.y >> 26) & 3u;
                    if(kind == 3){
                        const int len1 = (int)(dc.y & 0x1FFFu), len2 = (int)((dc.y >> 13) & 0x1FFFu), pmax = (int)len;
                        const int c0 = 4 * w, lo = max(len1, c0), hi = min(len1 + len2, c0 + 4);
                        if(lo < hi){
                            const int addr = (int)off + pmax - c0 - 3;                  // tile position of the byte for cycle c0+3
                            const int a1 = addr & ~3, sh = addr & 3;
                            const uint32_t q4 = __funnelshift_r(lds32(sq0 + a1), lds32(sq0 + a1 + 4), sh * 8);
                            uint32_t q2 = __byte_perm(q4, 0, 0x0123);                  // reversed: byte j <-> cycle c0+j
                            const uint32_t cb = lds8(sb0 + (addr >> 2)) | (lds8(sb0 + (addr >> 2) + 1) << 8);
                            uint32_t b2 = (cb >> (2 * sh)) & 0xFF;
                            b2 = ((b2 & 0x03) << 6) | ((b2 & 0x0C) << 2) | ((b2 & 0x30) >> 2) | ((b2 & 0xC0) >> 6);
                            b2 ^= 0x55;                                                 // complement: A<->T, C<->G
                            const int jlo = lo - c0, jhi = hi - c0;
                            const uint32_t bm = (jhi >= 4 ? 0xFFFFFFFFu : ((1u << (8 * jhi)) - 1u)) & ~((1u << (8 * jlo)) - 1u);
                            q2 &= bm;
                            const uint32_t f = ((((q2 >> 7) & 0x01010101u) * 0x00204081u) >> 21) & 0xFu;   // N flags of the 4 bytes
                            uint32_t sp = (f & 1u) | ((f & 2u) << 1) | ((f & 4u) << 2) | ((f & 8u) << 3);
                            sp |= sp << 1;
                            b2 &= ((1u << (2 * jhi)) - 1u) & ~((1u << (2 * jlo)) - 1u) & ~sp;   // N keeps code 0
                            acc_word(posto, q2, b2);
                        }
                        if(--leftO == 0){ drain_l1(preo, smPre, nullptr, C, w); drain_l1(posto, smPost, nullptr, C, w); leftO = 255; }
                        continue;
                    }
                    if((uint32_t)(4 * w) < len) acc_word(preo, lds32(sqw + off), lds8(sbw + (off >> 2)));
                    if(kind == 2){
                        const uint32_t start = dc.y & 0x1FFFu, plen = (dc.y >> 13) & 0x1FFFu;
                        const int nv = (int)plen - 4 * w;
                        if(nv > 0){
                            const uint32_t p = off + start + 4 * w;
                            const uint32_t a1 = p & ~3u, sh = p & 3u;
                            const uint32_t w0 = lds32(sq0 + a1);
                            const uint32_t w1 = sh ? lds32(sq0 + a1 + 4) : 0u;
                            uint32_t q2 = __funnelshift_r(w0, w1, sh * 8);
                            const uint32_t c0 = lds8(sb0 + (p >> 2)), c1 = sh ? lds8(sb0 + (p >> 2) + 1) : 0u;
                            uint32_t b2 = ((c0 | (c1 << 8)) >> (2 * sh)) & 0xFF;
                            if(nv < 4){ q2 &= (1u << (8 * nv)) - 1u; b2 &= (1u << (2 * nv)) - 1u; }
                            acc_word(posto, q2, b2);
                        }
                    }
                    if(--leftO == 0){ drain_l1(preo, smPre, nullptr, C, w); drain_l1(posto, smPost, nullptr, C, w); leftO = 255; }
                }
                // ---- k-mer histograms (Stats::statRead, src/stats.cpp:266-274): a separate sweep over the two lists ----
                if(KMER){
                    uint64_t* kpre = flat + D.L.stats_off[M.pre_obj] + D.L.s_kmer;
                    uint64_t* kpost = flat + D.L.stats_off[M.post_obj] + D.L.s_kmer;
                    const uint32_t kh32 = kbins ? smem_u32(khist) : 0u;
                    for(int i = g; i < nc + nr; i += G){
                        const uint32_t e = lds32(cl + 4 * (i < nc ? i : T - 1 - (i - nc)));
                        kmers_word(sq0, sb0, e & 0x1FFF0u, 0, e >> 17, e & 1u, w, P.kmer_len, kpre, kpost, kh32);
                    }
                    for(int i = g; i < no; i += G){
                        const uint2 dc = lds64(ol + 8 * i);
                        const uint32_t off = (dc.x & 0x1FFFFu) << 4, len = dc.x >> 17, kind = (dc.y >> 26) & 3u, hn = (dc.y >> 28) & 1u;
                        if(kind == 3) continue;
                        kmers_word(sq0, sb0, off, 0, len, hn, w, P.kmer_len, kpre, nullptr, kh32 ? kh32 + 4 * kbins : 0u);
                        if(kind == 2) kmers_word(sq0, sb0, off, dc.y & 0x1FFFu, (dc.y >> 13) & 0x1FFFu, hn, w, P.kmer_len, kpost, nullptr,
                                                 kh32 ? kh32 + 8 * kbins : 0u);
                    }
                }
            }
            mbar_arrive(&freeb[s]);
        }
        if(active){
            quad_drain(qa, smPre, smPost, C, w);
            drain_l1(preo, smPre, nullptr, C, w); drain_l1(posto, smPost, nullptr, C, w);
        }
    }
    __syncthreads();
    if(kbins){
        uint64_t* kpre = flat + D.L.stats_off[M.pre_obj] + D.L.s_kmer;
        uint64_t* kpost = flat + D.L.stats_off[M.post_obj] + D.L.s_kmer;
        for(uint32_t i = threadIdx.x; i < kbins; i += WS_THREADS){
            const uint32_t c = khist[i], a = khist[kbins + i], b = khist[2 * kbins + i];
            if(c + a) add_u64(kpre + i, (uint64_t)c + a);
            if(c + b) add_u64(kpost + i, (uint64_t)c + b);
        }
    }
    if(MODE != 2) retire(smPre, C, D.L, flat + D.L.stats_off[M.pre_obj]);
    retire(smPost, C, D.L, flat + D.L.stats_off[M.post_obj]);
    counters_flush(bc, D.L, flat);
}

size_t ws_smem_bytes(int W, int T, int K = 0){
    const int C = 4 * W, Cp = (C + 15) & ~15;
    size_t s = (size_t)2 * (4 * NSETS * C + C + 1) * 4 + sizeof(BlockCounters) + 8 + (3 * NS + MS) * 8 + 16 + 16;
    if(K > 0 && K <= KMER_SMEM_MAX) s += (size_t)3 * 4 * ((size_t)1 << (2 * K));
    const size_t qcap = (size_t)T * Cp, bcap = ((qcap / 4 + 15) & ~(size_t)15) + 16;
    const size_t mo = (T * 4 + 15) & ~15, ml = (T * 2 + 15) & ~15;
    s += NS * (qcap + bcap + mo + 2 * mo + 16) + MS * (mo + ml);
    return s + 128;
}

size_t g_smem_optin = 0;

#include "fqk_se3.cuh"
#include "fqk_pe4.cuh"

// largest tile (multiple of 32 reads, whole quads per statistics group preferred) whose NS stages fit
int v3_tile(int W, int nstat = V3_STATS){
    const size_t budget = g_smem_optin > 4096 ? g_smem_optin - 1024 : 0;
    if(W > nstat) return 0;
    const int G = nstat / W;
    int fallback = 0;
    for(int T = 256; T >= 32; T -= 32){
        const V3L l = v3_layout(W, T);
        if(l.qcap >= (1u << 17) || l.total + 128 > budget) continue;   // exception entries hold 17-bit tile offsets
        if((T / 4) % G == 0) return T;
        if(!fallback) fallback = T;
    }
    return fallback;
}
bool v3_fast_options(const fq_params& P){
    return !P.polyg_enabled && !P.polyx_enabled && !(P.adapter_trim_enabled && P.adapter_r1_len > 0) && !P.complexity_enabled &&
           !(P.average_quality_limit > 0) && P.trim_max_len1 <= 0 && P.low_quality_limit >= 1 && P.low_quality_limit <= 128;
}

} // namespace

static int ws_tile(int W, int K = 0);

void fused_configure(const cudaDeviceProp& prop){
    g_smem_optin = prop.sharedMemPerBlockOptin;
    cudaFuncSetAttribute(k_fused_se_ws<0, false, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_fused_se_ws<0, true, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_fused_se_ws<0, false, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_fused_se_ws<0, true, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_fused_se_ws<1, false, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_fused_se_ws<1, true, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_fused_se_ws<2, false, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_se_v3<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_se_v3<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_se_v3<false, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_se_v3<true, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_pe_v4<V4Spec<true>, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_pe_v4<V4Spec<true>, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_pe_v4<V4Spec<false>, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
    cudaFuncSetAttribute(k_pe_v4<V4Spec<true>, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_smem_optin);
}

bool fused_supported_se(const DevParams& D, const DevBatch& B){
    if(D.P.kmer_len > 13) return false;                    // k-mers longer than the 16-base window: generic kernels
    const int W = (int)(D.L.max_cycles + 3) / 4;
    if(D.P.kmer_len && ws_tile(W, D.P.kmer_len) <= 0) return false;   // only the warp-specialised kernel counts k-mers
    if(W < 8 || W > 256) return false;
    if((reinterpret_cast<uintptr_t>(B.qual) | reinterpret_cast<uintptr_t>(B.bases2) | reinterpret_cast<uintptr_t>(B.offsets) |
        reinterpret_cast<uintptr_t>(B.lengths)) & 15) return false;          // bulk copies: 16-byte aligned sources
    if(B.total % FQK_BATCH_PAD) return false;
    return ws_tile(W, D.P.kmer_len) > 0 || v3_tile(W) > 0;
}

static int ws_tile(int W, int K){
    const size_t budget = g_smem_optin > 4096 ? g_smem_optin - 2048 : 0;
    if(W > 256) return 0;                                      // a read has to fit the statistics threads of a block
    for(int T = 512; T >= 32; T -= 16){
        const size_t qcap = (size_t)T * ((4 * W + 15) & ~15);
        if(qcap < (1u << 17) && ws_smem_bytes(W, T, K) <= budget) return T;      // list entries hold 17-bit tile offsets
    }
    return 0;
}

int fused_launch_se(const DevParams& D, const DevBatch& B, fq_read_result* out, uint64_t* flat, int sms, cudaStream_t st,
                    int variant){
    const int W = (int)(D.L.max_cycles + 3) / 4;
    if(variant == 0 && !D.P.kmer_len && v3_tile(W) > 0){
        // generic option sets (adapter by sequence, polyG / polyX, ...): the trim role is the long one (one lane per read,
        // dependent scalar chains), so it gets 11 working warps and the statistics 4 (FQK_V3_WIDE=0 keeps 7 + 8)
        static const bool wide_on = !(fqk_tuning_env("FQK_V3_WIDE") && atoi(fqk_tuning_env("FQK_V3_WIDE")) == 0);
        static const int wide_fast = fqk_tuning_env("FQK_V3_WIDE_FAST") ? atoi(fqk_tuning_env("FQK_V3_WIDE_FAST")) : 0;   // experiment: 4 + 11 warps for the fast option sets too
        const bool wide = wide_on && (!v3_fast_options(D.P) || wide_fast) && v3_tile(W, 128) > 0;
        const int T = wide ? v3_tile(W, 128) : v3_tile(W);
        int rpb = (int)((B.n + sms - 1) / sms);
        rpb = std::min(rpb, 1 << 24);                          // 32-bit per-block sums of quality characters
        rpb = (rpb + T - 1) / T * T;
        const int grid = (int)((B.n + rpb - 1) / rpb);
        const size_t sm = v3_layout(W, T).total + 128;
        static const int dbg = fqk_tuning_env("FQK_V3_DEBUG") ? atoi(fqk_tuning_env("FQK_V3_DEBUG")) : 0;   // timing experiments: 1 no trim work, 2 no statistics work
        if(v3_fast_options(D.P) && wide) k_se_v3<true, 128><<<grid, V3_THREADS, sm, st>>>(D, B, out, flat, W, T, rpb, dbg);
        else if(v3_fast_options(D.P)) k_se_v3<true><<<grid, V3_THREADS, sm, st>>>(D, B, out, flat, W, T, rpb, dbg);
        else if(wide) k_se_v3<false, 128><<<grid, V3_THREADS, sm, st>>>(D, B, out, flat, W, T, rpb, dbg);
        else k_se_v3<false><<<grid, V3_THREADS, sm, st>>>(D, B, out, flat, W, T, rpb, dbg);
        return cudaGetLastError() == cudaSuccess ? 1 : -1;
    }
    if((variant == 0 || variant == 3 || D.P.kmer_len) && ws_tile(W, D.P.kmer_len) > 0){
        const int T = ws_tile(W, D.P.kmer_len);
        int rpb = (int)((B.n + sms - 1) / sms);
        rpb = std::min(rpb, 1 << 24);                          // 32-bit per-block sums of quality characters
        rpb = (rpb + T - 1) / T * T;
        const int grid = (int)((B.n + rpb - 1) / rpb);
        const StatsMode M{0, 0, 0, 0, 2, nullptr, nullptr, nullptr, nullptr};
        const size_t sm = ws_smem_bytes(W, T, D.P.kmer_len);
        static const int split128 = fqk_tuning_env("FQK_WS_SPLIT128") ? 1 : 0;   // tuning aid: 4 + 16 warps measured slower than 8 + 8 on B200
        if(W <= 128 && split128){                              // a read fits the 128 statistics threads
            if(D.P.kmer_len) k_fused_se_ws<0, true, 128><<<grid, 640, sm, st>>>(D, B, out, flat, W, T, rpb, M);
            else k_fused_se_ws<0, false, 128><<<grid, 640, sm, st>>>(D, B, out, flat, W, T, rpb, M);
        }else{
            if(D.P.kmer_len) k_fused_se_ws<0, true, 256><<<grid, 512, sm, st>>>(D, B, out, flat, W, T, rpb, M);
            else k_fused_se_ws<0, false, 256><<<grid, 512, sm, st>>>(D, B, out, flat, W, T, rpb, M);
        }
        return cudaGetLastError() == cudaSuccess ? 1 : -1;
    }
    return -1;
}

// paired-end statistics from the records of k_trim_pe: read1 pass, read2 pass and (with -m) the merged tails
bool ws_stats_supported_pe(const DevParams& D, const DevBatch& B1, const DevBatch& B2){
    if(D.P.kmer_len && D.P.merge_enabled) return false;    // k-mers across the junction of a merged read: generic kernel
    if(D.P.correction_enabled) return false;               // -c: pre statistics must see the batch before it is corrected
    return fused_supported_se(D, B1) && fused_supported_se(D, B2) && ws_tile((int)(D.L.max_cycles + 3) / 4, D.P.kmer_len) > 0;
}

int ws_stats_launch_pe(const DevParams& D, const DevBatch& B1, const DevBatch& B2, const fq_read_result* r1,
                       const fq_read_result* r2, const fq_pair_result* rp, uint64_t* flat, int sms, cudaStream_t st,
                       const uint32_t* maxlen2){
    const int W = (int)(D.L.max_cycles + 3) / 4;
    const int K = D.P.kmer_len;
    const int T = ws_tile(W, K);
    int rpb = (int)((B1.n + sms - 1) / sms);
    rpb = std::min(rpb, 1 << 24);
    rpb = (rpb + T - 1) / T * T;
    const int grid = (int)((B1.n + rpb - 1) / rpb);
    const int merge = D.P.merge_enabled;
    const StatsMode M1{1, 0, merge, 0, 2, r1, r2, rp, maxlen2}, M2{1, 1, merge, 1, 3, r2, r1, rp, maxlen2 ? maxlen2 + 1 : nullptr},
                    M3{2, 1, merge, 1, 2, r2, r1, rp, nullptr};
    if(K){
        k_fused_se_ws<1, true, 256><<<grid, 512, ws_smem_bytes(W, T, K), st>>>(D, B1, nullptr, flat, W, T, rpb, M1);
        k_fused_se_ws<1, true, 256><<<grid, 512, ws_smem_bytes(W, T, K), st>>>(D, B2, nullptr, flat, W, T, rpb, M2);
    }else{
        k_fused_se_ws<1, false, 256><<<grid, 512, ws_smem_bytes(W, T, 0), st>>>(D, B1, nullptr, flat, W, T, rpb, M1);
        k_fused_se_ws<1, false, 256><<<grid, 512, ws_smem_bytes(W, T, 0), st>>>(D, B2, nullptr, flat, W, T, rpb, M2);
    }
    int n = 2;
    if(merge){ k_fused_se_ws<2, false, 256><<<grid, 512, ws_smem_bytes(W, T, 0), st>>>(D, B2, nullptr, flat, W, T, rpb, M3); ++n; }
    return cudaGetLastError() == cudaSuccess ? n : -1;
}

// fused paired-end kernel: one launch per pair of batches
static int pe4_maxlen(const DevBatch& B1, const DevBatch& B2){
    return (B1.max_len && B2.max_len) ? (int)std::max(B1.max_len, B2.max_len) : 0;
}
bool pe4_supported(const DevParams& D, const DevBatch& B1, const DevBatch& B2){
    if((reinterpret_cast<uintptr_t>(B1.qual) | reinterpret_cast<uintptr_t>(B1.bases2) | reinterpret_cast<uintptr_t>(B1.offsets) |
        reinterpret_cast<uintptr_t>(B1.lengths) | reinterpret_cast<uintptr_t>(B2.qual) | reinterpret_cast<uintptr_t>(B2.bases2) |
        reinterpret_cast<uintptr_t>(B2.offsets) | reinterpret_cast<uintptr_t>(B2.lengths)) & 15) return false;   // bulk copies: 16-byte aligned sources
    if(B1.total % FQK_BATCH_PAD || B2.total % FQK_BATCH_PAD) return false;
    const int ml = pe4_maxlen(B1, B2);
    return v4_plan(D, ml, g_smem_optin, 0).ok && v4_plan(D, ml, g_smem_optin, 1).ok;
}
// pairs per block: an equal share per SM, a whole number of tiles
static int pe4_ppb(uint32_t n, int sms, int T){
    int ppb = (int)((n + sms - 1) / sms);
    ppb = std::min(ppb, 1 << 23);                          // 32-bit per-block sums of quality characters
    return (ppb + T - 1) / T * T;
}
int pe4_launch(const DevParams& D, const DevBatch& B1, const DevBatch& B2, fq_read_result* o1, fq_read_result* o2,
               fq_pair_result* op, uint64_t* flat, int sms, cudaStream_t st){
    static const int dbg = fqk_tuning_env("FQK_V4_DEBUG") ? atoi(fqk_tuning_env("FQK_V4_DEBUG")) : 0;   // timing experiments: 1 no trim work, 2 no statistics work
    static const bool fusedOnly = fqk_tuning_env("FQK_V4_SPLIT") && atoi(fqk_tuning_env("FQK_V4_SPLIT")) == 0;   // A/B switch: the one-kernel form
    static const bool nolean = fqk_tuning_env("FQK_V4_GENERIC") != nullptr;       // A/B switch: the general code for every option set
    const int ml = pe4_maxlen(B1, B2);
    V4Plan p = v4_plan(D, ml, g_smem_optin, 0);
    if(!p.ok) return -1;
    int nl = 1;
    if(fusedOnly){
        const int ppb = pe4_ppb(B1.n, sms, p.T);
        k_pe_v4<V4Spec<true>, 0><<<(int)((B1.n + ppb - 1) / ppb), V4_THREADS, p.smem, st>>>(D, B1, B2, o1, o2, op, flat, p.Wr, p.Wm, p.T, p.NS, ppb, dbg);
    }else{
        // split path: the trim kernel writes the records, the statistics kernel reads them back
        const V4Plan q = v4_plan(D, ml, g_smem_optin, 1);
        if(!q.ok) return -1;
        const int ppb1 = pe4_ppb(B1.n, sms, q.T), grid1 = (int)((B1.n + ppb1 - 1) / ppb1);
        if(v4_lean_options(D.P) && !nolean)
            k_pe_v4<V4Spec<false>, 1><<<grid1, V4_THREADS, q.smem, st>>>(D, B1, B2, o1, o2, op, flat, q.Wr, q.Wm, q.T, q.NS, ppb1, dbg);
        else
            k_pe_v4<V4Spec<true>, 1><<<grid1, V4_THREADS, q.smem, st>>>(D, B1, B2, o1, o2, op, flat, q.Wr, q.Wm, q.T, q.NS, ppb1, dbg);
        const int ppb2 = pe4_ppb(B1.n, sms, p.T);
        k_pe_v4<V4Spec<true>, 2><<<(int)((B1.n + ppb2 - 1) / ppb2), V4_THREADS, p.smem, st>>>(D, B1, B2, o1, o2, op, flat, p.Wr, p.Wm, p.T, p.NS, ppb2, dbg & ~256);
        nl = 2;
        if(dbg & 256) p = q;
    }
    if(dbg & 256){                                         // section clocks of the trim warps (see V4_PROF)
        unsigned long long h[16];
        cudaStreamSynchronize(st);
        cudaMemcpyFromSymbol(h, g_v4prof, sizeof(h));
        fprintf(stderr, "v4prof T=%d NS=%d Wr=%d chunks=%llu |", p.T, p.NS, p.Wr, h[15]);
        for(int i = 0; i < 12; ++i) fprintf(stderr, " %d:%.0f", i, h[15] ? (double)h[i] / (double)h[15] : 0.0);
        fprintf(stderr, " | verifies/chunk %.2f\n", h[15] ? (double)h[9] / (double)h[15] : 0.0);
        static const unsigned long long z[16] = {};
        cudaMemcpyToSymbol(g_v4prof, z, sizeof(z));
    }
    return cudaGetLastError() == cudaSuccess ? nl : -1;
}
