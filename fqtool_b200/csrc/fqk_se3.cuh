// fqk_se3.cuh -- third single-end kernel (k_se_v3); included inside the anonymous namespace of fqk_fused.cu and
// built from its helpers (QuadAcc, Acc1, drains, retire, mbarrier wrappers).
//
// Same one-pass idea as k_fused_se_ws (trim + filter + pre/post Stats::statRead in one sweep over HBM), with the
// three things the profile of that kernel asked for (profiles/r01b_*):
//   * roles are decoupled at warp level.  A tile of T reads is split into chunks of 32 reads; seven trim warps
//     take chunks round-robin (a slow read delays its own warp only) and an eighth warp does nothing but issue
//     bulk copies NS-1 tiles ahead.  The eight statistics warps wait for a tile only once all its chunks are done.
//   * pre/post statistics are decomposed linearly:  pre = P,  post = P - D + Z  with P = every raw read (the
//     dp4a quad path, no lists, no per-read conditions), D = the parts of raw reads that do not reach the output
//     at their raw cycle (failed reads, trimmed tails; whole read if the front was cut) and Z = kept parts of
//     front-trimmed reads at their shifted cycle.  D and Z are short exception lists handled by the SWAR word path.
//   * the trim lanes are specialised for the common option set (quality cuts + -q/-l filters): one uniform 16-byte
//     sweep per read for passFilter's counts, four window positions per step in the cut scans.
#pragma once

#ifndef V3_NSTAGES
#define V3_NSTAGES 4
#endif
constexpr int V3_STATS = 256, V3_TRIM = 256, V3_THREADS = V3_STATS + V3_TRIM, V3_NS = V3_NSTAGES;
static_assert(2 * V3_NS + 1 <= 16, "named barriers: id 0 is __syncthreads, ids 1 .. 2*NS the stage handshakes");
#ifndef V3_STATS_HIGH
#define V3_STATS_HIGH 0
#endif
constexpr int V3_WORKERS = V3_TRIM / 32 - 1;                      // trim warps 0..6 work on chunks, warp 7 is the copy producer

struct V3L { uint32_t statw, o_bc, o_bar, o_zero, o_stage, qcap, o_b, o_off, o_len, o_ex, o_cnt, stage_bytes, total; };
__host__ __device__ inline V3L v3_layout(int W, int T){
    V3L l;
    const uint32_t C = 4u * (uint32_t)W, Cp = (C + 15u) & ~15u;
    l.statw = 4 * NSETS * C + C + 1;
    uint32_t o = 128 + 2 * l.statw * 4; o = (o + 7u) & ~7u;     // first 128 bytes: cycle accounting of FQK_V3_PROF builds
    l.o_bc = o; o += (uint32_t)sizeof(BlockCounters); o = (o + 7u) & ~7u;
    l.o_bar = o; o += 3 * V3_NS * 8;
    l.o_zero = o; o += 16;
    o = (o + 127u) & ~127u; l.o_stage = o;
    l.qcap = (uint32_t)T * Cp;
    uint32_t s = l.qcap; l.o_b = s; s += ((l.qcap / 4 + 15u) & ~15u) + 32;
    l.o_off = s; s += (uint32_t)T * 4 + 16;
    l.o_len = s; s += (((uint32_t)T * 2 + 15u) & ~15u) + 16;
    l.o_ex = s; s += (uint32_t)T * 8;
    l.o_cnt = s; s += 16;
    l.stage_bytes = (s + 127u) & ~127u;
    l.total = l.o_stage + V3_NS * l.stage_bytes;
    return l;
}

// SWAR word path without the N class (filled by the trim lanes)
__device__ __forceinline__ void acc_word_nf(Acc1& a, uint32_t qraw, uint32_t b){
    const uint32_t q7 = qraw & 0x7F7F7F7Fu;
    const uint32_t t7 = (q7 + 0x4A4A4A4Au) >> 7;
    const uint32_t s6 = q7 >> 6;
    uint32_t x = (b | (b << 12)) & 0x000F000Fu;
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
}
// level 1 -> shared arrays of ONE Stats object, added or subtracted
__device__ __noinline__ void drain1_signed(uint32_t* vals /*[NSETS*5]*/, uint32_t* sm, int C, int w, int neg){
    const int c0 = 4 * w;
    for(int s = 0; s < NSETS; ++s){
        const uint32_t n = vals[s * 5 + 0], n20 = vals[s * 5 + 1], n30 = vals[s * 5 + 2], qe = vals[s * 5 + 3], qo = vals[s * 5 + 4];
        for(int j = 0; j < 4; ++j){
            uint32_t v[4];
            v[KN] = s ? ((n >> (8 * j)) & 0xFF) : 0u;
            v[KN20] = (n20 >> (8 * j)) & 0xFF;
            v[KN30] = (n30 >> (8 * j)) & 0xFF;
            v[KQS] = (j & 1) ? ((qo >> (8 * (j - 1))) & 0xFFFF) : ((qe >> (8 * j)) & 0xFFFF);
            for(int k = 0; k < 4; ++k) if(v[k]) atomicAdd(sm_cell(sm, C, k, s) + c0 + j, neg ? 0u - v[k] : v[k]);
        }
    }
}
__device__ __forceinline__ void drain_l1_signed(Acc1& a, uint32_t* sm, int C, int w, int neg){
    uint32_t vals[NSETS * 5];
#pragma unroll
    for(int s = 0; s < NSETS; ++s){ vals[s * 5] = a.n[s]; vals[s * 5 + 1] = a.n20[s]; vals[s * 5 + 2] = a.n30[s]; vals[s * 5 + 3] = a.qe[s]; vals[s * 5 + 4] = a.qo[s]; }
    drain1_signed(vals, sm, C, w, neg);
    a.zero();
}

// Quad path without N: 60 accumulators (the N set is filled by the trim lanes).
struct QuadAccNF {
    uint32_t a20[4], a30[4], aqs[4];
    uint32_t n[3][4], n20[3][4], n30[3][4], qs[3][4];
    __device__ __forceinline__ void zero(){
#pragma unroll
        for(int j = 0; j < 4; ++j){
            a20[j] = a30[j] = aqs[j] = 0;
#pragma unroll
            for(int s = 0; s < 3; ++s) n[s][j] = n20[s][j] = n30[s][j] = qs[s][j] = 0;
        }
    }
};
__device__ __forceinline__ void quad_accumulate_nf(QuadAccNF& A, uint32_t q0, uint32_t q1, uint32_t q2, uint32_t q3,
                                                   uint32_t b0, uint32_t b1, uint32_t b2, uint32_t b3){
    const uint32_t t0 = __byte_perm(q0, q1, 0x5140), t1 = __byte_perm(q0, q1, 0x7362);
    const uint32_t t2 = __byte_perm(q2, q3, 0x5140), t3 = __byte_perm(q2, q3, 0x7362);
    uint32_t qT[4];
    qT[0] = __byte_perm(t0, t2, 0x5410); qT[1] = __byte_perm(t0, t2, 0x7632);
    qT[2] = __byte_perm(t1, t3, 0x5410); qT[3] = __byte_perm(t1, t3, 0x7632);
    const uint32_t wc = __byte_perm(__byte_perm(b0, b1, 0x3340), __byte_perm(b2, b3, 0x3340), 0x5410);
    const uint32_t wc1 = wc >> 1;
    const uint32_t ones = 0x01010101u;
#pragma unroll
    for(int j = 0; j < 4; ++j){
        const uint32_t one = ones << (2 * j);
        const uint32_t q7 = qT[j] & 0x7F7F7F7Fu;
        const uint32_t t7 = (q7 + 0x4A4A4A4Au) & 0x80808080u;
        const uint32_t s6 = q7 & 0x40404040u;
        uint32_t m[3];
        m[0] = wc & one; m[1] = wc & (one << 1); m[2] = wc & wc1 & one;
        A.a20[j] = dp4(t7, ones, A.a20[j]);
        A.a30[j] = dp4(s6, ones, A.a30[j]);
        A.aqs[j] = dp4(q7, ones, A.aqs[j]);
#pragma unroll
        for(int s = 0; s < 3; ++s){
            A.n[s][j] = dp4(m[s], ones, A.n[s][j]);
            A.n20[s][j] = dp4(t7, m[s], A.n20[s][j]);
            A.n30[s][j] = dp4(s6, m[s], A.n30[s][j]);
            A.qs[s][j] = dp4(q7, m[s], A.qs[s][j]);
        }
    }
}
__device__ __noinline__ void quad_drain1_nf(const uint32_t* v, uint32_t* smA, uint32_t* smB, int C, int w){
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
    }
}
__device__ __forceinline__ void quad_drain_nf(QuadAccNF& A, uint32_t* smA, uint32_t* smB, int C, int w){
    uint32_t v[60];
#pragma unroll
    for(int j = 0; j < 4; ++j){
        v[j] = A.a20[j]; v[4 + j] = A.a30[j]; v[8 + j] = A.aqs[j];
#pragma unroll
        for(int s = 0; s < 3; ++s){
            v[12 + s * 4 + j] = A.n[s][j]; v[24 + s * 4 + j] = A.n20[s][j]; v[36 + s * 4 + j] = A.n30[s][j]; v[48 + s * 4 + j] = A.qs[s][j];
        }
    }
    quad_drain1_nf(v, smA, smB, C, w);
    A.zero();
}

// The per-read sequence of processSingleEnd (src/seprocessor.cpp:313-340) when only quality cuts, fixed trims and
// the -q / -l filters are enabled (no polyG/polyX/adapter/complexity/average-quality/max-length).  passFilter's
// low-quality count is taken over the whole padded read in one uniform sweep (16 quality bytes per load;
// (byte | 0x80) - limit keeps bit 7 iff the quality is >= limit whatever the N flag; one dp4a per word sums the
// flags); for trimmed reads the bases that were cut off are counted back out.  The same sweep notes which 32-byte
// granules hold an N (nmask), from which passFilter's N count is taken.
struct Fast1 { uint32_t ge; int nn, start, plen; bool alive; };
// first half: the sweep, trimAndCut and the N count of the whole read
__device__ __forceinline__ Fast1 fast_phase1a(const fq_params& P, uint32_t qa, int len){
    const uint32_t nch = ((uint32_t)len + 15u) >> 4;
    uint32_t nmask = 0, ge = 0, ge2 = 0;
    PROF_T(t_sweep);
    if(P.qual_filter_enabled){
        const uint32_t L4 = (uint32_t)P.low_quality_limit * 0x01010101u;
#pragma unroll 2
        for(uint32_t c = 0; c < nch; ++c){
            const uint4 v = lds128(qa + 16u * c);
            const uint32_t fl = ((v.x | v.y) | (v.z | v.w)) & 0x80808080u;
            nmask |= (fl ? 1u : 0u) << (c >> 1);
            ge = dp4(((v.x | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge);
            ge2 = dp4(((v.y | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge2);
            ge = dp4(((v.z | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge);
            ge2 = dp4(((v.w | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge2);
        }
        ge += ge2;
    }else{
        for(uint32_t c = 0; c < nch; ++c){
            const uint4 v = lds128(qa + 16u * c);
            const uint32_t fl = ((v.x | v.y) | (v.z | v.w)) & 0x80808080u;
            nmask |= (fl ? 1u : 0u) << (c >> 1);
        }
    }
    PROF_ADD(9, t_sweep);
    Fast1 f;
    f.ge = ge; f.nn = 0;
    PROF_T(t_cut);
    f.alive = s_trim_and_cut(P, qa, len, P.trim_front1, P.trim_tail1, f.start, f.plen);
    PROF_ADD(10, t_cut);
    PROF_T(t_nb);
    if(nmask && f.alive && P.qual_filter_enabled) f.nn = n_count(qa, len, nmask);
    PROF_ADD(2, t_nb);
    return f;
}
// second half: passFilter's decision.  tail_done: the counts of the cut-off tail (tl, tn) were taken by the warp
__device__ __forceinline__ fq_read_result fast_phase1b(const fq_params& P, uint32_t qa, int len, const Fast1& f,
                                                       bool tail_done, int tl, int tn){
    fq_read_result rec = {};
    const int start = f.start, plen = f.plen;
    int nn = f.nn;
    PROF_T(t_res);
    int result = FQK_PASS_FILTER;
    if(!f.alive || plen == 0) result = FQK_FAIL_LENGTH;
    else{
        if(P.qual_filter_enabled){
            int lowq = len - (int)(f.ge >> 7);             // zero padding never reaches the limit
            if(plen < len){                                // take the bases that were cut off back out
                if(start > 0){ int l1, n1; range_counts(qa, 0, start, P.low_quality_limit, l1, n1); lowq -= l1; nn -= n1; }
                if(tail_done){ lowq -= tl; nn -= tn; }
                else if(start + plen < len){ int l2, n2; range_counts(qa, start + plen, len - start - plen, P.low_quality_limit, l2, n2); lowq -= l2; nn -= n2; }
            }
            if(lowq > P.low_quality_base_limit) result = FQK_FAIL_QUALITY;
            else if(nn > P.n_base_limit) result = FQK_FAIL_N_BASE;
        }
        if(result == FQK_PASS_FILTER && P.length_filter_enabled){
            if(plen < P.min_read_length) result = FQK_FAIL_LENGTH;
            else if(P.max_read_length > 0 && plen > P.max_read_length) result = FQK_FAIL_TOO_LONG;
        }
    }
    PROF_ADD(12, t_res);
    rec.result = (uint8_t)result;
    if(!f.alive){ rec.flags |= FQK_F_DROPPED; rec.start = 0; rec.len = 0; }
    else { rec.start = (uint16_t)start; rec.len = (uint16_t)plen; }
    return rec;
}

// one 4-cycle word of a front-trimmed read's kept part, added to post1 at cycles 4w.. by shared atomics (rare: the
// reference's cut_front / fixed front trim moved the read): the set representation of acc_word, one base at a time
__device__ __noinline__ void z_word_atomic(uint32_t qraw, uint32_t b, int nv, uint32_t* sm, int C, int w){
    for(int j = 0; j < 4 && j < nv; ++j){
        const uint32_t by = (qraw >> (8 * j)) & 0xFFu, q7 = by & 0x7Fu, code = (b >> (2 * j)) & 3u;
        const int c = 4 * w + j;
        uint32_t sets = 1u << S_ALL;
        if(by & 0x80u) sets |= 1u << S_N;
        else{ if(code & 1u) sets |= 1u << S_B0; if(code & 2u) sets |= 1u << S_B1; if(code == 3u) sets |= 1u << S_B01; }
        for(int s = 0; s < NSETS; ++s){
            if(!((sets >> s) & 1u)) continue;
            if(s) atomicAdd(sm_cell(sm, C, KN, s) + c, 1u);
            if(q7 > 53u) atomicAdd(sm_cell(sm, C, KN20, s) + c, 1u);
            if(q7 > 63u) atomicAdd(sm_cell(sm, C, KN30, s) + c, 1u);
            atomicAdd(sm_cell(sm, C, KQS, s) + c, q7);
        }
    }
}

// exception entry (8 bytes): x = (tile offset)/16 | raw length << 17,  y = kept start | kept length << 13 | kind << 26
// kind 1: the read is not in the output (D = whole read); kind 2: passing read, trimmed (D = what was cut off when the
// front is intact, else D = whole read and Z = kept part at its shifted cycle)
// NSTAT: statistics threads (256, or 128 for option sets whose trim role is the long one: 4 statistics warps, 11 trim
// warps + the producer)
template <bool FAST, int NSTAT = V3_STATS>
__global__ void __launch_bounds__(V3_THREADS, 1)
k_se_v3(const __grid_constant__ DevParams D, DevBatch B, fq_read_result* __restrict__ out,
        uint64_t* __restrict__ flat, int W, int T, int reads_per_block, int dbg){
    constexpr int VS = NSTAT, VT = V3_THREADS - NSTAT, VW = VT / 32 - 1;       // statistics threads, trim threads, working trim warps
    extern __shared__ __align__(128) uint8_t smraw[];
    const fq_params& P = D.P;
    const V3L L = v3_layout(W, T);
    const int C = 4 * W, Cp = (C + 15) & ~15;
    uint32_t* smPre = reinterpret_cast<uint32_t*>(smraw + 128);
    uint32_t* smPost = smPre + L.statw;
    BlockCounters* bc = reinterpret_cast<BlockCounters*>(smraw + L.o_bc);
    uint64_t* full = reinterpret_cast<uint64_t*>(smraw + L.o_bar);
    uint64_t* ready = full + V3_NS; uint64_t* freeb = ready + V3_NS;
    uint32_t* blkmax = reinterpret_cast<uint32_t*>(smraw + L.o_zero);

    uint8_t* sp = smraw + L.o_stage;
    const uint32_t sp32 = smem_u32(sp);
    const int CH = T / 32;                                       // chunks per tile (T is a multiple of 32)

    for(int i = threadIdx.x; i < (int)(2 * L.statw); i += V3_THREADS) smPre[i] = 0;
#ifdef FQK_V3_PROF
    if(threadIdx.x < 16) s_v3prof[threadIdx.x] = 0;
#endif
    counters_zero(bc);
    if(threadIdx.x == 0){
        for(int s = 0; s < V3_NS; ++s){
            mbar_init(&full[s], 1); mbar_init(&ready[s], (uint32_t)CH); mbar_init(&freeb[s], VS / 32);
            *reinterpret_cast<uint32_t*>(sp + (size_t)s * L.stage_bytes + L.o_cnt + 12) = 0xFFFFFFFFu;   // tile tag: nothing staged yet
        }
        blkmax[0] = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const uint32_t lo = (uint32_t)min((uint64_t)B.n, (uint64_t)blockIdx.x * reads_per_block);
    const uint32_t hi = (uint32_t)min((uint64_t)B.n, (uint64_t)lo + reads_per_block);
    const int ntiles = (int)((hi - lo + T - 1) / T);
    {   // threads per read: enough 4-cycle words for the longest raw read of this block's share
        uint32_t mx = 0;
        for(uint32_t i = lo + threadIdx.x; i < hi; i += V3_THREADS) mx = max(mx, (uint32_t)B.lengths[i]);
#pragma unroll
        for(int o = 16; o; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if((threadIdx.x & 31) == 0) atomicMax(blkmax, mx);
        __syncthreads();
    }
    const int Wr = min(W, max(8, (int)((blkmax[0] + 3u) >> 2)));
    const int lane = threadIdx.x & 31;
    PROF_T(t_block);

    // role of a thread: the statistics warps take the high warp ids (V3_STATS_HIGH) or the low ones
    const bool is_trim = V3_STATS_HIGH ? (threadIdx.x < VT) : (threadIdx.x >= VS);
    const int rtid = V3_STATS_HIGH ? (is_trim ? threadIdx.x : threadIdx.x - VT) : (is_trim ? threadIdx.x - VS : threadIdx.x);
    if(is_trim){
        asm volatile("setmaxnreg.dec.sync.aligned.u32 88;\n");
        const int tw = rtid >> 5;
        if(tw == VW){
            // ============================== copy producer ==============================
            uint32_t a = 0, last = 0, ll = 0;
            if(ntiles > 0){ const uint32_t nt = min((uint32_t)T, hi - lo); a = B.offsets[lo]; last = B.offsets[lo + nt - 1]; ll = B.lengths[lo + nt - 1]; }
            for(int k = 0; k < ntiles; ++k){
                const int s = k % V3_NS, n = k / V3_NS;
                const uint32_t t0 = lo + (uint32_t)k * T;
                const uint32_t nt = min((uint32_t)T, hi - t0);
                const uint32_t ca = a, clast = last, cll = ll;
                if(k + 1 < ntiles){                                // next tile's end points: in flight during the wait below
                    const uint32_t t1 = t0 + T, nt1 = min((uint32_t)T, hi - t1);
                    a = B.offsets[t1]; last = B.offsets[t1 + nt1 - 1]; ll = B.lengths[t1 + nt1 - 1];
                }
                PROF_T(t_pw);
                if(n >= 1) nb_sync(1 + V3_NS + s, VS + 32);
                PROF_ADD(6, t_pw);
                PROF_T(t_pi);
                uint8_t* S = sp + (size_t)s * L.stage_bytes;
                uint32_t* So = reinterpret_cast<uint32_t*>(S + L.o_off);
                uint16_t* Sl = reinterpret_cast<uint16_t*>(S + L.o_len);
                uint32_t* cnt = reinterpret_cast<uint32_t*>(S + L.o_cnt);
                const uint32_t nb = nt & ~7u;                      // bulk copies move multiples of 16 bytes
                for(uint32_t i = nb + lane; i < nt; i += 32){ So[i] = B.offsets[t0 + i]; Sl[i] = B.lengths[t0 + i]; }
                for(uint32_t i = nt + lane; i < (uint32_t)T; i += 32){ So[i] = ca; Sl[i] = 0; }
                __syncwarp();
                if(lane == 0){
                    uint64_t e = (uint64_t)clast + ((cll + 15u) & ~15u);
                    const bool bad = (ca & 15u) || e < ca || e - ca > L.qcap || e > B.total;
                    if(bad){ atomicOr(D.err, 2u); e = ca; }
                    const uint32_t qbytes = (uint32_t)(e - ca);
                    const uint32_t b0 = (ca >> 2) & ~15u;
                    const uint32_t bbytes = qbytes ? (((uint32_t)(e >> 2) + 15u) & ~15u) - b0 : 0u;
                    cnt[0] = 0;
                    cnt[1] = (nt == (uint32_t)T && qbytes == L.qcap) ? 1u : 0u;      // every read exactly Cp padded bytes
                    cnt[2] = bad ? 0u : nt;
                    *reinterpret_cast<volatile uint32_t*>(&cnt[3]) = (uint32_t)k;     // tile tag (wait_tile_tag, fqk_smem.cuh)
                    mbar_expect_tx(&full[s], qbytes + bbytes + nb * 6u);
                    if(qbytes){
                        bulk_g2s(S, B.qual + ca, qbytes, &full[s]);
                        bulk_g2s(S + L.o_b, B.bases2 + b0, bbytes, &full[s]);
                    }
                    if(nb){
                        bulk_g2s(So, B.offsets + t0, nb * 4u, &full[s]);
                        bulk_g2s(Sl, B.lengths + t0, nb * 2u, &full[s]);
                    }
                }
                PROF_ADD(7, t_pi);
            }
        }else{
            // ============================== trim warps ==============================
            Tally tl = {};
            const bool fast_filter = !P.complexity_enabled;
            const uint32_t pre32 = smem_u32(smPre), post32 = smem_u32(smPost);
            const uint32_t lh = 4u * NSETS * (uint32_t)C * 4u;         // byte offset of the length histogram
            const int nchunks = ntiles * CH;
            for(int c = tw; c < nchunks; c += VW){
                const int k = c / CH, j = c - k * CH, s = k % V3_NS;
                uint8_t* S = sp + (size_t)s * L.stage_bytes;
                const uint32_t S32 = sp32 + (uint32_t)s * L.stage_bytes;
                PROF_T(t_tw);
                wait_tile_tag(S32 + L.o_cnt + 12, (uint32_t)k);
                mbar_wait_sleep(&full[s], (uint32_t)(k / V3_NS) & 1u, 20000);
                PROF_ADD(0, t_tw);
                PROF_T(t_tc);
                const uint32_t* So = reinterpret_cast<const uint32_t*>(S + L.o_off);
                const uint16_t* Sl = reinterpret_cast<const uint16_t*>(S + L.o_len);
                const int nt = (int)lds32(S32 + L.o_cnt + 8);
                const int i = 32 * j + lane;
                const bool valid = i < nt;
                if(32 * j < nt && !(dbg & 1)){
                    const uint32_t a0 = So[0];
                    const uint32_t bshift = (a0 >> 2) & 15u;
                    const uint32_t t0 = lo + (uint32_t)k * T;
                    uint32_t doff = 0, dlen = 0;
                    fq_read_result rec = {};
                    Fast1 f1 = {0u, 0, 0, 0, false};
                    bool pass = false;
                    if(valid){
                        doff = So[i] - a0; dlen = Sl[i];
                        const uint32_t padded = (dlen + 15u) & ~15u;
                        bool ok = !(So[i] & 15u) && So[i] >= a0 && doff + padded <= L.qcap && dlen <= (uint32_t)C;
                        if(ok && i + 1 < nt) ok = (So[i + 1] == So[i] + padded);        // dense, ordered batch
                        if(!ok){
                            atomicOr(D.err, dlen > (uint32_t)C ? 1u : 2u); doff = 0; dlen = 0;
                            sts16(S32 + L.o_len + 2u * i, 0u);                          // the statistics warps skip it
                        }
                        const uint32_t qa = S32 + doff;
                        if(FAST) f1 = fast_phase1a(P, qa, (int)dlen);
                        else rec = phase1_read(D, S + doff, S + L.o_b + bshift + (doff >> 2), (int)dlen, fast_filter, tl);
                    }
                    bool coop = false;
                    if(FAST){
                        // Short cut-off tails of front-intact reads (the usual trimmed read) are handled by the whole warp,
                        // one lane per cut-off base: first passFilter's counts of the tail (two ballots instead of a
                        // per-lane range scan), then, for reads that pass, D = raw positions [plen, len) is taken out of
                        // post1 by shared reductions right here, so that the statistics warps never see these reads as
                        // exceptions.
                        coop = valid && f1.alive && f1.start == 0 && f1.plen > 0 && (uint32_t)f1.plen < dlen && dlen - (uint32_t)f1.plen <= 32u;
                        const unsigned mc0 = __ballot_sync(0xffffffffu, coop);
                        int tlow = 0, tn = 0;
                        const uint32_t lim = (uint32_t)P.low_quality_limit;
                        for(unsigned mc = mc0; mc; mc &= mc - 1){
                            const int src = __ffs((int)mc) - 1;
                            const uint32_t o = __shfl_sync(0xffffffffu, doff, src), pl = (uint32_t)__shfl_sync(0xffffffffu, f1.plen, src),
                                           ln = __shfl_sync(0xffffffffu, dlen, src);
                            const uint32_t pos = pl + (uint32_t)lane;
                            const uint32_t by = pos < ln ? lds8(S32 + o + pos) : 0u;
                            const unsigned bl = __ballot_sync(0xffffffffu, pos < ln && (by & 0x7Fu) < lim), bn = __ballot_sync(0xffffffffu, (by & 0x80u) != 0);
                            if(lane == src){ tlow = __popc(bl); tn = __popc(bn); }
                        }
                        if(valid) rec = fast_phase1b(P, S32 + doff, (int)dlen, f1, coop, tlow, tn);
                    }
                    if(valid){
                        out[t0 + i] = rec;
                        pass = !(rec.flags & FQK_F_DROPPED) && rec.result == FQK_PASS_FILTER;
                    }
                    if(FAST){
                        coop = coop && pass && !(dbg & 16);                 // (FQK_V3_DEBUG 16: every cut-off tail through the exception list)
                        const uint32_t sbase = S32 + L.o_b + bshift;
                        const uint32_t row = 4u * (uint32_t)C, krow = (uint32_t)NSETS * row;
                        for(unsigned mc = __ballot_sync(0xffffffffu, coop); mc; mc &= mc - 1){
                            const int src = __ffs((int)mc) - 1;
                            const uint32_t o = __shfl_sync(0xffffffffu, doff, src), pl = (uint32_t)__shfl_sync(0xffffffffu, (int)rec.len, src),
                                           ln = __shfl_sync(0xffffffffu, dlen, src);
                            const uint32_t pos = pl + (uint32_t)lane;
                            if(pos < ln){
                                const uint32_t by = lds8(S32 + o + pos), q7 = by & 0x7Fu;
                                const uint32_t code = (lds8(sbase + ((o + pos) >> 2)) >> (2u * (pos & 3u))) & 3u;
                                const bool isn = (by & 0x80u) != 0;
                                const uint32_t cell = post32 + 4u * pos;
                                const uint32_t m20 = q7 > 53u ? 0xFFFFFFFFu : 0u, m30 = q7 > 63u ? 0xFFFFFFFFu : 0u;
#pragma unroll
                                for(int st = 0; st < NSETS; ++st){
                                    const bool in = st == S_ALL ? true : st == S_N ? isn : (!isn && (st == S_B0 ? (code & 1u) != 0 : st == S_B1 ? (code & 2u) != 0 : code == 3u));
                                    if(!in) continue;
                                    const uint32_t a = cell + (uint32_t)st * row;
                                    if(st) reds_add(a + KN * krow, 0xFFFFFFFFu);
                                    if(m20) reds_add(a + KN20 * krow, 0xFFFFFFFFu);
                                    if(m30) reds_add(a + KN30 * krow, 0xFFFFFFFFu);
                                    reds_add(a + KQS * krow, 0u - q7);
                                }
                            }
                        }
                    }
                    PROF_T(t_tt);
                    // run counters, aggregated over the warp where every lane agrees (the usual case)
                    const unsigned mv = __ballot_sync(0xffffffffu, valid), mp = __ballot_sync(0xffffffffu, pass);
                    if(lane == 0 && mp) atomicAdd(&bc->filter[FQK_PASS_FILTER], (uint32_t)__popc(mp));
                    if(valid && !pass) atomicAdd(&bc->filter[rec.result], 1u);
                    const uint32_t dl0 = __shfl_sync(0xffffffffu, dlen, 0);
                    if(__all_sync(0xffffffffu, !valid || dlen == dl0)){ if(lane == 0) reds_add(pre32 + lh + 4u * dl0, (uint32_t)__popc(mv)); }
                    else if(valid) reds_add(pre32 + lh + 4u * dlen, 1u);
                    if(__all_sync(0xffffffffu, !pass || rec.len == dl0)){ if(lane == 0 && mp) reds_add(post32 + lh + 4u * dl0, (uint32_t)__popc(mp)); }
                    else if(pass) reds_add(post32 + lh + 4u * rec.len, 1u);
                    if(valid && !coop && !(pass && rec.start == 0 && rec.len == dlen)){
                        const Desc d = make_desc(doff, dlen, rec.start, rec.len, pass ? 2u : 1u);
                        const uint32_t idx = atomicAdd(reinterpret_cast<uint32_t*>(S + L.o_cnt), 1u);
                        reinterpret_cast<uint2*>(S + L.o_ex)[idx] = make_uint2(d.x, d.y);
                    }
                    PROF_ADD(11, t_tt);
                }
                __syncwarp();
                PROF_ADD(1, t_tc);
                __threadfence_block();
                nb_arrive(1 + s, CH * 32 + VS);
            }
            counters_add_tally(bc, tl);
        }
    }else{
        // ============================ statistics warps ============================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 168;\n");
        const int G = VS / Wr;
        const int g = rtid / Wr, w = rtid - g * Wr;
        const bool active = g < G;
        QuadAcc qa;
        Acc1 dsub;
        qa.zero(); dsub.zero();
        int usedQ = 0, leftO = 255;
        const int quads_per_tile = T / 4;
        for(int k = 0; k < ntiles; ++k){
            const int s = k % V3_NS;
            const uint32_t S32 = sp32 + (uint32_t)s * L.stage_bytes;
            PROF_T(t_sw);
            nb_sync(1 + s, CH * 32 + VS);
            PROF_ADD(3, t_sw);
            PROF_T(t_sp);
            if(active && !(dbg & 2)){
                const uint32_t a0 = lds32(S32 + L.o_off);
                const uint32_t bshift = (a0 >> 2) & 15u;
                const uint32_t sq0 = S32, sb0 = S32 + L.o_b + bshift;
                const uint32_t sqw = sq0 + 4 * w, sbw = sb0 + w;
                const int nex = (int)lds32(S32 + L.o_cnt), uniform = (int)lds32(S32 + L.o_cnt + 4), nt = (int)lds32(S32 + L.o_cnt + 8);
                const int mine = (quads_per_tile + G - 1) / G;
                if(usedQ + mine > QUAD_LIMIT){ quad_drain(qa, smPre, smPost, C, w); usedQ = 0; }
                usedQ += mine;
                uint32_t q[4], b[4];
                if(uniform){
                    // every read of the tile occupies exactly Cp bytes: addresses advance by constants, no range checks
                    const uint32_t qstep = 4u * (uint32_t)G * (uint32_t)Cp, bstep = (uint32_t)G * (uint32_t)Cp;
                    uint32_t aq = sqw + 4u * (uint32_t)g * (uint32_t)Cp, ab = sbw + (uint32_t)g * (uint32_t)Cp;
                    const uint32_t c1 = (uint32_t)Cp, c1b = (uint32_t)Cp >> 2;
                    // two register sets, loads of a quad issued one accumulate ahead (no register copies)
                    uint32_t q2[4], b2[4];
                    auto fetchA = [&](){
                        q[0] = lds32(aq); q[1] = lds32(aq + c1); q[2] = lds32(aq + 2 * c1); q[3] = lds32(aq + 3 * c1);
                        b[0] = lds8(ab); b[1] = lds8(ab + c1b); b[2] = lds8(ab + 2 * c1b); b[3] = lds8(ab + 3 * c1b);
                        aq += qstep; ab += bstep;
                    };
                    auto fetchB = [&](){
                        q2[0] = lds32(aq); q2[1] = lds32(aq + c1); q2[2] = lds32(aq + 2 * c1); q2[3] = lds32(aq + 3 * c1);
                        b2[0] = lds8(ab); b2[1] = lds8(ab + c1b); b2[2] = lds8(ab + 2 * c1b); b2[3] = lds8(ab + 3 * c1b);
                        aq += qstep; ab += bstep;
                    };
                    if(g < quads_per_tile) fetchA();
                    if(g + G < quads_per_tile) fetchB();
                    for(int qd = g; qd < quads_per_tile; qd += 2 * G){
                        quad_accumulate(qa, q[0], q[1], q[2], q[3], b[0], b[1], b[2], b[3]);
                        if(qd + 2 * G < quads_per_tile) fetchA();
                        if(qd + G < quads_per_tile){
                            quad_accumulate(qa, q2[0], q2[1], q2[2], q2[3], b2[0], b2[1], b2[2], b2[3]);
                            if(qd + 3 * G < quads_per_tile) fetchB();
                        }
                    }
                }else{
                    const int nq = (nt + 3) >> 2;
                    for(int qd = g; qd < nq; qd += G){
                        const uint4 o = lds128(S32 + L.o_off + 16u * qd);
                        const uint2 l2 = lds64(S32 + L.o_len + 8u * qd);
                        const uint32_t of[4] = {o.x - a0, o.y - a0, o.z - a0, o.w - a0};
                        const uint32_t ln[4] = {l2.x & 0xFFFFu, l2.x >> 16, l2.y & 0xFFFFu, l2.y >> 16};
#pragma unroll
                        for(int r = 0; r < 4; ++r){
                            const bool in = (uint32_t)(4 * w) < ln[r];
                            q[r] = in ? lds32(sqw + of[r]) : 0u;
                            b[r] = in ? lds8(sbw + (of[r] >> 2)) : 0u;
                        }
                        quad_accumulate(qa, q[0], q[1], q[2], q[3], b[0], b[1], b[2], b[3]);
                    }
                }
                PROF_ADD(4, t_sp);
                // ---- exceptions: D is subtracted from post1, Z is added to it ----
                const uint32_t ex = S32 + L.o_ex;
                const uint32_t c0 = 4u * (uint32_t)w;
                // descriptor and word of the next entry are loaded before the current one is reduced
                auto exfetch = [&](int i, uint2& d, uint32_t& qq, uint32_t& bb){
                    d = lds64(ex + 8 * i);
                    const uint32_t off = (d.x & 0x1FFFFu) << 4, len = d.x >> 17;
                    const uint32_t d0 = (((d.y >> 26) & 3u) == 2u && (d.y & 0x1FFFu) == 0u) ? ((d.y >> 13) & 0x1FFFu) : 0u;   // D = raw positions [d0, len)
                    const bool in = c0 < len && c0 + 4 > d0;
                    qq = in ? lds32(sqw + off) : 0u;
                    bb = in ? lds8(sbw + (off >> 2)) : 0u;
                    if(c0 < d0){ const uint32_t nb = min(d0 - c0, 4u); qq = nb >= 4 ? 0u : (qq & (0xFFFFFFFFu << (8 * nb))); bb &= 0xFFu << (2 * nb); }
                };
                uint2 dn = make_uint2(0u, 0u); uint32_t qn = 0, bn = 0;
                if(g < nex) exfetch(g, dn, qn, bn);
                for(int i = g; i < nex; i += G){
                    const uint2 dc = dn; const uint32_t qc = qn, bcur = bn;
                    if(i + G < nex) exfetch(i + G, dn, qn, bn);
                    const uint32_t off = (dc.x & 0x1FFFFu) << 4, kind = (dc.y >> 26) & 3u;
                    const uint32_t start = dc.y & 0x1FFFu, plen = (dc.y >> 13) & 0x1FFFu;
                    if(qc) acc_word(dsub, qc, bcur);                 // a word inside a read never has four zero quality bytes
                    if(kind == 2 && start > 0){
                        const int nv = (int)plen - 4 * w;
                        if(nv > 0){
                            const uint32_t p = off + start + 4 * w;
                            const uint32_t a1 = p & ~3u, sh = p & 3u;
                            const uint32_t w0 = lds32(sq0 + a1);
                            const uint32_t w1 = sh ? lds32(sq0 + a1 + 4) : 0u;
                            uint32_t q2 = __funnelshift_r(w0, w1, sh * 8);
                            const uint32_t e0 = lds8(sb0 + (p >> 2)), e1 = sh ? lds8(sb0 + (p >> 2) + 1) : 0u;
                            uint32_t b2 = ((e0 | (e1 << 8)) >> (2 * sh)) & 0xFF;
                            z_word_atomic(q2, b2, nv, smPost, C, w);
                        }
                    }
                    if(--leftO == 0){ drain_l1_signed(dsub, smPost, C, w, 1); leftO = 255; }
                }
            }
            __syncwarp();
            PROF_ADD(5, t_sp);
            if(k + V3_NS < ntiles) nb_arrive(1 + V3_NS + s, VS + 32);      // the producer waits only for stages it refills
        }
        if(active){
            quad_drain(qa, smPre, smPost, C, w);
            drain_l1_signed(dsub, smPost, C, w, 1);
        }
    }
    __syncthreads();
#ifdef FQK_V3_PROF
    if(threadIdx.x == 0) atomicAdd(&s_v3prof[8], (unsigned long long)(clock64() - t_block));
    __syncthreads();
    if(threadIdx.x < 16) atomicAdd(&g_v3prof[threadIdx.x], s_v3prof[threadIdx.x]);
#endif
    retire(smPre, C, D.L, flat + D.L.stats_off[0]);
    retire(smPost, C, D.L, flat + D.L.stats_off[2]);
    counters_flush(bc, D.L, flat);
}
