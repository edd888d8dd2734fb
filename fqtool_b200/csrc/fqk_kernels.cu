// fqk_kernels.cu -- generic (any length, any option) kernels of the hot path.
//   k_trim_se / k_trim_pe : one thread per read / read pair, records + run counters
//   k_stats_se / k_stats_pe: one warp per read, shared-memory privatised per-cycle
//                            histograms (global atomics when they do not fit)
//   k_dup                 : duplication table, order-independent (min, count, first) monoid
// The specialised fused SE kernel lives in fqk_fused.cu.
#include "fqk_trim.cuh"
#include "fqk_pe.cuh"
#include "fqk_kernels.h"

__global__ void __launch_bounds__(256) k_trim_se(const __grid_constant__ DevParams D, DevBatch B,
                                                  fq_read_result* __restrict__ out, uint64_t* __restrict__ flat){
    __shared__ BlockCounters bc;
    counters_zero(&bc);
    __syncthreads();
    Tally tl = {};
    for(uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < B.n; r += gridDim.x * blockDim.x){
        ReadView v = make_view(B, r);
        fq_read_result rec = {};
        se_read(D, v, rec, tl);
        out[r] = rec;
        atomicAdd(&bc.filter[rec.result], 1u);          // addFilterResult, mPaired == false
    }
    counters_add_tally(&bc, tl);
    __syncthreads();
    counters_flush(&bc, D.L, flat);
}

// ------------------------------------------------------------------ PE trim
__device__ inline int insert_size_of(const fq_params& P, const ReadView& a, const ReadView& b, const OvRes& ov){
    int isize = P.insert_size_max;                       // statInsertSize, src/peprocessor.cpp:510-523
    if(ov.overlapped) isize = ov.offset > 0 ? a.len + b.len - ov.len : ov.len;
    return min(isize, P.insert_size_max);
}

// base class / quality of base i of the merged read, src/overlapanalysis.cpp:77-93
struct MergedGet {
    const ReadView& a; const ReadView& b; int len1, ol;
    __device__ __forceinline__ void operator()(int i, int& c, int& q) const {
        if(i < len1){ c = a.cls(i); q = a.qual(i); }
        else { const int j = ol + i - len1; c = rc_cls(b, j); q = b.qual(b.len - 1 - j); }
    }
};

// in-place edits of the packed batch (a thread owns its pair; a base byte never spans two reads)
__device__ __forceinline__ void write_base_qual(const ReadView& v, int pos, int cls, int q){
    uint8_t* qp = const_cast<uint8_t*>(v.q) + v.start + pos;
    *qp = (uint8_t)((q & 0x7F) | (cls == CLS_N ? 0x80 : 0));
    const int p = v.start + pos;
    uint8_t* bp = const_cast<uint8_t*>(v.b2) + (p >> 2);
    const int sh = (p & 3) * 2;
    *bp = (uint8_t)((*bp & ~(3 << sh)) | ((cls == CLS_N ? 0 : cls) << sh));
}

// BaseCorrector::correctByOverlapAnalysis, src/basecorrector.cpp:14-70: inside the overlap a mismatching base
// with quality <= 14 is overwritten by the complement of its partner when that one has quality >= 30.
// The batch is corrected in place (the reference mutates its Reads); every edit is appended to `corr` for the host.
// addCorrection(from, to) runs after the overwrite through the same buffer (:40-45, :47-52), so from == to.
__device__ inline bool dev_correct(const DevParams& D, const ReadView& a, const ReadView& b, const OvRes& ov, uint32_t pair,
                                   fq_correction* corr, uint32_t* ncorr, uint32_t cap, uint64_t* flat){
    if(ov.diff == 0 || ov.diff > 5) return false;
    const int start1 = max(0, ov.offset), start2 = b.len - max(0, -ov.offset) - 1;
    int corrected = 0; bool r1c = false, r2c = false;
    for(int i = 0; i < ov.len; ++i){
        const int p1 = start1 + i, p2 = start2 - i;
        const int c1 = a.cls(p1), c2 = b.cls(p2);
        const int k1 = c1 == CLS_N ? CLS_N : (c1 ^ 1), k2 = c2 == CLS_N ? CLS_N : (c2 ^ 1);   // complements
        if(c1 == k2) continue;
        const int q1 = a.qual(p1), q2 = b.qual(p2);
        int mate, pos, ncls, nq;
        if(q1 >= 63 && q2 <= 47){ mate = 1; pos = p2; ncls = k1; nq = q1; r2c = true; write_base_qual(b, p2, ncls, nq); }
        else if(q2 >= 63 && q1 <= 47){ mate = 0; pos = p1; ncls = k2; nq = q2; r1c = true; write_base_qual(a, p1, ncls, nq); }
        else continue;
        ++corrected;
        const int bin = cls_to_bin(ncls);
        add_u64(flat + D.L.correction_off + bin * 8 + bin, 1);
        const uint32_t slot = atomicAdd(ncorr, 1u);
        if(slot < cap){
            fq_correction e; e.pair = pair; e.pos = (uint16_t)((mate ? b.start : a.start) + pos);
            e.mate_base = (uint8_t)((mate << 7) | "ATCGN"[ncls]); e.qual = (uint8_t)nq;
            corr[slot] = e;
        }
    }
    if(corrected) add_u64(flat + D.L.corrected_reads_off, (r1c && r2c) ? 2 : 1);
    return corrected > 0;
}

// OverlapAnalysis::analyze: bit-parallel on packed words when both reads fit, else the scalar closed form
__device__ inline OvRes analyze_pair(const ReadView& a, const ReadView& b, int limit, int require){
    if(a.start + a.len > 16 * PE_MAXW || b.start + b.len > 16 * PE_MAXW) return dev_analyze(a, b, limit, require);
    PackedSeq pa, pb, pr;
    pack_view(a, pa); pack_view(b, pb); revcomp_packed(pb, pr);
    return dev_analyze_packed(pa, pr, limit, require);
}

__global__ void __launch_bounds__(128, 8) k_trim_pe(const __grid_constant__ DevParams D, DevBatch B1, DevBatch B2,
                                                  fq_read_result* __restrict__ out1, fq_read_result* __restrict__ out2,
                                                  fq_pair_result* __restrict__ outp, uint64_t* __restrict__ flat,
                                                  uint32_t* __restrict__ maxlen2, fq_correction* __restrict__ corr,
                                                  uint32_t* __restrict__ ncorr, uint32_t corr_cap){
    __shared__ BlockCounters bc;
    __shared__ uint32_t smax[2];
    counters_zero(&bc);
    if(threadIdx.x < 2) smax[threadIdx.x] = 0;
    __syncthreads();
    const fq_params& P = D.P;
    Tally tl = {};
    uint32_t m1 = 0, m2 = 0;
    for(uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < B1.n; r += gridDim.x * blockDim.x){
        ReadView a = make_view(B1, r), b = make_view(B2, r);
        m1 = max(m1, (uint32_t)a.len); m2 = max(m2, (uint32_t)b.len);
        fq_read_result rec1 = {}, rec2 = {};
        fq_pair_result pr = {};
        pr.insert_size = 0xFFFF;
        // loop body of PairEndProcessor::processPairEnd, src/peprocessor.cpp:292-429 (threadId 0)
        const bool a1 = dev_trim_and_cut(P, a, P.trim_front1, P.trim_tail1);
        const bool a2 = dev_trim_and_cut(P, b, P.trim_front2, P.trim_tail2);
        const bool both = a1 && a2;
        if(both && P.polyg_enabled){
            // src/peprocessor.cpp:297 hands (maxMismatch, oneEach, minLen) to (compareReq, maxMismatch, oneEach)
            dev_trim_polyg(a, P.polyg_max_mismatch, P.polyg_one_mismatch_each, P.polyg_min_len, rec1, tl);
            dev_trim_polyg(b, P.polyg_max_mismatch, P.polyg_one_mismatch_each, P.polyg_min_len, rec2, tl);
        }
        bool have_ov = false; OvRes ov = {0, 0, 0, 0};
        if(both){
            // first analyze(): :303 when -a/-c, else :330 (insert size only); identical inputs either way
            ov = analyze_pair(a, b, P.overlap_diff_limit, P.overlap_require);
            have_ov = true;
            pr.flags |= FQK_P_OV1_DONE | (ov.overlapped ? FQK_P_OV1_OVERLAP : 0);
            pr.ov1_offset = (int16_t)ov.offset; pr.ov1_len = (uint16_t)ov.len; pr.ov1_diff = (uint16_t)ov.diff;
            const int isize = insert_size_of(P, a, b, ov);
            pr.insert_size = (uint16_t)isize;
            add_u64(flat + D.L.insert_hist_off + isize, 1);
        }
        bool modified = false;                            // reads changed since the first analyze()
        if(both && P.correction_enabled && corr)          // :310-312, before adapter trimming
            modified |= dev_correct(D, a, b, ov, r, corr, ncorr, corr_cap, flat);
        if(both && P.adapter_trim_enabled){
            // AdapterTrimmer::trimByOverlapAnalysis, src/adaptertrimmer.cpp:14-27
            if(ov.diff <= 5 && ov.overlapped && ov.offset < 0 && ov.len > a.len / 3){
                const int t1 = a.len - ov.len, t2 = b.len - ov.len;
                a.len = ov.len; b.len = ov.len;
                rec1.flags |= FQK_F_ADAPTER_OV; rec1.adapter_pos = (int16_t)ov.len; rec1.adapter_len = (uint16_t)t1;
                rec2.flags |= FQK_F_ADAPTER_OV; rec2.adapter_pos = (int16_t)ov.len; rec2.adapter_len = (uint16_t)t2;
                tl.adapter_reads += 2; tl.adapter_bases += (uint32_t)(t1 + t2);     // filterresult.cpp:159-161
                modified = true;
            }else{
                if(P.adapter_r1_len > 0) modified |= trim_by_sequence_any(D, a, 0, rec1, tl);
                if(P.adapter_r2_len > 0) modified |= trim_by_sequence_any(D, b, 1, rec2, tl);
            }
        }
        if(both && P.polyx_enabled){
            const int l1 = a.len, l2 = b.len;
            dev_trim_polyx(a, P.polyx_base_mask, P.polyx_min_len, P.polyx_max_mismatch, P.polyx_one_mismatch_each, rec1, tl);
            dev_trim_polyx(b, P.polyx_base_mask, P.polyx_min_len, P.polyx_max_mismatch, P.polyx_one_mismatch_each, rec2, tl);
            modified |= (a.len != l1) | (b.len != l2);
        }
        if(both){
            if(P.trim_max_len1 > 0 && P.trim_max_len1 < a.len){ view_resize(a, P.trim_max_len1); modified = true; }
            if(P.trim_max_len2 > 0 && P.trim_max_len2 < b.len){ view_resize(b, P.trim_max_len2); modified = true; }
        }
        bool merge_processed = false; int res1 = 0, res2 = 0;
        if(P.merge_enabled && both){
            // second analyze(), :354: same answer as the first when neither read changed
            OvRes ov2 = (have_ov && !modified) ? ov : analyze_pair(a, b, P.overlap_diff_limit, P.overlap_require);
            pr.flags |= FQK_P_OV2_DONE | (ov2.overlapped ? FQK_P_OV2_OVERLAP : 0);
            pr.ov2_offset = (int16_t)ov2.offset; pr.ov2_len = (uint16_t)ov2.len; pr.ov2_diff = (uint16_t)ov2.diff;
            if(ov2.overlapped){
                int result;
                if(ov2.len == 0){ result = FQK_FAIL_LENGTH; }          // merge() == NULL
                else{
                    const int len1 = ov2.len + max(0, ov2.offset);
                    const int len2 = ov2.offset > 0 ? b.len - ov2.len : 0;
                    pr.merge_len1 = (uint16_t)len1; pr.merge_len2 = (uint16_t)len2;
                    if(P.complexity_enabled){
                        MergedGet mg{a, b, len1, ov2.len};
                        result = dev_pass_filter(P, len1 + len2, false, mg);
                    }else{
                        // counts of the merged read = read1 prefix + the part of read2 that is appended (order is irrelevant)
                        int l1 = 0, n1 = 0, t1 = 0, l2 = 0, n2 = 0, t2 = 0;
                        if(P.qual_filter_enabled){
                            swar_counts_global(a.q, a.start, len1, P.low_quality_limit, l1, n1, t1);
                            swar_counts_global(b.q, b.start + (b.len - ov2.len - len2), len2, P.low_quality_limit, l2, n2, t2);
                        }
                        result = (len1 + len2 == 0) ? FQK_FAIL_LENGTH : filter_from_counts(P, len1 + len2, l1 + l2, n1 + n2, t1 + t2);
                    }
                    if(result == FQK_PASS_FILTER){ pr.flags |= FQK_P_MERGED; atomicAdd(&bc.merged_pairs, 1u); }
                }
                atomicAdd(&bc.filter[result], 2u);                      // :358
                pr.merged_result = (uint8_t)result;
                res1 = res2 = result;
                merge_processed = true;
            }else if(!P.merge_discard_unmerged){
                res1 = pass_filter_fast(P, a, false); atomicAdd(&bc.filter[res1], 1u);
                res2 = pass_filter_fast(P, b, false); atomicAdd(&bc.filter[res2], 1u);
                merge_processed = true;
            }
            if(merge_processed) pr.flags |= FQK_P_MERGE_PROC;
        }
        if(!merge_processed){
            res1 = pass_filter_fast(P, a, !a1);
            res2 = pass_filter_fast(P, b, !a2);
            atomicAdd(&bc.filter[max(res1, res2)], 2u);                 // :390, mPaired == true
        }
        rec1.result = (uint8_t)res1; rec2.result = (uint8_t)res2;
        if(!a1) rec1.flags |= FQK_F_DROPPED; else { rec1.start = (uint16_t)a.start; rec1.len = (uint16_t)a.len; }
        if(!a2) rec2.flags |= FQK_F_DROPPED; else { rec2.start = (uint16_t)b.start; rec2.len = (uint16_t)b.len; }
        out1[r] = rec1; out2[r] = rec2; outp[r] = pr;
    }
    counters_add_tally(&bc, tl);
    atomicMax(&smax[0], m1); atomicMax(&smax[1], m2);
    __syncthreads();
    counters_flush(&bc, D.L, flat);
    if(threadIdx.x < 2 && maxlen2) atomicMax(&maxlen2[threadIdx.x], smax[threadIdx.x]);   // longest raw read of each mate
}

// -------------------------------------------------------------------- stats
// Stats::statRead, src/stats.cpp:237-295.  A warp owns a read; lane j handles
// cycles j, j+32, ...  Per block the four per-cycle arrays of every Stats
// object are privatised in shared memory as u32 [kind][class][cycle] (kinds:
// content, q20, q30, qualsum; classes A T C G N) and flushed once with 64-bit
// global atomics, where total_base/total_qual are formed as sums over classes.
struct StatsAcc {
    uint32_t* sm;        // shared-memory privatised arrays or nullptr
    uint64_t* flat;      // global accumulators
    int Cs;              // cycles held in shared memory
    const fq_stats_layout* L;
    uint32_t* err;
    int se;              // SE: shared slots are {pre1, post1} = Stats objects {0, 2}
    unsigned long long* cnt;   // shared [4][2]: reads, length_sum per Stats object
};

__device__ __forceinline__ void stat_base(const StatsAcc& A, int obj, int i, int cls, int q){
    const int qv = q - 33;
    if(A.sm){
        uint32_t* s = A.sm + (size_t)(A.se ? obj >> 1 : obj) * 20 * A.Cs + i;
        atomicAdd(s + (0 * 5 + cls) * A.Cs, 1u);
        if(q > '5') atomicAdd(s + (1 * 5 + cls) * A.Cs, 1u);
        if(q > '?') atomicAdd(s + (2 * 5 + cls) * A.Cs, 1u);
        atomicAdd(s + (3 * 5 + cls) * A.Cs, (uint32_t)qv);
    }else{
        const fq_stats_layout& L = *A.L;
        uint64_t* S = A.flat + L.stats_off[obj];
        const uint64_t C = L.max_cycles;
        const int b = cls_to_bin(cls);
        add_u64(S + L.s_content + b * C + i, 1);
        if(q > '5') add_u64(S + L.s_q20 + b * C + i, 1);
        if(q > '?') add_u64(S + L.s_q30 + b * C + i, 1);
        add_u64(S + L.s_qualsum + b * C + i, (uint64_t)qv);
        add_u64(S + L.s_total_base + i, 1);
        add_u64(S + L.s_total_qual + i, (uint64_t)qv);
    }
}

// one warp accumulates one read (any accessor); k-mers via the rolling 2-bit
// code of Evaluator::seq2int (src/evaluator.cpp:3-49): counted iff all k bases are A/T/C/G
template <class Get>
__device__ inline void stat_stream(const StatsAcc& A, int obj, int len, int k, Get get){
    const int lane = threadIdx.x & 31;
    const fq_stats_layout& L = *A.L;
    if(len > (int)L.max_cycles){ if(lane == 0) atomicOr(A.err, 1u); return; }
    for(int i = lane; i < len; i += 32){
        int c, q; get(i, c, q);
        stat_base(A, obj, i, c, q);
        if(k && i >= k - 1){
            uint32_t code = 0; bool ok = true;
            for(int j = i - k + 1; j <= i; ++j){
                int cj, qj; get(j, cj, qj);
                ok &= (cj != CLS_N);
                code = (code << 2) | (uint32_t)(cj & 3);
            }
            if(ok) add_u64(A.flat + L.stats_off[obj] + L.s_kmer + code, 1);
        }
    }
    if(lane == 0){
        atomicAdd(A.cnt + obj * 2, 1ull);
        atomicAdd(A.cnt + obj * 2 + 1, (unsigned long long)len);
    }
}

__device__ inline void stats_cnt_flush(const StatsAcc& A){
    const fq_stats_layout& L = *A.L;
    if(threadIdx.x < 4 && A.cnt[threadIdx.x * 2]){
        add_u64(A.flat + L.stats_off[threadIdx.x] + L.s_reads, A.cnt[threadIdx.x * 2]);
        add_u64(A.flat + L.stats_off[threadIdx.x] + L.s_length_sum, A.cnt[threadIdx.x * 2 + 1]);
    }
}

__device__ inline void stats_smem_zero(uint32_t* sm, int words){
    for(int i = threadIdx.x; i < words; i += blockDim.x) sm[i] = 0;
}

__device__ inline void stats_smem_flush(const StatsAcc& A, int nobj, const int* objmap){
    const fq_stats_layout& L = *A.L;
    const uint64_t C = L.max_cycles;
    for(int o = 0; o < nobj; ++o){
        uint64_t* S = A.flat + L.stats_off[objmap[o]];
        const uint32_t* s = A.sm + (size_t)o * 20 * A.Cs;
        for(int i = threadIdx.x; i < A.Cs; i += blockDim.x){
            uint64_t tb = 0, tq = 0;
#pragma unroll
            for(int cls = 0; cls < 5; ++cls){
                const int b = cls_to_bin(cls);
                const uint32_t n = s[(0 * 5 + cls) * A.Cs + i];
                if(n){
                    const uint32_t n20 = s[(1 * 5 + cls) * A.Cs + i], n30 = s[(2 * 5 + cls) * A.Cs + i];
                    const uint32_t qs = s[(3 * 5 + cls) * A.Cs + i];
                    add_u64(S + L.s_content + b * C + i, n);
                    if(n20) add_u64(S + L.s_q20 + b * C + i, n20);
                    if(n30) add_u64(S + L.s_q30 + b * C + i, n30);
                    if(qs) add_u64(S + L.s_qualsum + b * C + i, qs);
                    tb += n; tq += qs;
                }
            }
            if(tb){ add_u64(S + L.s_total_base + i, tb); add_u64(S + L.s_total_qual + i, tq); }
        }
    }
}

struct ViewGet {
    const ReadView& v;
    __device__ __forceinline__ void operator()(int i, int& c, int& q) const { c = v.cls(i); q = v.qual(i); }
};

// SE: pre1 (obj 0) over the raw read, post1 (obj 2) over the kept part of passing reads
__global__ void __launch_bounds__(256) k_stats_se(const __grid_constant__ DevParams D, DevBatch B,
                                                   const fq_read_result* __restrict__ res,
                                                   uint64_t* __restrict__ flat, int Cs){
    extern __shared__ uint32_t sm[];
    __shared__ unsigned long long cnt[8];
    StatsAcc A{Cs ? sm : nullptr, flat, Cs, &D.L, D.err, 1, cnt};
    if(threadIdx.x < 8) cnt[threadIdx.x] = 0;
    if(Cs) stats_smem_zero(sm, 2 * 20 * Cs);
    __syncthreads();
    const int warps = blockDim.x >> 5, w = threadIdx.x >> 5;
    const int k = D.P.kmer_len;
    for(uint32_t r = blockIdx.x * warps + w; r < B.n; r += gridDim.x * warps){
        ReadView v = make_view(B, r);
        stat_stream(A, 0, v.len, k, ViewGet{v});
        const fq_read_result rec = res[r];
        if(!(rec.flags & FQK_F_DROPPED) && rec.result == FQK_PASS_FILTER){
            v.start = rec.start; v.len = rec.len;
            stat_stream(A, 2, v.len, k, ViewGet{v});
        }
    }
    __syncthreads();
    stats_cnt_flush(A);
    if(Cs){ const int om[2] = {0, 2}; stats_smem_flush(A, 2, om); }
}

// PE: pre1, pre2 raw; post1/post2 following src/peprocessor.cpp:353-429
__global__ void __launch_bounds__(256) k_stats_pe(const __grid_constant__ DevParams D, DevBatch B1, DevBatch B2,
                                                   const fq_read_result* __restrict__ res1,
                                                   const fq_read_result* __restrict__ res2,
                                                   const fq_pair_result* __restrict__ resp,
                                                   uint64_t* __restrict__ flat, int Cs, int phase){
    extern __shared__ uint32_t sm[];
    __shared__ unsigned long long cnt[8];
    StatsAcc A{Cs ? sm : nullptr, flat, Cs, &D.L, D.err, 0, cnt};
    if(threadIdx.x < 8) cnt[threadIdx.x] = 0;
    if(Cs) stats_smem_zero(sm, 4 * 20 * Cs);
    __syncthreads();
    const int warps = blockDim.x >> 5, w = threadIdx.x >> 5;
    const int k = D.P.kmer_len;
    const bool merge = D.P.merge_enabled;
    for(uint32_t r = blockIdx.x * warps + w; r < B1.n; r += gridDim.x * warps){
        ReadView a = make_view(B1, r), b = make_view(B2, r);
        if(phase & 1){                                                // pre: raw reads (before -c corrects the batch in place)
            stat_stream(A, 0, a.len, k, ViewGet{a});
            stat_stream(A, 1, b.len, k, ViewGet{b});
        }
        if(!(phase & 2)) continue;
        const fq_read_result r1 = res1[r], r2 = res2[r];
        const fq_pair_result pr = resp[r];
        const bool alive = !(r1.flags & FQK_F_DROPPED) && !(r2.flags & FQK_F_DROPPED);
        if(!alive) continue;
        a.start = r1.start; a.len = r1.len; b.start = r2.start; b.len = r2.len;
        if(!merge){
            if(r1.result == FQK_PASS_FILTER && r2.result == FQK_PASS_FILTER){          // :392-401
                stat_stream(A, 2, a.len, k, ViewGet{a});
                stat_stream(A, 3, b.len, k, ViewGet{b});
            }
        }else if(pr.flags & FQK_P_OV2_OVERLAP){
            if(pr.flags & FQK_P_MERGED){                                               // :359-362
                MergedGet mg{a, b, pr.merge_len1, pr.ov2_len};
                stat_stream(A, 2, pr.merge_len1 + pr.merge_len2, k, mg);
            }
        }else if(pr.flags & FQK_P_MERGE_PROC){                                         // :367-380
            if(r1.result == FQK_PASS_FILTER) stat_stream(A, 2, a.len, k, ViewGet{a});
            if(r2.result == FQK_PASS_FILTER) stat_stream(A, 3, b.len, k, ViewGet{b});
        }
    }
    __syncthreads();
    stats_cnt_flush(A);
    if(Cs){ const int om[4] = {0, 1, 2, 3}; stats_smem_flush(A, 4, om); }
}

// ---------------------------------------------------------------- duplicate
// Duplicate::statRead / statPair / addRecord, src/duplicate.cpp:46-129, as an
// order-independent reduction.  Per key:
//   tab128[key] = { kmer32 (64) , count (32) | minidx (32) } under the monoid
//      smaller kmer wins; equal kmers add counts and keep the smaller index
//   first64[key] = min over reads of (global index << 8 | gc byte)
// so that after the run  mDups = kmer32, mCounts = count and
//   mGC = gc(first read of the key) if that read carries the final minimum else 0.
__device__ __forceinline__ void cas128(unsigned long long* addr, unsigned long long clo, unsigned long long chi,
                                       unsigned long long slo, unsigned long long shi,
                                       unsigned long long& olo, unsigned long long& ohi){
    asm volatile("{\n\t.reg .b128 c, s, d;\n\tmov.b128 c, {%3, %4};\n\tmov.b128 s, {%5, %6};\n\t"
                 "atom.global.cas.b128 d, [%2], c, s;\n\tmov.b128 {%0, %1}, d;\n\t}"
                 : "=l"(olo), "=l"(ohi) : "l"(addr), "l"(clo), "l"(chi), "l"(slo), "l"(shi) : "memory");
}

// 2-bit code of `n` bases from `p` (Duplicate::seq2int); false if any base is N
__device__ inline bool dup_code(const ReadView& v, int p, int n, uint64_t& out){
    uint64_t r = 0;
    for(int i = 0; i < n; ++i){
        if(v.isN(p + i)) return false;
        r = (r << 2) | (uint64_t)v.code(p + i);
    }
    out = r;
    return true;
}

__device__ inline uint32_t gc_count(const ReadView& v){
    uint32_t g = 0;
    for(int i = 0; i < v.len; ++i){ const int c = v.cls(i); g += (c == 2 || c == 3); }
    return g;
}
// The same on whole words when the read's buffers are aligned (the batch layout guarantees it for dense batches):
// C and G are the codes with bit 1 set, N and the padding carry code 0, so the GC count of a raw read is a popcount per
// packed word; the k-mer codes are a 64-bit window of the packed bases with the 2-bit groups reversed (seq2int puts the
// first base on top); an N inside [p, p+n) is bit 7 of one of its quality bytes.
__device__ __forceinline__ uint32_t gc_count_words(const ReadView& v){
    const uint32_t* w = reinterpret_cast<const uint32_t*>(v.b2);
    const int nw = (v.len + 15) >> 4;
    uint32_t g = 0;
    for(int i = 0; i < nw; ++i) g += (uint32_t)__popc(w[i] & 0xAAAAAAAAu);
    return g;
}
__device__ __forceinline__ bool dup_code_words(const ReadView& v, int p, int n, uint64_t& out){      // n <= 32, p + n <= len
    for(int i = p & ~3; i < p + n; i += 4){                              // quality words covering [p, p+n)
        uint32_t qw = *reinterpret_cast<const uint32_t*>(v.q + i) & 0x80808080u;
        if(i < p) qw &= 0xFFFFFFFFu << (8 * (p - i));
        if(i + 4 > p + n) qw &= 0xFFFFFFFFu >> (8 * (i + 4 - p - n));
        if(qw) return false;
    }
    const uint32_t* w = reinterpret_cast<const uint32_t*>(v.b2);
    const int wi = p >> 4; const uint32_t sh = 2u * ((uint32_t)p & 15u);
    const uint32_t w0 = w[wi], w1 = w[wi + 1], w2 = sh ? w[wi + 2] : 0u;      // (words past the read exist: next read or the batch's tail pad)
    const uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);          // bases p .. p+31
    const uint64_t v64 = ((uint64_t)rev2(lo) << 32) | rev2(hi);                               // base p on top
    out = n >= 32 ? v64 : (v64 >> (2 * (32 - n)));
    return true;
}

__device__ inline void dup_insert(unsigned long long* tab128, unsigned long long* first64,
                                  uint32_t key, uint64_t kmer, uint64_t gidx, uint32_t gcbyte){
    atomicMin(first64 + key, (unsigned long long)((gidx << 8) | gcbyte));
    unsigned long long* slot = tab128 + 2ull * key;
    unsigned long long olo = slot[0], ohi = slot[1];       // racy peek; the CAS validates it
    while(true){
        unsigned long long nlo, nhi;
        const uint32_t ocnt = (uint32_t)(ohi >> 32), oidx = (uint32_t)ohi;
        if(ocnt == 0 || kmer < olo){ nlo = kmer; nhi = (1ull << 32) | (uint32_t)gidx; }
        else if(kmer == olo){ nlo = olo; nhi = ((unsigned long long)(ocnt + 1) << 32) | min(oidx, (uint32_t)gidx); }
        else return;                                        // larger kmer32: ignored (:55-59)
        unsigned long long rlo, rhi;
        cas128(slot, olo, ohi, nlo, nhi, rlo, rhi);
        if(rlo == olo && rhi == ohi) return;
        olo = rlo; ohi = rhi;
    }
}

__global__ void __launch_bounds__(256) k_dup(const __grid_constant__ DevParams D, DevBatch B1, DevBatch B2, int paired,
                                              unsigned long long* tab128, unsigned long long* first64){
    const int kl = D.P.dup_keylen;
    // word-wise form when every buffer is 4-byte aligned (reads start at multiples of 16 bases)
    const bool fast = ((reinterpret_cast<uintptr_t>(B1.qual) | reinterpret_cast<uintptr_t>(B1.bases2) |
                        (paired ? reinterpret_cast<uintptr_t>(B2.qual) | reinterpret_cast<uintptr_t>(B2.bases2) : 0)) & 3) == 0;
    for(uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < B1.n; r += gridDim.x * blockDim.x){
        ReadView a = make_view(B1, r);
        uint64_t key64, kmer; uint32_t gc; int tot;
        if(!paired){
            if(a.len < 32) continue;                                        // :65-67
            if(fast){
                if(!dup_code_words(a, 0, kl, key64)) continue;
                if(!dup_code_words(a, max(0, a.len - 37), 32, kmer)) continue;
                gc = gc_count_words(a) & 0xFF;
            }else{
                if(!dup_code(a, 0, kl, key64)) continue;
                if(!dup_code(a, max(0, a.len - 37), 32, kmer)) continue;    // :70,79
                gc = gc_count(a) & 0xFF;                                    // uint8_t counter wraps, :83-89
            }
            tot = a.len;
        }else{
            ReadView b = make_view(B2, r);
            if(a.len < 32 || b.len < 32) continue;                          // :96-98
            if(fast){
                if(!dup_code_words(a, 0, kl, key64)) continue;
                if(!dup_code_words(b, 0, 32, kmer)) continue;
                gc = (gc_count_words(a) + gc_count_words(b)) & 0xFF;
            }else{
                if(!dup_code(a, 0, kl, key64)) continue;
                if(!dup_code(b, 0, 32, kmer)) continue;                     // :108
                gc = (gc_count(a) + gc_count(b)) & 0xFF;
            }
            tot = a.len + b.len;
        }
        const uint32_t gcbyte = (uint32_t)(uint8_t)round(255.0 * (double)gc / (double)tot);   // :91,127
        dup_insert(tab128, first64, (uint32_t)key64, kmer, B1.first_index + r, gcbyte);
    }
}

// after the run (and after the cross-rank merge): unpack into the reference's three arrays
__global__ void k_dup_finish(uint64_t bins, const unsigned long long* tab128, const unsigned long long* first64,
                             uint64_t* dups, uint32_t* counts, uint8_t* gcs){
    for(uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; k < bins; k += (uint64_t)gridDim.x * blockDim.x){
        const unsigned long long lo = tab128[2 * k], hi = tab128[2 * k + 1], f = first64[k];
        const uint32_t cnt = (uint32_t)(hi >> 32), idx = (uint32_t)hi;
        dups[k] = cnt ? lo : 0;
        counts[k] = cnt;
        gcs[k] = (cnt && (uint32_t)(f >> 8) == idx) ? (uint8_t)(f & 0xFF) : 0;
    }
}


// ---------------------------------------------------------------- ASCII ingest
// FASTQ characters -> the packed batch layout, on the device (the host-side twin is fqk_pack_ascii).
// Sixteen lanes per read (40 groups at 150 bp fill 16-lane steps better than 32-lane ones), a lane per 4-base group of
// the padded read: four sequence and four quality characters are
// fetched as (at most two) aligned 32-bit words each and funnel-shifted to the read's byte position, classified
// four at a time (exact zero-byte tests against A, C, G, T, N), turned into 2-bit codes ((c >> 1) & 3 with its two
// bits swapped is the reference's A0 T1 C2 G3), gathered into one byte by a multiply, and written as one
// quality word (N flag in bit 7) + one base byte; the groups past the read's end write the zero padding.
// The staging buffers are 256-byte aligned with 16 spare bytes, so the second word of a straddling group exists.
__device__ __forceinline__ uint32_t eq_bytes(uint32_t v, uint32_t pat){                  // 0x80 in every byte of v equal to pat's
    const uint32_t z = v ^ pat;
    return ~(((z & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | z) & 0x80808080u;
}
__global__ void __launch_bounds__(256) k_pack_ascii(uint32_t n, const char* __restrict__ seq, const char* __restrict__ qual,
                                                     const uint32_t* __restrict__ src_off, const uint32_t* __restrict__ dst_off,
                                                     const uint16_t* __restrict__ lengths, uint8_t* __restrict__ bases2,
                                                     uint8_t* __restrict__ qout, uint32_t* __restrict__ err,
                                                     const uint32_t* __restrict__ qsrc_off){
    // qsrc_off: where the quality characters of read r start when they do not sit at the same offset as its bases (raw
    // FASTQ blocks: both lines live in one text buffer); nullptr = src_off
    const int lane = threadIdx.x & 15, warps = blockDim.x >> 4;                           // "warp" = half-warp here
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(seq);
    const uint32_t* q32 = reinterpret_cast<const uint32_t*>(qual);
    uint32_t bad = 0;
    for(uint32_t r = blockIdx.x * warps + (threadIdx.x >> 4); r < n; r += gridDim.x * warps){
        const uint32_t so = src_off[r], dof = dst_off[r], len = lengths[r];
        const uint32_t qo = qsrc_off ? qsrc_off[r] : so;
        const uint32_t groups = ((len + 15) >> 4) << 2;
        const uint32_t sh = 8u * (so & 3u), qsh = 8u * (qo & 3u);
        for(uint32_t g = lane; g < groups; g += 16){
            const uint32_t i0 = 4u * g;
            uint32_t q = 0, code = 0;
            if(i0 < len){
                const uint32_t w = (so + i0) >> 2, qw = (qo + i0) >> 2;
                const uint32_t a0 = s32[w], b0 = q32[qw];
                uint32_t a1 = 0, b1 = 0;
                if(sh) a1 = s32[w + 1];
                if(qsh) b1 = q32[qw + 1];
                const uint32_t sv = __funnelshift_r(a0, a1, sh), qv = __funnelshift_r(b0, b1, qsh);
                const uint32_t nv = len - i0;
                const uint32_t vm = nv >= 4u ? 0xFFFFFFFFu : ((1u << (8u * nv)) - 1u);
                const uint32_t hv = vm & 0x80808080u;
                const uint32_t fn = eq_bytes(sv, 0x4E4E4E4Eu) & hv;
                const uint32_t fa = (eq_bytes(sv, 0x41414141u) | eq_bytes(sv, 0x43434343u) | eq_bytes(sv, 0x47474747u) | eq_bytes(sv, 0x54545454u)) & hv;
                bad |= ((fa | fn) ^ hv) | (qv & hv);                                      // foreign character, or a quality byte with bit 7 set
                uint32_t x = (sv >> 1) & 0x03030303u;
                x = ((x & 0x01010101u) << 1) | ((x >> 1) & 0x01010101u);
                x &= (fa >> 7) * 3u;
                code = (x * 0x01041040u) >> 24;
                q = (qv & vm) | fn;
            }
            *reinterpret_cast<uint32_t*>(qout + dof + i0) = q;
            bases2[(dof >> 2) + g] = (uint8_t)code;
        }
    }
    if(bad) atomicOr(err, 4u);
}

// ---------------------------------------------------------------- k-mer counting as its own pass (single-end)
// Stats::statRead counts, for every position i >= k-1 of a read, the k-mer ending at i iff its k bases are all in ATCG
// (Evaluator::seq2int, src/evaluator.cpp:3-49; src/stats.cpp:266-274): once for the raw read (pre) and once for the kept
// part of a passing read (post), in the coordinates of the kept part.  As a pass of its own it reads 0.25 B/base of
// packed bases, the N flags (bit 7 of the quality bytes) and the records; the fused kernels then run without k-mers
// (cfg4: k_fused_se_ws with k-mers 4.9 ms, k_se_v3 without 2.5 ms per 4 M reads).
// A warp takes a read: its packed words and N masks go to shared memory, then a lane owns 4 positions and the 16 bases
// ending at its last position are one 32-bit window (funnel shift).  Two
// privatised histograms per block for k <= 6: both (window inside the kept part of a passing read), pre only -- a window that
// counts for both Stats objects costs one shared reduction; for k > 6 global atomics on the flat buffer.  (Two
// histograms suffice: a window of the kept part is always a window of the raw read.)
__device__ __forceinline__ uint32_t nflags16_g(const uint8_t* q){               // N flags (bit 7) of 16 quality bytes as a 16-bit mask
    const uint4 v = *reinterpret_cast<const uint4*>(q);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t m = 0;
#pragma unroll
    for(int k = 0; k < 4; ++k) m |= ((((w[k] >> 7) & 0x01010101u) * 0x00204081u) >> 21 & 0xFu) << (4 * k);
    return m;
}
__global__ void __launch_bounds__(512) k_kmer_se(const __grid_constant__ DevParams D, DevBatch B, const fq_read_result* __restrict__ recs,
                                                  uint64_t* __restrict__ flat, int use_smem){
    extern __shared__ uint32_t sh_hist[];                                        // [2][4^k] when use_smem: both, pre only
    const int K = D.P.kmer_len;
    const uint32_t bins = 1u << (2 * K), kmask = bins - 1u;
    if(use_smem){ for(uint32_t i = threadIdx.x; i < 2 * bins; i += blockDim.x) sh_hist[i] = 0; __syncthreads(); }
    uint64_t* pre = flat + D.L.stats_off[0] + D.L.s_kmer;
    uint64_t* post = flat + D.L.stats_off[2] + D.L.s_kmer;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    // per warp: the packed words and N masks of the read in hand, word w at index w + 1 (index 0 = "word -1" = nothing)
    __shared__ uint32_t s_c[16][34], s_n[16][34];
    const uint32_t wmask = (1u << K) - 1u;
    // the metadata of the warp's next read is fetched while it works on the current one (the loads at the top of the loop
    // were half of the stall samples, profiles/r02v_*)
    uint32_t noff = 0, nlen = 0; uint2 nrec = make_uint2(0u, 0u);
    if(warp < B.n){ noff = B.offsets[warp]; nlen = B.lengths[warp]; nrec = *reinterpret_cast<const uint2*>(recs + warp); }
    for(uint32_t r = warp; r < B.n; r += nwarps){
        const uint32_t off = noff, len = nlen;
        const uint2 xr = nrec;                                                  // start | len << 16, result | flags << 8 | ...
        if(r + nwarps < B.n){ noff = B.offsets[r + nwarps]; nlen = B.lengths[r + nwarps]; nrec = *reinterpret_cast<const uint2*>(recs + r + nwarps); }
        const uint32_t xstart = xr.x & 0xFFFFu, xlen = xr.x >> 16, xresult = xr.y & 0xFFu, xflags = (xr.y >> 8) & 0xFFu;
        const bool pass = !(xflags & FQK_F_DROPPED) && xresult == FQK_PASS_FILTER;
        const uint32_t vs = xstart, ve = xstart + xlen;                         // kept part [vs, ve)
        const uint32_t nw = (len + 15u) >> 4;
        for(uint32_t w0 = 0; w0 < nw; w0 += 32){                                // 512 positions per round
            __syncwarp();
            {   // stage: lane l loads word w0 + l (the word before the round, for windows that reach back, goes to index 0)
                const uint32_t w = w0 + lane;
                uint32_t cw = 0, nm = 0;
                if(w < nw){ cw = *reinterpret_cast<const uint32_t*>(B.bases2 + (off >> 2) + 4u * w); nm = nflags16_g(B.qual + off + 16u * w); }
                s_c[wib][lane + 1] = cw; s_n[wib][lane + 1] = nm;
                if(lane == 0){
                    uint32_t pw = 0, pn = 0;
                    if(w0 > 0){ pw = *reinterpret_cast<const uint32_t*>(B.bases2 + (off >> 2) + 4u * (w0 - 1)); pn = nflags16_g(B.qual + off + 16u * (w0 - 1)); }
                    s_c[wib][0] = pw; s_n[wib][0] = pn; s_c[wib][33] = 0; s_n[wib][33] = 0;
                }
            }
            __syncwarp();
            // count: a lane owns 4 positions; the 16 bases ending at its last position are one 32-bit window
#pragma unroll 1
            for(uint32_t g0 = 0; g0 < 128; g0 += 32){
                const uint32_t g = g0 + lane;                                   // group of 4 positions inside the round
                const uint32_t i0 = 16u * w0 + 4u * g;                          // raw position of its first base
                if(i0 >= len) break;                                            // (uniform: later groups lie further right)
                const int s0 = (int)(4u * g) + 3 - 15 + 16;                     // window start, in bases from index 0 of s_c (word -1)
                const uint32_t wi = (uint32_t)s0 >> 4, sh = (uint32_t)s0 & 15u;
                const uint32_t win = __funnelshift_r(s_c[wib][wi], s_c[wib][wi + 1], 2u * sh);       // bases s0 .. s0+15
                const uint32_t nwin = ((s_n[wib][wi] | (s_n[wib][wi + 1] << 16)) >> sh) & 0xFFFFu;   // their N flags
#pragma unroll
                for(int j = 0; j < 4; ++j){
                    const uint32_t i = i0 + (uint32_t)j;                        // the window ends at raw position i = index 12 + j
                    if(i >= len || i + 1u < (uint32_t)K) continue;
                    const int lo = 12 + j - (K - 1);
                    if((nwin >> lo) & wmask) continue;                          // an N inside: seq2int returns -1
                    uint32_t key = __brev((win >> (2 * lo)) & kmask) >> (32 - 2 * K);   // first base on top (seq2int)
                    key = ((key >> 1) & 0x55555555u) | ((key & 0x55555555u) << 1);
                    const bool inpost = pass && i + 1u >= vs + (uint32_t)K && i < ve;   // the window lies inside the kept part
                    if(use_smem) atomicAdd(&sh_hist[(inpost ? 0u : bins) + key], 1u);
                    else { add_u64(pre + key, 1); if(inpost) add_u64(post + key, 1); }
                }
            }
        }
    }
    if(use_smem){
        __syncthreads();
        for(uint32_t i = threadIdx.x; i < bins; i += blockDim.x){
            const uint32_t both = sh_hist[i], po = sh_hist[bins + i];
            if(both + po) add_u64(pre + i, both + po);
            if(both) add_u64(post + i, both);
        }
    }
}
