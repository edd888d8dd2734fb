// fqk_trim.cuh -- per-read device functions: quality cuts, polyG/polyX,
// adapter by sequence, passFilter, PE overlap analysis.  One thread owns one
// read (or read pair); all loops run over the packed layout of fqk_dev.cuh.
// Each function names the reference code whose results it reproduces
// bit-exactly (paths relative to /root/reference/).
#pragma once
#include "fqk_dev.cuh"

struct DevParams {
    fq_params P;
    uint8_t ad1[FQK_MAX_ADAPTER];   // adapter of read1 as base classes (0 A,1 T,2 C,3 G,4 N)
    uint8_t ad2[FQK_MAX_ADAPTER];
    uint32_t adc1[FQK_MAX_ADAPTER / 16 + 2], adn1[FQK_MAX_ADAPTER / 16 + 2];   // the adapters packed like reads: 2-bit codes, N flags
    uint32_t adc2[FQK_MAX_ADAPTER / 16 + 2], adn2[FQK_MAX_ADAPTER / 16 + 2];
    fq_stats_layout L;
    uint32_t* err;                  // device error flag: bit0 read longer than max_cycles, bit1 misaligned offset
};

// ---------------------------------------------------------------- trimAndCut
// Filter::trimAndCut, src/filter.cpp:69-189.  Returns false when the
// reference returns NULL.  Integer form of the window test:
// (double)sum/w >= 33+q  <=>  sum >= w*(33+q)   (sum, w small integers).
__device__ inline bool dev_trim_and_cut(const fq_params& P, ReadView& v, int F, int T){
    const bool ef = P.cut_front_enabled, er = P.cut_right_enabled, et = P.cut_tail_enabled;
    if(F == 0 && T == 0 && !ef && !er && !et) return true;
    int l = v.len;
    int rlen = l - F - T;
    if(rlen < 0) return false;
    if(!ef && !er && !et){                       // :80-87 (resize or substr)
        v.start += F; v.len = rlen;
        return true;
    }
    if(ef){                                      // :94-122
        const int w = P.cut_front_window;
        if(l - F - T - w <= 0) return false;
        const int thr = w * (33 + P.cut_front_qual);
        int s = F, sum = 0;
        for(int i = 0; i < w - 1; ++i) sum += v.qual(s + i);
        for(; s + w < l - T; ++s){
            sum += v.qual(s + w - 1);
            if(s > F) sum -= v.qual(s - 1);
            if(sum >= thr) break;
        }
        if(s > 0) s += w - 1;                    // drops w-1 bases of the good window
        while(s < l && v.isN(s)) ++s;
        F = s;
        rlen = l - F - T;
    }
    if(er){                                      // :125-152
        const int w = P.cut_right_window;
        if(l - F - T - w <= 0) return false;
        const int thr = w * (33 + P.cut_right_qual);
        int s = F, sum = 0;
        bool found = false;
        for(int i = 0; i < w - 1; ++i) sum += v.qual(s + i);
        for(; s + w < l - T; ++s){
            sum += v.qual(s + w - 1);
            if(s > F) sum -= v.qual(s - 1);
            if(sum < thr){ found = true; break; }
        }
        if(found){
            const int qr = 33 + P.cut_right_qual;
            while(s < l - 1 && v.qual(s) >= qr) ++s;
            rlen = s - F;
        }
    }
    if(!er && et){                               // :155-181
        const int w = P.cut_tail_window;
        if(l - F - T - w <= 0) return false;
        const int thr = w * (33 + P.cut_tail_qual);
        int t = l - T - 1, sum = 0;
        for(int i = 0; i < w - 1; ++i) sum += v.qual(t - i);
        for(; t - w >= F; --t){
            sum += v.qual(t - w + 1);
            if(t < l - T - 1) sum -= v.qual(t + 1);
            if(sum >= thr) break;
        }
        if(t < l - 1) t -= w - 1;
        while(t >= 0 && v.isN(t)) --t;
        rlen = t - F + 1;
    }
    if(rlen <= 0 || F >= l - 1) return false;    // :183-185
    v.start += F; v.len = rlen;
    return true;
}

// ---------------------------------------------------------------- passFilter
// Filter::passFilter + passLowComplexityFliter, src/filter.cpp:3-67.
// `get(i, cls, q)` yields base class and quality of base i (lets the merged
// read of OverlapAnalysis::merge reuse the same code).
template <class Get>
__device__ inline int dev_pass_filter(const fq_params& P, int rlen, bool isnull, Get get){
    if(isnull || rlen == 0) return FQK_FAIL_LENGTH;
    int lowq = 0, nn = 0, total = 0, diff = 0, prev = -1;
    const bool count = P.qual_filter_enabled || P.length_filter_enabled;
    if(count || P.complexity_enabled){
        for(int i = 0; i < rlen; ++i){
            int c, q; get(i, c, q);
            total += q - 33;
            nn += (c == CLS_N);
            lowq += (q < P.low_quality_limit);
            if(i > 0 && c != prev) ++diff;
            prev = c;
        }
    }
    if(P.qual_filter_enabled){
        if(lowq > P.low_quality_base_limit) return FQK_FAIL_QUALITY;
        if(P.average_quality_limit > 0 && P.average_quality_limit > (double)total / rlen) return FQK_FAIL_QUALITY;
        if(nn > P.n_base_limit) return FQK_FAIL_N_BASE;
    }
    if(P.length_filter_enabled){
        if(rlen < P.min_read_length) return FQK_FAIL_LENGTH;
        if(P.max_read_length > 0 && rlen > P.max_read_length) return FQK_FAIL_TOO_LONG;
    }
    if(P.complexity_enabled){
        if(rlen <= 1) return FQK_FAIL_COMPLEXITY;
        if(!((double)diff / (rlen - 1) >= P.complexity_threshold)) return FQK_FAIL_COMPLEXITY;
    }
    return FQK_PASS_FILTER;
}

__device__ inline int dev_pass_filter_view(const fq_params& P, const ReadView& v, bool isnull){
    return dev_pass_filter(P, v.len, isnull, [&](int i, int& c, int& q){ c = v.cls(i); q = v.qual(i); });
}

// per-thread tallies flushed once per block into the flat accumulator buffer
struct Tally {
    uint32_t polyx_reads[5];
    uint32_t polyx_bases[5];
    uint32_t adapter_reads, adapter_bases;
};

// --------------------------------------------------------------------- polyG
// PolyX::trimPolyG (single read), src/polyx.cpp:14-38
__device__ inline void dev_trim_polyg(ReadView& v, int compareReq, int maxMismatch, int oneEach,
                                      fq_read_result& rec, Tally& tl){
    const int rlen = v.len;
    int mismatch = 0, i = 0, firstG = rlen - 1;
    if(oneEach > 0){
        // (i + 1) / oneEach kept as a running quotient: the division was 4 % of k_trim_pe's instructions
        int quo = 0, rem = 0;
        for(i = 0; i < rlen; ++i){
            const int p = rlen - i - 1;
            if(v.cls(p) != 3) ++mismatch; else firstG = p;
            if(++rem == oneEach){ rem = 0; ++quo; }
            if(mismatch > min(maxMismatch, max(1, quo))) break;
        }
    }else{
        for(i = 0; i < rlen; ++i){
            const int p = rlen - i - 1;
            if(v.cls(p) != 3) ++mismatch; else firstG = p;
            const int allowed = min(maxMismatch, max(1, (i + 1) / oneEach));
            if(mismatch > allowed) break;
        }
    }
    if(i + 1 >= compareReq){
        view_resize(v, firstG);
        rec.flags |= FQK_F_POLYG;
        rec.polyg_len = (uint16_t)(rlen - firstG);
        tl.polyx_reads[3] += 1; tl.polyx_bases[3] += (uint32_t)(rlen - firstG);
    }
}

// --------------------------------------------------------------------- polyX
// PolyX::trimPolyX (single read), src/polyx.cpp:45-101; trimChr is a 5-bit mask
__device__ inline void dev_trim_polyx(ReadView& v, int mask, int compareReq, int maxMismatch, int oneEach,
                                      fq_read_result& rec, Tally& tl){
    const int rlen = v.len;
    int num[5] = {0, 0, 0, 0, 0};
    int pos = 0;
    for(pos = 0; pos < rlen; ++pos){
        const int c = v.cls(rlen - 1 - pos);
#pragma unroll
        for(int b = 0; b < 5; ++b) num[b] += (c == b);
        const int cmp = pos + 1;
        const int allowed = min(maxMismatch, max(1, cmp / oneEach));
        bool brk = true;
#pragma unroll
        for(int b = 0; b < 5; ++b) if(((mask >> b) & 1) && cmp - num[b] <= allowed) brk = false;
        if(brk) break;
    }
    if(pos + 1 >= compareReq){
        int poly = 0, maxc = -1;
#pragma unroll
        for(int b = 0; b < 5; ++b) if(((mask >> b) & 1) && num[b] > maxc){ maxc = num[b]; poly = b; }
        pos = min(rlen - 1, pos);
        // with rlen == 0 the reference compares '\0' (never equal) but pos > 0 is already false
        while(pos > 0 && v.cls(rlen - pos - 1) != poly) --pos;
        view_resize(v, rlen - pos - 1);
        rec.flags |= FQK_F_POLYX | (uint8_t)(poly << FQK_F_POLYX_SHIFT);
        rec.polyx_len = (uint16_t)(pos + 1);
        tl.polyx_reads[poly] += 1; tl.polyx_bases[poly] += (uint32_t)(pos + 1);
    }
}

// ------------------------------------------------------------ trimBySequence
// AdapterTrimmer::trimBySequence, src/adaptertrimmer.cpp:29-90
__device__ inline bool dev_trim_by_sequence(ReadView& v, const uint8_t* ad, int alen,
                                            fq_read_result& rec, Tally& tl){
    const int rlen = v.len;
    if(alen < 4) return false;
    int start = 0;
    if(alen >= 16) start = -4; else if(alen >= 12) start = -3; else if(alen >= 8) start = -2;
    int pos; bool found = false;
    for(pos = start; pos < rlen - 4; ++pos){
        const int cmplen = min(rlen - pos, alen);
        const int allowed = cmplen / 8;
        int mismatch = 0; bool matched = true;
        for(int i = max(0, -pos); i < cmplen; ++i){
            if(ad[i] != v.cls(i + pos)){
                if(++mismatch > allowed){ matched = false; break; }
            }
        }
        if(matched){ found = true; break; }
    }
    if(!found) return false;
    int textlen;
    if(pos < 0){ textlen = alen + pos; v.len = 0; }
    else { textlen = rlen - pos; v.len = pos; }
    rec.flags |= FQK_F_ADAPTER_SEQ;
    rec.adapter_pos = (int16_t)pos;
    rec.adapter_len = (uint16_t)textlen;
    if(textlen > 0){ tl.adapter_reads += 1; tl.adapter_bases += (uint32_t)textlen; }   // filterresult.cpp:138-142
    return true;
}

// ------------------------------------------------------------------- analyze
struct OvRes { int overlapped, offset, len, diff; };

// class of base i of reverseComplement(v), src/seq.h:24-48: A<->T (0<->1), C<->G (2<->3), N stays
__device__ __forceinline__ int rc_cls(const ReadView& v, int i){
    const int c = v.cls(v.len - 1 - i);
    return c == CLS_N ? CLS_N : (c ^ 1);
}

// OverlapAnalysis::analyze, src/overlapanalysis.cpp:7-72, in closed form
// (SURVEY.md section 8a-6): an offset is accepted iff the scan was not broken,
// i.e. fewer than max(limit,1) mismatches among the first min(50, ol) overlap
// positions, and (total diff < limit or ol > 50); the reported diff is the
// mismatch count over the whole overlap.
__device__ inline OvRes dev_analyze(const ReadView& a, const ReadView& b, int limit, int require){
    const int len1 = a.len, len2 = b.len;
    OvRes o;
    const int brk = max(limit, 1);
    for(int offset = 0; offset < len1 - require; ++offset){
        const int ol = min(len1 - offset, len2);
        int diff = 0, i = 0; bool broke = false;
        for(; i < ol; ++i){
            if(a.cls(offset + i) != rc_cls(b, i)){
                ++diff;
                if(i < 50 && diff >= brk){ broke = true; break; }
            }
        }
        if(!broke && (diff < limit || ol > 50)){
            o.overlapped = 1; o.offset = offset; o.len = ol; o.diff = diff; return o;
        }
    }
    for(int offset = 0; offset > require - len2; --offset){
        const int ol = min(len1, len2 + offset);
        int diff = 0, i = 0; bool broke = false;
        for(; i < ol; ++i){
            if(a.cls(i) != rc_cls(b, i - offset)){
                ++diff;
                if(i < 50 && diff >= brk){ broke = true; break; }
            }
        }
        if(!broke && (diff < limit || ol > 50)){
            o.overlapped = 1; o.offset = offset; o.len = ol; o.diff = diff; return o;
        }
    }
    o.overlapped = 0; o.offset = 0; o.len = 0; o.diff = 0;
    return o;
}

// ------------------------------------------------------------ block counters
// per-block shared counters flushed with one atomic each at block end
struct BlockCounters {
    uint32_t filter[FQK_FILTER_RESULT_TYPES];
    uint32_t polyx_reads[5], polyx_bases[5];
    uint32_t adapter_reads, adapter_bases, merged_pairs;
};

__device__ inline void counters_zero(BlockCounters* bc){
    uint32_t* p = reinterpret_cast<uint32_t*>(bc);
    for(int i = threadIdx.x; i < (int)(sizeof(BlockCounters) / 4); i += blockDim.x) p[i] = 0;
}
__device__ inline void counters_add_tally(BlockCounters* bc, const Tally& t){
#pragma unroll
    for(int b = 0; b < 5; ++b){
        if(t.polyx_reads[b]) atomicAdd(&bc->polyx_reads[b], t.polyx_reads[b]);
        if(t.polyx_bases[b]) atomicAdd(&bc->polyx_bases[b], t.polyx_bases[b]);
    }
    if(t.adapter_reads) atomicAdd(&bc->adapter_reads, t.adapter_reads);
    if(t.adapter_bases) atomicAdd(&bc->adapter_bases, t.adapter_bases);
}
__device__ inline void counters_flush(const BlockCounters* bc, const fq_stats_layout& L, uint64_t* flat){
    for(int i = threadIdx.x; i < FQK_FILTER_RESULT_TYPES; i += blockDim.x)
        if(bc->filter[i]) add_u64(flat + L.filter_off + i, bc->filter[i]);
    if(threadIdx.x < 5){
        if(bc->polyx_reads[threadIdx.x]) add_u64(flat + L.polyx_reads_off + threadIdx.x, bc->polyx_reads[threadIdx.x]);
        if(bc->polyx_bases[threadIdx.x]) add_u64(flat + L.polyx_bases_off + threadIdx.x, bc->polyx_bases[threadIdx.x]);
    }
    if(threadIdx.x == 0){
        if(bc->adapter_reads) add_u64(flat + L.adapter_reads_off, bc->adapter_reads);
        if(bc->adapter_bases) add_u64(flat + L.adapter_bases_off, bc->adapter_bases);
        if(bc->merged_pairs) add_u64(flat + L.merged_pairs_off, bc->merged_pairs);
    }
}


