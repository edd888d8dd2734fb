// fqk_pe.cuh -- bit-parallel paired-end helpers (one thread per pair).
//
// OverlapAnalysis::analyze (/root/reference/src/overlapanalysis.cpp:7-72) on 2-bit packed words:
// 16 bases per 32-bit word, mismatches of an alignment = popcount of the folded XOR of the
// shifted words (funnel shifts), N handled with 1-bit-per-base masks (N vs N matches, N vs
// A/C/G/T mismatches: the reference compares characters and revcomp(N) = N, src/seq.h:41-43).
// Same closed form as dev_analyze; a first-word quick reject removes
// almost every wrong offset with ~15 instructions.
#pragma once
#include "fqk_trim.cuh"

constexpr int PE_MAXW = 33;                 // words per read handled bit-parallel (528 bases); longer reads use dev_analyze

struct PackedSeq {
    uint32_t c[PE_MAXW + 2];                // 2-bit codes, base i of the view in bits 2*(i%16) of word i/16; N and padding are 0
    uint32_t n[PE_MAXW + 2];                // N flags at the even bit of each base
    int len;
    bool hasN;
};

__device__ __forceinline__ uint32_t rev2(uint32_t x){      // reverse the order of the sixteen 2-bit groups
    const uint32_t y = __brev(x);
    return ((y >> 1) & 0x55555555u) | ((y & 0x55555555u) << 1);
}
__device__ __forceinline__ uint32_t valid_mask(int v){     // even-bit mask of the first v bases of a word
    return v >= 16 ? 0x55555555u : (v <= 0 ? 0u : (((1u << (2 * v)) - 1u) & 0x55555555u));
}
// N flags of 16 bases starting at quality byte p (4-byte aligned) as even bits
__device__ __forceinline__ uint32_t nflags16(const uint32_t* q32){
    uint32_t bits = 0;
#pragma unroll
    for(int k = 0; k < 4; ++k){
        const uint32_t y = (q32[k] >> 7) & 0x01010101u;
        bits |= ((y * 0x00204081u) >> 21 & 0xFu) << (4 * k);         // gather the four flags
    }
    uint32_t x = bits;                                                 // spread 16 bits to the even positions
    x = (x | (x << 8)) & 0x00FF00FFu; x = (x | (x << 4)) & 0x0F0F0F0Fu;
    x = (x | (x << 2)) & 0x33333333u; x = (x | (x << 1)) & 0x55555555u;
    return x;
}

// pack the current view of a read (start/len inside the original, 16-base aligned original storage)
__device__ inline void pack_view(const ReadView& v, PackedSeq& s){
    const uint32_t* w32 = reinterpret_cast<const uint32_t*>(v.b2);
    const uint32_t* q32 = reinterpret_cast<const uint32_t*>(v.q);
    const int nw = (v.len + 15) >> 4;
    const int w0 = v.start >> 4, sh = 2 * (v.start & 15);
    s.len = v.len;
    // any N in the words that cover the view?
    uint32_t any = 0;
    if(v.len > 0) for(int i = (v.start >> 2); i <= ((v.start + v.len - 1) >> 2); ++i) any |= q32[i];
    s.hasN = (any & 0x80808080u) != 0;
    const int nsrc = ((v.start + v.len + 15) >> 4);                    // source words that exist for this view
    for(int j = 0; j < nw; ++j){
        const uint32_t lo = w32[w0 + j], hi = (w0 + j + 1 < nsrc) ? w32[w0 + j + 1] : 0u;
        uint32_t c = __funnelshift_r(lo, hi, sh);
        uint32_t n = 0;
        if(s.hasN){
            const uint32_t nlo = nflags16(q32 + 4 * (w0 + j));
            const uint32_t nhi = (w0 + j + 1 < nsrc) ? nflags16(q32 + 4 * (w0 + j + 1)) : 0u;
            n = __funnelshift_r(nlo, nhi, sh);
        }
        const uint32_t vm = valid_mask(v.len - 16 * j);
        n &= vm;
        c &= (vm | (vm << 1)) & ~(n | (n << 1));                       // bases outside the view and N bases carry code 0
        s.c[j] = c; s.n[j] = n;
    }
    for(int j = nw; j < nw + 2; ++j){ s.c[j] = 0; s.n[j] = 0; }        // words nw, nw+1 are read by the funnel shifts; nothing reads further
}

// reverse complement (Seq::reverseComplement, src/seq.h:24-48) of a packed view
__device__ inline void revcomp_packed(const PackedSeq& b, PackedSeq& r){
    const int len = b.len, nw = (len + 15) >> 4;
    const int pad = (16 - (len & 15)) & 15;                            // dummy bases in front so that the length is 16*nw
    r.len = len; r.hasN = b.hasN;
    for(int j = 0; j < nw; ++j){
        const int k = nw - 1 - j;                                      // source word of B' = [pad zeros] + b
        const uint32_t lo = k > 0 ? b.c[k - 1] : 0u, hi = b.c[k];
        const uint32_t nlo = k > 0 ? b.n[k - 1] : 0u, nhi = b.n[k];
        const uint32_t cw = pad ? __funnelshift_r(lo, hi, 32 - 2 * pad) : hi;
        const uint32_t nwd = pad ? __funnelshift_r(nlo, nhi, 32 - 2 * pad) : nhi;
        const uint32_t vm = valid_mask(len - 16 * j);
        const uint32_t rn = rev2(nwd) & vm;
        uint32_t rc = (rev2(cw) ^ 0x55555555u) & (vm | (vm << 1));     // complement: A<->T, C<->G is code ^ 1
        rc &= ~(rn | (rn << 1));                                       // N stays N (code 0)
        r.c[j] = rc; r.n[j] = rn;
    }
    for(int j = nw; j < nw + 2; ++j){ r.c[j] = 0; r.n[j] = 0; }
}

// mismatches between x[xo + i] and y[yo + i] for i in [from, to) ; from is a multiple of 16
__device__ __forceinline__ int mism_words(const uint32_t* xcw, const uint32_t* xnw, int xo, const uint32_t* ycw, const uint32_t* ynw, int yo,
                                          int from, int to){
    int d = 0;
    const int xs = 2 * (xo & 15), ys = 2 * (yo & 15);
    for(int i = from; i < to; i += 16){
        const int xw = (xo + i) >> 4, yw = (yo + i) >> 4;
        const uint32_t xc = __funnelshift_r(xcw[xw], xcw[xw + 1], xs), yc = __funnelshift_r(ycw[yw], ycw[yw + 1], ys);
        const uint32_t xn = __funnelshift_r(xnw[xw], xnw[xw + 1], xs), yn = __funnelshift_r(ynw[yw], ynw[yw + 1], ys);
        const uint32_t t = xc ^ yc;
        const uint32_t mm = (((t | (t >> 1)) & 0x55555555u) | (xn ^ yn)) & valid_mask(to - i);
        d += __popc(mm);
    }
    return d;
}
__device__ __forceinline__ int mism_range(const PackedSeq& x, int xo, const PackedSeq& y, int yo, int from, int to){
    return mism_words(x.c, x.n, xo, y.c, y.n, yo, from, to);
}

// one alignment: a[ao + i] vs r[ro + i], i < ol.  Returns true (and diff) when the reference accepts it.
__device__ __forceinline__ bool try_offset(const PackedSeq& a, int ao, const PackedSeq& r, int ro, int ol, int limit, int brk,
                                           int& diff){
    const int n16 = min(ol, 16);
    int d = mism_range(a, ao, r, ro, 0, n16);                          // quick reject on the first word
    if(d >= brk) return false;
    const int n50 = min(ol, 50);
    if(n50 > 16){
        d += mism_range(a, ao, r, ro, 16, n50);
        if(d >= brk) return false;                                     // the reference's scan breaks inside the first 50
    }
    if(ol > 50){
        // positions 50.. : the word holding 48,49 was counted up to 50; count its tail then whole words
        const int xs = 2 * (ao & 15), ys = 2 * (ro & 15);
        const int xw = (ao + 48) >> 4, yw = (ro + 48) >> 4;
        const uint32_t xc = __funnelshift_r(a.c[xw], a.c[xw + 1], xs), yc = __funnelshift_r(r.c[yw], r.c[yw + 1], ys);
        const uint32_t xn = __funnelshift_r(a.n[xw], a.n[xw + 1], xs), yn = __funnelshift_r(r.n[yw], r.n[yw + 1], ys);
        const uint32_t t = xc ^ yc;
        const uint32_t mm = (((t | (t >> 1)) & 0x55555555u) | (xn ^ yn)) & valid_mask(ol - 48) & ~valid_mask(2);
        d += __popc(mm);
        if(ol > 64) d += mism_range(a, ao, r, ro, 64, ol);
    }
    diff = d;
    return d < limit || ol > 50;
}

// out-of-line copy for the rare survivors of the sliding-window quick reject (keeps the scan loops small)
static __device__ __noinline__ bool try_offset_call(const PackedSeq& a, int ao, const PackedSeq& r, int ro, int ol, int limit, int brk, int& diff){
    return try_offset(a, ao, r, ro, ol, limit, brk, diff);
}

// OverlapAnalysis::analyze on packed views; r = revcomp of read2's view
__device__ inline OvRes dev_analyze_packed(const PackedSeq& a, const PackedSeq& r, int limit, int require){
    const int len1 = a.len, len2 = r.len;
    const int brk = max(limit, 1);
    OvRes o; int diff = 0;
    if(!a.hasN && !r.hasN && require >= 16 && len1 >= 16 && len2 >= 16){
        // Without N every overlap is at least 16 long, so the quick reject of try_offset (mismatches among the first
        // 16 positions >= brk) is one funnel shift of a sliding two-word window, XOR, fold and popc per offset; the
        // full test runs only for the few offsets that survive it.  Same scan order, same answers.
        const uint32_t y0 = r.c[0], x0 = a.c[0];
        const int nA = len1 - require, nB = len2 - require;
        uint32_t lo = 0, hi = 0;
        // four offsets per step, one branch for all four (a survivor is rare)
#define FQK_QR(sh, other) __popc(((__funnelshift_r(lo, hi, (sh)) ^ (other)) | ((__funnelshift_r(lo, hi, (sh)) ^ (other)) >> 1)) & 0x55555555u)
#pragma unroll 1
        for(int base = 0; base < nA; base += 4){                       // src/overlapanalysis.cpp:20-41
            if((base & 15) == 0){ lo = a.c[base >> 4]; hi = a.c[(base >> 4) + 1]; }
            const int sh = 2 * (base & 15);
            const int p0 = FQK_QR(sh, y0), p1 = FQK_QR(sh + 2, y0), p2 = FQK_QR(sh + 4, y0), p3 = FQK_QR(sh + 6, y0);
            if(min(min(p0, p1), min(p2, p3)) < brk){
                const int pk[4] = {p0, p1, p2, p3};
#pragma unroll 1
                for(int k = 0; k < 4; ++k){
                    const int offset = base + k;
                    if(pk[k] >= brk || offset >= nA) continue;
                    const int ol = min(len1 - offset, len2);
                    if(try_offset_call(a, offset, r, 0, ol, limit, brk, diff)){ o.overlapped = 1; o.offset = offset; o.len = ol; o.diff = diff; return o; }
                }
            }
        }
#pragma unroll 1
        for(int base = 0; base < nB; base += 4){                       // src/overlapanalysis.cpp:44-67 (offset = -d)
            if((base & 15) == 0){ lo = r.c[base >> 4]; hi = r.c[(base >> 4) + 1]; }
            const int sh = 2 * (base & 15);
            const int p0 = FQK_QR(sh, x0), p1 = FQK_QR(sh + 2, x0), p2 = FQK_QR(sh + 4, x0), p3 = FQK_QR(sh + 6, x0);
            if(min(min(p0, p1), min(p2, p3)) < brk){
                const int pk[4] = {p0, p1, p2, p3};
#pragma unroll 1
                for(int k = 0; k < 4; ++k){
                    const int d = base + k;
                    if(pk[k] >= brk || d >= nB) continue;
                    const int ol = min(len1, len2 - d);
                    if(try_offset_call(a, 0, r, d, ol, limit, brk, diff)){ o.overlapped = 1; o.offset = -d; o.len = ol; o.diff = diff; return o; }
                }
            }
        }
#undef FQK_QR
        o.overlapped = 0; o.offset = 0; o.len = 0; o.diff = 0;
        return o;
    }
    for(int offset = 0; offset < len1 - require; ++offset){           // src/overlapanalysis.cpp:20-41
        const int ol = min(len1 - offset, len2);
        if(try_offset(a, offset, r, 0, ol, limit, brk, diff)){ o.overlapped = 1; o.offset = offset; o.len = ol; o.diff = diff; return o; }
    }
    for(int d = 0; -d > require - len2; ++d){                          // src/overlapanalysis.cpp:44-67 (offset = -d)
        const int ol = min(len1, len2 - d);
        if(try_offset(a, 0, r, d, ol, limit, brk, diff)){ o.overlapped = 1; o.offset = -d; o.len = ol; o.diff = diff; return o; }
    }
    o.overlapped = 0; o.offset = 0; o.len = 0; o.diff = 0;
    return o;
}

// Filter::passFilter counts (src/filter.cpp:15-25) over [start, start+len) of a read in global memory
__device__ inline void swar_counts_global(const uint8_t* q, int start, int len, int lim, int& lowq, int& nn, int& total){
    lowq = nn = total = 0;
    if(len <= 0) return;
    const uint32_t* q32 = reinterpret_cast<const uint32_t*>(q);
    const int end = start + len;
    const uint32_t k = (uint32_t)(0x80 - lim) * 0x01010101u;
    int ge = 0;
    for(int wi = start >> 2; wi <= (end - 1) >> 2; ++wi){
        uint32_t v = q32[wi];
        const int p = 4 * wi;
        if(p < start || p + 4 > end){
            const int nlo = min(4, max(0, start - p)), nhi = min(4, max(0, p + 4 - end));
            v &= (nlo + nhi >= 4) ? 0u : ((0xFFFFFFFFu << (8 * nlo)) & (0xFFFFFFFFu >> (8 * nhi)));
        }
        nn += __popc(v & 0x80808080u);
        v &= 0x7F7F7F7Fu;
        ge += __popc((v + k) & 0x80808080u);
        total = (int)__dp4a(v, 0x01010101u, (uint32_t)total);
    }
    lowq = len - ge;
    total -= 33 * len;
}

// passFilter decision from the counts (same priority order as src/filter.cpp:26-45)
__device__ __forceinline__ int filter_from_counts(const fq_params& P, int rlen, int lowq, int nn, int total){
    if(P.qual_filter_enabled){
        if(lowq > P.low_quality_base_limit) return FQK_FAIL_QUALITY;
        if(P.average_quality_limit > 0 && P.average_quality_limit > (double)total / rlen) return FQK_FAIL_QUALITY;
        if(nn > P.n_base_limit) return FQK_FAIL_N_BASE;
    }
    if(P.length_filter_enabled){
        if(rlen < P.min_read_length) return FQK_FAIL_LENGTH;
        if(P.max_read_length > 0 && rlen > P.max_read_length) return FQK_FAIL_TOO_LONG;
    }
    return FQK_PASS_FILTER;
}
__device__ inline int pass_filter_fast(const fq_params& P, const ReadView& v, bool isnull){
    if(isnull || v.len == 0) return FQK_FAIL_LENGTH;
    if(P.complexity_enabled) return dev_pass_filter_view(P, v, false);
    if(!P.qual_filter_enabled && !P.length_filter_enabled) return FQK_PASS_FILTER;
    int lowq = 0, nn = 0, total = 0;
    if(P.qual_filter_enabled) swar_counts_global(v.q, v.start, v.len, P.low_quality_limit, lowq, nn, total);
    return filter_from_counts(P, v.len, lowq, nn, total);
}

// AdapterTrimmer::trimBySequence (src/adaptertrimmer.cpp:29-90) on packed words: the adapter is packed once on the
// host (DevParams::adc/adn), the read view once per call; every candidate position is one or a few XOR/popcount steps,
// with a first-word quick reject (the whole-alignment budget is cmplen/8 mismatches).
__device__ inline bool dev_trim_by_sequence_packed(ReadView& v, const uint32_t* adc, const uint32_t* adn, int alen,
                                                   fq_read_result& rec, Tally& tl){
    const int rlen = v.len;
    if(alen < 4) return false;
    if(v.start + v.len > 16 * PE_MAXW) return false;                   // caller falls back to the scalar form
    int start = 0;
    if(alen >= 16) start = -4; else if(alen >= 12) start = -3; else if(alen >= 8) start = -2;
    PackedSeq r;
    int pos; bool found = false;
    // Quick reject on the first word of the comparison, codes only, straight from the read's packed words (two words in
    // registers, one funnel shift per position): N bases carry code 0 in the batch and in the packed adapter, so the
    // code-only count never exceeds the true count -- a position the reference accepts is never dropped here -- and the
    // survivors go through the exact test below.  The view is packed (N masks, local memory) only when a position
    // survives, which most reads never see.  (Generic trim role of k_se_v3, adapter only, 4 M reads of 100 bp: 2.44 ms
    // with a packed compare per position, 2.18 ms with the reject on the packed view; profiles/r02q_*.)
    const bool aligned = (reinterpret_cast<uintptr_t>(v.b2) & 3u) == 0;
    const uint32_t* bw = reinterpret_cast<const uint32_t*>(v.b2);      // word i = bases 16i .. 16i+15 of the raw read
    bool packed = false;
    int cur = -1; uint32_t w0 = 0, w1 = 0;
    // the read's 16 bases from raw position p0, codes only
    auto window = [&](int p0) -> uint32_t {
        if((p0 >> 4) != cur){ cur = p0 >> 4; w0 = bw[cur]; w1 = bw[cur + 1]; }
        return __funnelshift_r(w0, w1, 2u * ((uint32_t)p0 & 15u));
    };
    // the exact test of one position (the reference's inner loop, on packed words)
    auto exact = [&](int p, int cmplen, int allowed) -> bool {
        if(!packed){ pack_view(v, r); packed = true; }
        const int ao = max(0, -p), ro = max(0, p), n = cmplen - ao;
        int d = mism_words(adc, adn, ao, r.c, r.n, ro, 0, min(n, 16));
        if(d > allowed) return false;
        if(n > 16) d += mism_words(adc, adn, ao, r.c, r.n, ro, 16, n);
        return d <= allowed;
    };
    // any position, with the quick reject when the words can be read directly
    auto general = [&](int p) -> bool {
        const int cmplen = min(rlen - p, alen), allowed = cmplen / 8;
        const int ao = max(0, -p), ro = max(0, p), n = cmplen - ao;
        if(aligned && n > 0){
            const uint32_t aw = ao ? __funnelshift_r(adc[0], adc[1], 2u * (uint32_t)ao) : adc[0];      // ao <= 4
            const uint32_t t = window(v.start + ro) ^ aw;
            if(__popc((t | (t >> 1)) & valid_mask(min(n, 16))) > allowed) return false;
        }
        return exact(p, cmplen, allowed);
    };
    pos = start;
    for(; pos < 0 && pos < rlen - 4 && !found; ++pos) found = general(pos);
    if(found) --pos;
    if(!found && aligned && alen >= 16){
        // the bulk: every position whose comparison covers a whole first word (cmplen >= 16).  Sixteen positions per step:
        // three words give the 32 bases from the step's first position, every window is one funnel shift by a constant;
        // the budget cmplen / 8 shrinks over the last alen positions.
        const int last = rlen - 16;
        const uint32_t A0 = adc[0];
        if(pos <= last){
            const int q0 = v.start + pos;
            const uint32_t sh = 2u * ((uint32_t)q0 & 15u);
            int wi = q0 >> 4;
            uint32_t wa = bw[wi], wb = bw[wi + 1];
            for(; pos <= last && !found; pos += 16, ++wi){
                const uint32_t wc = bw[wi + 2];
                const uint32_t B0 = __funnelshift_r(wa, wb, sh), B1 = __funnelshift_r(wb, wc, sh);     // bases pos .. pos+31 of the view
                wa = wb; wb = wc;
                const int room = rlen - pos;                                                          // cmplen of position pos + k = min(room - k, alen)
                uint32_t cm = 0;
#pragma unroll
                for(int k = 0; k < 16; ++k){
                    const uint32_t t = (k ? __funnelshift_r(B0, B1, 2 * k) : B0) ^ A0;
                    cm |= (__popc((t | (t >> 1)) & 0x55555555u) <= (min(room - k, alen) >> 3)) ? (1u << k) : 0u;
                }
                const int nv = last - pos + 1;
                if(nv < 16) cm &= (1u << nv) - 1u;
                while(cm){
                    const int k = __ffs((int)cm) - 1; cm &= cm - 1u;
                    const int cl = min(room - k, alen);
                    if(exact(pos + k, cl, cl >> 3)){ found = true; pos += k - 16; break; }             // (the loop adds 16 back)
                }
            }
            if(!found) pos = last + 1;
        }
    }
    if(!found){
        for(; pos < rlen - 4; ++pos) if(general(pos)){ found = true; break; }
    }
    if(!found) return false;
    int textlen;
    if(pos < 0){ textlen = alen + pos; v.len = 0; }
    else { textlen = rlen - pos; v.len = pos; }
    rec.flags |= FQK_F_ADAPTER_SEQ;
    rec.adapter_pos = (int16_t)pos;
    rec.adapter_len = (uint16_t)textlen;
    if(textlen > 0){ tl.adapter_reads += 1; tl.adapter_bases += (uint32_t)textlen; }
    return true;
}
// packed form when the read fits, scalar form otherwise
__device__ inline bool trim_by_sequence_any(const DevParams& D, ReadView& v, int which, fq_read_result& rec, Tally& tl){
    const int alen = which ? D.P.adapter_r2_len : D.P.adapter_r1_len;
    if(v.start + v.len <= 16 * PE_MAXW)
        return dev_trim_by_sequence_packed(v, which ? D.adc2 : D.adc1, which ? D.adn2 : D.adn1, alen, rec, tl);
    return dev_trim_by_sequence(v, which ? D.ad2 : D.ad1, alen, rec, tl);
}

// ------------------------------------------------------------------ SE trim
// loop body of SingleEndProcessor::processSingleEnd, src/seprocessor.cpp:313-340
__device__ inline void se_read(const DevParams& D, ReadView& v, fq_read_result& rec, Tally& tl){
    const fq_params& P = D.P;
    const bool alive = dev_trim_and_cut(P, v, P.trim_front1, P.trim_tail1);
    if(alive){
        if(P.polyg_enabled)
            dev_trim_polyg(v, P.polyg_min_len, P.polyg_max_mismatch, P.polyg_one_mismatch_each, rec, tl);
        if(P.adapter_trim_enabled && P.adapter_r1_len > 0)
            trim_by_sequence_any(D, v, 0, rec, tl);
        if(P.polyx_enabled)
            dev_trim_polyx(v, P.polyx_base_mask, P.polyx_min_len, P.polyx_max_mismatch, P.polyx_one_mismatch_each, rec, tl);
        if(P.trim_max_len1 > 0 && P.trim_max_len1 < v.len) view_resize(v, P.trim_max_len1);
    }
    rec.result = (uint8_t)pass_filter_fast(P, v, !alive);
    if(!alive){ rec.flags |= FQK_F_DROPPED; rec.start = 0; rec.len = 0; }
    else { rec.start = (uint16_t)v.start; rec.len = (uint16_t)v.len; }
}
