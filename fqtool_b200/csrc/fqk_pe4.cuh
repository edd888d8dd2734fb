// fqk_pe4.cuh -- fused paired-end kernel k_pe_v4 (included inside fqk_fused.cu; its trim role also compiles with
// g++ under -DFQK_EMU for the CPU test suite, tests/emu).
//
// One pass over HBM does the whole per-pair sequence of PairEndProcessor::processPairEnd
// (/root/reference/src/peprocessor.cpp:261-508) and the four Stats objects (pre1, pre2, post1, post2):
//   * a tile of T pairs = two contiguous byte ranges per mate (qualities, packed bases) of the dense batches,
//     staged into shared memory by one producer warp with cp.async.bulk (NS stages, mbarrier transaction count);
//   * seven trim warps take 32-pair chunks.  Per-pair scalar work runs one lane per pair (trimAndCut, polyG as a
//     bit mask of the last 32 bases, filters, records); OverlapAnalysis::analyze runs WARP-COOPERATIVELY per pair:
//     the two reads are packed to 2-bit words (lane j owns word j), read2 is reverse-complemented with shuffles,
//     every lane tests four offsets per step against the first word (funnel shift, XOR, fold, popc), a ballot picks
//     the first surviving offset in the reference's scan order and the warp verifies it (one 16-base word per
//     lane, two warp reductions).  Nothing is kept in local memory;
//   * eight statistics warps: groups of Wr threads own the 4-cycle words of a read; even groups take read1, odd
//     groups read2.  pre = P (every raw read, dp4a quad path), post = P - D + Z (+ M for post1):  D = raw parts
//     that do not reach the output at their raw cycle, Z = kept parts of front-trimmed reads at their shifted
//     cycle, M = reverse-complemented read2 tails that OverlapAnalysis::merge appends (overlapanalysis.cpp:80-93).
//     Short cut-off tails (polyG with the PE argument quirk trims every read) are taken out of post by the trim
//     warps themselves, cycle by cycle with ballots; everything else goes through per-tile exception lists.
#pragma once
#include "fqk_pe.cuh"
#include "fqk_smem.cuh"

#define V4_FULL 0xffffffffu
#ifndef EMU_MARK2
#define EMU_MARK2(k)
#endif
#ifndef EMU_MARK
#define EMU_MARK(k)
#endif
constexpr int V4_TAILMAX = 8;                 // cut-off tails up to this many bases are handled by the trim warps
constexpr int V4_MAXLEN = 496;                // reads up to this length: one packed word per lane

// ------------------------------------------------------------------ layout
struct V4L {
    uint32_t Cr, Cm, Cp;                      // cycles of pre1/pre2/post2, of post1, padded bytes per read
    uint32_t statw_r, statw_m;                // words per statistics array
    uint32_t o_pre[2], o_post[2], o_bc, o_ins, o_bar, o_misc, o_stage;
    uint32_t qcap, o_q[2], o_b[2], o_off[2], o_len[2], o_ex[2], o_exm, o_cnt, stage_bytes, total;
    int T, NS;
};
__host__ __device__ inline V4L v4_layout(int Wr, int Wm, int T, int NS, int ins_len){
    V4L l;
    l.T = T; l.NS = NS;
    l.Cr = 4u * (uint32_t)Wr; l.Cm = 4u * (uint32_t)Wm; l.Cp = (l.Cr + 15u) & ~15u;
    l.statw_r = 4 * NSETS * l.Cr + l.Cr + 1; l.statw_m = 4 * NSETS * l.Cm + l.Cm + 1;
    uint32_t o = 128;
    l.o_pre[0] = o; o += 4 * l.statw_r;
    l.o_pre[1] = o; o += 4 * l.statw_r;
    l.o_post[0] = o; o += 4 * l.statw_m;
    l.o_post[1] = o; o += 4 * l.statw_r;
    o = (o + 7u) & ~7u; l.o_bc = o; o += (uint32_t)sizeof(BlockCounters);
    o = (o + 7u) & ~7u; l.o_ins = o; o += 4u * (uint32_t)ins_len;
    o = (o + 7u) & ~7u; l.o_bar = o; o += 8u * (uint32_t)NS;
    l.o_misc = o; o += 16;
    o = (o + 127u) & ~127u; l.o_stage = o;
    l.qcap = (uint32_t)T * l.Cp;
    uint32_t s = 0;
    for(int m = 0; m < 2; ++m){
        s = (s + 127u) & ~127u; l.o_q[m] = s; s += l.qcap + 64;
        s = (s + 15u) & ~15u; l.o_b[m] = s; s += ((l.qcap / 4 + 15u) & ~15u) + 64;
    }
    for(int m = 0; m < 2; ++m){
        s = (s + 15u) & ~15u; l.o_off[m] = s; s += (uint32_t)T * 4 + 16;
        s = (s + 15u) & ~15u; l.o_len[m] = s; s += (((uint32_t)T * 2 + 15u) & ~15u) + 16;
    }
    for(int m = 0; m < 2; ++m){ s = (s + 15u) & ~15u; l.o_ex[m] = s; s += (uint32_t)T * 8; }
    l.o_exm = s; s += (uint32_t)T * 8;
    l.o_cnt = s; s += 32;                     // [0] nex1 [1] nex2 [2] nM [3] uniform flags [4] pairs in the tile
    l.stage_bytes = (s + 127u) & ~127u;
    l.total = l.o_stage + (uint32_t)NS * l.stage_bytes;
    return l;
}

// ------------------------------------------------------------------ per-read state of a trim lane
struct PeRead {
    uint32_t qa, ba;          // shared addresses of quality byte 0 / base byte 0 of the raw read
    uint32_t doff;            // tile offset of the read
    int rawlen, start, len;   // raw length and the current view
    uint32_t ge, nmask;       // sweep of the raw read: 128 x (bases with q >= limit); 32-byte granules holding an N
};

// one uniform 16-byte sweep of the padded raw read (see fast_phase1a in fqk_se3.cuh).  Lanes start at chunk
// (lane >> 2) & 1: with a read pitch of 160 bytes the eight lanes of a quarter warp would otherwise land on four
// 16-byte bank groups (two-way conflicts); the sums do not depend on the order.
__device__ __forceinline__ void v4_sweep(PeRead& R, bool count, uint32_t L4, int lane){
    const uint32_t nch = ((uint32_t)R.rawlen + 15u) >> 4;
    uint32_t nmask = 0, ge = 0, ge2 = 0;
    const uint32_t rot = ((uint32_t)lane >> 2) & 1u;
    for(uint32_t cc = 0; cc < nch; ++cc){
        uint32_t c = cc + rot; if(c >= nch) c = 0;
        const uint4 v = lds128(R.qa + 16u * c);
        const uint32_t fl = ((v.x | v.y) | (v.z | v.w)) & 0x80808080u;
        nmask |= (fl ? 1u : 0u) << (c >> 1);
        if(count){
            ge = dp4(((v.x | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge);
            ge2 = dp4(((v.y | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge2);
            ge = dp4(((v.z | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge);
            ge2 = dp4(((v.w | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge2);
        }
    }
    R.ge = ge + ge2; R.nmask = nmask;
}

// Filter::passFilter's counts (src/filter.cpp:15-25) of the current view: the sweep of the raw read minus what the
// trims removed; the exact range scan when the average-quality test needs the sum
__device__ __forceinline__ void v4_view_counts(const fq_params& P, const PeRead& R, int& lowq, int& nn, int& total){
    total = 0;
    if(P.average_quality_limit > 0){ swar_counts(R.qa, R.start, R.len, P.low_quality_limit, true, lowq, nn, total); return; }
    lowq = R.rawlen - (int)(R.ge >> 7);
    nn = R.nmask ? n_count(R.qa, R.rawlen, R.nmask) : 0;
    if(R.start > 0){ int l1, n1; range_counts(R.qa, 0, R.start, P.low_quality_limit, l1, n1); lowq -= l1; nn -= n1; }
    const int e = R.start + R.len;
    if(e < R.rawlen){ int l2, n2; range_counts(R.qa, e, R.rawlen - e, P.low_quality_limit, l2, n2); lowq -= l2; nn -= n2; }
}
__device__ __forceinline__ ReadView v4_view(const PeRead& R){
    ReadView v; v.q = smem_ptr(R.qa); v.b2 = smem_ptr(R.ba); v.start = R.start; v.len = R.len; return v;
}
// Filter::passFilter of the current view (same decisions as pass_filter_fast, fqk_pe.cuh)
__device__ __forceinline__ int v4_pass_filter(const fq_params& P, const PeRead& R, bool isnull){
    if(isnull || R.len == 0) return FQK_FAIL_LENGTH;
    if(P.complexity_enabled) return dev_pass_filter_view(P, v4_view(R), false);
    if(!P.qual_filter_enabled && !P.length_filter_enabled) return FQK_PASS_FILTER;
    int lowq = 0, nn = 0, total = 0;
    EMU_MARK2(10);
    if(P.qual_filter_enabled) v4_view_counts(P, R, lowq, nn, total);
    EMU_MARK2(11);
    return filter_from_counts(P, R.len, lowq, nn, total);
}

// ------------------------------------------------------------------ polyG as a bit mask
__device__ __forceinline__ uint32_t even_bits16(uint32_t t){           // gather the 16 even bits of a word
    t &= 0x55555555u; t = (t | (t >> 1)) & 0x33333333u; t = (t | (t >> 2)) & 0x0F0F0F0Fu;
    t = (t | (t >> 4)) & 0x00FF00FFu; t = (t | (t >> 8)) & 0x0000FFFFu; return t;
}
// bit t of the result: base e-1-t of the staged read is 'G' (code 3; N is stored as code 0), for t < n <= min(32, e)
__device__ __forceinline__ uint32_t tail_g_mask(uint32_t ba, int e, int n){
    const int p0 = e - n, wi = p0 >> 4;
    const uint32_t sh = 2u * ((uint32_t)p0 & 15u);
    const uint32_t w0 = lds32(ba + 4u * wi), w1 = lds32(ba + 4u * wi + 4u), w2 = lds32(ba + 4u * wi + 8u);
    const uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);          // bases p0 .. p0+31
    const uint32_t g = even_bits16(lo & (lo >> 1)) | (even_bits16(hi & (hi >> 1)) << 16);       // bit k: base p0+k is G
    return __brev(g) >> (32 - n);                                                               // base e-1-t = p0 + (n-1-t)
}
// PolyX::trimPolyG (single read), src/polyx.cpp:14-38, on the current view.  The reference's loop can only break at
// a non-G base, so it is walked mismatch by mismatch over a 32-base mask; firstGpos is the highest G below the break.
__device__ inline void v4_polyg(PeRead& R, int compareReq, int maxMismatch, int oneEach, fq_read_result& rec, Tally& tl){
    const int rlen = R.len;
    int mismatch = 0, lastG = -1, iexit = rlen;
    int quo = 0, nextq = oneEach;                         // quo = (i + 1) / oneEach for the increasing i of the walk
    bool broke = false;
    for(int b0 = 0; b0 < rlen && !broke; b0 += 32){
        const int n = min(32, rlen - b0);
        const uint32_t gm = tail_g_mask(R.ba, R.start + rlen - b0, n);
        const uint32_t valid = n >= 32 ? 0xFFFFFFFFu : ((1u << n) - 1u);
        uint32_t m = ~gm & valid, below = valid;
        while(m){
            const int t = __ffs((int)m) - 1; m &= m - 1;
            const int i = b0 + t;
            ++mismatch;
            while(i + 1 >= nextq){ ++quo; nextq += oneEach; }
            if(mismatch > min(maxMismatch, max(1, quo))){ broke = true; iexit = i; below = (1u << t) - 1u; break; }
        }
        const uint32_t gb = gm & below;
        if(gb) lastG = b0 + 31 - __clz((int)gb);
    }
    if(iexit + 1 >= compareReq){
        const int firstG = lastG >= 0 ? rlen - 1 - lastG : rlen - 1;
        if(firstG >= 0 && firstG <= R.len) R.len = firstG;             // Read::resize, src/read.h:181-187
        rec.flags |= FQK_F_POLYG;
        rec.polyg_len = (uint16_t)(rlen - firstG);
        tl.polyx_reads[3] += 1; tl.polyx_bases[3] += (uint32_t)(rlen - firstG);
    }
}

// ------------------------------------------------------------------ warp-cooperative overlap analysis
struct PkLane { uint32_t c, n; };             // lane j: 2-bit codes / N flags (even bits) of bases 16j .. 16j+15 of a view

__device__ __forceinline__ uint32_t nflags16_s(uint32_t qa16){          // N flags of 16 staged quality bytes as even bits
    const uint4 v = lds128(qa16);
    const uint32_t q[4] = {v.x, v.y, v.z, v.w};
    uint32_t bits = 0;
#pragma unroll
    for(int k = 0; k < 4; ++k){
        const uint32_t y = (q[k] >> 7) & 0x01010101u;
        bits |= ((y * 0x00204081u) >> 21 & 0xFu) << (4 * k);
    }
    uint32_t x = bits;
    x = (x | (x << 8)) & 0x00FF00FFu; x = (x | (x << 4)) & 0x0F0F0F0Fu;
    x = (x | (x << 2)) & 0x33333333u; x = (x | (x << 1)) & 0x55555555u;
    return x;
}
// pack the view [start, start+len) of a staged read (start + len <= 512): same words as pack_view (fqk_pe.cuh)
__device__ __forceinline__ PkLane v4_pack(uint32_t ba, uint32_t qa, int start, int len, bool hasN, int lane){
    const int w0 = start >> 4, nsrc = (start + len + 15) >> 4, idx = w0 + lane;
    const uint32_t sh = 2u * ((uint32_t)start & 15u);
    const uint32_t cw = idx < nsrc ? lds32(ba + 4u * idx) : 0u;
    uint32_t cn = __shfl_down_sync(V4_FULL, cw, 1); if(lane == 31) cn = 0u;
    PkLane r; r.c = __funnelshift_r(cw, cn, sh); r.n = 0u;
    if(hasN){
        const uint32_t nw = idx < nsrc ? nflags16_s(qa + 16u * idx) : 0u;
        uint32_t nn = __shfl_down_sync(V4_FULL, nw, 1); if(lane == 31) nn = 0u;
        r.n = __funnelshift_r(nw, nn, sh);
    }
    const uint32_t vm = valid_mask(len - 16 * lane);
    r.n &= vm;
    r.c &= (vm | (vm << 1)) & ~(r.n | (r.n << 1));                     // bases outside the view and N bases carry code 0
    return r;
}
// Seq::reverseComplement (src/seq.h:24-48) of a packed view: same words as revcomp_packed (fqk_pe.cuh)
__device__ __forceinline__ PkLane v4_revcomp(PkLane b, int len, bool hasN, int lane){
    const int nw = (len + 15) >> 4, pad = (16 - (len & 15)) & 15, k = nw - 1 - lane;
    uint32_t hic = __shfl_sync(V4_FULL, b.c, k & 31), loc = __shfl_sync(V4_FULL, b.c, (k - 1) & 31);
    uint32_t hin = 0u, lon = 0u;
    if(hasN){ hin = __shfl_sync(V4_FULL, b.n, k & 31); lon = __shfl_sync(V4_FULL, b.n, (k - 1) & 31); }
    if(k < 0){ hic = 0u; hin = 0u; }
    if(k < 1){ loc = 0u; lon = 0u; }
    const uint32_t cw = pad ? __funnelshift_r(loc, hic, 32 - 2 * pad) : hic;
    const uint32_t nwd = pad ? __funnelshift_r(lon, hin, 32 - 2 * pad) : hin;
    const uint32_t vm = valid_mask(len - 16 * lane);
    PkLane r;
    r.n = rev2(nwd) & vm;
    r.c = (rev2(cw) ^ 0x55555555u) & (vm | (vm << 1)) & ~(r.n | (r.n << 1));
    return r;
}
// this lane's 16 bases of `v` seen from base offset o (word lane + o/16, shifted)
__device__ __forceinline__ uint32_t v4_shiftw(uint32_t v, int o, int lane){
    const int w = (o >> 4) + lane;
    uint32_t lo = __shfl_sync(V4_FULL, v, w & 31), hi = __shfl_sync(V4_FULL, v, (w + 1) & 31);
    if(w > 31) lo = 0u;
    if(w + 1 > 31) hi = 0u;
    return __funnelshift_r(lo, hi, 2u * ((uint32_t)o & 15u));
}
// one alignment a[ao + i] vs r[ro + i], i < ol (ao == 0 or ro == 0): lane i counts the mismatches of overlap
// positions 16i .. 16i+15; the reference accepts iff fewer than brk mismatches among the first min(50, ol)
// positions and (diff < limit or ol > 50)  (closed form of src/overlapanalysis.cpp:20-67)
__device__ __forceinline__ bool v4_verify(PkLane a, int ao, PkLane r, int ro, int ol, bool anyN, int limit, int brk, int lane, int& diff){
    uint32_t xc = a.c, xn = a.n, yc = r.c, yn = r.n;
    if(ao){ xc = v4_shiftw(a.c, ao, lane); if(anyN) xn = v4_shiftw(a.n, ao, lane); }
    if(ro){ yc = v4_shiftw(r.c, ro, lane); if(anyN) yn = v4_shiftw(r.n, ro, lane); }
    const uint32_t t = xc ^ yc;
    uint32_t mm = (t | (t >> 1)) & 0x55555555u;
    if(anyN) mm |= xn ^ yn;
    mm &= valid_mask(ol - 16 * lane);
    const uint32_t tot = __reduce_add_sync(V4_FULL, (uint32_t)__popc(mm));
    const uint32_t t50 = __reduce_add_sync(V4_FULL, (uint32_t)__popc(mm & valid_mask(50 - 16 * lane)));
    diff = (int)tot;
    if((int)t50 >= brk) return false;
    return (int)tot < limit || ol > 50;
}
// OverlapAnalysis::analyze (src/overlapanalysis.cpp:7-72): a = read1's view, r = revcomp of read2's view.  Every lane
// tests four offsets per step against the first word of the other read (N ignored: an N has code 0, so the code-only
// count never exceeds the true count and the reject is conservative); survivors are verified in scan order.
__device__ inline OvRes v4_analyze(PkLane a, int len1, PkLane r, int len2, bool anyN, int limit, int require, int lane){
    const int brk = max(limit, 1);
    OvRes o; o.overlapped = 0; o.offset = 0; o.len = 0; o.diff = 0;
    const bool fast = require >= 16 && len1 >= 16 && len2 >= 16;       // then every overlap is at least 16 long
    const uint32_t y0 = __shfl_sync(V4_FULL, r.c, 0), x0 = __shfl_sync(V4_FULL, a.c, 0);
    const int nA = len1 - require, nB = len2 - require;
    for(int base = 0; base < nA; base += 128){                         // src/overlapanalysis.cpp:20-41
        const int o0 = base + 4 * lane, wi = o0 >> 4;
        uint32_t lo = __shfl_sync(V4_FULL, a.c, wi & 31), hi = __shfl_sync(V4_FULL, a.c, (wi + 1) & 31);
        if(wi > 31) lo = 0u;
        if(wi + 1 > 31) hi = 0u;
        uint32_t cm = 0;
#pragma unroll
        for(int k = 0; k < 4; ++k){
            const int off = o0 + k;
            const uint32_t t = __funnelshift_r(lo, hi, 2u * (((uint32_t)o0 & 15u) + k)) ^ y0;
            uint32_t mm = (t | (t >> 1)) & 0x55555555u;
            if(!fast) mm &= valid_mask(min(min(len1 - off, len2), 16));
            if(off < nA && __popc(mm) < brk) cm |= 1u << k;
        }
        for(;;){
            const uint32_t bal = __ballot_sync(V4_FULL, cm != 0u);
            if(!bal) break;
            const int src = __ffs((int)bal) - 1;
            const uint32_t c4 = __shfl_sync(V4_FULL, cm, src);
            const int off = base + 4 * src + (__ffs((int)c4) - 1);
            const int ol = min(len1 - off, len2);
            int diff;
            if(v4_verify(a, off, r, 0, ol, anyN, limit, brk, lane, diff)){ o.overlapped = 1; o.offset = off; o.len = ol; o.diff = diff; return o; }
            if(lane == src) cm &= cm - 1u;
        }
    }
    for(int base = 0; base < nB; base += 128){                         // src/overlapanalysis.cpp:44-67 (offset = -d)
        const int d0 = base + 4 * lane, wi = d0 >> 4;
        uint32_t lo = __shfl_sync(V4_FULL, r.c, wi & 31), hi = __shfl_sync(V4_FULL, r.c, (wi + 1) & 31);
        if(wi > 31) lo = 0u;
        if(wi + 1 > 31) hi = 0u;
        uint32_t cm = 0;
#pragma unroll
        for(int k = 0; k < 4; ++k){
            const int d = d0 + k;
            const uint32_t t = __funnelshift_r(lo, hi, 2u * (((uint32_t)d0 & 15u) + k)) ^ x0;
            uint32_t mm = (t | (t >> 1)) & 0x55555555u;
            if(!fast) mm &= valid_mask(min(min(len1, len2 - d), 16));
            if(d < nB && __popc(mm) < brk) cm |= 1u << k;
        }
        for(;;){
            const uint32_t bal = __ballot_sync(V4_FULL, cm != 0u);
            if(!bal) break;
            const int src = __ffs((int)bal) - 1;
            const uint32_t c4 = __shfl_sync(V4_FULL, cm, src);
            const int d = base + 4 * src + (__ffs((int)c4) - 1);
            const int ol = min(len1, len2 - d);
            int diff;
            if(v4_verify(a, 0, r, d, ol, anyN, limit, brk, lane, diff)){ o.overlapped = 1; o.offset = -d; o.len = ol; o.diff = diff; return o; }
            if(lane == src) cm &= cm - 1u;
        }
    }
    return o;
}
// analyze() for every lane of `need`, one pair at a time by the whole warp; each lane keeps its own pair's result
__device__ inline void v4_analyze_chunk(const PeRead& A, const PeRead& B, bool need, int limit, int require, int lane, OvRes& ov){
    const uint32_t pa = (uint32_t)A.start | ((uint32_t)A.len << 12) | (A.nmask ? 1u << 31 : 0u);
    const uint32_t pb = (uint32_t)B.start | ((uint32_t)B.len << 12) | (B.nmask ? 1u << 31 : 0u);
    for(uint32_t mk = __ballot_sync(V4_FULL, need); mk; mk &= mk - 1u){
        const int src = __ffs((int)mk) - 1;
        const uint32_t qa1 = __shfl_sync(V4_FULL, A.qa, src), ba1 = __shfl_sync(V4_FULL, A.ba, src), p1 = __shfl_sync(V4_FULL, pa, src);
        const uint32_t qa2 = __shfl_sync(V4_FULL, B.qa, src), ba2 = __shfl_sync(V4_FULL, B.ba, src), p2 = __shfl_sync(V4_FULL, pb, src);
        const int s1 = (int)(p1 & 0xFFFu), l1 = (int)((p1 >> 12) & 0xFFFu), s2 = (int)(p2 & 0xFFFu), l2 = (int)((p2 >> 12) & 0xFFFu);
        const bool n1 = (p1 >> 31) != 0u, n2 = (p2 >> 31) != 0u, anyN = n1 || n2;
        const PkLane a = v4_pack(ba1, qa1, s1, l1, n1, lane);
        const PkLane b = v4_pack(ba2, qa2, s2, l2, n2, lane);
        const PkLane r = v4_revcomp(b, l2, n2, lane);
        const OvRes o = v4_analyze(a, l1, r, l2, anyN, limit, require, lane);
        if(lane == src) ov = o;
    }
}

// ------------------------------------------------------------------ one chunk of 32 pairs (a trim warp)
struct V4X {
    const DevParams* D;
    const V4L* Lp;                 // layout of the block (kept in shared memory by the kernel)
    uint32_t S32;                  // shared address of the stage
    uint32_t pre32[2], post32[2];  // shared addresses of the statistics arrays: pre1, pre2 / post1, post2
    uint32_t ins32;                // block-private insert-size histogram
    BlockCounters* bc;
    fq_read_result* out1; fq_read_result* out2; fq_pair_result* outp;   // records of the tile's first pair
    int nt;                        // pairs in the tile
};

// append the entries of the lanes in `want` to a list of the stage (one shared atomic per chunk)
__device__ __forceinline__ void v4_push(bool want, uint32_t cnt_addr, uint32_t list_addr, Desc d, int lane){
    const uint32_t mk = __ballot_sync(V4_FULL, want);
    if(!mk) return;
    uint32_t base = 0;
    if(lane == 0) base = atoms_add(cnt_addr, (uint32_t)__popc(mk));
    base = __shfl_sync(V4_FULL, base, 0);
    if(want) sts64(list_addr + 8u * (base + (uint32_t)__popc(mk & ((1u << lane) - 1u))), d.x, d.y);
}
// one count per lane in `want` into a length histogram; a single reduction when the lanes agree (the usual case)
__device__ __forceinline__ void v4_hist(bool want, uint32_t len, uint32_t hist_addr, int lane){
    const uint32_t mk = __ballot_sync(V4_FULL, want);
    if(!mk) return;
    const uint32_t l0 = __shfl_sync(V4_FULL, len, __ffs((int)mk) - 1);
    if(__all_sync(V4_FULL, !want || len == l0)){ if(lane == 0) reds_add(hist_addr + 4u * l0, (uint32_t)__popc(mk)); }
    else if(want) reds_add(hist_addr + 4u * len, 1u);
}
// D of short cut-off tails: raw cycles [plen, dlen) of the lanes in `td` are taken out of the post array, one cycle
// at a time.  Lane k < 20 owns one (kind, set) counter of the cycle: its value is a popcount of two ballots or one
// of five warp sums.
__device__ inline void v4_tail_sub(bool td, const PeRead& R, int plen, uint32_t post32, uint32_t C, int lane){
    const uint32_t cmin = __reduce_min_sync(V4_FULL, td ? (uint32_t)plen : 0xFFFFFFFFu);
    const uint32_t cmax = __reduce_max_sync(V4_FULL, td ? (uint32_t)R.rawlen : 0u);
    const int myset = lane < 15 ? lane % 5 : lane - 15, mykind = lane < 15 ? lane / 5 : KQS;     // kinds KN, KN20, KN30 then KQS
    const bool lane_on = lane < 20 && !(mykind == KN && myset == S_ALL);                        // n of "all" comes from the length histogram
    const uint32_t cell = post32 + 4u * ((uint32_t)(mykind * NSETS + myset) * C);
    for(uint32_t c = cmin; c < cmax; ++c){
        const bool in = td && (uint32_t)plen <= c && c < (uint32_t)R.rawlen;
        const uint32_t by = in ? lds8(R.qa + c) : 0u, q7 = by & 0x7Fu;
        const bool isn = (by & 0x80u) != 0u;
        const uint32_t code = (in && !isn) ? (lds8(R.ba + (c >> 2)) >> (2u * (c & 3u))) & 3u : 0u;
        const uint32_t bIn = __ballot_sync(V4_FULL, in), b20 = __ballot_sync(V4_FULL, in && q7 > 53u), b30 = __ballot_sync(V4_FULL, in && q7 > 63u);
        const uint32_t bB0 = __ballot_sync(V4_FULL, (code & 1u) != 0u), bB1 = __ballot_sync(V4_FULL, (code & 2u) != 0u), bN = __ballot_sync(V4_FULL, in && isn);
        const uint32_t qall = __reduce_add_sync(V4_FULL, in ? q7 : 0u), qb0 = __reduce_add_sync(V4_FULL, (code & 1u) ? q7 : 0u),
                       qb1 = __reduce_add_sync(V4_FULL, (code & 2u) ? q7 : 0u), qb01 = __reduce_add_sync(V4_FULL, code == 3u ? q7 : 0u);
        const uint32_t qn = bN ? __reduce_add_sync(V4_FULL, (in && isn) ? q7 : 0u) : 0u;
        const uint32_t bS = myset == S_ALL ? bIn : myset == S_B0 ? bB0 : myset == S_B1 ? bB1 : myset == S_B01 ? (bB0 & bB1) : bN;
        const uint32_t bK = mykind == KN ? 0xFFFFFFFFu : mykind == KN20 ? b20 : b30;
        const uint32_t qS = myset == S_ALL ? qall : myset == S_B0 ? qb0 : myset == S_B1 ? qb1 : myset == S_B01 ? qb01 : qn;
        const uint32_t val = mykind == KQS ? qS : (uint32_t)__popc(bS & bK);
        if(lane_on && val) reds_add(cell + 4u * c, 0u - val);
    }
}

// The loop body of PairEndProcessor::processPairEnd (src/peprocessor.cpp:292-429, threadId 0) for pairs
// 32j .. 32j+31 of a staged tile, lane = pair; same statements as k_trim_pe (fqk_kernels.cu).
__device__ inline void v4_chunk(const V4X& X, const int j, const int lane, Tally& tl){
    const DevParams& D = *X.D;
    const fq_params& P = D.P;
    const V4L& L = *X.Lp;
    const int i = 32 * j + lane;
    const bool valid = i < X.nt;
    PeRead R[2];
#pragma unroll
    for(int m = 0; m < 2; ++m){
        const uint32_t So = X.S32 + L.o_off[m], Sl = X.S32 + L.o_len[m];
        const uint32_t a0 = lds32(So), bshift = (a0 >> 2) & 15u;
        uint32_t doff = 0, dlen = 0;
        if(valid){
            const uint32_t o = lds32(So + 4u * i);
            dlen = lds16(Sl + 2u * i); doff = o - a0;
            const uint32_t padded = (dlen + 15u) & ~15u;
            bool ok = !(o & 15u) && o >= a0 && doff + padded <= L.qcap && dlen <= L.Cr && dlen <= (uint32_t)V4_MAXLEN;
            if(ok && i + 1 < X.nt) ok = (lds32(So + 4u * (i + 1)) == o + padded);          // dense, ordered batch
            if(!ok){ atomicOr(D.err, (dlen > L.Cr || dlen > (uint32_t)V4_MAXLEN) ? 1u : 2u); doff = 0; dlen = 0; sts16(Sl + 2u * i, 0u); }
        }
        R[m].qa = X.S32 + L.o_q[m] + doff; R[m].ba = X.S32 + L.o_b[m] + bshift + (doff >> 2);
        R[m].doff = doff; R[m].rawlen = (int)dlen; R[m].start = 0; R[m].len = (int)dlen; R[m].ge = 0; R[m].nmask = 0;
        v4_sweep(R[m], P.qual_filter_enabled != 0, (uint32_t)P.low_quality_limit * 0x01010101u, lane);
    }
    EMU_MARK(0);
    PeRead& a = R[0]; PeRead& b = R[1];
    fq_read_result rec1 = {}, rec2 = {};
    fq_pair_result pr = {};
    pr.insert_size = 0xFFFF;
    const bool a1 = valid && s_trim_and_cut(P, a.qa, a.rawlen, P.trim_front1, P.trim_tail1, a.start, a.len);
    const bool a2 = valid && s_trim_and_cut(P, b.qa, b.rawlen, P.trim_front2, P.trim_tail2, b.start, b.len);
    const bool both = a1 && a2;
    if(both && P.polyg_enabled){
        // src/peprocessor.cpp:297 hands (maxMismatch, oneEach, minLen) to (compareReq, maxMismatch, oneEach)
        v4_polyg(a, P.polyg_max_mismatch, P.polyg_one_mismatch_each, P.polyg_min_len, rec1, tl);
        v4_polyg(b, P.polyg_max_mismatch, P.polyg_one_mismatch_each, P.polyg_min_len, rec2, tl);
    }
    EMU_MARK(1);
    OvRes ov; ov.overlapped = 0; ov.offset = 0; ov.len = 0; ov.diff = 0;
    // first analyze(): :303 when -a/-c, else :330 (insert size only); identical inputs either way
    v4_analyze_chunk(a, b, both, P.overlap_diff_limit, P.overlap_require, lane, ov);
    EMU_MARK(2);
    int isize = -1;
    if(both){
        pr.flags |= FQK_P_OV1_DONE | (ov.overlapped ? FQK_P_OV1_OVERLAP : 0);
        pr.ov1_offset = (int16_t)ov.offset; pr.ov1_len = (uint16_t)ov.len; pr.ov1_diff = (uint16_t)ov.diff;
        isize = P.insert_size_max;                           // statInsertSize, src/peprocessor.cpp:510-523
        if(ov.overlapped) isize = ov.offset > 0 ? a.len + b.len - ov.len : ov.len;
        isize = min(isize, P.insert_size_max);
        pr.insert_size = (uint16_t)isize;
    }
    {   // insert-size histogram of the block: the unknown-size bin is shared by most pairs, one reduction for all of them
        const uint32_t mx = __ballot_sync(V4_FULL, both && isize == P.insert_size_max);
        if(lane == 0 && mx) reds_add(X.ins32 + 4u * (uint32_t)P.insert_size_max, (uint32_t)__popc(mx));
        if(both && isize != P.insert_size_max) reds_add(X.ins32 + 4u * (uint32_t)isize, 1u);
    }
    EMU_MARK(3);
    bool modified = false;                                   // reads changed since the first analyze()
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
            if(P.adapter_r1_len > 0){ ReadView v = v4_view(a); modified |= trim_by_sequence_any(D, v, 0, rec1, tl); a.len = v.len; }
            if(P.adapter_r2_len > 0){ ReadView v = v4_view(b); modified |= trim_by_sequence_any(D, v, 1, rec2, tl); b.len = v.len; }
        }
    }
    if(both && P.polyx_enabled){
        const int l1 = a.len, l2 = b.len;
        ReadView v = v4_view(a);
        dev_trim_polyx(v, P.polyx_base_mask, P.polyx_min_len, P.polyx_max_mismatch, P.polyx_one_mismatch_each, rec1, tl); a.len = v.len;
        v = v4_view(b);
        dev_trim_polyx(v, P.polyx_base_mask, P.polyx_min_len, P.polyx_max_mismatch, P.polyx_one_mismatch_each, rec2, tl); b.len = v.len;
        modified |= (a.len != l1) | (b.len != l2);
    }
    if(both){
        if(P.trim_max_len1 > 0 && P.trim_max_len1 < a.len){ a.len = P.trim_max_len1; modified = true; }
        if(P.trim_max_len2 > 0 && P.trim_max_len2 < b.len){ b.len = P.trim_max_len2; modified = true; }
    }
    // second analyze(), :354: same answer as the first when neither read changed
    EMU_MARK(4);
    OvRes ov2 = ov;
    v4_analyze_chunk(a, b, P.merge_enabled && both && modified, P.overlap_diff_limit, P.overlap_require, lane, ov2);
    EMU_MARK(5);
    bool merge_processed = false; int res1 = 0, res2 = 0;
    int fcA = -1, fnA = 0, fcB = -1, fnB = 0;                // updates of mFilterReadStats: (code, amount) x 2
    bool merged = false;
    if(P.merge_enabled && both){
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
                    const ReadView va = v4_view(a), vb = v4_view(b);
                    const int ol = ov2.len;
                    result = dev_pass_filter(P, len1 + len2, false, [&](int t, int& c, int& q){
                        if(t < len1){ c = va.cls(t); q = va.qual(t); }
                        else { const int u = ol + t - len1; c = rc_cls(vb, u); q = vb.qual(vb.len - 1 - u); }
                    });
                }else{
                    // counts of the merged read = read1 prefix + the part of read2 that is appended (order is irrelevant)
                    int l1 = 0, n1 = 0, t1 = 0, l2 = 0, n2 = 0, t2 = 0;
                    if(P.qual_filter_enabled){
                        const bool wt = P.average_quality_limit > 0;
                        swar_counts(a.qa, a.start, len1, P.low_quality_limit, wt, l1, n1, t1);
                        swar_counts(b.qa, b.start + (b.len - ov2.len - len2), len2, P.low_quality_limit, wt, l2, n2, t2);
                    }
                    result = (len1 + len2 == 0) ? FQK_FAIL_LENGTH : filter_from_counts(P, len1 + len2, l1 + l2, n1 + n2, t1 + t2);
                }
                if(result == FQK_PASS_FILTER){ pr.flags |= FQK_P_MERGED; merged = true; }
            }
            fcA = result; fnA = 2;                                  // :358
            pr.merged_result = (uint8_t)result;
            res1 = res2 = result;
            merge_processed = true;
        }else if(!P.merge_discard_unmerged){
            res1 = v4_pass_filter(P, a, false); fcA = res1; fnA = 1;
            res2 = v4_pass_filter(P, b, false); fcB = res2; fnB = 1;
            merge_processed = true;
        }
        if(merge_processed) pr.flags |= FQK_P_MERGE_PROC;
    }
    if(valid && !merge_processed){
        res1 = v4_pass_filter(P, a, !a1);
        res2 = v4_pass_filter(P, b, !a2);
        fcA = max(res1, res2); fnA = 2;                             // :390, mPaired == true
    }
    EMU_MARK(6);
    rec1.result = (uint8_t)res1; rec2.result = (uint8_t)res2;
    if(!a1) rec1.flags |= FQK_F_DROPPED; else { rec1.start = (uint16_t)a.start; rec1.len = (uint16_t)a.len; }
    if(!a2) rec2.flags |= FQK_F_DROPPED; else { rec2.start = (uint16_t)b.start; rec2.len = (uint16_t)b.len; }
    if(valid){ X.out1[i] = rec1; X.out2[i] = rec2; X.outp[i] = pr; }
    {   // run counters: the pass bin is shared by most lanes
        const uint32_t n0 = __reduce_add_sync(V4_FULL, (uint32_t)((fcA == 0 ? fnA : 0) + (fcB == 0 ? fnB : 0)));
        if(lane == 0 && n0) atomicAdd(&X.bc->filter[FQK_PASS_FILTER], n0);
        if(fcA > 0) atomicAdd(&X.bc->filter[fcA], (uint32_t)fnA);
        if(fcB > 0) atomicAdd(&X.bc->filter[fcB], (uint32_t)fnB);
        const uint32_t nm = __ballot_sync(V4_FULL, merged);
        if(lane == 0 && nm) atomicAdd(&X.bc->merged_pairs, (uint32_t)__popc(nm));
    }
    // ---- what the statistics warps need: length histograms, exception lists, short tails ----
    EMU_MARK(7);
    const uint32_t cnt32 = X.S32 + L.o_cnt;
#pragma unroll
    for(int m = 0; m < 2; ++m){
        const PeRead& Rm = R[m];
        const fq_read_result& rm = m ? rec2 : rec1;
        const uint32_t C = m ? L.Cr : L.Cm;                       // cycles of post1 / post2
        const uint32_t dlen = (uint32_t)Rm.rawlen;
        bool own = false; uint32_t ostart = rm.start, olen = rm.len, hlen = rm.len;
        if(both){                                                   // src/peprocessor.cpp:353-429
            if(!P.merge_enabled) own = res1 == FQK_PASS_FILTER && res2 == FQK_PASS_FILTER;
            else if(pr.flags & FQK_P_OV2_OVERLAP){
                if(merged && m == 0){ own = true; olen = pr.merge_len1; hlen = (uint32_t)pr.merge_len1 + pr.merge_len2; }
            }else if(pr.flags & FQK_P_MERGE_PROC) own = (m ? res2 : res1) == FQK_PASS_FILTER;
        }
        if(own && hlen > C){ atomicOr(D.err, 1u); own = false; }   // merged read longer than max_cycles
        v4_hist(valid, dlen, X.pre32[m] + 16u * NSETS * L.Cr, lane);
        v4_hist(own, hlen, X.post32[m] + 16u * NSETS * C, lane);
        const bool whole = valid && dlen > 0 && (!own || ostart > 0);                 // D = the whole raw read
        bool td = own && ostart == 0 && olen < dlen;                                  // D = raw [olen, dlen)
        const uint32_t mt = __ballot_sync(V4_FULL, td);
        bool tlong = false;
        if(mt){
            const uint32_t cmin = __reduce_min_sync(V4_FULL, td ? olen : 0xFFFFFFFFu), cmax = __reduce_max_sync(V4_FULL, td ? dlen : 0u);
            if(cmax - cmin > 2u * V4_TAILMAX){ tlong = td; td = false; }              // ragged chunk: through the list
            else { tlong = td && dlen - olen > (uint32_t)V4_TAILMAX; td = td && !tlong; }
            if(__any_sync(V4_FULL, td)) v4_tail_sub(td, Rm, (int)olen, X.post32[m], C, lane);
        }
        const Desc d = make_desc(Rm.doff, dlen, whole ? (own ? ostart : 0u) : 0u, whole ? (own ? olen : 0u) : olen, (own ? 2u : 1u));
        v4_push(whole || tlong, cnt32 + 4u * m, X.S32 + L.o_ex[m], d, lane);
    }
    {   // M: tail of the merged read, revcomp(read2 kept part)[ol, ol+len2) on cycles [len1, len1+len2) of post1;
        // cycle c reads original position pmax - c of read2
        const bool wm = merged && pr.merge_len2 > 0 && (uint32_t)pr.merge_len1 + pr.merge_len2 <= L.Cm;
        const uint32_t pmax = (uint32_t)b.start + (uint32_t)b.len - 1u - pr.ov2_len + pr.merge_len1;
        v4_push(wm, cnt32 + 8u, X.S32 + L.o_exm, make_desc(b.doff, pmax, pr.merge_len1, pr.merge_len2, 3u), lane);
    }
}
