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
// Out-of-line helpers keep the kernel small (instruction-fetch stalls were 21 % of the samples of the fully inlined
// version, profiles/r02c_*).  ptxas 12.9 fails register allocation ("register count of 168") when v4_sweep,
// v4_polyg_ni, v4_trim_and_cut or v4_polyx are out of line inside this setmaxnreg kernel, so those four stay inline
// and are instantiated once, in a loop over the two mates.
#define V4_NI __noinline__
#define V4_IN __forceinline__
#ifndef EMU_MARK2
#define EMU_MARK2(k)
#endif
#ifndef EMU_MARK
#define EMU_MARK(k)
#endif
// FQK_V4_DEBUG bit 256: per-section clock totals of the trim warps (lane 0) into g_v4prof, printed by pe4_launch
#ifndef FQK_EMU
#define V4_GSYNC(X) do{ if((X).sync_id) nb_sync((X).sync_id, (X).sync_n); }while(0)
__device__ unsigned long long g_v4prof[16];
#define V4_PROF(on, k) do{ if(on){ const long long t_ = clock64(); if(lane == 0) atomicAdd(&g_v4prof[k], (unsigned long long)(t_ - t0_)); t0_ = clock64(); } }while(0)
#define V4_PROF0(on) long long t0_ = (on) ? clock64() : 0
#define V4_PROFC(on, k) do{ if((on) && lane == 0) atomicAdd(&g_v4prof[k], 1ull); }while(0)
#else
#define V4_GSYNC(X)
#define V4_PROF(on, k)
#define V4_PROF0(on)
#define V4_PROFC(on, k)
#endif
// Option specialisation.  V4Spec<true> is the general code.  V4Spec<false> is compiled for option sets WITHOUT forced
// trims, sliding-window cuts, polyX, adapter-by-sequence, complexity filter, average-quality limit and max-length trims
// (the BASELINE configurations cfg3 and cfg5): those statements and the helpers behind them vanish from the kernel, and
// every read keeps start == 0.  The trim role is bound by instruction fetch (the code a warp walks per chunk is larger
// than the instruction cache, profiles/r02_pe_v4_notes.md), so smaller code is faster code here.
template<bool GENERIC> struct V4Spec { static constexpr bool generic = GENERIC; };
inline bool v4_lean_options(const fq_params& P){
    return !P.trim_front1 && !P.trim_tail1 && !P.trim_front2 && !P.trim_tail2 && !P.trim_max_len1 && !P.trim_max_len2 &&
           !P.cut_front_enabled && !P.cut_tail_enabled && !P.cut_right_enabled && !P.polyx_enabled &&
           !(P.adapter_trim_enabled && (P.adapter_r1_len > 0 || P.adapter_r2_len > 0)) && !P.complexity_enabled &&
           !(P.average_quality_limit > 0);
}
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
__host__ __device__ inline V4L v4_layout(int Wr, int Wm, int T, int NS, int ins_len, bool stats = true){
    V4L l;
    l.T = T; l.NS = NS;
    l.Cr = 4u * (uint32_t)Wr; l.Cm = 4u * (uint32_t)Wm; l.Cp = (l.Cr + 15u) & ~15u;
    l.statw_r = stats ? 4 * NSETS * l.Cr + l.Cr + 1 : 0; l.statw_m = stats ? 4 * NSETS * l.Cm + l.Cm + 1 : 0;   // the trim-only kernel keeps no statistics
    uint32_t o = 256;                         // the head of the block's shared memory holds this struct
    l.o_pre[0] = o; o += 4 * l.statw_r;
    l.o_pre[1] = o; o += 4 * l.statw_r;
    l.o_post[0] = o; o += 4 * l.statw_m;
    l.o_post[1] = o; o += 4 * l.statw_r;
    o = (o + 7u) & ~7u; l.o_bc = o; o += (uint32_t)sizeof(BlockCounters);
    o = (o + 7u) & ~7u; l.o_ins = o; o += 4u * (uint32_t)ins_len;
    o = (o + 7u) & ~7u; l.o_bar = o; o += 16u * (uint32_t)NS;             // full[NS] + empty[NS]
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
    l.o_cnt = s; s += 32;                     // [0] nex1 [1] nex2 [2] nM [3] uniform flags [4] pairs in the tile [5] tile tag
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
__device__ V4_IN uint2 v4_sweep(uint32_t qa, int rawlen, bool count, uint32_t L4, int lane){
    const uint32_t nch = ((uint32_t)rawlen + 15u) >> 4;
    uint32_t nmask = 0, ge = 0, ge2 = 0;
    const uint32_t rot = ((uint32_t)lane >> 2) & 1u;
    for(uint32_t cc = 0; cc < nch; ++cc){
        uint32_t c = cc + rot; if(c >= nch) c = 0;
        const uint4 v = lds128(qa + 16u * c);
        const uint32_t fl = ((v.x | v.y) | (v.z | v.w)) & 0x80808080u;
        nmask |= (fl ? 1u : 0u) << (c >> 1);
        if(count){
            ge = dp4(((v.x | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge);
            ge2 = dp4(((v.y | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge2);
            ge = dp4(((v.z | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge);
            ge2 = dp4(((v.w | 0x80808080u) - L4) & 0x80808080u, 0x01010101u, ge2);
        }
    }
    return make_uint2(ge + ge2, nmask);
}
// Filter::trimAndCut out of line: len | start << 12 | alive << 31
__device__ V4_IN uint32_t v4_trim_and_cut(const fq_params* P, uint32_t qa, int l, int F, int T){
    int st, ln;
    const bool alive = s_trim_and_cut(*P, qa, l, F, T, st, ln);
    return (uint32_t)ln | ((uint32_t)st << 12) | (alive ? 0x80000000u : 0u);
}

// Filter::passFilter's counts (src/filter.cpp:15-25) of the current view: the sweep of the raw read minus what the
// trims removed; the exact range scan when the average-quality test needs the sum
template<class SP>
__device__ __forceinline__ void v4_view_counts(const fq_params& P, const PeRead& R, int& lowq, int& nn, int& total){
    total = 0;
    if(SP::generic && P.average_quality_limit > 0){ swar_counts(R.qa, R.start, R.len, P.low_quality_limit, true, lowq, nn, total); return; }
    lowq = R.rawlen - (int)(R.ge >> 7);
    nn = R.nmask ? n_count(R.qa, R.rawlen, R.nmask) : 0;
    if(SP::generic && R.start > 0){ int l1, n1; range_counts(R.qa, 0, R.start, P.low_quality_limit, l1, n1); lowq -= l1; nn -= n1; }
    const int e = R.start + R.len;
    if(e < R.rawlen){ int l2, n2; range_counts(R.qa, e, R.rawlen - e, P.low_quality_limit, l2, n2); lowq -= l2; nn -= n2; }
}
__device__ __forceinline__ ReadView v4_view(const PeRead& R){
    ReadView v; v.q = smem_ptr(R.qa); v.b2 = smem_ptr(R.ba); v.start = R.start; v.len = R.len; return v;
}
// Filter::passFilter of the current view (same decisions as pass_filter_fast, fqk_pe.cuh)
template<class SP>
__device__ V4_NI int v4_pass_filter_ni(const fq_params* Pp, uint32_t qa, uint32_t ba, int rawlen, int start, int len, uint32_t ge, uint32_t nmask, bool isnull){
    const fq_params& P = *Pp;
    PeRead R; R.qa = qa; R.ba = ba; R.doff = 0; R.rawlen = rawlen; R.start = start; R.len = len; R.ge = ge; R.nmask = nmask;
    if(isnull || R.len == 0) return FQK_FAIL_LENGTH;
    if(SP::generic && P.complexity_enabled) return dev_pass_filter_view(P, v4_view(R), false);
    if(!P.qual_filter_enabled && !P.length_filter_enabled) return FQK_PASS_FILTER;
    int lowq = 0, nn = 0, total = 0;
    if(P.qual_filter_enabled) v4_view_counts<SP>(P, R, lowq, nn, total);
    return filter_from_counts(P, R.len, lowq, nn, total);
}
template<class SP>
__device__ __forceinline__ int v4_pass_filter(const fq_params& P, const PeRead& R, bool isnull){
    return v4_pass_filter_ni<SP>(&P, R.qa, R.ba, R.rawlen, SP::generic ? R.start : 0, R.len, R.ge, R.nmask, isnull);
}
// passFilter of the merged read (OverlapAnalysis::merge, src/overlapanalysis.cpp:74-93): read1's first len1 bases + the
// last len2 bases of revcomp(read2's view)
template<class SP>
__device__ V4_NI int v4_merged_filter(const fq_params* Pp, uint32_t aqa, uint32_t aba, int astart, int alen,
                                             uint32_t bqa, uint32_t bba, int bstart, int blen, int ol, int len1, int len2){
    const fq_params& P = *Pp;
    if(SP::generic && P.complexity_enabled){
        ReadView va; va.q = smem_ptr(aqa); va.b2 = smem_ptr(aba); va.start = astart; va.len = alen;
        ReadView vb; vb.q = smem_ptr(bqa); vb.b2 = smem_ptr(bba); vb.start = bstart; vb.len = blen;
        return dev_pass_filter(P, len1 + len2, false, [&](int t, int& c, int& q){
            if(t < len1){ c = va.cls(t); q = va.qual(t); }
            else { const int u = ol + t - len1; c = rc_cls(vb, u); q = vb.qual(vb.len - 1 - u); }
        });
    }
    // counts of the merged read = read1 prefix + the part of read2 that is appended (order is irrelevant)
    int l1 = 0, n1 = 0, t1 = 0, l2 = 0, n2 = 0, t2 = 0;
    if(P.qual_filter_enabled){
        const bool wt = SP::generic && P.average_quality_limit > 0;
        swar_counts(aqa, astart, len1, P.low_quality_limit, wt, l1, n1, t1);
        swar_counts(bqa, bstart + (blen - ol - len2), len2, P.low_quality_limit, wt, l2, n2, t2);
    }
    return (len1 + len2 == 0) ? FQK_FAIL_LENGTH : filter_from_counts(P, len1 + len2, l1 + l2, n1 + n2, t1 + t2);
}
// AdapterTrimmer::trimBySequence / PolyX::trimPolyX out of line (rare option sets): the new length is returned, record
// fields and tallies go through temporaries so that the caller's stay in registers
__device__ V4_NI int v4_adapter_seq(const DevParams* D, uint32_t qa, uint32_t ba, int start, int len, int which, fq_read_result* rec, Tally* tl){
    ReadView v; v.q = smem_ptr(qa); v.b2 = smem_ptr(ba); v.start = start; v.len = len;
    const bool hit = trim_by_sequence_any(*D, v, which, *rec, *tl);
    return v.len | (hit ? 0x40000000 : 0);
}
__device__ V4_IN int v4_polyx(const fq_params* P, uint32_t qa, uint32_t ba, int start, int len, fq_read_result* rec, Tally* tl){
    ReadView v; v.q = smem_ptr(qa); v.b2 = smem_ptr(ba); v.start = start; v.len = len;
    dev_trim_polyx(v, P->polyx_base_mask, P->polyx_min_len, P->polyx_max_mismatch, P->polyx_one_mismatch_each, *rec, *tl);
    return v.len;
}
__device__ __forceinline__ void v4_merge_rec(fq_read_result& rec, Tally& tl, const fq_read_result& r2, const Tally& t2){
    rec.flags |= r2.flags; 
    if(r2.flags & FQK_F_ADAPTER_SEQ){ rec.adapter_pos = r2.adapter_pos; rec.adapter_len = r2.adapter_len; }
    if(r2.flags & FQK_F_POLYX) rec.polyx_len = r2.polyx_len;
#pragma unroll
    for(int b = 0; b < 5; ++b){ tl.polyx_reads[b] += t2.polyx_reads[b]; tl.polyx_bases[b] += t2.polyx_bases[b]; }
    tl.adapter_reads += t2.adapter_reads; tl.adapter_bases += t2.adapter_bases;
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
// Returns new length | tallied bases << 12 | trimmed << 31.
__device__ V4_IN uint32_t v4_polyg_ni(uint32_t ba, int start, int rlen, int compareReq, int maxMismatch, int oneEach){
    int mismatch = 0, lastG = -1, iexit = rlen;
    int quo = 0, nextq = oneEach;                         // quo = (i + 1) / oneEach for the increasing i of the walk
    bool broke = false;
    for(int b0 = 0; b0 < rlen && !broke; b0 += 32){
        const int n = min(32, rlen - b0);
        const uint32_t gm = tail_g_mask(ba, start + rlen - b0, n);
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
    if(iexit + 1 < compareReq) return (uint32_t)rlen;
    const int firstG = lastG >= 0 ? rlen - 1 - lastG : rlen - 1;
    const int nl = (firstG >= 0 && firstG <= rlen) ? firstG : rlen;     // Read::resize, src/read.h:181-187
    return (uint32_t)nl | ((uint32_t)(rlen - firstG) << 12) | 0x80000000u;
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
// positions and (diff < limit or ol > 50)  (closed form of src/overlapanalysis.cpp:20-67).
// Returns diff | accepted << 31.  Out of line: the kernel's code size matters (instruction-fetch stalls were 21 %
// of the first version's samples, profiles/r02c_*).
__device__ V4_NI uint32_t v4_verify(uint32_t ac, uint32_t an, int ao, uint32_t rc, uint32_t rn, int ro, int ol, bool anyN,
                                           int limit, int brk, int lane){
    uint32_t xc = ac, xn = an, yc = rc, yn = rn;
    if(ao){ xc = v4_shiftw(ac, ao, lane); if(anyN) xn = v4_shiftw(an, ao, lane); }
    if(ro){ yc = v4_shiftw(rc, ro, lane); if(anyN) yn = v4_shiftw(rn, ro, lane); }
    const uint32_t t = xc ^ yc;
    uint32_t mm = (t | (t >> 1)) & 0x55555555u;
    if(anyN) mm |= xn ^ yn;
    mm &= valid_mask(ol - 16 * lane);
    const uint32_t tot = __reduce_add_sync(V4_FULL, (uint32_t)__popc(mm));
    const uint32_t t50 = __reduce_add_sync(V4_FULL, (uint32_t)__popc(mm & valid_mask(50 - 16 * lane)));
    const bool acc = (int)t50 < brk && ((int)tot < limit || ol > 50);
    return tot | (acc ? 0x80000000u : 0u);
}
__device__ __forceinline__ uint2 ov_pack(int overlapped, int offset, int len, int diff){
    return make_uint2((uint32_t)offset, (uint32_t)len | ((uint32_t)diff << 12) | (overlapped ? 0x80000000u : 0u));
}
__device__ __forceinline__ OvRes ov_unpack(uint2 v){
    OvRes o; o.offset = (int)v.x; o.len = (int)(v.y & 0xFFFu); o.diff = (int)((v.y >> 12) & 0xFFFu); o.overlapped = (int)(v.y >> 31); return o;
}
// both views of one pair packed and read2 reverse-complemented by the whole warp
struct PkPair { PkLane a, r; int l1, l2; bool anyN; };
__device__ __forceinline__ PkPair v4_pack_pair(uint32_t qa1, uint32_t ba1, uint32_t p1, uint32_t qa2, uint32_t ba2, uint32_t p2, int lane){
    PkPair k;
    const int s1 = (int)(p1 & 0xFFFu), s2 = (int)(p2 & 0xFFFu);
    k.l1 = (int)((p1 >> 12) & 0xFFFu); k.l2 = (int)((p2 >> 12) & 0xFFFu);
    const bool n1 = (p1 >> 31) != 0u, n2 = (p2 >> 31) != 0u;
    k.anyN = n1 || n2;
    k.a = v4_pack(ba1, qa1, s1, k.l1, n1, lane);
    const PkLane b = v4_pack(ba2, qa2, s2, k.l2, n2, lane);
    k.r = v4_revcomp(b, k.l2, n2, lane);
    return k;
}
// OverlapAnalysis::analyze (src/overlapanalysis.cpp:7-72) of ONE pair by the whole warp: every lane tests four
// offsets per step against the first word of the other read (N ignored: an N has code 0 in both packed reads, so the
// code-only count never exceeds the true count and the reject is conservative); survivors are verified in scan order.
// Used when only a few pairs of a chunk need an analysis (the merge-time second call).
__device__ V4_NI uint2 v4_analyze_coop(uint32_t qa1, uint32_t ba1, uint32_t p1, uint32_t qa2, uint32_t ba2, uint32_t p2,
                                              int limit, int require, int lane){
    const PkPair K = v4_pack_pair(qa1, ba1, p1, qa2, ba2, p2, lane);
    const int len1 = K.l1, len2 = K.l2;
    const int brk = max(limit, 1);
    const bool fast = require >= 16 && len1 >= 16 && len2 >= 16;       // then every overlap is at least 16 long
    const uint32_t first[2] = {__shfl_sync(V4_FULL, K.r.c, 0), __shfl_sync(V4_FULL, K.a.c, 0)};
#pragma unroll 1
    for(int phase = 0; phase < 2; ++phase){                            // 0: :20-41 (read1 slides), 1: :44-67 (offset = -d)
        const int nmine = (phase ? len2 : len1) - require;
        const uint32_t sl = phase ? K.r.c : K.a.c, other = first[phase];
        for(int base = 0; base < nmine; base += 128){
            const int o0 = base + 4 * lane, wi = o0 >> 4;
            uint32_t lo = __shfl_sync(V4_FULL, sl, wi & 31), hi = __shfl_sync(V4_FULL, sl, (wi + 1) & 31);
            if(wi > 31) lo = 0u;
            if(wi + 1 > 31) hi = 0u;
            uint32_t cm = 0;
#pragma unroll
            for(int k = 0; k < 4; ++k){
                const int off = o0 + k;
                const uint32_t t = __funnelshift_r(lo, hi, 2u * (((uint32_t)o0 & 15u) + k)) ^ other;
                uint32_t mm = (t | (t >> 1)) & 0x55555555u;
                if(!fast) mm &= valid_mask(min(phase ? min(len1, len2 - off) : min(len1 - off, len2), 16));
                if(off < nmine && __popc(mm) < brk) cm |= 1u << k;
            }
            for(;;){
                const uint32_t bal = __ballot_sync(V4_FULL, cm != 0u);
                if(!bal) break;
                const int src = __ffs((int)bal) - 1;
                const uint32_t c4 = __shfl_sync(V4_FULL, cm, src);
                const int off = base + 4 * src + (__ffs((int)c4) - 1);
                const int ol = phase ? min(len1, len2 - off) : min(len1 - off, len2);
                const uint32_t v = v4_verify(K.a.c, K.a.n, phase ? 0 : off, K.r.c, K.r.n, phase ? off : 0, ol, K.anyN, limit, brk, lane);
                if(v >> 31) return ov_pack(1, phase ? -off : off, ol, (int)(v & 0x7FFFFFFFu));
                if(lane == src) cm &= cm - 1u;
            }
        }
    }
    return ov_pack(0, 0, 0, 0);
}
// 16 bases of a staged read starting at base position p (p may lie outside the read: the caller masks)
struct Win3 { uint32_t w0, w1, w2; uint32_t sh; };                     // three words from the word holding position q
__device__ __forceinline__ Win3 v4_win3(uint32_t ba, int q){
    Win3 r; const uint32_t a = ba + 4u * (uint32_t)(q >> 4);           // arithmetic shift: q may be negative
    r.w0 = lds32(a); r.w1 = lds32(a + 4u); r.w2 = lds32(a + 8u); r.sh = (uint32_t)q & 15u;
    return r;
}
__device__ __forceinline__ uint32_t v4_win16(const Win3& w, uint32_t k){   // bases q+k .. q+k+15, k <= 3
    const uint32_t s = w.sh + k;
    return s < 16u ? __funnelshift_r(w.w0, w.w1, 2u * s) : __funnelshift_r(w.w1, w.w2, 2u * (s - 16u));
}

// OverlapAnalysis::analyze (src/overlapanalysis.cpp:7-72) for the pairs of a chunk.
//   scan   every lane walks ITS pair's offsets in the reference's order, four per step, all lanes in lock step:
//          the first min(16, ol) overlap positions are one unaligned 32-bit window of the staged 2-bit bases against
//          the first word of the other read (loop A: read1 slides over revcomp(read2)[0..16); loop B: read2 slides,
//          compared in read2's orientation against the reverse complement of read1's first word), XOR, fold, popc.
//          N bases carry code 0, so the code-only count can exceed the true count only where BOTH reads have an N:
//          an offset is dropped when the count reaches brk + min(#N read1, #N read2)  (conservative);
//   verify survivors are checked exactly by the whole warp, one pair at a time, in each lane's scan order: both
//          views packed (lane j owns word j), read2 reverse-complemented with shuffles, one 16-base word per lane,
//          two warp sums (v4_verify).
__device__ V4_NI uint2 v4_overlap_lockstep(uint32_t Aqa, uint32_t Aba, uint32_t pa, uint32_t Bqa, uint32_t Bba, uint32_t pb,
                                           int nboth, bool need, int limit, int require, int lane){
    // (timing experiments: bits 16.. of nboth carry FQK_V4_DEBUG; 64 = skip the verification of survivors)
    // every pair of the chunk has overlaps of at least 16 bases (the caller checks): no length masks in the scan
    const int brk = max(limit, 1);
    const int s1 = (int)(pa & 0xFFFu), len1 = (int)((pa >> 12) & 0xFFFu), s2 = (int)(pb & 0xFFFu), len2 = (int)((pb >> 12) & 0xFFFu);
    const int dbgx = nboth >> 16; nboth &= 0xFFFF;
    const int qbrk = brk + nboth;
    const int nA = need ? max(len1 - require, 0) : 0, nB = need ? max(len2 - require, 0) : 0;
    // first word of the other read, in the orientation of the sliding read
    const uint32_t r0 = rev2(v4_win16(v4_win3(Bba, s2 + len2 - 16), 0)) ^ 0x55555555u;        // revcomp(read2)[0..16)
    const uint32_t a0rc = rev2(v4_win16(v4_win3(Aba, s1), 0)) ^ 0x55555555u;                  // revcomp of read1[0..16)
    uint2 res = ov_pack(0, 0, 0, 0);
    bool done = !need;
#pragma unroll 1
    for(int phase = 0; phase < 2; ++phase){                                // 0: src/overlapanalysis.cpp:20-41, 1: :44-67 (offset = -d)
        const int nmine = phase ? nB : nA;
        const int nmax = (int)__reduce_max_sync(V4_FULL, done ? 0u : (uint32_t)nmine);
        const uint32_t other = phase ? a0rc : r0;
        // Sixteen offsets per step.  Loop A: offset base+k is the window at read position s1+base+k.  Loop B: d = base+k is
        // the window that ENDS at len2-d, i.e. starts at q+15-k with q = s2+len2-base-31.  The step is 16 bases, so the
        // position of the lowest window modulo 16 never changes: two aligned words (32 bases from q) per step, and every
        // window is one funnel shift by a compile-time amount.
        const int q00 = phase ? s2 + len2 - 31 : s1;
        const uint32_t sh = 2u * ((uint32_t)q00 & 15u);
        uint32_t adr = (phase ? Bba : Aba) + 4u * (uint32_t)(q00 >> 4);
#pragma unroll 1
        for(int base = 0; base < nmax; base += 16, adr += phase ? 0xFFFFFFFCu : 4u){
            uint32_t cm = 0;
            if(!done && base < nmine){
                const uint32_t w0 = lds32(adr), w1 = lds32(adr + 4u), w2 = lds32(adr + 8u);
                const uint32_t A0 = __funnelshift_r(w0, w1, sh), A1 = __funnelshift_r(w1, w2, sh);   // bases q .. q+31
                const int lim = nmine - base;                                 // offsets base+k with k < lim exist
#pragma unroll
                for(int k = 0; k < 16; ++k){
                    const uint32_t win = phase ? (k == 15 ? A0 : __funnelshift_r(A0, A1, 2 * (15 - k))) : (k == 0 ? A0 : __funnelshift_r(A0, A1, 2 * k));
                    const uint32_t t = win ^ other;
                    const uint32_t mm = (t | (t >> 1)) & 0x55555555u;
                    cm |= (__popc(mm) < qbrk && k < lim) ? (1u << k) : 0u;
                }
            }
            if(dbgx & 64) cm = 0u;
            V4_PROF0(dbgx & 256);
            while(__any_sync(V4_FULL, cm != 0u)){
                V4_PROFC(dbgx & 256, 9);
                const uint32_t bal = __ballot_sync(V4_FULL, cm != 0u);
                const int src = __ffs((int)bal) - 1;
                const PkPair K = v4_pack_pair(__shfl_sync(V4_FULL, Aqa, src), __shfl_sync(V4_FULL, Aba, src), __shfl_sync(V4_FULL, pa, src),
                                              __shfl_sync(V4_FULL, Bqa, src), __shfl_sync(V4_FULL, Bba, src), __shfl_sync(V4_FULL, pb, src), lane);
                const uint32_t c16 = __shfl_sync(V4_FULL, cm, src);
                const int off = base + (__ffs((int)c16) - 1);
                const int ol = phase ? min(K.l1, K.l2 - off) : min(K.l1 - off, K.l2);
                const uint32_t v = v4_verify(K.a.c, K.a.n, phase ? 0 : off, K.r.c, K.r.n, phase ? off : 0, ol, K.anyN, limit, brk, lane);
                if(lane == src){
                    if(v >> 31){ res = ov_pack(1, phase ? -off : off, ol, (int)(v & 0x7FFFFFFFu)); done = true; cm = 0u; }
                    else cm &= cm - 1u;
                }
            }
            V4_PROF(dbgx & 256, 8);
        }
    }
    return res;
}
// analyze() for the lanes of `need`: a few pairs go through the warp one at a time, a chunk full of them in lock step
constexpr int V4_COOP_MAX = 10;
__device__ __forceinline__ void v4_overlap_chunk(const PeRead& A, const PeRead& B, bool need, int limit, int require, int lane, OvRes& ov, int dbg = 0){
    const uint32_t mk0 = __ballot_sync(V4_FULL, need);
    if(!mk0) return;
    const uint32_t pa = (uint32_t)A.start | ((uint32_t)A.len << 12) | (A.nmask ? 1u << 31 : 0u);
    const uint32_t pb = (uint32_t)B.start | ((uint32_t)B.len << 12) | (B.nmask ? 1u << 31 : 0u);
    const bool fastl = require >= 16 && A.len >= 16 && B.len >= 16;        // then every overlap of this pair is at least 16 long
    if(__popc(mk0) <= V4_COOP_MAX || __any_sync(V4_FULL, need && !fastl)){
        for(uint32_t mk = mk0; mk; mk &= mk - 1u){
            const int src = __ffs((int)mk) - 1;
            const uint2 o = v4_analyze_coop(__shfl_sync(V4_FULL, A.qa, src), __shfl_sync(V4_FULL, A.ba, src), __shfl_sync(V4_FULL, pa, src),
                                            __shfl_sync(V4_FULL, B.qa, src), __shfl_sync(V4_FULL, B.ba, src), __shfl_sync(V4_FULL, pb, src),
                                            limit, require, lane);
            if(lane == src) ov = ov_unpack(o);
        }
        return;
    }
    int nboth = 0;
    if(need && A.nmask && B.nmask) nboth = min(n_count(A.qa, A.rawlen, A.nmask), n_count(B.qa, B.rawlen, B.nmask));
    const uint2 o = v4_overlap_lockstep(A.qa, A.ba, pa, B.qa, B.ba, pb, nboth | (dbg << 16), need, limit, require, lane);
    if(need) ov = ov_unpack(o);
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
    int sync_id, sync_n;           // named barrier of the warps that work on the same tile (0 = none), threads in it
    int dbg;                       // timing experiments (FQK_V4_DEBUG): 4 no first analyze, 8 no second, 16 no tail corrections, 32 no polyG
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
__device__ V4_NI void v4_tail_sub(bool td, uint32_t Rqa, uint32_t Rba, int Rrawlen, int plen, uint32_t post32, uint32_t C, int lane){
    const uint32_t cmin = __reduce_min_sync(V4_FULL, td ? (uint32_t)plen : 0xFFFFFFFFu);
    const uint32_t cmax = __reduce_max_sync(V4_FULL, td ? (uint32_t)Rrawlen : 0u);
    const int myset = lane < 15 ? lane % 5 : lane - 15, mykind = lane < 15 ? lane / 5 : KQS;     // kinds KN, KN20, KN30 then KQS
    const bool lane_on = lane < 20 && !(mykind == KN && myset == S_ALL);                        // n of "all" comes from the length histogram
    const uint32_t cell = post32 + 4u * ((uint32_t)(mykind * NSETS + myset) * C);
    for(uint32_t c = cmin; c < cmax; ++c){
        const bool in = td && (uint32_t)plen <= c && c < (uint32_t)Rrawlen;
        const uint32_t by = in ? lds8(Rqa + c) : 0u, q7 = by & 0x7Fu;
        const bool isn = (by & 0x80u) != 0u;
        const uint32_t code = (in && !isn) ? (lds8(Rba + (c >> 2)) >> (2u * (c & 3u))) & 3u : 0u;
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
// MODE 0: the whole body (fused kernel).  MODE 1: everything up to the records (trim kernel of the split path).
// MODE 2: views and results are read back from the records and only the statistics bookkeeping runs (statistics
// kernel of the split path).
template<class SP, int MODE>
__device__ inline void v4_chunk(const V4X& X, const int j, const int lane, Tally& tl){
    const DevParams& D = *X.D;
    const fq_params& P = D.P;
    const V4L& L = *X.Lp;
    const int i = 32 * j + lane;
    const bool valid = i < X.nt;
    PeRead a, b;
    bool a1 = false, a2 = false;
    fq_read_result rec1 = {}, rec2 = {};
    fq_pair_result pr = {};
    int res1 = 0, res2 = 0;
    bool merged = false;
    const bool prof = (X.dbg & 256) != 0;
    V4_PROF0(prof);
#pragma unroll 1
    for(int m = 0; m < 2; ++m){                                     // one instance of the per-read code for both mates
        V4_GSYNC(X);
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
        PeRead r;
        r.qa = X.S32 + L.o_q[m] + doff; r.ba = X.S32 + L.o_b[m] + bshift + (doff >> 2);
        r.doff = doff; r.rawlen = (int)dlen; r.start = 0; r.len = (int)dlen;
        bool alive = false;
        if(MODE == 2){                                              // the final view, from the record of the trim kernel
            r.ge = 0; r.nmask = 0;
            fq_read_result rr = {};
            if(valid) rr = m ? X.out2[i] : X.out1[i];
            alive = valid && !(rr.flags & FQK_F_DROPPED); r.start = rr.start; r.len = rr.len;
            if(m) rec2 = rr; else rec1 = rr;
            if(m){ b = r; a2 = alive; } else { a = r; a1 = alive; }
            continue;
        }
        const uint2 sw = v4_sweep(r.qa, (int)dlen, P.qual_filter_enabled != 0, (uint32_t)P.low_quality_limit * 0x01010101u, lane);
        r.ge = sw.x; r.nmask = sw.y;
        if(!SP::generic) alive = valid;                             // Filter::trimAndCut returns the read itself (src/filter.cpp:70-73)
        else if(valid){
            const uint32_t c = v4_trim_and_cut(&P, r.qa, r.rawlen, m ? P.trim_front2 : P.trim_front1, m ? P.trim_tail2 : P.trim_tail1);
            alive = (c >> 31) != 0u; r.len = (int)(c & 0xFFFu); r.start = (int)((c >> 12) & 0xFFFu);
        }
        if(m){ b = r; a2 = alive; } else { a = r; a1 = alive; }
    }
    V4_GSYNC(X);
    V4_PROF(prof, 0);
    const bool both = a1 && a2;
    if(MODE == 2){
        if(valid) pr = X.outp[i];
        res1 = rec1.result; res2 = rec2.result;
        merged = (pr.flags & FQK_P_MERGED) != 0;
    }else{
    pr.insert_size = 0xFFFF;
    if(both && P.polyg_enabled && !(X.dbg & 32)){
        // src/peprocessor.cpp:297 hands (maxMismatch, oneEach, minLen) to (compareReq, maxMismatch, oneEach)
#pragma unroll 1
        for(int m = 0; m < 2; ++m){
            const uint32_t v = v4_polyg_ni(m ? b.ba : a.ba, m ? b.start : a.start, m ? b.len : a.len,
                                           P.polyg_max_mismatch, P.polyg_one_mismatch_each, P.polyg_min_len);
            if(v >> 31){
                const uint32_t nb = (v >> 12) & 0x1FFFu;
                if(m){ b.len = (int)(v & 0xFFFu); rec2.flags |= FQK_F_POLYG; rec2.polyg_len = (uint16_t)nb; }
                else { a.len = (int)(v & 0xFFFu); rec1.flags |= FQK_F_POLYG; rec1.polyg_len = (uint16_t)nb; }
                tl.polyx_reads[3] += 1; tl.polyx_bases[3] += nb;
            }
        }
    }
    V4_GSYNC(X);
    V4_PROF(prof, 1);
    OvRes ov; ov.overlapped = 0; ov.offset = 0; ov.len = 0; ov.diff = 0;
    // first analyze(): :303 when -a/-c, else :330 (insert size only); identical inputs either way
    if(!(X.dbg & 4)) v4_overlap_chunk(a, b, both, P.overlap_diff_limit, P.overlap_require, lane, ov, X.dbg);
    V4_GSYNC(X);
    V4_PROF(prof, 2);
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
    bool modified = false;                                   // reads changed since the first analyze()
    bool known2 = false;                                     // ... in a way that makes the second analyze()'s answer known
    if(both && P.adapter_trim_enabled){
        // AdapterTrimmer::trimByOverlapAnalysis, src/adaptertrimmer.cpp:14-27
        if(ov.diff <= 5 && ov.overlapped && ov.offset < 0 && ov.len > a.len / 3){
            const int t1 = a.len - ov.len, t2 = b.len - ov.len;
            // Both reads are cut to the overlap.  When the overlap ran to the end of read 2's reverse complement
            // (ov.len == len2 + offset) the cut reads are reverse complements over their whole length: the second analyze()
            // tries offset 0 first (ov.len > overlapRequire holds because loop B only visits offsets > overlapRequire - len2),
            // compares exactly the base pairs the accepted alignment compared, and returns (offset 0, ov.len, same diff).
            known2 = ov.len == b.len + ov.offset;
            a.len = ov.len; b.len = ov.len;
            rec1.flags |= FQK_F_ADAPTER_OV; rec1.adapter_pos = (int16_t)ov.len; rec1.adapter_len = (uint16_t)t1;
            rec2.flags |= FQK_F_ADAPTER_OV; rec2.adapter_pos = (int16_t)ov.len; rec2.adapter_len = (uint16_t)t2;
            tl.adapter_reads += 2; tl.adapter_bases += (uint32_t)(t1 + t2);     // filterresult.cpp:159-161
            modified = true;
        }else if(SP::generic){
            if(P.adapter_r1_len > 0){
                fq_read_result rt = {}; Tally tt = {};
                const int v = v4_adapter_seq(&D, a.qa, a.ba, a.start, a.len, 0, &rt, &tt);
                a.len = v & 0xFFFF; modified |= (v >> 30) != 0; v4_merge_rec(rec1, tl, rt, tt);
            }
            if(P.adapter_r2_len > 0){
                fq_read_result rt = {}; Tally tt = {};
                const int v = v4_adapter_seq(&D, b.qa, b.ba, b.start, b.len, 1, &rt, &tt);
                b.len = v & 0xFFFF; modified |= (v >> 30) != 0; v4_merge_rec(rec2, tl, rt, tt);
            }
        }
    }
    if(SP::generic && both && P.polyx_enabled){
        const int l1 = a.len, l2 = b.len;
#pragma unroll 1
        for(int m = 0; m < 2; ++m){
            fq_read_result rt = {}; Tally tt = {};
            const int nl = v4_polyx(&P, m ? b.qa : a.qa, m ? b.ba : a.ba, m ? b.start : a.start, m ? b.len : a.len, &rt, &tt);
            if(m){ b.len = nl; v4_merge_rec(rec2, tl, rt, tt); } else { a.len = nl; v4_merge_rec(rec1, tl, rt, tt); }
        }
        modified |= (a.len != l1) | (b.len != l2);
    }
    if(SP::generic && both){
        if(P.trim_max_len1 > 0 && P.trim_max_len1 < a.len){ a.len = P.trim_max_len1; modified = true; }
        if(P.trim_max_len2 > 0 && P.trim_max_len2 < b.len){ b.len = P.trim_max_len2; modified = true; }
    }
    V4_GSYNC(X);
    V4_PROF(prof, 3);
    // second analyze(), :354: same answer as the first when neither read changed
    OvRes ov2 = ov;
    if(known2 && a.len == ov.len && b.len == ov.len){        // nothing touched the reads after the overlap trim
        ov2.overlapped = 1; ov2.offset = 0; ov2.len = ov.len; ov2.diff = ov.diff;
        modified = false;
    }
    if(!(X.dbg & 8)) v4_overlap_chunk(a, b, P.merge_enabled && both && modified, P.overlap_diff_limit, P.overlap_require, lane, ov2);
    V4_GSYNC(X);
    V4_PROF(prof, 4);
    bool merge_processed = false;
    int fcA = -1, fnA = 0, fcB = -1, fnB = 0;                // updates of mFilterReadStats: (code, amount) x 2
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
                result = v4_merged_filter<SP>(&P, a.qa, a.ba, a.start, a.len, b.qa, b.ba, b.start, b.len, ov2.len, len1, len2);
                if(result == FQK_PASS_FILTER){ pr.flags |= FQK_P_MERGED; merged = true; }
            }
            fcA = result; fnA = 2;                                  // :358
            pr.merged_result = (uint8_t)result;
            res1 = res2 = result;
            merge_processed = true;
        }else if(!P.merge_discard_unmerged){
            res1 = v4_pass_filter<SP>(P, a, false); fcA = res1; fnA = 1;
            res2 = v4_pass_filter<SP>(P, b, false); fcB = res2; fnB = 1;
            merge_processed = true;
        }
        if(merge_processed) pr.flags |= FQK_P_MERGE_PROC;
    }
    if(valid && !merge_processed){
        res1 = v4_pass_filter<SP>(P, a, !a1);
        res2 = v4_pass_filter<SP>(P, b, !a2);
        fcA = max(res1, res2); fnA = 2;                             // :390, mPaired == true
    }
    V4_GSYNC(X);
    V4_PROF(prof, 5);
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
    V4_GSYNC(X);
    V4_PROF(prof, 6);
    }   // MODE != 2
    if(MODE == 1) return;
    // ---- what the statistics warps need: length histograms, exception lists, short tails ----
    const uint32_t cnt32 = X.S32 + L.o_cnt;
#pragma unroll
    for(int m = 0; m < 2; ++m){
        const PeRead& Rm = m ? b : a;
        const fq_read_result& rm = m ? rec2 : rec1;
        const uint32_t C = m ? L.Cr : L.Cm;                       // cycles of post1 / post2
        const uint32_t dlen = (uint32_t)Rm.rawlen;
        bool own = false; uint32_t ostart = SP::generic ? rm.start : 0u, olen = rm.len, hlen = rm.len;
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
            if(!(X.dbg & 16) && __any_sync(V4_FULL, td)) v4_tail_sub(td, Rm.qa, Rm.ba, Rm.rawlen, (int)olen, X.post32[m], C, lane);
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
    V4_PROF(prof, 7);
    V4_PROFC(prof, 15);
}

#ifndef FQK_EMU
// ====================================================================== the kernel
constexpr int V4_STATS = 256, V4_TRIM = 256, V4_THREADS = V4_STATS + V4_TRIM;
constexpr int V4_WORKERS = V4_TRIM / 32 - 1;                      // trim warps 0..6 work on chunks, warp 7 is the copy producer
constexpr int V4_WORKERS_SPLIT = V4_THREADS / 32 - 1;             // trim kernel of the split path: 15 working warps + the producer
static_assert(sizeof(V4L) <= 256, "V4L lives in the first 256 bytes of the block's shared memory");

// accumulators of the quad path -> the two statistics arrays of a mate (different cycle counts: post1 holds merged reads)
__device__ __noinline__ void quad_drain1_ab(const uint32_t* v, uint32_t* smA, int CA, uint32_t* smB, int CB, int w){
    auto add2 = [&](int kind, int set, int c, uint32_t x){
        if(x){ atomicAdd(sm_cell(smA, CA, kind, set) + c, x); atomicAdd(sm_cell(smB, CB, kind, set) + c, x); }
    };
    for(int j = 0; j < 4; ++j){
        const int c = 4 * w + j;
        add2(KN20, S_ALL, c, v[0 + j] >> 7);
        add2(KN30, S_ALL, c, v[4 + j] >> 6);
        add2(KQS, S_ALL, c, v[8 + j]);
        for(int s = 0; s < 3; ++s){
            const int sh = 2 * j + (s == 1 ? 1 : 0);
            add2(KN, S_B0 + s, c, v[12 + s * 4 + j] >> sh);
            add2(KN20, S_B0 + s, c, v[24 + s * 4 + j] >> (7 + sh));
            add2(KN30, S_B0 + s, c, v[36 + s * 4 + j] >> (6 + sh));
            add2(KQS, S_B0 + s, c, v[48 + s * 4 + j] >> sh);
        }
        add2(KN, S_N, c, v[60 + j]);
        add2(KN20, S_N, c, v[64 + j] >> 7);
        add2(KN30, S_N, c, v[68 + j] >> 6);
        add2(KQS, S_N, c, v[72 + j]);
    }
}
__device__ __forceinline__ void quad_drain_ab(QuadAcc& A, uint32_t* smA, int CA, uint32_t* smB, int CB, int w){
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
    quad_drain1_ab(v, smA, CA, smB, CB, w);
    A.zero();
}

// Named barriers: ids 1..NS tile ready (CH trim-warp arrivals + the statistics warps), NS+1..2NS stage free (the
// statistics warps arrive, the producer warp waits); `full` (bulk copies landed) is an mbarrier with a transaction count.
// MODE 0: fused (8 statistics warps + 7 trim warps + producer).  The split path runs the same code as two kernels:
// MODE 1, the trim kernel: no statistics warps and no statistics arrays, 15 trim warps + producer, the shared memory
// goes to a deeper ring of tiles; stages are handed back to the producer through `empty` mbarriers.  MODE 2, the
// statistics kernel: the trim warps only rebuild the exception lists from the records, everything else as MODE 0.
// Why two kernels: the trim role is latency bound (one warp per scheduler and a half, dependent integer chains,
// profiles/r02n_*) and the register file of an SM cannot hold more than 7 trim warps next to 8 statistics warps of
// 168 registers; reading the batch twice costs 0.3 ms of HBM time per 4 M pairs, the idle issue slots cost 3 ms.
template<class SP, int MODE>
__global__ void __launch_bounds__(V4_THREADS, 1)
k_pe_v4(const __grid_constant__ DevParams D, DevBatch B1, DevBatch B2, fq_read_result* __restrict__ out1,
        fq_read_result* __restrict__ out2, fq_pair_result* __restrict__ outp, uint64_t* __restrict__ flat,
        int Wr, int Wm, int T, int NS, int pairs_per_block, int dbg){
    extern __shared__ __align__(128) uint8_t smraw[];
    const fq_params& P = D.P;
    constexpr int NSTAT = MODE == 1 ? 0 : V4_STATS;               // statistics threads
    constexpr int NWORK = (V4_THREADS - NSTAT) / 32 - 1;          // working trim warps (the last trim warp is the copy producer)
    if(threadIdx.x == 0) *reinterpret_cast<V4L*>(smraw) = v4_layout(Wr, Wm, T, NS, P.insert_size_max + 1, MODE != 1);
    __syncthreads();
    const V4L& L = *reinterpret_cast<const V4L*>(smraw);
    const uint32_t sm32 = smem_u32(smraw);
    uint32_t* const smPost1 = reinterpret_cast<uint32_t*>(smraw + L.o_post[0]);
    BlockCounters* bc = reinterpret_cast<BlockCounters*>(smraw + L.o_bc);
    uint32_t* ins = reinterpret_cast<uint32_t*>(smraw + L.o_ins);
    uint64_t* full = reinterpret_cast<uint64_t*>(smraw + L.o_bar);
    uint64_t* empty = full + NS;                                 // MODE 1: all chunks of the staged tile are done
    const int CH = T / 32;                                       // chunks per tile (T is a multiple of 32)

    for(uint32_t i = L.o_pre[0] / 4 + threadIdx.x; i < L.o_bar / 4; i += V4_THREADS) reinterpret_cast<uint32_t*>(smraw)[i] = 0;   // statistics, counters, insert sizes
    if(threadIdx.x == 0){
        *reinterpret_cast<uint32_t*>(smraw + L.o_misc) = 0u;         // next chunk of the trim warps
        for(int s = 0; s < NS; ++s){
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], (uint32_t)CH);
            *reinterpret_cast<uint32_t*>(smraw + L.o_stage + (size_t)s * L.stage_bytes + L.o_cnt + 20) = 0xFFFFFFFFu;   // tile tag: nothing staged yet
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const uint32_t lo = (uint32_t)min((uint64_t)B1.n, (uint64_t)blockIdx.x * pairs_per_block);
    const uint32_t hi = (uint32_t)min((uint64_t)B1.n, (uint64_t)lo + pairs_per_block);
    const int ntiles = (int)((hi - lo + T - 1) / T);
    const int lane = threadIdx.x & 31;

    const bool is_trim = MODE == 1 || threadIdx.x >= NSTAT;
    const int rtid = is_trim ? threadIdx.x - NSTAT : threadIdx.x;
    if(is_trim){
        if(MODE != 1) asm volatile("setmaxnreg.dec.sync.aligned.u32 88;\n");
        const int tw = rtid >> 5;
        if(tw == NWORK){
            // ============================== copy producer ==============================
            auto BM = [&](int m) -> const DevBatch& { return m ? B2 : B1; };
            uint32_t a[2] = {0, 0}, last[2] = {0, 0}, ll[2] = {0, 0};
            if(ntiles > 0){
                const uint32_t nt = min((uint32_t)T, hi - lo);
#pragma unroll
                for(int m = 0; m < 2; ++m){ a[m] = BM(m).offsets[lo]; last[m] = BM(m).offsets[lo + nt - 1]; ll[m] = BM(m).lengths[lo + nt - 1]; }
            }
            for(int k = 0; k < ntiles; ++k){
                const int s = k % NS, n = k / NS;
                const uint32_t t0 = lo + (uint32_t)k * T;
                const uint32_t nt = min((uint32_t)T, hi - t0);
                const uint32_t ca[2] = {a[0], a[1]}, clast[2] = {last[0], last[1]}, cll[2] = {ll[0], ll[1]};
                if(k + 1 < ntiles){                                // next tile's end points: in flight during the wait below
                    const uint32_t t1 = t0 + T, nt1 = min((uint32_t)T, hi - t1);
#pragma unroll
                    for(int m = 0; m < 2; ++m){ a[m] = BM(m).offsets[t1]; last[m] = BM(m).offsets[t1 + nt1 - 1]; ll[m] = BM(m).lengths[t1 + nt1 - 1]; }
                }
                if(n >= 1){
                    if(MODE == 1) mbar_wait_lane0(&empty[s], (uint32_t)(n - 1) & 1u, 20000);
                    else nb_sync(1 + NS + s, NSTAT + 32);
                }
                uint8_t* S = smraw + L.o_stage + (size_t)s * L.stage_bytes;
                const uint32_t nb = nt & ~7u;                      // bulk copies move multiples of 16 bytes
#pragma unroll
                for(int m = 0; m < 2; ++m){
                    uint32_t* So = reinterpret_cast<uint32_t*>(S + L.o_off[m]);
                    uint16_t* Sl = reinterpret_cast<uint16_t*>(S + L.o_len[m]);
                    for(uint32_t i = nb + lane; i < nt; i += 32){ So[i] = BM(m).offsets[t0 + i]; Sl[i] = BM(m).lengths[t0 + i]; }
                    for(uint32_t i = nt + lane; i < (uint32_t)T; i += 32){ So[i] = ca[m]; Sl[i] = 0; }
                }
                __syncwarp();
                if(lane == 0){
                    uint32_t* cnt = reinterpret_cast<uint32_t*>(S + L.o_cnt);
                    uint32_t qbytes[2], b0[2], bbytes[2], uni = 0; bool bad = false;
#pragma unroll
                    for(int m = 0; m < 2; ++m){
                        uint64_t e = (uint64_t)clast[m] + ((cll[m] + 15u) & ~15u);
                        const bool bd = (ca[m] & 15u) || e < ca[m] || e - ca[m] > L.qcap || e > BM(m).total;
                        if(bd){ bad = true; e = ca[m]; }
                        qbytes[m] = (uint32_t)(e - ca[m]);
                        b0[m] = (ca[m] >> 2) & ~15u;
                        bbytes[m] = qbytes[m] ? (((uint32_t)(e >> 2) + 15u) & ~15u) - b0[m] : 0u;
                        if(nt == (uint32_t)T && qbytes[m] == L.qcap) uni |= 1u << m;                // every read exactly Cp padded bytes
                    }
                    if(bad){ atomicOr(D.err, 2u); qbytes[0] = qbytes[1] = bbytes[0] = bbytes[1] = 0; uni = 0; }
                    cnt[0] = 0; cnt[1] = 0; cnt[2] = 0; cnt[3] = uni; cnt[4] = bad ? 0u : nt;
                    *reinterpret_cast<volatile uint32_t*>(&cnt[5]) = (uint32_t)k;   // tile tag (wait_tile_tag)
                    mbar_expect_tx(&full[s], qbytes[0] + bbytes[0] + qbytes[1] + bbytes[1] + 2u * nb * 6u);
#pragma unroll
                    for(int m = 0; m < 2; ++m){
                        if(qbytes[m]){
                            bulk_g2s(S + L.o_q[m], BM(m).qual + ca[m], qbytes[m], &full[s]);
                            bulk_g2s(S + L.o_b[m], BM(m).bases2 + b0[m], bbytes[m], &full[s]);
                        }
                        if(nb){
                            bulk_g2s(S + L.o_off[m], BM(m).offsets + t0, nb * 4u, &full[s]);
                            bulk_g2s(S + L.o_len[m], BM(m).lengths + t0, nb * 2u, &full[s]);
                        }
                    }
                }
            }
        }else{
            // ============================== trim warps ==============================
            Tally tl = {};
            V4X X;
            X.D = &D; X.Lp = &L;
            X.pre32[0] = sm32 + L.o_pre[0]; X.pre32[1] = sm32 + L.o_pre[1];
            X.post32[0] = sm32 + L.o_post[0]; X.post32[1] = sm32 + L.o_post[1];
            X.ins32 = sm32 + L.o_ins; X.bc = bc; X.dbg = dbg;
            const int nchunks = ntiles * CH;
            // chunks are handed out in order from a shared counter: a chunk that takes long (many candidates to verify)
            // does not hold back the chunks behind it, and at most V4_WORKERS consecutive chunks are in flight
            const uint32_t next32 = sm32 + L.o_misc;
            const int idle_from = NWORK - ((dbg >> 12) & 15);              // timing experiment: fewer working warps
            // Tile groups: the CH warps of a group work on the CH chunks of one tile and meet at a named barrier between the
            // sections of v4_chunk, so that they run the same code at about the same time.  The trim role walks ~48 KB of
            // code per chunk, more than the SM's instruction cache holds; unsynchronised, every warp takes its own misses
            // and the role does not speed up with more warps (3 warps 8.4 ms, 7 warps 6.5 ms on cfg3).  In step, one miss
            // serves the group: 6.7 -> 5.1 ms with two groups of three (profiles/r02_pe_v4_notes.md).  Groups take tiles
            // round robin.  FQK_V4_DEBUG bit 512 switches back to chunks handed out one by one from a shared counter.
            // Groups need a spare stage to prefetch into (NS > groups); without one the hand-out from the counter is as fast.
            // The trim kernel (MODE 1) has no ready/free named barriers, so group g syncs on id 1 + g, and it runs as many
            // groups as there are stages to work on while one is being refilled.
            const bool gs = !(dbg & 512) && CH >= 2 && CH <= NWORK && (MODE == 1 ? NS >= 2 : NS > NWORK / CH);
            const int NG = gs ? (MODE == 1 ? min(min(NWORK / CH, NS - 1), 14) : NWORK / CH) : 1, grp = tw / CH, jst = tw - grp * CH;
            X.sync_id = gs ? (MODE == 1 ? 1 + grp : 2 * NS + 1 + grp) : 0; X.sync_n = CH * 32;
            int kk = grp;
            for(;;){
                if(tw >= idle_from) break;
                int k, j;
                if(gs){
                    if(grp >= NG || kk >= ntiles) break;
                    k = kk; j = jst; kk += NG;
                }else{
                    int c = 0;
                    if(lane == 0) c = (int)atoms_add(next32, 1u);
                    c = __shfl_sync(V4_FULL, c, 0);
                    if(c >= nchunks) break;
                    k = c / CH; j = c - k * CH;
                }
                const int s = k % NS;
                X.S32 = sm32 + L.o_stage + (uint32_t)s * L.stage_bytes;
                V4_PROF0(dbg & 256);
                wait_tile_tag(X.S32 + L.o_cnt + 20, (uint32_t)k);
                mbar_wait_sleep(&full[s], (uint32_t)(k / NS) & 1u, 20000);
                X.nt = (int)lds32(X.S32 + L.o_cnt + 16);
                V4_PROF(dbg & 256, 10);
                const uint32_t t0 = lo + (uint32_t)k * T;
                X.out1 = out1 + t0; X.out2 = out2 + t0; X.outp = outp + t0;
                if((gs || 32 * j < X.nt) && !(dbg & 1)) v4_chunk<SP, MODE>(X, j, lane, tl);
                V4_PROF(dbg & 256, 11);
                __syncwarp();
                __threadfence_block();
                if(MODE == 1){ if(lane == 0 && k + NS < ntiles) mbar_arrive(&empty[s]); }   // the producer waits only for stages it refills
                else nb_arrive(1 + s, CH * 32 + NSTAT);
            }
            if(MODE != 2) counters_add_tally(bc, tl);
        }
    }else{
        // ============================ statistics warps ============================
        if(MODE != 1) asm volatile("setmaxnreg.inc.sync.aligned.u32 168;\n");
        const int C1 = (int)L.Cr, CM = (int)L.Cm, Cp = (int)L.Cp;
        const int G = V4_STATS / Wr, Gm = G >> 1;                // groups of Wr threads; even groups read1, odd groups read2
        const int g = rtid / Wr, w = rtid - g * Wr;
        const bool active = g < 2 * Gm;
        const int mate = g & 1, gi = g >> 1;
        uint32_t* smPre = reinterpret_cast<uint32_t*>(smraw + (mate ? L.o_pre[1] : L.o_pre[0]));
        uint32_t* smPost = reinterpret_cast<uint32_t*>(smraw + (mate ? L.o_post[1] : L.o_post[0]));
        const int Cpost = mate ? C1 : CM;
        // the tails of merged reads land on post1 at cycles up to Cm: a second mapping, Wm threads per entry
        const int GM = V4_STATS / Wm, gM = rtid / Wm, wM = rtid - gM * Wm;
        const bool activeM = P.merge_enabled && gM < GM;
        QuadAcc qa;
        Acc1 dsub, madd;
        qa.zero(); dsub.zero(); madd.zero();
        int usedQ = 0, leftO = 255, leftM = 255;
        const int quads_per_tile = T / 4;
        for(int k = 0; k < ntiles; ++k){
            const int s = k % NS;
            const uint32_t S32 = sm32 + L.o_stage + (uint32_t)s * L.stage_bytes;
            nb_sync(1 + s, CH * 32 + V4_STATS);
            if(!(dbg & 2)){
                const int nt = (int)lds32(S32 + L.o_cnt + 16);
                if(active){
                    const uint32_t a0 = lds32(S32 + L.o_off[mate]);
                    const uint32_t bshift = (a0 >> 2) & 15u;
                    const uint32_t sq0 = S32 + L.o_q[mate], sb0 = S32 + L.o_b[mate] + bshift;
                    const uint32_t sqw = sq0 + 4 * w, sbw = sb0 + w;
                    const int nex = (int)lds32(S32 + L.o_cnt + 4 * mate), uniform = (int)((lds32(S32 + L.o_cnt + 12) >> mate) & 1u);
                    const int mine = (quads_per_tile + Gm - 1) / Gm;
                    if(usedQ + mine > QUAD_LIMIT){ quad_drain_ab(qa, smPre, C1, smPost, Cpost, w); usedQ = 0; }
                    usedQ += mine;
                    uint32_t q[4], b[4];
                    if(uniform){
                        // every read of the tile occupies exactly Cp bytes: addresses advance by constants, no range checks
                        const uint32_t qstep = 4u * (uint32_t)Gm * (uint32_t)Cp, bstep = (uint32_t)Gm * (uint32_t)Cp;
                        uint32_t aq = sqw + 4u * (uint32_t)gi * (uint32_t)Cp, ab = sbw + (uint32_t)gi * (uint32_t)Cp;
                        const uint32_t c1 = (uint32_t)Cp, c1b = (uint32_t)Cp >> 2;
                        for(int qd = gi; qd < quads_per_tile; qd += Gm){
                            q[0] = lds32(aq); q[1] = lds32(aq + c1); q[2] = lds32(aq + 2 * c1); q[3] = lds32(aq + 3 * c1);
                            b[0] = lds8(ab); b[1] = lds8(ab + c1b); b[2] = lds8(ab + 2 * c1b); b[3] = lds8(ab + 3 * c1b);
                            aq += qstep; ab += bstep;
                            quad_accumulate(qa, q[0], q[1], q[2], q[3], b[0], b[1], b[2], b[3]);
                        }
                    }else{
                        const int nq = (nt + 3) >> 2;
                        for(int qd = gi; qd < nq; qd += Gm){
                            const uint4 o = lds128(S32 + L.o_off[mate] + 16u * qd);
                            const uint2 l2 = lds64(S32 + L.o_len[mate] + 8u * qd);
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
                    // ---- exceptions of this mate: D is subtracted from its post array, Z is added to it ----
                    const uint32_t ex = S32 + L.o_ex[mate];
                    const uint32_t c0 = 4u * (uint32_t)w;
                    for(int i = gi; i < nex; i += Gm){
                        const uint2 dc = lds64(ex + 8 * i);
                        const uint32_t off = (dc.x & 0x1FFFFu) << 4, len = dc.x >> 17, kind = (dc.y >> 26) & 3u;
                        const uint32_t start = dc.y & 0x1FFFu, plen = (dc.y >> 13) & 0x1FFFu;
                        const uint32_t d0 = (kind == 2u && start == 0u) ? plen : 0u;              // D = raw positions [d0, len)
                        if(c0 < len && c0 + 4 > d0){
                            uint32_t qq = lds32(sqw + off), bb = lds8(sbw + (off >> 2));
                            if(c0 < d0){ const uint32_t nbz = d0 - c0; qq &= 0xFFFFFFFFu << (8 * nbz); bb &= 0xFFu << (2 * nbz); }
                            if(qq) acc_word(dsub, qq, bb);             // a word inside a read never has four zero quality bytes
                        }
                        if(kind == 2u && start > 0u){
                            const int nv = (int)plen - 4 * w;
                            if(nv > 0){
                                const uint32_t p = off + start + 4 * w;
                                const uint32_t a1 = p & ~3u, sh = p & 3u;
                                const uint32_t w0 = lds32(sq0 + a1);
                                const uint32_t w1 = sh ? lds32(sq0 + a1 + 4) : 0u;
                                const uint32_t q2 = __funnelshift_r(w0, w1, sh * 8);
                                const uint32_t e0 = lds8(sb0 + (p >> 2)), e1 = sh ? lds8(sb0 + (p >> 2) + 1) : 0u;
                                const uint32_t b2 = ((e0 | (e1 << 8)) >> (2 * sh)) & 0xFF;
                                z_word_atomic(q2, b2, nv, smPost, Cpost, w);
                            }
                        }
                        if(--leftO == 0){ drain_l1_signed(dsub, smPost, Cpost, w, 1); leftO = 255; }
                    }
                }
                if(activeM){
                    // ---- M: reverse-complemented tails of read2 appended to merged reads (post1, cycles [len1, len1+len2)) ----
                    const uint32_t a0 = lds32(S32 + L.o_off[1]);
                    const uint32_t sq0 = S32 + L.o_q[1], sb0 = S32 + L.o_b[1] + ((a0 >> 2) & 15u);
                    const int nm = (int)lds32(S32 + L.o_cnt + 8);
                    const uint32_t exm = S32 + L.o_exm;
                    for(int i = gM; i < nm; i += GM){
                        const uint2 dc = lds64(exm + 8 * i);
                        const uint32_t off = (dc.x & 0x1FFFFu) << 4;
                        const int pmax = (int)(dc.x >> 17), len1 = (int)(dc.y & 0x1FFFu), len2 = (int)((dc.y >> 13) & 0x1FFFu);
                        const int c0 = 4 * wM, lo2 = max(len1, c0), hi2 = min(len1 + len2, c0 + 4);
                        if(lo2 < hi2){
                            const int addr = (int)off + pmax - c0 - 3;                  // tile position of the byte for cycle c0+3
                            const int a1 = addr & ~3, sh = addr & 3;
                            const uint32_t q4 = __funnelshift_r(lds32(sq0 + a1), lds32(sq0 + a1 + 4), sh * 8);
                            uint32_t q2 = __byte_perm(q4, 0, 0x0123);                  // reversed: byte j <-> cycle c0+j
                            const uint32_t cb = lds8(sb0 + (addr >> 2)) | (lds8(sb0 + (addr >> 2) + 1) << 8);
                            uint32_t b2 = (cb >> (2 * sh)) & 0xFF;
                            b2 = ((b2 & 0x03) << 6) | ((b2 & 0x0C) << 2) | ((b2 & 0x30) >> 2) | ((b2 & 0xC0) >> 6);
                            b2 ^= 0x55;                                                 // complement: A<->T, C<->G
                            const int jlo = lo2 - c0, jhi = hi2 - c0;
                            const uint32_t bm = (jhi >= 4 ? 0xFFFFFFFFu : ((1u << (8 * jhi)) - 1u)) & ~((1u << (8 * jlo)) - 1u);
                            q2 &= bm;
                            const uint32_t f = ((((q2 >> 7) & 0x01010101u) * 0x00204081u) >> 21) & 0xFu;   // N flags of the 4 bytes
                            uint32_t spn = (f & 1u) | ((f & 2u) << 1) | ((f & 4u) << 2) | ((f & 8u) << 3);
                            spn |= spn << 1;
                            b2 &= ((1u << (2 * jhi)) - 1u) & ~((1u << (2 * jlo)) - 1u) & ~spn;   // N keeps code 0
                            acc_word(madd, q2, b2);
                        }
                        if(--leftM == 0){ drain_l1(madd, smPost1, nullptr, CM, wM); leftM = 255; }
                    }
                }
            }
            __syncwarp();
            if(k + NS < ntiles) nb_arrive(1 + NS + s, V4_STATS + 32);      // the producer waits only for stages it refills
        }
        if(active){
            quad_drain_ab(qa, smPre, C1, smPost, Cpost, w);
            drain_l1_signed(dsub, smPost, Cpost, w, 1);
        }
        if(activeM) drain_l1(madd, smPost1, nullptr, CM, wM);
    }
    __syncthreads();
    if(MODE != 1){
        retire(reinterpret_cast<uint32_t*>(smraw + L.o_pre[0]), (int)L.Cr, D.L, flat + D.L.stats_off[0]);
        retire(reinterpret_cast<uint32_t*>(smraw + L.o_pre[1]), (int)L.Cr, D.L, flat + D.L.stats_off[1]);
        retire(smPost1, (int)L.Cm, D.L, flat + D.L.stats_off[2]);
        retire(reinterpret_cast<uint32_t*>(smraw + L.o_post[1]), (int)L.Cr, D.L, flat + D.L.stats_off[3]);
    }
    if(MODE != 2){                                               // run counters and insert sizes belong to the kernel that trims
        counters_flush(bc, D.L, flat);
        for(int i = threadIdx.x; i <= P.insert_size_max; i += V4_THREADS) if(ins[i]) add_u64(flat + D.L.insert_hist_off + i, ins[i]);
    }
}

// ---------------------------------------------------------------------- host side
struct V4Plan { int Wr, Wm, T, NS; size_t smem; bool ok; };
// tile size and stage count for reads up to maxlen bases (0 = unknown: the capacity max_cycles)
inline V4Plan v4_plan(const DevParams& D, int maxlen, size_t smem_optin, int mode = 0){
    V4Plan p = {0, 0, 0, 0, 0, false};
    const fq_params& P = D.P;
    const int cap = (int)D.L.max_cycles;
    if(maxlen <= 0 || maxlen > cap) maxlen = cap;
    if(maxlen > V4_MAXLEN) return p;
    if(P.correction_enabled || P.kmer_len) return p;         // -c mutates the batch, k-mers: the two-phase path
    p.Wr = std::max(8, (maxlen + 3) / 4);
    p.Wm = P.merge_enabled ? std::max(p.Wr, std::min((cap + 3) / 4, (2 * maxlen + 3) / 4)) : p.Wr;
    if(p.Wm > V4_STATS || V4_STATS / p.Wr < 2) return p;
    const size_t budget = smem_optin > 4096 ? smem_optin - 1024 : 0;
    const int Gm = (V4_STATS / p.Wr) >> 1;
    int bestT = 0, bestNS = 0; long best = -1;
    const bool st = mode != 1;
    for(int NS = (st ? 4 : 12); NS >= 2; --NS)
        for(int T = 256; T >= 32; T -= 32){
            const V4L l = v4_layout(p.Wr, p.Wm, T, NS, P.insert_size_max + 1, st);
            if(l.qcap >= (1u << 21) || (size_t)l.total + 128 > budget) continue;
            long score;
            if(st) score = (long)T * NS + (((T / 4) % Gm == 0) ? 100000 : 0) + (NS >= 3 ? 50000 : 0);   // whole quads per statistics group, then bytes in flight
            else{
                // trim kernel: as many working warps as possible, in groups that walk the code together (CH >= 2)
                const int CH = T / 32;
                const bool gs = CH >= 2 && CH <= V4_WORKERS_SPLIT;
                const int busy = gs ? std::min(std::min(V4_WORKERS_SPLIT / CH, NS - 1), 14) * CH : std::min(V4_WORKERS_SPLIT, (NS - 1) * CH);
                // measured on cfg3 (4 M pairs, trim + statistics kernel): T=64 NS=8 (7 groups of 2) 5.65 ms, T=96 NS=5 (4 x 3)
                // 5.27 ms, T=128 NS=4 (3 x 4) 5.24 ms: warps that walk the code together count for more than warps
                score = (long)busy * 1000 + (gs ? (long)std::min(CH, 4) * 1200 : 0) + NS;
            }
            if(score > best){ best = score; bestT = T; bestNS = NS; }
        }
    if(best < 0) return p;
    static const int envT = fqk_tuning_env("FQK_V4_T") ? atoi(fqk_tuning_env("FQK_V4_T")) : 0, envNS = fqk_tuning_env("FQK_V4_NS") ? atoi(fqk_tuning_env("FQK_V4_NS")) : 0;
    if(mode == (fqk_tuning_env("FQK_V4_TMODE") ? atoi(fqk_tuning_env("FQK_V4_TMODE")) : 0) && envT > 0 && envNS >= 2 && envNS <= (st ? 4 : 12) && envT % 32 == 0 &&
       (size_t)v4_layout(p.Wr, p.Wm, envT, envNS, P.insert_size_max + 1, st).total + 128 <= budget){ bestT = envT; bestNS = envNS; }   // timing experiments
    p.T = bestT; p.NS = bestNS;
    p.smem = v4_layout(p.Wr, p.Wm, p.T, p.NS, P.insert_size_max + 1, st).total + 128;
    p.ok = true;
    return p;
}
#endif  // !FQK_EMU
