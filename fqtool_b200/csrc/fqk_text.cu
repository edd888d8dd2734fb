// fqk_text.cu -- device-side output compaction and FASTQ record assembly.
//
// What the reference does per read on the host after the hot path: Read::toString / toStringWithTag
// (/root/reference/src/read.h:166-176), the routing ladders of SingleEndProcessor::processSingleEnd
// (src/seprocessor.cpp:340-349) and PairEndProcessor::processPairEnd (src/peprocessor.cpp:353-429) and the assembly
// of merged reads (OverlapAnalysis::merge, src/overlapanalysis.cpp:74-104, with its name quirk).  Here the records
// of the trim kernels and the FASTQ characters already staged on the device become, per output channel, ONE
// contiguous block of FASTQ text in input order:
//   k_text_sizes   a thread per read (pair): the routing ladder -> bytes per channel
//   k_scan_*       exclusive prefix sums per channel (three small kernels, 2048 items per block)
//   k_text_copy    a warp per read (pair): the same ladder -> the pieces, copied with aligned word stores (funnel-shifted source words)
// so that the host's writer threads only hand blocks to gzwrite; no per-read string is built on the host.
#include "fqk_text.h"

namespace {

// failure tags, /root/reference/src/common.h:21-30; index 32 = "paired_read_is_failing" (src/peprocessor.cpp:409,424)
__constant__ char c_tags[33][24] = {
    "passed", "", "", "", "failed_polyx_filter", "", "", "", "failed_bad_overlap", "", "", "",
    "failed_too_many_n_bases", "", "", "", "failed_too_short", "failed_too_long", "", "",
    "failed_quality_filter", "", "", "", "failed_low_complexity", "", "", "", "", "", "", "", "paired_read_is_failing"};
__device__ __forceinline__ uint32_t tag_len(int t){ uint32_t n = 0; while(c_tags[t][n]) ++n; return n; }
__device__ __forceinline__ uint32_t dec_digits(uint32_t v){ uint32_t n = 1; while(v >= 10){ v /= 10; ++n; } return n; }

struct Piece { int ch; int kind; int tag; };          // kind 0 read1, 1 read2, 2 merged read; tag -1 none
constexpr int TAG_PAIR = 32;

// the if/else ladder of src/peprocessor.cpp:353-429 (and fqtool_b200/host/fqk_host.cpp) -> at most two pieces
__device__ __forceinline__ int route_pe(const fq_read_result& x1, const fq_read_result& x2, const fq_pair_result& pr, uint32_t writers,
                                        Piece* pc, bool& passed){
    const bool a1 = !(x1.flags & FQK_F_DROPPED), a2 = !(x2.flags & FQK_F_DROPPED);
    const int r1 = x1.result, r2 = x2.result;
    const bool wf = writers & FQK_W_FAILED, wu = writers & FQK_W_UNPAIRED1;
    int n = 0; passed = false;
    if(pr.flags & FQK_P_MERGE_PROC){
        if(pr.flags & FQK_P_OV2_OVERLAP){
            if((pr.flags & FQK_P_MERGED) && pr.ov2_len){ pc[n++] = {FQK_CH_MERGED, 2, -1}; passed = true; }
        }else{
            if(r1 == FQK_PASS_FILTER) pc[n++] = {FQK_CH_MERGED, 0, -1};
            if(r2 == FQK_PASS_FILTER) pc[n++] = {FQK_CH_MERGED, 1, -1};
            passed = r1 == FQK_PASS_FILTER && r2 == FQK_PASS_FILTER;
        }
    }else if(a1 && r1 == FQK_PASS_FILTER && a2 && r2 == FQK_PASS_FILTER){
        if(writers & FQK_W_STDOUT){ pc[n++] = {FQK_CH_SINGLE, 0, -1}; pc[n++] = {FQK_CH_SINGLE, 1, -1}; }
        else { pc[n++] = {FQK_CH_OUT1, 0, -1}; pc[n++] = {FQK_CH_OUT2, 1, -1}; }
        passed = true;
    }else if(a1 && r1 == FQK_PASS_FILTER){
        if(wu){ pc[n++] = {FQK_CH_UNPAIRED1, 0, -1}; if(wf) pc[n++] = {FQK_CH_FAILED, 1, r2}; }
        else if(wf){ pc[n++] = {FQK_CH_FAILED, 0, TAG_PAIR}; pc[n++] = {FQK_CH_FAILED, 1, r2}; }
    }else if(a2 && r2 == FQK_PASS_FILTER){
        // as written in the reference: the LEFT unpaired writer is tested and read1 is tagged with result2
        if(wu){ pc[n++] = {FQK_CH_UNPAIRED2, 1, -1}; if(wf) pc[n++] = {FQK_CH_FAILED, 0, r2}; }
        else if(wf){ pc[n++] = {FQK_CH_FAILED, 0, r1}; pc[n++] = {FQK_CH_FAILED, 1, TAG_PAIR}; }
    }
    return n;
}
__device__ __forceinline__ int route_se(const fq_read_result& x, uint32_t writers, Piece* pc, bool& passed){
    passed = !(x.flags & FQK_F_DROPPED) && x.result == FQK_PASS_FILTER;
    if(passed){ pc[0] = {FQK_CH_OUT1, 0, -1}; return 1; }
    if(writers & FQK_W_FAILED){ pc[0] = {FQK_CH_FAILED, 0, (int)x.result}; return 1; }
    return 0;
}

// the part of a read that its Read object holds after the trimming chain: the view, or everything when trimAndCut dropped it
__device__ __forceinline__ void view_of(const fq_read_result& x, uint32_t rawlen, uint32_t& start, uint32_t& len){
    if(x.flags & FQK_F_DROPPED){ start = 0; len = rawlen; } else { start = x.start; len = x.len; }
}
// position of the first ' ' of a name (or the length when there is none)
__device__ __forceinline__ uint32_t first_space(const char* s, uint32_t n){ uint32_t p = 0; while(p < n && s[p] != ' ') ++p; return p; }

// bytes of the merged read's name: OverlapAnalysis::merge builds "_merged_<len1>_<len2>" and splices it before the first
// space of read1's name MINUS ONE character (substr(0, pos - 1)); without a space the name itself is lost
__device__ __forceinline__ uint32_t merged_name_len(const char* name, uint32_t nl, uint32_t len1, uint32_t len2){
    const uint32_t tagl = 8 + dec_digits(len1) + 1 + dec_digits(len2);
    const uint32_t sp = first_space(name, nl);
    if(sp >= nl) return tagl;
    // a name that BEGINS with a space: substr(0, pos - 1) is substr(0, npos), the whole name, so the name appears twice
    return (sp ? sp - 1 : nl) + tagl + (nl - sp);
}

__device__ __forceinline__ uint32_t piece_bytes(const Piece& p, const TextMate& m1, const TextMate& m2, uint32_t i,
                                                const fq_read_result& x1, const fq_read_result& x2, const fq_pair_result& pr){
    if(p.kind == 2){
        const uint32_t nl = m1.nlen(i), sl = m1.slen(i);
        const uint32_t L = (uint32_t)pr.merge_len1 + pr.merge_len2;
        return merged_name_len(m1.names + m1.name_off[i], nl, pr.merge_len1, pr.merge_len2) + 1 + L + 1 + sl + 1 + L + 1;
    }
    const TextMate& m = p.kind ? m2 : m1;
    const fq_read_result& x = p.kind ? x2 : x1;
    uint32_t st, ln; view_of(x, m.len[i], st, ln);
    const uint32_t nl = m.nlen(i), sl = m.slen(i);
    return nl + (p.tag >= 0 ? 1 + tag_len(p.tag) : 0) + 1 + ln + 1 + sl + 1 + ln + 1;
}

__global__ void k_text_sizes(bool paired, uint32_t n, uint32_t writers, TextMate m1, TextMate m2, const fq_read_result* __restrict__ r1,
                             const fq_read_result* __restrict__ r2, const fq_pair_result* __restrict__ rp,
                             uint32_t* __restrict__ sizes, uint32_t* __restrict__ counters){
    const int lane = threadIdx.x & 31;
    for(uint32_t base = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; base < n; base += gridDim.x * blockDim.x){
        const uint32_t i = base + lane;
        uint32_t sz[FQK_CH_COUNT] = {0, 0, 0, 0, 0, 0, 0};
        uint32_t rc[FQK_CH_COUNT] = {0, 0, 0, 0, 0, 0, 0};
        bool passed = false;
        if(i < n){
            Piece pc[2]; int np;
            const fq_read_result x1 = r1[i];
            fq_read_result x2 = {}; fq_pair_result pr = {};
            if(paired){ x2 = r2[i]; pr = rp[i]; np = route_pe(x1, x2, pr, writers, pc, passed); }
            else np = route_se(x1, writers, pc, passed);
            for(int k = 0; k < np; ++k){
                const uint32_t b = piece_bytes(pc[k], m1, m2, i, x1, x2, pr);
#pragma unroll
                for(int c = 0; c < FQK_CH_COUNT; ++c) if(pc[k].ch == c){ sz[c] += b; rc[c] += 1; }
            }
#pragma unroll
            for(int c = 0; c < FQK_CH_COUNT; ++c) sizes[(size_t)c * n + i] = sz[c];
        }
#pragma unroll
        for(int c = 0; c < FQK_CH_COUNT; ++c){
            const uint32_t t = __reduce_add_sync(0xffffffffu, rc[c]);
            if(lane == 0 && t) atomicAdd(&counters[c], t);
        }
        const uint32_t np_ = __reduce_add_sync(0xffffffffu, passed ? 1u : 0u);
        if(lane == 0 && np_) atomicAdd(&counters[FQK_CH_COUNT], np_);
    }
}

// ------------------------------------------------------------------ exclusive scans (one per channel)
constexpr int SCAN_T = 256, SCAN_ITEMS = 8, SCAN_BLOCK = SCAN_T * SCAN_ITEMS;

__device__ __forceinline__ uint64_t block_excl_scan(uint64_t v, uint64_t* total){     // exclusive scan of one value per thread
    __shared__ uint64_t wsum[SCAN_T / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint64_t x = v;
#pragma unroll
    for(int d = 1; d < 32; d <<= 1){ const uint64_t y = __shfl_up_sync(0xffffffffu, x, d); if(lane >= d) x += y; }
    if(lane == 31) wsum[w] = x;
    __syncthreads();
    uint64_t before = 0, all = 0;
#pragma unroll
    for(int k = 0; k < SCAN_T / 32; ++k){ if(k < w) before += wsum[k]; all += wsum[k]; }
    __syncthreads();
    if(total) *total = all;
    return before + x - v;
}
// blocksum[ch][b] = sum of block b's items
__global__ void k_scan_reduce(uint32_t n, const uint32_t* __restrict__ sizes, uint64_t* __restrict__ blocksum, uint32_t nblocks){
    const uint32_t ch = blockIdx.y, b = blockIdx.x;
    const uint32_t* s = sizes + (size_t)ch * n;
    uint64_t v = 0;
    for(int k = 0; k < SCAN_ITEMS; ++k){
        const uint64_t i = (uint64_t)b * SCAN_BLOCK + (uint64_t)k * SCAN_T + threadIdx.x;
        if(i < n) v += s[i];
    }
    uint64_t tot;
    block_excl_scan(v, &tot);
    if(threadIdx.x == 0) blocksum[(size_t)ch * nblocks + b] = tot;
}
// exclusive scan of the block sums of a channel (one block per channel), totals[ch] = bytes of the channel
__global__ void k_scan_tops(uint64_t* __restrict__ blocksum, uint32_t nblocks, uint64_t* __restrict__ totals){
    const uint32_t ch = blockIdx.x;
    uint64_t* s = blocksum + (size_t)ch * nblocks;
    uint64_t carry = 0;
    for(uint32_t b0 = 0; b0 < nblocks; b0 += SCAN_T){
        const uint32_t i = b0 + threadIdx.x;
        const uint64_t v = i < nblocks ? s[i] : 0;
        uint64_t tot;
        const uint64_t ex = block_excl_scan(v, &tot);
        if(i < nblocks) s[i] = carry + ex;
        carry += tot;
    }
    if(threadIdx.x == 0) totals[ch] = carry;
}
// sizes -> offsets in place.  Offsets are 32 bit: the caller rejects channels of 4 GiB and more (totals are 64 bit)
__global__ void k_scan_apply(uint32_t n, uint32_t* __restrict__ sizes, const uint64_t* __restrict__ blocksum, uint32_t nblocks){
    const uint32_t ch = blockIdx.y, b = blockIdx.x;
    uint32_t* s = sizes + (size_t)ch * n;
    // thread t owns SCAN_ITEMS consecutive items
    const uint64_t i0 = (uint64_t)b * SCAN_BLOCK + (uint64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS]; uint64_t sum = 0;
#pragma unroll
    for(int k = 0; k < SCAN_ITEMS; ++k){ v[k] = (i0 + k < n) ? s[i0 + k] : 0u; sum += v[k]; }
    uint64_t run = blocksum[(size_t)ch * nblocks + b] + block_excl_scan(sum, nullptr);
#pragma unroll
    for(int k = 0; k < SCAN_ITEMS; ++k){ if(i0 + k < n) s[i0 + k] = (uint32_t)run; run += v[k]; }
}

// ------------------------------------------------------------------ copy
__device__ __forceinline__ void put_bytes(char* dst, uint32_t& o, const char* src, uint32_t n, int lane){     // warp copy, advances o
    for(uint32_t k = lane; k < n; k += 32) dst[o + k] = src[k];
    o += n;
}
// The same for the long pieces (names, bases, qualities) with aligned 32-bit stores: bytes up to the destination's next
// word boundary, then whole words assembled from two aligned source words by a funnel shift, then the last bytes.  The
// source buffers carry at least 16 spare bytes, so the word that straddles the end of a piece can be read.
__device__ __forceinline__ void put(char* dst, uint32_t& o, const char* src, uint32_t n, int lane){
    char* d = dst + o;
    const uint32_t head = min((4u - (uint32_t)(reinterpret_cast<uintptr_t>(d) & 3u)) & 3u, n);
    if((uint32_t)lane < head) d[lane] = src[lane];
    const uint32_t nwords = (n - head) >> 2;
    const char* s = src + head;
    const uint32_t sa = (uint32_t)(reinterpret_cast<uintptr_t>(s) & 3u);
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(s - sa);
    uint32_t* dw = reinterpret_cast<uint32_t*>(d + head);
    for(uint32_t k = lane; k < nwords; k += 32){
        const uint32_t w0 = sw[k], w1 = sa ? sw[k + 1] : 0u;
        dw[k] = __funnelshift_r(w0, w1, 8u * sa);
    }
    const uint32_t done = head + 4u * nwords;
    if((uint32_t)lane < n - done) d[done + lane] = src[done + lane];
    o += n;
}
__device__ __forceinline__ void put_ch(char* dst, uint32_t& o, char c, int lane){ if(lane == 0) dst[o] = c; o += 1; }
__device__ __forceinline__ void put_dec(char* dst, uint32_t& o, uint32_t v, int lane){
    const uint32_t nd = dec_digits(v);
    if(lane == 0){ uint32_t x = v; for(uint32_t k = 0; k < nd; ++k){ dst[o + nd - 1 - k] = (char)('0' + x % 10); x /= 10; } }
    o += nd;
}
__device__ __forceinline__ char comp_base(char c){          // Seq::reverseComplement, src/seq.h:24-48
    switch(c){ case 'A': case 'a': return 'T'; case 'T': case 't': return 'A'; case 'C': case 'c': return 'G';
               case 'G': case 'g': return 'C'; default: return 'N'; }
}

__device__ void write_piece(const Piece& p, char* dst, uint32_t o, const TextMate& m1, const TextMate& m2, uint32_t i,
                            const fq_read_result& x1, const fq_read_result& x2, const fq_pair_result& pr, int lane){
    if(p.kind == 2){
        // OverlapAnalysis::merge: read1[0, len1) + revcomp(read2 view)[ol, ol + len2), qualities reversed alike
        const char* name = m1.names + m1.name_off[i]; const uint32_t nl = m1.nlen(i);
        const char* strand = m1.strands + m1.strand_off[i]; const uint32_t sl = m1.slen(i);
        const uint32_t len1 = pr.merge_len1, len2 = pr.merge_len2, ol = pr.ov2_len;
        const uint32_t sp = first_space(name, nl);
        if(sp < nl) put(dst, o, name, sp ? sp - 1 : nl, lane);    // sp == 0: see merged_name_len
        put_bytes(dst, o, "_merged_", 8, lane); put_dec(dst, o, len1, lane); put_ch(dst, o, '_', lane); put_dec(dst, o, len2, lane);
        if(sp < nl) put(dst, o, name + sp, nl - sp, lane);
        put_ch(dst, o, '\n', lane);
        uint32_t s1, l1, s2, l2; view_of(x1, m1.len[i], s1, l1); view_of(x2, m2.len[i], s2, l2);
        const char* a = m1.seq + m1.src[i] + s1; const char* aq = m1.qual + m1.qoff(i) + s1;
        const char* b = m2.seq + m2.src[i] + s2; const char* bq = m2.qual + m2.qoff(i) + s2;
        put(dst, o, a, len1, lane);
        for(uint32_t t = lane; t < len2; t += 32) dst[o + t] = comp_base(b[l2 - 1 - ol - t]);
        o += len2;
        put_ch(dst, o, '\n', lane);
        put(dst, o, strand, sl, lane); put_ch(dst, o, '\n', lane);
        put(dst, o, aq, len1, lane);
        for(uint32_t t = lane; t < len2; t += 32) dst[o + t] = bq[l2 - 1 - ol - t];
        o += len2;
        put_ch(dst, o, '\n', lane);
        return;
    }
    const TextMate& m = p.kind ? m2 : m1;
    const fq_read_result& x = p.kind ? x2 : x1;
    uint32_t st, ln; view_of(x, m.len[i], st, ln);
    put(dst, o, m.names + m.name_off[i], m.nlen(i), lane);
    if(p.tag >= 0){ put_ch(dst, o, ' ', lane); put_bytes(dst, o, c_tags[p.tag], tag_len(p.tag), lane); }
    put_ch(dst, o, '\n', lane);
    put(dst, o, m.seq + m.src[i] + st, ln, lane); put_ch(dst, o, '\n', lane);
    put(dst, o, m.strands + m.strand_off[i], m.slen(i), lane); put_ch(dst, o, '\n', lane);
    put(dst, o, m.qual + m.qoff(i) + st, ln, lane); put_ch(dst, o, '\n', lane);
}

__global__ void k_text_copy(bool paired, uint32_t n, uint32_t writers, TextMate m1, TextMate m2, const fq_read_result* __restrict__ r1,
                            const fq_read_result* __restrict__ r2, const fq_pair_result* __restrict__ rp,
                            const uint32_t* __restrict__ offsets, TextOut out){
    const int lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for(uint32_t i = warp; i < n; i += nwarps){
        Piece pc[2]; int np; bool passed;
        const fq_read_result x1 = r1[i];
        fq_read_result x2 = {}; fq_pair_result pr = {};
        if(paired){ x2 = r2[i]; pr = rp[i]; np = route_pe(x1, x2, pr, writers, pc, passed); }
        else np = route_se(x1, writers, pc, passed);
        uint32_t used[FQK_CH_COUNT] = {0, 0, 0, 0, 0, 0, 0};          // a second piece of the same channel follows the first
        for(int k = 0; k < np; ++k){
            const int ch = pc[k].ch;
            uint32_t u = 0;
#pragma unroll
            for(int c = 0; c < FQK_CH_COUNT; ++c) if(ch == c) u = used[c];
            if(out.p[ch]) write_piece(pc[k], out.p[ch], offsets[(size_t)ch * n + i] + u, m1, m2, i, x1, x2, pr, lane);   // nullptr: channel not wanted
            const uint32_t b = piece_bytes(pc[k], m1, m2, i, x1, x2, pr);
#pragma unroll
            for(int c = 0; c < FQK_CH_COUNT; ++c) if(ch == c) used[c] += b;
        }
    }
}

// BaseCorrector::correctByOverlapAnalysis rewrote bases of the PACKED batch; the same edits on the staged characters
__global__ void k_text_apply_edits(const fq_correction* __restrict__ edits, const uint32_t* __restrict__ n_edits, uint32_t cap,
                                   char* seq1, char* qual1, const uint32_t* __restrict__ src1, const uint32_t* __restrict__ qsrc1,
                                   char* seq2, char* qual2, const uint32_t* __restrict__ src2, const uint32_t* __restrict__ qsrc2){
    const uint32_t n = min(*n_edits, cap);
    for(uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x){
        const fq_correction c = edits[e];
        const bool m2 = (c.mate_base & 0x80) != 0;
        (m2 ? seq2 : seq1)[(m2 ? src2 : src1)[c.pair] + c.pos] = (char)(c.mate_base & 0x7F);
        (m2 ? qual2 : qual1)[(m2 ? qsrc2 : qsrc1)[c.pair] + c.pos] = (char)c.qual;
    }
}

// ------------------------------------------------------------------ raw FASTQ blocks: the newline scan on the device
// FqReader::getLine (src/fqreader.cpp:90-148) walks the decompressed text byte by byte.  Here: newlines are counted
// per 2 KiB chunk, the counts are scanned, every chunk writes the positions of its newlines, and record r is lines
// 4r .. 4r+3 (name, sequence, third line, quality; a trailing '\r' is dropped as clearLineBreaks does).  The reference
// also skips stray lines that do not start with '@' (src/fqreader.cpp:167-170); a block that needs that is refused
// (error bit 8): the GPU path takes well-formed four-line FASTQ only.
__device__ __forceinline__ uint32_t nl_in8(const char* text, uint64_t bytes, uint64_t at, uint32_t& mask){
    mask = 0;
#pragma unroll
    for(int k = 0; k < 8; ++k) if(at + k < bytes && text[at + k] == '\n') mask |= 1u << k;
    return (uint32_t)__popc(mask);
}
__global__ void k_nl_count(const char* __restrict__ text, uint64_t bytes, uint32_t* __restrict__ counts){
    const uint64_t at = (uint64_t)blockIdx.x * NL_CHUNK + threadIdx.x * 8u;
    uint32_t m; const uint32_t c = nl_in8(text, bytes, at, m);
    uint64_t tot;
    block_excl_scan(c, &tot);
    if(threadIdx.x == 0) counts[blockIdx.x] = (uint32_t)tot;
}
__global__ void k_nl_positions(const char* __restrict__ text, uint64_t bytes, const uint32_t* __restrict__ chunk_off, uint32_t* __restrict__ nlpos){
    const uint64_t at = (uint64_t)blockIdx.x * NL_CHUNK + threadIdx.x * 8u;
    uint32_t m; const uint32_t c = nl_in8(text, bytes, at, m);
    uint32_t o = chunk_off[blockIdx.x] + (uint32_t)block_excl_scan(c, nullptr);
    while(m){ const int k = __ffs((int)m) - 1; m &= m - 1; nlpos[o++] = (uint32_t)(at + k); }
}
__global__ void k_fastq_index(const char* __restrict__ text, const uint32_t* __restrict__ nlpos, uint32_t n, uint32_t* __restrict__ src_seq,
                              uint32_t* __restrict__ src_qual, uint16_t* __restrict__ len, uint32_t* __restrict__ name_off,
                              uint32_t* __restrict__ name_len, uint32_t* __restrict__ strand_off, uint32_t* __restrict__ strand_len,
                              uint32_t* __restrict__ padlen, uint32_t* __restrict__ misc, uint32_t* __restrict__ err){
    uint32_t mx = 0, bad = 0;
    for(uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x){
        uint32_t s[4], e[4];
#pragma unroll
        for(int k = 0; k < 4; ++k){
            const uint32_t li = 4u * r + k;
            s[k] = li ? nlpos[li - 1] + 1u : 0u;
            e[k] = nlpos[li];
            if(e[k] > s[k] && text[e[k] - 1] == '\r') --e[k];
        }
        const uint32_t l = e[1] - s[1];
        if(e[0] == s[0] || text[s[0]] != '@' || e[2] == s[2] || text[s[2]] != '+' || e[3] - s[3] != l) bad |= 8u;
        const bool malformed = e[0] == s[0] || text[s[0]] != '@' || e[2] == s[2] || text[s[2]] != '+' || e[3] - s[3] != l;
        if(l > FQK_MAX_READ_LEN){ bad |= 1u; }
        // a malformed record gets length 0 and no name / third line: the block is refused when the error word is read at
        // the end, but the packer, the trim kernels and the text assembly still run over it first and must not follow a
        // sequence length the quality line does not have (that would read past the uploaded block)
        const uint32_t lc = malformed ? 0u : min(l, (uint32_t)FQK_MAX_READ_LEN);
        src_seq[r] = s[1]; src_qual[r] = malformed ? s[1] : s[3]; len[r] = (uint16_t)lc;
        name_off[r] = s[0]; name_len[r] = malformed ? 0u : e[0] - s[0]; strand_off[r] = s[2]; strand_len[r] = malformed ? 0u : e[2] - s[2];
        padlen[r] = (lc + FQK_PAD - 1u) / FQK_PAD * FQK_PAD;
        mx = max(mx, lc);
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    bad = __reduce_or_sync(0xffffffffu, bad);
    if((threadIdx.x & 31) == 0){ if(mx) atomicMax(&misc[0], mx); if(bad) atomicOr(err, bad); }
}

__global__ void k_ascii_index(uint32_t n, const uint64_t* __restrict__ off, uint32_t* __restrict__ src, uint16_t* __restrict__ len,
                              uint32_t* __restrict__ padlen, uint32_t* __restrict__ err){
    const uint64_t o0 = off[0];
    for(uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x){
        const uint64_t a = off[r], l = off[r + 1] - a;
        if(l > FQK_MAX_READ_LEN) atomicOr(err, 1u);
        const uint32_t lc = (uint32_t)(l > FQK_MAX_READ_LEN ? FQK_MAX_READ_LEN : l);
        src[r] = (uint32_t)(a - o0); len[r] = (uint16_t)lc;
        padlen[r] = (lc + FQK_PAD - 1u) / FQK_PAD * FQK_PAD;
    }
}
__global__ void k_zero_tail(const uint64_t* __restrict__ total, uint8_t* __restrict__ qual, uint8_t* __restrict__ bases2){
    const uint64_t t = *total;                                  // a multiple of 16
    if(threadIdx.x < 64) qual[t + threadIdx.x] = 0;
    if(threadIdx.x < 16) bases2[t / 4 + threadIdx.x] = 0;
}

__global__ void k_narrow(uint32_t n, const uint64_t* __restrict__ in, uint32_t* __restrict__ out){
    const uint64_t b = in[0];
    for(uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = (uint32_t)(in[i] - b);
}

} // namespace

// ------------------------------------------------------------------ launchers
static int text_grid(uint32_t n, int per_block){ const long long g = ((long long)n + per_block - 1) / per_block; return (int)(g < 1 ? 1 : (g > 148 * 16 ? 148 * 16 : g)); }

void text_launch_sizes(cudaStream_t st, bool paired, uint32_t n, uint32_t writers, TextMate m1, TextMate m2, const fq_read_result* r1,
                       const fq_read_result* r2, const fq_pair_result* rp, uint32_t* sizes, uint32_t* counters){
    k_text_sizes<<<text_grid(n, 256), 256, 0, st>>>(paired, n, writers, m1, m2, r1, r2, rp, sizes, counters);
}
size_t text_scan_scratch_u64(uint32_t n){ return (size_t)FQK_CH_COUNT * (((size_t)n + SCAN_BLOCK - 1) / SCAN_BLOCK) + 8; }
void text_launch_scan(cudaStream_t st, uint32_t n, uint32_t* sizes, uint64_t* totals, uint64_t* scratch, int nch){
    const uint32_t nb = (uint32_t)(((size_t)n + SCAN_BLOCK - 1) / SCAN_BLOCK);
    if(n == 0){ cudaMemsetAsync(totals, 0, (size_t)nch * 8, st); return; }
    k_scan_reduce<<<dim3(nb, nch), SCAN_T, 0, st>>>(n, sizes, scratch, nb);
    k_scan_tops<<<nch, SCAN_T, 0, st>>>(scratch, nb, totals);
    k_scan_apply<<<dim3(nb, nch), SCAN_T, 0, st>>>(n, sizes, scratch, nb);
}
void text_launch_copy(cudaStream_t st, bool paired, uint32_t n, uint32_t writers, TextMate m1, TextMate m2, const fq_read_result* r1,
                      const fq_read_result* r2, const fq_pair_result* rp, const uint32_t* offsets, TextOut out){
    k_text_copy<<<text_grid(n, 8), 256, 0, st>>>(paired, n, writers, m1, m2, r1, r2, rp, offsets, out);
}
void text_launch_apply_edits(cudaStream_t st, const fq_correction* edits, const uint32_t* n_edits, uint32_t cap,
                             char* seq1, char* qual1, const uint32_t* src1, const uint32_t* qsrc1,
                             char* seq2, char* qual2, const uint32_t* src2, const uint32_t* qsrc2){
    k_text_apply_edits<<<148, 256, 0, st>>>(edits, n_edits, cap, seq1, qual1, src1, qsrc1 ? qsrc1 : src1, seq2, qual2, src2, qsrc2 ? qsrc2 : src2);
}

void fastq_launch_count(cudaStream_t st, const char* text, uint64_t bytes, uint32_t* counts){
    const uint32_t nc = (uint32_t)((bytes + NL_CHUNK - 1) / NL_CHUNK);
    if(nc) k_nl_count<<<nc, NL_CHUNK / 8, 0, st>>>(text, bytes, counts);
}
void fastq_launch_positions(cudaStream_t st, const char* text, uint64_t bytes, const uint32_t* chunk_off, uint32_t* nlpos){
    const uint32_t nc = (uint32_t)((bytes + NL_CHUNK - 1) / NL_CHUNK);
    if(nc) k_nl_positions<<<nc, NL_CHUNK / 8, 0, st>>>(text, bytes, chunk_off, nlpos);
}
void fastq_launch_index(cudaStream_t st, const char* text, const uint32_t* nlpos, uint32_t n, uint32_t* src_seq, uint32_t* src_qual,
                        uint16_t* len, uint32_t* name_off, uint32_t* name_len, uint32_t* strand_off, uint32_t* strand_len,
                        uint32_t* padlen, uint32_t* misc, uint32_t* err){
    if(n) k_fastq_index<<<text_grid(n, 256), 256, 0, st>>>(text, nlpos, n, src_seq, src_qual, len, name_off, name_len, strand_off, strand_len,
                                                         padlen, misc, err);
}

void ascii_launch_index(cudaStream_t st, uint32_t n, const uint64_t* off, uint32_t* src, uint16_t* len, uint32_t* padlen, uint32_t* err){
    if(n) k_ascii_index<<<text_grid(n, 256), 256, 0, st>>>(n, off, src, len, padlen, err);
}
void ascii_launch_zero_tail(cudaStream_t st, const uint64_t* total, uint8_t* qual, uint8_t* bases2){
    k_zero_tail<<<1, 64, 0, st>>>(total, qual, bases2);
}
void text_launch_narrow(cudaStream_t st, uint32_t n, const uint64_t* in, uint32_t* out){
    if(n) k_narrow<<<text_grid(n, 256), 256, 0, st>>>(n, in, out);
}
