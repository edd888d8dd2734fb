/*
 * fq_oracle.c -- TEST INFRASTRUCTURE ONLY (see fq_oracle.h).
 *
 * Plain-C restatement of the reference's per-read hot path.  Every function
 * cites the reference file:line (relative to /root/reference/) it follows.
 * Reads are handled as views (pointer + length) into the caller's ASCII
 * buffers: all of the reference's trims are prefix/suffix cuts.
 *
 * PARITY PINNED: this port is checked against the reference itself, built here from its own sources
 * (oracle/_ref/libfqref.so, oracle/_ref/fqtool_ref): tests/test_oracle_vs_ref.py (records, accumulators, duplication
 * tables on 22 option sets), tests/test_known_answers.py (whole-binary known answers), tests/test_golden.py (fixtures
 * generated from the reference build), tests/test_prepass.py (the pre-pass loops with the reference's Evaluator::seq2int
 * and getAdapterWithSeed).  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline leg may use it; the product
 * (fqtool_b200/, include/, integration/) never does.
 */
#include "fq_oracle.h"
#include <stdlib.h>
#include <string.h>
#include <math.h>

typedef struct { const char* seq; const char* qual; int len; } rview;

struct fqo_ctx {
    fq_params P;
    fq_stats_layout L;
    uint64_t* flat;
    uint64_t* dups; uint32_t* counts; uint8_t* gc;   /* src/duplicate.h:64-66 */
    char* tmp; int tmp_cap;                          /* revcomp scratch       */
    char* cbuf; size_t cbuf_cap;                     /* mutable copy of a pair (base correction) */
    fq_correction* corr; uint32_t ncorr, corr_cap;   /* edits of the last process_pe call */
};

/* ------------------------------------------------------------------ layout */
int fqo_layout(const fq_params* p, fq_stats_layout* L){
    memset(L, 0, sizeof(*L));
    uint64_t C = p->max_cycles > 0 ? (uint64_t)p->max_cycles : 512;
    L->max_cycles = C;
    L->kmer_bins = p->kmer_len > 0 ? (1ull << (2 * p->kmer_len)) : 0;
    L->s_q20 = 0; L->s_q30 = 8 * C; L->s_content = 16 * C; L->s_qualsum = 24 * C;
    L->s_total_base = 32 * C; L->s_total_qual = 33 * C;
    L->s_reads = 34 * C; L->s_length_sum = 34 * C + 1; L->s_kmer = 34 * C + 2;
    L->stats_stride = 34 * C + 2 + L->kmer_bins;
    for(int s = 0; s < 4; ++s) L->stats_off[s] = (uint64_t)s * L->stats_stride;
    uint64_t o = 4 * L->stats_stride;
    L->filter_off = o;        o += FQK_FILTER_RESULT_TYPES;
    L->adapter_reads_off = o; o += 1;
    L->adapter_bases_off = o; o += 1;
    L->polyx_reads_off = o;   o += 5;
    L->polyx_bases_off = o;   o += 5;
    L->merged_pairs_off = o;  o += 1;
    L->insert_hist_off = o;   L->insert_hist_len = (uint64_t)p->insert_size_max + 1; o += L->insert_hist_len;
    L->correction_off = o; o += 64;
    L->corrected_reads_off = o; o += 1;
    L->total_u64 = o;
    L->dup_bins = p->dup_enabled ? (1ull << (2 * p->dup_keylen)) : 0;
    return 0;
}

fqo_ctx* fqo_create(const fq_params* p){
    fqo_ctx* c = (fqo_ctx*)calloc(1, sizeof(fqo_ctx));
    if(!c) return NULL;
    c->P = *p;
    fqo_layout(p, &c->L);
    c->flat = (uint64_t*)calloc(c->L.total_u64, sizeof(uint64_t));
    if(c->L.dup_bins){
        c->dups = (uint64_t*)calloc(c->L.dup_bins, sizeof(uint64_t));
        c->counts = (uint32_t*)calloc(c->L.dup_bins, sizeof(uint32_t));
        c->gc = (uint8_t*)calloc(c->L.dup_bins, 1);
    }
    c->tmp_cap = 1 << 16;
    c->tmp = (char*)malloc(c->tmp_cap);
    return c;
}

void fqo_destroy(fqo_ctx* c){
    if(!c) return;
    free(c->flat); free(c->dups); free(c->counts); free(c->gc); free(c->tmp); free(c->cbuf); free(c->corr); free(c);
}

static inline int imin(int a, int b){ return a < b ? a : b; }
static inline int imax(int a, int b){ return a > b ? a : b; }

/* Read::resize, src/read.h:181-187: ignores len > length and len < 0 */
static inline void rv_resize(rview* r, int len){
    if(len > r->len || len < 0) return;
    r->len = len;
}

/* ------------------------------------------------------------- trimAndCut */
/* Filter::trimAndCut, src/filter.cpp:69-189.  Returns 0 when the reference
 * returns NULL (read untouched), 1 otherwise (view updated in place). */
static int trim_and_cut(const fq_params* P, rview* r, int F, int T){
    int ef = P->cut_front_enabled, er = P->cut_right_enabled, et = P->cut_tail_enabled;
    if(F == 0 && T == 0 && !ef && !er && !et) return 1;              /* :71-73 */
    int rlen = r->len - F - T;                                        /* :75 */
    if(rlen < 0) return 0;                                            /* :76-78 */
    if(F == 0 && !ef && !er && !et){ rv_resize(r, rlen); return 1; }  /* :80-82 */
    if(!ef && !er && !et){ r->seq += F; r->qual += F; r->len = rlen; return 1; } /* :83-87 */
    int l = r->len;                                                   /* :90 */
    const char* qual = r->qual; const char* seq = r->seq;
    if(ef){                                                           /* :94-122 */
        int w = P->cut_front_window, s = F;
        if(l - F - T - w <= 0) return 0;
        int tq = 0;
        for(int i = 0; i < w - 1; ++i) tq += qual[s + i];
        for(s = F; s + w < l - T; ++s){
            tq += qual[s + w - 1];
            if(s > F) tq -= qual[s - 1];
            if((double)tq / (double)w >= 33 + P->cut_front_qual) break;
        }
        if(s > 0) s = s + w - 1;
        while(s < l && seq[s] == 'N') ++s;
        F = s;
        rlen = l - F - T;
    }
    if(er){                                                           /* :125-152 */
        int w = P->cut_right_window, s = F;
        if(l - F - T - w <= 0) return 0;
        int tq = 0, found = 0;
        for(int i = 0; i < w - 1; ++i) tq += qual[s + i];
        for(s = F; s + w < l - T; ++s){
            tq += qual[s + w - 1];
            if(s > F) tq -= qual[s - 1];
            if((double)tq / (double)w < 33 + P->cut_right_qual){ found = 1; break; }
        }
        if(found){
            while(s < l - 1 && qual[s] >= 33 + P->cut_right_qual) ++s;
            rlen = s - F;
        }
    }
    if(!er && et){                                                    /* :155-181 */
        int w = P->cut_tail_window;
        if(l - F - T - w <= 0) return 0;
        int tq = 0, t = l - T - 1;
        for(int i = 0; i < w - 1; ++i) tq += qual[t - i];
        for(t = l - T - 1; t - w >= F; --t){
            tq += qual[t - w + 1];
            if(t < l - T - 1) tq -= qual[t + 1];
            if((double)tq / (double)w >= 33 + P->cut_tail_qual) break;
        }
        if(t < l - 1) t = t - w + 1;
        while(t >= 0 && seq[t] == 'N') --t;
        rlen = t - F + 1;
    }
    if(rlen <= 0 || F >= l - 1) return 0;                             /* :183-185 */
    r->seq += F; r->qual += F; r->len = rlen;                         /* :186-187 */
    return 1;
}

/* ------------------------------------------------------------- passFilter */
/* Filter::passLowComplexityFliter, src/filter.cpp:54-67 */
static int pass_low_complexity(const fq_params* P, const rview* r){
    int diff = 0, rlen = r->len;
    if(rlen <= 1) return 0;
    for(int i = 0; i < rlen - 1; ++i) if(r->seq[i] != r->seq[i + 1]) ++diff;
    return (double)diff / (rlen - 1) >= P->complexity_threshold;
}

/* Filter::passFilter, src/filter.cpp:3-52; r == NULL is passed as isnull */
static int pass_filter(const fq_params* P, const rview* r, int isnull){
    if(isnull || r->len == 0) return FQK_FAIL_LENGTH;
    int rlen = r->len, lowQualNum = 0, nBaseNum = 0, totalQual = 0;
    if(P->qual_filter_enabled || P->length_filter_enabled){
        for(int i = 0; i < rlen; ++i){
            totalQual += r->qual[i] - 33;
            if(r->seq[i] == 'N') ++nBaseNum;
            if(r->qual[i] < P->low_quality_limit) ++lowQualNum;
        }
    }
    if(P->qual_filter_enabled){
        if(lowQualNum > P->low_quality_base_limit) return FQK_FAIL_QUALITY;
        else if(P->average_quality_limit > 0 && P->average_quality_limit > (double)totalQual / rlen)
            return FQK_FAIL_QUALITY;
    }
    if(P->qual_filter_enabled && nBaseNum > P->n_base_limit) return FQK_FAIL_N_BASE;
    if(P->length_filter_enabled){
        if(rlen < P->min_read_length) return FQK_FAIL_LENGTH;
        if(P->max_read_length > 0 && rlen > P->max_read_length) return FQK_FAIL_TOO_LONG;
    }
    if(P->complexity_enabled && !pass_low_complexity(P, r)) return FQK_FAIL_COMPLEXITY;
    return FQK_PASS_FILTER;
}

/* ------------------------------------------------------------ FilterResult */
/* FilterResult::addPolyXTrimmed, src/filterresult.cpp:43-46 */
static void add_polyx(fqo_ctx* c, int base, int length){
    c->flat[c->L.polyx_reads_off + base] += 1;
    c->flat[c->L.polyx_bases_off + base] += (uint64_t)(int64_t)length;
}
/* FilterResult::addFilterResult(result, n), src/filterresult.cpp:25-41 */
static void add_filter_result(fqo_ctx* c, int result, int n){
    if(result < FQK_PASS_FILTER || result >= FQK_FILTER_RESULT_TYPES) return;
    c->flat[c->L.filter_off + result] += (uint64_t)n;
}

/* ------------------------------------------------------------------ PolyX */
/* PolyX::trimPolyG (single read), src/polyx.cpp:14-38 */
static void trim_polyg(fqo_ctx* c, rview* r, int compareReq, int maxMismatch, int oneEach,
                       fq_read_result* rec){
    const char* data = r->seq;
    int rlen = r->len, mismatch = 0, i = 0, firstGpos = rlen - 1;
    for(i = 0; i < rlen; ++i){
        if(data[rlen - i - 1] != 'G') ++mismatch;
        else firstGpos = rlen - i - 1;
        int allowed = imin(maxMismatch, imax(1, (i + 1) / oneEach));
        if(mismatch > allowed) break;
    }
    if(i + 1 >= compareReq){
        rv_resize(r, firstGpos);
        add_polyx(c, 3, rlen - firstGpos);
        rec->flags |= FQK_F_POLYG;
        rec->polyg_len = (uint16_t)(rlen - firstGpos);
    }
}

/* PolyX::trimPolyX (single read), src/polyx.cpp:45-101; trimChr as a 5-bit mask */
static void trim_polyx(fqo_ctx* c, rview* r, int mask, int compareReq, int maxMismatch, int oneEach,
                       fq_read_result* rec){
    static const char bases[5] = {'A', 'T', 'C', 'G', 'N'};
    const char* data = r->seq;
    int rlen = r->len, num[5] = {0, 0, 0, 0, 0}, pos = 0;
    for(pos = 0; pos < rlen; ++pos){
        switch(data[rlen - 1 - pos]){
            case 'A': ++num[0]; break;
            case 'T': ++num[1]; break;
            case 'C': ++num[2]; break;
            case 'G': ++num[3]; break;
            default:  ++num[4]; break;
        }
        int cmp = pos + 1;
        int allowed = imin(maxMismatch, imax(1, cmp / oneEach));
        int needToBreak = 1;
        for(int b = 0; b < 5; ++b)
            if(((mask >> b) & 1) && cmp - num[b] <= allowed) needToBreak = 0;
        if(needToBreak) break;
    }
    if(pos + 1 >= compareReq){
        int poly = 0, maxCount = -1;
        for(int b = 0; b < 5; ++b)
            if(((mask >> b) & 1) && num[b] > maxCount){ maxCount = num[b]; poly = b; }
        char polyBase = bases[poly];
        pos = imin(rlen - 1, pos);
        /* with rlen == 0 the reference reads data[0] ('\0' of the empty string) */
        while((rlen - pos - 1 >= 0 && rlen - pos - 1 < rlen ? data[rlen - pos - 1] : '\0') != polyBase && pos > 0) --pos;
        rv_resize(r, rlen - pos - 1);
        add_polyx(c, poly, pos + 1);
        rec->flags |= FQK_F_POLYX | (uint8_t)(poly << FQK_F_POLYX_SHIFT);
        rec->polyx_len = (uint16_t)(pos + 1);
    }
}

/* ---------------------------------------------------------- AdapterTrimmer */
/* FilterResult::addAdapterTrimmed(adapter, isR2), src/filterresult.cpp:138-157
 * (the std::map tallies keyed by text are rebuilt on the host from the record) */
static void add_adapter_trimmed1(fqo_ctx* c, int textlen){
    if(textlen == 0) return;
    c->flat[c->L.adapter_reads_off] += 1;
    c->flat[c->L.adapter_bases_off] += (uint64_t)textlen;
}
/* FilterResult::addAdapterTrimmed(adapterR1, adapterR2), src/filterresult.cpp:159-177 */
static void add_adapter_trimmed2(fqo_ctx* c, int len1, int len2){
    c->flat[c->L.adapter_reads_off] += 2;
    c->flat[c->L.adapter_bases_off] += (uint64_t)(len1 + len2);
}

/* AdapterTrimmer::trimBySequence, src/adaptertrimmer.cpp:29-90 */
static int trim_by_sequence(fqo_ctx* c, rview* r, const char* adata, int alen, fq_read_result* rec){
    const int matchRequired = 4, oneEach = 8;
    int rlen = r->len;
    const char* rdata = r->seq;
    if(alen < matchRequired) return 0;
    int pos = 0, found = 0, start = 0;
    if(alen >= 16) start = -4;
    else if(alen >= 12) start -= 3;
    else if(alen >= 8) start = -2;
    for(pos = start; pos < rlen - matchRequired; ++pos){
        int cmplen = imin(rlen - pos, alen);
        int allowed = cmplen / oneEach;
        int mismatch = 0, matched = 1;
        for(int i = imax(0, -pos); i < cmplen; ++i){
            if(adata[i] != rdata[i + pos]){
                ++mismatch;
                if(mismatch > allowed){ matched = 0; break; }
            }
        }
        if(matched){ found = 1; break; }
    }
    if(found){
        int textlen;
        if(pos < 0){ textlen = alen + pos; r->len = 0; }      /* :72-78 */
        else { textlen = rlen - pos; r->len = pos; }          /* :79-86 */
        add_adapter_trimmed1(c, textlen);
        rec->flags |= FQK_F_ADAPTER_SEQ;
        rec->adapter_pos = (int16_t)pos;
        rec->adapter_len = (uint16_t)textlen;
        return 1;
    }
    return 0;
}

/* --------------------------------------------------------- OverlapAnalysis */
typedef struct { int overlapped, offset, overlapLen, diff; } ovres;

/* Seq::reverseComplement, src/seq.h:24-48 */
static void revcomp(const char* s, int len, char* out){
    for(int i = 0; i < len; ++i){
        char ch;
        switch(s[i]){
            case 'A': case 'a': ch = 'T'; break;
            case 'T': case 't': ch = 'A'; break;
            case 'C': case 'c': ch = 'G'; break;
            case 'G': case 'g': ch = 'C'; break;
            default: ch = 'N'; break;
        }
        out[len - 1 - i] = ch;
    }
}

/* OverlapAnalysis::analyze, src/overlapanalysis.cpp:7-72 (sequential form) */
static ovres analyze(fqo_ctx* c, const rview* r1, const rview* r2){
    int limit = c->P.overlap_diff_limit, require = c->P.overlap_require;
    int len1 = r1->len, len2 = r2->len;
    if(len2 + 1 > c->tmp_cap){ c->tmp_cap = len2 + 1; c->tmp = (char*)realloc(c->tmp, c->tmp_cap); }
    revcomp(r2->seq, len2, c->tmp);
    const char* p1 = r1->seq; const char* p2 = c->tmp;
    const int ccr = 50;
    int overlapLen = 0, offset = 0, diff = 0;
    ovres o;
    while(offset < len1 - require){                                   /* :20-41 */
        overlapLen = imin(len1 - offset, len2);
        diff = 0;
        int i = 0;
        for(i = 0; i < overlapLen; ++i){
            if(p1[offset + i] != p2[i]){
                ++diff;
                if(diff >= limit && i < ccr) break;
            }
        }
        if(diff < limit || (diff >= limit && i > ccr)){
            o.overlapped = 1; o.offset = offset; o.overlapLen = overlapLen; o.diff = diff;
            return o;
        }
        ++offset;
    }
    offset = 0;
    while(offset > require - len2){                                   /* :44-67 */
        overlapLen = imin(len1, len2 - abs(offset));
        diff = 0;
        int i = 0;
        for(i = 0; i < overlapLen; ++i){
            if(p1[i] != p2[-offset + i]){
                ++diff;
                if(diff >= limit && i < ccr) break;
            }
        }
        if(diff < limit || (diff >= limit && i > ccr)){
            o.overlapped = 1; o.offset = offset; o.overlapLen = overlapLen; o.diff = diff;
            return o;
        }
        --offset;
    }
    o.overlapped = 0; o.offset = o.overlapLen = o.diff = 0;           /* :68-71 */
    return o;
}

/* ------------------------------------------------------------------ Stats */
/* Evaluator::seq2int restated for one window, src/evaluator.cpp:3-49: the
 * rolling form returns -1 on a non-ACGT base and then recomputes the whole
 * window next time, i.e. the value is -1 iff any base in the window is not
 * A/T/C/G, else the 2-bit code (A0 T1 C2 G3) of the window. */
static int64_t kmer_code(const char* s, int pos, int k){
    int64_t key = 0;
    for(int i = pos; i < pos + k; ++i){
        key <<= 2;
        switch(s[i]){
            case 'A': key += 0; break;
            case 'T': key += 1; break;
            case 'C': key += 2; break;
            case 'G': key += 3; break;
            default: return -1;
        }
    }
    return key;
}

/* Stats::statRead, src/stats.cpp:237-295 (ORA sampling :277-293 is out of scope) */
static int stat_read(fqo_ctx* c, int which, const char* seq, const char* qual, int len){
    const fq_stats_layout* L = &c->L;
    uint64_t* S = c->flat + L->stats_off[which];
    uint64_t C = L->max_cycles;
    if((uint64_t)len > C) return FQK_E_LIMIT;
    S[L->s_length_sum] += (uint64_t)len;
    int k = c->P.kmer_len;
    for(int i = 0; i < len; ++i){
        char base = seq[i], q = qual[i];
        int b = base & 0x07;
        if(q > '?'){ S[L->s_q20 + b * C + i]++; S[L->s_q30 + b * C + i]++; }
        else if(q > '5'){ S[L->s_q20 + b * C + i]++; }
        S[L->s_content + b * C + i]++;
        S[L->s_qualsum + b * C + i] += (uint64_t)(q - 33);
        S[L->s_total_base + i]++;
        S[L->s_total_qual + i] += (uint64_t)(q - 33);
        if(k){
            if(i < k - 1) continue;
            int64_t kv = kmer_code(seq, i - k + 1, k);
            if(kv >= 0) S[L->s_kmer + kv]++;
        }
    }
    S[L->s_reads]++;
    return 0;
}

/* -------------------------------------------------------------- Duplicate */
/* Duplicate::seq2int, src/duplicate.cpp:21-44 */
static uint64_t dup_seq2int(const char* s, int start, int keylen, int* valid){
    uint64_t ret = 0;
    for(int i = 0; i < keylen; ++i){
        ret <<= 2;
        switch(s[start + i]){
            case 'A': ret += 0; break;
            case 'T': ret += 1; break;
            case 'C': ret += 2; break;
            case 'G': ret += 3; break;
            default: *valid = 0; return 0;
        }
    }
    return ret;
}
/* Duplicate::addRecord, src/duplicate.cpp:46-62 */
static void dup_add(fqo_ctx* c, uint32_t key, uint64_t kmer32, uint8_t gc){
    if(c->counts[key] == 0){ c->counts[key] = 1; c->dups[key] = kmer32; c->gc[key] = gc; }
    else if(c->dups[key] == kmer32){ ++c->counts[key]; }
    else if(c->dups[key] > kmer32){ c->dups[key] = kmer32; c->counts[key] = 1; c->gc[key] = gc; }
}
/* Duplicate::statRead, src/duplicate.cpp:64-93 */
static void dup_stat_read(fqo_ctx* c, const char* s, int len){
    if(len < 32) return;
    int start2 = imax(0, len - 32 - 5), valid = 1;
    uint64_t ret = dup_seq2int(s, 0, c->P.dup_keylen, &valid);
    uint32_t key = (uint32_t)ret;
    if(!valid) return;
    uint64_t kmer32 = dup_seq2int(s, start2, 32, &valid);
    if(!valid) return;
    uint8_t gc = 0;
    if(c->counts[key] == 0)
        for(int i = 0; i < len; ++i) if(s[i] == 'C' || s[i] == 'G') ++gc;   /* uint8 wraps */
    gc = (uint8_t)round(255.0 * (double)gc / (double)len);
    dup_add(c, key, kmer32, gc);
}
/* Duplicate::statPair, src/duplicate.cpp:95-129 */
static void dup_stat_pair(fqo_ctx* c, const char* s1, int len1, const char* s2, int len2){
    if(len1 < 32 || len2 < 32) return;
    int valid = 1;
    uint64_t ret = dup_seq2int(s1, 0, c->P.dup_keylen, &valid);
    uint32_t key = (uint32_t)ret;
    if(!valid) return;
    uint64_t kmer32 = dup_seq2int(s2, 0, 32, &valid);
    if(!valid) return;
    uint8_t gc = 0;
    if(c->counts[key] == 0){
        for(int i = 0; i < len1; ++i) if(s1[i] == 'C' || s1[i] == 'G') ++gc;
        for(int i = 0; i < len2; ++i) if(s2[i] == 'C' || s2[i] == 'G') ++gc;
    }
    gc = (uint8_t)round(255.0 * (double)gc / (double)(len1 + len2));
    dup_add(c, key, kmer32, gc);
}

/* ---------------------------------------------------------------- SE flow */
static void rec_init(fq_read_result* r){ memset(r, 0, sizeof(*r)); }

/* body of the loop of SingleEndProcessor::processSingleEnd, src/seprocessor.cpp:294-354 */
int fqo_process_se(fqo_ctx* c, uint32_t n, uint64_t first_index,
                   const char* seq, const char* qual, const uint64_t* off, fq_read_result* out){
    (void)first_index;
    const fq_params* P = &c->P;
    for(uint32_t p = 0; p < n; ++p){
        const char* s0 = seq + off[p]; const char* q0 = qual + off[p];
        int len0 = (int)(off[p + 1] - off[p]);
        fq_read_result* rec = &out[p];
        rec_init(rec);
        int rc = stat_read(c, 0, s0, q0, len0);                       /* :298 */
        if(rc) return rc;
        if(P->dup_enabled) dup_stat_read(c, s0, len0);                /* :300-302 */
        rview r = { s0, q0, len0 };
        int alive = trim_and_cut(P, &r, P->trim_front1, P->trim_tail1);   /* :313 */
        if(alive){
            if(P->polyg_enabled)                                      /* :315-319 */
                trim_polyg(c, &r, P->polyg_min_len, P->polyg_max_mismatch, P->polyg_one_mismatch_each, rec);
            if(P->adapter_trim_enabled && P->adapter_r1_len > 0)      /* :321-323 */
                trim_by_sequence(c, &r, P->adapter_r1, P->adapter_r1_len, rec);
            if(P->polyx_enabled)                                      /* :325-330 */
                trim_polyx(c, &r, P->polyx_base_mask, P->polyx_min_len, P->polyx_max_mismatch,
                           P->polyx_one_mismatch_each, rec);
            if(P->trim_max_len1 > 0 && P->trim_max_len1 < r.len) rv_resize(&r, P->trim_max_len1); /* :332-336 */
        }
        int result = pass_filter(P, &r, !alive);                      /* :339 */
        add_filter_result(c, result, 1);                              /* :340, mPaired false */
        if(alive && result == FQK_PASS_FILTER){                       /* :342-345 */
            rc = stat_read(c, 2, r.seq, r.qual, r.len);
            if(rc) return rc;
        }
        rec->result = (uint8_t)result;
        if(!alive){ rec->flags |= FQK_F_DROPPED; rec->start = 0; rec->len = 0; }
        else { rec->start = (uint16_t)(r.seq - s0); rec->len = (uint16_t)r.len; }
    }
    return 0;
}

/* ------------------------------------------------------------ BaseCorrector */
/* util::complement, src/util.h:438-451 */
static char complement(char b){
    switch(b){ case 'A': case 'a': return 'T'; case 'T': case 't': return 'A';
               case 'C': case 'c': return 'G'; case 'G': case 'g': return 'C'; default: return 'N'; }
}
static void push_corr(fqo_ctx* c, uint32_t pair, int pos, int mate, char base, char qual){
    if(c->ncorr == c->corr_cap){
        c->corr_cap = c->corr_cap ? c->corr_cap * 2 : 1024;
        c->corr = (fq_correction*)realloc(c->corr, c->corr_cap * sizeof(fq_correction));
    }
    fq_correction e; e.pair = pair; e.pos = (uint16_t)pos; e.mate_base = (uint8_t)((mate << 7) | (base & 0x7F)); e.qual = (uint8_t)qual;
    c->corr[c->ncorr++] = e;
}
/* BaseCorrector::correctByOverlapAnalysis, src/basecorrector.cpp:14-70.  seq/qual are the mutable current
 * views; o1/o2 = offset of each view inside the original read.  addCorrection(from, to) is called AFTER the
 * base was overwritten through the same buffer (:40-45, :47-52), so `from` already equals `to`. */
static void correct_by_overlap(fqo_ctx* c, char* seq1, char* qual1, int len1, char* seq2, char* qual2, int len2,
                               const ovres* ov, uint32_t pair, int o1, int o2){
    (void)len1;
    if(ov->diff == 0 || ov->diff > 5) return;
    const int ol = ov->overlapLen;
    const int start1 = imax(0, ov->offset);
    const int start2 = len2 - imax(0, -ov->offset) - 1;
    const char GOOD = 30 + 33, BAD = 14 + 33;
    int corrected = 0, r1c = 0, r2c = 0;
    for(int i = 0; i < ol; ++i){
        const int p1 = start1 + i, p2 = start2 - i;
        if(seq1[p1] != complement(seq2[p2])){
            if(qual1[p1] >= GOOD && qual2[p2] <= BAD){
                seq2[p2] = complement(seq1[p1]); qual2[p2] = qual1[p1];
                ++corrected; r2c = 1;
                c->flat[c->L.correction_off + (seq2[p2] & 7) * 8 + (complement(seq1[p1]) & 7)]++;
                push_corr(c, pair, o2 + p2, 1, seq2[p2], qual2[p2]);
            }else if(qual2[p2] >= GOOD && qual1[p1] <= BAD){
                seq1[p1] = complement(seq2[p2]); qual1[p1] = qual2[p2];
                ++corrected; r1c = 1;
                c->flat[c->L.correction_off + (seq1[p1] & 7) * 8 + (complement(seq2[p2]) & 7)]++;
                push_corr(c, pair, o1 + p1, 0, seq1[p1], qual1[p1]);
            }
        }
    }
    if(corrected > 0) c->flat[c->L.corrected_reads_off] += (r1c && r2c) ? 2 : 1;
}

int fqo_get_corrections(fqo_ctx* c, fq_correction* out, uint32_t cap, uint32_t* n){
    *n = c->ncorr;
    if(out) memcpy(out, c->corr, (size_t)(c->ncorr < cap ? c->ncorr : cap) * sizeof(fq_correction));
    return 0;
}

/* ---------------------------------------------------------------- PE flow */
/* PairEndProcessor::statInsertSize, src/peprocessor.cpp:510-523 */
static int stat_insert_size(fqo_ctx* c, const rview* r1, const rview* r2, const ovres* ov){
    int isize = c->P.insert_size_max;
    if(ov->overlapped){
        if(ov->offset > 0) isize = r1->len + r2->len - ov->overlapLen;
        else isize = ov->overlapLen;
    }
    if(isize > c->P.insert_size_max) isize = c->P.insert_size_max;
    c->flat[c->L.insert_hist_off + isize]++;
    return isize;
}

/* body of the loop of PairEndProcessor::processPairEnd, src/peprocessor.cpp:271-438,
 * with threadId == 0 (the -w 1 oracle order) */
int fqo_process_pe(fqo_ctx* c, uint32_t n, uint64_t first_index,
                   const char* seq1, const char* qual1, const uint64_t* off1,
                   const char* seq2, const char* qual2, const uint64_t* off2,
                   fq_read_result* out1, fq_read_result* out2, fq_pair_result* outp){
    (void)first_index;
    const fq_params* P = &c->P;
    c->ncorr = 0;
    for(uint32_t p = 0; p < n; ++p){
        const char* s1 = seq1 + off1[p]; const char* q1 = qual1 + off1[p];
        const char* s2 = seq2 + off2[p]; const char* q2 = qual2 + off2[p];
        int l1 = (int)(off1[p + 1] - off1[p]), l2 = (int)(off2[p + 1] - off2[p]);
        if(P->correction_enabled){                                    /* the reference corrects its Reads in place */
            const size_t need = 2 * ((size_t)l1 + l2) + 4;
            if(need > c->cbuf_cap){ c->cbuf_cap = need * 2; c->cbuf = (char*)realloc(c->cbuf, c->cbuf_cap); }
            char* m = c->cbuf;
            memcpy(m, s1, l1); memcpy(m + l1 + 1, q1, l1); memcpy(m + 2 * l1 + 2, s2, l2); memcpy(m + 2 * l1 + l2 + 3, q2, l2);
        }
        fq_read_result* rec1 = &out1[p]; fq_read_result* rec2 = &out2[p];
        fq_pair_result* pr = &outp[p];
        rec_init(rec1); rec_init(rec2); memset(pr, 0, sizeof(*pr));
        pr->insert_size = 0xFFFF;
        int rc;
        if((rc = stat_read(c, 0, s1, q1, l1))) return rc;             /* :276 */
        if((rc = stat_read(c, 1, s2, q2, l2))) return rc;             /* :277 */
        if(P->dup_enabled) dup_stat_pair(c, s1, l1, s2, l2);          /* :279-281 */
        if(P->correction_enabled){ char* m = c->cbuf; s1 = m; q1 = m + l1 + 1; s2 = m + 2 * l1 + 2; q2 = m + 2 * l1 + l2 + 3; }
        rview r1 = { s1, q1, l1 }, r2 = { s2, q2, l2 };
        int a1 = trim_and_cut(P, &r1, P->trim_front1, P->trim_tail1); /* :292 */
        int a2 = trim_and_cut(P, &r2, P->trim_front2, P->trim_tail2); /* :293 */
        if(a1 && a2 && P->polyg_enabled){
            /* :297 passes (maxMismatch, oneEach, minLen) into (compareReq, maxMismatch, oneEach) */
            trim_polyg(c, &r1, P->polyg_max_mismatch, P->polyg_one_mismatch_each, P->polyg_min_len, rec1);
            trim_polyg(c, &r2, P->polyg_max_mismatch, P->polyg_one_mismatch_each, P->polyg_min_len, rec2);
        }
        int insertEvaluated = 0;
        if(a1 && a2 && (P->adapter_trim_enabled || P->correction_enabled)){   /* :302 */
            ovres ov = analyze(c, &r1, &r2);                          /* :303 */
            pr->flags |= FQK_P_OV1_DONE | (ov.overlapped ? FQK_P_OV1_OVERLAP : 0);
            pr->ov1_offset = (int16_t)ov.offset; pr->ov1_len = (uint16_t)ov.overlapLen; pr->ov1_diff = (uint16_t)ov.diff;
            pr->insert_size = (uint16_t)stat_insert_size(c, &r1, &r2, &ov);   /* :305-308 */
            insertEvaluated = 1;
            if(P->correction_enabled)                                 /* :310-312 */
                correct_by_overlap(c, (char*)r1.seq, (char*)r1.qual, r1.len, (char*)r2.seq, (char*)r2.qual, r2.len, &ov, p,
                                   (int)(r1.seq - s1), (int)(r2.seq - s2));
            if(P->adapter_trim_enabled){                              /* :314-326 */
                int trimmed = 0;
                int ol = ov.overlapLen;
                /* AdapterTrimmer::trimByOverlapAnalysis, src/adaptertrimmer.cpp:14-27 */
                if(ov.diff <= 5 && ov.overlapped && ov.offset < 0 && ol > r1.len / 3){
                    int t1 = r1.len - ol, t2 = r2.len - ol;
                    r1.len = ol; r2.len = ol;
                    add_adapter_trimmed2(c, t1, t2);
                    rec1->flags |= FQK_F_ADAPTER_OV; rec1->adapter_pos = (int16_t)ol; rec1->adapter_len = (uint16_t)t1;
                    rec2->flags |= FQK_F_ADAPTER_OV; rec2->adapter_pos = (int16_t)ol; rec2->adapter_len = (uint16_t)t2;
                    trimmed = 1;
                }
                if(!trimmed){
                    if(P->adapter_r1_len > 0) trim_by_sequence(c, &r1, P->adapter_r1, P->adapter_r1_len, rec1);
                    if(P->adapter_r2_len > 0) trim_by_sequence(c, &r2, P->adapter_r2, P->adapter_r2_len, rec2);
                }
            }
        }
        if(!insertEvaluated && a1 && a2){                             /* :329-333 */
            ovres ov = analyze(c, &r1, &r2);
            pr->flags |= FQK_P_OV1_DONE | (ov.overlapped ? FQK_P_OV1_OVERLAP : 0);
            pr->ov1_offset = (int16_t)ov.offset; pr->ov1_len = (uint16_t)ov.overlapLen; pr->ov1_diff = (uint16_t)ov.diff;
            pr->insert_size = (uint16_t)stat_insert_size(c, &r1, &r2, &ov);
        }
        if(a1 && a2 && P->polyx_enabled){                             /* :335-340 */
            trim_polyx(c, &r1, P->polyx_base_mask, P->polyx_min_len, P->polyx_max_mismatch, P->polyx_one_mismatch_each, rec1);
            trim_polyx(c, &r2, P->polyx_base_mask, P->polyx_min_len, P->polyx_max_mismatch, P->polyx_one_mismatch_each, rec2);
        }
        if(a1 && a2){                                                 /* :342-349 */
            if(P->trim_max_len1 > 0 && P->trim_max_len1 < r1.len) rv_resize(&r1, P->trim_max_len1);
            if(P->trim_max_len2 > 0 && P->trim_max_len2 < r2.len) rv_resize(&r2, P->trim_max_len2);
        }
        int mergeProcessed = 0, res1 = 0, res2 = 0;
        if(P->merge_enabled && a1 && a2){                             /* :353-385 */
            ovres ov = analyze(c, &r1, &r2);
            pr->flags |= FQK_P_OV2_DONE | (ov.overlapped ? FQK_P_OV2_OVERLAP : 0);
            pr->ov2_offset = (int16_t)ov.offset; pr->ov2_len = (uint16_t)ov.overlapLen; pr->ov2_diff = (uint16_t)ov.diff;
            if(ov.overlapped){
                /* OverlapAnalysis::merge, src/overlapanalysis.cpp:74-104 */
                int result;
                if(!ov.overlapLen){
                    result = FQK_FAIL_LENGTH;                          /* merge() == NULL -> passFilter(NULL) */
                }else{
                    int ol = ov.overlapLen;
                    int len1 = ol + imax(0, ov.offset);
                    int len2 = ov.offset > 0 ? r2.len - ol : 0;
                    int ml = len1 + len2;
                    char* ms = (char*)malloc((size_t)ml * 2 + 2 + (size_t)r2.len);
                    char* mq = ms + ml + 1;
                    char* rc2 = mq + ml + 1;
                    revcomp(r2.seq, r2.len, rc2);
                    memcpy(ms, r1.seq, (size_t)len1); memcpy(mq, r1.qual, (size_t)len1);
                    for(int i = 0; i < len2; ++i){
                        ms[len1 + i] = rc2[ol + i];
                        mq[len1 + i] = r2.qual[r2.len - 1 - (ol + i)];    /* reversed quality, read.h:95-100 */
                    }
                    rview m = { ms, mq, ml };
                    result = pass_filter(P, &m, 0);
                    pr->merge_len1 = (uint16_t)len1; pr->merge_len2 = (uint16_t)len2;
                    if(result == FQK_PASS_FILTER){
                        rc = stat_read(c, 2, ms, mq, ml);                 /* :361 postStats1 */
                        if(rc){ free(ms); return rc; }
                        pr->flags |= FQK_P_MERGED;
                    }
                    free(ms);
                }
                add_filter_result(c, result, 2);                      /* :358 */
                if(result == FQK_PASS_FILTER) c->flat[c->L.merged_pairs_off]++;   /* :363, :499-501 */
                pr->merged_result = (uint8_t)result;
                mergeProcessed = 1;
            }else if(!P->merge_discard_unmerged){                     /* :367-384 */
                res1 = pass_filter(P, &r1, 0);
                add_filter_result(c, res1, 1);
                if(res1 == FQK_PASS_FILTER) if((rc = stat_read(c, 2, r1.seq, r1.qual, r1.len))) return rc;
                res2 = pass_filter(P, &r2, 0);
                add_filter_result(c, res2, 1);
                if(res2 == FQK_PASS_FILTER) if((rc = stat_read(c, 3, r2.seq, r2.qual, r2.len))) return rc;
                mergeProcessed = 1;
            }
            if(mergeProcessed) pr->flags |= FQK_P_MERGE_PROC;
        }
        if(!mergeProcessed){                                          /* :387-429 */
            res1 = pass_filter(P, &r1, !a1);
            res2 = pass_filter(P, &r2, !a2);
            add_filter_result(c, imax(res1, res2), 2);                /* :390, mPaired true */
            if(a1 && res1 == FQK_PASS_FILTER && a2 && res2 == FQK_PASS_FILTER && !P->merge_enabled){
                if((rc = stat_read(c, 2, r1.seq, r1.qual, r1.len))) return rc;
                if((rc = stat_read(c, 3, r2.seq, r2.qual, r2.len))) return rc;
            }
        }else if(pr->flags & FQK_P_OV2_OVERLAP){
            /* merged path: per-read codes are not evaluated by the reference */
            res1 = res2 = pr->merged_result;
        }
        rec1->result = (uint8_t)res1; rec2->result = (uint8_t)res2;
        if(!a1){ rec1->flags |= FQK_F_DROPPED; } else { rec1->start = (uint16_t)(r1.seq - s1); rec1->len = (uint16_t)r1.len; }
        if(!a2){ rec2->flags |= FQK_F_DROPPED; } else { rec2->start = (uint16_t)(r2.seq - s2); rec2->len = (uint16_t)r2.len; }
    }
    return 0;
}

/* ------------------------------------------------------------ finalize etc */
int fqo_finalize(fqo_ctx* c, uint64_t* flat){
    memcpy(flat, c->flat, c->L.total_u64 * sizeof(uint64_t));
    return 0;
}
int fqo_dup_table(fqo_ctx* c, uint64_t* dups, uint32_t* counts, uint8_t* gc){
    if(!c->L.dup_bins) return FQK_E_INVALID;
    if(dups) memcpy(dups, c->dups, c->L.dup_bins * sizeof(uint64_t));
    if(counts) memcpy(counts, c->counts, c->L.dup_bins * sizeof(uint32_t));
    if(gc) memcpy(gc, c->gc, c->L.dup_bins);
    return 0;
}
/* Stats::merge src/stats.cpp:815-863 and FilterResult::merge src/filterresult.cpp:52-102
 * are plain sums; used only by the multi-thread CPU baseline (dup tables are
 * not merged there: the reference shares one mutex-guarded table). */
int fqo_merge(fqo_ctx* c, const fqo_ctx* o){
    if(c->L.total_u64 != o->L.total_u64) return FQK_E_INVALID;
    for(uint64_t i = 0; i < c->L.total_u64; ++i) c->flat[i] += o->flat[i];
    return 0;
}

/* ------------------------------------------------ adapter detection pre-pass */
/* Evaluator::seq2int, src/evaluator.cpp:3-49: rolling 2-bit key, -1 while the window holds a character outside ATCG */
static int ad_code(char c){ return c == 'A' ? 0 : c == 'T' ? 1 : c == 'C' ? 2 : c == 'G' ? 3 : -1; }
static int ad_seq2int(const char* s, int pos, int keylen, int last){
    if(last >= 0){
        const int mask = (1 << (keylen * 2)) - 1;
        const int c = ad_code(s[pos + keylen - 1]);
        return c < 0 ? -1 : (((last << 2) & mask) + c);
    }
    int key = 0;
    for(int i = pos; i < pos + keylen; ++i){
        const int c = ad_code(s[i]);
        if(c < 0) return -1;
        key = (key << 2) + c;
    }
    return key;
}
/* the counting loop of Evaluator::evaluateAdapterSeq, src/evaluator.cpp:271-283 */
int fqo_adapter_kmer_counts(uint32_t n, const char* seq, const uint64_t* off, int keylen, int first_pos, int shift_tail,
                            uint64_t* counts){
    if(keylen < 1 || keylen > 15) return FQK_E_INVALID;
    memset(counts, 0, sizeof(uint64_t) << (2 * keylen));
    for(uint32_t r = 0; r < n; ++r){
        const char* s = seq + off[r];
        const int len = (int)(off[r + 1] - off[r]);
        int key = -1;
        for(int pos = first_pos; pos <= len - keylen - shift_tail; ++pos){
            key = ad_seq2int(s, pos, keylen, key);
            if(key >= 0) ++counts[key];
        }
    }
    return 0;
}
/* the scan of Evaluator::getAdapterWithSeed, src/evaluator.cpp:396-407 */
int fqo_adapter_seed_hits(uint32_t n, const char* seq, const uint64_t* off, int keylen, int first_pos, int shift_tail,
                          uint32_t seed, uint32_t* read_idx, uint32_t* pos_out, uint32_t cap, uint32_t* nhits){
    uint32_t k = 0;
    for(uint32_t r = 0; r < n; ++r){
        const char* s = seq + off[r];
        const int len = (int)(off[r + 1] - off[r]);
        int key = -1;
        for(int pos = first_pos; pos <= len - keylen - shift_tail; ++pos){
            key = ad_seq2int(s, pos, keylen, key);
            if(key >= 0 && (uint32_t)key == seed){
                if(k < cap){ read_idx[k] = r; pos_out[k] = (uint32_t)pos; }
                ++k;
            }
        }
    }
    *nhits = k;
    return 0;
}
