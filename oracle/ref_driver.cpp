/*
 * ref_driver.cpp -- TEST INFRASTRUCTURE ONLY.
 *
 * Thin C API over the reference's own, unmodified hot-path sources.  This file
 * is ours; it is compiled together with /root/reference/src/{filter,polyx,
 * adaptertrimmer,overlapanalysis,stats,duplicate,filterresult,evaluator,
 * options,read,fqreader,nucleotidetree}.cpp where they lie (oracle/Makefile)
 * into oracle/_ref/libfqref.so.  It runs the exact call sequence of
 *   SingleEndProcessor::processSingleEnd  src/seprocessor.cpp:294-354
 *   PairEndProcessor::processPairEnd      src/peprocessor.cpp:271-438
 * by calling the reference's functions, observes the Read before/after each
 * call to fill the same records as the C ABI, and dumps the raw Stats /
 * FilterResult / Duplicate arrays in the fq_stats_layout order.
 *
 * Uses: (1) pin oracle/fq_oracle.c, (2) generate tests/golden/, (3) the timed
 * multi-threaded CPU baseline ("kind": "reference") of bench.py.
 */
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#include <zlib.h>
#include "json.hpp"
#include "ctml.hpp"
/* the raw per-cycle arrays of Stats and the tables of Duplicate are private
 * (src/stats.h:21-50, src/duplicate.h:61-69); expose them to this TU only */
#define private public
#include "stats.h"
#include "duplicate.h"
#undef private
#include "filter.h"
#include "polyx.h"
#include "adaptertrimmer.h"
#include "overlapanalysis.h"
#include "filterresult.h"
#include "basecorrector.h"
#include "fq_oracle.h"

namespace {

struct RefCtx {
    fq_params P;
    fq_stats_layout L;
    Options* opt;
    Filter* filter;
    Stats* st[4];
    FilterResult* fr;
    Duplicate* dup;
    std::vector<uint64_t> insertHist;
    std::string adapter1, adapter2;
    std::vector<fq_correction> corr;
};

void fill_options(Options* o, const fq_params* p){
    if(p->is_paired) o->in2 = "paired";                 // Options::isPaired() src/options.cpp:73
    o->trim.front1 = p->trim_front1; o->trim.tail1 = p->trim_tail1; o->trim.maxLen1 = p->trim_max_len1;
    o->trim.front2 = p->trim_front2; o->trim.tail2 = p->trim_tail2; o->trim.maxLen2 = p->trim_max_len2;
    o->qualitycut.enableFront = p->cut_front_enabled; o->qualitycut.enableTail = p->cut_tail_enabled;
    o->qualitycut.enableRright = p->cut_right_enabled;
    o->qualitycut.windowSizeFront = p->cut_front_window; o->qualitycut.windowSizeTail = p->cut_tail_window;
    o->qualitycut.windowSizeRight = p->cut_right_window;
    o->qualitycut.qualityFront = p->cut_front_qual; o->qualitycut.qualityTail = p->cut_tail_qual;
    o->qualitycut.qualityRight = p->cut_right_qual;
    o->polyGTrim.enabled = p->polyg_enabled; o->polyGTrim.minLen = p->polyg_min_len;
    o->polyGTrim.maxMismatch = p->polyg_max_mismatch; o->polyGTrim.allowedOneMismatchForEach = p->polyg_one_mismatch_each;
    o->polyXTrim.enabled = p->polyx_enabled;
    o->polyXTrim.trimChr = "";
    const char* b = "ATCGN";
    for(int i = 0; i < 5; ++i) if((p->polyx_base_mask >> i) & 1) o->polyXTrim.trimChr += b[i];
    o->polyXTrim.minLen = p->polyx_min_len; o->polyXTrim.maxMismatch = p->polyx_max_mismatch;
    o->polyXTrim.allowedOneMismatchForEach = p->polyx_one_mismatch_each;
    o->qualFilter.enabled = p->qual_filter_enabled; o->qualFilter.lowQualityLimit = p->low_quality_limit;
    o->qualFilter.lowQualityBaseLimit = p->low_quality_base_limit; o->qualFilter.nBaseLimit = p->n_base_limit;
    o->qualFilter.averageQualityLimit = p->average_quality_limit;
    o->lengthFilter.enabled = p->length_filter_enabled; o->lengthFilter.minReadLength = p->min_read_length;
    o->lengthFilter.maxReadLength = p->max_read_length;
    o->complexityFilter.enabled = p->complexity_enabled; o->complexityFilter.threshold = p->complexity_threshold;
    o->adapter.enableTriming = p->adapter_trim_enabled;
    o->adapter.inputAdapterSeqR1 = std::string(p->adapter_r1, p->adapter_r1_len);
    o->adapter.inputAdapterSeqR2 = std::string(p->adapter_r2, p->adapter_r2_len);
    o->adapter.adapterSeqR1Provided = p->adapter_r1_len > 0;
    o->adapter.adapterSeqR2Provided = p->adapter_r2_len > 0;
    o->overlapRequire = p->overlap_require; o->overlapDiffLimit = p->overlap_diff_limit;
    o->insertSizeMax = p->insert_size_max;
    o->correction.enabled = p->correction_enabled;
    o->mergePE.enabled = p->merge_enabled; o->mergePE.discardUnmerged = p->merge_discard_unmerged;
    o->kmer.enabled = p->kmer_len > 0; o->kmer.kmerLen = p->kmer_len;
    o->duplicate.enabled = p->dup_enabled; o->duplicate.keylen = p->dup_keylen;
    o->overRepAna.enabled = false;
    o->indexFilter.enabled = false; o->umi.enabled = false;
    int C = p->max_cycles > 0 ? p->max_cycles : 512;
    o->est.seqLen1 = C; o->est.seqLen2 = C;             // Stats buffer = seqLen + 1024, src/stats.cpp:12-17
}

/* position markers: a clone of the read whose bases are replaced by symbols that
 * encode their own index (never 'N', never 0); Filter::trimAndCut only tests
 * seq[i] == 'N' (src/filter.cpp:117,177), so the clone is cut exactly like the
 * original and its first surviving symbol reveals the cut position */
inline char marker(int v){ int c = v + 1; if(c >= 'N') c += 1; return (char)(unsigned char)c; }
inline int unmarker(char ch){ int c = (unsigned char)ch; return c > 'N' ? c - 2 : c - 1; }

/* Filter::trimAndCut on the real read + start recovery; returns the reference's pointer */
Read* ref_trim_and_cut(RefCtx* c, Read* r, int F, int T, int* start){
    *start = 0;
    if(!c->P.cut_front_enabled){
        Read* out = c->filter->trimAndCut(r, F, T);
        if(out) *start = F;          // src/filter.cpp:71-87 and the F' == F paths of :125-188
        return out;
    }
    int L = r->length();
    std::string lo(L, 'A'), hi(L, 'A');
    for(int i = 0; i < L; ++i){
        if(r->seq.seqStr[i] == 'N'){ lo[i] = 'N'; hi[i] = 'N'; }
        else { lo[i] = marker(i % 200); hi[i] = marker(i / 200); }
    }
    Read clo(r->name, lo, r->strand, r->quality), chi(r->name, hi, r->strand, r->quality);
    Read* a = c->filter->trimAndCut(&clo, F, T);
    Read* b = c->filter->trimAndCut(&chi, F, T);
    Read* out = c->filter->trimAndCut(r, F, T);
    if((a == NULL) != (out == NULL) || (b == NULL) != (out == NULL)){ fprintf(stderr, "ref_driver: marker clone diverged\n"); abort(); }
    if(out){
        if(a->length() != out->length() || a->length() == 0 || a->seq.seqStr[0] == 'N'){ fprintf(stderr, "ref_driver: marker decode failed\n"); abort(); }
        *start = unmarker(a->seq.seqStr[0]) + 200 * unmarker(b->seq.seqStr[0]);
    }
    return out;
}

void set_ov(fq_pair_result* pr, const OverlapResult& ov, bool first){
    if(first){
        pr->flags |= FQK_P_OV1_DONE | (ov.overlapped ? FQK_P_OV1_OVERLAP : 0);
        pr->ov1_offset = (int16_t)ov.offset; pr->ov1_len = (uint16_t)ov.overlapLen; pr->ov1_diff = (uint16_t)ov.diff;
    }else{
        pr->flags |= FQK_P_OV2_DONE | (ov.overlapped ? FQK_P_OV2_OVERLAP : 0);
        pr->ov2_offset = (int16_t)ov.offset; pr->ov2_len = (uint16_t)ov.overlapLen; pr->ov2_diff = (uint16_t)ov.diff;
    }
}

/* PolyX::trimPolyG with the tally observed through FilterResult's public counters */
void ref_polyg(RefCtx* c, Read* r, int a, int b, int d, fq_read_result* rec){
    size_t reads0 = c->fr->mTrimmedPolyXReads[3], bases0 = c->fr->mTrimmedPolyXBases[3];
    PolyX::trimPolyG(r, a, b, d, c->fr);
    if(c->fr->mTrimmedPolyXReads[3] != reads0){
        rec->flags |= FQK_F_POLYG;
        rec->polyg_len = (uint16_t)(c->fr->mTrimmedPolyXBases[3] - bases0);
    }
}

void ref_polyx(RefCtx* c, Read* r, fq_read_result* rec){
    size_t reads0[5], bases0[5];
    for(int i = 0; i < 5; ++i){ reads0[i] = c->fr->mTrimmedPolyXReads[i]; bases0[i] = c->fr->mTrimmedPolyXBases[i]; }
    PolyX::trimPolyX(r, c->opt->polyXTrim.trimChr, c->opt->polyXTrim.minLen, c->opt->polyXTrim.maxMismatch,
                     c->opt->polyXTrim.allowedOneMismatchForEach, c->fr);
    for(int i = 0; i < 5; ++i) if(c->fr->mTrimmedPolyXReads[i] != reads0[i]){
        rec->flags |= FQK_F_POLYX | (uint8_t)(i << FQK_F_POLYX_SHIFT);
        rec->polyx_len = (uint16_t)(c->fr->mTrimmedPolyXBases[i] - bases0[i]);
    }
}

/* AdapterTrimmer::trimBySequence through a scratch FilterResult so that the tallied
 * text (src/adaptertrimmer.cpp:73,80) is visible; totals are folded back */
void ref_trim_by_seq(RefCtx* c, Read* r, std::string& adapter, bool isR2, fq_read_result* rec){
    FilterResult tmp(c->opt, c->P.is_paired);
    int oldlen = r->length();
    bool hit = AdapterTrimmer::trimBySequence(r, &tmp, adapter, isR2);
    if(!hit) return;
    c->fr->mTrimmedAdapterReads += tmp.mTrimmedAdapterReads;
    c->fr->mTrimmedAdapterBases += tmp.mTrimmedAdapterBases;
    std::map<std::string, size_t>& m = isR2 ? tmp.mAdapter2Count : tmp.mAdapter1Count;
    std::string text = m.empty() ? std::string() : m.begin()->first;
    int textlen = (int)text.length(), alen = (int)adapter.length();
    int pos = r->length();
    int minstart = alen >= 16 ? -4 : (alen >= 12 ? -3 : (alen >= 8 ? -2 : 0));   // src/adaptertrimmer.cpp:43-51
    if(r->length() == 0 && textlen < alen && textlen - alen >= minstart && text == adapter.substr(alen - textlen))
        pos = textlen - alen;      // the scan reaches negative positions first (src/adaptertrimmer.cpp:52)
    (void)oldlen;
    rec->flags |= FQK_F_ADAPTER_SEQ;
    rec->adapter_pos = (int16_t)pos;
    rec->adapter_len = (uint16_t)textlen;
}

int stat_insert(RefCtx* c, Read* r1, Read* r2, OverlapResult& ov){   // src/peprocessor.cpp:510-523
    int isize = c->opt->insertSizeMax;
    if(ov.overlapped){
        if(ov.offset > 0) isize = r1->length() + r2->length() - ov.overlapLen;
        else isize = ov.overlapLen;
    }
    if(isize > c->opt->insertSizeMax) isize = c->opt->insertSizeMax;
    ++c->insertHist[isize];
    return isize;
}

} // namespace

/* ------------------------------------------------------------------ C API */
extern "C" {

void* fqr_create(const fq_params* p){
    RefCtx* c = new RefCtx();
    c->P = *p;
    fqo_layout(p, &c->L);
    c->opt = new Options();
    fill_options(c->opt, p);
    c->filter = new Filter(c->opt);
    c->st[0] = new Stats(c->opt, false); c->st[2] = new Stats(c->opt, false);
    c->st[1] = new Stats(c->opt, true);  c->st[3] = new Stats(c->opt, true);
    c->fr = new FilterResult(c->opt, p->is_paired);
    c->dup = p->dup_enabled ? new Duplicate(c->opt) : NULL;
    c->insertHist.assign((size_t)p->insert_size_max + 1, 0);
    c->adapter1 = c->opt->adapter.inputAdapterSeqR1; c->adapter2 = c->opt->adapter.inputAdapterSeqR2;
    return c;
}

void fqr_destroy(void* h){
    RefCtx* c = (RefCtx*)h;
    if(!c) return;
    for(int i = 0; i < 4; ++i) delete c->st[i];
    delete c->fr; delete c->dup; delete c->filter; delete c->opt; delete c;
}

/* loop body of SingleEndProcessor::processSingleEnd, src/seprocessor.cpp:294-354 */
int fqr_process_se(void* h, uint32_t n, uint64_t first_index, const char* seq, const char* qual,
                   const uint64_t* off, fq_read_result* out){
    (void)first_index;
    RefCtx* c = (RefCtx*)h; Options* o = c->opt;
    for(uint32_t p = 0; p < n; ++p){
        size_t len0 = off[p + 1] - off[p];
        Read* or1 = new Read("@r", std::string(seq + off[p], len0), "+", std::string(qual + off[p], len0));
        fq_read_result* rec = &out[p]; memset(rec, 0, sizeof(*rec));
        c->st[0]->statRead(or1);
        if(o->duplicate.enabled) c->dup->statRead(or1);
        int start = 0;
        Read* r1 = ref_trim_and_cut(c, or1, o->trim.front1, o->trim.tail1, &start);
        if(r1 != NULL && o->polyGTrim.enabled)
            ref_polyg(c, r1, o->polyGTrim.minLen, o->polyGTrim.maxMismatch, o->polyGTrim.allowedOneMismatchForEach, rec);
        if(r1 != NULL && o->adapter.enableTriming && o->adapter.adapterSeqR1Provided)
            ref_trim_by_seq(c, r1, o->adapter.inputAdapterSeqR1, false, rec);
        if(r1 != NULL && o->polyXTrim.enabled) ref_polyx(c, r1, rec);
        if(r1 != NULL && o->trim.maxLen1 > 0 && o->trim.maxLen1 < r1->length()) r1->resize(o->trim.maxLen1);
        int result = c->filter->passFilter(r1);
        c->fr->addFilterResult(result);
        if(r1 != NULL && result == COMMONCONST::PASS_FILTER) c->st[2]->statRead(r1);
        rec->result = (uint8_t)result;
        if(r1 == NULL) rec->flags |= FQK_F_DROPPED;
        else { rec->start = (uint16_t)start; rec->len = (uint16_t)r1->length(); }
        delete or1;
    }
    return 0;
}

/* loop body of PairEndProcessor::processPairEnd, src/peprocessor.cpp:271-438, threadId 0 */
int fqr_process_pe(void* h, uint32_t n, uint64_t first_index,
                   const char* seq1, const char* qual1, const uint64_t* off1,
                   const char* seq2, const char* qual2, const uint64_t* off2,
                   fq_read_result* out1, fq_read_result* out2, fq_pair_result* outp){
    (void)first_index;
    RefCtx* c = (RefCtx*)h; Options* o = c->opt;
    c->corr.clear();
    for(uint32_t p = 0; p < n; ++p){
        size_t l1 = off1[p + 1] - off1[p], l2 = off2[p + 1] - off2[p];
        Read* or1 = new Read("@r 1", std::string(seq1 + off1[p], l1), "+", std::string(qual1 + off1[p], l1));
        Read* or2 = new Read("@r 2", std::string(seq2 + off2[p], l2), "+", std::string(qual2 + off2[p], l2));
        fq_read_result* rec1 = &out1[p]; fq_read_result* rec2 = &out2[p]; fq_pair_result* pr = &outp[p];
        memset(rec1, 0, sizeof(*rec1)); memset(rec2, 0, sizeof(*rec2)); memset(pr, 0, sizeof(*pr));
        pr->insert_size = 0xFFFF;
        c->st[0]->statRead(or1); c->st[1]->statRead(or2);
        if(o->duplicate.enabled) c->dup->statPair(or1, or2);
        int start1 = 0, start2 = 0;
        Read* r1 = ref_trim_and_cut(c, or1, o->trim.front1, o->trim.tail1, &start1);
        Read* r2 = ref_trim_and_cut(c, or2, o->trim.front2, o->trim.tail2, &start2);
        if(r1 && r2 && o->polyGTrim.enabled){
            /* same (swapped) argument order as src/peprocessor.cpp:297 */
            ref_polyg(c, r1, o->polyGTrim.maxMismatch, o->polyGTrim.allowedOneMismatchForEach, o->polyGTrim.minLen, rec1);
            ref_polyg(c, r2, o->polyGTrim.maxMismatch, o->polyGTrim.allowedOneMismatchForEach, o->polyGTrim.minLen, rec2);
        }
        bool insertSizeEvaluated = false;
        if(r1 && r2 && (o->adapter.enableTriming || o->correction.enabled)){
            OverlapResult ov = OverlapAnalysis::analyze(r1, r2, o->overlapDiffLimit, o->overlapRequire);
            set_ov(pr, ov, true);
            pr->insert_size = (uint16_t)stat_insert(c, r1, r2, ov);
            insertSizeEvaluated = true;
            if(o->correction.enabled){
                // BaseCorrector mutates the reads in place; the edits are found by comparing before/after
                const std::string s1b = r1->seq.seqStr, s2b = r2->seq.seqStr, q1b = r1->quality, q2b = r2->quality;
                BaseCorrector::correctByOverlapAnalysis(r1, r2, c->fr, ov);
                for(size_t i = 0; i < s1b.size(); ++i) if(s1b[i] != r1->seq.seqStr[i] || q1b[i] != r1->quality[i]){
                    fq_correction e; e.pair = p; e.pos = (uint16_t)(start1 + i); e.mate_base = (uint8_t)(r1->seq.seqStr[i] & 0x7F); e.qual = (uint8_t)r1->quality[i];
                    c->corr.push_back(e);
                }
                for(size_t i = 0; i < s2b.size(); ++i) if(s2b[i] != r2->seq.seqStr[i] || q2b[i] != r2->quality[i]){
                    fq_correction e; e.pair = p; e.pos = (uint16_t)(start2 + i); e.mate_base = (uint8_t)(0x80 | (r2->seq.seqStr[i] & 0x7F)); e.qual = (uint8_t)r2->quality[i];
                    c->corr.push_back(e);
                }
            }
            if(o->adapter.enableTriming){
                int b1 = r1->length(), b2 = r2->length();
                bool trimmed = AdapterTrimmer::trimByOverlapAnalysis(r1, r2, c->fr, ov);
                if(trimmed){
                    rec1->flags |= FQK_F_ADAPTER_OV; rec1->adapter_pos = (int16_t)ov.overlapLen; rec1->adapter_len = (uint16_t)(b1 - r1->length());
                    rec2->flags |= FQK_F_ADAPTER_OV; rec2->adapter_pos = (int16_t)ov.overlapLen; rec2->adapter_len = (uint16_t)(b2 - r2->length());
                }else{
                    if(o->adapter.adapterSeqR1Provided) ref_trim_by_seq(c, r1, o->adapter.inputAdapterSeqR1, false, rec1);
                    if(o->adapter.adapterSeqR2Provided) ref_trim_by_seq(c, r2, o->adapter.inputAdapterSeqR2, true, rec2);
                }
            }
        }
        if(!insertSizeEvaluated && r1 && r2){
            OverlapResult ov = OverlapAnalysis::analyze(r1, r2, o->overlapDiffLimit, o->overlapRequire);
            set_ov(pr, ov, true);
            pr->insert_size = (uint16_t)stat_insert(c, r1, r2, ov);
            insertSizeEvaluated = true;
        }
        if(r1 && r2 && o->polyXTrim.enabled){ ref_polyx(c, r1, rec1); ref_polyx(c, r2, rec2); }
        if(r1 && r2){
            if(o->trim.maxLen1 > 0 && o->trim.maxLen1 < r1->length()) r1->resize(o->trim.maxLen1);
            if(o->trim.maxLen2 > 0 && o->trim.maxLen2 < r2->length()) r2->resize(o->trim.maxLen2);
        }
        bool mergeProcessed = false; int result1 = 0, result2 = 0;
        if(o->mergePE.enabled && r1 && r2){
            OverlapResult ov = OverlapAnalysis::analyze(r1, r2, o->overlapDiffLimit, o->overlapRequire);
            set_ov(pr, ov, false);
            if(ov.overlapped){
                Read* merged = OverlapAnalysis::merge(r1, r2, ov);
                int result = c->filter->passFilter(merged);
                c->fr->addFilterResult(result, 2);
                if(merged){
                    int ol = ov.overlapLen;
                    pr->merge_len1 = (uint16_t)(ol + std::max(0, ov.offset));
                    pr->merge_len2 = (uint16_t)(merged->length() - (ol + std::max(0, ov.offset)));
                }
                if(result == COMMONCONST::PASS_FILTER){
                    c->st[2]->statRead(merged);
                    c->fr->addMergedPairs(1);
                    pr->flags |= FQK_P_MERGED;
                }
                delete merged;
                pr->merged_result = (uint8_t)result;
                result1 = result2 = result;
                mergeProcessed = true;
            }else if(!o->mergePE.discardUnmerged){
                result1 = c->filter->passFilter(r1);
                c->fr->addFilterResult(result1, 1);
                if(result1 == COMMONCONST::PASS_FILTER) c->st[2]->statRead(r1);
                result2 = c->filter->passFilter(r2);
                c->fr->addFilterResult(result2, 1);
                if(result2 == COMMONCONST::PASS_FILTER) c->st[3]->statRead(r2);
                mergeProcessed = true;
            }
            if(mergeProcessed) pr->flags |= FQK_P_MERGE_PROC;
        }
        if(!mergeProcessed){
            result1 = c->filter->passFilter(r1);
            result2 = c->filter->passFilter(r2);
            c->fr->addFilterResult(std::max(result1, result2));
            if(r1 && result1 == COMMONCONST::PASS_FILTER && r2 && result2 == COMMONCONST::PASS_FILTER){
                if(!o->mergePE.enabled){ c->st[2]->statRead(r1); c->st[3]->statRead(r2); }
            }
        }
        rec1->result = (uint8_t)result1; rec2->result = (uint8_t)result2;
        if(!r1) rec1->flags |= FQK_F_DROPPED; else { rec1->start = (uint16_t)start1; rec1->len = (uint16_t)r1->length(); }
        if(!r2) rec2->flags |= FQK_F_DROPPED; else { rec2->start = (uint16_t)start2; rec2->len = (uint16_t)r2->length(); }
        delete or1; delete or2;
    }
    return 0;
}

static void dump_stats(RefCtx* c, int which, uint64_t* flat){
    const fq_stats_layout* L = &c->L;
    Stats* s = c->st[which];
    uint64_t* S = flat + L->stats_off[which];
    uint64_t C = L->max_cycles;
    uint64_t nc = std::min<uint64_t>(C, (uint64_t)s->mBufLen);
    for(int b = 0; b < 8; ++b) for(uint64_t i = 0; i < nc; ++i){
        S[L->s_q20 + b * C + i] = s->mCycleQ20Bases[b][i];
        S[L->s_q30 + b * C + i] = s->mCycleQ30Bases[b][i];
        S[L->s_content + b * C + i] = s->mCycleBaseContents[b][i];
        S[L->s_qualsum + b * C + i] = s->mCycleBaseQuality[b][i];
    }
    for(uint64_t i = 0; i < nc; ++i){
        S[L->s_total_base + i] = s->mCycleTotalBase[i];
        S[L->s_total_qual + i] = s->mCycleTotalQuality[i];
    }
    S[L->s_reads] = s->mReads;
    S[L->s_length_sum] = s->mLengthSum;
    for(uint64_t i = 0; i < L->kmer_bins; ++i) S[L->s_kmer + i] = s->mKmer[i];
}

int fqr_finalize(void* h, uint64_t* flat){
    RefCtx* c = (RefCtx*)h; const fq_stats_layout* L = &c->L;
    memset(flat, 0, L->total_u64 * sizeof(uint64_t));
    for(int i = 0; i < 4; ++i) dump_stats(c, i, flat);
    for(int i = 0; i < FQK_FILTER_RESULT_TYPES; ++i) flat[L->filter_off + i] = c->fr->mFilterReadStats[i];
    flat[L->adapter_reads_off] = c->fr->mTrimmedAdapterReads;
    flat[L->adapter_bases_off] = c->fr->mTrimmedAdapterBases;
    for(int i = 0; i < 5; ++i){
        flat[L->polyx_reads_off + i] = c->fr->mTrimmedPolyXReads[i];
        flat[L->polyx_bases_off + i] = c->fr->mTrimmedPolyXBases[i];
    }
    flat[L->merged_pairs_off] = c->fr->mMergedPairs;
    for(uint64_t i = 0; i < L->insert_hist_len; ++i) flat[L->insert_hist_off + i] = c->insertHist[i];
    for(int i = 0; i < 64; ++i) flat[L->correction_off + i] = c->fr->getCorrectionMatrix()[i];
    flat[L->corrected_reads_off] = c->fr->mCorrectedReads;
    return 0;
}

int fqr_get_corrections(void* h, fq_correction* out, uint32_t cap, uint32_t* n){
    RefCtx* c = (RefCtx*)h;
    *n = (uint32_t)c->corr.size();
    if(out) memcpy(out, c->corr.data(), std::min<size_t>(cap, c->corr.size()) * sizeof(fq_correction));
    return 0;
}

int fqr_dup_table(void* h, uint64_t* dups, uint32_t* counts, uint8_t* gc){
    RefCtx* c = (RefCtx*)h;
    if(!c->dup) return FQK_E_INVALID;
    uint64_t n = c->L.dup_bins;
    if(dups) memcpy(dups, c->dup->mDups, n * sizeof(uint64_t));
    if(counts) memcpy(counts, c->dup->mCounts, n * sizeof(uint32_t));
    if(gc) memcpy(gc, c->dup->mGC, n);
    return 0;
}

} // extern "C"

/* ------------------------------------------------ timed multi-thread baseline */
/*
 * The reference's hot path on T host threads: reads are pre-parsed into Read
 * objects (outside the timed region, as the reference's producer thread does),
 * then every thread runs the per-read call sequence of processSingleEnd /
 * processPairEnd on its contiguous share with a private pre/post Stats and
 * FilterResult (src/threadconfig.cpp:3-20) and the shared mutex-guarded
 * Duplicate table (src/duplicate.cpp:46-62).  Output string assembly and the
 * writer hand-off (src/seprocessor.cpp:343,356-379) are NOT included, which
 * favours the reference.  Returns the wall seconds of the slowest thread span.
 */
extern "C" int fqr_bench(const fq_params* p, int threads, uint32_t n,
                         const char* seq1, const char* qual1, const uint64_t* off1,
                         const char* seq2, const char* qual2, const uint64_t* off2,
                         double* seconds, uint64_t* passed_out){
    Options opt; fill_options(&opt, p);
    Filter filter(&opt);
    Duplicate* dup = p->dup_enabled ? new Duplicate(&opt) : NULL;
    bool paired = p->is_paired != 0;
    std::vector<Read*> R1(n), R2(paired ? n : 0);
    for(uint32_t i = 0; i < n; ++i){
        size_t l1 = off1[i + 1] - off1[i];
        R1[i] = new Read("@r 1", std::string(seq1 + off1[i], l1), "+", std::string(qual1 + off1[i], l1));
        if(paired){
            size_t l2 = off2[i + 1] - off2[i];
            R2[i] = new Read("@r 2", std::string(seq2 + off2[i], l2), "+", std::string(qual2 + off2[i], l2));
        }
    }
    std::vector<std::thread> th;
    std::vector<uint64_t> passed(threads, 0);
    std::atomic<int> ready(0); std::atomic<bool> go(false);
    auto worker = [&](int t){
        Stats pre1(&opt, false), post1(&opt, false), pre2(&opt, true), post2(&opt, true);
        FilterResult fr(&opt, paired);
        std::vector<long> ins((size_t)opt.insertSizeMax + 1, 0);
        uint32_t lo = (uint32_t)((uint64_t)n * t / threads), hi = (uint32_t)((uint64_t)n * (t + 1) / threads);
        ready++; while(!go.load()) std::this_thread::yield();
        uint64_t np = 0;
        for(uint32_t i = lo; i < hi; ++i){
            if(!paired){
                Read* or1 = R1[i];
                pre1.statRead(or1);
                if(dup) dup->statRead(or1);
                Read* r1 = filter.trimAndCut(or1, opt.trim.front1, opt.trim.tail1);
                if(r1 && opt.polyGTrim.enabled) PolyX::trimPolyG(r1, opt.polyGTrim.minLen, opt.polyGTrim.maxMismatch, opt.polyGTrim.allowedOneMismatchForEach, &fr);
                if(r1 && opt.adapter.enableTriming && opt.adapter.adapterSeqR1Provided) AdapterTrimmer::trimBySequence(r1, &fr, opt.adapter.inputAdapterSeqR1);
                if(r1 && opt.polyXTrim.enabled) PolyX::trimPolyX(r1, opt.polyXTrim.trimChr, opt.polyXTrim.minLen, opt.polyXTrim.maxMismatch, opt.polyXTrim.allowedOneMismatchForEach, &fr);
                if(r1 && opt.trim.maxLen1 > 0 && opt.trim.maxLen1 < r1->length()) r1->resize(opt.trim.maxLen1);
                int result = filter.passFilter(r1);
                fr.addFilterResult(result);
                if(r1 && result == COMMONCONST::PASS_FILTER){ post1.statRead(r1); ++np; }
            }else{
                Read* or1 = R1[i]; Read* or2 = R2[i];
                pre1.statRead(or1); pre2.statRead(or2);
                if(dup) dup->statPair(or1, or2);
                Read* r1 = filter.trimAndCut(or1, opt.trim.front1, opt.trim.tail1);
                Read* r2 = filter.trimAndCut(or2, opt.trim.front2, opt.trim.tail2);
                if(r1 && r2 && opt.polyGTrim.enabled)
                    PolyX::trimPolyG(r1, r2, opt.polyGTrim.maxMismatch, opt.polyGTrim.allowedOneMismatchForEach, opt.polyGTrim.minLen, &fr);
                bool ise = false;
                if(r1 && r2 && (opt.adapter.enableTriming || opt.correction.enabled)){
                    OverlapResult ov = OverlapAnalysis::analyze(r1, r2, opt.overlapDiffLimit, opt.overlapRequire);
                    if(t == 0){ int is = opt.insertSizeMax; if(ov.overlapped) is = ov.offset > 0 ? r1->length() + r2->length() - ov.overlapLen : ov.overlapLen; if(is > opt.insertSizeMax) is = opt.insertSizeMax; ++ins[is]; ise = true; }
                    if(opt.adapter.enableTriming){
                        bool trimmed = AdapterTrimmer::trimByOverlapAnalysis(r1, r2, &fr, ov);
                        if(!trimmed){
                            if(opt.adapter.adapterSeqR1Provided) AdapterTrimmer::trimBySequence(r1, &fr, opt.adapter.inputAdapterSeqR1, false);
                            if(opt.adapter.adapterSeqR2Provided) AdapterTrimmer::trimBySequence(r2, &fr, opt.adapter.inputAdapterSeqR2, true);
                        }
                    }
                }
                if(t == 0 && !ise && r1 && r2){
                    OverlapResult ov = OverlapAnalysis::analyze(r1, r2, opt.overlapDiffLimit, opt.overlapRequire);
                    int is = opt.insertSizeMax; if(ov.overlapped) is = ov.offset > 0 ? r1->length() + r2->length() - ov.overlapLen : ov.overlapLen; if(is > opt.insertSizeMax) is = opt.insertSizeMax; ++ins[is];
                }
                if(r1 && r2 && opt.polyXTrim.enabled) PolyX::trimPolyX(r1, r2, opt.polyXTrim.trimChr, opt.polyXTrim.minLen, opt.polyXTrim.maxMismatch, opt.polyXTrim.allowedOneMismatchForEach, &fr);
                if(r1 && r2){
                    if(opt.trim.maxLen1 > 0 && opt.trim.maxLen1 < r1->length()) r1->resize(opt.trim.maxLen1);
                    if(opt.trim.maxLen2 > 0 && opt.trim.maxLen2 < r2->length()) r2->resize(opt.trim.maxLen2);
                }
                bool mp = false;
                if(opt.mergePE.enabled && r1 && r2){
                    OverlapResult ov = OverlapAnalysis::analyze(r1, r2, opt.overlapDiffLimit, opt.overlapRequire);
                    if(ov.overlapped){
                        Read* merged = OverlapAnalysis::merge(r1, r2, ov);
                        int result = filter.passFilter(merged);
                        fr.addFilterResult(result, 2);
                        if(result == COMMONCONST::PASS_FILTER){ post1.statRead(merged); ++np; }
                        delete merged; mp = true;
                    }else if(!opt.mergePE.discardUnmerged){
                        int a = filter.passFilter(r1); fr.addFilterResult(a, 1); if(a == COMMONCONST::PASS_FILTER) post1.statRead(r1);
                        int b = filter.passFilter(r2); fr.addFilterResult(b, 1); if(b == COMMONCONST::PASS_FILTER) post2.statRead(r2);
                        if(a == COMMONCONST::PASS_FILTER && b == COMMONCONST::PASS_FILTER) ++np;
                        mp = true;
                    }
                }
                if(!mp){
                    int a = filter.passFilter(r1), b = filter.passFilter(r2);
                    fr.addFilterResult(std::max(a, b));
                    if(r1 && a == COMMONCONST::PASS_FILTER && r2 && b == COMMONCONST::PASS_FILTER){
                        if(!opt.mergePE.enabled){ post1.statRead(r1); post2.statRead(r2); }
                        ++np;
                    }
                }
            }
        }
        passed[t] = np;
    };
    for(int t = 0; t < threads; ++t) th.emplace_back(worker, t);
    while(ready.load() < threads) std::this_thread::yield();
    auto t0 = std::chrono::steady_clock::now();
    go.store(true);
    for(auto& x : th) x.join();
    auto t1 = std::chrono::steady_clock::now();
    *seconds = std::chrono::duration<double>(t1 - t0).count();
    uint64_t tot = 0; for(auto v : passed) tot += v;
    if(passed_out) *passed_out = tot;
    for(auto r : R1) delete r;
    for(auto r : R2) delete r;
    delete dup;
    return 0;
}

// ---------------------------------------------------------------- adapter detection pre-pass (reference functions)
// counts with the reference's own Evaluator::seq2int in the loop of Evaluator::evaluateAdapterSeq (src/evaluator.cpp:271-283)
extern "C" int fqr_adapter_kmer_counts(uint32_t n, const char* seq, const uint64_t* off, int keylen, int first_pos, int shift_tail,
                                       uint64_t* counts){
    memset(counts, 0, sizeof(uint64_t) << (2 * keylen));
    for(uint32_t r = 0; r < n; ++r){
        const std::string s(seq + off[r], (size_t)(off[r + 1] - off[r]));
        int key = -1;
        for(int pos = first_pos; pos <= (int)s.length() - keylen - shift_tail; ++pos){
            key = Evaluator::seq2int(s, pos, keylen, key);
            if(key >= 0) ++counts[key];
        }
    }
    return 0;
}
// Evaluator::getAdapterWithSeed (src/evaluator.cpp:392-425) itself on a set of reads: the adapter the reference derives from a seed
extern "C" int fqr_adapter_with_seed(uint32_t n, const char* seq, const uint64_t* off, int seed, int keylen, int trim_tail1,
                                     char* out, uint32_t cap){
    Options opt;
    Evaluator ev(&opt);
    std::vector<Read*> reads(n);
    for(uint32_t r = 0; r < n; ++r){
        const std::string s(seq + off[r], (size_t)(off[r + 1] - off[r]));
        reads[r] = new Read("@r", s, "+", std::string(s.size(), 'I'));
    }
    const std::string a = ev.getAdapterWithSeed(seed, reads.data(), (long)n, keylen, trim_tail1);
    for(auto r : reads) delete r;
    if(a.size() + 1 > cap) return -1;
    memcpy(out, a.c_str(), a.size() + 1);
    return 0;
}
