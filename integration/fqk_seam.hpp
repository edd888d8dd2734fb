// fqk_seam.hpp -- the reference-side binding of the B200 hot path (include/fqk.h).
//
// This header is what a fqtool maintainer adds to the reference tree.  It is compiled INSIDE the reference's own
// seprocessor.cpp / peprocessor.cpp (see integration/make_seam.py: the two worker bodies
//     SingleEndProcessor::processSingleEnd   /root/reference/src/seprocessor.cpp:290-388
//     PairEndProcessor::processPairEnd       /root/reference/src/peprocessor.cpp:261-508
// are replaced by integration/seprocessor_body.inc / peprocessor_body.inc, and one call to fqkseam::finishSE/PE is
// inserted in process() before Stats::merge), with the reference's Options / Read / ReadPack / ThreadConfig / Stats /
// FilterResult / Duplicate types.  Everything per base (trimming, filtering, statistics, duplication keys) runs on
// the GPU through the C ABI; what stays here is what the reference does with std::string: slicing the kept part of
// a read, assembling merged reads (OverlapAnalysis::merge, src/overlapanalysis.cpp:74-104), routing records to the
// writers and keying the adapter tallies by their text (FilterResult::addAdapterTrimmed, src/filterresult.cpp:138-177).
// At the end of the run the accumulators of fqk_finalize are written into the reference's own Stats / FilterResult /
// Duplicate objects, so JsonReporter / HtmlReporter run unchanged.
//
// No CPU fallback: if the GPU context cannot be created, or a pack violates the input contract (bases outside
// {A,C,G,T,N}, reads longer than the Stats capacity), the program stops with the library's error text.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>
#include "fqk.h"

namespace fqkseam {

[[noreturn]] inline void die(const std::string& what, fqk_ctx* ctx = nullptr){
    std::fprintf(stderr, "fqk seam: %s: %s\n", what.c_str(), fqk_last_error(ctx));
    std::exit(3);
}

#ifdef FQK_SEAM_PROCESSOR   /* the parts that need Stats / FilterResult / ThreadConfig: seprocessor.cpp, peprocessor.cpp */
// Options (after Options::update and the Evaluator pre-passes of main.cpp:124-142) -> fq_params
inline fq_params paramsFromOptions(Options* o){
    fq_params p;
    fqk_default_params(&p);
    p.is_paired = o->isPaired();
    p.trim_front1 = o->trim.front1; p.trim_tail1 = o->trim.tail1; p.trim_max_len1 = o->trim.maxLen1;
    p.trim_front2 = o->trim.front2; p.trim_tail2 = o->trim.tail2; p.trim_max_len2 = o->trim.maxLen2;
    p.cut_front_enabled = o->qualitycut.enableFront; p.cut_tail_enabled = o->qualitycut.enableTail;
    p.cut_right_enabled = o->qualitycut.enableRright;
    p.cut_front_window = o->qualitycut.windowSizeFront; p.cut_tail_window = o->qualitycut.windowSizeTail;
    p.cut_right_window = o->qualitycut.windowSizeRight;
    p.cut_front_qual = o->qualitycut.qualityFront; p.cut_tail_qual = o->qualitycut.qualityTail;
    p.cut_right_qual = o->qualitycut.qualityRight;
    p.polyg_enabled = o->polyGTrim.enabled; p.polyg_min_len = o->polyGTrim.minLen;
    p.polyg_max_mismatch = o->polyGTrim.maxMismatch; p.polyg_one_mismatch_each = o->polyGTrim.allowedOneMismatchForEach;
    p.polyx_enabled = o->polyXTrim.enabled; p.polyx_base_mask = 0;
    for(size_t i = 0; i < o->polyXTrim.trimChr.size(); ++i){
        const char* order = "ATCGN"; const char* f = std::strchr(order, o->polyXTrim.trimChr[i]);
        if(f) p.polyx_base_mask |= 1 << (int)(f - order);
    }
    p.polyx_min_len = o->polyXTrim.minLen; p.polyx_max_mismatch = o->polyXTrim.maxMismatch;
    p.polyx_one_mismatch_each = o->polyXTrim.allowedOneMismatchForEach;
    p.qual_filter_enabled = o->qualFilter.enabled; p.low_quality_limit = o->qualFilter.lowQualityLimit;
    p.low_quality_base_limit = o->qualFilter.lowQualityBaseLimit; p.n_base_limit = o->qualFilter.nBaseLimit;
    p.average_quality_limit = o->qualFilter.averageQualityLimit;
    p.length_filter_enabled = o->lengthFilter.enabled; p.min_read_length = o->lengthFilter.minReadLength;
    p.max_read_length = o->lengthFilter.maxReadLength;
    p.complexity_enabled = o->complexityFilter.enabled; p.complexity_threshold = o->complexityFilter.threshold;
    p.adapter_trim_enabled = o->adapter.enableTriming;
    if(o->adapter.adapterSeqR1Provided){
        if(o->adapter.inputAdapterSeqR1.size() > FQK_MAX_ADAPTER) die("adapter of read1 longer than FQK_MAX_ADAPTER");
        p.adapter_r1_len = (int)o->adapter.inputAdapterSeqR1.size();
        std::memcpy(p.adapter_r1, o->adapter.inputAdapterSeqR1.data(), p.adapter_r1_len);
    }
    if(o->adapter.adapterSeqR2Provided){
        if(o->adapter.inputAdapterSeqR2.size() > FQK_MAX_ADAPTER) die("adapter of read2 longer than FQK_MAX_ADAPTER");
        p.adapter_r2_len = (int)o->adapter.inputAdapterSeqR2.size();
        std::memcpy(p.adapter_r2, o->adapter.inputAdapterSeqR2.data(), p.adapter_r2_len);
    }
    p.overlap_require = o->overlapRequire; p.overlap_diff_limit = o->overlapDiffLimit; p.insert_size_max = o->insertSizeMax;
    p.correction_enabled = o->correction.enabled;
    p.merge_enabled = o->mergePE.enabled; p.merge_discard_unmerged = o->mergePE.discardUnmerged;
    p.kmer_len = o->kmer.enabled ? o->kmer.kmerLen : 0;
    p.dup_enabled = o->duplicate.enabled; p.dup_keylen = o->duplicate.keylen;
    // capacity of the per-cycle arrays: what the reference's Stats objects hold without growing (seqLen + 1024,
    // src/stats.cpp:12-17) is far more than needed; the estimated read length with a margin, merged reads included
    int need = std::max(o->est.seqLen1, o->est.seqLen2) + 16;
    if(p.is_paired && p.merge_enabled) need = o->est.seqLen1 + o->est.seqLen2 + 16;
    p.max_cycles = (need + 31) / 32 * 32;
    return p;
}

// pinned staging of one mate of a pack: flat characters + offsets (fq_ascii_batch), records
struct Mate {
    char* seq = nullptr; char* qual = nullptr; uint64_t* off = nullptr; fq_read_result* res = nullptr;
    size_t capChars = 0, capReads = 0;
    void reserve(size_t n, size_t chars){
        if(n + 1 > capReads){
            fqk_host_free(off); fqk_host_free(res);
            capReads = n + n / 4 + 16;
            off = (uint64_t*)fqk_host_alloc(capReads * sizeof(uint64_t));
            res = (fq_read_result*)fqk_host_alloc(capReads * sizeof(fq_read_result));
            if(!off || !res) die("pinned allocation failed");
        }
        if(chars > capChars){
            fqk_host_free(seq); fqk_host_free(qual);
            capChars = chars + chars / 4 + 64;
            seq = (char*)fqk_host_alloc(capChars); qual = (char*)fqk_host_alloc(capChars);
            if(!seq || !qual) die("pinned allocation failed");
        }
    }
    template<class GetRead> void load(size_t n, GetRead get, uint64_t firstIndex, fq_ascii_batch& b){
        size_t chars = 0;
        for(size_t i = 0; i < n; ++i) chars += get(i)->seq.seqStr.size();
        reserve(n, chars);
        uint64_t pos = 0;
        for(size_t i = 0; i < n; ++i){
            Read* r = get(i);
            const size_t l = r->seq.seqStr.size();
            if(r->quality.size() != l) die("sequence and quality differ in length: " + r->name);
            off[i] = pos;
            std::memcpy(seq + pos, r->seq.seqStr.data(), l); std::memcpy(qual + pos, r->quality.data(), l);
            pos += l;
        }
        off[n] = pos;
        b.n_reads = (uint32_t)n; b.reserved = 0; b.first_index = firstIndex; b.off = off; b.seq = seq; b.qual = qual;
    }
};

// names and third lines of one mate as flat buffers (fq_text_batch), for the text entry points
struct MateText {
    std::string names, strands; std::vector<uint64_t> noff, soff;
    size_t textBytes = 0;          // upper bound of the FASTQ text of the pack (untrimmed records)
    template<class GetRead> void load(size_t n, GetRead get, const fq_ascii_batch& reads, fq_text_batch& b){
        names.clear(); strands.clear(); noff.assign(n + 1, 0); soff.assign(n + 1, 0);
        for(size_t i = 0; i < n; ++i){
            Read* r = get(i);
            noff[i] = names.size(); names += r->name;
            soff[i] = strands.size(); strands += r->strand;
        }
        noff[n] = names.size(); soff[n] = strands.size();
        textBytes = names.size() + strands.size() + 2 * (size_t)reads.off[n] + 4 * n;
        b.reads = reads; b.name_off = noff.data(); b.names = names.data(); b.strand_off = soff.data(); b.strands = strands.data();
    }
};
// FQK_SEAM_TEXT=0 switches the worker bodies back to one std::string per read on the host (A/B and tests)
inline bool useText(){ static const bool on = !(std::getenv("FQK_SEAM_TEXT") && std::atoi(std::getenv("FQK_SEAM_TEXT")) == 0); return on; }
// a text channel the caller wants: a heap block that is handed to the WriterThread as it is (WriterThread frees it)
inline void wantChannel(fq_text_out* outs, int ch, size_t cap){ outs[ch].data = new char[cap ? cap : 1]; outs[ch].cap = cap; }
inline void dropChannels(fq_text_out* outs){ for(int ch = 0; ch < FQK_CH_COUNT; ++ch){ delete[] outs[ch].data; outs[ch].data = nullptr; } }

struct Seam {
    std::mutex mtx;                 // one GPU context per process; worker threads take turns.  With -w 1 packs arrive in input
                                    // order; with -w > 1 in lock order, so firstIndex is NOT the read's position in the file then
                                    // (only Duplicate's mean-GC-of-the-first-read depends on it)
    fqk_ctx* ctx = nullptr;
    fq_params P;
    fq_stats_layout L;
    uint64_t firstIndex = 0;        // global index of the next read (Duplicate's first-read-wins rule needs input order)
    Mate m[2];
    MateText t[2];
    std::vector<fq_pair_result> pairs;
    std::vector<fq_correction> edits;

    static Seam& get(Options* opt){
        static Seam S;
        static std::once_flag once;
        std::call_once(once, [&](){
            if(opt->indexFilter.enabled || opt->umi.enabled || opt->overRepAna.enabled || opt->phred64)
                die("index filter, UMI, overrepresentation analysis and phred64 input stay on the reference's CPU path (SURVEY.md section 8: out of scope)");
            S.P = paramsFromOptions(opt);
            fqk_stats_layout(&S.P, &S.L);
            const char* dev = std::getenv("FQK_DEVICE");
            if(fqk_create(&S.ctx, &S.P, dev ? std::atoi(dev) : 0, 0, 1, nullptr) != FQK_OK) die("fqk_create");
        });
        return S;
    }
};

// src/common.h:21-30
inline const char* failedType(int result){ return COMMONCONST::FAILED_TYPES[result]; }

// FilterResult keys its adapter tallies by the trimmed TEXT; the device returns where the text is
inline std::string adapterText(const std::string& originalSeq, const fq_read_result& x, const std::string& adapterSeq){
    if(!(x.flags & (FQK_F_ADAPTER_SEQ | FQK_F_ADAPTER_OV)) || x.adapter_len == 0) return std::string();
    if((x.flags & FQK_F_ADAPTER_SEQ) && x.adapter_pos < 0) return adapterSeq.substr(-x.adapter_pos);     // src/adaptertrimmer.cpp:73
    return originalSeq.substr((size_t)x.start + x.adapter_pos, x.adapter_len);                             // :80 / :17-18
}
inline void countAdapter(std::map<std::string, size_t>& m, const std::string& text){ if(!text.empty()) ++m[text]; }

// the kept part of a read, in place (what the chain of trimAndCut / polyG / adapter / polyX / maxLen leaves in or1)
inline void applyTrim(Read* r, const fq_read_result& x){
    if(x.flags & FQK_F_DROPPED) return;                            // trimAndCut returned NULL: the original is untouched
    if(x.start == 0){ r->seq.seqStr.resize(x.len); r->quality.resize(x.len); }
    else { r->seq.seqStr = r->seq.seqStr.substr(x.start, x.len); r->quality = r->quality.substr(x.start, x.len); }
}

// ---- end of run: accumulators -> the reference's own objects (thread 0's; the others stay zero and merge as zero) ----
inline void fillStats(Stats* s, const uint64_t* flat, const fq_stats_layout& L, int which){
    const uint64_t* S = flat + L.stats_off[which];
    const uint64_t C = L.max_cycles;
    if((uint64_t)s->mBufLen < C) die("Stats buffer smaller than max_cycles");
    for(int b = 0; b < 8; ++b) for(uint64_t i = 0; i < C; ++i){
        s->mCycleQ20Bases[b][i] = S[L.s_q20 + b * C + i];
        s->mCycleQ30Bases[b][i] = S[L.s_q30 + b * C + i];
        s->mCycleBaseContents[b][i] = S[L.s_content + b * C + i];
        s->mCycleBaseQuality[b][i] = S[L.s_qualsum + b * C + i];
    }
    for(uint64_t i = 0; i < C; ++i){
        s->mCycleTotalBase[i] = S[L.s_total_base + i];
        s->mCycleTotalQuality[i] = S[L.s_total_qual + i];
    }
    s->mReads = S[L.s_reads];
    s->mLengthSum = S[L.s_length_sum];
    for(uint64_t i = 0; i < L.kmer_bins; ++i) s->mKmer[i] = S[L.s_kmer + i];
}

inline void fillFilterResult(FilterResult* fr, const uint64_t* flat, const fq_stats_layout& L){
    for(int i = 0; i < COMMONCONST::FILTER_RESULT_TYPES; ++i) fr->mFilterReadStats[i] = flat[L.filter_off + i];
    fr->mTrimmedAdapterReads = flat[L.adapter_reads_off];
    fr->mTrimmedAdapterBases = flat[L.adapter_bases_off];
    for(int i = 0; i < 5; ++i){
        fr->mTrimmedPolyXReads[i] = flat[L.polyx_reads_off + i];
        fr->mTrimmedPolyXBases[i] = flat[L.polyx_bases_off + i];
    }
    fr->mMergedPairs = flat[L.merged_pairs_off];
    for(int i = 0; i < 64; ++i) fr->mCorrectionMatrix[i] = flat[L.correction_off + i];
    fr->mCorrectedReads = flat[L.corrected_reads_off];
}

inline void fillDuplicate(Seam& S, Duplicate* dup){
    if(!dup) return;
    const uint64_t n = S.L.dup_bins;
    if(fqk_dup_table(S.ctx, (uint64_t*)dup->mDups, dup->mCounts, dup->mGC) != FQK_OK) die("fqk_dup_table", S.ctx);
    (void)n;
}

inline void finishCommon(Options* opt, ThreadConfig** configs, Duplicate* dup, long* insertHist){
    Seam& S = Seam::get(opt);
    std::lock_guard<std::mutex> lk(S.mtx);
    std::vector<uint64_t> flat(S.L.total_u64);
    if(fqk_finalize(S.ctx, flat.data()) != FQK_OK) die("fqk_finalize", S.ctx);
    ThreadConfig* c0 = configs[0];
    fillStats(c0->getPreStats1(), flat.data(), S.L, 0);
    fillStats(c0->getPostStats1(), flat.data(), S.L, 2);
    if(opt->isPaired()){
        fillStats(c0->getPreStats2(), flat.data(), S.L, 1);
        fillStats(c0->getPostStats2(), flat.data(), S.L, 3);
    }
    fillFilterResult(c0->getFilterResult(), flat.data(), S.L);
    if(insertHist) for(uint64_t i = 0; i < S.L.insert_hist_len; ++i) insertHist[i] = (long)flat[S.L.insert_hist_off + i];
    fillDuplicate(S, dup);
    fqk_destroy(S.ctx); S.ctx = nullptr;
}
inline void finishSE(Options* opt, ThreadConfig** configs, Duplicate* dup){ finishCommon(opt, configs, dup, nullptr); }
inline void finishPE(Options* opt, ThreadConfig** configs, Duplicate* dup, long* insertHist){ finishCommon(opt, configs, dup, insertHist); }

#endif /* FQK_SEAM_PROCESSOR */

} // namespace fqkseam
