// fqk_host.cpp -- see fqk_host.hpp.  Host-side string work only; the hot path is fqk_process_*.
#include "fqk_host.hpp"
#include <algorithm>
#include <cstring>
#include <stdexcept>
#include <zlib.h>

namespace fqk {

const char* const FAILED_TYPES[FQK_FILTER_RESULT_TYPES] = {      // src/common.h:21-30
    "passed", "", "", "", "failed_polyx_filter", "", "", "", "failed_bad_overlap", "", "", "",
    "failed_too_many_n_bases", "", "", "", "failed_too_short", "failed_too_long", "", "",
    "failed_quality_filter", "", "", "", "failed_low_complexity", "", "", "", "", "", "", ""};

// pinned staging buffers of one mate, grown on demand and reused across packs
struct HotPath::Buffers {
    uint32_t* offsets = nullptr; uint16_t* lengths = nullptr; uint8_t* bases2 = nullptr; uint8_t* qual = nullptr;
    fq_read_result* results = nullptr;
    size_t capReads = 0, capBases = 0;
    std::string seq, qualAscii; std::vector<uint64_t> off;
    void reserve(size_t n, size_t total){
        if(n > capReads){
            fqk_host_free(offsets); fqk_host_free(lengths); fqk_host_free(results);
            capReads = n + n / 4 + 16;
            offsets = (uint32_t*)fqk_host_alloc(capReads * 4); lengths = (uint16_t*)fqk_host_alloc(capReads * 2);
            results = (fq_read_result*)fqk_host_alloc(capReads * sizeof(fq_read_result));
            if(!offsets || !lengths || !results) throw std::runtime_error("fqk_host_alloc failed");
        }
        if(total > capBases){
            fqk_host_free(bases2); fqk_host_free(qual);
            capBases = total + total / 4 + 64;
            bases2 = (uint8_t*)fqk_host_alloc(capBases / 4 + 16); qual = (uint8_t*)fqk_host_alloc(capBases + 16);
            if(!bases2 || !qual) throw std::runtime_error("fqk_host_alloc failed");
        }
    }
    ~Buffers(){ fqk_host_free(offsets); fqk_host_free(lengths); fqk_host_free(bases2); fqk_host_free(qual); fqk_host_free(results); }
};

HotPath::HotPath(const fq_params& p, const Channels& ch, int device, int rank, int nranks, const void* id)
    : mParams(p), mCh(ch){
    mBuf[0] = new Buffers(); mBuf[1] = new Buffers();
    fqk_stats_layout(&mParams, &mLayout);
    if(fqk_create(&mCtx, &mParams, device, rank, nranks, id) != FQK_OK){
        std::string why = fqk_last_error(nullptr);
        delete mBuf[0]; delete mBuf[1];
        throw std::runtime_error("fqk_create: " + why);          // no CPU fallback
    }
}

HotPath::~HotPath(){
    fqk_destroy(mCtx);
    delete mBuf[0]; delete mBuf[1];
}

void HotPath::pack(const std::vector<Read>& reads, Buffers& b, fq_batch& batch){
    const uint32_t n = (uint32_t)reads.size();
    b.off.assign(n + 1, 0);
    for(uint32_t i = 0; i < n; ++i) b.off[i + 1] = b.off[i] + reads[i].seq.size();
    b.seq.clear(); b.qualAscii.clear(); b.seq.reserve(b.off[n]); b.qualAscii.reserve(b.off[n]);
    for(uint32_t i = 0; i < n; ++i){
        if(reads[i].quality.size() != reads[i].seq.size()) throw std::runtime_error("sequence and quality differ in length: " + reads[i].name);
        b.seq += reads[i].seq; b.qualAscii += reads[i].quality;
    }
    const uint64_t total = fqk_packed_bases(n, b.off.data());
    b.reserve(n, total);
    const int rc = fqk_pack_ascii(n, b.seq.data(), b.qualAscii.data(), b.off.data(), b.offsets, b.lengths, b.bases2, b.qual);
    if(rc == FQK_E_ALPHABET) throw std::runtime_error("read with a base outside {A,C,G,T,N}: outside the input contract of the GPU path");
    if(rc) throw std::runtime_error("read longer than FQK_MAX_READ_LEN");
    batch.n_reads = n; batch.max_len = 0; batch.first_index = mFirstIndex; batch.total_bases = total;
    batch.offsets = b.offsets; batch.lengths = b.lengths; batch.bases2 = b.bases2; batch.qual = b.qual;
}

// FilterResult::addAdapterTrimmed keys its maps by the trimmed text (src/filterresult.cpp:138-177); the
// device returns where the text is, the host slices it from the untrimmed read
void HotPath::tallyAdapter(const Read& o, const fq_read_result& x, bool isR2){
    if(!(x.flags & (FQK_F_ADAPTER_SEQ | FQK_F_ADAPTER_OV)) || x.adapter_len == 0) return;
    std::string text;
    if((x.flags & FQK_F_ADAPTER_SEQ) && x.adapter_pos < 0){
        const std::string ad(isR2 ? mParams.adapter_r2 : mParams.adapter_r1, isR2 ? mParams.adapter_r2_len : mParams.adapter_r1_len);
        text = ad.substr(-x.adapter_pos);                          // src/adaptertrimmer.cpp:73
    }else text = o.seq.substr((size_t)x.start + x.adapter_pos, x.adapter_len);   // :80 / :17-18
    ++(isR2 ? mAdapter2Count : mAdapter1Count)[text];
}

static inline void applyTrim(Read& r, const fq_read_result& x){
    if(x.flags & FQK_F_DROPPED) return;                            // trimAndCut returned NULL: read untouched
    r.seq = r.seq.substr(x.start, x.len);
    r.quality = r.quality.substr(x.start, x.len);
}

// loop body of SingleEndProcessor::processSingleEnd, src/seprocessor.cpp:294-354 (routing part)
void HotPath::processSingleEnd(std::vector<Read>& packReads, SEOutput& out){
    fq_batch b;
    pack(packReads, *mBuf[0], b);
    if(fqk_process_se(mCtx, &b, mBuf[0]->results) != FQK_OK) throw std::runtime_error(std::string("fqk_process_se: ") + fqk_last_error(mCtx));
    mFirstIndex += packReads.size();
    for(size_t i = 0; i < packReads.size(); ++i){
        Read& r = packReads[i]; const fq_read_result& x = mBuf[0]->results[i];
        tallyAdapter(r, x, false);
        applyTrim(r, x);
        const bool alive = !(x.flags & FQK_F_DROPPED);
        if(alive && x.result == FQK_PASS_FILTER){ out.out1 += r.toString(); ++out.readPassed; }      // :342-345
        else if(mCh.failed) out.failed += r.toStringWithTag(FAILED_TYPES[x.result]);                 // :346-348
    }
}

static std::string revcomp(const std::string& s){                 // Seq::reverseComplement, src/seq.h:24-48
    std::string r(s.size(), 0);
    for(size_t i = 0; i < s.size(); ++i){
        char c;
        switch(s[i]){ case 'A': case 'a': c = 'T'; break; case 'T': case 't': c = 'A'; break;
                      case 'C': case 'c': c = 'G'; break; case 'G': case 'g': c = 'C'; break; default: c = 'N'; }
        r[s.size() - 1 - i] = c;
    }
    return r;
}

// OverlapAnalysis::merge, src/overlapanalysis.cpp:74-104, from the geometry the device returns
static Read mergeReads(const Read& r1, const Read& r2, const fq_pair_result& p){
    const int ol = p.ov2_len, len1 = p.merge_len1, len2 = p.merge_len2;
    Read m;
    m.seq = r1.seq.substr(0, len1); m.quality = r1.quality.substr(0, len1);
    if(p.ov2_offset > 0){
        m.seq += revcomp(r2.seq).substr(ol, len2);
        m.quality += std::string(r2.quality.rbegin(), r2.quality.rend()).substr(ol, len2);
    }
    const std::string tag = "_merged_" + std::to_string(len1) + "_" + std::to_string(len2);
    const std::string::size_type pos = r1.name.find_first_of(" ");
    if(pos == std::string::npos) m.name = tag;                                       // :94-96 (the prefix is lost)
    else m.name = r1.name.substr(0, pos - 1) + tag + r1.name.substr(pos);            // :97-99 (pos - 1 as written)
    m.strand = r1.strand;
    return m;
}

// loop body of PairEndProcessor::processPairEnd, src/peprocessor.cpp:271-438 (routing part)
void HotPath::processPairEnd(std::vector<Read>& p1, std::vector<Read>& p2, PEOutput& out){
    if(p1.size() != p2.size()) throw std::runtime_error("read1/read2 packs differ in size");
    fq_batch b1, b2;
    pack(p1, *mBuf[0], b1); pack(p2, *mBuf[1], b2);
    mPairRes.resize(p1.size());
    if(fqk_process_pe(mCtx, &b1, &b2, mBuf[0]->results, mBuf[1]->results, mPairRes.data()) != FQK_OK)
        throw std::runtime_error(std::string("fqk_process_pe: ") + fqk_last_error(mCtx));
    mFirstIndex += p1.size();
    if(mParams.correction_enabled){                                                  // BaseCorrector edits (src/peprocessor.cpp:310-312)
        uint32_t n = 0;
        if(fqk_get_corrections(mCtx, nullptr, 0, &n) != FQK_OK) throw std::runtime_error(std::string("fqk_get_corrections: ") + fqk_last_error(mCtx));
        std::vector<fq_correction> ed(n);
        if(n && fqk_get_corrections(mCtx, ed.data(), n, &n) != FQK_OK) throw std::runtime_error(std::string("fqk_get_corrections: ") + fqk_last_error(mCtx));
        for(const fq_correction& e : ed){
            Read& r = (e.mate_base & 0x80) ? p2[e.pair] : p1[e.pair];
            r.seq[e.pos] = (char)(e.mate_base & 0x7F); r.quality[e.pos] = (char)e.qual;
        }
    }
    const int PASS = FQK_PASS_FILTER;
    for(size_t i = 0; i < p1.size(); ++i){
        Read& or1 = p1[i]; Read& or2 = p2[i];
        const fq_read_result& x1 = mBuf[0]->results[i]; const fq_read_result& x2 = mBuf[1]->results[i];
        const fq_pair_result& pr = mPairRes[i];
        tallyAdapter(or1, x1, false); tallyAdapter(or2, x2, true);
        applyTrim(or1, x1); applyTrim(or2, x2);
        const bool a1 = !(x1.flags & FQK_F_DROPPED), a2 = !(x2.flags & FQK_F_DROPPED);
        const int result1 = x1.result, result2 = x2.result;
        if(pr.flags & FQK_P_MERGE_PROC){                                             // :353-385
            if(pr.flags & FQK_P_OV2_OVERLAP){
                if((pr.flags & FQK_P_MERGED) && pr.ov2_len){
                    out.merged += mergeReads(or1, or2, pr).toString();
                    ++out.readPassed; ++out.mergedCount;
                }
            }else{
                if(result1 == PASS) out.merged += or1.toString();
                if(result2 == PASS) out.merged += or2.toString();
                if(result1 == PASS && result2 == PASS) ++out.readPassed;
            }
            continue;
        }
        if(a1 && result1 == PASS && a2 && result2 == PASS){                          // :392-402
            if(mCh.toStdout) out.single += or1.toString() + or2.toString();
            else { out.out1 += or1.toString(); out.out2 += or2.toString(); }
            ++out.readPassed;
        }else if(a1 && result1 == PASS){                                             // :403-414
            if(mCh.unpaired1){
                out.unpaired1 += or1.toString();
                if(mCh.failed) out.failed += or2.toStringWithTag(FAILED_TYPES[result2]);
            }else if(mCh.failed){
                out.failed += or1.toStringWithTag("paired_read_is_failing");
                out.failed += or2.toStringWithTag(FAILED_TYPES[result2]);
            }
        }else if(a2 && result2 == PASS){                                             // :415-427 (tests the LEFT writer, tags with result2)
            if(mCh.unpaired1){
                out.unpaired2 += or2.toString();
                if(mCh.failed) out.failed += or1.toStringWithTag(FAILED_TYPES[result2]);
            }else if(mCh.failed){
                out.failed += or1.toStringWithTag(FAILED_TYPES[result1]);
                out.failed += or2.toStringWithTag("paired_read_is_failing");
            }
        }
    }
}

void HotPath::finalize(std::vector<uint64_t>& flat){
    flat.assign(mLayout.total_u64, 0);
    if(fqk_finalize(mCtx, flat.data()) != FQK_OK) throw std::runtime_error(std::string("fqk_finalize: ") + fqk_last_error(mCtx));
}

void HotPath::dupTable(std::vector<uint64_t>& dups, std::vector<uint32_t>& counts, std::vector<uint8_t>& gc){
    dups.assign(mLayout.dup_bins, 0); counts.assign(mLayout.dup_bins, 0); gc.assign(mLayout.dup_bins, 0);
    if(fqk_dup_table(mCtx, dups.data(), counts.data(), gc.data()) != FQK_OK) throw std::runtime_error(std::string("fqk_dup_table: ") + fqk_last_error(mCtx));
}

// ------------------------------------------------------------------ CLI subset
// flag names and defaults of /root/reference/src/main.cpp:14-122; Options::update (src/options.cpp:24-58)
bool CliOptions::parse(int argc, char** argv, std::string& err){
    fqk_default_params(&params);
    int lowQ = 20; double ratio = 0.15; bool kmerOn = false; int kmerLen = 0; std::string trimChr = "ATCGN";
    auto need = [&](int& i) -> const char* { if(i + 1 >= argc){ err = std::string("missing value for ") + argv[i]; return nullptr; } return argv[++i]; };
    for(int i = 1; i < argc; ++i){
        const std::string a = argv[i];
        const char* v = nullptr;
#define FLAG(name, field) if(a == name){ params.field = 1; continue; }
#define IVAL(name, field) if(a == name){ if(!(v = need(i))) return false; params.field = atoi(v); continue; }
#define SVAL(name, var) if(a == name){ if(!(v = need(i))) return false; var = v; continue; }
        FLAG("-q", qual_filter_enabled) FLAG("-a", adapter_trim_enabled) FLAG("-g", polyg_enabled) FLAG("-x", polyx_enabled)
        FLAG("-d", dup_enabled) FLAG("-m", merge_enabled) FLAG("--discard_unmerged", merge_discard_unmerged)
        FLAG("-l", length_filter_enabled) FLAG("-y", complexity_enabled) FLAG("-c", correction_enabled)
        FLAG("--enable_cut_front", cut_front_enabled) FLAG("--enable_cut_tail", cut_tail_enabled) FLAG("--enable_cut_right", cut_right_enabled)
        IVAL("-f", trim_front1) IVAL("-t", trim_tail1) IVAL("-b", trim_max_len1) IVAL("-F", trim_front2) IVAL("-T", trim_tail2) IVAL("-B", trim_max_len2)
        IVAL("--min_len_detect_polyG", polyg_min_len) IVAL("--max_mismatches_polyG", polyg_max_mismatch) IVAL("--one_mismatch_each_polyG", polyg_one_mismatch_each)
        IVAL("--min_len_detect_polyX", polyx_min_len) IVAL("--max_mismatches_polyX", polyx_max_mismatch) IVAL("--one_mismatch_each_polyX", polyx_one_mismatch_each)
        IVAL("--cut_front_window", cut_front_window) IVAL("--cut_tail_window", cut_tail_window) IVAL("--cut_right_window", cut_right_window)
        IVAL("--cut_front_mean_qual", cut_front_qual) IVAL("--cut_tail_mean_qual", cut_tail_qual) IVAL("--cut_right_mean_qual", cut_right_qual)
        IVAL("-N", n_base_limit) IVAL("--min_length", min_read_length) IVAL("--max_length", max_read_length)
        IVAL("--dup_ana_key_len", dup_keylen) IVAL("--min_overlap_len", overlap_require) IVAL("--max_diff_for_overlap", overlap_diff_limit)
        IVAL("--max_cycles", max_cycles)
        SVAL("-i", in1) SVAL("-I", in2) SVAL("-o", out1) SVAL("-O", out2) SVAL("--merge_output", mergeOut) SVAL("--failed_out", failedOut)
        SVAL("--unpaired_read1", unpaired1) SVAL("--unpaired_read2", unpaired2) SVAL("--base_to_trim", trimChr)
        if(a == "-W" || a == "-M" || a == "-w" || a == "-J" || a == "-H" || a == "-z"){ if(!need(i)) return false; continue; }   // -W/-M are dead flags
        if(a == "-Q"){ if(!(v = need(i))) return false; lowQ = atoi(v); continue; }
        if(a == "-U"){ if(!(v = need(i))) return false; ratio = atof(v); continue; }
        if(a == "-e"){ if(!(v = need(i))) return false; params.average_quality_limit = atof(v); continue; }
        if(a == "-Y"){ if(!(v = need(i))) return false; params.complexity_threshold = atof(v); continue; }
        if(a == "--kmer"){ kmerOn = true; continue; }
        if(a == "--kmer_length"){ if(!(v = need(i))) return false; kmerLen = atoi(v); continue; }
        if(a == "--max_item_in_pack"){ if(!(v = need(i))) return false; packSize = atoi(v); continue; }
        if(a == "--adapter_of_read1" || a == "--adapter_of_read2"){
            if(!(v = need(i))) return false;
            const size_t n = strlen(v);
            if(n > FQK_MAX_ADAPTER){ err = "adapter too long"; return false; }
            if(a == "--adapter_of_read1"){ memcpy(params.adapter_r1, v, n); params.adapter_r1_len = (int)n; }
            else { memcpy(params.adapter_r2, v, n); params.adapter_r2_len = (int)n; }
            continue;
        }
#undef FLAG
#undef IVAL
#undef SVAL
        err = "flag " + a + " is outside the hot-path scope of this harness";
        return false;
    }
    if(in1.empty()){ err = "-i is required"; return false; }
    params.is_paired = !in2.empty();
    params.low_quality_limit = lowQ + 33;                          // options.cpp:26
    params.low_quality_base_limit = (int)(ratio * 151);            // options.cpp:44 (est.seqLen1 is still 151 there)
    params.kmer_len = kmerOn ? kmerLen : 0;
    params.polyx_base_mask = 0;
    for(char& ch : trimChr) ch = (char)toupper(ch);
    for(int b = 0; b < 5; ++b) if(trimChr.find("ATCGN"[b]) != std::string::npos) params.polyx_base_mask |= 1 << b;
    if(trimChr.find_first_not_of("ATCGN") != std::string::npos){ err = "Can only trim nucleotides ATCGN"; return false; }   // options.cpp:67-70
    if(params.merge_enabled && mergeOut.empty()){ err = "merged file output must be provided!"; return false; }          // options.cpp:62-66
    return true;
}

// ------------------------------------------------------------------ FASTQ input
bool readFastq(const std::string& path, std::vector<Read>& out, std::string& err){
    gzFile f = gzopen(path.c_str(), "rb");                         // gzread also reads plain files
    if(!f){ err = "cannot open " + path; return false; }
    gzbuffer(f, 1 << 20);
    std::string data; char buf[1 << 16]; int n;
    while((n = gzread(f, buf, sizeof(buf))) > 0) data.append(buf, n);
    gzclose(f);
    size_t p = 0;
    auto line = [&](std::string& s) -> bool {
        if(p >= data.size()) return false;
        size_t e = data.find('\n', p);
        if(e == std::string::npos) e = data.size();
        s.assign(data, p, e - p);
        if(!s.empty() && s.back() == '\r') s.pop_back();
        p = e + 1;
        return true;
    };
    std::string name;
    while(line(name)){
        if(name.empty() || name[0] != '@') continue;               // src/fqreader.cpp:168-171
        Read r; r.name = name;
        if(!line(r.seq) || !line(r.strand) || !line(r.quality)) break;
        if(r.quality.size() != r.seq.size()){ err = "sequence and quality differ in length: " + name; return false; }   // :184-191
        out.push_back(std::move(r));
    }
    return true;
}

} // namespace fqk
