// fqk_host.hpp -- C++ host adapter above the C ABI (include/fqk.h).
//
// Mirrors the part of the reference that sits around the hot path, with the same names and
// behaviour, so that a pack of Reads handed to processSingleEnd / processPairEnd is a drop-in:
//   fqk::Read        <-> Read          /root/reference/src/read.h:13-198 (name, seq, strand, quality, toString...)
//   fqk::HotPath     <-> the loop bodies of SingleEndProcessor::processSingleEnd
//                        (src/seprocessor.cpp:290-388) and PairEndProcessor::processPairEnd
//                        (src/peprocessor.cpp:261-508): output routing, failed tags, merged read
//                        assembly (OverlapAnalysis::merge, src/overlapanalysis.cpp:74-104) and the
//                        adapter text tallies of FilterResult (src/filterresult.cpp:138-177)
// The trimming / filtering / statistics themselves run on the GPU through fqk_process_*; there
// is no CPU implementation of them here.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>
#include "../../include/fqk.h"

namespace fqk {

struct Read {                                   // src/read.h:13-19
    std::string name, seq, strand, quality;
    std::string toString() const { return name + "\n" + seq + "\n" + strand + "\n" + quality + "\n"; }               // read.h:166-168
    std::string toStringWithTag(const std::string& tag) const {                                                    // read.h:174-176
        return name + " " + tag + "\n" + seq + "\n" + strand + "\n" + quality + "\n";
    }
    int length() const { return (int)seq.length(); }
};

// which output channels exist (the reference tests its WriterThread pointers, e.g. mFailedWriter)
struct Channels {
    bool failed = false;            // --failed_out
    bool unpaired1 = false;         // --unpaired_read1 (mUnPairedLeftWriter)
    bool unpaired2 = false;         // --unpaired_read2
    bool toStdout = false;          // Options::outputToSTDOUT
};

struct SEOutput { std::string out1, failed; int readPassed = 0; };
struct PEOutput {
    std::string out1, out2, merged, failed, unpaired1, unpaired2, single;
    int readPassed = 0, mergedCount = 0;
};

extern const char* const FAILED_TYPES[FQK_FILTER_RESULT_TYPES];   // src/common.h:21-30

class HotPath {
public:
    HotPath(const fq_params& p, const Channels& ch, int device = 0, int rank = 0, int nranks = 1,
            const void* ncclUniqueId = nullptr);
    ~HotPath();
    HotPath(const HotPath&) = delete;
    HotPath& operator=(const HotPath&) = delete;

    // the pack is consumed like the reference does (reads are trimmed in place)
    void processSingleEnd(std::vector<Read>& pack, SEOutput& out);
    void processPairEnd(std::vector<Read>& pack1, std::vector<Read>& pack2, PEOutput& out);

    // end of run: merged accumulators in fq_stats_layout order (NCCL all-reduce across ranks inside)
    void finalize(std::vector<uint64_t>& flat);
    void dupTable(std::vector<uint64_t>& dups, std::vector<uint32_t>& counts, std::vector<uint8_t>& gc);
    const fq_stats_layout& layout() const { return mLayout; }
    const std::map<std::string, size_t>& adapter1Count() const { return mAdapter1Count; }   // FilterResult::mAdapter1Count
    const std::map<std::string, size_t>& adapter2Count() const { return mAdapter2Count; }

private:
    struct Buffers;                              // pinned staging of one mate
    void pack(const std::vector<Read>& reads, Buffers& b, fq_batch& batch);
    void tallyAdapter(const Read& original, const fq_read_result& x, bool isR2);

    fq_params mParams;
    Channels mCh;
    fq_stats_layout mLayout;
    fqk_ctx* mCtx = nullptr;
    uint64_t mFirstIndex = 0;
    Buffers* mBuf[2];
    std::vector<fq_pair_result> mPairRes;
    std::map<std::string, size_t> mAdapter1Count, mAdapter2Count;
};

// Options::update + the hot-path subset of the CLI (src/main.cpp:14-122), used by the harness
struct CliOptions {
    fq_params params;
    std::string in1, in2, out1, out2, mergeOut, failedOut, unpaired1, unpaired2;
    int packSize = 100000;                       // BufferSizeOptions::maxReadsInPack, src/options.h:21
    bool parse(int argc, char** argv, std::string& err);
};

// minimal FASTQ(.gz) reader / writer for the harness (the reference keeps its own FqReader / WriterThread)
bool readFastq(const std::string& path, std::vector<Read>& out, std::string& err);

} // namespace fqk
