// fqk_run -- small command-line harness around fqk::HotPath with the reference's flag names:
//   fqk_run -i r1.fq[.gz] [-I r2.fq[.gz]] -o out1.fq [-O out2.fq] [--merge_output m.fq] [--failed_out f.fq] <hot-path flags>
// It plays the host role of SingleEndProcessor / PairEndProcessor for one process and one GPU
// (packs of --max_item_in_pack reads in input order = the reference's -w 1 order) and writes plain FASTQ,
// so that its output files can be compared byte for byte with the reference binary's.
#include <chrono>
#include <cstdio>
#include <fstream>
#include <iostream>
#include "fqk_host.hpp"

static bool writeFile(const std::string& path, const std::string& s){
    if(path.empty()) return true;
    std::ofstream f(path, std::ios::binary);
    f.write(s.data(), (std::streamsize)s.size());
    return (bool)f;
}

int main(int argc, char** argv){
    fqk::CliOptions o; std::string err;
    if(!o.parse(argc, argv, err)){ std::cerr << "fqk_run: " << err << "\n"; return 2; }
    std::vector<fqk::Read> r1, r2;
    if(!fqk::readFastq(o.in1, r1, err) || (!o.in2.empty() && !fqk::readFastq(o.in2, r2, err))){ std::cerr << "fqk_run: " << err << "\n"; return 2; }
    const bool paired = !o.in2.empty();
    if(paired && r1.size() != r2.size()) r1.resize(std::min(r1.size(), r2.size())), r2.resize(r1.size());   // src/fqreader.cpp:254-266
    fqk::Channels ch; ch.failed = !o.failedOut.empty(); ch.unpaired1 = !o.unpaired1.empty(); ch.unpaired2 = !o.unpaired2.empty();
    try{
        fqk::HotPath hp(o.params, ch, 0);
        std::string out1, out2, merged, failed, un1, un2;
        size_t passed = 0, mergedCount = 0;
        double hotSeconds = 0;                         // packing + fqk_process_* + slicing/concatenation (file I/O excluded)
        for(size_t lo = 0; lo < r1.size(); lo += (size_t)o.packSize){
            const size_t hi = std::min(r1.size(), lo + (size_t)o.packSize);
            std::vector<fqk::Read> p1(r1.begin() + lo, r1.begin() + hi);
            if(!paired){
                const auto t0 = std::chrono::steady_clock::now();
                fqk::SEOutput so; hp.processSingleEnd(p1, so);
                hotSeconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                out1 += so.out1; failed += so.failed; passed += so.readPassed;
            }else{
                std::vector<fqk::Read> p2(r2.begin() + lo, r2.begin() + hi);
                const auto t0 = std::chrono::steady_clock::now();
                fqk::PEOutput po; hp.processPairEnd(p1, p2, po);
                hotSeconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                out1 += po.out1; out2 += po.out2; merged += po.merged; failed += po.failed; un1 += po.unpaired1; un2 += po.unpaired2;
                passed += po.readPassed; mergedCount += po.mergedCount;
            }
        }
        std::vector<uint64_t> flat; hp.finalize(flat);
        const fq_stats_layout& L = hp.layout();
        if(!writeFile(o.out1, out1) || !writeFile(o.out2, out2) || !writeFile(o.mergeOut, merged) || !writeFile(o.failedOut, failed) ||
           !writeFile(o.unpaired1, un1) || (ch.unpaired2 && !writeFile(o.unpaired2, un2))){ std::cerr << "fqk_run: write failed\n"; return 3; }
        // summary in the spirit of FilterResult::reportJsonBasic (src/filterresult.cpp): raw counters only
        printf("{\"reads_in\": %zu, \"passed\": %zu, \"merged\": %zu, \"filter\": {\"pass\": %llu, \"n_base\": %llu, \"length\": %llu, "
               "\"too_long\": %llu, \"quality\": %llu, \"complexity\": %llu}, \"adapter_reads\": %llu, \"adapter_bases\": %llu, "
               "\"adapter1_kinds\": %zu, \"adapter2_kinds\": %zu, \"pre1_reads\": %llu, \"post1_reads\": %llu, \"host_path_seconds\": %.6f}\n",
               r1.size() * (paired ? 2 : 1), passed, mergedCount,
               (unsigned long long)flat[L.filter_off + 0], (unsigned long long)flat[L.filter_off + 12], (unsigned long long)flat[L.filter_off + 16],
               (unsigned long long)flat[L.filter_off + 17], (unsigned long long)flat[L.filter_off + 20], (unsigned long long)flat[L.filter_off + 24],
               (unsigned long long)flat[L.adapter_reads_off], (unsigned long long)flat[L.adapter_bases_off],
               hp.adapter1Count().size(), hp.adapter2Count().size(),
               (unsigned long long)flat[L.stats_off[0] + L.s_reads], (unsigned long long)flat[L.stats_off[2] + L.s_reads], hotSeconds);
    }catch(const std::exception& e){ std::cerr << "fqk_run: " << e.what() << "\n"; return 1; }
    return 0;
}
