// fqk_text.h -- device-side output compaction and FASTQ record assembly, see fqk_text.cu
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/fqk.h"

// one mate of a pack as it sits on the device after the ASCII staging (fqk_api.cu: stage_ascii + stage_text)
struct TextMate {
    const char* seq; const char* qual;        // raw FASTQ characters of the pack
    const uint32_t* src;                      // [n] offset of read r's characters
    const uint16_t* len;                      // [n] raw length
    const char* names; const uint32_t* name_off;      // name lines, [n+1]
    const char* strands; const uint32_t* strand_off;  // third lines, [n+1]
};

constexpr int TEXT_SLOTS = FQK_CH_COUNT + 1;  // per-channel record counters + "readPassed"

// sizes[ch][i] = bytes pair/read i writes to channel ch; counters[ch] = records, counters[FQK_CH_COUNT] = readPassed
void text_launch_sizes(cudaStream_t st, bool paired, uint32_t n, uint32_t writers, TextMate m1, TextMate m2,
                       const fq_read_result* r1, const fq_read_result* r2, const fq_pair_result* rp,
                       uint32_t* sizes, uint32_t* counters);
// exclusive scan of every channel in place (sizes -> offsets); totals[ch] = bytes of the channel; scratch: TEXT_SCAN_SCRATCH(n) u64
size_t text_scan_scratch_u64(uint32_t n);
void text_launch_scan(cudaStream_t st, uint32_t n, uint32_t* sizes, uint64_t* totals, uint64_t* scratch);
// write the record texts: out[ch] + offsets[ch][i]
struct TextOut { char* p[FQK_CH_COUNT]; };
void text_launch_copy(cudaStream_t st, bool paired, uint32_t n, uint32_t writers, TextMate m1, TextMate m2,
                      const fq_read_result* r1, const fq_read_result* r2, const fq_pair_result* rp,
                      const uint32_t* offsets, TextOut out);
// BaseCorrector edits (-c) applied to the staged characters before the texts are written
void text_launch_apply_edits(cudaStream_t st, const fq_correction* edits, const uint32_t* n_edits, uint32_t cap,
                             char* seq1, char* qual1, const uint32_t* src1, char* seq2, char* qual2, const uint32_t* src2);
