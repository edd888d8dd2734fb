// fqk_text.h -- device-side output compaction and FASTQ record assembly, see fqk_text.cu
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/fqk.h"

// one mate of a pack as it sits on the device after the ASCII staging (fqk_api.cu: stage_ascii + stage_text)
struct TextMate {
    const char* seq; const char* qual;        // raw FASTQ characters of the pack
    const uint32_t* src;                      // [n] offset of read r's sequence characters
    const uint32_t* qsrc;                     // [n] offset of its quality characters, nullptr = src (parsed records)
    const uint16_t* len;                      // [n] raw length
    const char* names; const uint32_t* name_off;      // name lines: [n+1] offsets, or [n] offsets + name_len (raw FASTQ blocks)
    const char* strands; const uint32_t* strand_off;  // third lines, likewise
    const uint32_t* name_len; const uint32_t* strand_len;
    __host__ __device__ uint32_t nlen(uint32_t i) const { return name_len ? name_len[i] : name_off[i + 1] - name_off[i]; }
    __host__ __device__ uint32_t slen(uint32_t i) const { return strand_len ? strand_len[i] : strand_off[i + 1] - strand_off[i]; }
    __host__ __device__ uint32_t qoff(uint32_t i) const { return qsrc ? qsrc[i] : src[i]; }
};

constexpr int TEXT_SLOTS = FQK_CH_COUNT + 1;  // per-channel record counters + "readPassed"

// sizes[ch][i] = bytes pair/read i writes to channel ch; counters[ch] = records, counters[FQK_CH_COUNT] = readPassed
void text_launch_sizes(cudaStream_t st, bool paired, uint32_t n, uint32_t writers, TextMate m1, TextMate m2,
                       const fq_read_result* r1, const fq_read_result* r2, const fq_pair_result* rp,
                       uint32_t* sizes, uint32_t* counters);
// exclusive scan of every channel in place (sizes -> offsets); totals[ch] = bytes of the channel; scratch: TEXT_SCAN_SCRATCH(n) u64
size_t text_scan_scratch_u64(uint32_t n);
void text_launch_scan(cudaStream_t st, uint32_t n, uint32_t* sizes, uint64_t* totals, uint64_t* scratch, int nch = FQK_CH_COUNT);
// write the record texts: out[ch] + offsets[ch][i]
struct TextOut { char* p[FQK_CH_COUNT]; };
void text_launch_copy(cudaStream_t st, bool paired, uint32_t n, uint32_t writers, TextMate m1, TextMate m2,
                      const fq_read_result* r1, const fq_read_result* r2, const fq_pair_result* rp,
                      const uint32_t* offsets, TextOut out);
// BaseCorrector edits (-c) applied to the staged characters before the texts are written
void text_launch_apply_edits(cudaStream_t st, const fq_correction* edits, const uint32_t* n_edits, uint32_t cap,
                             char* seq1, char* qual1, const uint32_t* src1, const uint32_t* qsrc1,
                             char* seq2, char* qual2, const uint32_t* src2, const uint32_t* qsrc2);

// ---- raw FASTQ blocks (the newline scan of FqReader, src/fqreader.cpp:90-195, on the device) ----
constexpr uint32_t NL_CHUNK = 2048;                        // bytes per block of the newline kernels
// counts[c] = newlines in chunk c of text[0, bytes)
void fastq_launch_count(cudaStream_t st, const char* text, uint64_t bytes, uint32_t* counts);
// nlpos[k] = position of the k-th newline (chunk_off = exclusive scan of counts)
void fastq_launch_positions(cudaStream_t st, const char* text, uint64_t bytes, const uint32_t* chunk_off, uint32_t* nlpos);
// records 0..n-1 = lines 4r..4r+3: offsets/lengths of the four lines, padded length for the packed layout;
// misc[0] = max length (atomicMax), err |= 8 when a block is not 4-line FASTQ, |= 1 for over-long reads
void fastq_launch_index(cudaStream_t st, const char* text, const uint32_t* nlpos, uint32_t n, uint32_t* src_seq, uint32_t* src_qual,
                        uint16_t* len, uint32_t* name_off, uint32_t* name_len, uint32_t* strand_off, uint32_t* strand_len,
                        uint32_t* padlen, uint32_t* misc, uint32_t* err);

// ---- ASCII batches indexed on the device (pipelined ingest): src / len / padded length from the caller's offsets ----
// off: device copy of fq_ascii_batch.off[n+1]; err |= 1 for a read longer than FQK_MAX_READ_LEN
void ascii_launch_index(cudaStream_t st, uint32_t n, const uint64_t* off, uint32_t* src, uint16_t* len, uint32_t* padlen, uint32_t* err);
// zero the 64 bytes (16 base bytes) after the last padded read: *total = end of the packed data
void ascii_launch_zero_tail(cudaStream_t st, const uint64_t* total, uint8_t* qual, uint8_t* bases2);
// out[i] = (uint32_t)(in[i] - in[0]), i < n (offset arrays of a text batch narrowed on the device: pipelined text path)
void text_launch_narrow(cudaStream_t st, uint32_t n, const uint64_t* in, uint32_t* out);
