// emu_pe.cpp -- TEST INFRASTRUCTURE (CPU test suite only; never on the product path).
// Runs the trim role of the fused paired-end kernel k_pe_v4 (v4_chunk, fqtool_b200/csrc/fqk_pe4.cuh: the very
// source nvcc compiles for the GPU) for whole batches on the CPU, one emulated warp at a time (cuda_emu.h), fed by
// a scalar stand-in for the copy producer, and consumes its outputs (records, length histograms, exception lists,
// tail corrections) with a scalar restatement of the statistics role's contract:
//     pre = P,  post = P - D + Z (+ M),
// so that records AND the flat accumulator buffer can be compared bit for bit with the oracle without a GPU.
#include "cuda_emu.h"
static struct { unsigned x, y, z; } threadIdx = {0, 0, 0}, blockIdx = {0, 0, 0}, blockDim = {1, 1, 1}, gridDim = {1, 1, 1};
#define EMU_MARK2(k) do{ if(getenv("EMU_TRACE")) fprintf(stderr, "  mark2 %d lane %d rawlen %d start %d len %d nmask %x qa %u\n", k, emu::cur, R.rawlen, R.start, R.len, R.nmask, R.qa); }while(0)
#define EMU_MARK(k) do{ if(getenv("EMU_TRACE") && lane == 0) fprintf(stderr, "  mark %d\n", k); }while(0)
#include "../../fqtool_b200/csrc/fqk_pe4.cuh"
#include "../../fqtool_b200/csrc/fqk_hostutil.h"
#include <vector>

namespace {
struct ChunkArg { const V4X* X; int j; Tally* tl; };
void chunk_body(int lane, void* p){ ChunkArg* a = (ChunkArg*)p; v4_chunk<V4Spec<true>, 0>(*a->X, a->j, lane, a->tl[lane]); }

// statistics arrays in the kernel's representation: u32 [kind][set][C] then lenhist[C+1]
struct SmStats { uint32_t base; uint32_t C; };
inline uint32_t cell(const SmStats& s, int kind, int set, uint32_t c){ return s.base + 4u * ((uint32_t)(kind * NSETS + set) * s.C + c); }
void add_base(const SmStats& s, uint32_t c, uint32_t by, uint32_t code, int sign){
    const uint32_t q7 = by & 0x7F; const bool isn = by & 0x80;
    const uint32_t v = (uint32_t)sign;
    bool in[NSETS] = {true, !isn && (code & 1), !isn && (code & 2), !isn && code == 3, isn};
    for(int st = 0; st < NSETS; ++st){
        if(!in[st]) continue;
        if(st) reds_add(cell(s, KN, st, c), v);
        if(q7 > 53) reds_add(cell(s, KN20, st, c), v);
        if(q7 > 63) reds_add(cell(s, KN30, st, c), v);
        reds_add(cell(s, KQS, st, c), v * q7);
    }
}
// sets -> bins, quality characters -> Phred, into the flat u64 block of one Stats object (retire() of fqk_fused.cu)
void retire_host(const SmStats& s, const fq_stats_layout& L, uint64_t* S){
    const uint32_t C = s.C;
    const uint32_t lh = s.base + 16u * NSETS * C;
    const uint64_t GC = L.max_cycles;
    for(uint32_t i = 0; i < C; ++i){
        int32_t nall = 0;
        for(uint32_t l = i + 1; l <= C; ++l) nall += (int32_t)lds32(lh + 4u * l);
        int64_t tb = 0, tq = 0;
        for(int cls = 0; cls < 5; ++cls){
            int64_t r[4];
            for(int k = 0; k < 4; ++k){
                const int64_t all = k == KN ? (int64_t)nall : (int64_t)lds32(cell(s, k, S_ALL, i));
                const int64_t b0 = lds32(cell(s, k, S_B0, i)), b1 = lds32(cell(s, k, S_B1, i)), b01 = lds32(cell(s, k, S_B01, i)), nn = lds32(cell(s, k, S_N, i));
                // the arrays hold wrapping 32-bit sums of signed contributions
                auto sx = [](int64_t v){ return (int64_t)(int32_t)(uint32_t)v; };
                r[k] = cls == 0 ? sx(all) - sx(b0) - sx(b1) + sx(b01) - sx(nn) : cls == 1 ? sx(b0) - sx(b01) : cls == 2 ? sx(b1) - sx(b01) : cls == 3 ? sx(b01) : sx(nn);
            }
            const int64_t qs = r[KQS] - 33 * r[KN];
            const int bin = (0x67341 >> (cls * 4)) & 0xF;
            S[L.s_content + bin * GC + i] += (uint64_t)r[KN]; S[L.s_q20 + bin * GC + i] += (uint64_t)r[KN20];
            S[L.s_q30 + bin * GC + i] += (uint64_t)r[KN30]; S[L.s_qualsum + bin * GC + i] += (uint64_t)qs;
            tb += r[KN]; tq += qs;
        }
        S[L.s_total_base + i] += (uint64_t)tb; S[L.s_total_qual + i] += (uint64_t)tq;
    }
    int64_t nr = 0, ls = 0;
    for(uint32_t l = 0; l <= C; ++l){ const int64_t h = (int32_t)lds32(lh + 4u * l); nr += h; ls += h * l; }
    S[L.s_reads] += (uint64_t)nr; S[L.s_length_sum] += (uint64_t)ls;
}
}  // namespace

// returns 0, or the device error flag (bit0 too long, bit1 malformed batch) as a positive number
extern "C" __attribute__((visibility("default"))) int emu_pe_run(const fq_params* p, const fq_batch* b1, const fq_batch* b2, fq_read_result* o1, fq_read_result* o2,
                          fq_pair_result* op, uint64_t* flat, int T, int maxlen){
    DevParams D; fqk_dev_params_fill(D, p);
    const bool trace = getenv("EMU_TRACE") != nullptr;
    if(trace) fprintf(stderr, "emu_pe_run n=%u T=%d maxlen=%d\n", b1->n_reads, T, maxlen);
    uint32_t err = 0; D.err = &err;
    const fq_params& P = D.P;
    const int Wr = std::max(8, (maxlen + 3) / 4);
    int Wm = Wr;
    if(P.merge_enabled) Wm = std::max(Wr, (int)std::min<uint64_t>((D.L.max_cycles + 3) / 4, (uint64_t)(2 * maxlen + 3) / 4));
    const V4L L = v4_layout(Wr, Wm, T, 1, P.insert_size_max + 1);
    std::vector<uint8_t> sm(L.total + 256, 0);
    emu_smem = sm.data();
    const uint32_t S32 = L.o_stage;
    const fq_batch* B[2] = {b1, b2};
    const uint32_t n = b1->n_reads;
    SmStats pre[2] = {{L.o_pre[0], L.Cr}, {L.o_pre[1], L.Cr}}, post[2] = {{L.o_post[0], L.Cm}, {L.o_post[1], L.Cr}};
    Tally tl[32]; memset(tl, 0, sizeof(tl));
    for(uint32_t t0 = 0; t0 < n; t0 += (uint32_t)T){
        const uint32_t nt = std::min<uint32_t>((uint32_t)T, n - t0);
        memset(sm.data() + S32, 0xEE, L.stage_bytes);                    // stale bytes of an earlier tile
        bool bad = false;
        for(int m = 0; m < 2; ++m){                                      // the copy producer
            const fq_batch& b = *B[m];
            const uint32_t a = b.offsets[t0], last = b.offsets[t0 + nt - 1], ll = b.lengths[t0 + nt - 1];
            uint64_t e = (uint64_t)last + ((ll + 15u) & ~15u);
            if((a & 15u) || e < a || e - a > L.qcap || e > b.total_bases){ err |= 2u; bad = true; e = a; }
            const uint32_t qbytes = (uint32_t)(e - a), b0 = (a >> 2) & ~15u;
            const uint32_t bbytes = qbytes ? (((uint32_t)(e >> 2) + 15u) & ~15u) - b0 : 0u;
            memcpy(sm.data() + S32 + L.o_q[m], b.qual + a, qbytes);
            memcpy(sm.data() + S32 + L.o_b[m], b.bases2 + b0, std::min<uint64_t>(bbytes, b.total_bases / 4 - b0));
            for(uint32_t i = 0; i < (uint32_t)T; ++i){
                sts32(S32 + L.o_off[m] + 4u * i, i < nt ? b.offsets[t0 + i] : a);
                sts16(S32 + L.o_len[m] + 2u * i, i < nt ? b.lengths[t0 + i] : 0u);
            }
        }
        for(int k = 0; k < 8; ++k) sts32(S32 + L.o_cnt + 4u * k, 0u);
        sts32(S32 + L.o_cnt + 16u, bad ? 0u : nt);
        V4X X; X.D = &D; X.Lp = &L; X.S32 = S32;
        for(int m = 0; m < 2; ++m){ X.pre32[m] = L.o_pre[m]; X.post32[m] = L.o_post[m]; }
        X.dbg = 0; X.ins32 = L.o_ins; X.bc = reinterpret_cast<BlockCounters*>(sm.data() + L.o_bc);
        X.out1 = o1 + t0; X.out2 = o2 + t0; X.outp = op + t0; X.nt = bad ? 0 : (int)nt;
        for(int j = 0; j < T / 32; ++j){
            if(32 * j >= X.nt) break;
            if(trace) fprintf(stderr, " tile %u chunk %d\n", t0, j);
            ChunkArg arg{&X, j, tl};
            emu::run_warp(chunk_body, &arg);
        }
        // ---- the statistics role's contract, scalar ----
        for(int m = 0; m < 2; ++m){
            const uint32_t a0 = lds32(S32 + L.o_off[m]), bshift = (a0 >> 2) & 15u;
            const uint32_t q0 = S32 + L.o_q[m], bb0 = S32 + L.o_b[m] + bshift;
            auto byte_at = [&](uint32_t off, uint32_t pos){ return lds8(q0 + off + pos); };
            auto code_at = [&](uint32_t off, uint32_t pos){ const uint32_t by = lds8(q0 + off + pos); return (by & 0x80) ? 0u : (lds8(bb0 + ((off + pos) >> 2)) >> (2 * ((off + pos) & 3))) & 3u; };
            for(uint32_t i = 0; i < (uint32_t)X.nt; ++i){                // P: every raw read, to pre and post
                const uint32_t off = lds32(S32 + L.o_off[m] + 4u * i) - a0, len = lds16(S32 + L.o_len[m] + 2u * i);
                for(uint32_t c = 0; c < len; ++c){ add_base(pre[m], c, byte_at(off, c), code_at(off, c), 1); add_base(post[m], c, byte_at(off, c), code_at(off, c), 1); }
            }
            const uint32_t nex = lds32(S32 + L.o_cnt + 4u * m);
            for(uint32_t k = 0; k < nex; ++k){
                const uint2 d = lds64(S32 + L.o_ex[m] + 8u * k);
                const uint32_t off = (d.x & 0x1FFFFu) << 4, len = d.x >> 17, kind = (d.y >> 26) & 3u, start = d.y & 0x1FFFu, plen = (d.y >> 13) & 0x1FFFu;
                const uint32_t d0 = (kind == 2 && start == 0) ? plen : 0u;
                for(uint32_t c = d0; c < len; ++c) add_base(post[m], c, byte_at(off, c), code_at(off, c), -1);          // D
                if(kind == 2 && start > 0) for(uint32_t c = 0; c < plen; ++c) add_base(post[m], c, byte_at(off, start + c), code_at(off, start + c), 1);   // Z
            }
            if(m == 1){
                const uint32_t nm = lds32(S32 + L.o_cnt + 8u);
                for(uint32_t k = 0; k < nm; ++k){                        // M: cycle c of post1 = complement of read2 position pmax - c
                    const uint2 d = lds64(S32 + L.o_exm + 8u * k);
                    const uint32_t off = (d.x & 0x1FFFFu) << 4, pmax = d.x >> 17, len1 = d.y & 0x1FFFu, len2 = (d.y >> 13) & 0x1FFFu;
                    for(uint32_t c = len1; c < len1 + len2; ++c){
                        const uint32_t by = byte_at(off, pmax - c);
                        add_base(post[0], c, by, (by & 0x80) ? 0u : (code_at(off, pmax - c) ^ 1u), 1);
                    }
                }
            }
        }
    }
    // block retirement
    const fq_stats_layout& FL = D.L;
    retire_host(pre[0], FL, flat + FL.stats_off[0]); retire_host(pre[1], FL, flat + FL.stats_off[1]);
    retire_host(post[0], FL, flat + FL.stats_off[2]); retire_host(post[1], FL, flat + FL.stats_off[3]);
    BlockCounters* bc = reinterpret_cast<BlockCounters*>(sm.data() + L.o_bc);
    for(int l = 0; l < 32; ++l){
        for(int b = 0; b < 5; ++b){ bc->polyx_reads[b] += tl[l].polyx_reads[b]; bc->polyx_bases[b] += tl[l].polyx_bases[b]; }
        bc->adapter_reads += tl[l].adapter_reads; bc->adapter_bases += tl[l].adapter_bases;
    }
    for(int i = 0; i < FQK_FILTER_RESULT_TYPES; ++i) flat[FL.filter_off + i] += bc->filter[i];
    for(int b = 0; b < 5; ++b){ flat[FL.polyx_reads_off + b] += bc->polyx_reads[b]; flat[FL.polyx_bases_off + b] += bc->polyx_bases[b]; }
    flat[FL.adapter_reads_off] += bc->adapter_reads; flat[FL.adapter_bases_off] += bc->adapter_bases; flat[FL.merged_pairs_off] += bc->merged_pairs;
    for(uint32_t i = 0; i <= (uint32_t)P.insert_size_max; ++i) flat[FL.insert_hist_off + i] += lds32(L.o_ins + 4u * i);
    emu_smem = nullptr;
    return (int)err;
}
