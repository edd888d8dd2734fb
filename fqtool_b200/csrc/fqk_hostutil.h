// fqk_hostutil.h -- host-side helpers shared by fqk_api.cu and the CPU emulation harness (tests/emu):
// the accumulator layout of include/fqk.h and the device copy of the parameters.
#pragma once
#include <cstring>
#include "fqk_trim.cuh"

inline int fqk_layout_fill(const fq_params* p, fq_stats_layout* L){
    if(!p || !L) return FQK_E_INVALID;
    memset(L, 0, sizeof(*L));
    const uint64_t C = p->max_cycles > 0 ? (uint64_t)p->max_cycles : 512;
    L->max_cycles = C;
    L->kmer_bins = p->kmer_len > 0 ? (1ull << (2 * p->kmer_len)) : 0;
    L->s_q20 = 0; L->s_q30 = 8 * C; L->s_content = 16 * C; L->s_qualsum = 24 * C;
    L->s_total_base = 32 * C; L->s_total_qual = 33 * C;
    L->s_reads = 34 * C; L->s_length_sum = 34 * C + 1; L->s_kmer = 34 * C + 2;
    L->stats_stride = 34 * C + 2 + L->kmer_bins;
    for(int s = 0; s < 4; ++s) L->stats_off[s] = (uint64_t)s * L->stats_stride;
    uint64_t o = 4 * L->stats_stride;
    L->filter_off = o; o += FQK_FILTER_RESULT_TYPES;
    L->adapter_reads_off = o++; L->adapter_bases_off = o++;
    L->polyx_reads_off = o; o += 5; L->polyx_bases_off = o; o += 5;
    L->merged_pairs_off = o++;
    L->insert_hist_off = o; L->insert_hist_len = (uint64_t)p->insert_size_max + 1; o += L->insert_hist_len;
    L->correction_off = o; o += 64;
    L->corrected_reads_off = o; o += 1;
    L->total_u64 = o;
    L->dup_bins = p->dup_enabled ? (1ull << (2 * p->dup_keylen)) : 0;
    return 0;
}

inline int fqk_cls_of_char(char ch){
    switch(ch){ case 'A': return 0; case 'T': return 1; case 'C': return 2; case 'G': return 3; case 'N': return 4; default: return -1; }
}

// DevParams from validated fq_params: adapters as base classes and packed like reads, accumulator layout
inline void fqk_dev_params_fill(DevParams& D, const fq_params* p){
    memset(&D, 0, sizeof(D));
    D.P = *p;
    if(D.P.max_cycles == 0) D.P.max_cycles = 512;
    for(int i = 0; i < p->adapter_r1_len; ++i) D.ad1[i] = (uint8_t)fqk_cls_of_char(p->adapter_r1[i]);
    for(int i = 0; i < p->adapter_r2_len; ++i) D.ad2[i] = (uint8_t)fqk_cls_of_char(p->adapter_r2[i]);
    for(int i = 0; i < p->adapter_r1_len; ++i){
        const int cl = D.ad1[i];
        if(cl == 4) D.adn1[i >> 4] |= 1u << (2 * (i & 15)); else D.adc1[i >> 4] |= (uint32_t)cl << (2 * (i & 15));
    }
    for(int i = 0; i < p->adapter_r2_len; ++i){
        const int cl = D.ad2[i];
        if(cl == 4) D.adn2[i >> 4] |= 1u << (2 * (i & 15)); else D.adc2[i >> 4] |= (uint32_t)cl << (2 * (i & 15));
    }
    fqk_layout_fill(&D.P, &D.L);
}
