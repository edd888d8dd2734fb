// fqk_api.cu -- the C ABI of include/fqk.h: context, batch staging, kernel
// launches, run-level merge (NCCL) and the host-side packer.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <dlfcn.h>
#include "fqk_kernels.h"
#include "fqk_fused.h"
#include "fqk_hostutil.h"
#include "fqk_text.h"

// ------------------------------------------------------------------- NCCL
// NCCL is bound at run time (dlopen) so that the library also loads on hosts
// without it; only ncclUniqueId crosses the ABI (as 128 opaque bytes).
namespace nccl {
struct UniqueId { char internal[128]; };
typedef void* Comm;
typedef int (*GetUniqueId_t)(UniqueId*);
typedef int (*CommInitRank_t)(Comm*, int, UniqueId, int);
typedef int (*AllReduce_t)(const void*, void*, size_t, int, int, Comm, cudaStream_t);
typedef int (*CommDestroy_t)(Comm);
typedef const char* (*GetErrorString_t)(int);
enum { Uint8 = 1, Uint32 = 3, Uint64 = 5 };      // ncclDataType_t, nccl.h
enum { Sum = 0, Max = 2, Min = 3 };              // ncclRedOp_t, nccl.h
static void* handle = nullptr;
static GetUniqueId_t GetUniqueId; static CommInitRank_t CommInitRank; static AllReduce_t AllReduce;
static CommDestroy_t CommDestroy; static GetErrorString_t GetErrorString;
static bool load(){
    if(handle) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for(const char* n : names){ handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if(handle) break; }
    if(!handle) return false;
    GetUniqueId = (GetUniqueId_t)dlsym(handle, "ncclGetUniqueId");
    CommInitRank = (CommInitRank_t)dlsym(handle, "ncclCommInitRank");
    AllReduce = (AllReduce_t)dlsym(handle, "ncclAllReduce");
    CommDestroy = (CommDestroy_t)dlsym(handle, "ncclCommDestroy");
    GetErrorString = (GetErrorString_t)dlsym(handle, "ncclGetErrorString");
    return GetUniqueId && CommInitRank && AllReduce && CommDestroy;
}
}

// ---------------------------------------------------------------- context
struct DevBuf {
    void* p = nullptr; size_t cap = 0;
    int ensure(size_t bytes){
        if(bytes <= cap) return 0;
        if(p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        if(cudaMalloc(&p, want) != cudaSuccess){ p = nullptr; return FQK_E_NOMEM; }
        cap = want;
        return 0;
    }
    void release(){ if(p) cudaFree(p); p = nullptr; cap = 0; }
};

struct StagedBatch { DevBuf off, len, b2, q; DevBatch dev; };

struct fqk_ctx {
    DevParams D;
    int device = 0, rank = 0, nranks = 1, sms = 148;
    cudaStream_t stream = nullptr;
    uint64_t* flat = nullptr;                 // device accumulators (fq_stats_layout)
    uint32_t* err = nullptr;                  // device error flag
    uint32_t* maxlen2 = nullptr;              // device: [0..1] longest raw read of each mate, [2] number of corrections
    DevBuf corr; uint32_t corr_cap = 0;       // device list of fq_correction of the current PE batch
    // ASCII ingest staging (fqk_process_*_ascii): raw characters and source offsets per mate, pinned index arrays
    DevBuf aseq[2], aqual[2], asrc[2];
    std::vector<uint32_t> hsrc[2], hdst[2]; std::vector<uint16_t> hlen[2];
    // text entry points (fqk_process_*_text): names and third lines per mate, sizes/offsets, per-channel text blocks
    DevBuf tnames[2], tnoff[2], tstr[2], tsoff[2], tsizes, tscratch, ttot, tout[FQK_CH_COUNT];
    // raw FASTQ blocks (fqk_process_*_fastq): the text, newline counts/positions, quality offsets, line lengths
    DevBuf traw[2], tnlc[2], tnl[2], aqsrc[2], tnlen[2], tslen[2];
    // pipelined text path (fqk_submit_se_text / fqk_wait_text): per slot the 64-bit name / third-line offsets, sizes, totals, blocks
    DevBuf pno[2], pso[2], psz[2], ptt[2], pout[2][FQK_CH_COUNT];
    cudaEvent_t evTot[2] = {nullptr, nullptr};
    cudaStream_t tdl[2] = {nullptr, nullptr}; // one download stream per slot: the small downloads of pack k+1 (which wait for its
                                              // kernels) must not sit in front of the text download of pack k
    uint64_t* ptoth = nullptr;                // pinned: [2][16] totals + counters of a slot
    struct TextSlot { uint32_t n = 0, writers = 0; fq_text_out* outs = nullptr; bool open = false; } tslot[2];
    struct RawSlot { fq_raw_block blk; bool open = false; } rslot[2];      // fqk_submit_se_fastq / fqk_wait_fastq
    DevBuf poff[2], pscr[2], ptot[2];         // pipelined ASCII ingest (fqk_submit_se_ascii): offsets, scan scratch, totals per slot
    bool raw_mode = false;                    // the staged mates come from raw blocks (TextMate then points into traw)
    std::vector<uint32_t> hnoff[2], hsoff[2];
    unsigned long long* tab128 = nullptr; unsigned long long* first64 = nullptr;   // dup table
    uint64_t* d_dups = nullptr; uint32_t* d_counts = nullptr; uint8_t* d_gc = nullptr; bool dup_final = false;
    // cross-rank merge scratch, allocated once in fqk_create when nranks > 1 (no allocation between collectives)
    uint64_t* m_flat = nullptr; uint32_t* m_idx = nullptr; unsigned long long* m_first = nullptr; uint32_t* m_status = nullptr;
    int sticky = 0; std::string sticky_msg;   // first error of a pipelined pack (fqk_submit_se): reported by every later wait/finalize
    StagedBatch st[2];
    DevBuf res1, res2, resp;
    // pipelined single-end path (fqk_submit_se / fqk_wait): two slots, copies on their own streams
    cudaStream_t h2d = nullptr, d2h = nullptr;
    StagedBatch pst[2]; DevBuf pres[2];
    StagedBatch pst2[2]; DevBuf pres2[2], presp[2];   // read 2 and the pair records of a pipelined paired-end pack (fqk_submit_pe)
    cudaEvent_t evH2D[2] = {nullptr, nullptr}, evK[2] = {nullptr, nullptr}, evD2H[2] = {nullptr, nullptr};
    uint32_t* perr_host = nullptr;            // pinned: error flag snapshot of each slot
    uint64_t submitted = 0;
    nccl::Comm comm = nullptr;
    uint64_t launches = 0;
    int impl = 0;
    int stats_cs_se = 0, stats_cs_pe = 0;     // cycles privatised in shared memory (0 = global atomics)
    bool timing = false;
    struct Ev { cudaEvent_t a, b; int slot; };
    std::vector<Ev> evs;
    std::vector<cudaEvent_t> evpool;
    double t_ms[FQK_T_SLOTS] = {0}; uint64_t t_n[FQK_T_SLOTS] = {0};
    std::string error;
};

static thread_local std::string g_create_error;

// error text for the context-free entry points of other translation units (read back with fqk_last_error(NULL))
void fqk_set_thread_error(const char* msg){ g_create_error = msg ? msg : ""; }

static int fail(fqk_ctx* c, int code, const std::string& msg){
    if(c) c->error = msg; else g_create_error = msg;
    return code;
}
#define CU(call) do{ cudaError_t e_ = (call); if(e_ != cudaSuccess) return fail(c, FQK_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); }while(0)

static cudaEvent_t ev_get(fqk_ctx* c){
    if(!c->evpool.empty()){ cudaEvent_t e = c->evpool.back(); c->evpool.pop_back(); return e; }
    cudaEvent_t e; cudaEventCreate(&e); return e;
}
struct Timer {
    fqk_ctx* c; int slot; cudaEvent_t a = nullptr;
    Timer(fqk_ctx* c_, int s) : c(c_), slot(s) { if(c->timing){ a = ev_get(c); cudaEventRecord(a, c->stream); } }
    ~Timer(){ if(a){ cudaEvent_t b = ev_get(c); cudaEventRecord(b, c->stream); c->evs.push_back({a, b, slot}); } }
};
static void timers_collect(fqk_ctx* c){
    for(auto& e : c->evs){
        cudaEventSynchronize(e.b);
        float ms = 0; cudaEventElapsedTime(&ms, e.a, e.b);
        c->t_ms[e.slot] += ms; c->t_n[e.slot] += 1;
        c->evpool.push_back(e.a); c->evpool.push_back(e.b);
    }
    c->evs.clear();
}

// ----------------------------------------------------------------- params
extern "C" void fqk_default_params(fq_params* p){
    memset(p, 0, sizeof(*p));
    p->abi_version = FQK_ABI_VERSION;
    // defaults of /root/reference/src/options.h after CLI11's flag reset and Options::update()
    p->cut_front_window = p->cut_tail_window = p->cut_right_window = 4;
    p->cut_front_qual = p->cut_tail_qual = p->cut_right_qual = 20;
    p->polyg_min_len = 10; p->polyg_max_mismatch = 1; p->polyg_one_mismatch_each = 10;
    p->polyx_base_mask = 31; p->polyx_min_len = 10; p->polyx_max_mismatch = 1; p->polyx_one_mismatch_each = 10;
    p->low_quality_limit = 20 + 33; p->low_quality_base_limit = 22; p->n_base_limit = 5;
    p->min_read_length = 15; p->complexity_threshold = 0.3;
    p->overlap_require = 30; p->overlap_diff_limit = 5; p->insert_size_max = 512;
    p->dup_keylen = 12; p->max_cycles = 512;
}

extern "C" int fqk_stats_layout(const fq_params* p, fq_stats_layout* L){ return fqk_layout_fill(p, L); }

static int cls_of_char(char ch){ return fqk_cls_of_char(ch); }

static int validate_params(const fq_params* p, std::string& why){
    if(p->abi_version != FQK_ABI_VERSION){ why = "abi_version mismatch"; return FQK_E_INVALID; }
    if((p->cut_front_enabled && p->cut_front_window < 1) || (p->cut_tail_enabled && p->cut_tail_window < 1) ||
       (p->cut_right_enabled && p->cut_right_window < 1)){ why = "cut window must be >= 1"; return FQK_E_INVALID; }
    if(p->polyg_enabled && (p->polyg_one_mismatch_each < 1 || (p->is_paired && p->polyg_min_len < 1))){ why = "polyG one_mismatch_each must be >= 1"; return FQK_E_INVALID; }
    if(p->polyx_enabled && p->polyx_one_mismatch_each < 1){ why = "polyX one_mismatch_each must be >= 1"; return FQK_E_INVALID; }
    if(p->qual_filter_enabled && (p->low_quality_limit < 33 || p->low_quality_limit > 126)){ why = "low_quality_limit must be Phred+33 in 33..126"; return FQK_E_INVALID; }
    if(p->kmer_len != 0 && (p->kmer_len < 4 || p->kmer_len > 15)){ why = "kmer_len must be 0 or 4..15"; return FQK_E_INVALID; }
    if(p->dup_enabled && (p->dup_keylen < 12 || p->dup_keylen > 14)){ why = "dup_keylen must be 12..14 on the device"; return FQK_E_INVALID; }
    if(p->adapter_r1_len < 0 || p->adapter_r1_len > FQK_MAX_ADAPTER || p->adapter_r2_len < 0 || p->adapter_r2_len > FQK_MAX_ADAPTER){ why = "adapter too long"; return FQK_E_INVALID; }
    for(int i = 0; i < p->adapter_r1_len; ++i) if(cls_of_char(p->adapter_r1[i]) < 0){ why = "adapter_r1 outside ACGTN"; return FQK_E_ALPHABET; }
    for(int i = 0; i < p->adapter_r2_len; ++i) if(cls_of_char(p->adapter_r2[i]) < 0){ why = "adapter_r2 outside ACGTN"; return FQK_E_ALPHABET; }
    if(p->max_cycles < 0 || p->max_cycles > 65535){ why = "max_cycles out of range"; return FQK_E_INVALID; }
    if(p->insert_size_max < 1 || p->insert_size_max > 65534){ why = "insert_size_max out of range"; return FQK_E_INVALID; }
    if(p->trim_front1 < 0 || p->trim_tail1 < 0 || p->trim_front2 < 0 || p->trim_tail2 < 0){ why = "negative trim"; return FQK_E_INVALID; }
    return 0;
}

// ----------------------------------------------------------------- create
extern "C" int fqk_nccl_unique_id(void* out128){
    if(!out128) return FQK_E_INVALID;
    if(!nccl::load()) return FQK_E_NCCL;
    nccl::UniqueId id;
    if(nccl::GetUniqueId(&id) != 0) return FQK_E_NCCL;
    memcpy(out128, &id, sizeof(id));
    return 0;
}

extern "C" int fqk_get_corrections(fqk_ctx* c, fq_correction* out, uint32_t cap, uint32_t* n){
    if(!c || !n) return FQK_E_INVALID;
    *n = 0;
    if(!c->D.P.correction_enabled || !c->corr.p) return 0;
    CU(cudaSetDevice(c->device));
    uint32_t cnt = 0;
    CU(cudaMemcpyAsync(&cnt, c->maxlen2 + 2, sizeof(cnt), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *n = cnt;
    const uint32_t m = cnt < cap ? cnt : cap;
    if(out && m){
        CU(cudaMemcpyAsync(out, c->corr.p, (size_t)m * sizeof(fq_correction), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
    }
    return 0;
}

extern "C" int fqk_reset(fqk_ctx* c){
    if(!c) return FQK_E_INVALID;
    CU(cudaSetDevice(c->device));
    CU(cudaMemsetAsync(c->flat, 0, c->D.L.total_u64 * sizeof(uint64_t), c->stream));
    CU(cudaMemsetAsync(c->err, 0, sizeof(uint32_t), c->stream));
    c->sticky = 0; c->sticky_msg.clear();
    if(c->perr_host) c->perr_host[0] = c->perr_host[1] = 0;
    if(c->D.L.dup_bins){
        CU(cudaMemsetAsync(c->tab128, 0, c->D.L.dup_bins * 16, c->stream));
        CU(cudaMemsetAsync(c->first64, 0xFF, c->D.L.dup_bins * 8, c->stream));
        c->dup_final = false;
    }
    CU(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int fqk_create(fqk_ctx** out, const fq_params* p, int device_ordinal, int rank, int nranks,
                          const void* nccl_unique_id){
    if(!out || !p) return fail(nullptr, FQK_E_INVALID, "null argument");
    *out = nullptr;
    std::string why;
    int rc = validate_params(p, why);
    if(rc) return fail(nullptr, rc, why);
    if(nranks < 1 || rank < 0 || rank >= nranks) return fail(nullptr, FQK_E_INVALID, "bad rank/nranks");
    if(nranks > 1 && !nccl_unique_id) return fail(nullptr, FQK_E_INVALID, "nccl_unique_id required when nranks > 1");
    int ndev = 0;
    if(cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(nullptr, FQK_E_CUDA, "no CUDA device: the hot path has no CPU fallback");
    if(device_ordinal < 0 || device_ordinal >= ndev) return fail(nullptr, FQK_E_INVALID, "bad device ordinal");
    fqk_ctx* c = new fqk_ctx();
    c->device = device_ordinal; c->rank = rank; c->nranks = nranks;
    fqk_dev_params_fill(c->D, p);
    auto bail = [&](int code, const std::string& m){ std::string mm = m; fqk_destroy(c); return fail(nullptr, code, mm); };
    if(cudaSetDevice(device_ordinal) != cudaSuccess) return bail(FQK_E_CUDA, "cudaSetDevice failed");
    cudaDeviceProp prop;
    if(cudaGetDeviceProperties(&prop, device_ordinal) != cudaSuccess) return bail(FQK_E_CUDA, "cudaGetDeviceProperties failed");
    c->sms = prop.multiProcessorCount;
    if(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) return bail(FQK_E_CUDA, "stream create failed");
    if(cudaMalloc(&c->flat, c->D.L.total_u64 * sizeof(uint64_t)) != cudaSuccess) return bail(FQK_E_NOMEM, "accumulator alloc failed");
    if(cudaMalloc(&c->err, sizeof(uint32_t)) != cudaSuccess) return bail(FQK_E_NOMEM, "flag alloc failed");
    c->D.err = c->err;
    if(cudaMalloc(&c->maxlen2, 4 * sizeof(uint32_t)) != cudaSuccess) return bail(FQK_E_NOMEM, "scalar alloc failed");
    if(c->D.L.dup_bins){
        const uint64_t n = c->D.L.dup_bins;
        if(cudaMalloc(&c->tab128, n * 16) != cudaSuccess || cudaMalloc(&c->first64, n * 8) != cudaSuccess ||
           cudaMalloc(&c->d_dups, n * 8) != cudaSuccess || cudaMalloc(&c->d_counts, n * 4) != cudaSuccess ||
           cudaMalloc(&c->d_gc, n) != cudaSuccess) return bail(FQK_E_NOMEM, "dup table alloc failed");
    }
    // shared-memory budget of the generic stats kernels
    const size_t smem_max = prop.sharedMemPerBlockOptin > 2048 ? prop.sharedMemPerBlockOptin - 2048 : 0;
    const size_t C = c->D.L.max_cycles;
    c->stats_cs_se = (2 * 20 * C * 4 <= smem_max) ? (int)C : 0;
    c->stats_cs_pe = (4 * 20 * C * 4 <= smem_max) ? (int)C : 0;
    if(c->stats_cs_se) cudaFuncSetAttribute(k_stats_se, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 20 * (int)C * 4);
    if(c->stats_cs_pe) cudaFuncSetAttribute(k_stats_pe, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 20 * (int)C * 4);
    fused_configure(prop);
    if(nranks > 1){
        if(cudaMalloc(&c->m_flat, c->D.L.total_u64 * 8) != cudaSuccess || cudaMalloc(&c->m_status, 16) != cudaSuccess ||
           (c->D.L.dup_bins && (cudaMalloc(&c->m_idx, c->D.L.dup_bins * 4) != cudaSuccess ||
                                cudaMalloc(&c->m_first, c->D.L.dup_bins * 8) != cudaSuccess)))
            return bail(FQK_E_NOMEM, "merge scratch alloc failed");
        if(!nccl::load()) return bail(FQK_E_NCCL, "libnccl.so.2 not found");
        nccl::UniqueId id; memcpy(&id, nccl_unique_id, sizeof(id));
        int e = nccl::CommInitRank(&c->comm, nranks, id, rank);
        if(e != 0) return bail(FQK_E_NCCL, std::string("ncclCommInitRank: ") + (nccl::GetErrorString ? nccl::GetErrorString(e) : "?"));
    }
    rc = fqk_reset(c);
    if(rc){ std::string m = c->error; fqk_destroy(c); return fail(nullptr, rc, m); }
    *out = c;
    return 0;
}

extern "C" void fqk_destroy(fqk_ctx* c){
    if(!c) return;
    cudaSetDevice(c->device);
    if(c->stream) cudaStreamSynchronize(c->stream);
    if(c->comm) nccl::CommDestroy(c->comm);
    for(auto& e : c->evs){ cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    for(auto& e : c->evpool) cudaEventDestroy(e);
    for(int i = 0; i < 2; ++i){
        c->pst[i].off.release(); c->pst[i].len.release(); c->pst[i].b2.release(); c->pst[i].q.release(); c->pres[i].release();
        c->pst2[i].off.release(); c->pst2[i].len.release(); c->pst2[i].b2.release(); c->pst2[i].q.release();
        c->pres2[i].release(); c->presp[i].release();
        if(c->evH2D[i]) cudaEventDestroy(c->evH2D[i]);
        if(c->evK[i]) cudaEventDestroy(c->evK[i]);
        if(c->evD2H[i]) cudaEventDestroy(c->evD2H[i]);
    }
    if(c->h2d) cudaStreamDestroy(c->h2d);
    if(c->d2h) cudaStreamDestroy(c->d2h);
    if(c->perr_host) cudaFreeHost(c->perr_host);
    for(int i = 0; i < 2; ++i){ c->st[i].off.release(); c->st[i].len.release(); c->st[i].b2.release(); c->st[i].q.release(); }
    c->res1.release(); c->res2.release(); c->resp.release(); c->corr.release();
    for(int i = 0; i < 2; ++i){ c->aseq[i].release(); c->aqual[i].release(); c->asrc[i].release(); }
    for(int i = 0; i < 2; ++i){ c->tnames[i].release(); c->tnoff[i].release(); c->tstr[i].release(); c->tsoff[i].release(); }
    c->tsizes.release(); c->tscratch.release(); c->ttot.release();
    for(int i = 0; i < 2; ++i){ c->poff[i].release(); c->pscr[i].release(); c->ptot[i].release(); }
    for(int i = 0; i < 2; ++i){
        c->pno[i].release(); c->pso[i].release(); c->psz[i].release(); c->ptt[i].release();
        for(int ch = 0; ch < FQK_CH_COUNT; ++ch) c->pout[i][ch].release();
        if(c->evTot[i]) cudaEventDestroy(c->evTot[i]);
        if(c->tdl[i]) cudaStreamDestroy(c->tdl[i]);
    }
    if(c->ptoth) cudaFreeHost(c->ptoth);
    for(int i = 0; i < 2; ++i){ c->traw[i].release(); c->tnlc[i].release(); c->tnl[i].release(); c->aqsrc[i].release(); c->tnlen[i].release(); c->tslen[i].release(); }
    for(int i = 0; i < FQK_CH_COUNT; ++i) c->tout[i].release();
    cudaFree(c->flat); cudaFree(c->err); cudaFree(c->maxlen2); cudaFree(c->tab128); cudaFree(c->first64);
    cudaFree(c->d_dups); cudaFree(c->d_counts); cudaFree(c->d_gc);
    cudaFree(c->m_flat); cudaFree(c->m_idx); cudaFree(c->m_first); cudaFree(c->m_status);
    if(c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" const char* fqk_last_error(const fqk_ctx* c){
    return c ? c->error.c_str() : g_create_error.c_str();
}
extern "C" void* fqk_stream(fqk_ctx* c){ return c ? (void*)c->stream : nullptr; }
extern "C" uint64_t fqk_launch_count(const fqk_ctx* c){ return c ? c->launches : 0; }
extern "C" int fqk_set_impl(fqk_ctx* c, int impl){ if(!c || impl < 0 || impl > 3 || impl == 2) return FQK_E_INVALID; c->impl = impl; return 0; }
extern "C" int fqk_set_kernel_timing(fqk_ctx* c, int enabled){ if(!c) return FQK_E_INVALID; c->timing = enabled != 0; return 0; }
extern "C" int fqk_kernel_times(fqk_ctx* c, double* ms, uint64_t* launches){
    if(!c) return FQK_E_INVALID;
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    timers_collect(c);
    for(int i = 0; i < FQK_T_SLOTS; ++i){
        if(ms) ms[i] = c->t_ms[i];
        if(launches) launches[i] = c->t_n[i];
        c->t_ms[i] = 0; c->t_n[i] = 0;
    }
    return 0;
}

// --------------------------------------------------------------- launches
static int grid_for(fqk_ctx* c, uint32_t n, int per_block, int blocks_per_sm){
    long long need = ((long long)n + per_block - 1) / per_block;
    long long cap = (long long)c->sms * blocks_per_sm;
    return (int)std::max(1LL, std::min(need, cap));
}

static DevBatch to_dev(const fq_batch* b){
    DevBatch d; d.n = b->n_reads; d.first_index = b->first_index; d.total = b->total_bases; d.offsets = b->offsets; d.lengths = b->lengths;
    d.bases2 = b->bases2; d.qual = b->qual; d.max_len = b->max_len;
    return d;
}

static int launch_se(fqk_ctx* c, const DevBatch& B, fq_read_result* out){
    if(B.n == 0) return 0;
    Timer all(c, FQK_T_ALL);
    // k-mers as a pass of their own (k_kmer_se) after a fused kernel that runs WITHOUT them: k_se_v3 has no k-mer code and
    // is twice as fast as k_fused_se_ws, whose k-mer counting costs another 2.3 ms per 4 M reads (tools/prof_cfg4_parts.py)
    static const bool kmer_fused = fqk_tuning_env("FQK_KMER_FUSED") != nullptr;       // A/B switch: k-mers inside k_fused_se_ws as in round 1
    DevParams Dn = c->D; Dn.P.kmer_len = 0;
    const bool ksplit = c->impl == 0 && c->D.P.kmer_len >= 4 && c->D.P.kmer_len <= 13 && !kmer_fused && fused_supported_se(Dn, B);
    if(ksplit){
        { Timer t(c, FQK_T_FUSED);
          int nl = fused_launch_se(Dn, B, out, c->flat, c->sms, c->stream, c->impl);
          if(nl < 0) return fail(c, FQK_E_CUDA, "fused SE launch failed");
          c->launches += nl; }
        Timer t(c, FQK_T_STATS);
        const int K = c->D.P.kmer_len;
        const int use_smem = K <= 6;
        k_kmer_se<<<c->sms * 4, 512, use_smem ? (size_t)2 * 4 * ((size_t)1 << (2 * K)) : 0, c->stream>>>(c->D, B, out, c->flat, use_smem);
        c->launches++;
    }else if(c->impl != 1 && fused_supported_se(c->D, B)){
        Timer t(c, FQK_T_FUSED);
        int nl = fused_launch_se(c->D, B, out, c->flat, c->sms, c->stream, c->impl);
        if(nl < 0) return fail(c, FQK_E_CUDA, "fused SE launch failed");
        c->launches += nl;
    }else{
        { Timer t(c, FQK_T_TRIM);
          k_trim_se<<<grid_for(c, B.n, 256, 8), 256, 0, c->stream>>>(c->D, B, out, c->flat); c->launches++; }
        {
            Timer t(c, FQK_T_STATS);
            const int Cs = c->stats_cs_se;
            k_stats_se<<<grid_for(c, B.n, 8, Cs ? 1 : 4), 256, (size_t)2 * 20 * Cs * 4, c->stream>>>(c->D, B, out, c->flat, Cs);
            c->launches++;
        }
    }
    if(c->D.P.dup_enabled){
        Timer t(c, FQK_T_DUP);
        k_dup<<<grid_for(c, B.n, 256, 8), 256, 0, c->stream>>>(c->D, B, B, 0, c->tab128, c->first64); c->launches++;
        c->dup_final = false;
    }
    CU(cudaGetLastError());
    return 0;
}

static int launch_pe(fqk_ctx* c, const DevBatch& B1, const DevBatch& B2, fq_read_result* o1, fq_read_result* o2,
                     fq_pair_result* op){
    if(B1.n == 0) return 0;
    Timer all(c, FQK_T_ALL);
    const bool correct = c->D.P.correction_enabled != 0;
    const int Cs = c->stats_cs_pe;
    if(c->impl == 0 && pe4_supported(c->D, B1, B2)){
        // one fused launch: trim + filter + the four Stats objects (the duplication table keeps its own kernel)
        if(c->D.P.dup_enabled){
            Timer t(c, FQK_T_DUP);
            k_dup<<<grid_for(c, B1.n, 256, 8), 256, 0, c->stream>>>(c->D, B1, B2, 1, c->tab128, c->first64); c->launches++;
            c->dup_final = false;
        }
        Timer t(c, FQK_T_FUSED);
        int nl = pe4_launch(c->D, B1, B2, o1, o2, op, c->flat, c->sms, c->stream);
        if(nl < 0) return fail(c, FQK_E_CUDA, "fused PE launch failed");
        c->launches += nl;
        return 0;
    }
    // everything that reads the RAW pair runs before k_trim_pe when -c corrects the batch in place
    if(c->D.P.dup_enabled){
        Timer t(c, FQK_T_DUP);
        k_dup<<<grid_for(c, B1.n, 256, 8), 256, 0, c->stream>>>(c->D, B1, B2, 1, c->tab128, c->first64); c->launches++;
        c->dup_final = false;
    }
    if(correct){
        if(c->corr.ensure((size_t)5 * B1.n * sizeof(fq_correction) + 16)) return fail(c, FQK_E_NOMEM, "correction list allocation failed");
        c->corr_cap = 5u * B1.n;
        Timer t(c, FQK_T_STATS);
        k_stats_pe<<<grid_for(c, B1.n, 8, Cs ? 1 : 4), 256, (size_t)4 * 20 * Cs * 4, c->stream>>>(c->D, B1, B2, o1, o2, op, c->flat, Cs, 1);
        c->launches++;
    }
    { Timer t(c, FQK_T_TRIM);
      cudaMemsetAsync(c->maxlen2, 0, 3 * sizeof(uint32_t), c->stream);       // [0..1] longest raw reads, [2] corrections
      k_trim_pe<<<grid_for(c, B1.n, 128, 8), 128, 0, c->stream>>>(c->D, B1, B2, o1, o2, op, c->flat, c->maxlen2,
                                                                  correct ? (fq_correction*)c->corr.p : nullptr, c->maxlen2 + 2, c->corr_cap);
      c->launches++; }
    if(!correct && c->impl != 1 && ws_stats_supported_pe(c->D, B1, B2)){
        Timer t(c, FQK_T_FUSED);
        int nl = ws_stats_launch_pe(c->D, B1, B2, o1, o2, op, c->flat, c->sms, c->stream, c->maxlen2);
        if(nl < 0) return fail(c, FQK_E_CUDA, "PE statistics launch failed");
        c->launches += nl;
    }else{
        Timer t(c, FQK_T_STATS);
        k_stats_pe<<<grid_for(c, B1.n, 8, Cs ? 1 : 4), 256, (size_t)4 * 20 * Cs * 4, c->stream>>>(c->D, B1, B2, o1, o2, op, c->flat, Cs,
                                                                                                  correct ? 2 : 3);
        c->launches++;
    }
    CU(cudaGetLastError());
    return 0;
}

static int err_bits(uint32_t e, const char** msg){
    if(e & 1){ *msg = "a read (or merged read) is longer than max_cycles"; return FQK_E_LIMIT; }
    if(e & 2){ *msg = "batch is not dense/ordered/16-byte aligned (see fq_batch)"; return FQK_E_INVALID; }
    if(e & 4){ *msg = "base outside {A,C,G,T,N} or quality >= 128 in an ASCII batch"; return FQK_E_ALPHABET; }
    if(e & 8){ *msg = "raw block is not four-line FASTQ ('@' name, bases, '+' line, qualities of the same length)"; return FQK_E_INVALID; }
    return 0;
}
static int check_err_flag(fqk_ctx* c){
    uint32_t e = 0;
    CU(cudaMemcpyAsync(&e, c->err, sizeof(e), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    const char* m = nullptr;
    const int rc = err_bits(e, &m);
    return rc ? fail(c, rc, m) : 0;
}
// error flag snapshot of a pipelined pack: the first one sticks to the context until fqk_reset
static void note_pack_error(fqk_ctx* c, uint32_t e){
    const char* m = nullptr;
    const int rc = err_bits(e, &m);
    if(rc && !c->sticky){ c->sticky = rc; c->sticky_msg = m; }
}

extern "C" int fqk_sync(fqk_ctx* c){
    if(!c) return FQK_E_INVALID;
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    return check_err_flag(c);
}

extern "C" int fqk_process_se_device(fqk_ctx* c, const fq_batch* in, fq_read_result* out){
    if(!c || !in || (!out && in->n_reads)) return FQK_E_INVALID;
    if(c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for paired-end data");
    CU(cudaSetDevice(c->device));
    return launch_se(c, to_dev(in), out);
}

extern "C" int fqk_process_pe_device(fqk_ctx* c, const fq_batch* r1, const fq_batch* r2, fq_read_result* o1,
                                     fq_read_result* o2, fq_pair_result* op){
    if(!c || !r1 || !r2) return FQK_E_INVALID;
    if(!c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for single-end data");
    if(r1->n_reads != r2->n_reads) return fail(c, FQK_E_INVALID, "read1/read2 batches differ in size");
    CU(cudaSetDevice(c->device));
    return launch_pe(c, to_dev(r1), to_dev(r2), o1, o2, op);
}

// host batch -> device staging (4 H2D copies on the context's stream)
static int stage(fqk_ctx* c, StagedBatch& s, const fq_batch* b, cudaStream_t st = nullptr){
    if(!st) st = c->stream;
    const size_t n = b->n_reads;
    if(b->total_bases % FQK_BATCH_PAD) return fail(c, FQK_E_INVALID, "total_bases must be a multiple of FQK_BATCH_PAD");
    if(s.off.ensure(n * 4 + 16) || s.len.ensure(n * 2 + 16) || s.b2.ensure(b->total_bases / 4 + 16) || s.q.ensure(b->total_bases + 16))
        return fail(c, FQK_E_NOMEM, "device staging allocation failed");
    CU(cudaMemcpyAsync(s.off.p, b->offsets, n * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.len.p, b->lengths, n * 2, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.b2.p, b->bases2, b->total_bases / 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.q.p, b->qual, b->total_bases, cudaMemcpyHostToDevice, st));
    s.dev.n = b->n_reads; s.dev.first_index = b->first_index; s.dev.total = b->total_bases;
    s.dev.offsets = (const uint32_t*)s.off.p; s.dev.lengths = (const uint16_t*)s.len.p;
    s.dev.bases2 = (const uint8_t*)s.b2.p; s.dev.qual = (const uint8_t*)s.q.p;
    uint32_t mx = b->max_len;
    if(!mx && c->D.P.is_paired) for(size_t r = 0; r < n; ++r) mx = std::max<uint32_t>(mx, b->lengths[r]);   // the fused PE kernel sizes its tiles by it
    s.dev.max_len = mx;
    return 0;
}

extern "C" int fqk_process_se(fqk_ctx* c, const fq_batch* in, fq_read_result* out){
    if(!c || !in || (!out && in->n_reads)) return FQK_E_INVALID;
    if(c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for paired-end data");
    if(in->n_reads == 0) return 0;
    CU(cudaSetDevice(c->device));
    int rc = stage(c, c->st[0], in); if(rc) return rc;
    const size_t n = in->n_reads;
    if(c->res1.ensure(n * sizeof(fq_read_result))) return fail(c, FQK_E_NOMEM, "result allocation failed");
    rc = launch_se(c, c->st[0].dev, (fq_read_result*)c->res1.p); if(rc) return rc;
    CU(cudaMemcpyAsync(out, c->res1.p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    return check_err_flag(c);
}

// ASCII batch -> staged packed batch: index arrays on the host (O(reads)), characters packed by k_pack_ascii
static int stage_ascii(fqk_ctx* c, int m, const fq_ascii_batch* b){
    const uint32_t n = b->n_reads;
    StagedBatch& s = c->st[m];
    const uint64_t chars = b->off[n] - b->off[0];
    if(chars > 0xFFFFFFF0ull) return fail(c, FQK_E_LIMIT, "ASCII batch larger than 2^32 characters");
    // the characters go first: the two large copies (pinned source) run while the host builds the index arrays below
    if(c->aseq[m].ensure(chars + 16) || c->aqual[m].ensure(chars + 16))
        return fail(c, FQK_E_NOMEM, "device staging allocation failed");
    CU(cudaMemcpyAsync(c->aseq[m].p, b->seq + b->off[0], chars, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->aqual[m].p, b->qual + b->off[0], chars, cudaMemcpyHostToDevice, c->stream));
    c->hsrc[m].resize(n); c->hdst[m].resize(n); c->hlen[m].resize(n);
    uint64_t pos = 0, maxl = 0;
    for(uint32_t r = 0; r < n; ++r){
        const uint64_t l = b->off[r + 1] - b->off[r];
        if(l > FQK_MAX_READ_LEN){ cudaStreamSynchronize(c->stream); return fail(c, FQK_E_LIMIT, "read longer than FQK_MAX_READ_LEN"); }   // the copies read the caller's buffers
        c->hsrc[m][r] = (uint32_t)(b->off[r] - b->off[0]); c->hdst[m][r] = (uint32_t)pos; c->hlen[m][r] = (uint16_t)l;
        if(l > maxl) maxl = l;
        pos += (l + FQK_PAD - 1) / FQK_PAD * FQK_PAD;
        if(pos > 0xFFFFFFF0ull){ cudaStreamSynchronize(c->stream); return fail(c, FQK_E_LIMIT, "batch larger than 2^32 padded bases"); }
    }
    const uint64_t total = (pos + FQK_BATCH_PAD - 1) / FQK_BATCH_PAD * FQK_BATCH_PAD;
    if(s.off.ensure((size_t)n * 4 + 16) || s.len.ensure((size_t)n * 2 + 16) || s.b2.ensure(total / 4 + 16) || s.q.ensure(total + 16) ||
       c->asrc[m].ensure((size_t)n * 4 + 16)){
        cudaStreamSynchronize(c->stream);
        return fail(c, FQK_E_NOMEM, "device staging allocation failed");
    }
    CU(cudaMemcpyAsync(c->asrc[m].p, c->hsrc[m].data(), (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(s.off.p, c->hdst[m].data(), (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(s.len.p, c->hlen[m].data(), (size_t)n * 2, cudaMemcpyHostToDevice, c->stream));
    if(total > pos){                                                      // zero tail of the batch
        CU(cudaMemsetAsync((uint8_t*)s.q.p + pos, 0, total - pos, c->stream));
        CU(cudaMemsetAsync((uint8_t*)s.b2.p + pos / 4, 0, (total - pos) / 4, c->stream));
    }
    if(n){
        k_pack_ascii<<<grid_for(c, n, 16, 8), 256, 0, c->stream>>>(n, (const char*)c->aseq[m].p, (const char*)c->aqual[m].p,
            (const uint32_t*)c->asrc[m].p, (const uint32_t*)s.off.p, (const uint16_t*)s.len.p, (uint8_t*)s.b2.p, (uint8_t*)s.q.p, c->err);
        c->launches++;
    }
    s.dev.n = n; s.dev.first_index = b->first_index; s.dev.total = total;
    s.dev.offsets = (const uint32_t*)s.off.p; s.dev.lengths = (const uint16_t*)s.len.p;
    s.dev.bases2 = (const uint8_t*)s.b2.p; s.dev.qual = (const uint8_t*)s.q.p; s.dev.max_len = (uint32_t)maxl;
    return 0;
}

extern "C" int fqk_process_se_ascii(fqk_ctx* c, const fq_ascii_batch* in, fq_read_result* out){
    if(!c || !in || !in->off || (!out && in->n_reads)) return FQK_E_INVALID;
    if(c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for paired-end data");
    if(in->n_reads == 0) return 0;
    CU(cudaSetDevice(c->device));
    int rc = stage_ascii(c, 0, in); if(rc) return rc;
    const size_t n = in->n_reads;
    if(c->res1.ensure(n * sizeof(fq_read_result))) return fail(c, FQK_E_NOMEM, "result allocation failed");
    rc = launch_se(c, c->st[0].dev, (fq_read_result*)c->res1.p); if(rc) return rc;
    CU(cudaMemcpyAsync(out, c->res1.p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    return check_err_flag(c);
}

extern "C" int fqk_process_pe_ascii(fqk_ctx* c, const fq_ascii_batch* r1, const fq_ascii_batch* r2, fq_read_result* o1,
                                    fq_read_result* o2, fq_pair_result* op){
    if(!c || !r1 || !r2 || !r1->off || !r2->off) return FQK_E_INVALID;
    if(!c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for single-end data");
    if(r1->n_reads != r2->n_reads) return fail(c, FQK_E_INVALID, "read1/read2 batches differ in size");
    if(r1->n_reads == 0) return 0;
    if(!o1 || !o2 || !op) return FQK_E_INVALID;
    CU(cudaSetDevice(c->device));
    int rc = stage_ascii(c, 0, r1); if(rc) return rc;
    rc = stage_ascii(c, 1, r2); if(rc) return rc;
    const size_t n = r1->n_reads;
    if(c->res1.ensure(n * sizeof(fq_read_result)) || c->res2.ensure(n * sizeof(fq_read_result)) ||
       c->resp.ensure(n * sizeof(fq_pair_result))) return fail(c, FQK_E_NOMEM, "result allocation failed");
    rc = launch_pe(c, c->st[0].dev, c->st[1].dev, (fq_read_result*)c->res1.p, (fq_read_result*)c->res2.p, (fq_pair_result*)c->resp.p);
    if(rc) return rc;
    CU(cudaMemcpyAsync(o1, c->res1.p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(o2, c->res2.p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(op, c->resp.p, n * sizeof(fq_pair_result), cudaMemcpyDeviceToHost, c->stream));
    return check_err_flag(c);
}


// ---------------------------------------------------------------- text entry points (fqk_text.cu)
// names and third lines of one mate to the device (offsets rebased and narrowed to 32 bit)
static int stage_text(fqk_ctx* c, int m, const fq_text_batch* b){
    const uint32_t n = b->reads.n_reads;
    if(!b->name_off || !b->strand_off || (!b->names && b->name_off[n] > b->name_off[0]) || (!b->strands && b->strand_off[n] > b->strand_off[0]))
        return fail(c, FQK_E_INVALID, "text batch without names / third lines");
    const uint64_t nb = b->name_off[n] - b->name_off[0], sb = b->strand_off[n] - b->strand_off[0];
    if(nb > 0xFFFFFFF0ull || sb > 0xFFFFFFF0ull) return fail(c, FQK_E_LIMIT, "more than 2^32 name characters in a pack");
    if(c->tnames[m].ensure(nb + 16) || c->tstr[m].ensure(sb + 16) || c->tnoff[m].ensure((size_t)(n + 1) * 4) || c->tsoff[m].ensure((size_t)(n + 1) * 4))
        return fail(c, FQK_E_NOMEM, "device staging allocation failed");
    if(nb) CU(cudaMemcpyAsync(c->tnames[m].p, b->names + b->name_off[0], nb, cudaMemcpyHostToDevice, c->stream));
    if(sb) CU(cudaMemcpyAsync(c->tstr[m].p, b->strands + b->strand_off[0], sb, cudaMemcpyHostToDevice, c->stream));
    c->hnoff[m].resize(n + 1); c->hsoff[m].resize(n + 1);
    for(uint32_t i = 0; i <= n; ++i){
        c->hnoff[m][i] = (uint32_t)(b->name_off[i] - b->name_off[0]); c->hsoff[m][i] = (uint32_t)(b->strand_off[i] - b->strand_off[0]);
    }
    CU(cudaMemcpyAsync(c->tnoff[m].p, c->hnoff[m].data(), (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->tsoff[m].p, c->hsoff[m].data(), (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));                  // hnoff / hsoff are pageable and reused by the next call
    return 0;
}
static TextMate text_mate(fqk_ctx* c, int m){
    TextMate t;
    t.src = (const uint32_t*)c->asrc[m].p; t.len = (const uint16_t*)c->st[m].len.p;
    t.name_off = (const uint32_t*)c->tnoff[m].p; t.strand_off = (const uint32_t*)c->tsoff[m].p;
    if(c->raw_mode){                                   // all four lines of a record live in the raw text
        const char* raw = (const char*)c->traw[m].p;
        t.seq = t.qual = t.names = t.strands = raw;
        t.qsrc = (const uint32_t*)c->aqsrc[m].p; t.name_len = (const uint32_t*)c->tnlen[m].p; t.strand_len = (const uint32_t*)c->tslen[m].p;
    }else{
        t.seq = (const char*)c->aseq[m].p; t.qual = (const char*)c->aqual[m].p; t.qsrc = nullptr;
        t.names = (const char*)c->tnames[m].p; t.strands = (const char*)c->tstr[m].p; t.name_len = t.strand_len = nullptr;
    }
    return t;
}
// records on the device -> per-channel text blocks in the caller's buffers
static int text_finish(fqk_ctx* c, bool paired, uint32_t n, uint32_t writers, fq_text_out* outs, uint64_t* read_passed, int slot = 0){
    const TextMate m1 = text_mate(c, slot), m2 = paired ? text_mate(c, 1) : m1;      // slot: staging set of a single-end pack
    const fq_read_result* r1 = (const fq_read_result*)c->res1.p;
    const fq_read_result* r2 = paired ? (const fq_read_result*)c->res2.p : nullptr;
    const fq_pair_result* rp = paired ? (const fq_pair_result*)c->resp.p : nullptr;
    if(c->tsizes.ensure((size_t)FQK_CH_COUNT * n * 4 + 16) || c->tscratch.ensure(text_scan_scratch_u64(n) * 8) || c->ttot.ensure(256))
        return fail(c, FQK_E_NOMEM, "text scratch allocation failed");
    uint64_t* totals = (uint64_t*)c->ttot.p;                               // [FQK_CH_COUNT] bytes, then TEXT_SLOTS u32 counters
    uint32_t* counters = (uint32_t*)(totals + FQK_CH_COUNT);
    CU(cudaMemsetAsync(c->ttot.p, 0, 256, c->stream));
    if(paired && c->D.P.correction_enabled && c->corr.p)
        text_launch_apply_edits(c->stream, (const fq_correction*)c->corr.p, c->maxlen2 + 2, c->corr_cap, (char*)m1.seq, (char*)m1.qual,
                                m1.src, m1.qsrc, (char*)m2.seq, (char*)m2.qual, m2.src, m2.qsrc);
    text_launch_sizes(c->stream, paired, n, writers, m1, m2, r1, r2, rp, (uint32_t*)c->tsizes.p, counters);
    text_launch_scan(c->stream, n, (uint32_t*)c->tsizes.p, totals, (uint64_t*)c->tscratch.p);
    c->launches += paired && c->D.P.correction_enabled ? 5 : 4;
    struct { uint64_t tot[FQK_CH_COUNT]; uint32_t cnt[TEXT_SLOTS]; } h;
    CU(cudaMemcpyAsync(&h, c->ttot.p, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    int rc = 0;
    TextOut to;
    for(int ch = 0; ch < FQK_CH_COUNT; ++ch){
        outs[ch].len = h.tot[ch]; outs[ch].records = h.cnt[ch];
        to.p[ch] = nullptr;
        if(!h.tot[ch]) continue;
        if(!outs[ch].data) continue;                                      // channel not wanted: sizes and counts only
        if(h.tot[ch] > 0xFFFFFFF0ull) rc = fail(c, FQK_E_LIMIT, "a text channel of a pack exceeds 4 GiB: use smaller packs");
        else if(outs[ch].cap < h.tot[ch]) rc = fail(c, FQK_E_LIMIT, "text output buffer too small (fq_text_out.len holds the size needed)");
        else if(c->tout[ch].ensure(h.tot[ch] + 16)) rc = fail(c, FQK_E_NOMEM, "text block allocation failed");
        to.p[ch] = (char*)c->tout[ch].p;
    }
    if(read_passed) *read_passed = h.cnt[FQK_CH_COUNT];
    if(rc) return rc;
    text_launch_copy(c->stream, paired, n, writers, m1, m2, r1, r2, rp, (const uint32_t*)c->tsizes.p, to);
    c->launches++;
    for(int ch = 0; ch < FQK_CH_COUNT; ++ch)
        if(h.tot[ch] && outs[ch].data) CU(cudaMemcpyAsync(outs[ch].data, c->tout[ch].p, h.tot[ch], cudaMemcpyDeviceToHost, c->stream));
    CU(cudaGetLastError());
    return 0;
}

extern "C" int fqk_process_se_text(fqk_ctx* c, const fq_text_batch* in, uint32_t writers, fq_read_result* out, fq_text_out* outs,
                                   uint64_t* read_passed){
    if(!c || !in || !in->reads.off || !outs) return FQK_E_INVALID;
    if(c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for paired-end data");
    for(int ch = 0; ch < FQK_CH_COUNT; ++ch){ outs[ch].len = 0; outs[ch].records = 0; }
    if(read_passed) *read_passed = 0;
    const size_t n = in->reads.n_reads;
    if(n == 0) return 0;
    CU(cudaSetDevice(c->device));
    c->raw_mode = false;
    int rc = stage_ascii(c, 0, &in->reads); if(rc) return rc;
    rc = stage_text(c, 0, in); if(rc) return rc;
    if(c->res1.ensure(n * sizeof(fq_read_result))) return fail(c, FQK_E_NOMEM, "result allocation failed");
    rc = launch_se(c, c->st[0].dev, (fq_read_result*)c->res1.p); if(rc) return rc;
    if(out) CU(cudaMemcpyAsync(out, c->res1.p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    rc = text_finish(c, false, (uint32_t)n, writers, outs, read_passed); if(rc) return rc;
    return check_err_flag(c);
}

extern "C" int fqk_process_pe_text(fqk_ctx* c, const fq_text_batch* r1, const fq_text_batch* r2, uint32_t writers, fq_read_result* o1,
                                   fq_read_result* o2, fq_pair_result* op, fq_text_out* outs, uint64_t* read_passed){
    if(!c || !r1 || !r2 || !r1->reads.off || !r2->reads.off || !outs) return FQK_E_INVALID;
    if(!c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for single-end data");
    if(r1->reads.n_reads != r2->reads.n_reads) return fail(c, FQK_E_INVALID, "read1/read2 batches differ in size");
    for(int ch = 0; ch < FQK_CH_COUNT; ++ch){ outs[ch].len = 0; outs[ch].records = 0; }
    if(read_passed) *read_passed = 0;
    const size_t n = r1->reads.n_reads;
    if(n == 0) return 0;
    CU(cudaSetDevice(c->device));
    c->raw_mode = false;
    int rc = stage_ascii(c, 0, &r1->reads); if(rc) return rc;
    rc = stage_ascii(c, 1, &r2->reads); if(rc) return rc;
    rc = stage_text(c, 0, r1); if(rc) return rc;
    rc = stage_text(c, 1, r2); if(rc) return rc;
    if(c->res1.ensure(n * sizeof(fq_read_result)) || c->res2.ensure(n * sizeof(fq_read_result)) ||
       c->resp.ensure(n * sizeof(fq_pair_result))) return fail(c, FQK_E_NOMEM, "result allocation failed");
    rc = launch_pe(c, c->st[0].dev, c->st[1].dev, (fq_read_result*)c->res1.p, (fq_read_result*)c->res2.p, (fq_pair_result*)c->resp.p);
    if(rc) return rc;
    if(o1) CU(cudaMemcpyAsync(o1, c->res1.p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    if(o2) CU(cudaMemcpyAsync(o2, c->res2.p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    if(op) CU(cudaMemcpyAsync(op, c->resp.p, n * sizeof(fq_pair_result), cudaMemcpyDeviceToHost, c->stream));
    rc = text_finish(c, true, (uint32_t)n, writers, outs, read_passed); if(rc) return rc;
    return check_err_flag(c);
}


// ---------------------------------------------------------------- raw FASTQ blocks
// step 1: the block to the device, newlines counted and scanned; *lines = complete lines in the block
// the asynchronous half: upload, newline counts, their prefix sum (tot[0] = newlines of the block)
static int fastq_upload(fqk_ctx* c, int m, const fq_raw_block* b, cudaStream_t st, DevBuf& scratch, DevBuf& tot){
    if(b->bytes > 0xFFFFFFF0ull) return fail(c, FQK_E_LIMIT, "raw FASTQ block of 4 GiB or more: use smaller blocks");
    const uint32_t nc = (uint32_t)((b->bytes + NL_CHUNK - 1) / NL_CHUNK);
    if(c->traw[m].ensure(b->bytes + 64) || c->tnlc[m].ensure((size_t)nc * 4 + 16) || scratch.ensure(text_scan_scratch_u64(nc) * 8) || tot.ensure(256))
        return fail(c, FQK_E_NOMEM, "device staging allocation failed");
    if(b->bytes) CU(cudaMemcpyAsync(c->traw[m].p, b->text, b->bytes, cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync((char*)c->traw[m].p + b->bytes, 0, 32, st));                 // the pack kernel reads whole words
    fastq_launch_count(st, (const char*)c->traw[m].p, b->bytes, (uint32_t*)c->tnlc[m].p);
    text_launch_scan(st, nc, (uint32_t*)c->tnlc[m].p, (uint64_t*)tot.p, (uint64_t*)scratch.p, 1);
    c->launches += 4;
    return 0;
}
static int fastq_lines(fqk_ctx* c, int m, const fq_raw_block* b, uint32_t* lines){
    int rc = fastq_upload(c, m, b, c->stream, c->tscratch, c->ttot); if(rc) return rc;
    uint64_t nl = 0;
    CU(cudaMemcpyAsync(&nl, c->ttot.p, 8, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *lines = (uint32_t)nl;
    return 0;
}
// step 2: records 0..n-1 of the block indexed, destination offsets by a device prefix sum, bases packed; *consumed = bytes of those records
static int fastq_stage(fqk_ctx* c, int m, const fq_raw_block* b, uint32_t lines, uint32_t n, uint64_t* consumed){
    StagedBatch& s = c->st[m];
    *consumed = 0;
    if(n == 0){ s.dev = DevBatch(); s.dev.n = 0; return 0; }
    if(c->tnl[m].ensure((size_t)lines * 4 + 16) || c->asrc[m].ensure((size_t)n * 4 + 16) || c->aqsrc[m].ensure((size_t)n * 4 + 16) ||
       s.len.ensure((size_t)n * 2 + 16) || s.off.ensure((size_t)n * 4 + 16) || c->tnoff[m].ensure((size_t)n * 4 + 16) || c->tnlen[m].ensure((size_t)n * 4 + 16) ||
       c->tsoff[m].ensure((size_t)n * 4 + 16) || c->tslen[m].ensure((size_t)n * 4 + 16) || c->tscratch.ensure(text_scan_scratch_u64(n) * 8) ||
       c->ttot.ensure(256))
        return fail(c, FQK_E_NOMEM, "device staging allocation failed");
    const char* raw = (const char*)c->traw[m].p;
    uint32_t* misc = (uint32_t*)((uint64_t*)c->ttot.p + 8);                          // [0] longest read
    CU(cudaMemsetAsync(c->ttot.p, 0, 256, c->stream));
    fastq_launch_positions(c->stream, raw, b->bytes, (const uint32_t*)c->tnlc[m].p, (uint32_t*)c->tnl[m].p);
    fastq_launch_index(c->stream, raw, (const uint32_t*)c->tnl[m].p, n, (uint32_t*)c->asrc[m].p, (uint32_t*)c->aqsrc[m].p, (uint16_t*)s.len.p,
                       (uint32_t*)c->tnoff[m].p, (uint32_t*)c->tnlen[m].p, (uint32_t*)c->tsoff[m].p, (uint32_t*)c->tslen[m].p,
                       (uint32_t*)s.off.p, misc, c->err);
    text_launch_scan(c->stream, n, (uint32_t*)s.off.p, (uint64_t*)c->ttot.p, (uint64_t*)c->tscratch.p, 1);   // padded lengths -> offsets
    c->launches += 5;
    struct { uint64_t total; uint64_t pad[7]; uint32_t maxlen; } h;
    uint32_t last_nl = 0;
    CU(cudaMemcpyAsync(&h, c->ttot.p, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(&last_nl, (uint32_t*)c->tnl[m].p + (4 * (size_t)n - 1), 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *consumed = (uint64_t)last_nl + 1;
    if(h.total > 0xFFFFFFF0ull) return fail(c, FQK_E_LIMIT, "batch larger than 2^32 padded bases");
    const uint64_t total = (h.total + FQK_BATCH_PAD - 1) / FQK_BATCH_PAD * FQK_BATCH_PAD;
    if(s.b2.ensure(total / 4 + 16) || s.q.ensure(total + 16)) return fail(c, FQK_E_NOMEM, "device staging allocation failed");
    if(total > h.total){
        CU(cudaMemsetAsync((uint8_t*)s.q.p + h.total, 0, total - h.total, c->stream));
        CU(cudaMemsetAsync((uint8_t*)s.b2.p + h.total / 4, 0, (total - h.total) / 4, c->stream));
    }
    k_pack_ascii<<<grid_for(c, n, 16, 8), 256, 0, c->stream>>>(n, raw, raw, (const uint32_t*)c->asrc[m].p, (const uint32_t*)s.off.p,
        (const uint16_t*)s.len.p, (uint8_t*)s.b2.p, (uint8_t*)s.q.p, c->err, (const uint32_t*)c->aqsrc[m].p);
    c->launches++;
    s.dev.n = n; s.dev.first_index = b->first_index; s.dev.total = total;
    s.dev.offsets = (const uint32_t*)s.off.p; s.dev.lengths = (const uint16_t*)s.len.p;
    s.dev.bases2 = (const uint8_t*)s.b2.p; s.dev.qual = (const uint8_t*)s.q.p; s.dev.max_len = h.maxlen;
    return 0;
}

extern "C" int fqk_process_se_fastq(fqk_ctx* c, const fq_raw_block* in, uint32_t writers, fq_read_result* out, uint32_t out_cap,
                                    fq_text_out* outs, uint64_t* read_passed, uint32_t* n_reads, uint64_t* consumed){
    if(!c || !in || (!in->text && in->bytes) || !outs || !n_reads || !consumed) return FQK_E_INVALID;
    if(c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for paired-end data");
    for(int ch = 0; ch < FQK_CH_COUNT; ++ch){ outs[ch].len = 0; outs[ch].records = 0; }
    if(read_passed) *read_passed = 0;
    *n_reads = 0; *consumed = 0;
    CU(cudaSetDevice(c->device));
    c->raw_mode = true;
    uint32_t lines = 0;
    int rc = fastq_lines(c, 0, in, &lines); if(rc) return rc;
    const uint32_t n = lines / 4;
    if(n == 0) return 0;
    if(out && out_cap < n) return fail(c, FQK_E_LIMIT, "record array smaller than the number of reads in the block");
    rc = fastq_stage(c, 0, in, lines, n, consumed); if(rc) return rc;
    if(c->res1.ensure((size_t)n * sizeof(fq_read_result))) return fail(c, FQK_E_NOMEM, "result allocation failed");
    rc = launch_se(c, c->st[0].dev, (fq_read_result*)c->res1.p); if(rc) return rc;
    if(out) CU(cudaMemcpyAsync(out, c->res1.p, (size_t)n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    rc = text_finish(c, false, n, writers, outs, read_passed); if(rc) return rc;
    *n_reads = n;
    return check_err_flag(c);
}

extern "C" int fqk_process_pe_fastq(fqk_ctx* c, const fq_raw_block* in1, const fq_raw_block* in2, uint32_t writers, fq_read_result* o1,
                                    fq_read_result* o2, fq_pair_result* op, uint32_t out_cap, fq_text_out* outs, uint64_t* read_passed,
                                    uint32_t* n_pairs, uint64_t* consumed1, uint64_t* consumed2){
    if(!c || !in1 || !in2 || (!in1->text && in1->bytes) || (!in2->text && in2->bytes) || !outs || !n_pairs || !consumed1 || !consumed2) return FQK_E_INVALID;
    if(!c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for single-end data");
    for(int ch = 0; ch < FQK_CH_COUNT; ++ch){ outs[ch].len = 0; outs[ch].records = 0; }
    if(read_passed) *read_passed = 0;
    *n_pairs = 0; *consumed1 = *consumed2 = 0;
    CU(cudaSetDevice(c->device));
    c->raw_mode = true;
    uint32_t l1 = 0, l2 = 0;
    int rc = fastq_lines(c, 0, in1, &l1); if(rc) return rc;
    rc = fastq_lines(c, 1, in2, &l2); if(rc) return rc;
    const uint32_t n = std::min(l1, l2) / 4;                               // pairs go by position: as many as both blocks hold
    if(n == 0) return 0;
    if((o1 || o2 || op) && out_cap < n) return fail(c, FQK_E_LIMIT, "record arrays smaller than the number of pairs in the blocks");
    rc = fastq_stage(c, 0, in1, l1, n, consumed1); if(rc) return rc;
    rc = fastq_stage(c, 1, in2, l2, n, consumed2); if(rc) return rc;
    if(c->res1.ensure((size_t)n * sizeof(fq_read_result)) || c->res2.ensure((size_t)n * sizeof(fq_read_result)) ||
       c->resp.ensure((size_t)n * sizeof(fq_pair_result))) return fail(c, FQK_E_NOMEM, "result allocation failed");
    rc = launch_pe(c, c->st[0].dev, c->st[1].dev, (fq_read_result*)c->res1.p, (fq_read_result*)c->res2.p, (fq_pair_result*)c->resp.p);
    if(rc) return rc;
    if(o1) CU(cudaMemcpyAsync(o1, c->res1.p, (size_t)n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    if(o2) CU(cudaMemcpyAsync(o2, c->res2.p, (size_t)n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    if(op) CU(cudaMemcpyAsync(op, c->resp.p, (size_t)n * sizeof(fq_pair_result), cudaMemcpyDeviceToHost, c->stream));
    rc = text_finish(c, true, n, writers, outs, read_passed); if(rc) return rc;
    *n_pairs = n;
    return check_err_flag(c);
}

static int pipe_init(fqk_ctx* c){
    if(c->h2d) return 0;
    CU(cudaStreamCreateWithFlags(&c->h2d, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&c->d2h, cudaStreamNonBlocking));
    for(int i = 0; i < 2; ++i){
        CU(cudaEventCreateWithFlags(&c->evH2D[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->evK[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->evD2H[i], cudaEventDisableTiming));
    }
    CU(cudaHostAlloc((void**)&c->perr_host, 2 * sizeof(uint32_t), cudaHostAllocDefault));
    c->perr_host[0] = c->perr_host[1] = 0;
    return 0;
}

// Pipelined form of fqk_process_se_ascii: two packs in flight, and NOTHING per read on the host -- the
// per-read source offsets, lengths and destination offsets (a prefix sum of the padded lengths) are computed on the
// device from the caller's offset array, so the call only enqueues copies and kernels.  Slot s of the pipeline uses the
// staging buffers of "mate" s (a single-end context has no second mate).
extern "C" int fqk_submit_se_ascii(fqk_ctx* c, const fq_ascii_batch* in, fq_read_result* out, uint64_t* ticket){
    if(!c || !in || !in->off || !ticket || (!out && in->n_reads)) return FQK_E_INVALID;
    if(c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for paired-end data");
    CU(cudaSetDevice(c->device));
    int rc = pipe_init(c); if(rc) return rc;
    const uint64_t t = c->submitted;
    const int s = (int)(t & 1);
    const uint32_t n = in->n_reads;
    if(t >= 2){ CU(cudaEventSynchronize(c->evD2H[s])); note_pack_error(c, c->perr_host[s]); }
    if(n == 0){ CU(cudaEventRecord(c->evD2H[s], c->d2h)); *ticket = t; c->submitted = t + 1; return 0; }
    const uint64_t chars = in->off[n] - in->off[0];
    const uint64_t cap = ((chars + 15ull * n + 64) + FQK_BATCH_PAD - 1) / FQK_BATCH_PAD * FQK_BATCH_PAD;   // upper bound of the padded extent
    if(chars > 0xFFFFFFF0ull || cap > 0xFFFFFFF0ull) return fail(c, FQK_E_LIMIT, "ASCII batch larger than 2^32 (padded) characters");
    StagedBatch& S = c->st[s];
    if(c->aseq[s].ensure(chars + 16) || c->aqual[s].ensure(chars + 16) || c->asrc[s].ensure((size_t)n * 4 + 16) || c->poff[s].ensure((size_t)(n + 1) * 8) ||
       c->pscr[s].ensure(text_scan_scratch_u64(n) * 8) || c->ptot[s].ensure(64) || S.off.ensure((size_t)n * 4 + 16) || S.len.ensure((size_t)n * 2 + 16) ||
       S.b2.ensure(cap / 4 + 32) || S.q.ensure(cap + 80) || c->pres[s].ensure((size_t)n * sizeof(fq_read_result)))
        return fail(c, FQK_E_NOMEM, "device staging allocation failed");
    CU(cudaStreamWaitEvent(c->h2d, c->evK[s], 0));                      // the slot's previous pack is through the kernels
    CU(cudaMemcpyAsync(c->aseq[s].p, in->seq + in->off[0], chars, cudaMemcpyHostToDevice, c->h2d));
    CU(cudaMemcpyAsync(c->aqual[s].p, in->qual + in->off[0], chars, cudaMemcpyHostToDevice, c->h2d));
    CU(cudaMemcpyAsync(c->poff[s].p, in->off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, c->h2d));
    ascii_launch_index(c->h2d, n, (const uint64_t*)c->poff[s].p, (uint32_t*)c->asrc[s].p, (uint16_t*)S.len.p, (uint32_t*)S.off.p, c->err);
    text_launch_scan(c->h2d, n, (uint32_t*)S.off.p, (uint64_t*)c->ptot[s].p, (uint64_t*)c->pscr[s].p, 1);
    ascii_launch_zero_tail(c->h2d, (const uint64_t*)c->ptot[s].p, (uint8_t*)S.q.p, (uint8_t*)S.b2.p);
    k_pack_ascii<<<grid_for(c, n, 16, 8), 256, 0, c->h2d>>>(n, (const char*)c->aseq[s].p, (const char*)c->aqual[s].p, (const uint32_t*)c->asrc[s].p,
        (const uint32_t*)S.off.p, (const uint16_t*)S.len.p, (uint8_t*)S.b2.p, (uint8_t*)S.q.p, c->err);
    c->launches += 6;
    S.dev.n = n; S.dev.first_index = in->first_index; S.dev.total = cap; S.dev.max_len = 0;
    S.dev.offsets = (const uint32_t*)S.off.p; S.dev.lengths = (const uint16_t*)S.len.p;
    S.dev.bases2 = (const uint8_t*)S.b2.p; S.dev.qual = (const uint8_t*)S.q.p;
    CU(cudaEventRecord(c->evH2D[s], c->h2d));
    CU(cudaStreamWaitEvent(c->stream, c->evH2D[s], 0));
    rc = launch_se(c, S.dev, (fq_read_result*)c->pres[s].p); if(rc) return rc;
    CU(cudaEventRecord(c->evK[s], c->stream));
    CU(cudaStreamWaitEvent(c->d2h, c->evK[s], 0));
    CU(cudaMemcpyAsync(out, c->pres[s].p, (size_t)n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->d2h));
    CU(cudaMemcpyAsync(&c->perr_host[s], c->err, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->d2h));
    CU(cudaEventRecord(c->evD2H[s], c->d2h));
    *ticket = t; c->submitted = t + 1;
    return 0;
}


// Pipelined form of fqk_process_se_text: submit enqueues the upload, the per-read index, the packing, the hot path and the
// sizes / prefix sums of the text; wait reads the totals, launches the copy kernel and downloads the blocks.  With two
// packs in flight (submit k+1, then wait k) the upload of pack k+1 overlaps the assembly and the download of pack k.
extern "C" int fqk_submit_se_text(fqk_ctx* c, const fq_text_batch* in, uint32_t writers, fq_read_result* out, fq_text_out* outs,
                                  uint64_t* ticket){
    if(!c || !in || !in->reads.off || !in->name_off || !in->strand_off || !outs || !ticket) return FQK_E_INVALID;
    if(c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for paired-end data");
    CU(cudaSetDevice(c->device));
    int rc = pipe_init(c); if(rc) return rc;
    if(!c->ptoth){
        CU(cudaHostAlloc((void**)&c->ptoth, 2 * 16 * sizeof(uint64_t), cudaHostAllocDefault));
        for(int i = 0; i < 2; ++i){
            CU(cudaEventCreateWithFlags(&c->evTot[i], cudaEventDisableTiming));
            CU(cudaStreamCreateWithFlags(&c->tdl[i], cudaStreamNonBlocking));
        }
    }
    const uint64_t t = c->submitted;
    const int s = (int)(t & 1);
    const uint32_t n = in->reads.n_reads;
    if(c->tslot[s].open) return fail(c, FQK_E_INVALID, "fqk_wait_text of the pack two submissions back is missing");
    if(t >= 2){ CU(cudaEventSynchronize(c->evD2H[s])); note_pack_error(c, c->perr_host[s]); }
    for(int ch = 0; ch < FQK_CH_COUNT; ++ch){ outs[ch].len = 0; outs[ch].records = 0; }
    c->tslot[s].n = n; c->tslot[s].writers = writers; c->tslot[s].outs = outs; c->tslot[s].open = true;
    *ticket = t; c->submitted = t + 1;
    if(n == 0){ CU(cudaEventRecord(c->evTot[s], c->tdl[s])); return 0; }
    const uint64_t chars = in->reads.off[n] - in->reads.off[0];
    const uint64_t nb = in->name_off[n] - in->name_off[0], sb = in->strand_off[n] - in->strand_off[0];
    const uint64_t cap = ((chars + 15ull * n + 64) + FQK_BATCH_PAD - 1) / FQK_BATCH_PAD * FQK_BATCH_PAD;
    if(chars > 0xFFFFFFF0ull || cap > 0xFFFFFFF0ull || nb > 0xFFFFFFF0ull || sb > 0xFFFFFFF0ull){
        c->tslot[s].open = false; c->submitted = t;
        return fail(c, FQK_E_LIMIT, "text batch larger than 2^32 characters");
    }
    StagedBatch& S = c->st[s];
    if(c->aseq[s].ensure(chars + 16) || c->aqual[s].ensure(chars + 16) || c->asrc[s].ensure((size_t)n * 4 + 16) || c->poff[s].ensure((size_t)(n + 1) * 8) ||
       c->pscr[s].ensure(text_scan_scratch_u64(n) * 8) || c->ptot[s].ensure(64) || S.off.ensure((size_t)n * 4 + 16) || S.len.ensure((size_t)n * 2 + 16) ||
       S.b2.ensure(cap / 4 + 32) || S.q.ensure(cap + 80) || c->pres[s].ensure((size_t)n * sizeof(fq_read_result)) ||
       c->tnames[s].ensure(nb + 16) || c->tstr[s].ensure(sb + 16) || c->pno[s].ensure((size_t)(n + 1) * 8) || c->pso[s].ensure((size_t)(n + 1) * 8) ||
       c->tnoff[s].ensure((size_t)(n + 1) * 4) || c->tsoff[s].ensure((size_t)(n + 1) * 4) || c->psz[s].ensure((size_t)FQK_CH_COUNT * n * 4 + 16) ||
       c->ptt[s].ensure(256)){
        c->tslot[s].open = false; c->submitted = t;
        return fail(c, FQK_E_NOMEM, "device staging allocation failed");
    }
    cudaStream_t h = c->h2d;
    CU(cudaStreamWaitEvent(h, c->evK[s], 0));
    CU(cudaMemcpyAsync(c->aseq[s].p, in->reads.seq + in->reads.off[0], chars, cudaMemcpyHostToDevice, h));
    CU(cudaMemcpyAsync(c->aqual[s].p, in->reads.qual + in->reads.off[0], chars, cudaMemcpyHostToDevice, h));
    CU(cudaMemcpyAsync(c->poff[s].p, in->reads.off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, h));
    if(nb) CU(cudaMemcpyAsync(c->tnames[s].p, in->names + in->name_off[0], nb, cudaMemcpyHostToDevice, h));
    if(sb) CU(cudaMemcpyAsync(c->tstr[s].p, in->strands + in->strand_off[0], sb, cudaMemcpyHostToDevice, h));
    CU(cudaMemcpyAsync(c->pno[s].p, in->name_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, h));
    CU(cudaMemcpyAsync(c->pso[s].p, in->strand_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, h));
    ascii_launch_index(h, n, (const uint64_t*)c->poff[s].p, (uint32_t*)c->asrc[s].p, (uint16_t*)S.len.p, (uint32_t*)S.off.p, c->err);
    text_launch_scan(h, n, (uint32_t*)S.off.p, (uint64_t*)c->ptot[s].p, (uint64_t*)c->pscr[s].p, 1);
    ascii_launch_zero_tail(h, (const uint64_t*)c->ptot[s].p, (uint8_t*)S.q.p, (uint8_t*)S.b2.p);
    k_pack_ascii<<<grid_for(c, n, 16, 8), 256, 0, h>>>(n, (const char*)c->aseq[s].p, (const char*)c->aqual[s].p, (const uint32_t*)c->asrc[s].p,
        (const uint32_t*)S.off.p, (const uint16_t*)S.len.p, (uint8_t*)S.b2.p, (uint8_t*)S.q.p, c->err);
    text_launch_narrow(h, n + 1, (const uint64_t*)c->pno[s].p, (uint32_t*)c->tnoff[s].p);
    text_launch_narrow(h, n + 1, (const uint64_t*)c->pso[s].p, (uint32_t*)c->tsoff[s].p);
    c->launches += 8;
    S.dev.n = n; S.dev.first_index = in->reads.first_index; S.dev.total = cap; S.dev.max_len = 0;
    S.dev.offsets = (const uint32_t*)S.off.p; S.dev.lengths = (const uint16_t*)S.len.p;
    S.dev.bases2 = (const uint8_t*)S.b2.p; S.dev.qual = (const uint8_t*)S.q.p;
    CU(cudaEventRecord(c->evH2D[s], h));
    CU(cudaStreamWaitEvent(c->stream, c->evH2D[s], 0));
    rc = launch_se(c, S.dev, (fq_read_result*)c->pres[s].p); if(rc) return rc;
    // sizes and prefix sums of the text on the kernel stream; the scan scratch of the slot is free again by then
    TextMate m; 
    m.seq = (const char*)c->aseq[s].p; m.qual = (const char*)c->aqual[s].p; m.src = (const uint32_t*)c->asrc[s].p; m.qsrc = nullptr;
    m.len = (const uint16_t*)S.len.p; m.names = (const char*)c->tnames[s].p; m.name_off = (const uint32_t*)c->tnoff[s].p;
    m.strands = (const char*)c->tstr[s].p; m.strand_off = (const uint32_t*)c->tsoff[s].p; m.name_len = m.strand_len = nullptr;
    uint64_t* totals = (uint64_t*)c->ptt[s].p; uint32_t* counters = (uint32_t*)(totals + FQK_CH_COUNT);
    CU(cudaMemsetAsync(c->ptt[s].p, 0, 256, c->stream));
    if(c->pscr[s].ensure(text_scan_scratch_u64(n) * 8)) return fail(c, FQK_E_NOMEM, "text scratch allocation failed");
    text_launch_sizes(c->stream, false, n, writers, m, m, (const fq_read_result*)c->pres[s].p, nullptr, nullptr, (uint32_t*)c->psz[s].p, counters);
    text_launch_scan(c->stream, n, (uint32_t*)c->psz[s].p, totals, (uint64_t*)c->pscr[s].p);
    c->launches += 4;
    CU(cudaEventRecord(c->evK[s], c->stream));
    cudaStream_t dl = c->tdl[s];
    CU(cudaStreamWaitEvent(dl, c->evK[s], 0));
    if(out) CU(cudaMemcpyAsync(out, c->pres[s].p, (size_t)n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, dl));
    CU(cudaMemcpyAsync(c->ptoth + 16 * s, c->ptt[s].p, 16 * sizeof(uint64_t), cudaMemcpyDeviceToHost, dl));
    CU(cudaMemcpyAsync(&c->perr_host[s], c->err, sizeof(uint32_t), cudaMemcpyDeviceToHost, dl));
    CU(cudaEventRecord(c->evTot[s], dl));
    return 0;
}

extern "C" int fqk_wait_text(fqk_ctx* c, uint64_t ticket, uint64_t* read_passed){
    if(!c || ticket >= c->submitted) return FQK_E_INVALID;
    const int s = (int)(ticket & 1);
    if(!c->tslot[s].open || ticket + 2 < c->submitted) return fail(c, FQK_E_INVALID, "no text pack with this ticket is in flight");
    CU(cudaSetDevice(c->device));
    fqk_ctx::TextSlot& ts = c->tslot[s];
    ts.open = false;
    if(read_passed) *read_passed = 0;
    CU(cudaEventSynchronize(c->evTot[s]));
    if(ts.n == 0){ CU(cudaEventRecord(c->evD2H[s], c->tdl[s])); return c->sticky ? fail(c, c->sticky, c->sticky_msg) : 0; }
    note_pack_error(c, c->perr_host[s]);
    const uint64_t* tot = c->ptoth + 16 * s;
    const uint32_t* cnt = reinterpret_cast<const uint32_t*>(tot + FQK_CH_COUNT);
    int rc = 0;
    TextOut to;
    for(int ch = 0; ch < FQK_CH_COUNT; ++ch){
        ts.outs[ch].len = tot[ch]; ts.outs[ch].records = cnt[ch];
        to.p[ch] = nullptr;
        if(!tot[ch] || !ts.outs[ch].data) continue;
        if(tot[ch] > 0xFFFFFFF0ull) rc = fail(c, FQK_E_LIMIT, "a text channel of a pack exceeds 4 GiB: use smaller packs");
        else if(ts.outs[ch].cap < tot[ch]) rc = fail(c, FQK_E_LIMIT, "text output buffer too small (fq_text_out.len holds the size needed)");
        else if(c->pout[s][ch].ensure(tot[ch] + 16)) rc = fail(c, FQK_E_NOMEM, "text block allocation failed");
        to.p[ch] = (char*)c->pout[s][ch].p;
    }
    if(read_passed) *read_passed = cnt[FQK_CH_COUNT];
    if(!rc && !c->sticky){
        StagedBatch& S = c->st[s];
        TextMate m;
        m.seq = (const char*)c->aseq[s].p; m.qual = (const char*)c->aqual[s].p; m.src = (const uint32_t*)c->asrc[s].p; m.qsrc = nullptr;
        m.len = (const uint16_t*)S.len.p; m.names = (const char*)c->tnames[s].p; m.name_off = (const uint32_t*)c->tnoff[s].p;
        m.strands = (const char*)c->tstr[s].p; m.strand_off = (const uint32_t*)c->tsoff[s].p; m.name_len = m.strand_len = nullptr;
        // the copy kernel and the download run on the download stream: the kernel stream may already hold the next pack
        text_launch_copy(c->tdl[s], false, ts.n, ts.writers, m, m, (const fq_read_result*)c->pres[s].p, nullptr, nullptr, (const uint32_t*)c->psz[s].p, to);
        c->launches++;
        for(int ch = 0; ch < FQK_CH_COUNT; ++ch)
            if(to.p[ch]) CU(cudaMemcpyAsync(ts.outs[ch].data, to.p[ch], tot[ch], cudaMemcpyDeviceToHost, c->tdl[s]));
    }
    // the slot's staging buffers are read by the copy kernel: the next upload into this slot waits for it
    CU(cudaEventRecord(c->evD2H[s], c->tdl[s]));
    CU(cudaEventRecord(c->evK[s], c->tdl[s]));
    CU(cudaEventSynchronize(c->evD2H[s]));
    if(rc) return rc;
    return c->sticky ? fail(c, c->sticky, c->sticky_msg) : 0;
}


// Pipelined form of fqk_process_se_fastq.  submit: the block goes up and its newlines are counted (asynchronous, on the
// upload stream); wait: everything that needs the number of records on the host (index, packing, hot path, assembly,
// download).  With two blocks in flight (submit k+1, then wait k) the upload of block k+1 overlaps the kernels and the
// download of block k.
extern "C" int fqk_submit_se_fastq(fqk_ctx* c, const fq_raw_block* in, uint64_t* ticket){
    if(!c || !in || (!in->text && in->bytes) || !ticket) return FQK_E_INVALID;
    if(c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for paired-end data");
    CU(cudaSetDevice(c->device));
    int rc = pipe_init(c); if(rc) return rc;
    if(!c->ptoth){
        CU(cudaHostAlloc((void**)&c->ptoth, 2 * 16 * sizeof(uint64_t), cudaHostAllocDefault));
        for(int i = 0; i < 2; ++i){
            CU(cudaEventCreateWithFlags(&c->evTot[i], cudaEventDisableTiming));
            CU(cudaStreamCreateWithFlags(&c->tdl[i], cudaStreamNonBlocking));
        }
    }
    const uint64_t t = c->submitted;
    const int s = (int)(t & 1);
    if(c->rslot[s].open) return fail(c, FQK_E_INVALID, "fqk_wait_fastq of the block two submissions back is missing");
    rc = fastq_upload(c, s, in, c->h2d, c->pscr[s], c->ptot[s]); if(rc) return rc;
    CU(cudaMemcpyAsync(c->ptoth + 16 * s, c->ptot[s].p, 8, cudaMemcpyDeviceToHost, c->h2d));
    CU(cudaEventRecord(c->evH2D[s], c->h2d));
    c->rslot[s].blk = *in; c->rslot[s].open = true;
    *ticket = t; c->submitted = t + 1;
    return 0;
}

extern "C" int fqk_wait_fastq(fqk_ctx* c, uint64_t ticket, uint32_t writers, fq_read_result* out, uint32_t out_cap, fq_text_out* outs,
                              uint64_t* read_passed, uint32_t* n_reads, uint64_t* consumed){
    if(!c || ticket >= c->submitted || !outs || !n_reads || !consumed) return FQK_E_INVALID;
    const int s = (int)(ticket & 1);
    if(!c->rslot[s].open || ticket + 2 < c->submitted) return fail(c, FQK_E_INVALID, "no FASTQ block with this ticket is in flight");
    CU(cudaSetDevice(c->device));
    c->rslot[s].open = false;
    for(int ch = 0; ch < FQK_CH_COUNT; ++ch){ outs[ch].len = 0; outs[ch].records = 0; }
    if(read_passed) *read_passed = 0;
    *n_reads = 0; *consumed = 0;
    CU(cudaEventSynchronize(c->evH2D[s]));                              // the block is on the device, its newlines are counted
    CU(cudaStreamWaitEvent(c->stream, c->evH2D[s], 0));
    const uint32_t lines = (uint32_t)c->ptoth[16 * s];
    const uint32_t n = lines / 4;
    if(n == 0) return 0;
    if(out && out_cap < n) return fail(c, FQK_E_LIMIT, "record array smaller than the number of reads in the block");
    c->raw_mode = true;
    int rc = fastq_stage(c, s, &c->rslot[s].blk, lines, n, consumed); if(rc) return rc;
    if(c->res1.ensure((size_t)n * sizeof(fq_read_result))) return fail(c, FQK_E_NOMEM, "result allocation failed");
    rc = launch_se(c, c->st[s].dev, (fq_read_result*)c->res1.p); if(rc) return rc;
    if(out) CU(cudaMemcpyAsync(out, c->res1.p, (size_t)n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    rc = text_finish(c, false, n, writers, outs, read_passed, s); if(rc) return rc;
    *n_reads = n;
    return check_err_flag(c);
}

// Pipelined form of fqk_process_se: the upload of pack k+1 overlaps the kernels and the download of pack k
// (PCIe is full duplex).  At most two packs are in flight; buffers of a pack must stay valid until fqk_wait.
extern "C" int fqk_submit_se(fqk_ctx* c, const fq_batch* in, fq_read_result* out, uint64_t* ticket){
    if(!c || !in || !ticket || (!out && in->n_reads)) return FQK_E_INVALID;
    if(c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for paired-end data");
    CU(cudaSetDevice(c->device));
    int rcp = pipe_init(c); if(rcp) return rcp;
    const uint64_t t = c->submitted;                                    // the ticket is handed out only when the pack is enqueued
    const int s = (int)(t & 1);
    const size_t n = in->n_reads;
    if(t >= 2){                                                         // the slot's previous pack has left the device
        CU(cudaEventSynchronize(c->evD2H[s]));
        note_pack_error(c, c->perr_host[s]);
    }
    if(n == 0){ CU(cudaEventRecord(c->evD2H[s], c->d2h)); *ticket = t; c->submitted = t + 1; return 0; }
    if(c->pres[s].ensure(n * sizeof(fq_read_result))) return fail(c, FQK_E_NOMEM, "result allocation failed");
    CU(cudaStreamWaitEvent(c->h2d, c->evK[s], 0));                      // staging of this slot is free again
    int rc = stage(c, c->pst[s], in, c->h2d); if(rc) return rc;
    CU(cudaEventRecord(c->evH2D[s], c->h2d));
    CU(cudaStreamWaitEvent(c->stream, c->evH2D[s], 0));
    rc = launch_se(c, c->pst[s].dev, (fq_read_result*)c->pres[s].p); if(rc) return rc;
    CU(cudaEventRecord(c->evK[s], c->stream));
    CU(cudaStreamWaitEvent(c->d2h, c->evK[s], 0));
    CU(cudaMemcpyAsync(out, c->pres[s].p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->d2h));
    CU(cudaMemcpyAsync(&c->perr_host[s], c->err, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->d2h));
    CU(cudaEventRecord(c->evD2H[s], c->d2h));
    *ticket = t; c->submitted = t + 1;
    return 0;
}

// Pipelined form of fqk_process_pe (the reference runs up to 16 workers through processPairEnd side by side,
// /root/reference/src/peprocessor.cpp:546-566; here ONE thread keeps two pair packs in flight): same slots, streams and
// events as fqk_submit_se, both mates and three record arrays per slot.  fqk_wait(ticket) completes it.
extern "C" int fqk_submit_pe(fqk_ctx* c, const fq_batch* r1, const fq_batch* r2, fq_read_result* o1, fq_read_result* o2,
                             fq_pair_result* op, uint64_t* ticket){
    if(!c || !r1 || !r2 || !ticket) return FQK_E_INVALID;
    if(!c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for single-end data");
    if(r1->n_reads != r2->n_reads) return fail(c, FQK_E_INVALID, "read1/read2 batches differ in size");
    if(c->D.P.correction_enabled)
        return fail(c, FQK_E_INVALID, "base correction (-c) returns a per-pack edit list (fqk_get_corrections): use fqk_process_pe");
    if(r1->n_reads && (!o1 || !o2 || !op)) return FQK_E_INVALID;
    CU(cudaSetDevice(c->device));
    int rcp = pipe_init(c); if(rcp) return rcp;
    const uint64_t t = c->submitted;
    const int s = (int)(t & 1);
    const size_t n = r1->n_reads;
    if(t >= 2){                                                         // the slot's previous pack has left the device
        CU(cudaEventSynchronize(c->evD2H[s]));
        note_pack_error(c, c->perr_host[s]);
    }
    if(n == 0){ CU(cudaEventRecord(c->evD2H[s], c->d2h)); *ticket = t; c->submitted = t + 1; return 0; }
    if(c->pres[s].ensure(n * sizeof(fq_read_result)) || c->pres2[s].ensure(n * sizeof(fq_read_result)) ||
       c->presp[s].ensure(n * sizeof(fq_pair_result))) return fail(c, FQK_E_NOMEM, "result allocation failed");
    CU(cudaStreamWaitEvent(c->h2d, c->evK[s], 0));                      // staging of this slot is free again
    int rc = stage(c, c->pst[s], r1, c->h2d); if(rc) return rc;
    rc = stage(c, c->pst2[s], r2, c->h2d); if(rc) return rc;
    CU(cudaEventRecord(c->evH2D[s], c->h2d));
    CU(cudaStreamWaitEvent(c->stream, c->evH2D[s], 0));
    rc = launch_pe(c, c->pst[s].dev, c->pst2[s].dev, (fq_read_result*)c->pres[s].p, (fq_read_result*)c->pres2[s].p,
                   (fq_pair_result*)c->presp[s].p);
    if(rc) return rc;
    CU(cudaEventRecord(c->evK[s], c->stream));
    CU(cudaStreamWaitEvent(c->d2h, c->evK[s], 0));
    CU(cudaMemcpyAsync(o1, c->pres[s].p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->d2h));
    CU(cudaMemcpyAsync(o2, c->pres2[s].p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->d2h));
    CU(cudaMemcpyAsync(op, c->presp[s].p, n * sizeof(fq_pair_result), cudaMemcpyDeviceToHost, c->d2h));
    CU(cudaMemcpyAsync(&c->perr_host[s], c->err, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->d2h));
    CU(cudaEventRecord(c->evD2H[s], c->d2h));
    *ticket = t; c->submitted = t + 1;
    return 0;
}

extern "C" int fqk_wait(fqk_ctx* c, uint64_t ticket){
    if(!c || ticket >= c->submitted) return FQK_E_INVALID;
    CU(cudaSetDevice(c->device));
    if(ticket + 2 >= c->submitted){                                     // not recycled yet: wait for its download
        const int s = (int)(ticket & 1);
        CU(cudaEventSynchronize(c->evD2H[s]));
        note_pack_error(c, c->perr_host[s]);
    }
    return c->sticky ? fail(c, c->sticky, c->sticky_msg) : 0;           // sticky: an older pack's error is not lost either
}

extern "C" int fqk_process_pe(fqk_ctx* c, const fq_batch* r1, const fq_batch* r2, fq_read_result* o1,
                              fq_read_result* o2, fq_pair_result* op){
    if(!c || !r1 || !r2) return FQK_E_INVALID;
    if(!c->D.P.is_paired) return fail(c, FQK_E_INVALID, "context was created for single-end data");
    if(r1->n_reads != r2->n_reads) return fail(c, FQK_E_INVALID, "read1/read2 batches differ in size");
    if(r1->n_reads == 0) return 0;
    if(!o1 || !o2 || !op) return FQK_E_INVALID;
    CU(cudaSetDevice(c->device));
    int rc = stage(c, c->st[0], r1); if(rc) return rc;
    rc = stage(c, c->st[1], r2); if(rc) return rc;
    const size_t n = r1->n_reads;
    if(c->res1.ensure(n * sizeof(fq_read_result)) || c->res2.ensure(n * sizeof(fq_read_result)) ||
       c->resp.ensure(n * sizeof(fq_pair_result))) return fail(c, FQK_E_NOMEM, "result allocation failed");
    rc = launch_pe(c, c->st[0].dev, c->st[1].dev, (fq_read_result*)c->res1.p, (fq_read_result*)c->res2.p, (fq_pair_result*)c->resp.p);
    if(rc) return rc;
    CU(cudaMemcpyAsync(o1, c->res1.p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(o2, c->res2.p, n * sizeof(fq_read_result), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(op, c->resp.p, n * sizeof(fq_pair_result), cudaMemcpyDeviceToHost, c->stream));
    return check_err_flag(c);
}

// ---------------------------------------------------------------- finalize
// cross-rank merge helpers for the duplication table
__global__ void k_dup_split(uint64_t bins, const unsigned long long* tab128, unsigned long long* kmin){
    for(uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; k < bins; k += (uint64_t)gridDim.x * blockDim.x)
        kmin[k] = (uint32_t)(tab128[2 * k + 1] >> 32) ? tab128[2 * k] : ~0ull;
}
__global__ void k_dup_mask(uint64_t bins, const unsigned long long* tab128, const unsigned long long* kmin,
                           uint32_t* cnt, uint32_t* idx){
    for(uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; k < bins; k += (uint64_t)gridDim.x * blockDim.x){
        const unsigned long long lo = tab128[2 * k], hi = tab128[2 * k + 1];
        const uint32_t c = (uint32_t)(hi >> 32);
        const bool keep = c && lo == kmin[k];
        cnt[k] = keep ? c : 0; idx[k] = keep ? (uint32_t)hi : 0xFFFFFFFFu;
    }
}
// merged (kmer32 minimum, count, index of the first read carrying it, first read of the key) -> mDups / mCounts / mGC;
// dups aliases kmin.  The local table (tab128, first64) is left untouched, so a later finalize merges it again correctly.
__global__ void k_dup_finish_merged(uint64_t bins, unsigned long long* kmin_dups, const uint32_t* cnt, const uint32_t* idx,
                                    const unsigned long long* first, uint8_t* gcs){
    for(uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; k < bins; k += (uint64_t)gridDim.x * blockDim.x){
        const uint32_t c = cnt[k];
        const unsigned long long f = first[k];
        kmin_dups[k] = c ? kmin_dups[k] : 0ull;
        gcs[k] = (c && (uint32_t)(f >> 8) == idx[k]) ? (uint8_t)(f & 0xFF) : 0;
    }
}

// Every rank runs the same sequence of collectives or none of them (fqk_finalize agrees on that first), nothing is
// allocated and no rank can leave between two all-reduces; launch and NCCL errors are collected and reported at the end.
static int dup_finalize(fqk_ctx* c, bool collective = false){
    const uint64_t n = c->D.L.dup_bins;
    if(!n || (c->dup_final && !collective)) return 0;            // ranks must not disagree on whether the collectives run
    const int g = c->sms * 8;
    if(c->nranks > 1){
        // min over kmer32, then sum of the counts / min of the indices that carry the global minimum,
        // and min over the packed (first index, gc) word: 4 all-reduces, once per run
        unsigned long long* kmin = (unsigned long long*)c->d_dups;       // the output arrays double as merge buffers
        uint32_t* cnt = c->d_counts; uint32_t* idx = c->m_idx;
        int e = 0; cudaError_t ce = cudaSuccess;
        auto chk = [&](){ cudaError_t x = cudaGetLastError(); if(ce == cudaSuccess) ce = x; };
        k_dup_split<<<g, 256, 0, c->stream>>>(n, c->tab128, kmin); chk();
        { int x = nccl::AllReduce(kmin, kmin, n, nccl::Uint64, nccl::Min, c->comm, c->stream); if(!e) e = x; }
        k_dup_mask<<<g, 256, 0, c->stream>>>(n, c->tab128, kmin, cnt, idx); chk();
        { int x = nccl::AllReduce(cnt, cnt, n, nccl::Uint32, nccl::Sum, c->comm, c->stream); if(!e) e = x; }
        { int x = nccl::AllReduce(idx, idx, n, nccl::Uint32, nccl::Min, c->comm, c->stream); if(!e) e = x; }
        { int x = nccl::AllReduce(c->first64, c->m_first, n, nccl::Uint64, nccl::Min, c->comm, c->stream); if(!e) e = x; }
        k_dup_finish_merged<<<g, 256, 0, c->stream>>>(n, kmin, cnt, idx, c->m_first, c->d_gc); chk();
        c->launches += 3;
        { cudaError_t x = cudaStreamSynchronize(c->stream); if(ce == cudaSuccess) ce = x; }
        if(e) return fail(c, FQK_E_NCCL, "ncclAllReduce (dup table) failed");
        if(ce != cudaSuccess) return fail(c, FQK_E_CUDA, std::string("dup table merge: ") + cudaGetErrorString(ce));
    }else{
        k_dup_finish<<<g, 256, 0, c->stream>>>(n, c->tab128, c->first64, c->d_dups, c->d_counts, c->d_gc);
        c->launches++;
        CU(cudaGetLastError());
    }
    c->dup_final = true;
    return 0;
}

extern "C" int fqk_finalize(fqk_ctx* c, uint64_t* flat){
    if(!c || !flat) return FQK_E_INVALID;
    CU(cudaSetDevice(c->device));
    int rc = c->sticky ? fail(c, c->sticky, c->sticky_msg) : check_err_flag(c);
    const size_t n = c->D.L.total_u64;
    if(c->nranks > 1){
        // Agree first: one all-reduce(max) of a status word.  A rank whose shard failed (over-length read, bad batch)
        // must not leave the others blocked inside ncclAllReduce: either every rank runs all collectives or none does.
        uint32_t st = rc ? 1u : 0u, gst = 1u;
        int e = 0;
        cudaError_t ce = cudaMemcpyAsync(c->m_status, &st, 4, cudaMemcpyHostToDevice, c->stream);
        if(ce == cudaSuccess) ce = cudaStreamSynchronize(c->stream);      // st is a stack variable
        e = nccl::AllReduce(c->m_status, c->m_status + 1, 1, nccl::Uint32, nccl::Max, c->comm, c->stream);
        if(ce == cudaSuccess) ce = cudaMemcpyAsync(&gst, c->m_status + 1, 4, cudaMemcpyDeviceToHost, c->stream);
        if(ce == cudaSuccess) ce = cudaStreamSynchronize(c->stream);
        if(rc) return rc;
        if(e) return fail(c, FQK_E_NCCL, "ncclAllReduce (status) failed");
        if(ce != cudaSuccess) return fail(c, FQK_E_CUDA, std::string("finalize: ") + cudaGetErrorString(ce));
        if(gst) return fail(c, FQK_E_INVALID, "another rank reported an error: the run-level merge was skipped on every rank");
        // Stats::merge / FilterResult::merge are plain sums: one all-reduce of the flat buffer.
        // The result is left in a scratch copy so that finalize can be called again.
        e = nccl::AllReduce(c->flat, c->m_flat, n, nccl::Uint64, nccl::Sum, c->comm, c->stream);
        ce = cudaMemcpyAsync(flat, c->m_flat, n * 8, cudaMemcpyDeviceToHost, c->stream);
        { cudaError_t x = cudaStreamSynchronize(c->stream); if(ce == cudaSuccess) ce = x; }
        const int rd = dup_finalize(c, true);                                  // runs on every rank whatever happened above
        if(e) return fail(c, FQK_E_NCCL, "ncclAllReduce (stats) failed");
        if(ce != cudaSuccess) return fail(c, FQK_E_CUDA, std::string("finalize: ") + cudaGetErrorString(ce));
        return rd;
    }
    if(rc) return rc;
    CU(cudaMemcpyAsync(flat, c->flat, n * 8, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return dup_finalize(c);
}

extern "C" int fqk_dup_table(fqk_ctx* c, uint64_t* dups, uint32_t* counts, uint8_t* gc){
    if(!c) return FQK_E_INVALID;
    const uint64_t n = c->D.L.dup_bins;
    if(!n) return fail(c, FQK_E_INVALID, "duplication analysis is not enabled");
    CU(cudaSetDevice(c->device));
    if(c->nranks > 1 && !c->dup_final) return fail(c, FQK_E_INVALID, "call fqk_finalize first: the cross-rank merge is collective");
    int rc = dup_finalize(c); if(rc) return rc;
    if(dups) CU(cudaMemcpyAsync(dups, c->d_dups, n * 8, cudaMemcpyDeviceToHost, c->stream));
    if(counts) CU(cudaMemcpyAsync(counts, c->d_counts, n * 4, cudaMemcpyDeviceToHost, c->stream));
    if(gc) CU(cudaMemcpyAsync(gc, c->d_gc, n, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return 0;
}

// ------------------------------------------------------------------ packer
extern "C" uint64_t fqk_packed_bases(uint32_t n, const uint64_t* src_off){
    uint64_t t = 0;
    for(uint32_t r = 0; r < n; ++r){ uint64_t l = src_off[r + 1] - src_off[r]; t += (l + FQK_PAD - 1) / FQK_PAD * FQK_PAD; }
    return (t + FQK_BATCH_PAD - 1) / FQK_BATCH_PAD * FQK_BATCH_PAD;
}

extern "C" int fqk_pack_ascii(uint32_t n, const char* seq, const char* qual, const uint64_t* src_off,
                              uint32_t* offsets, uint16_t* lengths, uint8_t* bases2, uint8_t* qual_out){
    static int8_t lut[256]; static bool init = false;
    if(!init){ for(int i = 0; i < 256; ++i) lut[i] = -1; lut['A'] = 0; lut['T'] = 1; lut['C'] = 2; lut['G'] = 3; lut['N'] = 4; init = true; }
    uint64_t pos = 0;
    for(uint32_t r = 0; r < n; ++r){
        const uint64_t l = src_off[r + 1] - src_off[r];
        if(l > FQK_MAX_READ_LEN) return FQK_E_LIMIT;
        if(pos > 0xFFFFFFFFull) return FQK_E_LIMIT;
        offsets[r] = (uint32_t)pos; lengths[r] = (uint16_t)l;
        const uint64_t padded = (l + FQK_PAD - 1) / FQK_PAD * FQK_PAD;
        const char* s = seq + src_off[r]; const char* q = qual + src_off[r];
        uint8_t* qo = qual_out + pos; uint8_t* bo = bases2 + pos / 4;
        memset(bo, 0, padded / 4);
        for(uint64_t i = 0; i < l; ++i){
            const int c = lut[(uint8_t)s[i]];
            if(c < 0) return FQK_E_ALPHABET;
            const uint8_t qq = (uint8_t)q[i];
            if(qq & 0x80) return FQK_E_ALPHABET;
            if(c == 4) qo[i] = qq | 0x80;
            else { qo[i] = qq; bo[i >> 2] |= (uint8_t)(c << ((i & 3) * 2)); }
        }
        memset(qo + l, 0, padded - l);
        pos += padded;
    }
    const uint64_t total = (pos + FQK_BATCH_PAD - 1) / FQK_BATCH_PAD * FQK_BATCH_PAD;   // zero tail of the batch
    memset(qual_out + pos, 0, total - pos);
    memset(bases2 + pos / 4, 0, (total - pos) / 4);
    return 0;
}

extern "C" void* fqk_host_alloc(size_t bytes){
    void* p = nullptr;
    if(cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess){ cudaGetLastError(); return nullptr; }
    return p;
}
extern "C" void fqk_host_free(void* p){ if(p) cudaFreeHost(p); }
