// bvc_ctx.h -- the context behind the opaque `bvc_ctx` of include/bvc_gpu.h, shared by bvc_api.cu (tile path)
// and bvc_shard_api.cu (sample-sharded path).  Private to the library.
#pragma once
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/bvc_gpu.h"
#include "bvc_kernels.h"
#include "bvc_pipeline_args.h"

namespace bvc {

constexpr size_t kCtrWords = 16 + kNBins + 16;
constexpr int kCtrErr = 6;         // d_ctr word of the device error bits
constexpr int kMaxClsStride = 4 * BVC_PACK8_MAX_QUALS + 4;   // pool words per site of site_hist8_kernel's fixed layout (pad4(major) + the rest)

struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    uint8_t *d_ref = nullptr, *d_code = nullptr, *d_qual = nullptr, *d_mapq = nullptr, *d_rpr = nullptr;
    uint8_t *d_pk8 = nullptr;      // a tile's packed8 form: what K1 / K5 read (expanded into d_code / d_qual only for the paths that need the arrays)
    bool packed = false;           // the slot's tile is resident in the one-byte form only (d_code / d_qual not filled)
    QualDict dict;                 // its qual dictionary
    int n_dict = 0;
    uint32_t* h_err = nullptr;     // pinned: the device error word of the slot's last launch (kErr* bits)
    uint32_t *d_off = nullptr, *d_sample = nullptr;
    bvc_site_result* d_out = nullptr;
    unsigned int* d_counter = nullptr;
    int32_t* d_gdepth = nullptr;   // --group outputs [S][n_groups][4]
    double* d_gaf = nullptr;
    // staged-pipeline intermediates (bvc_pipeline.cuh)
    SiteSum* d_sum = nullptr;
    uint32_t* d_cls = nullptr;
    Ladder* d_lad = nullptr;
    FitRes* d_fit = nullptr;
    uint8_t* d_state = nullptr;
    uint32_t *d_qa = nullptr, *d_qb = nullptr, *d_qs = nullptr, *d_ql = nullptr, *d_vcf_list = nullptr, *d_legacy_list = nullptr;
    uint32_t* d_ctr = nullptr;     // [kCtrWords]: cls_cursor, qa_n, qb_n, -, vcf_n, legacy_n, -, -, work[4] (u64 at word 8), q_head / long_n / long_head at 17..19, bins at word 20, deep_n behind them
    const uint8_t *k_mapq = nullptr, *k_rpr = nullptr;   // what the kernels read: d_mapq/d_rpr or zero-copy host pointers
    // split result (bvc_enqueue_split): device core records, compacted full records, their count; host side of the ticket
    bvc_site_core* d_core = nullptr;
    bvc_site_result* d_vcfc = nullptr;
    uint32_t* d_vcfc_n = nullptr;
    uint32_t* h_vcfc_n = nullptr;          // pinned
    bvc_site_result* split_vcf = nullptr;  // caller's vcf[] of the ticket in flight (nullptr: not a split ticket)
    int32_t split_cap = 0;
    bool is_split = false;                 // the slot's last ticket came from bvc_enqueue_split
    uint64_t retired_ticket = 0;           // last split ticket completed on this slot, its count and status
    int32_t retired_n = 0;
    int retired_rc = 0;
    uint64_t ticket = 0;       // 0 = idle
    uint64_t last_ticket = 0;  // ticket of the tile whose outputs the slot's device buffers hold
    int n_sites = 0;
    int64_t n_entries = 0;
};

struct ShardState;                 // bvc_shard_api.cu

}  // namespace bvc

struct bvc_ctx {
    int device = 0;
    int n_sm = 0;
    bvc_params prm;
    std::vector<int32_t> sample_group;
    bvc::Slot slot[2];
    double* d_lut = nullptr;
    double* d_lgam = nullptr;
    int32_t* d_sample_group = nullptr;
    void* d_flush = nullptr;
    size_t flush_bytes = 0;
    uint64_t next_ticket = 1;
    int64_t launches = 0;
    int64_t h2d_bytes = 0, d2h_bytes = 0;   // bytes moved by cudaMemcpyAsync on the host-buffer path
    bool zero_copy = true;         // BVC_NO_ZERO_COPY=1 forces the full mapq/rpr copy (A/B checks)
    bool wide_hist = false;        // BVC_WIDE_HIST=1: K1 as site_hist_kernel<G> even for small cohorts (A/B checks)
    bool expand8 = false;          // BVC_EXPAND8=1: a packed8 tile is expanded into the (code, qual) arrays and K1 runs in its two-array form (A/B checks)
    int em_min_fits = 0;           // BVC_EM_MIN_FITS: stages below this many fits run entirely in em_long_kernel; 0 = one fit per K3 lane
    int long_exact = 0;            // BVC_LONG_EXACT=1: em_long_kernel with the exact delta on every pass (A/B checks)
    int refill_min = 4;            // BVC_REFILL_MIN: em_fit_kernel batches its refills (C2: em_fit 0.79 -> 0.73 ms at 4..8, 0.78 at 16; C5 neutral)
    int pass_budget = 64, drain_min_pass = 16, drain_lanes = 8;  // em_fit_kernel -> em_long_kernel hand-over (BVC_PASS_BUDGET / BVC_DRAIN_MIN_PASS)
    bool legacy_only = false;      // BVC_LEGACY_KERNEL=1: the single warp-per-site kernel for every site (A/B checks)
    std::string err;
    bool resident = false;
    const bvc_coded_tile* pending_coded = nullptr;   // set by bvc_enqueue_split_coded for the upload of ITS tile only
    bvc::ShardState* shard = nullptr;   // sample-sharded mode (bvc_shard_init / bvc_shard_init_group)
};

namespace bvc {

void set_err(bvc_ctx* ctx, const char* m);     // ctx == NULL: the thread's create-time error
int  check_tile(bvc_ctx* ctx, const bvc_tile* t, int64_t* n_entries);
// H2D copies of a tile into the slot (code, qual, offsets, ref; mapq / rpr zero-copy when pinned) on the slot's stream
// need_arrays: the caller's kernels read the (code, qual) arrays (sample-sharded mode), so a packed8 tile is expanded
int  upload(bvc_ctx* ctx, Slot& s, const bvc_tile* t, int64_t R, bool resident, bool need_arrays = false);
size_t pool_words(const bvc_ctx* ctx);         // capacity of a slot's class pool
// the argument blocks of the kernels over the slot's buffers
void fill_args(const bvc_ctx* ctx, const Slot& s, SiteArgs& a, PipeArgs& p);
void shard_free(bvc_ctx* ctx);                 // called by bvc_destroy
// Device allocations of the library.  BVC_GUARD=1 (debug; compute-sanitizer is not available on every pool): every allocation
// gets a 4 KB band of 0xA5 on both sides and the whole of it is pre-filled with 0xA5; guard_verify() reads all bands of the
// process back and names the first one a kernel wrote into (bvc_wait / bvc_wait_split / bvc_shard_* / bvc_destroy call it).
cudaError_t gmalloc_bytes(void** p, size_t bytes);
template <class T> inline cudaError_t gmalloc(T** p, size_t bytes) { return gmalloc_bytes(reinterpret_cast<void**>(p), bytes); }
cudaError_t gfree(void* p);
int guard_verify(bvc_ctx* ctx);                // BVC_OK, or BVC_ERR_CUDA with the error text set

}  // namespace bvc

#define CK(call)                                                                               \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            char b_[512];                                                                      \
            snprintf(b_, sizeof b_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            bvc::set_err(ctx, b_);                                                             \
            return BVC_ERR_CUDA;                                                               \
        }                                                                                      \
    } while (0)
