// bvc_shard_api.cu -- host side of the SAMPLE-SHARDED mode of include/bvc_gpu.h (bvc_shard_*): one rank per GPU,
// each holding the entries of its sample range for every site of a tile; per-site integer histograms are exchanged
// once per tile (reduce-scatter), the owner of a site slice runs EM / LRT, and the VCF statistics of the
// LRT-positive sites take one all-gather + one all-reduce.  Kernels and the phase list: bvc_shard.cuh.
//
// Exchange back ends:
//   * NCCL (bvc_shard_init): resolved at run time with dlopen("libnccl.so.2") -- the tile path of the library has
//     no NCCL dependency, and a process that already loaded torch's NCCL shares that copy.
//   * in-process (bvc_shard_init_group / bvc_shard_call_group): W contexts of ONE process on ONE device, summed with
//     a plain kernel.  Same phases, no NCCL: the single-GPU parity check of the sharded path.
#include <dlfcn.h>
#include <nccl.h>
#include <string.h>

#include "bvc_ctx.h"

namespace bvc {

struct NcclApi {
    void* h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*ReduceScatter)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string err;
};

static NcclApi* nccl_api()
{
    static NcclApi api;
    static bool tried = false;
    if (tried) return &api;
    tried = true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) { api.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (api.h) break; }
    if (!api.h) { api.err = std::string("dlopen(libnccl.so.2) failed: ") + dlerror(); return &api; }
    bool ok = true;
    auto sym = [&](const char* name) { void* p = dlsym(api.h, name); if (!p) { ok = false; api.err = std::string("NCCL symbol missing: ") + name; } return p; };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.ReduceScatter = reinterpret_cast<decltype(api.ReduceScatter)>(sym("ncclReduceScatter"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    if (!ok) { dlclose(api.h); api.h = nullptr; }
    return &api;
}

constexpr int kShardPhases = BVC_SHARD_N_PHASES;

struct ShardState {
    int rank = 0, world = 1, qlo = 1, qhi = 63, B = 0, sr_cap = 0;
    bool group = false;                 // in-process exchange (no NCCL)
    ncclComm_t comm = nullptr;
    uint32_t* d_dense = nullptr;        // [world * sr_cap][B]
    uint8_t*  d_mask_all = nullptr;     // [world * sr_cap]
    uint32_t* d_pos_list = nullptr;     // [world * sr_cap]
    uint32_t* d_flags = nullptr;        // [0] pos_n, [1] err, [2] max(127 - qual), [3] max(qual) of the tile (shard_qual_range_kernel)
    uint32_t* d_hv = nullptr;           // [hv_cap][kVcfHistWords]
    size_t hv_cap = 0;
    uint32_t* h_flags = nullptr;        // pinned [8]: [0..3] read back from d_flags, [4..5] this rank's qual range of the tile in flight
    bool hist8 = false;                 // the tile in flight came in its one-byte form: with BVC_SHARD_HIST8=1 phase A runs site_hist8_kernel's streaming
    bool fit_window = true;             // shrink the exchange window to the tile's quals (BVC_SHARD_FIXED_WINDOW=1: keep the one of bvc_shard_init)
    cudaEvent_t ev[kShardPhases + 1] = {};
    float phase_ms[kShardPhases] = {};
    int64_t exchange_bytes = 0;         // payload bytes this rank handed to the collectives (cumulative)
    ShardArgs a;                        // the tile in flight
    bool have_tile = false;
    uint32_t n_pos = 0;
};

#define NC(call)                                                                                         \
    do {                                                                                                 \
        ncclResult_t r_ = (call);                                                                        \
        if (r_ != ncclSuccess) {                                                                         \
            char b_[512];                                                                                \
            snprintf(b_, sizeof b_, "%s failed: %s (%s:%d)", #call, nccl_api()->GetErrorString(r_), __FILE__, __LINE__); \
            set_err(ctx, b_);                                                                            \
            return BVC_ERR_CUDA;                                                                         \
        }                                                                                                \
    } while (0)

void shard_free(bvc_ctx* ctx)
{
    ShardState* st = ctx->shard;
    if (!st) return;
    if (st->comm) nccl_api()->CommDestroy(st->comm);
    bvc::gfree(st->d_dense); bvc::gfree(st->d_mask_all); bvc::gfree(st->d_pos_list); bvc::gfree(st->d_flags); bvc::gfree(st->d_hv);
    if (st->h_flags) cudaFreeHost(st->h_flags);
    for (auto& e : st->ev) if (e) cudaEventDestroy(e);
    delete st;
    ctx->shard = nullptr;
}

static int shard_alloc(bvc_ctx* ctx, int rank, int world, int qlo, int qhi, bool group)
{
    if (ctx->shard) { set_err(ctx, "bvc_shard_init: context is already in sharded mode"); return BVC_ERR_STATE; }
    if (world < 1 || rank < 0 || rank >= world) { set_err(ctx, "bvc_shard_init: bad rank / world"); return BVC_ERR_ARG; }
    if (qlo < 1 || qhi < qlo || qhi > 127) { set_err(ctx, "bvc_shard_init: qual window must satisfy 1 <= qual_lo <= qual_hi <= 127 (qual 0 is not served in sharded mode)"); return BVC_ERR_ARG; }
    if (ctx->prm.n_groups > 0) { set_err(ctx, "bvc_shard_init: the --group path is not served in sharded mode"); return BVC_ERR_ARG; }
    if (ctx->legacy_only) { set_err(ctx, "bvc_shard_init: BVC_LEGACY_KERNEL=1 does not serve sharded mode"); return BVC_ERR_ARG; }
    CK(cudaSetDevice(ctx->device));
    ShardState* st = new ShardState();
    ctx->shard = st;
    st->rank = rank; st->world = world; st->qlo = qlo; st->qhi = qhi; st->group = group;
    { const char* e = getenv("BVC_SHARD_FIXED_WINDOW"); st->fit_window = !(e && e[0] == '1'); }
    st->B = kShardHdr + 8 * (qhi - qlo + 1);
    st->sr_cap = (ctx->prm.max_sites + world - 1) / world;
    const size_t n_pad = (size_t)world * st->sr_cap;
    CK(bvc::gmalloc(&st->d_dense, n_pad * st->B * sizeof(uint32_t)));
    CK(bvc::gmalloc(&st->d_mask_all, n_pad + 16));
    CK(bvc::gmalloc(&st->d_pos_list, (n_pad + 16) * sizeof(uint32_t)));
    CK(bvc::gmalloc(&st->d_flags, 4 * sizeof(uint32_t)));
    CK(cudaHostAlloc(&st->h_flags, 8 * sizeof(uint32_t), cudaHostAllocDefault));
    memset(st->h_flags, 0, 8 * sizeof(uint32_t));
    for (auto& e : st->ev) CK(cudaEventCreate(&e));
    // the class pool must hold every (base, qual) class of the owned sites, whatever the local entry capacity
    Slot& s = ctx->slot[0];
    const size_t S = (size_t)ctx->prm.max_sites, R = (size_t)ctx->prm.max_entries;
    const size_t have = (R < S * (size_t)kNBins ? R : S * (size_t)kNBins) + 8 * S;
    const size_t need = (size_t)st->sr_cap * (4 * (size_t)(qhi - qlo + 1) + 8);
    if (need > have) {
        CK(bvc::gfree(s.d_cls)); s.d_cls = nullptr;
        CK(bvc::gmalloc(&s.d_cls, (need + 64) * sizeof(uint32_t)));
    }
    return BVC_OK;
}

// ---- phases (all work on slot 0 and its stream) -------------------------------------------------------------
static int shard_upload(bvc_ctx* ctx, const bvc_tile* tile)
{
    ShardState* st = ctx->shard;
    if (!st) { set_err(ctx, "bvc_shard_*: bvc_shard_init was not called on this context"); return BVC_ERR_STATE; }
    int64_t R = 0;
    int rc = check_tile(ctx, tile, &R);
    if (rc != BVC_OK) return rc;
    CK(cudaSetDevice(ctx->device));
    Slot& s = ctx->slot[0];
    if (s.ticket) { CK(cudaStreamSynchronize(s.stream)); s.ticket = 0; }
    CK(cudaEventRecord(st->ev[0], s.stream));
    if ((rc = upload(ctx, s, tile, R, false, true)) != BVC_OK) return rc;     // the shard kernels read the (code, qual) arrays
    ctx->resident = false;
    SiteArgs sa;
    ShardArgs& a = st->a;
    fill_args(ctx, s, sa, a.p);
    const int S = tile->n_sites;
    a.world = st->world; a.rank = st->rank; a.B = st->B; a.qlo = st->qlo; a.qhi = st->qhi;
    a.Sr = (S + st->world - 1) / st->world;
    a.n_tile_sites = S;
    const int lo = a.rank * a.Sr;
    a.n_own = S - lo < 0 ? 0 : (S - lo < a.Sr ? S - lo : a.Sr);
    a.tile_ref = s.d_ref; a.tile_off = s.d_off;
    a.dense = st->d_dense; a.err = st->d_flags + 1; a.mask_all = st->d_mask_all;
    a.pos_list = st->d_pos_list; a.pos_n = st->d_flags; a.hv = st->d_hv;
    // the owner's pipeline sees its slice only: local site index s <-> tile site lo + s
    a.p.s.n_sites = a.n_own;
    a.p.s.ref_base = s.d_ref + (a.n_own > 0 ? lo : 0);
    // this rank's qual range of the tile (h_flags[4] = max(127 - qual), [5] = max(qual); 0, 0 = no counted entry): from the
    // dictionary when the tile came in its one-byte form (a superset of the quals in use is fine), else one pass on the device
    st->hist8 = R > 0 && tile->packed8 && tile->qual_dict && tile->n_qual_dict > 0 && getenv("BVC_SHARD_HIST8") && !getenv("BVC_SHARD_HIST_ARRAYS");   // opt-in: measured equal to the two-array path
    st->h_flags[4] = st->h_flags[5] = 0;
    if (st->fit_window && R > 0) {
        if (tile->packed8 && tile->qual_dict && tile->n_qual_dict > 0) {
            for (int i = 0; i < tile->n_qual_dict; ++i) {
                const uint32_t q = tile->qual_dict[i] & 127u;
                if (127u - q > st->h_flags[4]) st->h_flags[4] = 127u - q;
                if (q > st->h_flags[5]) st->h_flags[5] = q;
            }
        } else {
            CK(cudaMemsetAsync(st->d_flags + 2, 0, 2 * sizeof(uint32_t), s.stream));
            CK(launch_shard_qual_range(s.d_code, s.d_qual, (size_t)R, st->d_flags + 2, ctx->n_sm, s.stream));
            ctx->launches += 1;
            CK(cudaMemcpyAsync(st->h_flags + 4, st->d_flags + 2, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
            CK(cudaStreamSynchronize(s.stream));
        }
    }
    st->have_tile = true;
    return BVC_OK;
}

// Phase A in three steps so that the ranks can agree on the tile's qual window between them: (1) this rank's qual range,
// (2) [the caller combines the ranges: a 2-word MAX all-reduce, or a host loop for an in-process group] shard_apply_window,
// (3) the histogram kernel over the agreed window.
static int shard_phase_range(bvc_ctx* ctx)
{
    ShardState* st = ctx->shard;
    Slot& s = ctx->slot[0];
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventRecord(st->ev[1], s.stream));
    CK(cudaMemsetAsync(st->d_flags, 0, 4 * sizeof(uint32_t), s.stream));
    return BVC_OK;
}

// the agreed range [127 - lo_neg, hi] (0, 0: no counted entry anywhere) -> this tile's window inside the one of bvc_shard_init
static void shard_apply_window(ShardState* st, uint32_t lo_neg, uint32_t hi)
{
    ShardArgs& a = st->a;
    a.qlo = st->qlo; a.qhi = st->qhi;
    if (st->fit_window) {
        int lo = 127 - (int)lo_neg, h = (int)hi;
        if (lo > h) { lo = st->qlo; h = st->qlo; }                 // no counted entry: the narrowest window
        // a qual outside the window of bvc_shard_init stays outside: the histogram kernel reports it (BVC_ERR_ARG)
        a.qlo = lo < st->qlo ? st->qlo : (lo > st->qhi ? st->qhi : lo);
        a.qhi = h > st->qhi ? st->qhi : (h < a.qlo ? a.qlo : h);
    }
    a.B = kShardHdr + 8 * (a.qhi - a.qlo + 1);
}

static int shard_phase_hist(bvc_ctx* ctx)
{
    ShardState* st = ctx->shard;
    Slot& s = ctx->slot[0];
    CK(cudaSetDevice(ctx->device));
    int n = 0;
    if (st->hist8) CK(launch_shard_hist8(st->a, s.d_pk8, ctx->n_sm, s.stream, &n));
    else CK(launch_shard_hist(st->a, ctx->n_sm, s.stream, &n));
    ctx->launches += n;
    CK(cudaEventRecord(st->ev[2], s.stream));
    return BVC_OK;
}

static int shard_phase_fit(bvc_ctx* ctx)
{
    ShardState* st = ctx->shard;
    Slot& s = ctx->slot[0];
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventRecord(st->ev[3], s.stream));
    CK(cudaMemsetAsync(s.d_ctr, 0, kCtrWords * sizeof(uint32_t), s.stream));
    int n = 0;
    CK(launch_shard_merge_and_fit(st->a, ctx->n_sm, s.stream, &n));
    ctx->launches += n;
    CK(cudaEventRecord(st->ev[4], s.stream));
    return BVC_OK;
}

static int shard_phase_vcf_hist(bvc_ctx* ctx)
{
    ShardState* st = ctx->shard;
    Slot& s = ctx->slot[0];
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventRecord(st->ev[5], s.stream));
    int n = 0;
    CK(launch_shard_poslist(st->a, s.stream, &n));
    ctx->launches += n;
    // the host needs P (the collective's element count) and the error flag: one 8-byte read per tile
    CK(cudaMemcpyAsync(st->h_flags, st->d_flags, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
    CK(cudaStreamSynchronize(s.stream));
    st->n_pos = st->h_flags[0];
    if (st->h_flags[1] & 1u) {
        // every rank must still take part in the remaining collectives with the same counts: carry on, fail at the end
        set_err(ctx, "bvc_shard_call: an entry's qual lies outside the exchange window given to bvc_shard_init");
    }
    if (st->n_pos > st->hv_cap) {
        size_t cap = st->hv_cap ? st->hv_cap : 1024;
        while (cap < st->n_pos) cap *= 2;
        CK(bvc::gfree(st->d_hv)); st->d_hv = nullptr; st->hv_cap = 0;
        CK(bvc::gmalloc(&st->d_hv, cap * kVcfHistWords * sizeof(uint32_t)));
        st->hv_cap = cap;
    }
    st->a.hv = st->d_hv;
    CK(launch_shard_vcf_hist(st->a, ctx->n_sm, st->n_pos, s.stream, &n));
    ctx->launches += n;
    CK(cudaEventRecord(st->ev[6], s.stream));
    return BVC_OK;
}

static int shard_phase_finish(bvc_ctx* ctx, bvc_site_result* out, int32_t* first_site, int32_t* n_out)
{
    ShardState* st = ctx->shard;
    Slot& s = ctx->slot[0];
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventRecord(st->ev[7], s.stream));
    int n = 0;
    CK(launch_shard_vcf_finish(st->a, ctx->n_sm, st->n_pos, s.stream, &n));
    ctx->launches += n;
    CK(cudaEventRecord(st->ev[8], s.stream));
    if (out && st->a.n_own > 0) {
        CK(cudaMemcpyAsync(out, s.d_out, (size_t)st->a.n_own * sizeof(bvc_site_result), cudaMemcpyDeviceToHost, s.stream));
        ctx->d2h_bytes += (int64_t)st->a.n_own * (int64_t)sizeof(bvc_site_result);
    }
    CK(cudaEventRecord(st->ev[9], s.stream));
    CK(cudaStreamSynchronize(s.stream));
    for (int i = 0; i < kShardPhases; ++i) CK(cudaEventElapsedTime(&st->phase_ms[i], st->ev[i], st->ev[i + 1]));
    if (first_site) *first_site = st->a.rank * st->a.Sr;
    if (n_out) *n_out = st->a.n_own;
    if (st->h_flags[1] & 1u) return BVC_ERR_ARG;
    return guard_verify(ctx);
}

// ---- NCCL run: the phases with the three collectives between them --------------------------------------------
static int shard_run_nccl(bvc_ctx* ctx, bvc_site_result* out, int32_t* first_site, int32_t* n_out)
{
    ShardState* st = ctx->shard;
    NcclApi* nc = nccl_api();
    Slot& s = ctx->slot[0];
    const ShardArgs& a = st->a;
    int rc;
    if ((rc = shard_phase_range(ctx)) != BVC_OK) return rc;
    if (st->fit_window) {                                         // exchange 0: two words, MAX
        CK(cudaMemcpyAsync(st->d_flags + 2, st->h_flags + 4, 2 * sizeof(uint32_t), cudaMemcpyHostToDevice, s.stream));
        NC(nc->AllReduce(st->d_flags + 2, st->d_flags + 2, 2, ncclUint32, ncclMax, st->comm, s.stream));
        CK(cudaMemcpyAsync(st->h_flags + 2, st->d_flags + 2, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
        CK(cudaStreamSynchronize(s.stream));
    }
    shard_apply_window(st, st->h_flags[2], st->h_flags[3]);
    if ((rc = shard_phase_hist(ctx)) != BVC_OK) return rc;
    const size_t seg = (size_t)a.Sr * a.B;
    NC(nc->ReduceScatter(a.dense, a.dense + (size_t)a.rank * seg, seg, ncclUint32, ncclSum, st->comm, s.stream));
    st->exchange_bytes += (int64_t)(seg * a.world * sizeof(uint32_t));
    if ((rc = shard_phase_fit(ctx)) != BVC_OK) return rc;
    NC(nc->AllGather(a.mask_all + (size_t)a.rank * a.Sr, a.mask_all, (size_t)a.Sr, ncclUint8, st->comm, s.stream));
    st->exchange_bytes += (int64_t)a.Sr;
    if ((rc = shard_phase_vcf_hist(ctx)) != BVC_OK) return rc;
    if (st->n_pos) {
        NC(nc->AllReduce(st->d_hv, st->d_hv, (size_t)st->n_pos * kVcfHistWords, ncclUint32, ncclSum, st->comm, s.stream));
        st->exchange_bytes += (int64_t)((size_t)st->n_pos * kVcfHistWords * sizeof(uint32_t));
    }
    return shard_phase_finish(ctx, out, first_site, n_out);
}

}  // namespace bvc

using namespace bvc;

extern "C" {

int bvc_shard_unique_id(void* id128)
{
    bvc_ctx* ctx = nullptr;
    if (!id128) return BVC_ERR_ARG;
    NcclApi* nc = nccl_api();
    if (!nc->h) { set_err(nullptr, nc->err.c_str()); return BVC_ERR_CUDA; }
    static_assert(sizeof(ncclUniqueId) == BVC_SHARD_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId id;
    NC(nc->GetUniqueId(&id));
    memcpy(id128, &id, sizeof id);
    return BVC_OK;
}

int bvc_shard_init(bvc_ctx* ctx, int rank, int world, const void* id128, int qual_lo, int qual_hi)
{
    if (!ctx || !id128) return BVC_ERR_ARG;
    NcclApi* nc = nccl_api();
    if (!nc->h) { set_err(ctx, nc->err.c_str()); return BVC_ERR_CUDA; }
    int rc = shard_alloc(ctx, rank, world, qual_lo, qual_hi, false);
    if (rc != BVC_OK) return rc;
    ncclUniqueId id;
    memcpy(&id, id128, sizeof id);
    NC(nc->CommInitRank(&ctx->shard->comm, world, id, rank));
    return BVC_OK;
}

int bvc_shard_init_group(bvc_ctx** ctxs, int world, int qual_lo, int qual_hi)
{
    if (!ctxs || world < 1) return BVC_ERR_ARG;
    for (int r = 0; r < world; ++r) {
        if (!ctxs[r]) return BVC_ERR_ARG;
        if (ctxs[r]->device != ctxs[0]->device || ctxs[r]->prm.max_sites != ctxs[0]->prm.max_sites) {
            set_err(ctxs[r], "bvc_shard_init_group: the contexts of a group must share one device and one max_sites");
            return BVC_ERR_ARG;
        }
        int rc = shard_alloc(ctxs[r], r, world, qual_lo, qual_hi, true);
        if (rc != BVC_OK) return rc;
    }
    return BVC_OK;
}

int bvc_shard_upload(bvc_ctx* ctx, const bvc_tile* tile)
{
    if (!ctx) return BVC_ERR_ARG;
    int rc = shard_upload(ctx, tile);
    if (rc != BVC_OK) return rc;
    CK(cudaStreamSynchronize(ctx->slot[0].stream));
    return BVC_OK;
}

int bvc_shard_run(bvc_ctx* ctx, bvc_site_result* out, int32_t* first_site, int32_t* n_out)
{
    if (!ctx) return BVC_ERR_ARG;
    ShardState* st = ctx->shard;
    if (!st || !st->have_tile) { set_err(ctx, "bvc_shard_run: no uploaded tile (bvc_shard_init + bvc_shard_upload first)"); return BVC_ERR_STATE; }
    if (st->group) { set_err(ctx, "bvc_shard_run: this context belongs to an in-process group (use bvc_shard_call_group)"); return BVC_ERR_STATE; }
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventRecord(st->ev[0], ctx->slot[0].stream));     // resident run: the h2d phase is empty
    return shard_run_nccl(ctx, out, first_site, n_out);
}

int bvc_shard_call(bvc_ctx* ctx, const bvc_tile* tile, bvc_site_result* out, int32_t* first_site, int32_t* n_out)
{
    if (!ctx) return BVC_ERR_ARG;
    if (ctx->shard && ctx->shard->group) { set_err(ctx, "bvc_shard_call: this context belongs to an in-process group (use bvc_shard_call_group)"); return BVC_ERR_STATE; }
    int rc = shard_upload(ctx, tile);
    if (rc != BVC_OK) return rc;
    return shard_run_nccl(ctx, out, first_site, n_out);
}

int bvc_shard_call_group(bvc_ctx** ctxs, int world, const bvc_tile* const* tiles, bvc_site_result* const* outs,
                         int32_t* first_site, int32_t* n_out)
{
    if (!ctxs || !tiles || !outs || world < 1) return BVC_ERR_ARG;
    for (int r = 0; r < world; ++r) {
        if (!ctxs[r] || !tiles[r]) return BVC_ERR_ARG;
        bvc_ctx* ctx = ctxs[r];
        if (!ctx->shard || !ctx->shard->group || ctx->shard->world != world || ctx->shard->rank != r) {
            set_err(ctx, "bvc_shard_call_group: contexts must come from bvc_shard_init_group, in rank order");
            return BVC_ERR_STATE;
        }
        if (tiles[r]->n_sites != tiles[0]->n_sites) { set_err(ctx, "bvc_shard_call_group: every rank's tile must cover the same sites"); return BVC_ERR_ARG; }
    }
    int rc, worst = BVC_OK;
    auto sync_all = [&]() { for (int r = 0; r < world; ++r) cudaStreamSynchronize(ctxs[r]->slot[0].stream); };
    for (int r = 0; r < world; ++r) {
        if ((rc = shard_upload(ctxs[r], tiles[r])) != BVC_OK) return rc;
        if ((rc = shard_phase_range(ctxs[r])) != BVC_OK) return rc;
    }
    {   // exchange 0: the tile's qual range over all ranks
        uint32_t lo_neg = 0, hi = 0;
        for (int r = 0; r < world; ++r) {
            ShardState* q = ctxs[r]->shard;
            if (!q->fit_window) continue;
            if (q->h_flags[4] > lo_neg) lo_neg = q->h_flags[4];
            if (q->h_flags[5] > hi) hi = q->h_flags[5];
        }
        for (int r = 0; r < world; ++r) {
            shard_apply_window(ctxs[r]->shard, lo_neg, hi);
            if ((rc = shard_phase_hist(ctxs[r])) != BVC_OK) return rc;
        }
    }
    sync_all();
    {   // exchange 1: rank r's segment <- sum over ranks
        bvc_ctx* ctx = ctxs[0];
        for (int r = 0; r < world; ++r) {
            const ShardArgs& a = ctxs[r]->shard->a;
            const size_t seg = (size_t)a.Sr * a.B;
            for (int k = 0; k < world; ++k) if (k != r)
                CK(launch_sum_u32(a.dense + (size_t)r * seg, ctxs[k]->shard->a.dense + (size_t)r * seg, seg, ctxs[r]->n_sm, ctxs[r]->slot[0].stream));
            ctxs[r]->shard->exchange_bytes += (int64_t)(seg * world * sizeof(uint32_t));
        }
    }
    sync_all();
    for (int r = 0; r < world; ++r) if ((rc = shard_phase_fit(ctxs[r])) != BVC_OK) return rc;
    sync_all();
    {   // exchange 2: every rank gets every owner's mask bytes
        bvc_ctx* ctx = ctxs[0];
        for (int r = 0; r < world; ++r) {
            const ShardArgs& a = ctxs[r]->shard->a;
            for (int k = 0; k < world; ++k) if (k != r)
                CK(cudaMemcpyAsync(ctxs[k]->shard->a.mask_all + (size_t)r * a.Sr, a.mask_all + (size_t)r * a.Sr, (size_t)a.Sr,
                                   cudaMemcpyDeviceToDevice, ctxs[r]->slot[0].stream));
            ctxs[r]->shard->exchange_bytes += (int64_t)a.Sr;
        }
    }
    sync_all();
    for (int r = 0; r < world; ++r) if ((rc = shard_phase_vcf_hist(ctxs[r])) != BVC_OK) return rc;
    sync_all();
    {   // exchange 3: all-reduce of the VCF histograms through rank 0
        bvc_ctx* ctx = ctxs[0];
        const size_t n = (size_t)ctxs[0]->shard->n_pos * kVcfHistWords;
        for (int k = 1; k < world; ++k) {
            if (ctxs[k]->shard->n_pos != ctxs[0]->shard->n_pos) { set_err(ctx, "bvc_shard_call_group: ranks disagree on the LRT-positive list"); return BVC_ERR_STATE; }
            CK(launch_sum_u32(ctxs[0]->shard->d_hv, ctxs[k]->shard->d_hv, n, ctxs[0]->n_sm, ctxs[0]->slot[0].stream));
        }
        CK(cudaStreamSynchronize(ctxs[0]->slot[0].stream));
        for (int k = 1; k < world && n; ++k)
            CK(cudaMemcpyAsync(ctxs[k]->shard->d_hv, ctxs[0]->shard->d_hv, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctxs[k]->slot[0].stream));
        for (int r = 0; r < world; ++r) ctxs[r]->shard->exchange_bytes += (int64_t)(n * sizeof(uint32_t));
    }
    sync_all();
    for (int r = 0; r < world; ++r) {
        rc = shard_phase_finish(ctxs[r], outs[r], first_site ? first_site + r : nullptr, n_out ? n_out + r : nullptr);
        if (rc != BVC_OK) worst = rc;
    }
    return worst;
}

int bvc_shard_stats(const bvc_ctx* ctx, float* phase_ms, int64_t* exchange_bytes, int32_t* n_positive)
{
    if (!ctx || !ctx->shard) return BVC_ERR_STATE;
    if (phase_ms) for (int i = 0; i < kShardPhases; ++i) phase_ms[i] = ctx->shard->phase_ms[i];
    if (exchange_bytes) *exchange_bytes = ctx->shard->exchange_bytes;
    if (n_positive) *n_positive = (int32_t)ctx->shard->n_pos;
    return BVC_OK;
}

const char* bvc_shard_phase_name(int i)
{
    static const char* names[kShardPhases] = {"h2d", "shard_hist", "reduce_scatter", "merge_em_lrt", "all_gather",
                                              "vcf_hist", "all_reduce", "vcf_finish", "d2h"};
    return (i >= 0 && i < kShardPhases) ? names[i] : "?";
}

}  // extern "C"
