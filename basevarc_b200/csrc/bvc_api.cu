// bvc_api.cu -- the C-ABI of include/bvc_gpu.h: context, double-buffered tile pipeline, resident
// (device-timed) path.  No CPU fallback: every entry point needs a CUDA device.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "bvc_ctx.h"
#include "bvc_codes.h"

using bvc::Slot;
using bvc::kCtrWords;
using bvc::set_err;
using bvc::check_tile;
using bvc::upload;

namespace {
thread_local std::string g_create_error;
}

namespace bvc {
void set_err(bvc_ctx* ctx, const char* m)
{
    if (ctx) ctx->err = m; else g_create_error = m;
}

// ---- guard bands (BVC_GUARD=1) ----------------------------------------------------------------------------------------
namespace {
constexpr size_t kGuardBytes = 4096;
struct GuardRec { unsigned char* base; size_t bytes; int serial; };
std::mutex g_guard_mu;
std::unordered_map<void*, GuardRec> g_guard;
int g_guard_serial = 0;
bool guard_on()
{
    static const bool on = [] { const char* e = getenv("BVC_GUARD"); return e && e[0] == '1'; }();
    return on;
}
}  // namespace

cudaError_t gmalloc_bytes(void** p, size_t bytes)
{
    if (!guard_on()) return cudaMalloc(p, bytes);
    unsigned char* b = nullptr;
    cudaError_t e = cudaMalloc(&b, bytes + 2 * kGuardBytes);
    if (e != cudaSuccess) return e;
    e = cudaMemset(b, 0xA5, bytes + 2 * kGuardBytes);
    if (e != cudaSuccess) { cudaFree(b); return e; }
    *p = b + kGuardBytes;
    std::lock_guard<std::mutex> lk(g_guard_mu);
    g_guard[*p] = GuardRec{b, bytes, g_guard_serial++};
    return cudaSuccess;
}

cudaError_t gfree(void* p)
{
    if (!guard_on() || !p) return cudaFree(p);
    unsigned char* b = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_guard_mu);
        auto it = g_guard.find(p);
        if (it == g_guard.end()) return cudaFree(p);
        b = it->second.base;
        g_guard.erase(it);
    }
    return cudaFree(b);
}

int guard_verify(bvc_ctx* ctx)
{
    if (!guard_on()) return BVC_OK;
    std::lock_guard<std::mutex> lk(g_guard_mu);
    std::vector<unsigned char> h(2 * kGuardBytes);
    for (auto& kv : g_guard) {
        const GuardRec& r = kv.second;
        if (cudaMemcpy(h.data(), r.base, kGuardBytes, cudaMemcpyDeviceToHost) != cudaSuccess ||
            cudaMemcpy(h.data() + kGuardBytes, r.base + kGuardBytes + r.bytes, kGuardBytes, cudaMemcpyDeviceToHost) != cudaSuccess) {
            cudaGetLastError();
            set_err(ctx, "BVC_GUARD: cannot read a guard band back");
            return BVC_ERR_CUDA;
        }
        for (size_t i = 0; i < 2 * kGuardBytes; ++i)
            if (h[i] != 0xA5) {
                char b[200];
                snprintf(b, sizeof b, "BVC_GUARD: allocation #%d (%zu bytes) was written out of bounds: %s band, byte %zu", r.serial, r.bytes,
                         i < kGuardBytes ? "front" : "back", i < kGuardBytes ? kGuardBytes - i : i - kGuardBytes);
                set_err(ctx, b);
                fprintf(stderr, "%s\n", b);
                return BVC_ERR_CUDA;
            }
    }
    return BVC_OK;
}
}  // namespace bvc

namespace {

int alloc_slot(bvc_ctx* ctx, Slot& s)
{
    const size_t S = (size_t)ctx->prm.max_sites, R = (size_t)ctx->prm.max_entries;
    CK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
    CK(bvc::gmalloc(&s.d_ref, S + 16));
    CK(bvc::gmalloc(&s.d_off, (S + 1) * sizeof(uint32_t)));
    CK(bvc::gmalloc(&s.d_code, R + 16));
    CK(bvc::gmalloc(&s.d_qual, R + 16));
    CK(bvc::gmalloc(&s.d_mapq, R + 16));
    CK(bvc::gmalloc(&s.d_rpr, R + 16));
    CK(bvc::gmalloc(&s.d_pk8, R + 16));
    if (ctx->prm.n_groups > 0) {
        CK(bvc::gmalloc(&s.d_sample, (R + 4) * sizeof(uint32_t)));
        CK(bvc::gmalloc(&s.d_gdepth, (S + 1) * (size_t)ctx->prm.n_groups * 4 * sizeof(int32_t)));
        CK(bvc::gmalloc(&s.d_gaf, (S + 1) * (size_t)ctx->prm.n_groups * 4 * sizeof(double)));
    }
    CK(bvc::gmalloc(&s.d_out, (S + 1) * sizeof(bvc_site_result)));
    CK(bvc::gmalloc(&s.d_counter, sizeof(unsigned int)));
    const size_t pool = bvc::pool_words(ctx);
    CK(bvc::gmalloc(&s.d_sum, (S + 1) * sizeof(bvc::SiteSum)));
    CK(bvc::gmalloc(&s.d_cls, (pool + 64) * sizeof(uint32_t)));
    CK(bvc::gmalloc(&s.d_lad, (S + 1) * sizeof(bvc::Ladder)));
    CK(bvc::gmalloc(&s.d_fit, (S + 1) * 4 * sizeof(bvc::FitRes)));
    CK(bvc::gmalloc(&s.d_state, S + 16));
    CK(bvc::gmalloc(&s.d_qa, (4 * S + 16) * sizeof(uint32_t)));
    CK(bvc::gmalloc(&s.d_qb, (4 * S + 16) * sizeof(uint32_t)));
    CK(bvc::gmalloc(&s.d_qs, (4 * S + 16) * sizeof(uint32_t)));
    CK(bvc::gmalloc(&s.d_ql, (4 * S + 16) * sizeof(uint32_t)));
    CK(bvc::gmalloc(&s.d_vcf_list, (S + 16) * sizeof(uint32_t)));
    CK(bvc::gmalloc(&s.d_legacy_list, (S + 16) * sizeof(uint32_t)));
    CK(bvc::gmalloc(&s.d_ctr, kCtrWords * sizeof(uint32_t)));
    CK(cudaHostAlloc(&s.h_err, sizeof(uint32_t), cudaHostAllocDefault));
    *s.h_err = 0;
    return BVC_OK;
}

void free_slot(Slot& s)
{
    bvc::gfree(s.d_ref); bvc::gfree(s.d_off); bvc::gfree(s.d_code); bvc::gfree(s.d_qual); bvc::gfree(s.d_mapq);
    bvc::gfree(s.d_rpr); bvc::gfree(s.d_pk8); bvc::gfree(s.d_sample); bvc::gfree(s.d_core); bvc::gfree(s.d_vcfc); bvc::gfree(s.d_vcfc_n);
    if (s.h_vcfc_n) cudaFreeHost(s.h_vcfc_n); if (s.h_err) cudaFreeHost(s.h_err); bvc::gfree(s.d_gdepth); bvc::gfree(s.d_gaf); bvc::gfree(s.d_out); bvc::gfree(s.d_counter);
    bvc::gfree(s.d_sum); bvc::gfree(s.d_cls); bvc::gfree(s.d_lad); bvc::gfree(s.d_fit); bvc::gfree(s.d_state);
    bvc::gfree(s.d_qa); bvc::gfree(s.d_qb); bvc::gfree(s.d_qs); bvc::gfree(s.d_ql); bvc::gfree(s.d_vcf_list); bvc::gfree(s.d_legacy_list); bvc::gfree(s.d_ctr);
    if (s.done) cudaEventDestroy(s.done);
    if (s.stream) cudaStreamDestroy(s.stream);
    s = Slot();
}

}  // namespace

// class pool: one word per (base, qual) class + up to 6 alignment / padding words per site (pad4 layout) when K1 allocates
// with its cursor; site_hist8_kernel's fixed layout needs kMaxClsStride words per site
size_t bvc::pool_words(const bvc_ctx* ctx)
{
    const size_t S = (size_t)ctx->prm.max_sites, R = (size_t)ctx->prm.max_entries;
    const size_t cursor = (R < S * (size_t)bvc::kNBins ? R : S * (size_t)bvc::kNBins) + 8 * S;
    const size_t fixed = S * (size_t)bvc::kMaxClsStride;
    return cursor > fixed ? cursor : fixed;
}

int bvc::check_tile(bvc_ctx* ctx, const bvc_tile* t, int64_t* n_entries)
{
    if (!t || t->n_sites < 0 || !t->entry_off || (t->n_sites > 0 && !t->ref_base)) { set_err(ctx, "bad tile: NULL array or n_sites < 0"); return BVC_ERR_ARG; }
    if (t->entry_off[0] != 0) { set_err(ctx, "bad tile: entry_off[0] != 0"); return BVC_ERR_ARG; }
    const int64_t R = t->entry_off[t->n_sites];
    const bool coded = ctx->pending_coded != nullptr;
    if (R > 0 && (!t->mapq || !t->rpr || (!coded && !t->packed8 && (!t->code || !t->qual)))) { set_err(ctx, "bad tile: NULL entry array"); return BVC_ERR_ARG; }
    if (R > 0 && (t->packed8 || coded) && (!t->qual_dict || t->n_qual_dict < 1 || t->n_qual_dict > BVC_PACK8_MAX_QUALS)) { set_err(ctx, "bad tile: packed8 needs qual_dict with 1..30 entries"); return BVC_ERR_ARG; }
    if (ctx->prm.n_groups > 0 && R > 0 && !t->sample) { set_err(ctx, "bad tile: groups need the sample array"); return BVC_ERR_ARG; }
    if (t->n_sites > ctx->prm.max_sites || R > ctx->prm.max_entries) { set_err(ctx, "tile exceeds the capacity given at bvc_create"); return BVC_ERR_CAPACITY; }
    *n_entries = R;
    return BVC_OK;
}

namespace {

// Device-visible alias of a pinned, 4-byte-aligned host array, or nullptr (pageable memory / BVC_NO_ZERO_COPY=1).
const uint8_t* zero_copy_ptr(bvc_ctx* ctx, const uint8_t* host)
{
    if (!ctx->zero_copy || (reinterpret_cast<uintptr_t>(host) & 3u)) return nullptr;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (at.type != cudaMemoryTypeHost || !at.devicePointer) return nullptr;
    return static_cast<const uint8_t*>(at.devicePointer);
}

}  // namespace

int bvc::upload(bvc_ctx* ctx, Slot& s, const bvc_tile* t, int64_t R, bool resident, bool need_arrays)
{
    const size_t S = (size_t)t->n_sites;
    s.n_sites = t->n_sites; s.n_entries = R;
    s.packed = false; s.n_dict = 0;
    if (S) CK(cudaMemcpyAsync(s.d_ref, t->ref_base, S, cudaMemcpyHostToDevice, s.stream));
    CK(cudaMemcpyAsync(s.d_off, t->entry_off, (S + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, s.stream));
    ctx->h2d_bytes += (int64_t)(S + (S + 1) * sizeof(uint32_t));
    if (R) {
        const bvc_coded_tile* cd = ctx->pending_coded;
        if (t->packed8 || cd) {
            // compact wire form: R bytes over PCIe instead of 2R, and the form the tile STAYS in: K1 (site_hist8_kernel) and K5
            // read the byte directly.  Only the paths that still read the two arrays get them expanded: --group, the
            // sample-sharded kernels, the IEEE-faithful legacy kernel (a qual-0 read anywhere in the tile, BVC_LEGACY_KERNEL=1),
            // cohorts of >= 65,536 samples (K1's packed depths are 16-bit) and the A/B switch BVC_EXPAND8=1.
            bvc::QualDict d;
            bool q0 = false;
            for (int i = 0; i < 32; ++i) { d.q[i] = i < t->n_qual_dict ? t->qual_dict[i] : (uint8_t)0; q0 = q0 || (i < t->n_qual_dict && d.q[i] == 0); }
            for (int i = 0; i < t->n_qual_dict; ++i) if (d.q[i] > 127) { set_err(ctx, "bad tile: qual_dict entry > 127"); return BVC_ERR_ARG; }
            if (cd) {
                // coded wire form: the bit stream rides in d_code, the escapes and their per-chunk offsets in d_qual (both unused
                // by a tile that stays in its one-byte form; expand8 below runs AFTER the decode if the arrays are needed)
                const size_t wbytes = bvc_coded_words(R, cd->bits) * sizeof(uint32_t), nch = (size_t)((R + bvc::kCodedChunk - 1) / bvc::kCodedChunk) + 1;
                const size_t esc_room = ((size_t)cd->n_escape + 15) & ~(size_t)15;
                if (wbytes > (size_t)R + 16 || esc_room + nch * sizeof(uint32_t) > (size_t)R + 16) { set_err(ctx, "bad coded tile: more escapes than n/8 (send packed8)"); return BVC_ERR_ARG; }
                CK(cudaMemcpyAsync(s.d_code, cd->words, wbytes, cudaMemcpyHostToDevice, s.stream));
                if (cd->n_escape) CK(cudaMemcpyAsync(s.d_qual, cd->escapes, (size_t)cd->n_escape, cudaMemcpyHostToDevice, s.stream));
                CK(cudaMemcpyAsync(s.d_qual + esc_room, cd->chunk_escapes, nch * sizeof(uint32_t), cudaMemcpyHostToDevice, s.stream));
                ctx->h2d_bytes += (int64_t)(wbytes + (size_t)cd->n_escape + nch * sizeof(uint32_t));
                CK(bvc::launch_decode_codes(cd->bits, reinterpret_cast<const uint32_t*>(s.d_code), s.d_qual, reinterpret_cast<const uint32_t*>(s.d_qual + esc_room),
                                            cd->code_byte, s.d_off, s.d_ref, t->n_sites, s.d_pk8, R, ctx->n_sm, s.stream));
                ctx->launches += 1;
            } else {
                CK(cudaMemcpyAsync(s.d_pk8, t->packed8, (size_t)R, cudaMemcpyHostToDevice, s.stream));
                ctx->h2d_bytes += R;
            }
            s.dict = d; s.n_dict = t->n_qual_dict;
            if (need_arrays || q0 || ctx->expand8 || ctx->legacy_only || ctx->wide_hist || ctx->prm.n_groups > 0 || ctx->prm.n_samples >= 65536) {
                CK(bvc::launch_expand8(s.d_pk8, s.d_code, s.d_qual, R, d, ctx->n_sm, s.stream));
                ctx->launches += 1;
            } else s.packed = true;
        } else {
            CK(cudaMemcpyAsync(s.d_code, t->code, (size_t)R, cudaMemcpyHostToDevice, s.stream));
            CK(cudaMemcpyAsync(s.d_qual, t->qual, (size_t)R, cudaMemcpyHostToDevice, s.stream));
            ctx->h2d_bytes += 2 * R;
        }
        // mapq / rpr are read only at LRT-positive sites (K5: ~0.3 % of a low-pass window).  When the caller's
        // arrays are pinned (bvc_host_alloc / cudaHostRegister) the kernels read them in place over PCIe
        // instead of paying a full H2D copy; pageable arrays are copied as before.
        s.k_mapq = resident ? nullptr : zero_copy_ptr(ctx, t->mapq);
        s.k_rpr = resident ? nullptr : zero_copy_ptr(ctx, t->rpr);
        if (!s.k_mapq) { CK(cudaMemcpyAsync(s.d_mapq, t->mapq, (size_t)R, cudaMemcpyHostToDevice, s.stream)); s.k_mapq = s.d_mapq; ctx->h2d_bytes += R; }
        if (!s.k_rpr) { CK(cudaMemcpyAsync(s.d_rpr, t->rpr, (size_t)R, cudaMemcpyHostToDevice, s.stream)); s.k_rpr = s.d_rpr; ctx->h2d_bytes += R; }
        if (s.d_sample && t->sample) {
            CK(cudaMemcpyAsync(s.d_sample, t->sample, (size_t)R * sizeof(uint32_t), cudaMemcpyHostToDevice, s.stream));
            ctx->h2d_bytes += (int64_t)R * 4;
        }
    }
    return BVC_OK;
}

void bvc::fill_args(const bvc_ctx* ctx, const Slot& s, SiteArgs& a, PipeArgs& p)
{
    a.ref_base = s.d_ref; a.entry_off = s.d_off; a.code = s.d_code; a.qual = s.d_qual; a.mapq = s.k_mapq ? s.k_mapq : s.d_mapq; a.rpr = s.k_rpr ? s.k_rpr : s.d_rpr;
    a.pk8 = s.packed ? s.d_pk8 : nullptr; a.dict = s.dict; a.n_dict = s.n_dict; a.err = s.d_ctr + bvc::kCtrErr;
    a.out = s.d_out; a.counter = s.d_counter; a.lut = ctx->d_lut; a.lgam = ctx->d_lgam;
    a.n_sites = s.n_sites; a.max_iter = ctx->prm.em_max_iter;
    a.min_af = ctx->prm.min_af; a.lrt_thr = ctx->prm.lrt_threshold; a.eps = ctx->prm.em_epsilon;
    a.site_list = nullptr; a.list_n = nullptr;
    p.s = a;
    p.mean_entries = (int)(s.n_entries / (s.n_sites > 0 ? s.n_sites : 1));
    p.pass_budget = ctx->pass_budget; p.drain_min_pass = ctx->drain_min_pass; p.drain_lanes = ctx->drain_lanes; p.em_min_fits = ctx->em_min_fits; p.refill_min = ctx->refill_min; p.long_exact = ctx->long_exact;
    p.packed_hist = (ctx->prm.n_samples < 65536 && !ctx->wide_hist) ? 1 : 0;
    p.cls_stride = 4 * s.n_dict + 4;
    p.sum = s.d_sum; p.cls = s.d_cls; p.cls_cursor = s.d_ctr + 0;
    p.lad = s.d_lad; p.fit = s.d_fit; p.state = s.d_state;
    p.q_in = s.d_qb; p.q_in_n = s.d_ctr + 2;          // swapped before the first EM stage
    p.q_out = s.d_qa; p.q_out_n = s.d_ctr + 1;
    p.q_in_head = s.d_ctr + 17;                        // words 17..19 and the bins behind them are zeroed per EM stage by ONE memset
    p.vcf_list = s.d_vcf_list; p.vcf_n = s.d_ctr + 4;
    p.legacy_list = s.d_legacy_list; p.legacy_n = s.d_ctr + 5;
    p.q_sorted = s.d_qs; p.q_long = s.d_ql; p.q_long_n = s.d_ctr + 18; p.q_long_head = s.d_ctr + 19;
    p.deep_list = s.d_ql; p.deep_n = s.d_ctr + 16 + bvc::kNBins + 8;
    p.work = reinterpret_cast<unsigned long long*>(s.d_ctr + 8);
    p.bins = s.d_ctr + 20;
}

namespace {

int launch(bvc_ctx* ctx, Slot& s, cudaEvent_t* ev = nullptr, int* kinds = nullptr, int* n_out = nullptr)
{
    if (s.n_sites == 0) return BVC_OK;
    CK(cudaMemsetAsync(s.d_counter, 0, sizeof(unsigned int), s.stream));
    bvc::SiteArgs a;
    bvc::PipeArgs p;
    bvc::fill_args(ctx, s, a, p);
    if (ctx->legacy_only) {
        if (ev) { CK(cudaEventRecord(ev[0], s.stream)); kinds[0] = bvc::kKLegacy; }
        CK(bvc::launch_site_call(a, ctx->n_sm, s.stream));
        if (ev) CK(cudaEventRecord(ev[1], s.stream));
        if (n_out) *n_out = 1;
        ctx->launches += 1;
        return BVC_OK;
    }
    CK(cudaMemsetAsync(s.d_ctr, 0, kCtrWords * sizeof(uint32_t), s.stream));
    int nl = 0;
    CK(bvc::launch_pipeline(p, ctx->n_sm, s.stream, &nl, ev, kinds));
    if (n_out) *n_out = nl;
    ctx->launches += nl;
    if (ctx->prm.n_groups > 0) {
        CK(bvc::launch_groups(a, s.d_sample, ctx->d_sample_group, ctx->prm.n_groups, s.d_gdepth, s.d_gaf, ctx->n_sm, s.stream));
        ctx->launches += 2;
    }
    // what K1 found of the input the ABI rejects (qual > 127, non-indel base > 3, dictionary index out of range)
    CK(cudaMemcpyAsync(s.h_err, s.d_ctr + bvc::kCtrErr, sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
    return BVC_OK;
}

}  // namespace

extern "C" {

int bvc_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { set_err(nullptr, "cudaGetDeviceCount failed: no usable CUDA driver/device"); cudaGetLastError(); return 0; }
    return n;
}

void bvc_default_params(bvc_params* p, int32_t n_samples, double maf)
{
    memset(p, 0, sizeof(*p));
    p->abi_version = BVC_ABI_VERSION;
    p->n_samples = n_samples;
    double min_af = 100.0 / n_samples;             // src/BaseVarC.cpp:541-543
    if (min_af > 0.001) min_af = 0.001;
    if (maf < min_af) min_af = maf;
    p->min_af = min_af;
    p->lrt_threshold = 24.0;                       // src/BaseType.h:9
    p->em_epsilon = 0.001;                         // src/BaseType.cpp:45
    p->em_max_iter = 100;                          // src/BaseType.cpp:46
    p->max_sites = 1 << 16;
    p->max_entries = (int64_t)p->max_sites * (n_samples > 0 ? n_samples : 1);
    if (p->max_entries > ((int64_t)1 << 31)) p->max_entries = (int64_t)1 << 31;
}

int bvc_create(bvc_ctx** out, int device, const bvc_params* prm)
{
    bvc_ctx* ctx = nullptr;
    if (!out || !prm) { set_err(nullptr, "bvc_create: NULL argument"); return BVC_ERR_ARG; }
    *out = nullptr;
    if (prm->abi_version != BVC_ABI_VERSION) { set_err(nullptr, "bvc_create: abi_version mismatch"); return BVC_ERR_ARG; }
    if (prm->n_samples <= 0 || !(prm->min_af > 0) || prm->em_max_iter < 0 || prm->max_sites <= 0 || prm->max_entries < 0 ||
        prm->max_entries >= ((int64_t)1 << 32) || prm->n_groups < 0 || (prm->n_groups > 0 && !prm->sample_group)) {
        set_err(nullptr, "bvc_create: bad parameter (n_samples, min_af > 0, capacities, groups)");
        return BVC_ERR_ARG;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        set_err(nullptr, "bvc_create: no CUDA device (this library has no CPU path)");
        cudaGetLastError();
        return BVC_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) { set_err(nullptr, "bvc_create: device index out of range"); return BVC_ERR_ARG; }
    ctx = new bvc_ctx();
    ctx->device = device;
    ctx->prm = *prm;
    if (prm->n_groups > 0) {
        ctx->sample_group.assign(prm->sample_group, prm->sample_group + prm->n_samples);
        ctx->prm.sample_group = ctx->sample_group.data();
    } else ctx->prm.sample_group = nullptr;

    auto fail = [&](int rc) { g_create_error = ctx->err; bvc_destroy(ctx); return rc; };
    cudaDeviceProp prop;
    if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
        ctx->err = "bvc_create: cudaSetDevice/cudaGetDeviceProperties failed"; return fail(BVC_ERR_CUDA);
    }
    if (prop.major < 10) { ctx->err = "bvc_create: device is not sm_100 class (kernels are built for sm_100a only)"; return fail(BVC_ERR_CUDA); }
    ctx->n_sm = prop.multiProcessorCount;
    { const char* e = getenv("BVC_LEGACY_KERNEL"); ctx->legacy_only = (e && e[0] == '1'); }
    { const char* e = getenv("BVC_NO_ZERO_COPY"); ctx->zero_copy = !(e && e[0] == '1'); }
    { const char* e = getenv("BVC_WIDE_HIST"); ctx->wide_hist = (e && e[0] == '1'); }
    { const char* e = getenv("BVC_EXPAND8"); ctx->expand8 = (e && e[0] == '1'); }                        // A/B: K1 in its one-site-per-warp form
    { const char* e = getenv("BVC_PASS_BUDGET"); if (e && atoi(e) > 0) ctx->pass_budget = atoi(e); }       // tuning knobs (A/B runs)
    { const char* e = getenv("BVC_DRAIN_MIN_PASS"); if (e && atoi(e) > 0) ctx->drain_min_pass = atoi(e); }
    { const char* e = getenv("BVC_DRAIN_LANES"); if (e && atoi(e) > 0) ctx->drain_lanes = atoi(e); }
    { const char* e = getenv("BVC_LONG_EXACT"); ctx->long_exact = (e && e[0] == '1'); }
    { const char* e = getenv("BVC_REFILL_MIN"); if (e && atoi(e) > 0) ctx->refill_min = atoi(e); }
    { const char* e = getenv("BVC_EM_MIN_FITS"); if (e && atoi(e) >= 0) ctx->em_min_fits = atoi(e); }
    if (!(prm->em_epsilon <= 0.0019)) { ctx->err = "bvc_create: em_epsilon must be <= 0.0019 (see log1p_tiny in bvc_pipeline.cuh)"; return fail(BVC_ERR_ARG); }

    // host-built tables (glibc): per-qual likelihoods (src/BaseType.cpp:13-15) and lgamma(k)
    std::vector<double> lut(5 * bvc::kNQ), lgam((size_t)prm->n_samples + 3);
    for (int q = 0; q < bvc::kNQ; ++q) {
        const double e = exp(-0.23025850929940458 * q);      // MLN10TO10, src/BaseType.h:10
        lut[q] = e / 3.0; lut[bvc::kNQ + q] = 1.0 - e;
        lut[2 * bvc::kNQ + q] = log(e / 3.0); lut[3 * bvc::kNQ + q] = log(1.0 - e);
        lut[4 * bvc::kNQ + q] = (1.0 - e) - e / 3.0;
    }
    for (size_t k = 0; k < lgam.size(); ++k) lgam[k] = k == 0 ? 0.0 : lgamma((double)k);
    auto ck = [&](cudaError_t e, const char* what) {
        if (e == cudaSuccess) return true;
        ctx->err = std::string(what) + ": " + cudaGetErrorString(e); return false;
    };
    if (!ck(bvc::gmalloc(&ctx->d_lut, lut.size() * sizeof(double)), "cudaMalloc lut") ||
        !ck(bvc::gmalloc(&ctx->d_lgam, lgam.size() * sizeof(double)), "cudaMalloc lgam") ||
        !ck(cudaMemcpy(ctx->d_lut, lut.data(), lut.size() * sizeof(double), cudaMemcpyHostToDevice), "copy lut") ||
        !ck(cudaMemcpy(ctx->d_lgam, lgam.data(), lgam.size() * sizeof(double), cudaMemcpyHostToDevice), "copy lgam"))
        return fail(BVC_ERR_CUDA);
    if (prm->n_groups > 0) {
        for (int i = 0; i < prm->n_samples; ++i)
            if (ctx->sample_group[(size_t)i] >= prm->n_groups) { ctx->err = "bvc_create: sample_group id out of range"; return fail(BVC_ERR_ARG); }
        if (ctx->legacy_only) { ctx->err = "bvc_create: BVC_LEGACY_KERNEL=1 does not serve the --group path"; return fail(BVC_ERR_ARG); }
        if (!ck(bvc::gmalloc(&ctx->d_sample_group, (size_t)prm->n_samples * sizeof(int32_t)), "cudaMalloc sample_group") ||
            !ck(cudaMemcpy(ctx->d_sample_group, ctx->sample_group.data(), (size_t)prm->n_samples * sizeof(int32_t), cudaMemcpyHostToDevice), "copy sample_group"))
            return fail(BVC_ERR_CUDA);
    }
    for (int i = 0; i < 2; ++i) { int rc = alloc_slot(ctx, ctx->slot[i]); if (rc != BVC_OK) return fail(rc); }
    { const char* e = getenv("BVC_GUARD_SELFTEST");             // positive control of BVC_GUARD=1: one byte behind slot 0's site counter
      if (e && e[0] == '1') cudaMemset(reinterpret_cast<unsigned char*>(ctx->slot[0].d_counter) + sizeof(unsigned int), 0, 1); }
    *out = ctx;
    return BVC_OK;
}

void bvc_destroy(bvc_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    for (int i = 0; i < 2; ++i) if (ctx->slot[i].stream) cudaStreamSynchronize(ctx->slot[i].stream);
    if (bvc::guard_verify(nullptr) != BVC_OK) abort();          // BVC_GUARD=1 only: an out-of-bounds write nobody waited for
    for (int i = 0; i < 2; ++i) free_slot(ctx->slot[i]);
    bvc::shard_free(ctx);
    bvc::gfree(ctx->d_lut); bvc::gfree(ctx->d_lgam); bvc::gfree(ctx->d_flush); bvc::gfree(ctx->d_sample_group);
    delete ctx;
}

const char* bvc_last_error(const bvc_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

// Completes the slot's in-flight ticket before the slot is re-used or waited for: a plain ticket only needs the stream
// drained; a split ticket still has to fetch its compacted records (their number is known only now).
static int retire_slot(bvc_ctx* ctx, bvc::Slot& s)
{
    if (!s.ticket) return BVC_OK;
    CK(cudaStreamSynchronize(s.stream));
    { const int g = bvc::guard_verify(ctx); if (g != BVC_OK) { s.ticket = 0; return g; } }
    if (s.n_sites && *s.h_err) {
        s.retired_ticket = s.ticket; s.retired_n = 0; s.retired_rc = BVC_ERR_ARG;
        s.ticket = 0;
        set_err(ctx, (*s.h_err & bvc::kErrQual) ? "bad tile: an entry with qual > 127 (or a packed8 dictionary index >= n_qual_dict)"
                                                : "bad tile: a non-indel entry with base > 3");
        return BVC_ERR_ARG;
    }
    if (s.is_split) {
        const int32_t n = (int32_t)*s.h_vcfc_n;
        s.retired_ticket = s.ticket; s.retired_n = n; s.retired_rc = BVC_OK;
        if (n > s.split_cap) s.retired_rc = BVC_ERR_CAPACITY;
        else if (n > 0) {
            CK(cudaMemcpyAsync(s.split_vcf, s.d_vcfc, (size_t)n * sizeof(bvc_site_result), cudaMemcpyDeviceToHost, s.stream));
            CK(cudaStreamSynchronize(s.stream));
            ctx->d2h_bytes += (int64_t)n * (int64_t)sizeof(bvc_site_result);
        }
    }
    s.ticket = 0;
    return BVC_OK;
}

int bvc_enqueue(bvc_ctx* ctx, const bvc_tile* tile, bvc_site_result* out, uint64_t* ticket)
{
    if (!ctx) return BVC_ERR_ARG;
    if (!out || !ticket) { set_err(ctx, "bvc_enqueue: NULL out/ticket"); return BVC_ERR_ARG; }
    int64_t R = 0;
    int rc = check_tile(ctx, tile, &R);
    if (rc != BVC_OK) return rc;
    CK(cudaSetDevice(ctx->device));
    const uint64_t tk = ctx->next_ticket;
    Slot& s = ctx->slot[tk & 1];
    if ((rc = retire_slot(ctx, s)) != BVC_OK) return rc;
    if ((rc = upload(ctx, s, tile, R, false)) != BVC_OK) return rc;
    if ((rc = launch(ctx, s)) != BVC_OK) return rc;
    if (s.n_sites) CK(cudaMemcpyAsync(out, s.d_out, (size_t)s.n_sites * sizeof(bvc_site_result), cudaMemcpyDeviceToHost, s.stream));
    ctx->d2h_bytes += (int64_t)s.n_sites * (int64_t)sizeof(bvc_site_result);
    CK(cudaEventRecord(s.done, s.stream));
    s.ticket = tk; s.last_ticket = tk; s.is_split = false;
    ctx->next_ticket += 1;
    ctx->resident = false;
    *ticket = tk;
    return BVC_OK;
}

int bvc_enqueue_split(bvc_ctx* ctx, const bvc_tile* tile, bvc_site_core* core, bvc_site_result* vcf, int32_t vcf_cap, uint64_t* ticket)
{
    if (!ctx) return BVC_ERR_ARG;
    if (!core || !ticket || vcf_cap < 0 || (vcf_cap > 0 && !vcf)) { set_err(ctx, "bvc_enqueue_split: NULL core/ticket/vcf"); return BVC_ERR_ARG; }
    int64_t R = 0;
    int rc = check_tile(ctx, tile, &R);
    if (rc != BVC_OK) return rc;
    CK(cudaSetDevice(ctx->device));
    const uint64_t tk = ctx->next_ticket;
    Slot& s = ctx->slot[tk & 1];
    if ((rc = retire_slot(ctx, s)) != BVC_OK) return rc;
    if (!s.d_core) {                                   // first split call on this slot
        const size_t S = (size_t)ctx->prm.max_sites;
        CK(bvc::gmalloc(&s.d_core, (S + 1) * sizeof(bvc_site_core)));
        CK(bvc::gmalloc(&s.d_vcfc, (S + 1) * sizeof(bvc_site_result)));
        CK(bvc::gmalloc(&s.d_vcfc_n, sizeof(uint32_t)));
        CK(cudaHostAlloc(&s.h_vcfc_n, sizeof(uint32_t), cudaHostAllocDefault));
    }
    if ((rc = upload(ctx, s, tile, R, false)) != BVC_OK) return rc;
    if ((rc = launch(ctx, s)) != BVC_OK) return rc;
    CK(bvc::launch_split_result(s.d_out, s.n_sites, s.d_core, s.d_vcfc, s.d_vcfc_n, s.stream));
    ctx->launches += s.n_sites > 0 ? 1 : 0;
    if (s.n_sites) CK(cudaMemcpyAsync(core, s.d_core, (size_t)s.n_sites * sizeof(bvc_site_core), cudaMemcpyDeviceToHost, s.stream));
    CK(cudaMemcpyAsync(s.h_vcfc_n, s.d_vcfc_n, sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
    ctx->d2h_bytes += (int64_t)s.n_sites * (int64_t)sizeof(bvc_site_core) + 4;
    CK(cudaEventRecord(s.done, s.stream));
    s.split_vcf = vcf; s.split_cap = vcf_cap; s.is_split = true;
    s.ticket = tk; s.last_ticket = tk;
    ctx->next_ticket += 1;
    ctx->resident = false;
    *ticket = tk;
    return BVC_OK;
}

size_t bvc_coded_words(int64_t n_entries, int bits) { return n_entries > 0 ? (size_t)((n_entries + 31) / 32) * (size_t)bits : 0; }

int bvc_encode_codes(const bvc_tile* t, uint32_t* words, uint8_t* escapes, uint32_t* chunk_escapes, bvc_coded_tile* out)
{
    if (!out || !t || t->n_sites < 0 || !t->entry_off) return BVC_ERR_ARG;
    const uint8_t* pk = t->packed8;
    const int64_t n = t->entry_off[t->n_sites];
    if (n > 0 && (!pk || !t->ref_base || !words || !escapes || !chunk_escapes)) return BVC_ERR_ARG;
    memset(out, 0, sizeof *out);
    if (n < 4096) return BVC_OK;                                    // not worth a kernel: send packed8
    // relative bytes (base field -> (base - ref) & 3) into a scratch copy: one pass over the sites
    std::vector<uint8_t> rel((size_t)n);
    for (int32_t s = 0; s < t->n_sites; ++s) {
        const unsigned ref = t->ref_base[s] & 3u;
        if (t->entry_off[s] > t->entry_off[s + 1]) return BVC_ERR_ARG;
        for (uint32_t e = t->entry_off[s]; e < t->entry_off[s + 1]; ++e) rel[e] = bvc::abs_to_rel(pk[e], ref);
    }
    int64_t hist[256] = {0};
    for (int64_t e = 0; e < n; ++e) hist[rel[e]]++;
    int order[256];
    for (int i = 0; i < 256; ++i) order[i] = i;
    std::stable_sort(order, order + 256, [&](int a, int b) { return hist[a] > hist[b]; });
    int best = 0; int64_t best_cost = n - n / 10, best_esc = 0;      // must save >= 10 % of the packed8 bytes
    for (int W = 4; W <= 6; ++W) {
        const int nd = (1 << W) - 1;
        int64_t direct = 0;
        for (int i = 0; i < nd; ++i) direct += hist[order[i]];
        const int64_t esc = n - direct, cost = (int64_t)bvc_coded_words(n, W) * 4 + esc;
        if (esc <= n / 8 && cost < best_cost) { best = W; best_cost = cost; best_esc = esc; }
    }
    if (!best) return BVC_OK;
    const int W = best, nd = (1 << W) - 1;
    const uint32_t ESC = (uint32_t)nd;
    uint8_t rank[256];
    memset(rank, (int)ESC, sizeof rank);
    for (int i = 0; i < nd; ++i) if (hist[order[i]] > 0) { rank[order[i]] = (uint8_t)i; out->code_byte[i] = (uint8_t)order[i]; out->n_codes = i + 1; }
    const int64_t n_units = (n + 31) / 32;
    int64_t ne = 0;
    for (int64_t u = 0; u < n_units; ++u) {
        if ((u & 255) == 0) chunk_escapes[u >> 8] = (uint32_t)ne;
        uint32_t w[7] = {0, 0, 0, 0, 0, 0, 0};
        for (int j = 0; j < 32; ++j) {
            const int64_t e = u * 32 + j;
            const uint32_t c = e < n ? rank[rel[(size_t)e]] : 0u;
            if (e < n && c == ESC) escapes[ne++] = pk[e];
            const int bit = j * W, wi = bit >> 5, sh = bit & 31;
            w[wi] |= c << sh;
            if (sh + W > 32) w[wi + 1] |= c >> (32 - sh);
        }
        for (int i = 0; i < W; ++i) words[u * W + i] = w[i];
    }
    chunk_escapes[(n + bvc::kCodedChunk - 1) / bvc::kCodedChunk] = (uint32_t)ne;
    if (ne != best_esc) return BVC_ERR_STATE;
    out->bits = W; out->words = words; out->escapes = escapes; out->chunk_escapes = chunk_escapes; out->n_escape = ne;
    return BVC_OK;
}

int bvc_decode_codes_host(const bvc_coded_tile* c, const bvc_tile* t, uint8_t* dst)
{
    if (!t || t->n_sites < 0 || !t->entry_off) return BVC_ERR_ARG;
    const int64_t n = t->entry_off[t->n_sites];
    if (n > 0 && !t->ref_base) return BVC_ERR_ARG;
    int32_t site = 0;
    if (!c || n < 0 || (n > 0 && (!dst || c->bits < 4 || c->bits > 6 || !c->words || !c->chunk_escapes))) return BVC_ERR_ARG;
    int64_t ne = 0;
    for (int64_t u = 0; u * 32 < n; ++u) {
        if ((u & 255) == 0 && c->chunk_escapes[u >> 8] != (uint32_t)ne) return BVC_ERR_STATE;      // the per-chunk offsets the kernel starts from
        uint32_t w[7] = {0, 0, 0, 0, 0, 0, 0};
        for (int i = 0; i < c->bits; ++i) w[i] = c->words[u * c->bits + i];
        uint8_t out[32]; uint32_t em = 0;
        if (c->bits == 4) bvc::decode_unit<4>(w, c->code_byte, out, &em);
        else if (c->bits == 5) bvc::decode_unit<5>(w, c->code_byte, out, &em);
        else bvc::decode_unit<6>(w, c->code_byte, out, &em);
        for (int j = 0; j < 32 && u * 32 + j < n; ++j) {
            const int64_t e = u * 32 + j;
            while (e >= (int64_t)t->entry_off[site + 1]) ++site;
            dst[e] = (em >> j & 1u) ? c->escapes[ne++] : bvc::rel_to_abs(out[j], t->ref_base[site] & 3u);
        }
    }
    return ne == c->n_escape ? BVC_OK : BVC_ERR_STATE;
}

int bvc_enqueue_split_coded(bvc_ctx* ctx, const bvc_tile* tile, const bvc_coded_tile* coded, bvc_site_core* core, bvc_site_result* vcf,
                            int32_t vcf_cap, uint64_t* ticket)
{
    if (!ctx) return BVC_ERR_ARG;
    if (!tile || !coded || coded->bits < 4 || coded->bits > 6 || !coded->words || !coded->chunk_escapes || coded->n_escape < 0 ||
        (coded->n_escape > 0 && !coded->escapes) || tile->packed8) {
        set_err(ctx, "bvc_enqueue_split_coded: bad coded tile (bits 4..6, arrays, packed8 must be NULL)");
        return BVC_ERR_ARG;
    }
    ctx->pending_coded = coded;
    const int rc = bvc_enqueue_split(ctx, tile, core, vcf, vcf_cap, ticket);
    ctx->pending_coded = nullptr;
    return rc;
}

int bvc_wait_split(bvc_ctx* ctx, uint64_t ticket, int32_t* n_vcf)
{
    if (!ctx || !n_vcf) return BVC_ERR_ARG;
    Slot& s = ctx->slot[ticket & 1];
    if (ticket == 0 || ticket >= ctx->next_ticket) { set_err(ctx, "bvc_wait_split: unknown ticket"); return BVC_ERR_STATE; }
    CK(cudaSetDevice(ctx->device));
    if (s.ticket == ticket) {
        if (!s.is_split) { set_err(ctx, "bvc_wait_split: not a split ticket"); return BVC_ERR_STATE; }
        const int rc = retire_slot(ctx, s);
        if (rc != BVC_OK) return rc;
    }
    if (s.retired_ticket != ticket) { set_err(ctx, "bvc_wait_split: not a split ticket, or an older one than the slot remembers"); return BVC_ERR_STATE; }
    *n_vcf = s.retired_n;
    if (s.retired_rc == BVC_ERR_CAPACITY) set_err(ctx, "bvc_wait_split: more LRT-positive sites than vcf_cap");
    return s.retired_rc;
}

int bvc_wait(bvc_ctx* ctx, uint64_t ticket)
{
    if (!ctx) return BVC_ERR_ARG;
    Slot& s = ctx->slot[ticket & 1];
    if (ticket == 0 || ticket >= ctx->next_ticket) { set_err(ctx, "bvc_wait: unknown ticket"); return BVC_ERR_STATE; }
    if (s.ticket != ticket) {                   // already retired by a later enqueue on the same slot
        if (s.retired_ticket == ticket && s.retired_rc == BVC_ERR_ARG) { set_err(ctx, "bad tile: rejected on the device (qual > 127, non-indel base > 3 or dictionary index out of range)"); return BVC_ERR_ARG; }
        return BVC_OK;
    }
    CK(cudaSetDevice(ctx->device));
    return retire_slot(ctx, s);                 // a split ticket also fetches its records (count: bvc_wait_split)
}

int bvc_fetch_groups(bvc_ctx* ctx, uint64_t ticket, int32_t* group_depth, double* group_af)
{
    if (!ctx) return BVC_ERR_ARG;
    if (ctx->prm.n_groups <= 0 || !group_depth || !group_af) { set_err(ctx, "bvc_fetch_groups: context has no groups / NULL output"); return BVC_ERR_ARG; }
    if (ticket == 0 || ticket >= ctx->next_ticket) { set_err(ctx, "bvc_fetch_groups: unknown ticket"); return BVC_ERR_STATE; }
    Slot& s = ctx->slot[ticket & 1];
    if (s.last_ticket != ticket) { set_err(ctx, "bvc_fetch_groups: the slot of this ticket was re-used by a later bvc_enqueue"); return BVC_ERR_STATE; }
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(s.stream));
    const size_t n = (size_t)s.n_sites * (size_t)ctx->prm.n_groups * 4;
    if (n) {
        CK(cudaMemcpy(group_depth, s.d_gdepth, n * sizeof(int32_t), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(group_af, s.d_gaf, n * sizeof(double), cudaMemcpyDeviceToHost));
        ctx->d2h_bytes += (int64_t)(n * 12);
    }
    return BVC_OK;
}

void* bvc_host_alloc(size_t bytes)
{
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void bvc_host_free(void* p) { if (p) cudaFreeHost(p); }

int bvc_host_register(void* p, size_t bytes)
{
    if (!p || !bytes) return BVC_ERR_ARG;
    if (cudaHostRegister(p, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped) != cudaSuccess) { cudaGetLastError(); return BVC_ERR_CUDA; }
    return BVC_OK;
}
void bvc_host_unregister(void* p) { if (p && cudaHostUnregister(p) != cudaSuccess) cudaGetLastError(); }

int bvc_resident_upload(bvc_ctx* ctx, const bvc_tile* tile)
{
    if (!ctx) return BVC_ERR_ARG;
    int64_t R = 0;
    int rc = check_tile(ctx, tile, &R);
    if (rc != BVC_OK) return rc;
    CK(cudaSetDevice(ctx->device));
    Slot& s = ctx->slot[0];
    if ((rc = retire_slot(ctx, s)) != BVC_OK) return rc;
    CK(cudaStreamSynchronize(s.stream));
    if ((rc = upload(ctx, s, tile, R, true)) != BVC_OK) return rc;
    CK(cudaStreamSynchronize(s.stream));
    ctx->resident = true;
    return BVC_OK;
}

int bvc_resident_run(bvc_ctx* ctx, int iters, int flush_l2, float* ms_per_launch, float* per_kernel_ms, int* n_kernels_out)
{
    if (!ctx) return BVC_ERR_ARG;
    if (!ctx->resident || iters <= 0) { set_err(ctx, "bvc_resident_run: no resident tile / iters <= 0"); return BVC_ERR_STATE; }
    CK(cudaSetDevice(ctx->device));
    Slot& s = ctx->slot[0];
    if (flush_l2 && !ctx->d_flush) { ctx->flush_bytes = (size_t)256 << 20; CK(bvc::gmalloc(&ctx->d_flush, ctx->flush_bytes)); }
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    // per-kernel breakdown: an event before every launch (only when asked for: the records sit between kernels)
    cudaEvent_t ev[bvc::kMaxPipeLaunches + 1];
    int kinds[bvc::kMaxPipeLaunches];
    double kind_ms[bvc::kNumKernelKinds];
    for (int k = 0; k < bvc::kNumKernelKinds; ++k) kind_ms[k] = 0.0;
    if (per_kernel_ms) for (int i = 0; i <= bvc::kMaxPipeLaunches; ++i) CK(cudaEventCreate(&ev[i]));
    double total = 0.0;
    for (int i = 0; i < iters; ++i) {
        if (flush_l2) CK(cudaMemsetAsync(ctx->d_flush, i & 0xff, ctx->flush_bytes, s.stream));
        CK(cudaEventRecord(e0, s.stream));
        int nl = 0;
        int rc = launch(ctx, s, per_kernel_ms ? ev : nullptr, per_kernel_ms ? kinds : nullptr, &nl);
        if (rc != BVC_OK) return rc;
        CK(cudaEventRecord(e1, s.stream));
        CK(cudaEventSynchronize(e1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        total += ms;
        if (per_kernel_ms) for (int l = 0; l < nl; ++l) {
            float k_ms = 0.f;
            CK(cudaEventElapsedTime(&k_ms, ev[l], ev[l + 1]));
            kind_ms[kinds[l]] += k_ms;
        }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (per_kernel_ms) for (int i = 0; i <= bvc::kMaxPipeLaunches; ++i) cudaEventDestroy(ev[i]);
    CK(cudaStreamSynchronize(s.stream));
    if (s.n_sites && *s.h_err) { set_err(ctx, "bad tile: qual > 127, non-indel base > 3 or dictionary index out of range"); return BVC_ERR_ARG; }
    if (ms_per_launch) *ms_per_launch = (float)(total / iters);
    if (per_kernel_ms) for (int k = 0; k < bvc::kNumKernelKinds; ++k) per_kernel_ms[k] = (float)(kind_ms[k] / iters);
    if (n_kernels_out) *n_kernels_out = bvc::kNumKernelKinds;
    return BVC_OK;
}

const char* bvc_kernel_name(int kind) { return bvc::kernel_kind_name(kind); }

int bvc_work_counters(bvc_ctx* ctx, uint64_t* out4)
{
    if (!ctx || !out4) return BVC_ERR_ARG;
    if (!ctx->resident) { set_err(ctx, "bvc_work_counters: no resident tile"); return BVC_ERR_STATE; }
    CK(cudaSetDevice(ctx->device));
    Slot& s = ctx->slot[0];
    CK(cudaStreamSynchronize(s.stream));
    CK(cudaMemcpy(out4, s.d_ctr + 8, 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return BVC_OK;
}

void bvc_transfer_bytes(const bvc_ctx* ctx, int64_t* h2d, int64_t* d2h)
{
    if (h2d) *h2d = ctx ? ctx->h2d_bytes : 0;
    if (d2h) *d2h = ctx ? ctx->d2h_bytes : 0;
}

int bvc_resident_download(bvc_ctx* ctx, bvc_site_result* out)
{
    if (!ctx || !out) return BVC_ERR_ARG;
    if (!ctx->resident) { set_err(ctx, "bvc_resident_download: no resident tile"); return BVC_ERR_STATE; }
    CK(cudaSetDevice(ctx->device));
    Slot& s = ctx->slot[0];
    CK(cudaStreamSynchronize(s.stream));
    if (s.n_sites) CK(cudaMemcpy(out, s.d_out, (size_t)s.n_sites * sizeof(bvc_site_result), cudaMemcpyDeviceToHost));
    return bvc::guard_verify(ctx);
}

int64_t bvc_kernel_launch_count(const bvc_ctx* ctx) { return ctx ? ctx->launches : 0; }

int bvc_measure_fp64_peak(int device, double* tflops)
{
    bvc_ctx* ctx = nullptr;
    if (!tflops) return BVC_ERR_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) { cudaGetLastError(); set_err(nullptr, "bvc_measure_fp64_peak: no such CUDA device"); return BVC_ERR_CUDA; }
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 15;
    double* d = nullptr;
    CK(bvc::gmalloc(&d, (size_t)blocks * threads * sizeof(double)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CK(cudaEventRecord(e0, 0));
        CK(bvc::launch_fp64_peak(d, blocks, threads, iters, 0));
        CK(cudaEventRecord(e1, 0));
        CK(cudaEventSynchronize(e1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); bvc::gfree(d);
    *tflops = 2.0 * 8.0 * (double)iters * blocks * threads / (best * 1e-3) / 1e12;
    return BVC_OK;
}

}  // extern "C"
