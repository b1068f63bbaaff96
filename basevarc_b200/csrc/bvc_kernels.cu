// bvc_kernels.cu -- kernel entry points (sm_100a) and their launchers.
// See bvc_site_kernel.cuh for the design and the reference code each phase replaces.
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <mutex>
#include "bvc_kernels.h"
#include "bvc_ctx.h"
#include "bvc_codes.h"
#include "bvc_site_kernel.cuh"
#include "bvc_pipeline.cuh"
#include "bvc_groups.cuh"
#include "bvc_shard.cuh"

namespace bvc {

__global__ void __launch_bounds__(kWarpsPerBlock * 32)
site_call_kernel(SiteArgs a)
{
    __shared__ __align__(16) unsigned char smem_w[kWarpsPerBlock][kWarpSmemBytes];
    __shared__ double lut_s[4 * kNQ];

    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 4 * kNQ; i += blockDim.x) lut_s[i] = a.lut[i];

    WarpSmem w;
    w.lml = reinterpret_cast<double*>(smem_w[wid]);
    w.hist = reinterpret_cast<uint32_t*>(smem_w[wid] + kNBins * 8);
    w.keys = reinterpret_cast<uint16_t*>(smem_w[wid] + kNBins * 12);
    w.codehist = reinterpret_cast<uint32_t*>(smem_w[wid] + kNBins * 14);
    for (int i = lane; i < kNBins; i += 32) w.hist[i] = 0;
    w.codehist[lane] = 0;
    __syncthreads();

    for (;;) {
        int site = 0;
        if (lane == 0) site = (int)atomicAdd(a.counter, 1u);
        site = __shfl_sync(0xffffffffu, site, 0);
        if (a.site_list) {
            if (site >= (int)*a.list_n) break;
            site = (int)a.site_list[site];
        } else if (site >= a.n_sites) break;

        const uint32_t b0 = a.entry_off[site], b1 = a.entry_off[site + 1];
        const int ref = a.ref_base[site] & 3;

        // ---- Phase A: one coalesced pass over the packed entries -> histograms ------------------
        unsigned qmin = 127, qmax = 0;
        for (uint32_t e = b0 + lane; e < b1; e += 32) {
            const unsigned c = a.code[e];
            const unsigned q = a.qual[e] & (kNQ - 1);
            atomicAdd(&w.codehist[c & 31], 1u);
            if (!(c & 16)) {
                atomicAdd(&w.hist[((c & 3) << 7) | q], 1u);
                qmin = min(qmin, q); qmax = max(qmax, q);
            }
        }
        __syncwarp();
        const int cnt = (int)w.codehist[lane];      // lane = base | strand<<3 | indel<<4
        w.codehist[lane] = 0;
        qmin = __reduce_min_sync(0xffffffffu, qmin);
        qmax = __reduce_max_sync(0xffffffffu, qmax);

        SiteState st;
        st.em_fits = 0; st.em_passes = 0; st.flags = 0; st.n_reads = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {               // src/BaseVarC.cpp:550-559
            st.depth[b] = __shfl_sync(0xffffffffu, cnt, b) + __shfl_sync(0xffffffffu, cnt, b | 8);
            st.n_reads += st.depth[b];
        }
        st.has_q0 = (st.n_reads > 0 && qmin == 0);

        // src/BaseVarC.cpp:576-583: descending argsort, ties in index order; top non-ref base
        int order[4] = {0, 1, 2, 3};
#pragma unroll
        for (int i = 1; i < 4; ++i)
#pragma unroll
            for (int j = i; j > 0; --j)
                if (st.depth[order[j]] > st.depth[order[j - 1]]) { int x = order[j]; order[j] = order[j - 1]; order[j - 1] = x; }
        const int alt = (order[0] != ref) ? order[0] : order[1];

        // src/BaseVarC.cpp:587-602: 2x2 table over ALL entries (indel entries carry stale base/strand)
        int cs[4];
        cs[0] = __shfl_sync(0xffffffffu, cnt, ref | 8) + __shfl_sync(0xffffffffu, cnt, ref | 8 | 16);
        cs[1] = __shfl_sync(0xffffffffu, cnt, ref)     + __shfl_sync(0xffffffffu, cnt, ref | 16);
        cs[2] = __shfl_sync(0xffffffffu, cnt, alt | 8) + __shfl_sync(0xffffffffu, cnt, alt | 8 | 16);
        cs[3] = __shfl_sync(0xffffffffu, cnt, alt)     + __shfl_sync(0xffffffffu, cnt, alt | 16);
        const double cvg_fs = dm_fisher(a.lgam, cs[0], cs[1], cs[2], cs[3]);
        const double cvg_sor = dm_sor(cs[0], cs[1], cs[2], cs[3]);

        // ---- Phase B: compact the non-empty (base, qual) classes in place ------------------------
        int ncls = 0;
        if (st.n_reads > 0) {
            for (int b = 0; b < 4; ++b) {
                if (st.depth[b] == 0) continue;
                for (unsigned q0 = qmin & ~31u; q0 <= qmax; q0 += 32) {
                    const unsigned bin = (b << 7) + q0 + lane;
                    const uint32_t v = w.hist[bin];
                    const unsigned m = __ballot_sync(0xffffffffu, v != 0);
                    const int pos = ncls + __popc(m & ((1u << lane) - 1));
                    __syncwarp();
                    if (v) w.hist[bin] = 0;
                    __syncwarp();
                    if (v) { w.hist[pos] = v; w.keys[pos] = (uint16_t)bin; }
                    ncls += __popc(m);
                    __syncwarp();
                }
            }
        }
        st.ncls = ncls;

        // ---- Phase C: LRT ladder over EM fits ----------------------------------------------------
        const int comb[4] = {0, 1, 2, 3};
        LrtOut L;
        lrt_ladder(a, w, lut_s, st, ref, comb, 4, lane, L);

        // ---- Phase D: VCF-side summaries for LRT-positive sites ----------------------------------
        int vs[4] = {0, 0, 0, 0};
        double vcf_fs = 0.0, vcf_sor = 0.0, rs[3] = {0.0, 0.0, 0.0};
        __syncwarp();
        if (L.ok) {
            unsigned altmask = 0;
            for (int i = 0; i < L.n_alt; ++i) altmask |= 1u << L.alt_bases[i];
            // src/BaseType.cpp:174-187: non-indel entries only, alt = any surviving alt base
            vs[0] = __shfl_sync(0xffffffffu, cnt, ref | 8);
            vs[1] = __shfl_sync(0xffffffffu, cnt, ref);
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int f = __shfl_sync(0xffffffffu, cnt, b | 8), r = __shfl_sync(0xffffffffu, cnt, b);
                if (altmask >> b & 1) { vs[2] += f; vs[3] += r; }
            }
            vcf_fs = dm_fisher(a.lgam, vs[0], vs[1], vs[2], vs[3]);
            vcf_sor = dm_sor(vs[0], vs[1], vs[2], vs[3]);
            // six 256-bin histograms alias lml+hist (6144 B): [feature][ref|alt][256]
            uint32_t* H = reinterpret_cast<uint32_t*>(w.lml);
            for (int i = lane; i < 6 * 256; i += 32) H[i] = 0;
            __syncwarp();
            for (uint32_t e = b0 + lane; e < b1; e += 32) {
                const unsigned c = a.code[e];
                if (c & 16) continue;
                const int base = c & 7;
                int g = -1;
                if (base == ref) g = 0; else if (base < 4 && (altmask >> base & 1)) g = 1;
                if (g < 0) continue;
                atomicAdd(&H[(0 * 2 + g) * 256 + a.mapq[e]], 1u);
                atomicAdd(&H[(1 * 2 + g) * 256 + a.qual[e]], 1u);
                atomicAdd(&H[(2 * 2 + g) * 256 + a.rpr[e]], 1u);
            }
            __syncwarp();
#pragma unroll
            for (int f = 0; f < 3; ++f) rs[f] = ranksum_hist(H + f * 512, H + f * 512 + 256, lane);
            __syncwarp();
            for (int i = lane; i < kNBins; i += 32) w.hist[i] = 0;
        } else {
            for (int i = lane; i < ncls; i += 32) w.hist[i] = 0;
        }
        __syncwarp();

        // ---- write the 176-byte record: stage in shared memory, store as 11 x 16 B ----------------
        bvc_site_result* rec = reinterpret_cast<bvc_site_result*>(w.keys);
        if (lane == 0) {
            rec->cvg_fs = cvg_fs; rec->cvg_sor = cvg_sor;
            for (int j = 0; j < 4; ++j) rec->af_lrt[j] = L.af[j];
            rec->chi = L.chi; rec->var_qual = L.var_qual;
            rec->vcf_fs = vcf_fs; rec->vcf_sor = vcf_sor;
            for (int f = 0; f < 3; ++f) rec->ranksum[f] = rs[f];
            rec->n_entries = (int32_t)(b1 - b0);
            for (int j = 0; j < 4; ++j) { rec->depth[j] = st.depth[j]; rec->cvg_strand[j] = cs[j]; rec->vcf_strand[j] = vs[j]; }
            rec->em_fits = st.em_fits; rec->em_passes = st.em_passes; rec->flags = st.flags;
            rec->cvg_alt_base = (int8_t)alt; rec->lrt_ok = (int8_t)L.ok; rec->n_alt = (int8_t)L.n_alt;
            for (int i = 0; i < 3; ++i) rec->alt_bases[i] = (int8_t)L.alt_bases[i];
            rec->pad[0] = rec->pad[1] = 0;
        }
        __syncwarp();
        if (lane < (int)(sizeof(bvc_site_result) / 16))
            reinterpret_cast<uint4*>(a.out + site)[lane] = reinterpret_cast<const uint4*>(rec)[lane];
        __syncwarp();
    }
}

// Dependent-free DFMA chains: 8 independent accumulators per thread.
__global__ void fp64_peak_kernel(double* out, int iters)
{
    double x0 = threadIdx.x * 1e-9, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    const double m = 1.0000001, c = 1e-7;
    for (int i = 0; i < iters; ++i) {
        x0 = fma(x0, m, c); x1 = fma(x1, m, c); x2 = fma(x2, m, c); x3 = fma(x3, m, c);
        x4 = fma(x4, m, c); x5 = fma(x5, m, c); x6 = fma(x6, m, c); x7 = fma(x7, m, c);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

cudaError_t launch_site_call(const SiteArgs& a, int n_sm, cudaStream_t stream)
{
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, site_call_kernel, kWarpsPerBlock * 32, 0);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    long long want = ((long long)a.n_sites + kWarpsPerBlock - 1) / kWarpsPerBlock;
    long long grid = (long long)n_sm * per_sm;          // persistent: a whole number of waves
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    site_call_kernel<<<(unsigned)grid, kWarpsPerBlock * 32, 0, stream>>>(a);
    return cudaGetLastError();
}

const char* kernel_kind_name(int kind)
{
    static const char* names[kNumKernelKinds] = {"site_hist_kernel", "site_prepare_kernel", "queue_sort_kernels", "em_fit_kernel",
                                                 "em_long_kernel", "fit_loglik_kernel", "ladder_kernel", "vcf_stats_kernel",
                                                 "site_call_kernel"};
    return (kind >= 0 && kind < kNumKernelKinds) ? names[kind] : "?";
}

cudaError_t launch_pipeline(const PipeArgs& a0, int n_sm, cudaStream_t stream, int* n_launches, cudaEvent_t* ev, int* kinds, int parts)
{
    PipeArgs a = a0;
    const int S = a.s.n_sites;
    int launches = 0;
    // occupancies: computed once per process (a dispatcher thread per GPU may arrive here concurrently)
    static std::once_flag occ_once;
    static int occ_hist8 = 1, occ_hist4 = 1, occ_hist = 1, occ_hist16 = 1, occ_em = 1, occ_vcf = 1, occ_leg = 1, occ_long = 1, occ_ll = 1, occ_lad = 1;
    std::call_once(occ_once, [] {
        auto occ = [](int* out, auto kernel, int threads) {
            int v = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, kernel, threads, 0) != cudaSuccess) cudaGetLastError();
            *out = v < 1 ? 1 : v;
        };
        occ(&occ_long, em_long_kernel, kLongThreads);
        occ(&occ_ll, fit_loglik_kernel, kEmThreads);
        // K4 is a static grid-stride loop: with more blocks than fit on the device (it ran 8 per SM where 72 registers allow 7)
        // the blocks of the second, nearly empty wave start only when the first wave's equal shares are done -- twice the time
        occ(&occ_lad, ladder_kernel, 128);
        occ(&occ_hist8, site_hist8_kernel, kHistWarps * 32);
        occ(&occ_hist4, site_hist4_kernel, kHistWarps * 32);
        occ(&occ_hist, site_hist_kernel<32>, kHistWarps * 32);
        occ(&occ_hist16, site_hist_kernel<16>, kHistWarps * 32);
        occ(&occ_em, em_fit_kernel, kEmThreads);
        occ(&occ_vcf, vcf_stats_kernel, kVcfWarps * 32);
        occ(&occ_leg, site_call_kernel, kWarpsPerBlock * 32);
    });
    auto capped = [&](long long want, int per_sm) { long long g = (long long)n_sm * per_sm; if (g > want) g = want; return (unsigned)(g < 1 ? 1 : g); };
    const unsigned site_blocks = (unsigned)((S + 127) / 128);
    // optional per-launch timing: ev[i] is recorded BEFORE launch i, ev[launches] after the last one
    auto mark = [&](int kind) { if (ev) { cudaEventRecord(ev[launches], stream); kinds[launches] = kind; } ++launches; };

    if (parts & kPartHist) {
    mark(kKHist);
    // lanes per site by mean pileup depth of the tile: a full warp from 256 entries per site (measured: C2, 393 entries, 1.11 ms vs 1.23 ms), half a warp below
    if (a.s.pk8)
        site_hist8_kernel<<<capped((S + kHistWarps * kH8Sites - 1) / (kHistWarps * kH8Sites), occ_hist8), kHistWarps * 32, 0, stream>>>(a);
    else if (a.packed_hist)
        site_hist4_kernel<<<capped((S + kHistWarps * kH4Sites - 1) / (kHistWarps * kH4Sites), occ_hist4), kHistWarps * 32, 0, stream>>>(a);
    else if (a.mean_entries >= 256)
        site_hist_kernel<32><<<capped((S + kHistWarps - 1) / kHistWarps, occ_hist), kHistWarps * 32, 0, stream>>>(a);
    else
        site_hist_kernel<16><<<capped((S + 2 * kHistWarps - 1) / (2 * kHistWarps), occ_hist16), kHistWarps * 32, 0, stream>>>(a);
    }
    if (parts & kPartCore) {
    mark(kKPrepare);
    site_prepare_kernel<<<site_blocks, 128, 0, stream>>>(a);
    mark(kKPrepare);
    cvg_fisher_deep_kernel<<<(unsigned)(n_sm * 4), 128, 0, stream>>>(a);
    // at most three EM stages: the full model, the (n-1)-subsets (n >= 3), the pairs of a reduced 4-candidate site
    for (int stage = 0; stage < 3; ++stage) {
        uint32_t* t;
        t = a.q_in; a.q_in = a.q_out; a.q_out = t;
        t = a.q_in_n; a.q_in_n = a.q_out_n; a.q_out_n = t;
        cudaMemsetAsync(a.q_out_n, 0, sizeof(uint32_t), stream);
        if (a.q_long_n == a.q_in_head + 1 && a.q_long_head == a.q_in_head + 2 && a.bins == a.q_in_head + 3)
            cudaMemsetAsync(a.q_in_head, 0, (3 + kNBins + 1) * sizeof(uint32_t), stream);   // q_in_head, q_long_n, q_long_head, bins: one range
        else {
            cudaMemsetAsync(a.q_in_head, 0, sizeof(uint32_t), stream);
            cudaMemsetAsync(a.q_long_n, 0, 2 * sizeof(uint32_t), stream);   // q_long_n, q_long_head
            cudaMemsetAsync(a.bins, 0, (kNBins + 1) * sizeof(uint32_t), stream);
        }
        mark(kKSort);
        queue_hist_kernel<<<(unsigned)(n_sm * 2), kSortThreads, 0, stream>>>(a);
        mark(kKSort);
        queue_scan_kernel<<<1, kSortThreads, 0, stream>>>(a);
        mark(kKSort);
        queue_scatter_kernel<<<(unsigned)(n_sm * 2), kSortThreads, 0, stream>>>(a);
        mark(kKEm);
        em_fit_kernel<<<(unsigned)(n_sm * occ_em), kEmThreads, 0, stream>>>(a);
        mark(kKEmLong);
        em_long_kernel<<<(unsigned)(n_sm * occ_long), kLongThreads, 0, stream>>>(a);
        mark(kKLoglik);
        fit_loglik_kernel<<<(unsigned)(n_sm * occ_ll), kEmThreads, 0, stream>>>(a);
        mark(kKLadder);
        ladder_kernel<<<(unsigned)(n_sm * occ_lad), 128, 0, stream>>>(a);
    }
    }
    if (parts & kPartVcf) {
    mark(kKVcf);
    vcf_stats_kernel<<<capped(S, occ_vcf), kVcfWarps * 32, 0, stream>>>(a);
    // qual-0 sites: the IEEE-faithful warp-per-site kernel over the legacy list
    SiteArgs leg = a.s;
    leg.site_list = a.legacy_list; leg.list_n = a.legacy_n;
    mark(kKLegacy);
    site_call_kernel<<<(unsigned)(n_sm * occ_leg), kWarpsPerBlock * 32, 0, stream>>>(leg);
    }
    if (ev) cudaEventRecord(ev[launches], stream);
    if (n_launches) *n_launches = launches;
    return cudaGetLastError();
}

cudaError_t launch_groups(const SiteArgs& s, const uint32_t* sample, const int32_t* sample_group, int n_groups,
                          int32_t* gdepth, double* gaf, int n_sm, cudaStream_t stream)
{
    GroupArgs g;
    g.s = s; g.sample = sample; g.sample_group = sample_group; g.gdepth = gdepth; g.gaf = gaf; g.n_groups = n_groups;
    const size_t n = (size_t)s.n_sites * n_groups * 4;
    cudaMemsetAsync(gdepth, 0, n * sizeof(int32_t), stream);
    cudaMemsetAsync(gaf, 0, n * sizeof(double), stream);
    static std::once_flag g_once;
    static int occ_gd = 1, occ_gl = 1;                      // one resident wave each (both are static grid-stride loops)
    std::call_once(g_once, [] {
        int v = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, group_depth_kernel, kGroupWarps * 32, 0) == cudaSuccess && v > 0) occ_gd = v;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, group_lrt_kernel, kGroupWarps * 32, 0) == cudaSuccess && v > 0) occ_gl = v;
        cudaGetLastError();
    });
    group_depth_kernel<<<(unsigned)(n_sm * occ_gd), kGroupWarps * 32, 0, stream>>>(g);
    group_lrt_kernel<<<(unsigned)(n_sm * occ_gl), kGroupWarps * 32, 0, stream>>>(g);
    return cudaGetLastError();
}

// ---- sample-sharded mode ---------------------------------------------------------------------------------
static unsigned shard_grid(long long groups_wanted, int n_sm, int per_sm)
{
    long long g = (long long)n_sm * (per_sm < 1 ? 1 : per_sm);
    if (g > groups_wanted) g = groups_wanted;
    return (unsigned)(g < 1 ? 1 : g);
}

cudaError_t launch_shard_hist(const ShardArgs& a, int n_sm, cudaStream_t stream, int* n_launches)
{
    static std::once_flag once;
    static int occ32 = 1, occ16 = 1;
    std::call_once(once, [] {
        int v = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, shard_hist_kernel<32>, kHistWarps * 32, 0) == cudaSuccess && v > 0) occ32 = v;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, shard_hist_kernel<16>, kHistWarps * 32, 0) == cudaSuccess && v > 0) occ16 = v;
        cudaGetLastError();
    });
    const long long n_pad = (long long)a.world * a.Sr;
    if (a.p.mean_entries >= 256)
        shard_hist_kernel<32><<<shard_grid((n_pad + kHistWarps - 1) / kHistWarps, n_sm, occ32), kHistWarps * 32, 0, stream>>>(a);
    else
        shard_hist_kernel<16><<<shard_grid((n_pad + 2 * kHistWarps - 1) / (2 * kHistWarps), n_sm, occ16), kHistWarps * 32, 0, stream>>>(a);
    if (n_launches) *n_launches = 1;
    return cudaGetLastError();
}

// phase A on the one-byte form: site_hist8_kernel's streaming with the dense exchange record as its output
cudaError_t launch_shard_hist8(const ShardArgs& a, const uint8_t* pk8, int n_sm, cudaStream_t stream, int* n_launches)
{
    static std::once_flag once;
    static int occ = 1;
    std::call_once(once, [] {
        int v = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, shard_hist8_kernel, kHistWarps * 32, 0) == cudaSuccess && v > 0) occ = v;
        cudaGetLastError();
    });
    const size_t n_pad = (size_t)a.world * a.Sr;
    cudaError_t e = cudaMemsetAsync(a.dense, 0, n_pad * (size_t)a.B * sizeof(uint32_t), stream);     // a lane writes only its own qual's words
    if (e != cudaSuccess) return e;
    if (n_launches) *n_launches = 0;
    if (a.n_tile_sites <= 0) return cudaSuccess;
    PipeArgs p = a.p;
    p.s.entry_off = a.tile_off; p.s.ref_base = a.tile_ref; p.s.n_sites = a.n_tile_sites; p.s.pk8 = pk8;
    const DenseArgs dn{a.dense, a.err, a.B, a.qlo, a.qhi};
    const long long batches = ((long long)a.n_tile_sites + kHistWarps * kH8Sites - 1) / (kHistWarps * kH8Sites);
    shard_hist8_kernel<<<shard_grid(batches, n_sm, occ), kHistWarps * 32, 0, stream>>>(p, dn);
    if (n_launches) *n_launches = 1;
    return cudaGetLastError();
}

cudaError_t launch_shard_qual_range(const uint8_t* code, const uint8_t* qual, size_t n_entries, uint32_t* out, int n_sm, cudaStream_t stream)
{
    size_t blocks = (n_entries / 4 + 255) / 256;
    if (blocks > (size_t)n_sm * 8) blocks = (size_t)n_sm * 8;
    if (blocks < 1) blocks = 1;
    shard_qual_range_kernel<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const uint32_t*>(code), reinterpret_cast<const uint32_t*>(qual), n_entries, out);
    return cudaGetLastError();
}

// owner: merged histograms -> SiteSum + classes, then K2 / EM stages / K4 on the owned sites, then the mask bytes
cudaError_t launch_shard_merge_and_fit(const ShardArgs& a, int n_sm, cudaStream_t stream, int* n_launches)
{
    static std::once_flag once;
    static int occ = 1;
    std::call_once(once, [] {
        int v = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, shard_merge_kernel<32>, kHistWarps * 32, 0) == cudaSuccess && v > 0) occ = v;
        cudaGetLastError();
    });
    int n = 0;
    if (a.n_own > 0) {
        shard_merge_kernel<32><<<shard_grid((a.n_own + kHistWarps - 1) / kHistWarps, n_sm, occ), kHistWarps * 32, 0, stream>>>(a);
        cudaError_t e = launch_pipeline(a.p, n_sm, stream, &n, nullptr, nullptr, kPartCore);
        if (e != cudaSuccess) return e;
        n += 1;
    }
    shard_posmask_kernel<<<(unsigned)((a.Sr + 255) / 256), 256, 0, stream>>>(a);
    if (n_launches) *n_launches = n + 1;
    return cudaGetLastError();
}

cudaError_t launch_shard_poslist(const ShardArgs& a, cudaStream_t stream, int* n_launches)
{
    shard_poslist_kernel<<<1, 1024, 0, stream>>>(a);
    if (n_launches) *n_launches = 1;
    return cudaGetLastError();
}

cudaError_t launch_shard_vcf_hist(const ShardArgs& a, int n_sm, uint32_t n_pos, cudaStream_t stream, int* n_launches)
{
    if (n_launches) *n_launches = 0;
    if (!n_pos) return cudaSuccess;
    shard_vcf_hist_kernel<<<shard_grid((n_pos + kVcfWarps - 1) / kVcfWarps, n_sm, 4), kVcfWarps * 32, 0, stream>>>(a);
    if (n_launches) *n_launches = 1;
    return cudaGetLastError();
}

cudaError_t launch_shard_vcf_finish(const ShardArgs& a, int n_sm, uint32_t n_pos, cudaStream_t stream, int* n_launches)
{
    if (n_launches) *n_launches = 0;
    if (!n_pos) return cudaSuccess;
    shard_vcf_finish_kernel<<<shard_grid((n_pos + kVcfWarps - 1) / kVcfWarps, n_sm, 4), kVcfWarps * 32, 0, stream>>>(a);
    if (n_launches) *n_launches = 1;
    return cudaGetLastError();
}

cudaError_t launch_sum_u32(uint32_t* dst, const uint32_t* src, size_t n, int n_sm, cudaStream_t stream)
{
    if (!n) return cudaSuccess;
    sum_u32_kernel<<<(unsigned)(n_sm * 8), 256, 0, stream>>>(dst, src, n);
    return cudaGetLastError();
}

// ---- split result: thread per site ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
split_result_kernel(const bvc_site_result* __restrict__ rec, int n_sites, bvc_site_core* __restrict__ core,
                    bvc_site_result* __restrict__ vcf, uint32_t* vcf_n)
{
    const int site = blockIdx.x * blockDim.x + threadIdx.x;
    if (site >= n_sites) return;
    const bvc_site_result r = rec[site];
    bvc_site_core c;
    c.cvg_fs = r.cvg_fs; c.cvg_sor = r.cvg_sor; c.n_entries = r.n_entries;
#pragma unroll
    for (int b = 0; b < 4; ++b) { c.depth[b] = r.depth[b]; c.cvg_strand[b] = r.cvg_strand[b]; }
    c.em_fits = r.em_fits; c.em_passes = r.em_passes; c.flags = r.flags;
    c.cvg_alt_base = r.cvg_alt_base; c.lrt_ok = r.lrt_ok; c.n_alt = r.n_alt; c.reserved = 0;
    c.vcf_index = -1;
    if (r.lrt_ok) {
        const uint32_t i = atomicAdd(vcf_n, 1u);
        vcf[i] = r;
        c.vcf_index = (int32_t)i;
    }
    core[site] = c;
}

cudaError_t launch_split_result(const bvc_site_result* rec, int n_sites, bvc_site_core* core, bvc_site_result* vcf, uint32_t* vcf_n,
                                cudaStream_t stream)
{
    cudaMemsetAsync(vcf_n, 0, sizeof(uint32_t), stream);
    if (n_sites > 0) split_result_kernel<<<(unsigned)((n_sites + 255) / 256), 256, 0, stream>>>(rec, n_sites, core, vcf, vcf_n);
    return cudaGetLastError();
}

// ---- packed8 expansion: 16 entries per thread (one 16-byte load, two 16-byte stores) -------------------------------------
__global__ void __launch_bounds__(256)
expand8_kernel(const uint4* __restrict__ pk, uint4* __restrict__ code, uint4* __restrict__ qual, int64_t n16, QualDict dict)
{
    __shared__ uint8_t code_lut[256], qual_lut[256];
    {
        const unsigned b = threadIdx.x, qi = BVC_PACK8_QI(b), strand = BVC_PACK8_STRAND(b), base = BVC_PACK8_BASE(b);
        code_lut[b] = qi < 30u ? (uint8_t)(base | (strand << 3)) : qi == 31u ? (uint8_t)(base | (strand << 3) | 0x10u) : (uint8_t)(4u | (strand << 3) | 0x10u);
        qual_lut[b] = dict.q[qi < 30u ? qi : 0u];
    }
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (int64_t)gridDim.x * blockDim.x) {
        const uint4 v = __ldg(pk + i);
        const uint32_t in[4] = {v.x, v.y, v.z, v.w};
        uint32_t c[4], q[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t w = in[k];
            c[k] = (uint32_t)code_lut[w & 0xffu] | ((uint32_t)code_lut[(w >> 8) & 0xffu] << 8) | ((uint32_t)code_lut[(w >> 16) & 0xffu] << 16) | ((uint32_t)code_lut[w >> 24] << 24);
            q[k] = (uint32_t)qual_lut[w & 0xffu] | ((uint32_t)qual_lut[(w >> 8) & 0xffu] << 8) | ((uint32_t)qual_lut[(w >> 16) & 0xffu] << 16) | ((uint32_t)qual_lut[w >> 24] << 24);
        }
        code[i] = make_uint4(c[0], c[1], c[2], c[3]);
        qual[i] = make_uint4(q[0], q[1], q[2], q[3]);
    }
}

cudaError_t launch_expand8(const uint8_t* pk8, uint8_t* code, uint8_t* qual, int64_t n, const QualDict& dict, int n_sm, cudaStream_t stream)
{
    const int64_t n16 = (n + 15) / 16;
    if (n16 <= 0) return cudaSuccess;
    long long blocks = (n16 + 255) / 256;
    if (blocks > (long long)n_sm * 8) blocks = (long long)n_sm * 8;
    expand8_kernel<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const uint4*>(pk8), reinterpret_cast<uint4*>(code),
                                                         reinterpret_cast<uint4*>(qual), n16, dict);
    return cudaGetLastError();
}

cudaError_t launch_fp64_peak(double* out, int blocks, int threads, int iters, cudaStream_t stream)
{
    fp64_peak_kernel<<<blocks, threads, 0, stream>>>(out, iters);
    return cudaGetLastError();
}

}  // namespace bvc

// ---- coded wire form -> resident packed8 bytes (bvc_codes.h): a block decodes one chunk of 8,192 entries ---------------
namespace bvc {

struct CodeLut { uint8_t b[64]; };

template <int W>
__global__ void __launch_bounds__(256)
decode_codes_kernel(const uint32_t* __restrict__ words, const uint8_t* __restrict__ esc, const uint32_t* __restrict__ chunk_esc,
                    CodeLut lut, const uint32_t* __restrict__ entry_off, const uint8_t* __restrict__ ref_base, int n_sites,
                    uint8_t* __restrict__ pk8, long long n)
{
    __shared__ uint8_t lut_s[64];
    __shared__ uint32_t warp_tot[8];
    if (threadIdx.x < 64) lut_s[threadIdx.x] = lut.b[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long long n_chunks = (n + kCodedChunk - 1) / kCodedChunk;
    for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        const long long unit = chunk * 256 + threadIdx.x, e0 = unit * 32;
        uint32_t w[W + 1];
        uint8_t out[32];
        uint32_t em = 0;
        const bool on = e0 < n;
        if (on) {
#pragma unroll
            for (int i = 0; i < W; ++i) w[i] = __ldg(words + unit * W + i);
            w[W] = 0;
            decode_unit<W>(w, lut_s, out, &em);
            if (e0 + 32 > n) em &= (1u << (int)(n - e0)) - 1u;           // padding codes behind the last entry are not escapes
            // relative -> absolute base: the site of entry e0 (largest s with entry_off[s] <= e0), then walk on with the entries
            int lo = 0, hi = n_sites;                                    // entry_off[lo] <= e0 < entry_off[hi] (= n)
            while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if ((long long)__ldg(entry_off + mid) <= e0) lo = mid; else hi = mid; }
            int site = lo;
            long long next = (long long)__ldg(entry_off + site + 1);
            unsigned ref = __ldg(ref_base + site) & 3u;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const long long e = e0 + j;
                if (e < n) {
                    while (e >= next) { ++site; next = (long long)__ldg(entry_off + site + 1); ref = __ldg(ref_base + site) & 3u; }
                    out[j] = rel_to_abs(out[j], ref);
                }
            }
        }
        // exclusive prefix of the escape counts over the block: the chunk's escapes lie in entry order behind chunk_esc[chunk]
        const uint32_t cnt = (uint32_t)__popc(em);
        uint32_t incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += up; }
        if (lane == 31) warp_tot[wid] = incl;
        __syncthreads();
        uint32_t base = __ldg(chunk_esc + chunk) + incl - cnt;
        for (int k = 0; k < wid; ++k) base += warp_tot[k];
        __syncthreads();                                              // warp_tot is rewritten by the next chunk
        if (on) {
            for (uint32_t m = em; m; m &= m - 1) out[__ffs(m) - 1] = __ldg(esc + base++);
            if (e0 + 32 <= n) {
                uint4 v[2];
                memcpy(v, out, 32);
                reinterpret_cast<uint4*>(pk8 + e0)[0] = v[0];
                reinterpret_cast<uint4*>(pk8 + e0)[1] = v[1];
            } else {
                for (int j = 0; j < 32 && e0 + j < n; ++j) pk8[e0 + j] = out[j];
            }
        }
    }
}

cudaError_t launch_decode_codes(int bits, const uint32_t* words, const uint8_t* esc, const uint32_t* chunk_esc, const uint8_t* code_byte,
                                const uint32_t* entry_off, const uint8_t* ref_base, int n_sites, uint8_t* pk8, int64_t n, int n_sm, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    CodeLut lut;
    memcpy(lut.b, code_byte, 64);
    long long blocks = (n + kCodedChunk - 1) / kCodedChunk;
    if (blocks > (long long)n_sm * 8) blocks = (long long)n_sm * 8;
    if (bits == 4) decode_codes_kernel<4><<<(unsigned)blocks, 256, 0, stream>>>(words, esc, chunk_esc, lut, entry_off, ref_base, n_sites, pk8, (long long)n);
    else if (bits == 5) decode_codes_kernel<5><<<(unsigned)blocks, 256, 0, stream>>>(words, esc, chunk_esc, lut, entry_off, ref_base, n_sites, pk8, (long long)n);
    else if (bits == 6) decode_codes_kernel<6><<<(unsigned)blocks, 256, 0, stream>>>(words, esc, chunk_esc, lut, entry_off, ref_base, n_sites, pk8, (long long)n);
    else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

}  // namespace bvc

// ---- bvc_call_sites_noalloc: the allocation-free call (include/bvc_gpu.h) ---------------------------------------------
// No counterpart in the reference.  site_call_kernel with everything it
// needs besides the caller's tile and records in ONE file-scope device array: no context, no cudaMalloc.
namespace bvc {
namespace {

constexpr size_t kNaCtrlBytes = 256;                                            // counter, list_n, err (words 0..2)
constexpr size_t kNaLutBytes  = 5 * kNQ * sizeof(double);
constexpr size_t kNaLgamN     = (size_t)BVC_NOALLOC_MAX_SAMPLES + 3;
constexpr size_t kNaLgamBytes = (kNaLgamN * sizeof(double) + 255) / 256 * 256;
constexpr size_t kNaListBytes = (size_t)BVC_NOALLOC_MAX_SITES * sizeof(uint32_t);
constexpr size_t kNaLutOff = kNaCtrlBytes, kNaLgamOff = kNaLutOff + kNaLutBytes, kNaListOff = kNaLgamOff + kNaLgamBytes;
constexpr size_t kNaBytes = kNaListOff + kNaListBytes;                           // 398,848 bytes

// In HOST code this name is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress (noalloc_setup), once per device.  Every call on a device shares it.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaBytes];

constexpr int kNaMaxDev = 64;
std::mutex g_na_mutex;
unsigned char* g_na_ptr[kNaMaxDev];        // device address of g_noalloc_scratch on device d (nullptr: not set up yet)
int g_na_sm[kNaMaxDev];
double g_na_lut[5 * kNQ], g_na_lgam[kNaLgamN];   // host copies of the tables (static: no host allocation either)

// thread per site: the offsets; grid-stride over the entries: what K1 rejects on the pipeline path.  Writes the site list.
__global__ void __launch_bounds__(256)
noalloc_check_kernel(const uint32_t* __restrict__ off, const uint8_t* __restrict__ code, const uint8_t* __restrict__ qual,
                     int n_sites, unsigned long long n_entries, uint32_t max_per_site, uint32_t* list, uint32_t* err)
{
    const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, nt = (unsigned long long)gridDim.x * blockDim.x;
    uint32_t bad = 0;
    for (unsigned long long s = tid; s < (unsigned long long)n_sites; s += nt) {
        const uint32_t a = off[s], b = off[s + 1];
        if ((s == 0 && a != 0) || b < a || b - a > max_per_site || (unsigned long long)b > n_entries) bad |= BVC_NOALLOC_BAD_OFFSETS;
        list[s] = (uint32_t)s;
    }
    for (unsigned long long e = tid; e < n_entries; e += nt) {
        const unsigned c = code[e];
        if (qual[e] > 127) bad |= BVC_NOALLOC_BAD_QUAL;
        if (!(c & 16u) && (c & 7u) > 3u) bad |= BVC_NOALLOC_BAD_BASE;
    }
    bad = __reduce_or_sync(0xffffffffu, bad);
    if (bad && (threadIdx.x & 31) == 0) atomicOr(err, bad);
}

// one thread: the number of sites site_call_kernel may touch (all or none) and the caller's status word
__global__ void noalloc_finish_kernel(const uint32_t* err, uint32_t* list_n, uint32_t n_sites, uint32_t* status)
{
    const uint32_t e = *err;
    *list_n = e ? 0u : n_sites;
    if (status) *status = e;
}

int noalloc_fail(int rc, const char* what, cudaError_t e = cudaSuccess)
{
    char b[256];
    if (e != cudaSuccess) { snprintf(b, sizeof b, "bvc_call_sites_noalloc: %s: %s", what, cudaGetErrorString(e)); cudaGetLastError(); }
    else snprintf(b, sizeof b, "bvc_call_sites_noalloc: %s", what);
    set_err(nullptr, b);
    return rc;
}

// once per device, with that device current
int noalloc_setup(int device)
{
    std::lock_guard<std::mutex> lk(g_na_mutex);
    if (g_na_ptr[device]) return BVC_OK;
    int major = 0, sm = 0;
    cudaError_t e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) return noalloc_fail(BVC_ERR_CUDA, "cudaDeviceGetAttribute", e);
    if (major < 10) return noalloc_fail(BVC_ERR_CUDA, "device is not sm_100 class (kernels are built for sm_100a only)");
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) return noalloc_fail(BVC_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch)", e);
    // the tables of bvc_create (glibc exp / log / lgamma; src/BaseType.cpp:13-15, src/BaseType.h:10), same expressions
    for (int q = 0; q < kNQ; ++q) {
        const double er = exp(-0.23025850929940458 * q);
        g_na_lut[q] = er / 3.0; g_na_lut[kNQ + q] = 1.0 - er;
        g_na_lut[2 * kNQ + q] = log(er / 3.0); g_na_lut[3 * kNQ + q] = log(1.0 - er);
        g_na_lut[4 * kNQ + q] = (1.0 - er) - er / 3.0;
    }
    for (size_t k = 0; k < kNaLgamN; ++k) g_na_lgam[k] = k == 0 ? 0.0 : lgamma((double)k);
    unsigned char* d = static_cast<unsigned char*>(p);
    e = cudaMemcpy(d + kNaLutOff, g_na_lut, sizeof g_na_lut, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d + kNaLgamOff, g_na_lgam, sizeof g_na_lgam, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return noalloc_fail(BVC_ERR_CUDA, "copy of the tables into the static array (first call on a device must be outside stream capture)", e);
    g_na_sm[device] = sm;
    g_na_ptr[device] = d;
    return BVC_OK;
}

}  // namespace
}  // namespace bvc

extern "C" int bvc_call_sites_noalloc(int device, const bvc_params* prm, const bvc_tile* t, int64_t n_entries,
                                      bvc_site_result* d_out, uint32_t* d_status, void* stream)
{
    using namespace bvc;
    if (!prm || !t) return noalloc_fail(BVC_ERR_ARG, "NULL argument");
    if (prm->abi_version != BVC_ABI_VERSION) return noalloc_fail(BVC_ERR_ARG, "abi_version mismatch");
    if (prm->n_samples <= 0 || !(prm->min_af > 0) || prm->em_max_iter < 0 || prm->n_groups != 0)
        return noalloc_fail(BVC_ERR_ARG, "bad parameter (n_samples, min_af > 0, em_max_iter >= 0, n_groups == 0)");
    if (t->n_sites < 0 || n_entries < 0) return noalloc_fail(BVC_ERR_ARG, "n_sites < 0 or n_entries < 0");
    if (t->n_sites > BVC_NOALLOC_MAX_SITES || prm->n_samples > BVC_NOALLOC_MAX_SAMPLES || n_entries >= ((int64_t)1 << 32)) {
        char b[200];
        snprintf(b, sizeof b, "tile beyond the static scratch array: n_sites %d (max %d), n_samples %d (max %d), n_entries %lld (max 2^32-1)",
                 t->n_sites, BVC_NOALLOC_MAX_SITES, prm->n_samples, BVC_NOALLOC_MAX_SAMPLES, (long long)n_entries);
        return noalloc_fail(BVC_ERR_CAPACITY, b);
    }
    if (t->packed8) return noalloc_fail(BVC_ERR_ARG, "packed8 tiles are not served (pass code / qual)");
    if (t->n_sites > 0 && (!t->ref_base || !t->entry_off || !d_out)) return noalloc_fail(BVC_ERR_ARG, "NULL ref_base / entry_off / d_out");
    if (n_entries > 0 && (!t->code || !t->qual || !t->mapq || !t->rpr)) return noalloc_fail(BVC_ERR_ARG, "NULL entry array");
    int ndev = 0, prev = -1;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return noalloc_fail(BVC_ERR_CUDA, "no CUDA device (this library has no CPU path)"); }
    if (device < 0 || device >= ndev || device >= kNaMaxDev) return noalloc_fail(BVC_ERR_ARG, "device index out of range");
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess && prev != device) e = cudaSetDevice(device);
    if (e != cudaSuccess) return noalloc_fail(BVC_ERR_CUDA, "cudaSetDevice", e);
    int rc = noalloc_setup(device);
    if (rc == BVC_OK && t->n_sites > 0) {
        cudaStream_t st = static_cast<cudaStream_t>(stream);
        unsigned char* d = g_na_ptr[device];
        uint32_t* ctrl = reinterpret_cast<uint32_t*>(d);
        uint32_t* list = reinterpret_cast<uint32_t*>(d + kNaListOff);
        const int n_sm = g_na_sm[device];
        // the array keeps the last call's contents: the counter, the list length and the error word start at zero
        e = cudaMemsetAsync(ctrl, 0, kNaCtrlBytes, st);
        if (e == cudaSuccess) {
            long long work = n_entries > t->n_sites ? n_entries : t->n_sites, blocks = (work + 255) / 256;
            if (blocks > (long long)n_sm * 8) blocks = (long long)n_sm * 8;
            noalloc_check_kernel<<<(unsigned)blocks, 256, 0, st>>>(t->entry_off, t->code, t->qual, t->n_sites, (unsigned long long)n_entries,
                                                                  (uint32_t)prm->n_samples, list, ctrl + 2);
            noalloc_finish_kernel<<<1, 1, 0, st>>>(ctrl + 2, ctrl + 1, (uint32_t)t->n_sites, d_status);
            e = cudaGetLastError();
        }
        if (e == cudaSuccess) {
            SiteArgs a;
            a.ref_base = t->ref_base; a.entry_off = t->entry_off; a.code = t->code; a.qual = t->qual; a.mapq = t->mapq; a.rpr = t->rpr;
            a.pk8 = nullptr; memset(&a.dict, 0, sizeof a.dict); a.n_dict = 0; a.err = ctrl + 2;
            a.out = d_out; a.counter = ctrl; a.site_list = list; a.list_n = ctrl + 1;
            a.lut = reinterpret_cast<const double*>(d + kNaLutOff); a.lgam = reinterpret_cast<const double*>(d + kNaLgamOff);
            a.n_sites = t->n_sites; a.max_iter = prm->em_max_iter;
            a.min_af = prm->min_af; a.lrt_thr = prm->lrt_threshold; a.eps = prm->em_epsilon;
            e = launch_site_call(a, n_sm, st);
        }
        if (e != cudaSuccess) rc = noalloc_fail(BVC_ERR_CUDA, "launch", e);
    }
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
    return rc;
}
