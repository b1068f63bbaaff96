// bvc_shard.cuh -- kernels of the SAMPLE-SHARDED mode (cohorts of >= 100k samples).
//
// The reference merges the sample batches of a site on one host thread: the inner loop over the batch files
// (src/BaseVarC.cpp:405-441 under /root/reference) appends every batch's tokens to the site's one
// AlleleInfoVector before bt_f runs.  Here rank r of W holds the entries of ITS sample range for all sites of
// a tile, and the per-site sufficient statistics are exchanged ONCE per tile -- integer histograms, so the merged
// counts are bit-identical whatever the rank order:
//
//   A  shard_hist_kernel      every rank: stream the local entries once (K1's hist_stream) and dump the site's
//                             (base, qual, strand) histogram over the agreed qual window + the 8 indel bins +
//                             n_entries as B = 16 + 8*Qw u32 words into a dense [W*Sr][B] buffer.
//      -- exchange 1: reduce-scatter (sum, u32) -> rank r owns the merged histograms of sites [r*Sr, (r+1)*Sr)
//   B  shard_merge_kernel     owner: dense words -> shared-memory histogram -> K1's hist_emit (SiteSum + classes);
//                             then the unchanged K2 / EM stages / K4 of bvc_pipeline.cuh on the owned sites.
//      shard_posmask_kernel   owner: one byte per owned site (0x10 | ALT mask) for the LRT-positive ones.
//      -- exchange 2: all-gather of the mask bytes
//   C  shard_poslist_kernel   every rank: the same ascending list of LRT-positive sites (deterministic scan).
//      shard_vcf_hist_kernel  every rank: for each listed site the six 256-bin {ref, alt} x {mapq, qual, rpr}
//                             histograms of the LOCAL entries (the statistics of WriteVcf, src/BaseType.cpp:149-197).
//      -- exchange 3: all-reduce (sum, u32) of [P][6*256]
//   D  shard_vcf_finish_kernel owner: exact Mann-Whitney tests from the merged histograms, VCF Fisher / SOR.
//
// Not served in this mode (stated in DESIGN.md): qual-0 entries (the IEEE-faithful legacy kernel needs the raw
// entries; the window must start at qual >= 1) and the --group path.  An entry whose qual lies outside the window
// sets a device error flag and the call fails with BVC_ERR_ARG.
#pragma once
#include "bvc_pipeline.cuh"

namespace bvc {

// ---- A0: the quals this rank's entries actually carry ---------------------------------------------------------------
// out[0] = max(127 - qual), out[1] = max(qual) over the non-indel entries (both start at 0): after a MAX all-reduce over the
// ranks the cohort's qual range of this tile is [127 - out[0], out[1]], and the exchange window shrinks to it -- the dense
// record is 16 + 8 words per qual of the window, and the reduce-scatter of those records is what bounds the mode at 8 GPUs.
__global__ void __launch_bounds__(256)
shard_qual_range_kernel(const uint32_t* __restrict__ code4, const uint32_t* __restrict__ qual4, size_t n_entries, uint32_t* out)
{
    const size_t n4 = n_entries / 4, tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (size_t)gridDim.x * blockDim.x;
    uint32_t lo = 127, hi = 0, seen = 0;
    auto one = [&](uint32_t c, uint32_t q) { if (!(c & 16u)) { q &= 127u; lo = min(lo, q); hi = max(hi, q); seen = 1; } };
    for (size_t i = tid; i < n4; i += nt) {
        const uint32_t c = __ldg(code4 + i), q = __ldg(qual4 + i);
#pragma unroll
        for (int k = 0; k < 4; ++k) one((c >> (8 * k)) & 0xffu, (q >> (8 * k)) & 0xffu);
    }
    if (tid == 0) {
        const uint8_t* c1 = reinterpret_cast<const uint8_t*>(code4); const uint8_t* q1 = reinterpret_cast<const uint8_t*>(qual4);
        for (size_t e = n4 * 4; e < n_entries; ++e) one(c1[e], q1[e]);
    }
    uint32_t a = seen ? 127u - lo : 0u, b = seen ? hi : 0u;
    a = __reduce_max_sync(0xffffffffu, a); b = __reduce_max_sync(0xffffffffu, b);
    if ((threadIdx.x & 31) == 0 && (a | b)) { atomicMax(out, a); atomicMax(out + 1, b); }
}

// ---- A: local histogram of every tile site -> dense words ------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(kHistWarps * 32)
shard_hist_kernel(ShardArgs a)
{
    constexpr int kGroups = 32 / G;
    __shared__ __align__(16) uint32_t hist_s[kHistWarps * kGroups][kHistWords];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int g = lane & (G - 1), gi = lane / G;
    uint32_t* hist = hist_s[wid * kGroups + gi];
    uint32_t* ind = hist + kHistBins;
    for (int i = g; i < kHistWords; i += G) hist[i] = 0;
    __syncwarp();
    const uint32_t* code4 = reinterpret_cast<const uint32_t*>(a.p.s.code);
    const uint32_t* qual4 = reinterpret_cast<const uint32_t*>(a.p.s.qual);
    const int span = 2 * (a.qhi - a.qlo + 1);          // words per base
    const int n_pad = a.world * a.Sr;                  // sites >= S are padding: all-zero records

    const int n_groups = gridDim.x * kHistWarps * kGroups;
    const int first = (blockIdx.x * kHistWarps + wid) * kGroups + gi;
    const int n_iter = (n_pad + n_groups - 1) / n_groups;
    for (int it = 0; it < n_iter; ++it) {
        const int site = first + it * n_groups;
        const bool live = site < a.n_tile_sites;
        const uint32_t b0 = live ? a.tile_off[site] : 0u, b1 = live ? a.tile_off[site + 1] : 0u;
        uint32_t qmax = 0, qmin = 127, badin = 0;
        hist_stream<G>(code4, qual4, b0, b1, g, hist, qmin, qmax, badin);
        if (badin) atomicOr(a.err, 1u);                                             // qual > 127 / non-indel base > 3: the call fails (BVC_ERR_ARG)
        qmax = group_max<G>(qmax, gi);
        qmin = group_min<G>(qmin, gi);
        const bool bad = qmin <= qmax && ((int)qmin < a.qlo || (int)qmax > a.qhi);   // a non-indel qual outside the window
        if (__any_sync(0xffffffffu, bad)) {
            if (bad) {
                if (g == 0) atomicOr(a.err, 1u);
                for (int i = g; i < kHistBins; i += G) hist[i] = 0;                  // keep the bins clean; the call fails anyway
            }
            __syncwarp();
        }
        if (site < n_pad) {
            uint32_t* dst = a.dense + (size_t)site * a.B;
            if (g < kShardHdr) {
                uint32_t w = 0;
                if (g < 8) { w = ind[g]; ind[g] = 0; }
                else if (g == 8) w = b1 - b0;
                dst[g] = w;
            }
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                uint32_t* hb = hist + (b << 8) + (a.qlo << 1);
                for (int i = g; i < span; i += G) { dst[kShardHdr + b * span + i] = hb[i]; hb[i] = 0; }
            }
        }
        __syncwarp();
    }
}

// ---- B: merged dense words of an owned site -> SiteSum + class pool ---------------------------------------------
template <int G>
__global__ void __launch_bounds__(kHistWarps * 32)
shard_merge_kernel(ShardArgs a)
{
    constexpr int kGroups = 32 / G;
    __shared__ __align__(16) uint32_t hist_s[kHistWarps * kGroups][kHistWords];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int g = lane & (G - 1), gi = lane / G;
    uint32_t* hist = hist_s[wid * kGroups + gi];
    uint32_t* ind = hist + kHistBins;
    for (int i = g; i < kHistWords; i += G) hist[i] = 0;
    __syncwarp();
    const int span = 2 * (a.qhi - a.qlo + 1);

    const int n_groups = gridDim.x * kHistWarps * kGroups;
    const int first = (blockIdx.x * kHistWarps + wid) * kGroups + gi;
    const int n_iter = (a.n_own + n_groups - 1) / n_groups;
    for (int it = 0; it < n_iter; ++it) {
        const int s = first + it * n_groups;                    // owned-site index
        const bool live = s < a.n_own;
        const uint32_t* src = a.dense + (size_t)(a.rank * a.Sr + (live ? s : 0)) * a.B;
        const int ref = live ? (a.p.s.ref_base[s] & 3) : 0;
        uint32_t qmax = 0, qmin = 127, hdr = 0;
        if (live && g < kShardHdr) hdr = __ldg(src + g);
        if (live && g < 8) ind[g] = hdr;
        const int n_entries = (int)__shfl_sync(0xffffffffu, hdr, gi * G + 8);
        if (live) {
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                uint32_t* hb = hist + (b << 8) + (a.qlo << 1);
                for (int i = g; i < span; i += G) {
                    const uint32_t v = __ldg(src + kShardHdr + b * span + i);
                    hb[i] = v;
                    if (v) { const uint32_t q = (uint32_t)(a.qlo + (i >> 1)); qmin = min(qmin, q); qmax = max(qmax, q); }
                }
            }
        }
        __syncwarp();
        hist_emit<G>(a.p, s, live, ref, n_entries, qmin, qmax, hist, g, gi);
    }
}

// one byte per site of the owner's padded slice: 0x10 | ALT mask for an LRT-positive site, else 0
__global__ void __launch_bounds__(256)
shard_posmask_kernel(ShardArgs a)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= a.Sr) return;
    uint8_t m = 0;
    if (s < a.n_own) {
        const bvc_site_result* r = a.p.s.out + s;
        if (r->lrt_ok) {
            m = 0x10;
            for (int t = 0; t < r->n_alt; ++t) m |= (uint8_t)(1u << r->alt_bases[t]);
        }
    }
    a.mask_all[(size_t)a.rank * a.Sr + s] = m;
}

// ---- C: the ascending list of LRT-positive sites, identical on every rank (single block, two-level scan) ---------
__global__ void __launch_bounds__(1024)
shard_poslist_kernel(ShardArgs a)
{
    __shared__ uint32_t part[1024];
    const int n = a.world * a.Sr;
    const int chunk = (n + 1023) / 1024;
    const int i0 = min(n, (int)threadIdx.x * chunk), i1 = min(n, i0 + chunk);
    uint32_t c = 0;
    for (int i = i0; i < i1; ++i) c += a.mask_all[i] != 0;
    part[threadIdx.x] = c;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {                       // Hillis-Steele inclusive scan
        const uint32_t v = threadIdx.x >= (unsigned)o ? part[threadIdx.x - o] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t pos = part[threadIdx.x] - c;
    for (int i = i0; i < i1; ++i) if (a.mask_all[i]) a.pos_list[pos++] = (uint32_t)i;
    if (threadIdx.x == 1023) *a.pos_n = part[1023];
}

// warp per listed site: {mapq, qual, rpr} x {ref, alt} histograms of the LOCAL non-indel entries
__global__ void __launch_bounds__(kVcfWarps * 32)
shard_vcf_hist_kernel(ShardArgs a)
{
    __shared__ uint32_t H_s[kVcfWarps][kVcfHistWords];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t* H = H_s[wid];
    const uint32_t n_list = *a.pos_n;
    const uint32_t* code4 = reinterpret_cast<const uint32_t*>(a.p.s.code);
    const uint32_t* qual4 = reinterpret_cast<const uint32_t*>(a.p.s.qual);
    const uint32_t* mapq4 = reinterpret_cast<const uint32_t*>(a.p.s.mapq);
    const uint32_t* rpr4 = reinterpret_cast<const uint32_t*>(a.p.s.rpr);
    for (uint32_t i = blockIdx.x * kVcfWarps + wid; i < n_list; i += gridDim.x * kVcfWarps) {
        const int site = (int)a.pos_list[i];
        const int ref = a.tile_ref[site] & 3;
        const unsigned altmask = a.mask_all[site] & 15u;
        for (int j = lane; j < kVcfHistWords; j += 32) H[j] = 0;
        __syncwarp();
        const uint32_t b0 = a.tile_off[site], b1 = a.tile_off[site + 1];
        const uint32_t w0 = b0 >> 2, w1 = (b1 + 3) >> 2;
        for (uint32_t wi = w0 + lane; wi < w1; wi += 32) {
            const uint32_t cw = __ldg(code4 + wi);
            int grp[4]; bool any = false;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t e = wi * 4 + k;
                const unsigned c = (cw >> (8 * k)) & 31u;
                const int base = c & 7;
                grp[k] = -1;
                if (e >= b0 && e < b1 && !(c & 16)) {
                    if (base == ref) grp[k] = 0; else if (base < 4 && (altmask >> base & 1)) grp[k] = 1;
                }
                any |= grp[k] >= 0;
            }
            if (!any) continue;
            const uint32_t qw = __ldg(qual4 + wi), mw = mapq4[wi], rw = rpr4[wi];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (grp[k] < 0) continue;
                atomicAdd(&H[(0 * 2 + grp[k]) * 256 + ((mw >> (8 * k)) & 255u)], 1u);
                atomicAdd(&H[(1 * 2 + grp[k]) * 256 + ((qw >> (8 * k)) & 255u)], 1u);
                atomicAdd(&H[(2 * 2 + grp[k]) * 256 + ((rw >> (8 * k)) & 255u)], 1u);
            }
        }
        __syncwarp();
        uint32_t* dst = a.hv + (size_t)i * kVcfHistWords;
        for (int j = lane; j < kVcfHistWords; j += 32) dst[j] = H[j];
        __syncwarp();
    }
}

// ---- D: owner finishes the VCF statistics of its LRT-positive sites from the merged histograms -----------------
__global__ void __launch_bounds__(kVcfWarps * 32)
shard_vcf_finish_kernel(ShardArgs a)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t n_list = *a.pos_n;
    const int lo = a.rank * a.Sr, hi = lo + a.n_own;
    for (uint32_t i = blockIdx.x * kVcfWarps + wid; i < n_list; i += gridDim.x * kVcfWarps) {
        const int site = (int)a.pos_list[i];
        if (site < lo || site >= hi) continue;
        const int s = site - lo;
        const SiteSum su = a.p.sum[s];
        bvc_site_result* r = a.p.s.out + s;
        const int ref = a.p.s.ref_base[s] & 3;
        const unsigned altmask = a.mask_all[site] & 15u;
        int vs[4] = {0, 0, 0, 0};                          // src/BaseType.cpp:174-187
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            if (b == ref) { vs[0] = su.fwd[b]; vs[1] = su.depth[b] - su.fwd[b]; }
            else if (altmask >> b & 1) { vs[2] += su.fwd[b]; vs[3] += su.depth[b] - su.fwd[b]; }
        }
        const double vcf_fs = dm_fisher_warp(a.p.s.lgam, vs[0], vs[1], vs[2], vs[3], lane);
        const double vcf_sor = dm_sor(vs[0], vs[1], vs[2], vs[3]);
        const uint32_t* H = a.hv + (size_t)i * kVcfHistWords;
        double rs[3];
        for (int f = 0; f < 3; ++f) rs[f] = ranksum_hist(H + f * 512, H + f * 512 + 256, lane);
        if (lane == 0) {
            r->vcf_fs = vcf_fs; r->vcf_sor = vcf_sor;
            r->ranksum[0] = rs[0]; r->ranksum[1] = rs[1]; r->ranksum[2] = rs[2];
            r->vcf_strand[0] = vs[0]; r->vcf_strand[1] = vs[1]; r->vcf_strand[2] = vs[2]; r->vcf_strand[3] = vs[3];
        }
        __syncwarp();
    }
}

// dst[i] += src[i]: the in-process exchange of bvc_shard_call_group (contexts of one process on one device)
__global__ void __launch_bounds__(256)
sum_u32_kernel(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] += src[i];
}

}  // namespace bvc
