// bvc_pipeline.cuh -- the staged B200 pipeline of the per-site `basetype` calling core.
//
// One tile of sites goes through these kernels instead of the reference's one-site-at-a-time loop
// (bt_s -> bt_f -> BaseType::LRT -> EM -> WriteVcf; paths under /root/reference):
//
//   K1 site_hist_kernel    warp per site.  Streams the packed entries ONCE (code + qual bytes, 32-bit
//                          coalesced loads) into a shared-memory (base, qual) histogram; compacts the
//                          non-empty classes to a global pool; derives depth / strand tables from a
//                          32-bin code histogram.  HBM-bound by construction.
//                          (bt_f counting src/BaseVarC.cpp:548-602; BaseType ctor src/BaseType.cpp:5-23)
//   K2 site_prepare_kernel thread per site.  CVG Fisher + SOR, candidate filter, first ladder step,
//                          emits the full-model fit.  (src/BaseVarC.cpp:603-608, src/BaseType.cpp:73-90)
//   Qs queue_{hist,scan,scatter}_kernel   counting sort of the stage's fit queue by descending class count.
//   K3 em_fit_kernel       THREAD PER FIT, pass-aligned: the sorted queue gives the 32 lanes of a warp fits of
//                          (nearly) equal class count, so the class loop has a warp-uniform trip count, every
//                          pass ends for all lanes together and lanes refill at pass boundaries.  No shuffles,
//                          no per-class state.  FP64-pipe-bound.  A fit that needs more than pass_budget passes
//                          is parked for K3b.  (EM / singleEM / delta_bylog src/Algorithm.cpp:69-130;
//                          UpdateF / SetAlleleFreq src/BaseType.cpp:25-71)
//   K3b em_long_kernel     warp per parked fit (classes across lanes + butterfly sums): bounds the stage's tail.
//   K3c fit_loglik_kernel  thread per fit: lr = sum_c cnt_c log ml_c(f_last).  (src/BaseType.cpp:57-60)
//   K4 ladder_kernel       thread per site.  LRT decision from the fits' lr, next-stage subsets (single-base
//                          models are closed-form) or final ALT set / AF / chi-square QUAL.
//                          (src/BaseType.cpp:93-135, combs_ :237-255, chisf src/Algorithm.cpp:3-7)
//   K5 vcf_stats_kernel    warp per LRT-positive site.  VCF strand table Fisher/SOR and the three
//                          Mann-Whitney tests from 256-bin histograms.  (src/BaseType.cpp:149-197)
//
// EM restatement used by K3/K3b/K3c (all reads of a site with equal (base, qual) share one E-step row):
//   ml_c   = sum_j f_j L_cj = e3_q * F + f_b * (ome_q - e3_q),  F = sum_j f_j          (Algorithm.cpp:76-79)
//   eap_j  = f_j * (sum_c cnt_c e3_q / ml_c + sum_{c: b=j} cnt_c (ome_q - e3_q) / ml_c) / n   (:80-91)
//   delta  = sum_c cnt_c |log ml'_c - log ml_c| = sum_c cnt_c |log1p((ml_c - ml'_c) / ml'_c)|  (:103-113)
// with ml_c recomputed from the previous pass's f (no per-class state).  delta only ever enters the test
// delta < epsilon: log1p is a 5-term series when |x| <= 2^-9, and a class with |x| > 2^-9 alone contributes
// more than epsilon, so that pass is "not converged" without evaluating the sum (see log1p_tiny).
// Differences from the reference are pure rounding (<= a few ulp per pass); the hard thresholds are flagged
// when met within 1e-9 (BVC_FLAG_*).
#pragma once
#include <stdint.h>
#include "../../include/bvc_gpu.h"
#include "bvc_device_math.cuh"
#include "bvc_kernels.h"
#include "bvc_pipeline_args.h"

namespace bvc {

// ---- shared per-class arithmetic ----------------------------------------------------------------------
__device__ __forceinline__ double sel4(int b, double v0, double v1, double v2, double v3)
{
    return b == 0 ? v0 : (b == 1 ? v1 : (b == 2 ? v2 : v3));
}

// =====================================================================================================
// K1: warp per site -- stream entries, histogram, compact classes, strand tables
// =====================================================================================================
constexpr int kHistWarps = 4;
// a class word that contributes exactly nothing (count 0) with finite arithmetic for any valid f: (base 0, qual 30)
constexpr uint32_t kNullClass = 30u;
// Pool layout of a site: [major-base classes, padded with null classes to a multiple of 4][other classes]; the site's
// first word is 16-byte aligned, so K3 / K3b / K3c read the major segment as uint4 (one L1 request per 4 classes).
__host__ __device__ __forceinline__ int pad4(int n) { return (n + 3) & ~3; }
constexpr int kHistBins = 2 * kNBins;      // per site in flight: [base][qual][strand]
constexpr int kHistWords = kHistBins + 8;  // + the 8 indel bins (strand << 2 | stale base) right behind them

// shared-memory increment without a result and without a branch around it
__device__ __forceinline__ void red_inc_if(uint32_t* p, bool pred)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %1, 0;\n\t@q red.shared.add.u32 [%0], 1;\n\t}"
                 :: "r"((uint32_t)__cvta_generic_to_shared(p)), "r"((uint32_t)pred) : "memory");
}

// Reductions over the G lanes of a group.  The *_sync reduce intrinsics with a run-time partial mask compile to a
// ~140-instruction emulation; one full-warp REDUX per group on pre-masked values is 2 instructions.
template <int G> __device__ __forceinline__ int group_add(int v, int gi)
{
    if (G == 32) return __reduce_add_sync(0xffffffffu, v);
    int r = 0;
#pragma unroll
    for (int k = 0; k < 32 / G; ++k) { const int t = __reduce_add_sync(0xffffffffu, gi == k ? v : 0); if (gi == k) r = t; }
    return r;
}
template <int G> __device__ __forceinline__ unsigned group_max(unsigned v, int gi)
{
    if (G == 32) return __reduce_max_sync(0xffffffffu, v);
    unsigned r = 0;
#pragma unroll
    for (int k = 0; k < 32 / G; ++k) { const unsigned t = __reduce_max_sync(0xffffffffu, gi == k ? v : 0u); if (gi == k) r = t; }
    return r;
}
template <int G> __device__ __forceinline__ unsigned group_min(unsigned v, int gi)
{
    if (G == 32) return __reduce_min_sync(0xffffffffu, v);
    unsigned r = 0;
#pragma unroll
    for (int k = 0; k < 32 / G; ++k) { const unsigned t = __reduce_min_sync(0xffffffffu, gi == k ? v : 0xffffffffu); if (gi == k) r = t; }
    return r;
}

// ---- K1's two halves, shared with the sample-sharded kernels of bvc_shard.cuh --------------------------------
// hist_stream: 4 entries per lane per 32-bit load, both arrays in flight.  One shared-memory increment per entry
// does all the counting: the bin carries base, qual and strand.  Bytes outside [b0, b1) in the first / last word are
// turned into code 0x14 (indel, base 4), which nothing counts.  hist: non-indel entries, base << 8 | qual << 1 | strand;
// hist[kHistBins + (strand << 2 | base)]: indel entries (stale base < 4).
template <int G>
__device__ __forceinline__ void hist_stream(const uint32_t* __restrict__ code4, const uint32_t* __restrict__ qual4,
                                            uint32_t b0, uint32_t b1, int g, uint32_t* hist,
                                            uint32_t& qmin, uint32_t& qmax, uint32_t& bad)
{
    const uint32_t w0 = b0 >> 2, w1 = (b1 + 3) >> 2;
    const uint32_t trips = (__reduce_max_sync(0xffffffffu, w1 - w0) + G - 1) / G;
    for (uint32_t t = 0; t < trips; ++t) {
        const uint32_t wi = w0 + t * G + g;
        if (wi < w1) {
            uint32_t cw = __ldg(code4 + wi);
            const uint32_t qraw = __ldg(qual4 + wi), qw = qraw & 0x7f7f7f7fu;
            if (wi == w0 || wi + 1 == w1) {
                const uint32_t e0 = wi * 4;
                const uint32_t lo = b0 > e0 ? b0 - e0 : 0u, hi = b1 < e0 + 4 ? b1 - e0 : 4u;     // valid bytes [lo, hi)
                const uint32_t vm = (hi >= 4 ? 0xffffffffu : ((1u << (8 * hi)) - 1u)) & ~((1u << (8 * lo)) - 1u);
                cw = (cw & vm) | (0x14141414u & ~vm);
            }
            bad |= (((qraw >> 7) & ~((cw >> 2) | (cw >> 4)) & 0x01010101u) ? kErrQual : 0u) | (((cw >> 2) & ~(cw >> 4) & 0x01010101u) ? kErrBase : 0u);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t c = cw >> (8 * k), q = (qw >> (8 * k)) & 0x7fu;
                const bool ok = !(c & 0x14u);
                qmin = min(qmin, ok ? q : 127u); qmax = max(qmax, ok ? q : 0u);
                // one increment per entry with a base < 4: a (base, qual, strand) bin, or for an indel entry (which
                // keeps only its stale base / strand) one of the 8 bins behind the histogram -- no branch for the ~0.5 %
                const uint32_t bin = (c & 0x10u) ? (uint32_t)kHistBins + (((c >> 1) & 4u) | (c & 3u))
                                                 : (((c & 3u) << 8) | (q << 1) | ((c >> 3) & 1u));
                red_inc_if(&hist[bin], !(c & 4u));
            }
        }
    }
    __syncwarp();
}

// hist_emit: from the group's shared-memory histogram to the 64-byte SiteSum + the site's classes in the pool
// (major base first); clears the bins it visited.  qmin / qmax are the lane-local extremes of the quals present.
template <int G>
__device__ __forceinline__ void hist_emit(const PipeArgs& a, int site, bool live, int ref, int n_entries,
                                          uint32_t qmin, uint32_t qmax, uint32_t* hist, int g, int gi)
{
    const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (gi * G));
    const unsigned lt = (1u << g) - 1u;
    // qual range present in this site, scanned G quals at a time per base
    qmax = group_max<G>(qmax, gi);
    qmin = group_min<G>(qmin, gi);
    const int q_lo = (int)min(qmin, qmax);
    const int my_steps = live && n_entries > 0 ? ((int)qmax - q_lo) / G + 1 : 0;
    const int nsteps = __reduce_max_sync(0xffffffffu, my_steps);

    // ---- pass 1 over the bins: merged counts, per-base depth / forward depth / class count -------------------
    int dsum[4], fsum[4], ncb[4];
    unsigned lowq = 0;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        dsum[b] = 0; fsum[b] = 0; ncb[b] = 0;
        for (int st = 0; st < nsteps; ++st) {
            const int q = q_lo + st * G + g;
            uint2 h = make_uint2(0u, 0u);
            if (st < my_steps && q < kNQ) h = *reinterpret_cast<const uint2*>(&hist[(b << 8) | (q << 1)]);
            dsum[b] += (int)(h.x + h.y); fsum[b] += (int)h.y;
            const unsigned m = (__ballot_sync(0xffffffffu, (h.x + h.y) != 0) & gmask) >> (gi * G);
            ncb[b] += __popc(m);
        }
    }
    if (__any_sync(0xffffffffu, q_lo <= 1 && my_steps > 0)) {          // rare: bit 0 = a qual-0 class, bit 1 = a qual-1 class
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const uint4 h = *reinterpret_cast<const uint4*>(&hist[b << 8]);     // quals 0 and 1 of base b, both strands
            lowq |= ((h.x + h.y) ? 1u : 0u) | ((h.z + h.w) ? 2u : 0u);
        }
    }
    int depth[4], fwd[4], n_reads = 0;
#pragma unroll
    for (int b = 0; b < 4; ++b) {               // src/BaseVarC.cpp:550-559
        depth[b] = group_add<G>(dsum[b], gi);
        fwd[b] = group_add<G>(fsum[b], gi);
        n_reads += depth[b];
    }
    const int icnt = g < 8 ? (int)hist[kHistBins + g] : 0;      // g = strand << 2 | base of the indel entries
    if (icnt) hist[kHistBins + g] = 0;
    // src/BaseVarC.cpp:576-583: descending argsort, ties in index order; top non-ref base
    int best = -1, bestd = -1, second = -1, secondd = -1;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        if (depth[b] > bestd) { second = best; secondd = bestd; best = b; bestd = depth[b]; }
        else if (depth[b] > secondd) { second = b; secondd = depth[b]; }
    }
    const int alt = (best != ref) ? best : second;
    // src/BaseVarC.cpp:587-602: 2x2 table over ALL entries (indel entries carry stale base/strand)
    const int fr_ = ref == 0 ? fwd[0] : ref == 1 ? fwd[1] : ref == 2 ? fwd[2] : fwd[3];
    const int dr_ = ref == 0 ? depth[0] : ref == 1 ? depth[1] : ref == 2 ? depth[2] : depth[3];
    const int fa_ = alt == 0 ? fwd[0] : alt == 1 ? fwd[1] : alt == 2 ? fwd[2] : fwd[3];
    const int da_ = alt == 0 ? depth[0] : alt == 1 ? depth[1] : alt == 2 ? depth[2] : depth[3];
    const int gl = gi * G;                      // first lane of my group
    const int c0 = fr_ + __shfl_sync(0xffffffffu, icnt, gl + (ref | 4));
    const int c1 = (dr_ - fr_) + __shfl_sync(0xffffffffu, icnt, gl + ref);
    const int c2 = fa_ + __shfl_sync(0xffffffffu, icnt, gl + (alt | 4));
    const int c3 = (da_ - fa_) + __shfl_sync(0xffffffffu, icnt, gl + alt);

    // ---- pass 2: write the non-empty classes to the pool (and clear the bins).  Order: the MAJOR base (top
    // depth, ties -> lowest code) first, then the other bases ascending; K3 runs the major segment select-free.
    const int ncls = ncb[0] + ncb[1] + ncb[2] + ncb[3];
    const int n_major = best == 0 ? ncb[0] : best == 1 ? ncb[1] : best == 2 ? ncb[2] : ncb[3];
    uint32_t off = 0;
    if (g == 0 && ncls) off = atomicAdd(a.cls_cursor, (uint32_t)pad4(pad4(n_major) + ncls - n_major));
    off = __shfl_sync(0xffffffffu, off, gl);
    if (g < 4 && n_major + g < pad4(n_major)) a.cls[off + n_major + g] = kNullClass;
    uint32_t run = off + (uint32_t)pad4(n_major);
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        uint32_t pos;
        if (b == best) pos = off; else { pos = run; run += (uint32_t)ncb[b]; }
        for (int st = 0; st < nsteps; ++st) {
            const int q = q_lo + st * G + g;
            uint32_t x = 0;
            if (st < my_steps && q < kNQ) {
                uint2* hp = reinterpret_cast<uint2*>(&hist[(b << 8) | (q << 1)]);
                const uint2 h = *hp;
                x = h.x + h.y;
                if (x) *hp = make_uint2(0u, 0u);
            }
            const unsigned m = (__ballot_sync(0xffffffffu, x != 0) & gmask) >> (gi * G);
            if (x) a.cls[pos + __popc(m & lt)] = (x << 9) | (uint32_t)((b << 7) | q);
            pos += __popc(m);
        }
    }
    // ---- 64-byte summary: lanes 0..3 of the group store 16 bytes each (every lane holds all the values) ------------
    uint4 v;
    if (g == 0) v = make_uint4((uint32_t)depth[0], (uint32_t)depth[1], (uint32_t)depth[2], (uint32_t)depth[3]);
    else if (g == 1) v = make_uint4((uint32_t)fwd[0], (uint32_t)fwd[1], (uint32_t)fwd[2], (uint32_t)fwd[3]);
    else if (g == 2) v = make_uint4((uint32_t)c0, (uint32_t)c1, (uint32_t)c2, (uint32_t)c3);
    else v = make_uint4(off, (uint32_t)(ncls | (alt << 16) | ((lowq & 1u) ? (1 << 24) : 0) | (lowq ? (2 << 24) : 0)),
                        (uint32_t)n_entries, (uint32_t)(n_major | (best << 16)));
    if (live && g < 4) reinterpret_cast<uint4*>(a.sum + site)[g] = v;
    __syncwarp();
}

// G lanes per site (32 / G sites per warp): G = 32 for deep pileups, 16 for a few hundred entries per site.
template <int G>
__global__ void __launch_bounds__(kHistWarps * 32)
site_hist_kernel(PipeArgs a)
{
    constexpr int kGroups = 32 / G;
    __shared__ __align__(16) uint32_t hist_s[kHistWarps * kGroups][kHistWords];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int g = lane & (G - 1), gi = lane / G;
    uint32_t* hist = hist_s[wid * kGroups + gi];
    for (int i = g; i < kHistWords; i += G) hist[i] = 0;
    __syncwarp();
    const uint32_t* code4 = reinterpret_cast<const uint32_t*>(a.s.code);
    const uint32_t* qual4 = reinterpret_cast<const uint32_t*>(a.s.qual);

    const int n_groups = gridDim.x * kHistWarps * kGroups;
    const int first = (blockIdx.x * kHistWarps + wid) * kGroups + gi;
    uint32_t bad = 0;
    // warp-uniform trip count: every lane runs the same number of site iterations, idle groups just predicate off
    const int n_iter = (a.s.n_sites + n_groups - 1) / n_groups;
    for (int it = 0; it < n_iter; ++it) {
        const int site = first + it * n_groups;
        const bool live = site < a.s.n_sites;
        const uint32_t b0 = live ? a.s.entry_off[site] : 0u, b1 = live ? a.s.entry_off[site + 1] : 0u;
        const int ref = live ? (a.s.ref_base[site] & 3) : 0;
        uint32_t qmax = 0, qmin = 127;
        hist_stream<G>(code4, qual4, b0, b1, g, hist, qmin, qmax, bad);
        hist_emit<G>(a, site, live, ref, (int)(b1 - b0), qmin, qmax, hist, g, gi);
    }
    const uint32_t err_all = __reduce_or_sync(0xffffffffu, bad);
    if (lane == 0 && err_all) atomicOr(a.s.err, err_all);
}

// =====================================================================================================
// K1 (packed form, cohorts of < 65,536 samples): a warp takes FOUR consecutive sites at a time.
//  * Their entries are one contiguous range of the CSR arrays, streamed by all 32 lanes as ONE range (no
//    per-site partial trips); an entry finds its site by comparing its index with the three inner offsets.
//  * A bin is ONE 32-bit word per (base, qual): forward count in the high half, reverse count in the low half
//    (a site has at most n_samples < 2^16 entries), so a site's histogram is 2 KB + 8 indel words and four
//    sites per warp fit the shared memory of six 4-warp blocks per SM.
//  * The four sites are then emitted side by side, 8 lanes each: the per-site scalar work (argsort, strand
//    table, pool allocation, summary) costs a quarter of the one-site-per-warp form.
// =====================================================================================================
constexpr int kH4Sites = 4;
constexpr int kH4Words = kNBins + 8;       // (base << 7 | qual) words + the 8 indel bins (strand << 2 | stale base)

__device__ __forceinline__ void red_add_if(uint32_t* p, uint32_t v, bool pred)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %2, 0;\n\t@q red.shared.add.u32 [%0], %1;\n\t}"
                 :: "r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v), "r"((uint32_t)pred) : "memory");
}
__device__ __forceinline__ void red_add_s(uint32_t saddr, uint32_t v, bool pred)      // saddr: shared-window byte address
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %2, 0;\n\t@q red.shared.add.u32 [%0], %1;\n\t}"
                 :: "r"(saddr), "r"(v), "r"((uint32_t)pred) : "memory");
}
__device__ __forceinline__ void red_s(uint32_t saddr, uint32_t v)
{
    asm volatile("red.shared.add.u32 [%0], %1;" :: "r"(saddr), "r"(v) : "memory");
}
// code bytes of the word at entry e0 that lie outside [lo_e, hi_e) become 0x14 (indel, base 4): nothing counts them
__device__ __forceinline__ uint32_t mask_outside(uint32_t cw, uint32_t e0, uint32_t lo_e, uint32_t hi_e)
{
    const uint32_t lo = lo_e > e0 ? min(lo_e - e0, 4u) : 0u, hi = hi_e < e0 + 4 ? (hi_e > e0 ? hi_e - e0 : 0u) : 4u;
    const uint32_t vm = (hi >= 4 ? 0xffffffffu : ((1u << (8 * hi)) - 1u)) & ~(lo >= 4 ? 0xffffffffu : ((1u << (8 * lo)) - 1u));
    return (cw & vm) | (0x14141414u & ~vm);
}
__device__ __forceinline__ int bfly8(int v)
{
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    return v;
}

// the words of the first two trips of the range [o0, o4) (lane's word, lane's word + 32), loaded ahead of their use
__device__ __forceinline__ void first_trips(const uint32_t* __restrict__ code4, const uint32_t* __restrict__ qual4, uint32_t o0, uint32_t o4,
                                            int lane, uint32_t& c0, uint32_t& q0, uint32_t& c1, uint32_t& q1)
{
    const uint32_t w1 = (o4 + 3) >> 2;
    uint32_t wi = (o0 >> 2) + (uint32_t)lane;
    if (wi < w1) { c0 = __ldg(code4 + wi); q0 = __ldg(qual4 + wi); }
    wi += 32;
    if (wi < w1) { c1 = __ldg(code4 + wi); q1 = __ldg(qual4 + wi); }
}

__global__ void __launch_bounds__(kHistWarps * 32)
site_hist4_kernel(PipeArgs a)
{
    // + one dump word per warp: entries that count nowhere (indel / N-base codes, bytes outside the range) are added THERE,
    // so the four adds of a word are unconditional (ptxas turns a predicated red into a branch region: 3 more instructions)
    __shared__ __align__(16) uint32_t hist_s[kHistWarps][kH4Sites * kH4Words + 4];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int g = lane & 7, gi = lane >> 3;
    uint32_t* hw = hist_s[wid];
    for (int i = lane; i < kH4Sites * kH4Words; i += 32) hw[i] = 0;
    __syncwarp();
    uint32_t* hist = hw + gi * kH4Words;                        // my group's site during the emit
    const uint32_t hw_s = (uint32_t)__cvta_generic_to_shared(hw);
    const uint32_t dump_s = hw_s + (uint32_t)(kH4Sites * kH4Words * 4);
    const uint32_t hs1 = hw_s + (uint32_t)(kH4Words * 4), hs2 = hs1 + (uint32_t)(kH4Words * 4), hs3 = hs2 + (uint32_t)(kH4Words * 4);
    const uint32_t* code4 = reinterpret_cast<const uint32_t*>(a.s.code);
    const uint32_t* qual4 = reinterpret_cast<const uint32_t*>(a.s.qual);
    const int S = a.s.n_sites;
    const int n_batches = (S + kH4Sites - 1) / kH4Sites;
    const unsigned lt = (1u << g) - 1u;

    const int stride = gridDim.x * kHistWarps;
    int batch = blockIdx.x * kHistWarps + wid;
    uint32_t ov = 0, bad_q = 0, bad_b = 0;                      // bit 0 of each byte
    if (batch < n_batches && lane <= kH4Sites) ov = a.s.entry_off[min(batch * kH4Sites + lane, S)];
    uint32_t pc0 = 0, pq0 = 0, pc1 = 0, pq1 = 0;
    first_trips(code4, qual4, __shfl_sync(0xffffffffu, ov, 0), __shfl_sync(0xffffffffu, ov, kH4Sites), lane, pc0, pq0, pc1, pq1);
    for (; batch < n_batches; batch += stride) {
        const int s0 = batch * kH4Sites;
        // the offsets of this warp's NEXT batch: loaded now, used (a) before the emit to pull that batch's bytes into L2
        // while this one is emitted, (b) as the next iteration's offsets -- no dependent load at the top of a batch
        uint32_t nov = 0;
        if (batch + stride < n_batches && lane <= kH4Sites) nov = a.s.entry_off[min((batch + stride) * kH4Sites + lane, S)];
        const uint32_t o0 = __shfl_sync(0xffffffffu, ov, 0), o1 = __shfl_sync(0xffffffffu, ov, 1), o2 = __shfl_sync(0xffffffffu, ov, 2),
                       o3 = __shfl_sync(0xffffffffu, ov, 3), o4 = __shfl_sync(0xffffffffu, ov, 4);
        // ---- stream [o0, o4): 4 entries per lane per 32-bit load.  A word goes to the site of its FIRST entry; the
        // <= 3 words that straddle a site boundary are corrected afterwards.  Bin byte offsets (base << 9 | qual << 2)
        // are formed two entries at a time in 16-bit lanes; qmin / qmax run over every byte of the range's words
        // (pad and indel bytes carry real quals of neighbouring entries: a superset range is all the emit needs).
        uint32_t qmx2 = 0, qmn2 = 0x007f007fu;                  // packed 16-bit halves (one VIMNMX3.U16x2 per word each)
        const uint32_t w0 = o0 >> 2, w1 = (o4 + 3) >> 2;
        const uint32_t A1 = (o1 + 3) >> 2, A2 = (o2 + 3) >> 2, A3 = (o3 + 3) >> 2;     // first word wholly behind boundary k
        auto proc = [&](uint32_t wi, uint32_t cw, uint32_t qw) {
            if (wi == w0 || wi + 1 == w1) cw = mask_outside(cw, wi * 4, o0, o4);
            // input the ABI rejects (include/bvc_gpu.h BVC_ERR_ARG): a counted entry with qual > 127, a non-indel base > 3
            bad_q |= (qw >> 7) & ~((cw >> 2) | (cw >> 4));
            bad_b |= (cw >> 2) & ~(cw >> 4);
            qw &= 0x7f7f7f7fu;
            const uint32_t hb = wi >= A2 ? (wi >= A3 ? hs3 : hs2) : (wi >= A1 ? hs1 : hw_s);    // A1 <= A2 <= A3
            const uint32_t ce = cw & 0x00ff00ffu, co = (cw >> 8) & 0x00ff00ffu;
            const uint32_t qe = qw & 0x00ff00ffu, qo = (qw >> 8) & 0x00ff00ffu;
            const uint32_t be = ((ce & 0x00030003u) << 9) | (qe << 2), bo = ((co & 0x00030003u) << 9) | (qo << 2);
            qmn2 = __vminu2(qmn2, __vminu2(qe, qo));
            qmx2 = __vmaxu2(qmx2, __vmaxu2(qe, qo));
            red_s((cw & 0x00000014u) ? dump_s : hb + (be & 0xffffu), (cw & 0x00000008u) ? 0x10000u : 1u);
            red_s((cw & 0x00001400u) ? dump_s : hb + (bo & 0xffffu), (cw & 0x00000800u) ? 0x10000u : 1u);
            red_s((cw & 0x00140000u) ? dump_s : hb + (be >> 16), (cw & 0x00080000u) ? 0x10000u : 1u);
            red_s((cw & 0x14000000u) ? dump_s : hb + (bo >> 16), (cw & 0x08000000u) ? 0x10000u : 1u);
            // indel entries (0.5 % of a pile-up) keep only (stale base, strand): a loop over the set indel bits of the word, so
            // a trip costs the warp its lanes' MAXIMUM number of indel entries (almost always 0 or 1) instead of four bodies
            for (uint32_t m = cw & 0x10101010u & ~((cw & 0x04040404u) << 2); m; m &= m - 1) {
                const uint32_t c = cw >> (__ffs((int)m) - 5);
                red_s(hb + 4u * ((uint32_t)kNBins + (((c >> 1) & 4u) | (c & 3u))), 1u);
            }
        };
        // the first two trips were loaded while the previous batch was being emitted (pc / pq registers)
        uint32_t wi = w0 + lane;
        if (wi < w1) proc(wi, pc0, pq0);
        wi += 32;
        if (wi < w1) proc(wi, pc1, pq1);
        // the rest in groups of four trips, the next group's words loaded before the current group is processed
        wi += 32;
        uint32_t gc[4] = {0, 0, 0, 0}, gq[4] = {0, 0, 0, 0};
#pragma unroll
        for (int k = 0; k < 4; ++k) if (wi + 32 * k < w1) { gc[k] = __ldg(code4 + wi + 32 * k); gq[k] = __ldg(qual4 + wi + 32 * k); }
        for (; wi < w1; wi += 128) {
            uint32_t nc[4] = {0, 0, 0, 0}, nq[4] = {0, 0, 0, 0};
#pragma unroll
            for (int k = 0; k < 4; ++k) if (wi + 128 + 32 * k < w1) { nc[k] = __ldg(code4 + wi + 128 + 32 * k); nq[k] = __ldg(qual4 + wi + 128 + 32 * k); }
#pragma unroll
            for (int k = 0; k < 4; ++k) if (wi + 32 * k < w1) proc(wi + 32 * k, gc[k], gq[k]);
#pragma unroll
            for (int k = 0; k < 4; ++k) { gc[k] = nc[k]; gq[k] = nq[k]; }
        }
        // straddling words: lane k-1 <-> boundary k; one lane per distinct straddling word moves the entries behind the
        // boundary from the site the word was credited to to their own sites (atomic adds commute: no ordering needed)
        if (lane < 3) {
            const uint32_t ok_ = lane == 0 ? o1 : lane == 1 ? o2 : o3, op_ = lane == 0 ? o0 : lane == 1 ? o1 : o2;
            const bool dup = lane > 0 && (op_ & 3u) && (op_ >> 2) == (ok_ >> 2);      // the previous boundary's lane has this word
            if ((ok_ & 3u) && ok_ < o4 && !dup) {
                const uint32_t wi = ok_ >> 2, e0 = wi * 4;
                const uint32_t cw = mask_outside(__ldg(code4 + wi), e0, o0, o4), qw = __ldg(qual4 + wi) & 0x7f7f7f7fu;
                const uint32_t kf = (uint32_t)(wi >= A1) + (uint32_t)(wi >= A2) + (uint32_t)(wi >= A3);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t c = cw >> (8 * k), q = (qw >> (8 * k)) & 0x7fu, e = e0 + k;
                    const uint32_t ks = (uint32_t)(e >= o1) + (uint32_t)(e >= o2) + (uint32_t)(e >= o3);
                    const uint32_t bin = (c & 0x10u) ? (uint32_t)kNBins + (((c >> 1) & 4u) | (c & 3u)) : (((c & 3u) << 7) | q);
                    const uint32_t inc = (c & 0x10u) ? 1u : (1u << ((c & 8u) << 1));
                    const bool mv = !(c & 4u) && ks != kf;
                    red_add_s(hw_s + 4u * (kf * kH4Words + bin), 0u - inc, mv);
                    red_add_s(hw_s + 4u * (ks * kH4Words + bin), inc, mv);
                }
            }
        }
        __syncwarp();
        {
            const uint32_t n0 = __shfl_sync(0xffffffffu, nov, 0), n4 = __shfl_sync(0xffffffffu, nov, kH4Sites);
            first_trips(code4, qual4, n0, n4, lane, pc0, pq0, pc1, pq1);      // land while this batch is emitted
            for (uint32_t pa = (n0 & ~127u) + 256u + (uint32_t)lane * 128u; pa < n4; pa += 32u * 128u) {    // the rest: into L2
                asm volatile("prefetch.global.L2 [%0];" :: "l"(a.s.code + pa));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(a.s.qual + pa));
            }
        }
        // one qual range for the whole batch (a superset for each of its sites)
        const uint32_t qmax = __reduce_max_sync(0xffffffffu, max(qmx2 & 0xffffu, qmx2 >> 16));
        const uint32_t qmin = __reduce_min_sync(0xffffffffu, min(qmn2 & 0xffffu, qmn2 >> 16));
        const int q_lo = (int)min(qmin, qmax);
        const int nsteps = o4 > o0 ? ((int)qmax - q_lo) / 8 + 1 : 0;

        // ---- emit: group gi <-> site s0 + gi ---------------------------------------------------------------------
        const int site = s0 + gi;
        const bool live = site < S;
        const int ref = live ? (a.s.ref_base[site] & 3) : 0;
        const int n_entries = (int)(gi == 0 ? o1 - o0 : gi == 1 ? o2 - o1 : gi == 2 ? o3 - o2 : o4 - o3);
        // Fast form (the batch's quals span <= 32 values from a 4-aligned start -- any real base-quality scale): lane g of
        // a group owns quals q0 + 4g .. q0 + 4g + 3 of every base, read as ONE 16-byte word per base; depths and class
        // counts travel packed (two 16-bit depths / four 8-bit counts per word) through 3-step group butterflies / scans.
        const int q0 = min(q_lo & ~3, kNQ - 32);
        const bool fast = o4 > o0 && (int)qmax < q0 + 32;       // warp-uniform
        int depth[4], fwd[4], ncb[4];
        uint32_t pre = 0;                                       // fast form: classes of base b in the group's lower lanes, 8 bits per base
        unsigned lowq = 0;
        if (q_lo <= 1 && nsteps > 0) {                          // rare: bit 0 = a qual-0 class, bit 1 = a qual-1 class
#pragma unroll
            for (int b = 0; b < 4; ++b) lowq |= (hist[b << 7] ? 1u : 0u) | (hist[(b << 7) | 1] ? 2u : 0u);
        }
        if (fast) {
            uint32_t dpk[2] = {0, 0}, fpk[2] = {0, 0}, cpk = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const uint4 hv = *reinterpret_cast<const uint4*>(hist + (b << 7) + q0 + 4 * g);
                const uint32_t hx[4] = {hv.x, hv.y, hv.z, hv.w};
                uint32_t ds = 0, fs = 0, c = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t x = (hx[k] & 0xffffu) + (hx[k] >> 16);
                    ds += x; fs += hx[k] >> 16; c += (x != 0);
                }
                dpk[b >> 1] |= ds << (16 * (b & 1)); fpk[b >> 1] |= fs << (16 * (b & 1)); cpk |= c << (8 * b);
            }
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {                   // a site has < 65,536 entries: the 16-bit halves never carry
                dpk[0] += __shfl_xor_sync(0xffffffffu, dpk[0], o); dpk[1] += __shfl_xor_sync(0xffffffffu, dpk[1], o);
                fpk[0] += __shfl_xor_sync(0xffffffffu, fpk[0], o); fpk[1] += __shfl_xor_sync(0xffffffffu, fpk[1], o);
            }
            uint32_t inc = cpk;                                 // inclusive scan over the group's lanes (<= 32 classes per base)
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
                if (g >= o) inc += t;
            }
            pre = inc - cpk;
            const uint32_t tot = __shfl_sync(0xffffffffu, inc, gi * 8 + 7);
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                depth[b] = (int)((dpk[b >> 1] >> (16 * (b & 1))) & 0xffffu);
                fwd[b] = (int)((fpk[b >> 1] >> (16 * (b & 1))) & 0xffffu);
                ncb[b] = (int)((tot >> (8 * b)) & 0xffu);
            }
        } else {
            int dsum[4], fsum[4];
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                dsum[b] = 0; fsum[b] = 0; ncb[b] = 0;
                for (int st = 0; st < nsteps; ++st) {
                    const int q = q_lo + st * 8 + g;
                    const uint32_t h = q < kNQ ? hist[(b << 7) | q] : 0u;
                    const uint32_t x = (h & 0xffffu) + (h >> 16);
                    dsum[b] += (int)x; fsum[b] += (int)(h >> 16);
                    ncb[b] += __popc((__ballot_sync(0xffffffffu, x != 0) >> (gi * 8)) & 0xffu);
                }
            }
#pragma unroll
            for (int b = 0; b < 4; ++b) { depth[b] = bfly8(dsum[b]); fwd[b] = bfly8(fsum[b]); }   // src/BaseVarC.cpp:550-559
        }
        const int icnt = (int)hist[kNBins + g];                 // g = strand << 2 | base of the indel entries
        if (icnt) hist[kNBins + g] = 0;
        // src/BaseVarC.cpp:576-583: descending argsort, ties in index order; top non-ref base
        int best = -1, bestd = -1, second = -1, secondd = -1;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            if (depth[b] > bestd) { second = best; secondd = bestd; best = b; bestd = depth[b]; }
            else if (depth[b] > secondd) { second = b; secondd = depth[b]; }
        }
        const int alt = (best != ref) ? best : second;
        // src/BaseVarC.cpp:587-602: 2x2 table over ALL entries (indel entries carry stale base/strand)
        const int fr_ = ref == 0 ? fwd[0] : ref == 1 ? fwd[1] : ref == 2 ? fwd[2] : fwd[3];
        const int dr_ = ref == 0 ? depth[0] : ref == 1 ? depth[1] : ref == 2 ? depth[2] : depth[3];
        const int fa_ = alt == 0 ? fwd[0] : alt == 1 ? fwd[1] : alt == 2 ? fwd[2] : fwd[3];
        const int da_ = alt == 0 ? depth[0] : alt == 1 ? depth[1] : alt == 2 ? depth[2] : depth[3];
        const int gl = gi * 8;
        const int c0 = fr_ + __shfl_sync(0xffffffffu, icnt, gl + (ref | 4));
        const int c1 = (dr_ - fr_) + __shfl_sync(0xffffffffu, icnt, gl + ref);
        const int c2 = fa_ + __shfl_sync(0xffffffffu, icnt, gl + (alt | 4));
        const int c3 = (da_ - fa_) + __shfl_sync(0xffffffffu, icnt, gl + alt);
        // pool space for the four sites with one atomic
        const int ncls = ncb[0] + ncb[1] + ncb[2] + ncb[3];
        const int n_major = best == 0 ? ncb[0] : best == 1 ? ncb[1] : best == 2 ? ncb[2] : ncb[3];
        const int nalloc = ncls ? pad4(pad4(n_major) + ncls - n_major) : 0;
        const int n0 = __shfl_sync(0xffffffffu, nalloc, 0), n1 = __shfl_sync(0xffffffffu, nalloc, 8),
                  n2 = __shfl_sync(0xffffffffu, nalloc, 16), n3 = __shfl_sync(0xffffffffu, nalloc, 24);
        uint32_t off = 0;
        if (lane == 0 && n0 + n1 + n2 + n3) off = atomicAdd(a.cls_cursor, (uint32_t)(n0 + n1 + n2 + n3));
        off = __shfl_sync(0xffffffffu, off, 0) + (uint32_t)(gi > 0 ? n0 : 0) + (uint32_t)(gi > 1 ? n1 : 0) + (uint32_t)(gi > 2 ? n2 : 0);
        // classes to the pool, the MAJOR base first (K3 runs that segment select-free); clears the bins
        if (g < 4 && n_major + g < pad4(n_major)) a.cls[off + n_major + g] = kNullClass;
        uint32_t run = off + (uint32_t)pad4(n_major);
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            uint32_t pos;
            if (b == best) pos = off; else { pos = run; run += (uint32_t)ncb[b]; }
            if (fast) {
                uint4* hp = reinterpret_cast<uint4*>(hist + (b << 7) + q0 + 4 * g);
                const uint4 hv = *hp;
                const uint32_t hx[4] = {hv.x, hv.y, hv.z, hv.w};
                pos += (pre >> (8 * b)) & 0xffu;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const uint32_t x = (hx[k] & 0xffffu) + (hx[k] >> 16);
                    if (x) a.cls[pos++] = (x << 9) | (uint32_t)((b << 7) | (q0 + 4 * g + k));
                }
                *hp = make_uint4(0u, 0u, 0u, 0u);
                continue;
            }
            for (int st = 0; st < nsteps; ++st) {
                const int q = q_lo + st * 8 + g;
                uint32_t x = 0;
                if (q < kNQ) {
                    const uint32_t h = hist[(b << 7) | q];
                    x = (h & 0xffffu) + (h >> 16);
                    if (x) hist[(b << 7) | q] = 0;
                }
                const unsigned m = (__ballot_sync(0xffffffffu, x != 0) >> (gi * 8)) & 0xffu;
                if (x) a.cls[pos + __popc(m & lt)] = (x << 9) | (uint32_t)((b << 7) | q);
                pos += __popc(m);
            }
        }
        uint4 v;
        if (g == 0) v = make_uint4((uint32_t)depth[0], (uint32_t)depth[1], (uint32_t)depth[2], (uint32_t)depth[3]);
        else if (g == 1) v = make_uint4((uint32_t)fwd[0], (uint32_t)fwd[1], (uint32_t)fwd[2], (uint32_t)fwd[3]);
        else if (g == 2) v = make_uint4((uint32_t)c0, (uint32_t)c1, (uint32_t)c2, (uint32_t)c3);
        else v = make_uint4(off, (uint32_t)(ncls | (alt << 16) | ((lowq & 1u) ? (1 << 24) : 0) | (lowq ? (2 << 24) : 0)),
                            (uint32_t)n_entries, (uint32_t)(n_major | (best << 16)));
        if (live && g < 4) reinterpret_cast<uint4*>(a.sum + site)[g] = v;
        ov = nov;
        __syncwarp();
    }
    const uint32_t err_all = __reduce_or_sync(0xffffffffu, ((bad_q & 0x01010101u) ? kErrQual : 0u) | ((bad_b & 0x01010101u) ? kErrBase : 0u));
    if (lane == 0 && err_all) atomicOr(a.s.err, err_all);
}

// =====================================================================================================
// K1 on the ONE-byte form (BVC_PACK8: base:2 | qi:5 | strand:1; the form every tile with <= 30 distinct quals
// travels and stays resident in).  The byte IS the bin: a site's histogram is 256 words [qi][strand][base], so an
// entry costs a byte extract, an address add and one red.shared -- no qual/base arithmetic, no predicates (indel
// entries are qi 31 = the 8 [strand][stale base] bins the CVG table needs, qi 30 counts nowhere).
//  * A warp takes FOUR consecutive sites; their bytes are one contiguous range read as 16-byte words
//    (ld.global.nc.v4, the next words in flight while the current ones are counted).  A 16-byte word is credited
//    to the site of its first byte; the <= 5 words that straddle a site boundary or the ends of the range are
//    corrected byte-wise afterwards (adds commute, so a wrong credit is undone by the opposite add).
//  * Emit, one site at a time with lane = dictionary index: two 16-byte shared-memory reads give the lane its
//    (strand, base) counts, four REDUX sums the packed depths, four ballots the class sets; a class's position in
//    its base's segment is the number of present classes with a LOWER qual (the dictionary need not be sorted).
//  * The site's classes land at a FIXED pool offset (site * cls_stride): no allocation atomic, reproducible layout.
// Reference semantics: src/BaseVarC.cpp:548-602 (counting, argsort, CVG table), src/BaseType.cpp:5-23.
// =====================================================================================================
constexpr int kH8Sites = 4;
constexpr int kH8Words = 256;
constexpr int kH8Stage = 128;              // staging words per site: pad4(major classes) + the other classes <= 4 * 30 + 3

__device__ __forceinline__ uint4 ldg16(const uint4* p) { return __ldg(p); }

#ifndef BVC_H8_MINBLOCKS
#define BVC_H8_MINBLOCKS 7
#endif
// kDense (sample-sharded mode, bvc_shard.cuh): the same streaming, but a site's counts go out as the dense exchange record
// [8 indel bins | n_entries | 7 spare | 4 bases x (qual window) x 2 strands] instead of SiteSum + classes.  The caller zeroes
// the dense buffer first: a lane writes only the words of ITS dictionary qual.
struct DenseArgs { uint32_t* dense; uint32_t* err; int B, qlo, qhi; };

template <bool kDense>
__device__ __forceinline__ void site_hist8_body(const PipeArgs& a, const DenseArgs& dn)
{
    __shared__ __align__(16) uint32_t hist_s[kHistWarps][kH8Sites * kH8Words];
    __shared__ __align__(16) uint32_t stage_s[kHistWarps][kH8Sites * kH8Stage];
    __shared__ __align__(16) uint32_t zero8[8];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t* hw = hist_s[wid];
    uint32_t* stage = stage_s[wid];
    for (int i = lane; i < kH8Sites * kH8Words; i += 32) hw[i] = 0;
    if (threadIdx.x < 8) zero8[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t hs0 = (uint32_t)__cvta_generic_to_shared(hw);
    const uint32_t hs1 = hs0 + 4u * kH8Words, hs2 = hs1 + 4u * kH8Words, hs3 = hs2 + 4u * kH8Words;
    const int nd = a.s.n_dict;
    // my dictionary qual; lanes whose class sorts before mine (ascending qual, ties by index)
    const uint32_t dq = lane < nd ? (uint32_t)a.s.dict.q[lane] : 255u;
    uint32_t lower = 0;
    for (int j = 0; j < BVC_PACK8_MAX_QUALS; ++j) {
        const uint32_t qj = __shfl_sync(0xffffffffu, dq, j);
        if (j < nd && (qj < dq || (qj == dq && j < lane))) lower |= 1u << j;
    }
    const bool counted = lane < BVC_PACK8_MAX_QUALS;
    const unsigned q0m = __ballot_sync(0xffffffffu, dq == 0), q1m = __ballot_sync(0xffffffffu, dq <= 1);
    __syncwarp();
    const uint4* pk16 = reinterpret_cast<const uint4*>(a.s.pk8);
    const int S = a.s.n_sites;
    const int n_batches = (S + kH8Sites - 1) / kH8Sites;
    const int stride = gridDim.x * kHistWarps;
    uint32_t err = 0;

    int batch = blockIdx.x * kHistWarps + wid;
    uint32_t ov = 0, refv = 0;
    if (batch < n_batches && lane <= kH8Sites) { ov = a.s.entry_off[min(batch * kH8Sites + lane, S)]; refv = a.s.ref_base[min(batch * kH8Sites + lane, S - 1)]; }
    // the first two trips of a batch are loaded while the previous batch is emitted
    uint4 px0 = make_uint4(0, 0, 0, 0), px1 = px0;
    auto first_two = [&](uint32_t b0, uint32_t b4) {
        const uint32_t u1 = (b4 + 15) >> 4;
        uint32_t vi = (b0 >> 4) + (uint32_t)lane;
        if (vi < u1) px0 = ldg16(pk16 + vi);
        vi += 32;
        if (vi < u1) px1 = ldg16(pk16 + vi);
    };
    first_two(__shfl_sync(0xffffffffu, ov, 0), __shfl_sync(0xffffffffu, ov, kH8Sites));
    for (; batch < n_batches; batch += stride) {
        const int s0 = batch * kH8Sites;
        uint32_t nov = 0, nrefv = 0;                            // the next batch's offsets and reference bases: no dependent load in the emit
        if (batch + stride < n_batches && lane <= kH8Sites) {
            nov = a.s.entry_off[min((batch + stride) * kH8Sites + lane, S)];
            nrefv = a.s.ref_base[min((batch + stride) * kH8Sites + lane, S - 1)];
        }
        const uint32_t o0 = __shfl_sync(0xffffffffu, ov, 0), o1 = __shfl_sync(0xffffffffu, ov, 1), o2 = __shfl_sync(0xffffffffu, ov, 2),
                       o3 = __shfl_sync(0xffffffffu, ov, 3), o4 = __shfl_sync(0xffffffffu, ov, 4);
        const uint32_t v0 = o0 >> 4, v1 = (o4 + 15) >> 4;
        // 16-byte word v is credited to the site of its first byte: 16 v >= o_k  <=>  v >= B_k
        const uint32_t B1 = (o1 + 15) >> 4, B2 = (o2 + 15) >> 4, B3 = (o3 + 15) >> 4;
        auto proc = [&](uint32_t vi, const uint4& x) {
            const uint32_t hb = vi >= B2 ? (vi >= B3 ? hs3 : hs2) : (vi >= B1 ? hs1 : hs0);
            const uint32_t xw[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                red_s(hb + (__byte_perm(xw[j], 0, 0x4440) << 2), 1u);
                red_s(hb + (__byte_perm(xw[j], 0, 0x4441) << 2), 1u);
                red_s(hb + (__byte_perm(xw[j], 0, 0x4442) << 2), 1u);
                red_s(hb + (__byte_perm(xw[j], 0, 0x4443) << 2), 1u);
            }
        };
        if (o4 > o0) {
            uint32_t vi = v0 + (uint32_t)lane;
            if (vi < v1) proc(vi, px0);
            vi += 32;
            if (vi < v1) proc(vi, px1);
            vi += 32;
            // the rest two trips at a time, the next pair loaded before the current pair is counted
            uint4 g0 = make_uint4(0, 0, 0, 0), g1 = g0;
            if (vi < v1) g0 = ldg16(pk16 + vi);
            if (vi + 32 < v1) g1 = ldg16(pk16 + vi + 32);
            for (; vi < v1; vi += 64) {
                uint4 n0 = make_uint4(0, 0, 0, 0), n1 = n0;
                if (vi + 64 < v1) n0 = ldg16(pk16 + vi + 64);
                if (vi + 96 < v1) n1 = ldg16(pk16 + vi + 96);
                proc(vi, g0);
                if (vi + 32 < v1) proc(vi + 32, g1);
                g0 = n0; g1 = n1;
            }
            // ---- corrections: every byte of the <= 5 special words (first, last, the three straddlers) whose true site is not
            // the credited one moves; bytes outside [o0, o4) are taken out.  Half a warp per word, three rounds.
            const uint32_t sv1 = o1 >> 4, sv2 = o2 >> 4, sv3 = o3 >> 4, sv4 = v1 - 1;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const int sp = 2 * r + (lane >> 4);
                const uint32_t v = sp == 0 ? v0 : sp == 1 ? sv1 : sp == 2 ? sv2 : sp == 3 ? sv3 : sv4;
                bool on = sp <= 4 && v >= v0 && v < v1;
                if (sp >= 1 && v == v0) on = false;
                if (sp >= 2 && v == sv1) on = false;
                if (sp >= 3 && v == sv2) on = false;
                if (sp >= 4 && v == sv3) on = false;
                if (on) {
                    const uint32_t e = v * 16u + (uint32_t)(lane & 15);
                    const uint32_t byte = a.s.pk8[e];
                    const uint32_t kf = (uint32_t)(v >= B1) + (uint32_t)(v >= B2) + (uint32_t)(v >= B3);
                    const uint32_t ks = (uint32_t)(e >= o1) + (uint32_t)(e >= o2) + (uint32_t)(e >= o3);
                    const bool inside = e >= o0 && e < o4;
                    if (!inside || ks != kf) {
                        red_s(hs0 + 4u * (kf * kH8Words + byte), 0xffffffffu);
                        if (inside) red_s(hs0 + 4u * (ks * kH8Words + byte), 1u);
                    }
                }
            }
        }
        __syncwarp();
        {
            const uint32_t n0 = __shfl_sync(0xffffffffu, nov, 0), n4 = __shfl_sync(0xffffffffu, nov, kH8Sites);
            first_two(n0, n4);                                   // land while this batch is emitted
            for (uint32_t pa = (n0 & ~127u) + 1024u + (uint32_t)lane * 128u; pa < n4; pa += 32u * 128u)    // the rest: into L2
                asm volatile("prefetch.global.L2 [%0];" :: "l"(a.s.pk8 + pa));
        }
        // ---- emit: one site at a time (four independent chains for the scheduler), lane = dictionary index.  Lanes 30 / 31 (the
        // indel codes) read a block of zeros, so no count needs masking.
#pragma unroll
        for (int k = 0; k < kH8Sites; ++k) {
            const int site = s0 + k;
            if (site >= S) break;                               // warp-uniform
            uint32_t* h = hw + k * kH8Words;
            // bin = the byte: base << 6 | qi << 1 | strand -> my (reverse, forward) counts of base b are one 8-byte word
            const uint2* hp = reinterpret_cast<const uint2*>(counted ? h + lane * 2 : zero8);
            const int hstep = counted ? 32 : 0;                 // uint2 stride between bases
            const uint2 h0 = hp[0], h1 = hp[hstep], h2 = hp[2 * hstep], h3 = hp[3 * hstep];
            const uint4 rv = make_uint4(h0.x, h1.x, h2.x, h3.x), fw = make_uint4(h0.y, h1.y, h2.y, h3.y);
            const int n_entries = (int)(k == 0 ? o1 - o0 : k == 1 ? o2 - o1 : k == 2 ? o3 - o2 : o4 - o3);
            if constexpr (kDense) {
                const bool mine = ((h0.x | h0.y) | (h1.x | h1.y) | (h2.x | h2.y) | (h3.x | h3.y)) != 0;
                const bool in_win = lane < nd && (int)dq >= dn.qlo && (int)dq <= dn.qhi;
                if (mine && !in_win) err |= 1u;                 // a dictionary index >= n_qual_dict, or a qual outside the exchange window
                uint32_t* dst = dn.dense + (size_t)site * dn.B;
                // header: the 8 indel bins in K1w's order (strand << 2 | stale base), then n_entries
                const uint32_t hw8 = lane < 8 ? h[BVC_PACK8_INDEL * 2 + 64 * (lane & 3) + (lane >> 2)] : (uint32_t)n_entries;
                __syncwarp();
                if (counted) {
                    uint2* cp = reinterpret_cast<uint2*>(h + lane * 2);
                    cp[0] = make_uint2(0u, 0u); cp[32] = make_uint2(0u, 0u); cp[64] = make_uint2(0u, 0u); cp[96] = make_uint2(0u, 0u);
                }
                if (lane < 4) *reinterpret_cast<uint2*>(h + BVC_PACK8_INDEL * 2 + 64 * lane) = make_uint2(0u, 0u);
                if (lane < 9) dst[lane] = hw8;
                if (mine && in_win) {
                    const int half = dn.qhi - dn.qlo + 1;       // uint2 words per base
                    uint2* bp = reinterpret_cast<uint2*>(dst + kShardHdr) + ((int)dq - dn.qlo);
                    bp[0] = h0; bp[half] = h1; bp[2 * half] = h2; bp[3 * half] = h3;
                }
                __syncwarp();
                continue;
            }
            const int ref = (int)(__shfl_sync(0xffffffffu, refv, k) & 3u);
            const uint32_t x0 = rv.x + fw.x, x1 = rv.y + fw.y, x2 = rv.z + fw.z, x3 = rv.w + fw.w;
            // a site has < 65,536 entries: two depths per word through one REDUX
            const uint32_t d01 = __reduce_add_sync(0xffffffffu, __byte_perm(x0, x1, 0x5410)), d23 = __reduce_add_sync(0xffffffffu, __byte_perm(x2, x3, 0x5410));
            const uint32_t f01 = __reduce_add_sync(0xffffffffu, __byte_perm(fw.x, fw.y, 0x5410)),
                           f23 = __reduce_add_sync(0xffffffffu, __byte_perm(fw.z, fw.w, 0x5410));
            const unsigned m0 = __ballot_sync(0xffffffffu, x0 != 0), m1 = __ballot_sync(0xffffffffu, x1 != 0),
                           m2 = __ballot_sync(0xffffffffu, x2 != 0), m3 = __ballot_sync(0xffffffffu, x3 != 0);
            // (the indel bins -- qi 31, read below by every lane -- are cleared by lane 31 after that read; qi 30 is never read)
            if (counted) {
                uint2* cp = reinterpret_cast<uint2*>(h + lane * 2);
                cp[0] = make_uint2(0u, 0u); cp[32] = make_uint2(0u, 0u); cp[64] = make_uint2(0u, 0u); cp[96] = make_uint2(0u, 0u);
            }
            const unsigned many = m0 | m1 | m2 | m3;
            if (many >> nd) err |= kErrQual;                    // a counted entry with a dictionary index >= n_qual_dict
            const unsigned lowq = ((many & q0m) ? 1u : 0u) | ((many & q1m) ? 2u : 0u);
            const int d0 = (int)(d01 & 0xffffu), d1 = (int)(d01 >> 16), d2 = (int)(d23 & 0xffffu), d3 = (int)(d23 >> 16);
            const int f0 = (int)(f01 & 0xffffu), f1 = (int)(f01 >> 16), f2 = (int)(f23 & 0xffffu), f3 = (int)(f23 >> 16);
            const int n0 = __popc(m0), n1 = __popc(m1), n2 = __popc(m2), n3 = __popc(m3);
            // src/BaseVarC.cpp:576-583: descending argsort, ties in index order -> key = depth << 2 | (3 - base); top non-ref base
            const uint32_t k0 = ((uint32_t)d0 << 2) | 3u, k1 = ((uint32_t)d1 << 2) | 2u, k2 = ((uint32_t)d2 << 2) | 1u, k3 = (uint32_t)d3 << 2;
            const uint32_t kb = max(max(k0, k1), max(k2, k3));
            const int best = 3 - (int)(kb & 3u);
            const uint32_t ks = max(max(best == 0 ? 0u : k0, best == 1 ? 0u : k1), max(best == 2 ? 0u : k2, best == 3 ? 0u : k3));
            const int second = 3 - (int)(ks & 3u);
            const int alt = (best != ref) ? best : second;
            // src/BaseVarC.cpp:587-602: 2x2 table over ALL entries; an indel entry (qi 31) carries its stale base / strand
            const int df_r = ref == 0 ? f0 : ref == 1 ? f1 : ref == 2 ? f2 : f3, dd_r = ref == 0 ? d0 : ref == 1 ? d1 : ref == 2 ? d2 : d3;
            const int df_a = alt == 0 ? f0 : alt == 1 ? f1 : alt == 2 ? f2 : f3, dd_a = alt == 0 ? d0 : alt == 1 ? d1 : alt == 2 ? d2 : d3;
            uint2* ibin = reinterpret_cast<uint2*>(h + BVC_PACK8_INDEL * 2);       // + 32 per base: (reverse, forward) indel entries
            const uint2 i_r = ibin[32 * ref], i_a = ibin[32 * alt];
            const int c0 = df_r + (int)i_r.y, c1 = (dd_r - df_r) + (int)i_r.x, c2 = df_a + (int)i_a.y, c3 = (dd_a - df_a) + (int)i_a.x;
            // pool segments: the MAJOR base first (K3 runs that segment select-free), null-padded to a multiple of 4, then the others
            // ascending; a class's position in its segment = number of present classes with a lower qual
            const int n_major = best == 0 ? n0 : best == 1 ? n1 : best == 2 ? n2 : n3;
            int run = pad4(n_major);
            const int g0 = best == 0 ? 0 : run; run += best == 0 ? 0 : n0;
            const int g1 = best == 1 ? 0 : run; run += best == 1 ? 0 : n1;
            const int g2 = best == 2 ? 0 : run; run += best == 2 ? 0 : n2;
            const int g3 = best == 3 ? 0 : run; run += best == 3 ? 0 : n3;
            uint32_t* st = stage + k * kH8Stage;
            if (lane < 3 && n_major + lane < pad4(n_major)) st[n_major + lane] = kNullClass;
            if (x0) st[g0 + __popc(m0 & lower)] = (x0 << 9) | dq;
            if (x1) st[g1 + __popc(m1 & lower)] = (x1 << 9) | (1u << 7) | dq;
            if (x2) st[g2 + __popc(m2 & lower)] = (x2 << 9) | (2u << 7) | dq;
            if (x3) st[g3 + __popc(m3 & lower)] = (x3 << 9) | (3u << 7) | dq;
            const uint32_t off = (uint32_t)site * (uint32_t)a.cls_stride;
            if (lane == 0) {                                    // the 64-byte summary: every lane holds the values, one stores them
                uint4* sp = reinterpret_cast<uint4*>(a.sum + site);
                sp[0] = make_uint4((uint32_t)d0, (uint32_t)d1, (uint32_t)d2, (uint32_t)d3);
                sp[1] = make_uint4((uint32_t)f0, (uint32_t)f1, (uint32_t)f2, (uint32_t)f3);
                sp[2] = make_uint4((uint32_t)c0, (uint32_t)c1, (uint32_t)c2, (uint32_t)c3);
                sp[3] = make_uint4(off, (uint32_t)((n0 + n1 + n2 + n3) | (alt << 16) | ((lowq & 1u) ? (1 << 24) : 0) | (lowq ? (2 << 24) : 0)),
                                   (uint32_t)n_entries, (uint32_t)(n_major | (best << 16)));
            }
            __syncwarp();
            if (lane < 4) ibin[32 * lane] = make_uint2(0u, 0u);
            if (lane * 4 < run)                                 // run <= 4 * nd + 3 <= 123 words: one coalesced 16-byte store per four
                reinterpret_cast<uint4*>(a.cls + off)[lane] = reinterpret_cast<const uint4*>(st)[lane];
        }
        ov = nov; refv = nrefv;
        __syncwarp();
    }
    const uint32_t err_all = __reduce_or_sync(0xffffffffu, err);
    if (lane == 0 && err_all) atomicOr(kDense ? dn.err : a.s.err, err_all);
}

__global__ void __launch_bounds__(kHistWarps * 32, BVC_H8_MINBLOCKS)
site_hist8_kernel(PipeArgs a)
{
    site_hist8_body<false>(a, DenseArgs{nullptr, nullptr, 0, 0, 0});
}

// sample-sharded mode, phase A on the one-byte form (a.s: the LOCAL tile -- entry_off, n_sites, pk8, dict)
__global__ void __launch_bounds__(kHistWarps * 32, BVC_H8_MINBLOCKS)
shard_hist8_kernel(PipeArgs a, DenseArgs dn)
{
    site_hist8_body<true>(a, dn);
}

// =====================================================================================================
// shared: queue append, finalisation of a site's LRT
// =====================================================================================================
__device__ __forceinline__ uint32_t warp_agg_append(uint32_t* ctr, bool pred)
{
    const unsigned active = __activemask();
    const unsigned m = __ballot_sync(active, pred);
    uint32_t base = 0;
    if (!pred) return 0;
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    if (lane == leader) base = atomicAdd(ctr, (uint32_t)__popc(m));
    base = __shfl_sync(m, base, leader);
    return base + __popc(m & ((1u << lane) - 1));
}

// The 176-byte record as eleven 16-byte stores (offsets: include/bvc_gpu.h; a record is 16-byte aligned in d_out).
__device__ __forceinline__ void st_dd(bvc_site_result* r, int off, double x, double y)
{
    *reinterpret_cast<double2*>(reinterpret_cast<char*>(r) + off) = make_double2(x, y);
}
__device__ __forceinline__ void st_iiii(bvc_site_result* r, int off, int x, int y, int z, int w)
{
    *reinterpret_cast<int4*>(reinterpret_cast<char*>(r) + off) = make_int4(x, y, z, w);
}

// BaseType.cpp:111-135; writes the WHOLE record of the site once (K5 later fills the VCF statistics of the
// LRT-positive sites).  su / cvg_fs / cvg_sor: the CVG half computed by K1 / K2.
// Everything arrives by value and no array is indexed with a run-time subscript: the function inlines into K2 / K4 without
// a stack frame (as a call taking pointers it cost those kernels ~0.5 KB of local-memory traffic per site -- more DRAM
// write traffic than the records themselves).
__device__ __forceinline__ int cand_at(uint32_t cand4, int t) { return (int)((cand4 >> (8 * t)) & 0xffu); }

__device__ __forceinline__ void finalize_site(const PipeArgs& a, int site, int ref, const SiteSum& su, double cvg_fs, double cvg_sor,
                                              uint32_t cand4, int n, double frq0, double frq1, double frq2, double frq3, double chi_t,
                                              int em_fits, int em_passes, uint32_t flags)
{
    bvc_site_result* r = a.s.out + site;
    const int* depth = su.depth;
    const int n_reads = depth[0] + depth[1] + depth[2] + depth[3];
    int n_alt = 0, alt0 = -1, alt1 = -1, alt2 = -1;
    double af0 = 0.0, af1 = 0.0, af2 = 0.0, af3 = 0.0;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int b = cand_at(cand4, t);
        if (t < n && b != ref && n_alt < 3) {
            const double fv = sel4(b, frq0, frq1, frq2, frq3);
            if (n_alt == 0) alt0 = b; else if (n_alt == 1) alt1 = b; else alt2 = b;
            ++n_alt;
            if (b == 0) af0 = fv; else if (b == 1) af1 = fv; else if (b == 2) af2 = fv; else af3 = fv;
        }
    }
    double vq = 0.0;
    const int ok = n_alt > 0;
    if (ok) {
        const double total = (double)n_reads;
        const int b0 = cand_at(cand4, 0);
        const double rr = (double)(b0 == 0 ? depth[0] : b0 == 1 ? depth[1] : b0 == 2 ? depth[2] : depth[3]) / total;
        if (n == 1 && total > 10 && rr > 0.5) vq = 5000.0;
        else if (chi_t <= 0) vq = 0.0;
        else {
            const double p = dm_chisf(chi_t, 1.0);
            vq = (p != 0) ? -10.0 * log10(p) : 10000.0;
            if (vq == 0) vq = 0.0;
        }
        if (near_rel(vq, 60.0)) flags |= BVC_FLAG_QUAL60;
    }
    st_dd(r, 0, cvg_fs, cvg_sor);
    st_dd(r, 16, af0, af1);
    st_dd(r, 32, af2, af3);
    st_dd(r, 48, chi_t, vq);
    st_dd(r, 64, 0.0, 0.0);                                   // vcf_fs, vcf_sor
    st_dd(r, 80, 0.0, 0.0);                                   // ranksum[0..1]
    st_iiii(r, 96, 0, 0, su.n_entries, depth[0]);             // ranksum[2] = 0.0
    st_iiii(r, 112, depth[1], depth[2], depth[3], su.cvg[0]);
    st_iiii(r, 128, su.cvg[1], su.cvg[2], su.cvg[3], 0);
    st_iiii(r, 144, 0, 0, 0, em_fits);
    st_iiii(r, 160, em_passes, (int)flags,
            (int)((uint32_t)su.alt | ((uint32_t)ok << 8) | ((uint32_t)n_alt << 16) | (((uint32_t)alt0 & 0xffu) << 24)),
            (int)(((uint32_t)alt1 & 0xffu) | (((uint32_t)alt2 & 0xffu) << 8)));
    a.state[site] = kStDone;
    if (ok) a.vcf_list[atomicAdd(a.vcf_n, 1u)] = (uint32_t)site;
}

// =====================================================================================================
// K2: thread per site -- CVG Fisher/SOR, candidate filter, emit the full-model fit
// =====================================================================================================
constexpr int kDeepFisher = 48;            // tables with more terms than this are evaluated by a whole warp
__device__ __forceinline__ int fisher_range(int n11, int n12, int n21, int n22)
{
    const int n1_ = n11 + n12, n_1 = n11 + n21, n = n1_ + n21 + n22;
    const int hi = n_1 < n1_ ? n_1 : n1_, lo = n1_ + n_1 - n;
    return hi - (lo < 0 ? 0 : lo);
}

// warp per listed site: the CVG Fisher test K2 skipped.  The value lands where K2 would have put it: in the record of
// a site K2 already finalised, else in the site's Ladder (K4 stores the record).
__global__ void __launch_bounds__(128)
cvg_fisher_deep_kernel(PipeArgs a)
{
    const int lane = threadIdx.x & 31;
    const uint32_t n = *a.deep_n;
    for (uint32_t i = blockIdx.x * 4 + (threadIdx.x >> 5); i < n; i += gridDim.x * 4) {
        const int site = (int)a.deep_list[i];
        const SiteSum su = a.sum[site];
        const double fs = dm_fisher_warp(a.s.lgam, su.cvg[0], su.cvg[1], su.cvg[2], su.cvg[3], lane);
        if (lane == 0) {
            if (a.state[site] == kStDone) a.s.out[site].cvg_fs = fs; else a.lad[site].cvg_fs = fs;
        }
    }
}

#ifndef BVC_PREP_MINBLOCKS
#define BVC_PREP_MINBLOCKS 1                 // 72 registers; 8 blocks per SM (64 registers) measured: 0.137 against 0.134 ms, no gain
#endif
__global__ void __launch_bounds__(128, BVC_PREP_MINBLOCKS)
site_prepare_kernel(PipeArgs a)
{
    const int site = blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = site < a.s.n_sites;
    bool emit = false;
    if (in) {
        const SiteSum su = a.sum[site];
        const int ref = a.s.ref_base[site] & 3;
        const int n_reads = su.depth[0] + su.depth[1] + su.depth[2] + su.depth[3];
        if (su.has_q0 & 1) {
            a.state[site] = kStLegacy;
            a.legacy_list[atomicAdd(a.legacy_n, 1u)] = (uint32_t)site;
        } else {
            // kt_fisher_exact walks up to (max n11 - min n11) terms: a deep table would make this thread the kernel's
            // tail (C4: thousands of terms at a common variant), so it goes to cvg_fisher_deep_kernel (warp per site)
            double cvg_fs = 0.0;
            if (fisher_range(su.cvg[0], su.cvg[1], su.cvg[2], su.cvg[3]) > kDeepFisher) a.deep_list[atomicAdd(a.deep_n, 1u)] = (uint32_t)site;
            else cvg_fs = dm_fisher(a.s.lgam, su.cvg[0], su.cvg[1], su.cvg[2], su.cvg[3]);
            const double cvg_sor = dm_sor(su.cvg[0], su.cvg[1], su.cvg[2], su.cvg[3]);
            // BaseType.cpp:75-82
            uint32_t cand4 = 0xffffffffu; int n = 0; uint32_t flags = 0, full = 0;
            const double total = (double)n_reads;
            if (n_reads > 0) {
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const double fr = (double)su.depth[b] / total;
                    if (near_rel(fr, a.s.min_af)) flags |= BVC_FLAG_MIN_AF;
                    if (fr >= a.s.min_af) { cand4 = (cand4 & ~(0xffu << (8 * n))) | ((uint32_t)b << (8 * n)); ++n; full |= 1u << b; }
                }
            }
            if (n <= 1) {
                // no EM needed: 0 candidates -> false; 1 candidate -> one (closed-form) fit, chi stays 0
                finalize_site(a, site, ref, su, cvg_fs, cvg_sor, cand4, n, full & 1 ? 1.0 : 0.0, full & 2 ? 1.0 : 0.0, full & 4 ? 1.0 : 0.0,
                              full & 8 ? 1.0 : 0.0, 0.0, n, 2 * n, flags);
            } else {
                Ladder L;
                L.lr_alt = 0.0; L.frq[0] = L.frq[1] = L.frq[2] = L.frq[3] = 0.0; L.chi_t = 0.0;
                L.cvg_fs = cvg_fs; L.cvg_sor = cvg_sor;
                L.cand4 = cand4;
                L.n = (int8_t)n; L.k = (int8_t)n; L.nfit = 1; L.pad0 = 0;
                L.em_fits = 0; L.em_passes = 0; L.flags = flags; L.pad1 = 0;
                a.lad[site] = L;
                a.fit[(size_t)site * 4].mask = full;
                a.state[site] = kStActive;
                emit = true;
            }
        }
    }
    const uint32_t pos = warp_agg_append(a.q_out_n, emit);
    if (emit) a.q_out[pos] = ((uint32_t)site << 2);
}

// =====================================================================================================
// Qs: counting sort of a stage's fit queue by DESCENDING class count.  Lanes of a warp then hold fits of
// (nearly) equal length, so K3's class loop has a warp-uniform trip count and every pass ends for all
// lanes together; long fits start first (LPT order).
// =====================================================================================================
constexpr int kSortThreads = 256;
constexpr int kSortItems = 8;              // items per thread per chunk

__global__ void __launch_bounds__(kSortThreads)
queue_hist_kernel(PipeArgs a)
{
    __shared__ uint32_t h[kNBins + 1];
    for (int i = threadIdx.x; i <= kNBins; i += blockDim.x) h[i] = 0;
    __syncthreads();
    const uint32_t n = *a.q_in_n;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        atomicAdd(&h[a.sum[a.q_in[i] >> 2].ncls], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i <= kNBins; i += blockDim.x) if (h[i]) atomicAdd(&a.bins[i], h[i]);
}

// one block: bins[k] <- number of fits with MORE than k classes (exclusive scan in descending key order)
__global__ void __launch_bounds__(kSortThreads)
queue_scan_kernel(PipeArgs a)
{
    __shared__ uint32_t h[kNBins + 1];
    for (int i = threadIdx.x; i <= kNBins; i += blockDim.x) h[i] = a.bins[i];
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (int k = kNBins; k >= 0; --k) { const uint32_t c = h[k]; h[k] = run; run += c; }
    }
    __syncthreads();
    for (int i = threadIdx.x; i <= kNBins; i += blockDim.x) a.bins[i] = h[i];
}

__global__ void __launch_bounds__(kSortThreads)
queue_scatter_kernel(PipeArgs a)
{
    __shared__ uint32_t h[kNBins + 1], base[kNBins + 1];
    const uint32_t n = *a.q_in_n;
    constexpr uint32_t kChunk = kSortThreads * kSortItems;
    for (uint32_t c0 = blockIdx.x * kChunk; c0 < n; c0 += gridDim.x * kChunk) {
        for (int i = threadIdx.x; i <= kNBins; i += blockDim.x) h[i] = 0;
        __syncthreads();
        uint32_t item[kSortItems], rank[kSortItems]; uint16_t key[kSortItems];
#pragma unroll
        for (int k = 0; k < kSortItems; ++k) {
            const uint32_t i = c0 + k * kSortThreads + threadIdx.x;
            if (i < n) {
                item[k] = a.q_in[i];
                key[k] = a.sum[item[k] >> 2].ncls;
                rank[k] = atomicAdd(&h[key[k]], 1u);
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i <= kNBins; i += blockDim.x) if (h[i]) base[i] = atomicAdd(&a.bins[i], h[i]);
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kSortItems; ++k) {
            const uint32_t i = c0 + k * kSortThreads + threadIdx.x;
            if (i < n) a.q_sorted[base[key[k]] + rank[k]] = item[k];
        }
        __syncthreads();
    }
}

// =====================================================================================================
// K3: thread per fit, pass-aligned within a warp
// =====================================================================================================
constexpr int kEmThreads = 128;
// hand-over to K3b: PipeArgs::pass_budget (passes a fit may spend in K3) and PipeArgs::drain_min_pass (once the queue
// is empty, fits at or beyond this pass move to K3b); defaults in bvc_ctx.h
#ifndef BVC_LONG_GROUP
#define BVC_LONG_GROUP 8                   // measured with the bound-based passes: 4 lanes per fit run C3's em_long in the same 0.80 ms, 2 in 1.01 ms
#endif
constexpr int kLongGroup = BVC_LONG_GROUP; // lanes per fit in K3b

// 1/x for normal positive x: MUFU.RCP64H seed (relative error e <= 2^-23, PTX rcp.approx.ftz.f64) + ONE cubic step
// r (1 + e + e^2), e = 1 - x r: remaining error e^3 <= 2^-69, i.e. correctly rounded up to the last fma (<= 1 ulp)
__device__ __forceinline__ double rcp_nr2(double x)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    const double e = fma(-x, r, 1.0);
    const double t = fma(e, e, e);
    return fma(r, t, r);
}

// log1p(x) for |x| <= 2^-9: x * (1 - x/2 + x^2/3 - x^3/4 + x^4/5), truncation < |x|^5/6 <= 5e-15 relative.
// A class with |x| >= 2^-9 alone contributes cnt*|log1p(x)| >= |x|/(1+|x|) > 1.9e-3 > epsilon (= 0.001,
// enforced <= 0.0019 at bvc_create), so that pass is "not converged" without evaluating the sum exactly.
__device__ __forceinline__ double log1p_tiny(double x)
{
    double p = 1.0 / 5.0;
    p = fma(p, x, -1.0 / 4.0);
    p = fma(p, x, 1.0 / 3.0);
    p = fma(p, x, -1.0 / 2.0);
    p = fma(p, x, 1.0);
    return p * x;
}

// exact double of a 23-bit count without an I2F: 2^52 + cnt has cnt in its low mantissa bits
__device__ __forceinline__ double cnt_to_double(uint32_t cnt)
{
    return __hiloint2double(0x43300000, (int)cnt) - 4503599627370496.0;
}

struct EmAcc { double accT, V0, V1, V2, V3, dl; int xmax; };   // xmax: max over classes of the high word of |x|
constexpr int kBigXHi = 0x3f600000;       // high word of 2^-9: |x| >= 2^-9  <=>  hi(|x|) >= kBigXHi

// Branch-free 4-way selects by the class's base bits (bit 7 / bit 8 of the class word): nvcc turns the
// equivalent C ternaries into divergent branches, which serialises lanes whose classes differ in base.
__device__ __forceinline__ double sel4_w(uint32_t w, double f0, double f1, double f2, double f3)
{
    double fb;
    asm("{\n\t.reg .pred o, t;\n\t.reg .f64 lo, hi;\n\t"
        "setp.ne.u32 o, %5, 0;\n\tsetp.ne.u32 t, %6, 0;\n\t"
        "selp.f64 lo, %2, %1, o;\n\tselp.f64 hi, %4, %3, o;\n\tselp.f64 %0, hi, lo, t;\n\t}"
        : "=d"(fb) : "d"(f0), "d"(f1), "d"(f2), "d"(f3), "r"(w & 0x80u), "r"(w & 0x100u));
    return fb;
}
// V_b += v for the class's base b: four predicated adds, no branch
__device__ __forceinline__ void acc_by_base(uint32_t w, double v, double& V0, double& V1, double& V2, double& V3)
{
    asm("{\n\t.reg .pred q0, q1, q2, q3;\n\t"
        "setp.eq.u32 q0, %5, 0;\n\tsetp.eq.u32 q1, %5, 0x80;\n\tsetp.eq.u32 q2, %5, 0x100;\n\tsetp.eq.u32 q3, %5, 0x180;\n\t"
        "@q0 add.f64 %0, %0, %4;\n\t@q1 add.f64 %1, %1, %4;\n\t@q2 add.f64 %2, %2, %4;\n\t@q3 add.f64 %3, %3, %4;\n\t}"
        : "+d"(V0), "+d"(V1), "+d"(V2), "+d"(V3)
        : "d"(v), "r"(w & 0x180u));
}

// ---- K3's class bodies: E and M step only (Algorithm.cpp:76-91) in RATIO form.  With d = (1-e) - e/3 > 0 (qual >= 2;
// fits with a qual <= 1 class run in K3b) and t = (e/3) / d, a class's marginal is ml = d (f_b + t F), so
//   w = cnt / (f_b + t F)   gives   d cnt / ml = w   (the class's share of V_b)   and   (e/3) cnt / ml = t w   (of accT):
// one 8-byte table word and 7 FP64 instructions per class instead of a 16-byte word and 10.
// 1/x: MUFU.RCP64H seed (relative error <= 2^-23, PTX rcp.approx.ftz.f64) + one Newton step -> <= 2^-46 (1.4e-14): five
// orders below the parity tolerance; decisions within that distance of epsilon are K3b's (delta_decision's band).
__device__ __forceinline__ double rcp_nr1(double x)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    return fma(r, fma(-x, r, 1.0), r);
}
// major segment: every class has the lane's major base -> fb and the V accumulator are plain registers
__device__ __forceinline__ void em_class_major(uint32_t w, const double* lutT, double fb, double F, double& accT, double& Vm)
{
    const double t = lutT[w & (kNQ - 1)];
    const double wgt = cnt_to_double(w >> 9) * rcp_nr1(fma(t, F, fb));
    Vm += wgt;
    accT = fma(t, wgt, accT);
}
// other bases: select by the class word
__device__ __forceinline__ void em_class_minor(uint32_t w, const double* lutT, double f0, double f1, double f2, double f3, double F,
                                               double& accT, double& V0, double& V1, double& V2, double& V3)
{
    const double t = lutT[w & (kNQ - 1)];
    const double wgt = cnt_to_double(w >> 9) * rcp_nr1(fma(t, F, sel4_w(w, f0, f1, f2, f3)));
    accT = fma(t, wgt, accT);
    acc_by_base(w, wgt, V0, V1, V2, V3);
}

// ---- K3b's class body: E/M step plus the exact, stateless delta term (Algorithm.cpp:103-113) against the
// previous pass's frequencies p (pass 0: p == f -> x == 0)
__device__ __forceinline__ void em_class_exact(uint32_t w, const double2* lut2, double f0, double f1, double f2, double f3, double F,
                                               double p0, double p1, double p2, double p3, double Fp, EmAcc& A)
{
    const double cnt = cnt_to_double(w >> 9);
    const double2 ed = lut2[w & (kNQ - 1)];
    const double ml = fma(sel4_w(w, f0, f1, f2, f3), ed.y, ed.x * F);
    const double r = rcp_nr2(ml);
    const double wgt = cnt * r;
    A.accT = fma(ed.x, wgt, A.accT);
    acc_by_base(w, ed.y * wgt, A.V0, A.V1, A.V2, A.V3);
    const double mlp = fma(sel4_w(w, p0, p1, p2, p3), ed.y, ed.x * Fp);
    const double x = (mlp - ml) * r;
    A.dl = fma(cnt, fabs(log1p_tiny(x)), A.dl);
    // range test on the high word (integer pipe); count-0 padding classes never trip it
    A.xmax = max(A.xmax, (w >> 9) ? (__double2hiint(x) & 0x7fffffff) : 0);
}

// ---- K3b's class body without the delta term: the SAME E/M arithmetic as em_class_exact (so a fit's numbers do not depend on
// which body ran a pass); the pass is then decided by delta_decision's bound, and repeated with em_class_exact only if that is ambiguous
__device__ __forceinline__ void em_class_plain(uint32_t w, const double2* lut2, double f0, double f1, double f2, double f3, double F, EmAcc& A)
{
    const double cnt = cnt_to_double(w >> 9);
    const double2 ed = lut2[w & (kNQ - 1)];
    const double ml = fma(sel4_w(w, f0, f1, f2, f3), ed.y, ed.x * F);
    const double wgt = cnt * rcp_nr2(ml);
    A.accT = fma(ed.x, wgt, A.accT);
    acc_by_base(w, ed.y * wgt, A.V0, A.V1, A.V2, A.V3);
}

__device__ __forceinline__ void load_lut2(double2* lut2, const double* lut)
{
    for (int i = threadIdx.x; i < kNQ; i += blockDim.x) lut2[i] = make_double2(lut[i], lut[4 * kNQ + i]);
    __syncthreads();
}

__device__ __forceinline__ void add_work(unsigned long long* ctr, unsigned long long v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(ctr, v);
}

// Convergence test of K3 without per-class work.  With x_c = (ml_c(p) - ml_c(f)) / ml_c(f) =
// (d_c (p_b - f_b) + e3_c (Fp - F)) / ml_c(f), all classes of a base move the same way, so
//   P = sum_b |p_b - f_b| V_b,  V_b = sum_{c in b} cnt_c d_c / ml_c(f)   (already accumulated by the M step)
// equals sum_c cnt_c |x_c| up to the (Fp - F) ~ 1e-16 term, and delta = sum_c cnt_c |log1p(x_c)| lies in
// [P / (1 + P), P / (1 - P)] because every |x_c| <= cnt_c |x_c| <= P.  Outside a +-0.1 % band around
// epsilon the decision delta < epsilon (Algorithm.cpp:124) is therefore certain; inside it the fit is
// parked and K3b repeats the pass with the exact sum.   returns 0: not converged, 1: converged, 2: ambiguous
__device__ __forceinline__ int delta_decision(double P, double eps)
{
    if (!(P < 0.25)) return 0;
    if (P > eps * (1.0 + P) * (1.0 + 1e-6)) return 0;            // P / (1 + P) > eps
    if (P < eps * (1.0 - P) * (1.0 - 1e-6)) return 1;            // P / (1 - P) < eps
    return 2;
}

__global__ void __launch_bounds__(kEmThreads)
em_fit_kernel(PipeArgs a)
{
    __shared__ double lutT[kNQ];                                  // t(q) = (e/3) / ((1-e) - e/3)
    for (int i = threadIdx.x; i < kNQ; i += blockDim.x) lutT[i] = a.s.lut[i] / a.s.lut[4 * kNQ + i];
    __syncthreads();
    const uint32_t nfits = *a.q_in_n;
    const int lane = threadIdx.x & 31;
    const double eps = a.s.eps;
    const int max_iter = a.s.max_iter;
    // A stage with fewer fits than K3 has lanes gains little from thread-per-fit: K3b runs them 8 lanes each with a shorter
    // pass.  (Measured with the threshold at 16k fits instead: C3 -5 %, C2 +0.6 %, C5 +4 % per step -- PipeArgs::em_min_fits.)
    const uint32_t min_fits = a.em_min_fits > 0 ? (uint32_t)a.em_min_fits : (uint32_t)(gridDim.x * blockDim.x);
    const int budget = (nfits >= min_fits) ? a.pass_budget : 0;

    bool active = false, done = false, drain = false, lowq = false;
    const uint32_t* cp = a.cls;
    int nmaj = 0, nmin = 0, pass = 0, major = 0;
    uint32_t item = 0;
    double f0 = 0, f1 = 0, f2 = 0, f3 = 0, p0 = 0, p1 = 0, p2 = 0, p3 = 0, F = 0, inv_n = 0;
    unsigned long long work = 0, nfit = 0;

    for (;;) {
        // ---- refill at the pass boundary: lanes without a fit take the next items (one atomic per warp) ---
        const bool want = !active && !done;
        unsigned wm = __ballot_sync(0xffffffffu, want);
        // The refill path (~150 warp instructions and a chain of dependent loads) is paid by the whole warp whenever ANY lane
        // takes a fit, an idle lane costs 1/32 of a pass: lanes wait until refill_min of them want one (or none holds a fit)
        if (wm && __popc(wm) < a.refill_min && !drain && __any_sync(0xffffffffu, active)) wm = 0;
        if (wm) {
            uint32_t base = 0;
            const int leader = __ffs(wm) - 1;
            if (lane == leader) base = atomicAdd(a.q_in_head, (uint32_t)__popc(wm));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (want) {
                const uint32_t idx = base + __popc(wm & ((1u << lane) - 1));
                if (idx < nfits) {
                    item = a.q_sorted[idx];
                    const uint32_t site = item >> 2;
                    const SiteSum su = a.sum[site];
                    const uint32_t mask = a.fit[(size_t)site * 4 + (item & 3)].mask;
                    int dsum = 0;                                   // BaseType.cpp:25-39
                    if (mask & 1) dsum += su.depth[0];
                    if (mask & 2) dsum += su.depth[1];
                    if (mask & 4) dsum += su.depth[2];
                    if (mask & 8) dsum += su.depth[3];
                    // start vector depth_b / sum (BaseType.cpp:25-39) as depth_b * (1 / sum): within 1 ulp of the quotient, and
                    // one IEEE division per fit instead of four (each ~25 instructions of the warp's refill path)
                    const int ntot = su.depth[0] + su.depth[1] + su.depth[2] + su.depth[3];
                    inv_n = 1.0 / (double)ntot;
                    const double rds = dsum == ntot ? inv_n : 1.0 / (double)dsum;
                    f0 = (mask & 1) ? (double)su.depth[0] * rds : 0.0;
                    f1 = (mask & 2) ? (double)su.depth[1] * rds : 0.0;
                    f2 = (mask & 4) ? (double)su.depth[2] * rds : 0.0;
                    f3 = (mask & 8) ? (double)su.depth[3] * rds : 0.0;
                    F = ((f0 + f1) + f2) + f3;
                    p0 = f0; p1 = f1; p2 = f2; p3 = f3;
                    cp = a.cls + su.cls_off; nmaj = su.n_major; nmin = su.ncls - su.n_major; major = su.major;
                    pass = 0; lowq = (su.has_q0 & 2) != 0;
                    active = true; ++nfit;
                } else done = true;
            }
        }
        if (!__any_sync(0xffffffffu, active)) break;
        // the queue is empty: nothing is left to overlap a long fit with, so finish what the lanes hold in K3b
        drain = drain || __any_sync(0xffffffffu, done);
        // ... but only once this warp has thinned out: while most of its lanes still hold a fit, thread-per-fit is the
        // cheaper way to run their passes (C3: 1.6 fits per lane, the queue is empty after the first refill)
        const bool thin = drain && __popc(__ballot_sync(0xffffffffu, active)) <= a.drain_lanes;

        // a class with qual <= 1 has (1-e) - e/3 < 0, which voids delta_decision's sign argument: exact path
        if (active && (budget == 0 || lowq)) {
            FitRes* fr = a.fit + (size_t)(item >> 2) * 4 + (item & 3);
            fr->f[0] = f0; fr->f[1] = f1; fr->f[2] = f2; fr->f[3] = f3;
            fr->eap[0] = p0; fr->eap[1] = p1; fr->eap[2] = p2; fr->eap[3] = p3;
            fr->passes = 0; fr->flags = 0; fr->unfinished = 1;
            a.q_long[atomicAdd(a.q_long_n, 1u)] = item;
            active = false;
        }
        if (__any_sync(0xffffffffu, active)) {
            const int my_maj4 = active ? (nmaj + 3) >> 2 : 0, my_min = active ? nmin : 0;
            const int nmaj4_max = __reduce_max_sync(0xffffffffu, my_maj4);
            const int nmin_max = __reduce_max_sync(0xffffffffu, my_min);
            // ---- one pass.  Straight-line bodies: a lane past its last class (or without a fit) processes
            // count-0 null classes, so the unrolled bodies interleave (ILP) instead of sitting behind branches.
            // The major segment is read four classes per load (16-byte aligned, null-padded by K1).
            double accT = 0.0, Vm = 0.0, V0 = 0.0, V1 = 0.0, V2 = 0.0, V3 = 0.0;
            const double fm = sel4(major, f0, f1, f2, f3);
            const uint4* cp4 = reinterpret_cast<const uint4*>(cp);
            const uint32_t* cmin = cp + 4 * my_maj4;
            // class words are loaded one trip ahead of their use (a lane's list is its own: the loads do not coalesce and
            // 1 in 5 misses L1, so an unprefetched load is a long-scoreboard stall at the first use of the word)
            const uint4 null4 = make_uint4(kNullClass, kNullClass, kNullClass, kNullClass);
            uint4 wn = (0 < my_maj4) ? __ldg(cp4) : null4;
            uint32_t mn = (0 < my_min) ? __ldg(cmin) : kNullClass;
            for (int c = 0; c < nmaj4_max; ++c) {
                const uint4 w = wn;
                wn = (c + 1 < my_maj4) ? __ldg(cp4 + c + 1) : null4;
                em_class_major(w.x, lutT, fm, F, accT, Vm);
                em_class_major(w.y, lutT, fm, F, accT, Vm);
                em_class_major(w.z, lutT, fm, F, accT, Vm);
                em_class_major(w.w, lutT, fm, F, accT, Vm);
            }
            for (int c = 0; c < nmin_max; ++c) {
                const uint32_t w = mn;
                mn = (c + 1 < my_min) ? __ldg(cmin + c + 1) : kNullClass;
                em_class_minor(w, lutT, f0, f1, f2, f3, F, accT, V0, V1, V2, V3);
            }
            if (active) {
                work += (unsigned long long)(nmaj + nmin);
                if (major == 0) V0 += Vm; else if (major == 1) V1 += Vm; else if (major == 2) V2 += Vm; else V3 += Vm;
                bool stop = (pass == max_iter), park = false;
                if (pass > 0) {
                    const double P = fabs(p0 - f0) * V0 + fabs(p1 - f1) * V1 + fabs(p2 - f2) * V2 + fabs(p3 - f3) * V3;
                    const int dec = delta_decision(P, eps);
                    if (dec == 1) stop = true;
                    else if (dec == 2 && !stop) park = true;         // K3b repeats this pass with the exact delta
                }
                FitRes* fr = a.fit + (size_t)(item >> 2) * 4 + (item & 3);
                if (park) {
                    fr->f[0] = f0; fr->f[1] = f1; fr->f[2] = f2; fr->f[3] = f3;
                    fr->eap[0] = p0; fr->eap[1] = p1; fr->eap[2] = p2; fr->eap[3] = p3;
                    fr->passes = pass; fr->flags = 0; fr->unfinished = 1;
                    a.q_long[atomicAdd(a.q_long_n, 1u)] = item;
                    active = false;
                } else {
                    const double e0 = f0 * (accT + V0) * inv_n, e1 = f1 * (accT + V1) * inv_n;
                    const double e2 = f2 * (accT + V2) * inv_n, e3 = f3 * (accT + V3) * inv_n;
                    if (stop) {
                        fr->f[0] = f0; fr->f[1] = f1; fr->f[2] = f2; fr->f[3] = f3;
                        fr->eap[0] = e0; fr->eap[1] = e1; fr->eap[2] = e2; fr->eap[3] = e3;
                        fr->passes = pass + 1; fr->flags = 0; fr->unfinished = 0;
                        active = false;
                    } else {
                        p0 = f0; p1 = f1; p2 = f2; p3 = f3;
                        f0 = e0; f1 = e1; f2 = e2; f3 = e3;              // Algorithm.cpp:95-101
                        F = ((f0 + f1) + f2) + f3;
                        ++pass;
                        park = (pass >= budget) || (thin && pass >= a.drain_min_pass);
                    }
                }
                if (park && active) {                                // resume point: pass `pass` with (f, p)
                    fr->f[0] = f0; fr->f[1] = f1; fr->f[2] = f2; fr->f[3] = f3;
                    fr->eap[0] = p0; fr->eap[1] = p1; fr->eap[2] = p2; fr->eap[3] = p3;
                    fr->passes = pass; fr->flags = 0; fr->unfinished = 1;
                    a.q_long[atomicAdd(a.q_long_n, 1u)] = item;
                    active = false;
                }
            }
        }
    }
    add_work(a.work + 0, work);
    add_work(a.work + 3, nfit);
}

// =====================================================================================================
// K3b: kLongGroup lanes per parked fit -- classes across the group's lanes, butterfly sums, exact delta.
// A pass costs a few hundred cycles instead of ncls x (latency of the dependent FP64 chain), so the fits
// that need up to 101 passes, the fits in flight when K3's queue drains and the fits of a small stage do
// not set the stage's duration.  Passes are aligned across the four groups of a warp as in K3.
// =====================================================================================================
constexpr int kLongThreads = 128;

#ifndef BVC_LONG_MINBLOCKS
#define BVC_LONG_MINBLOCKS 5               // 87 registers; capped at 80 for 6 blocks per SM it spills 8 bytes and C3's em_long runs 0.86 instead of 0.80 ms
#endif
__global__ void __launch_bounds__(kLongThreads, BVC_LONG_MINBLOCKS)
em_long_kernel(PipeArgs a)
{
    __shared__ double2 lut2[kNQ];
    load_lut2(lut2, a.s.lut);
    const uint32_t n_long = *a.q_long_n;
    const int lane = threadIdx.x & 31, g = lane & (kLongGroup - 1);
    const unsigned gmask = ((1u << kLongGroup) - 1u) << (lane & ~(kLongGroup - 1));
    const int gleader = lane & ~(kLongGroup - 1);
    const double eps = a.s.eps;
    const int max_iter = a.s.max_iter;
    unsigned long long work = 0;

    bool active = false, done = false, lowq = false, exact = false;
    const uint32_t* cp = a.cls;
    int ncls = 0, nmaj = 0, pass = 0;
    uint32_t item = 0, fl = 0;
    double f0 = 0, f1 = 0, f2 = 0, f3 = 0, p0 = 0, p1 = 0, p2 = 0, p3 = 0, F = 0, Fp = 0, inv_n = 0;

    for (;;) {
        // ---- refill: one queue slot per idle group (claimed by its leader lane) ------------------------------
        const bool want = !active && !done;
        const unsigned wm = __ballot_sync(0xffffffffu, want && g == 0);
        if (wm) {
            uint32_t base = 0;
            const int leader = __ffs(wm) - 1;
            if (lane == leader) base = atomicAdd(a.q_long_head, (uint32_t)__popc(wm));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (want) {
                const uint32_t idx = base + __popc(wm & ((1u << gleader) - 1));
                if (idx < n_long) {
                    item = a.q_long[idx];
                    const SiteSum su = a.sum[item >> 2];
                    const FitRes* fr = a.fit + (size_t)(item >> 2) * 4 + (item & 3);
                    f0 = fr->f[0]; f1 = fr->f[1]; f2 = fr->f[2]; f3 = fr->f[3];
                    p0 = fr->eap[0]; p1 = fr->eap[1]; p2 = fr->eap[2]; p3 = fr->eap[3];
                    F = ((f0 + f1) + f2) + f3; Fp = ((p0 + p1) + p2) + p3;
                    pass = fr->passes; fl = fr->flags;
                    inv_n = 1.0 / (double)(su.depth[0] + su.depth[1] + su.depth[2] + su.depth[3]);
                    cp = a.cls + su.cls_off; ncls = su.ncls; nmaj = su.n_major;
                    // a class with qual <= 1 voids delta_decision's sign argument (see em_fit_kernel): exact delta on every pass
                    lowq = (su.has_q0 & 2) != 0 || a.long_exact != 0; exact = lowq;
                    active = true;
                } else done = true;
            }
        }
        if (!__any_sync(0xffffffffu, active)) break;
        const int nmine = active ? ncls : 0;
        const int trips = (__reduce_max_sync(0xffffffffu, nmine) + kLongGroup - 1) / kLongGroup;

        // The exact delta (a second marginal and a log1p per class, two more butterflies) doubles a pass: the warp evaluates it
        // only if one of its fits needs it -- a qual <= 1 class, or a pass whose bound-based decision came out ambiguous
        const bool warp_exact = __any_sync(0xffffffffu, active && exact);
        EmAcc A; A.accT = A.V0 = A.V1 = A.V2 = A.V3 = A.dl = 0.0; A.xmax = 0;
        if (warp_exact) {
#pragma unroll 2
            for (int k = 0; k < trips; ++k) {
                const int c = k * kLongGroup + g;           // class index -> pool word: the minor classes start behind the padding
                const uint32_t w = (c < nmine) ? __ldg(cp + (c < nmaj ? c : c + (pad4(nmaj) - nmaj))) : kNullClass;
                em_class_exact(w, lut2, f0, f1, f2, f3, F, p0, p1, p2, p3, Fp, A);
            }
        } else {
#pragma unroll 2
            for (int k = 0; k < trips; ++k) {
                const int c = k * kLongGroup + g;
                const uint32_t w = (c < nmine) ? __ldg(cp + (c < nmaj ? c : c + (pad4(nmaj) - nmaj))) : kNullClass;
                em_class_plain(w, lut2, f0, f1, f2, f3, F, A);
            }
        }
#pragma unroll
        for (int o = kLongGroup / 2; o > 0; o >>= 1) {               // butterfly within the group: all its lanes agree
            A.accT += __shfl_xor_sync(0xffffffffu, A.accT, o);
            A.V0 += __shfl_xor_sync(0xffffffffu, A.V0, o);
            A.V1 += __shfl_xor_sync(0xffffffffu, A.V1, o);
            A.V2 += __shfl_xor_sync(0xffffffffu, A.V2, o);
            A.V3 += __shfl_xor_sync(0xffffffffu, A.V3, o);
            if (warp_exact) {
                A.dl += __shfl_xor_sync(0xffffffffu, A.dl, o);
                A.xmax = max(A.xmax, __shfl_xor_sync(0xffffffffu, A.xmax, o));
            }
        }
        if (active) {
            if (g == 0) work += (unsigned long long)ncls;
            const double e0 = f0 * (A.accT + A.V0) * inv_n, e1 = f1 * (A.accT + A.V1) * inv_n;
            const double e2 = f2 * (A.accT + A.V2) * inv_n, e3 = f3 * (A.accT + A.V3) * inv_n;
            bool stop = (pass == max_iter), redo = false;
            if (pass > 0) {
                if (warp_exact) {
                    if (A.xmax < kBigXHi) {
                        if (near_rel(A.dl, eps)) fl |= BVC_FLAG_EM_DELTA;
                        if (A.dl < eps) stop = true;                 // Algorithm.cpp:124
                    }
                } else {
                    const double P = fabs(p0 - f0) * A.V0 + fabs(p1 - f1) * A.V1 + fabs(p2 - f2) * A.V2 + fabs(p3 - f3) * A.V3;
                    const int dec = delta_decision(P, eps);
                    if (dec == 1) stop = true;
                    else if (dec == 2 && !stop) redo = true;         // this pass again, with the exact sum
                }
            }
            if (redo) exact = true;
            else if (stop) {
                if (g == 0) {
                    FitRes* fr = a.fit + (size_t)(item >> 2) * 4 + (item & 3);
                    fr->f[0] = f0; fr->f[1] = f1; fr->f[2] = f2; fr->f[3] = f3;
                    fr->eap[0] = e0; fr->eap[1] = e1; fr->eap[2] = e2; fr->eap[3] = e3;
                    fr->passes = pass + 1; fr->flags = fl; fr->unfinished = 0;
                }
                active = false;
            } else {
                p0 = f0; p1 = f1; p2 = f2; p3 = f3; Fp = F;
                f0 = e0; f1 = e1; f2 = e2; f3 = e3;                  // Algorithm.cpp:95-101
                F = ((f0 + f1) + f2) + f3;
                ++pass; exact = lowq;
            }
        }
        (void)gmask;
    }
    add_work(a.work + 1, work);
}

// =====================================================================================================
// K3c: thread per fit -- lr = sum_c cnt_c log ml_c(f) with the f of the fit's LAST E-step
// (BaseType.cpp:57-60).  Same sorted queue, so a warp's 32 loops have (nearly) equal length.
// =====================================================================================================
__global__ void __launch_bounds__(kEmThreads)
fit_loglik_kernel(PipeArgs a)
{
    __shared__ double2 lut2[kNQ];
    __shared__ double2 lutlog[kNQ];                     // (log(e/3), log(1-e)): glibc values from the host table
    for (int i = threadIdx.x; i < kNQ; i += blockDim.x) lutlog[i] = make_double2(a.s.lut[2 * kNQ + i], a.s.lut[3 * kNQ + i]);
    load_lut2(lut2, a.s.lut);
    const uint32_t nfits = *a.q_in_n;
    unsigned long long work = 0;
    const uint32_t i0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u, lane = threadIdx.x & 31;
    for (uint32_t wbase = i0; wbase < nfits; wbase += gridDim.x * blockDim.x) {
        const uint32_t idx = wbase + lane;
        const bool on = idx < nfits;
        const uint32_t item = on ? a.q_sorted[idx] : 0u;
        FitRes* fr = a.fit + (size_t)(item >> 2) * 4 + (item & 3);
        double f0 = 0, f1 = 0, f2 = 0, f3 = 0;
        const uint32_t* cp = a.cls; int ncls = 0;
        uint32_t mask = 0;
        if (on) {
            const SiteSum su = a.sum[item >> 2];
            f0 = fr->f[0]; f1 = fr->f[1]; f2 = fr->f[2]; f3 = fr->f[3];
            cp = a.cls + su.cls_off; ncls = pad4(su.n_major) + (su.ncls - su.n_major); mask = fr->mask;
        }
        // a two-allele model is followed by the two single-base models of its alleles (BaseType.cpp:93-110 at k = 1):
        // closed form, lr = sum_c cnt_c log L_c,b -- accumulated here, in the same class loop (see ladder_site)
        const bool pair = __popc(mask) == 2;
        const uint32_t bl = (uint32_t)(__ffs(mask) - 1) << 7, bh = (uint32_t)(31 - __clz(mask | 1u)) << 7;
        const double F = ((f0 + f1) + f2) + f3;
        const int nmax = __reduce_max_sync(0xffffffffu, ncls);
        double s = 0.0, sl = 0.0, sh = 0.0;
        // straight-line bodies: a lane past its last class (or without a fit: ml == 0, result unused) evaluates count-0 null
        // classes, so the unrolled log evaluations interleave instead of sitting behind a divergent branch each
#pragma unroll 4
        for (int c = 0; c < nmax; ++c) {
            const uint32_t w = (c < ncls) ? __ldg(cp + c) : kNullClass;
            const int q = w & (kNQ - 1);
            const double2 ed = lut2[q];
            const double2 lg = lutlog[q];
            const double cnt = cnt_to_double(w >> 9);
            const double ml = fma(sel4_w(w, f0, f1, f2, f3), ed.y, ed.x * F);
            s = fma(cnt, log(ml), s);
            sl = fma(cnt, (w & 0x180u) == bl ? lg.y : lg.x, sl);
            sh = fma(cnt, (w & 0x180u) == bh ? lg.y : lg.x, sh);
        }
        if (on) {
            fr->lr = s; work += (unsigned long long)ncls;
            if (pair) { fr->f[0] = sl; fr->f[1] = sh; }     // f is dead once lr is known: it now carries the singles' lr
        }
    }
    add_work(a.work + 2, work);
}

// =====================================================================================================
// K4: thread per site -- log-likelihoods of the finished fits, LRT decision, next stage or final call
// =====================================================================================================
// closed form for a single-base model: f = e_b is a fixed point of EM (2 passes, delta == 0);
// lr = sum_c cnt_c * log L_cb from the host-built (glibc) log table
__device__ __forceinline__ double single_loglik(const PipeArgs& a, const SiteSum& su, int base)
{
    const uint32_t* cp = a.cls + su.cls_off;
    double s = 0.0;
    const int n = pad4(su.n_major) + (su.ncls - su.n_major);
    for (int c = 0; c < n; ++c) {
        const uint32_t w = __ldg(cp + c);
        const int q = w & (kNQ - 1), b = (w >> 7) & 3;
        s = fma((double)(w >> 9), __ldg(a.s.lut + (b == base ? 3 : 2) * kNQ + q), s);
    }
    return s;
}

__device__ __forceinline__ void ladder_site(const PipeArgs& a, uint32_t item)
{
    if (item & 3u) return;
    const int site = (int)(item >> 2);
    const SiteSum su = a.sum[site];
    Ladder L = a.lad[site];
    const int ref = a.s.ref_base[site] & 3;
    const FitRes* fr = a.fit + (size_t)site * 4;
    int n = L.n, k = L.k;
    bool cont = true;          // ladder still running
    const FitRes* src = nullptr;   // the two-allele fit that the current candidate pair came from

    // ---- results of the EM stage in flight -----------------------------------------------------------------
    if (k == n) {                                                  // the full model (BaseType.cpp:88-90)
        L.lr_alt = fr[0].lr;
        if (n == 2) src = fr;
#pragma unroll
        for (int j = 0; j < 4; ++j) L.frq[j] = fr[0].eap[j];
        L.em_fits += 1; L.em_passes += fr[0].passes; L.flags |= fr[0].flags;
    } else {                                                       // k-subsets, combs_ order (BaseType.cpp:93-110)
        double cmin = 0.0, lr_min = 0.0; int imin = 0;
        for (int i = 0; i < L.nfit; ++i) {
            const double lr = fr[i].lr;
            const double chi = 2.0 * (L.lr_alt - lr);
            if (i > 0 && near_tie(chi, cmin, L.lr_alt, lr)) L.flags |= BVC_FLAG_ARGMIN_TIE;
            if (i == 0 || chi < cmin) { cmin = chi; lr_min = lr; imin = i; }
            L.em_fits += 1; L.em_passes += fr[i].passes; L.flags |= fr[i].flags;
        }
        L.lr_alt = lr_min; L.chi_t = cmin;
        if (near_rel(cmin, a.s.lrt_thr)) L.flags |= BVC_FLAG_LRT_CHI;
        if (cmin < a.s.lrt_thr) {
            const uint32_t keep = fr[imin].mask;
            if (__popc(keep) == 2) src = fr + imin;
            int nn = 0;
            uint32_t c4 = 0xffffffffu;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int b = cand_at(L.cand4, t);
                if (t < n && (keep >> b & 1)) { c4 = (c4 & ~(0xffu << (8 * nn))) | ((uint32_t)b << (8 * nn)); ++nn; }
            }
            L.cand4 = c4; n = nn;
#pragma unroll
            for (int j = 0; j < 4; ++j) L.frq[j] = fr[imin].eap[j];
        } else cont = false;
    }
    // ---- following stages: singles are closed-form (done here), larger subsets go back to K3 ---------------
    while (cont) {
        k = n - 1;
        if (k < 1) break;
        if (k >= 2) {
            int slot = 0;
            for (unsigned m = (1u << n) - 1; m >= 1; --m) {            // combs_ order: bit (n-1-i) <=> cand[i]
                if (__popc(m) != k) continue;
                uint32_t sub = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i) if (i < n && (m >> (n - 1 - i) & 1)) sub |= 1u << cand_at(L.cand4, i);
                a.fit[(size_t)site * 4 + slot].mask = sub;
                ++slot;
            }
            L.n = (int8_t)n; L.k = (int8_t)k; L.nfit = (int8_t)slot;
            a.lad[site] = L;
            const uint32_t pos = atomicAdd(a.q_out_n, (uint32_t)slot);
            for (int i = 0; i < slot; ++i) a.q_out[pos + i] = ((uint32_t)site << 2) | (uint32_t)i;
            return;
        }
        // k == 1: n == 2 single-base models, in cand order; fit_loglik_kernel left their lr in the pair fit's f[0..1]
        double cmin = 0.0, lr_min = 0.0; int imin = 0;
        for (int i = 0; i < n; ++i) {
            const double lr = (src && n == 2) ? src->f[i] : single_loglik(a, su, cand_at(L.cand4, i));
            const double chi = 2.0 * (L.lr_alt - lr);
            if (i > 0 && near_tie(chi, cmin, L.lr_alt, lr)) L.flags |= BVC_FLAG_ARGMIN_TIE;
            if (i == 0 || chi < cmin) { cmin = chi; lr_min = lr; imin = i; }
            L.em_fits += 1; L.em_passes += 2;
        }
        L.lr_alt = lr_min; L.chi_t = cmin;
        if (near_rel(cmin, a.s.lrt_thr)) L.flags |= BVC_FLAG_LRT_CHI;
        if (cmin < a.s.lrt_thr) {
            const int b = cand_at(L.cand4, imin);
            L.cand4 = 0xffffff00u | (uint32_t)b; n = 1;
#pragma unroll
            for (int j = 0; j < 4; ++j) L.frq[j] = (j == b) ? 1.0 : 0.0;
        }
        cont = false;
    }
    finalize_site(a, site, ref, su, L.cvg_fs, L.cvg_sor, L.cand4, n, L.frq[0], L.frq[1], L.frq[2], L.frq[3], L.chi_t, L.em_fits, L.em_passes, L.flags);
}

#ifndef BVC_LADDER_MINBLOCKS
#define BVC_LADDER_MINBLOCKS 1               // 80 registers; 8 blocks per SM (64 registers) measured: 0.245 against 0.240 ms, no gain
#endif
__global__ void __launch_bounds__(128, BVC_LADDER_MINBLOCKS)
ladder_kernel(PipeArgs a)
{
    // the sites with fits in flight are exactly the slot-0 items of this stage's (unsorted) queue
    const uint32_t n_items = *a.q_in_n;
    for (uint32_t it = blockIdx.x * blockDim.x + threadIdx.x; it < n_items; it += gridDim.x * blockDim.x)
        ladder_site(a, a.q_in[it]);
}


// =====================================================================================================
// K5: warp per LRT-positive site -- VCF strand table, Fisher/SOR, three rank-sum tests
// =====================================================================================================
constexpr int kVcfWarps = 4;

#ifndef BVC_VCF_MINBLOCKS
#define BVC_VCF_MINBLOCKS 8                // 64 registers (132 B of spills) for 8 blocks per SM: the kernel is latency-bound (27 % of the warp slots at
#endif                                    // 88 registers); C5's 25k LRT-positive sites: 0.337 -> 0.283 ms (7 blocks: 0.291)
__global__ void __launch_bounds__(kVcfWarps * 32, BVC_VCF_MINBLOCKS)
vcf_stats_kernel(PipeArgs a)
{
    __shared__ uint32_t H_s[kVcfWarps][6 * 256];
    __shared__ uint8_t dict_s[32];                      // the qual dictionary: indexed per entry (a kernel parameter indexed by a
                                                        // lane-varying value is served one distinct address at a time)
    if (threadIdx.x < 32) dict_s[threadIdx.x] = a.s.dict.q[threadIdx.x];
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t* H = H_s[wid];
    const uint32_t n_list = *a.vcf_n;
    for (uint32_t i = blockIdx.x * kVcfWarps + wid; i < n_list; i += gridDim.x * kVcfWarps) {
        const int site = (int)a.vcf_list[i];
        const SiteSum su = a.sum[site];
        bvc_site_result* r = a.s.out + site;
        const int ref = a.s.ref_base[site] & 3;
        unsigned altmask = 0;
        for (int t = 0; t < r->n_alt; ++t) altmask |= 1u << r->alt_bases[t];
        // BaseType.cpp:174-187: non-indel entries only, alt = any surviving alt base
        int vs[4] = {0, 0, 0, 0};
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            if (b == ref) { vs[0] = su.fwd[b]; vs[1] = su.depth[b] - su.fwd[b]; }
            else if (altmask >> b & 1) { vs[2] += su.fwd[b]; vs[3] += su.depth[b] - su.fwd[b]; }
        }
        const double vcf_fs = dm_fisher_warp(a.s.lgam, vs[0], vs[1], vs[2], vs[3], lane);
        const double vcf_sor = dm_sor(vs[0], vs[1], vs[2], vs[3]);
        for (int j = lane; j < 6 * 256; j += 32) H[j] = 0;
        __syncwarp();
        // 4 entries per lane per 32-bit load.  mapq / rpr may live in the caller's PINNED host memory
        // (zero-copy over PCIe: only LRT-positive sites ever touch them, see upload() in bvc_api.cu),
        // so they are fetched only by lanes that hold at least one ref/alt entry.
        const uint32_t b0 = a.s.entry_off[site], b1 = a.s.entry_off[site + 1];
        const uint32_t w0 = b0 >> 2, w1 = (b1 + 3) >> 2;
        const uint32_t* code4 = reinterpret_cast<const uint32_t*>(a.s.code);
        const uint32_t* qual4 = reinterpret_cast<const uint32_t*>(a.s.qual);
        const uint32_t* mapq4 = reinterpret_cast<const uint32_t*>(a.s.mapq);
        const uint32_t* rpr4 = reinterpret_cast<const uint32_t*>(a.s.rpr);
        const uint32_t* pk4 = reinterpret_cast<const uint32_t*>(a.s.pk8);
        for (uint32_t wi = w0 + lane; wi < w1; wi += 32) {
            // the one-byte form (pk8 != NULL) carries the same (base, is_indel) and the qual as a dictionary index
            const uint32_t cw = pk4 ? __ldg(pk4 + wi) : __ldg(code4 + wi);
            int g[4]; bool any = false;
            uint32_t qw = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t e = wi * 4 + k;
                const unsigned c = (cw >> (8 * k)) & 255u;
                const bool cnt = pk4 ? BVC_PACK8_QI(c) < BVC_PACK8_MAX_QUALS : !(c & 16);
                const int base = pk4 ? (int)BVC_PACK8_BASE(c) : (int)(c & 7u);
                g[k] = -1;
                if (e >= b0 && e < b1 && cnt) {
                    if (base == ref) g[k] = 0; else if (base < 4 && (altmask >> base & 1)) g[k] = 1;
                }
                any |= g[k] >= 0;
            }
            if (!any) continue;
            if (pk4) {
#pragma unroll
                for (int k = 0; k < 4; ++k) if (g[k] >= 0) qw |= (uint32_t)dict_s[BVC_PACK8_QI((cw >> (8 * k)) & 255u)] << (8 * k);
            } else qw = __ldg(qual4 + wi);
            const uint32_t mw = mapq4[wi], rw = rpr4[wi];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (g[k] < 0) continue;
                atomicAdd(&H[(0 * 2 + g[k]) * 256 + ((mw >> (8 * k)) & 255u)], 1u);
                atomicAdd(&H[(1 * 2 + g[k]) * 256 + ((qw >> (8 * k)) & 255u)], 1u);
                atomicAdd(&H[(2 * 2 + g[k]) * 256 + ((rw >> (8 * k)) & 255u)], 1u);
            }
        }
        __syncwarp();
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
}  // namespace bvc
