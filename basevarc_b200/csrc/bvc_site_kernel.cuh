// bvc_site_kernel.cuh -- the per-site `basetype` calling kernel for sm_100a (one warp per site).
//
// Replaces, for a whole tile of sites per launch, what the reference does one site at a time on a
// host thread (paths under /root/reference):
//   bt_f counting + CVG strand table + Fisher/SOR      src/BaseVarC.cpp:548-608
//   BaseType ctor (likelihood table, depth)            src/BaseType.cpp:5-23
//   BaseType::LRT / UpdateF / SetAlleleFreq / combs_   src/BaseType.cpp:25-139,237-255
//   EM / singleEM / delta_bylog                        src/Algorithm.cpp:69-130
//   WriteVcf numeric part (VCF strand table, FS, SOR, three rank-sum tests)   src/BaseType.cpp:149-197
//
// Design (B200-first, not a translation):
//  * Phase A streams the site's packed entries once (code + qual bytes, coalesced) into a per-warp
//    shared-memory histogram over (base, qual) and a 32-bin histogram over the 5-bit code byte.
//    All reads of a site with equal (base, qual) have identical E-step rows (BaseType.cpp:13-15),
//    so the EM runs over <= 4*128 weighted CLASSES instead of n reads: same mathematics, only the
//    summation order differs (count-weighted + warp butterfly instead of a sequential loop).
//  * Phase B compacts the non-empty classes in place (ballot + popc).
//  * Phase C runs the LRT ladder warp-uniformly; each EM pass gives every lane ceil(ncls/32)
//    classes, FP64 throughout with true divisions, and one 5-value butterfly reduction per pass.
//    Single-base subsets are closed-form (f=1 is a fixed point: post=L/L=1, delta=0) and use a
//    host-built (glibc) log-likelihood table, so they are bit-identical per read to the reference.
//  * Phase D (LRT-positive sites only) re-reads the entries (now L2/L1 resident) to histogram
//    mapq/qual/rpr per {ref, alt} and evaluates the three Mann-Whitney tests exactly from the
//    256-bin histograms (rank sums are half-integers: exact in integer arithmetic).
//  * The per-read likelihoods come from a host-built table lut[q] (glibc exp), so L_ij is
//    bit-identical to BaseType.cpp:13-15; lgamma for Fisher comes from a host table.
#pragma once
#include <stdint.h>
#include "../../include/bvc_gpu.h"
#include "bvc_device_math.cuh"
#include "bvc_kernels.h"

namespace bvc {

struct WarpSmem {
    double*   lml;       // log marginal likelihood per class (state between EM passes)
    uint32_t* hist;      // (base<<7 | qual) counts; after compaction hist[0..ncls) = class counts
    uint16_t* keys;      // class keys after compaction
    uint32_t* codehist;  // 32 bins over the code byte
};

struct SiteState {
    int depth[4];
    int n_reads;          // non-indel entries
    int ncls;
    bool has_q0;          // a class with qual 0 exists (1-e == 0: the NaN paths of SURVEY Q8)
    int em_fits, em_passes;
    uint32_t flags;
};


// One EM fit over the classes for the allele subset `mask` (bit b = base b).  Warp-uniform result.
// Algorithm.cpp:115-130 with UpdateF's start vector and log-likelihood sum (BaseType.cpp:25-39,57-60).
__device__ __noinline__ void em_fit(const SiteArgs& a, const WarpSmem& w, const double* lut_s, SiteState& st,
                              unsigned mask, int lane, double& lr, double frq[4])
{
    st.em_fits += 1;
    const int nb = __popc(mask);
    if (nb == 1 && !st.has_q0) {
        // closed form: f = e_b is a fixed point, EM stops after its 2nd pass with delta == 0
        const int b = __ffs(mask) - 1;
        double s = 0.0;
        for (int c = lane; c < st.ncls; c += 32) {
            const unsigned key = w.keys[c];
            const int q = key & (kNQ - 1);
            const double l = ((int)(key >> 7) == b) ? lut_s[3 * kNQ + q] : lut_s[2 * kNQ + q];
            s += (double)w.hist[c] * l;
        }
        lr = warp_sum(s);
        for (int j = 0; j < 4; ++j) frq[j] = (j == b) ? 1.0 : 0.0;
        st.em_passes += 2;
        return;
    }
    int dsum = 0;
    for (int j = 0; j < 4; ++j) if (mask >> j & 1) dsum += st.depth[j];
    double f0, f1, f2, f3;
    f0 = (dsum > 0 && (mask & 1)) ? (double)st.depth[0] / dsum : 0.0;
    f1 = (dsum > 0 && (mask & 2)) ? (double)st.depth[1] / dsum : 0.0;
    f2 = (dsum > 0 && (mask & 4)) ? (double)st.depth[2] / dsum : 0.0;
    f3 = (dsum > 0 && (mask & 8)) ? (double)st.depth[3] / dsum : 0.0;
    double e0 = 0, e1 = 0, e2 = 0, e3v = 0;
    const double n = (double)st.n_reads;
    for (int it = 0; it <= a.max_iter; ++it) {
        double acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0, dl = 0;
        for (int c = lane; c < st.ncls; c += 32) {
            const unsigned key = w.keys[c];
            const int q = key & (kNQ - 1), b = key >> 7;
            const double cnt = (double)w.hist[c];
            const double e3 = lut_s[q], ome = lut_s[kNQ + q];
            const double l0 = f0 * (b == 0 ? ome : e3);
            const double l1 = f1 * (b == 1 ? ome : e3);
            const double l2 = f2 * (b == 2 ? ome : e3);
            const double l3 = f3 * (b == 3 ? ome : e3);
            const double ml = ((l0 + l1) + l2) + l3;                   // Algorithm.cpp:76-79
            // post_j = l_j / ml (:80-84) as l_j * (1/ml): one IEEE division per class instead of four
            // (<= 1.5 ulp from the reference; 0 * (1/0) = NaN keeps the 0/0 semantics of qual-0 reads).
            const double r = cnt / ml;
            acc0 += l0 * r; acc1 += l1 * r; acc2 += l2 * r; acc3 += l3 * r;   // :87-91, count-weighted
            const double lg = log(ml);
            if (it > 0) dl += cnt * fabs(lg - w.lml[c]);              // :103-113
            w.lml[c] = lg;
        }
        // only alleles inside the subset carry mass; the others stay exactly 0
        if (mask & 1) e0 = warp_sum(acc0) / n;
        if (mask & 2) e1 = warp_sum(acc1) / n;
        if (mask & 4) e2 = warp_sum(acc2) / n;
        if (mask & 8) e3v = warp_sum(acc3) / n;
        st.em_passes += 1;
        if (it > 0) {
            const double delta = warp_sum(dl);
            if (near_rel(delta, a.eps)) st.flags |= BVC_FLAG_EM_DELTA;
            if (delta < a.eps) break;                                  // :124
        }
        if (it == a.max_iter) break;
        f0 = e0; f1 = e1; f2 = e2; f3 = e3v;                            // :95-101
    }
    double s = 0.0;
    for (int c = lane; c < st.ncls; c += 32) s += (double)w.hist[c] * w.lml[c];   // BaseType.cpp:57-60
    lr = warp_sum(s);
    frq[0] = e0; frq[1] = e1; frq[2] = e2; frq[3] = e3v;
}

struct LrtOut {
    int ok, n_alt;
    int alt_bases[3];
    double af[4];
    double chi, var_qual;
};

// BaseType::LRT (BaseType.cpp:73-139) over the candidate list comb[0..n_comb) (base codes, in order).
__device__ __noinline__ void lrt_ladder(const SiteArgs& a, const WarpSmem& w, const double* lut_s, SiteState& st,
                                  int ref, const int* comb, int n_comb, int lane, LrtOut& o)
{
    o.ok = 0; o.n_alt = 0; o.chi = 0.0; o.var_qual = 0.0;
    o.alt_bases[0] = o.alt_bases[1] = o.alt_bases[2] = -1;
    for (int j = 0; j < 4; ++j) o.af[j] = 0.0;
    const double total = (double)st.n_reads;
    if (st.n_reads == 0) return;                                       // :75
    int cand[4], n = 0;
    for (int t = 0; t < n_comb; ++t) {                                 // :77-82
        const double fr = (double)st.depth[comb[t]] / total;
        if (near_rel(fr, a.min_af)) st.flags |= BVC_FLAG_MIN_AF;
        if (fr >= a.min_af) cand[n++] = comb[t];
    }
    if (n == 0) return;
    unsigned full = 0;
    for (int t = 0; t < n; ++t) full |= 1u << cand[t];
    double lr_alt, frq[4], chi_t = 0.0;
    em_fit(a, w, lut_s, st, full, lane, lr_alt, frq);                  // :88-90
    for (int k = n - 1; k > 0; --k) {                                  // :93-110
        // k-subsets of cand[0..n) in combs_ order (:237-255): bit (n-1-i) of m <=> index i, m descending
        double cmin = 0.0, lr_min = 0.0, bp_min[4] = {0, 0, 0, 0};
        unsigned m_min = 0; bool first = true;
        for (unsigned m = (1u << n) - 1; m >= 1; --m) {
            if (__popc(m) != k) continue;
            unsigned sub = 0;
            for (int i = 0; i < n; ++i) if (m >> (n - 1 - i) & 1) sub |= 1u << cand[i];
            double lr_c, bp[4];
            em_fit(a, w, lut_s, st, sub, lane, lr_c, bp);
            const double chi = 2.0 * (lr_alt - lr_c);
            if (!first && near_tie(chi, cmin, lr_alt, lr_c)) st.flags |= BVC_FLAG_ARGMIN_TIE;
            if (first || chi < cmin) {                                 // std::min_element: first minimum
                cmin = chi; lr_min = lr_c; m_min = m;
                for (int j = 0; j < 4; ++j) bp_min[j] = bp[j];
            }
            first = false;
        }
        lr_alt = lr_min; chi_t = cmin;
        if (near_rel(chi_t, a.lrt_thr)) st.flags |= BVC_FLAG_LRT_CHI;
        if (chi_t < a.lrt_thr) {
            int nn = 0, keep[4];
            for (int i = 0; i < n; ++i) if (m_min >> (n - 1 - i) & 1) keep[nn++] = cand[i];
            for (int i = 0; i < nn; ++i) cand[i] = keep[i];
            n = nn;
            for (int j = 0; j < 4; ++j) frq[j] = bp_min[j];
        } else break;
    }
    for (int t = 0; t < n; ++t)                                        // :111-116
        if (cand[t] != ref) { o.alt_bases[o.n_alt++] = cand[t]; o.af[cand[t]] = frq[cand[t]]; }
    o.chi = chi_t;
    if (o.n_alt == 0) return;
    o.ok = 1;
    const double r = (double)st.depth[cand[0]] / total;                // :118-134
    if (n == 1 && total > 10 && r > 0.5) { o.var_qual = 5000.0; return; }
    if (chi_t <= 0) { o.var_qual = 0.0; return; }
    const double p = dm_chisf(chi_t, 1.0);
    double vq = (p != 0) ? -10.0 * log10(p) : 10000.0;
    if (vq == 0) vq = 0.0;
    o.var_qual = vq;
}

// Exact Mann-Whitney from two 256-bin histograms (ref = x, alt = y), Algorithm.cpp:27-67.
__device__ __noinline__ double ranksum_hist(const uint32_t* hr, const uint32_t* ha, int lane)
{
    // lane L owns values 255-8L .. 248-8L (descending order overall)
    uint32_t r[8], t[8]; uint32_t tl = 0, rl = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int v = 255 - (lane * 8 + i);
        r[i] = hr[v]; t[i] = r[i] + ha[v];
        tl += t[i]; rl += r[i];
    }
    uint32_t incl = tl;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t up = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += up; }
    long long kb = (long long)(incl - tl), twice = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { twice += (long long)r[i] * (2 * kb + t[i] + 1); kb += t[i]; }
    twice = warp_sum_ll(twice);
    const unsigned long long n1 = (unsigned long long)warp_sum_ll((long long)rl);
    const unsigned long long nall = (unsigned long long)__shfl_sync(0xffffffffu, incl, 31);
    const unsigned long long n2 = nall - n1;
    const double r1 = (double)twice / 2.0;
    const double expected = (double)(n1 * (n1 + n2 + 1)) / 2.0;
    const double z = (r1 - expected) / sqrt((double)(n1 * n2 * (n1 + n2 + 1)) / 12.0);
    // n1 == 0 or n2 == 0: z = 0/0.  The reference's NaN goes through fabs() and keeps a clear sign bit down to the
    // "{:.3f}" (prints "nan"); the device's canonical NaN has the sign bit set, so clear it.
    const double v = dm_phred(2 * dm_normsf(fabs(z)));
    return (v != v) ? __longlong_as_double(0x7ff8000000000000LL) : v;   // abs.f64 of a NaN is unspecified in PTX
}

}  // namespace bvc
