// bvc_device_math.cuh -- FP64 statistical leaves of the calling core, device side (sm_100a).
//
// From-scratch device restatements of the arithmetic the reference takes from htslib 1.5 kfunc.c
// and its own Algorithm.cpp (paths under /root/reference):
//   dm_chisf     = chisf -> kf_gammaq(k/2, x/2)          src/Algorithm.cpp:3-7,   SeqLib/htslib/kfunc.c:97-141
//   dm_normsf    = normsf -> kf_erfc                     src/Algorithm.cpp:9-14,  kfunc.c:56-82
//   dm_fisher    = bt_fisher_exact -> kt_fisher_exact    src/Algorithm.cpp:16-25, kfunc.c:195-279
//   dm_phred     = the -10*log10(p) tail with inf -> 10000, -0 -> 0   src/Algorithm.cpp:20-23,62-65
// The operation order of the reference is kept; exp/log/log10 are CUDA libdevice (<= 1-2 ulp from
// glibc), and libm lgamma is replaced by a host-built table lgam[k] = lgamma((double)k) so that
// lbinom() is bit-identical to the reference's.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace bvc {

// a hard FP64 threshold met within 1e-9 relative: the site is reported as a threshold-boundary site
__device__ __forceinline__ bool near_rel(double x, double thr) { return fabs(x - thr) <= 1e-9 * fabs(thr); }
// Two null models of one ladder step "tie" (BVC_FLAG_ARGMIN_TIE, src/BaseType.cpp:99 std::min_element) when their chi agree
// within 1e-9 relative OR within 1e-12 of the log-likelihood sums they are differences of: chi = 2 (lr_alt - lr) cancels two
// sums of magnitude n |log L|, so a tiny chi is only known to ~1e-12 |lr| (seen: two depth-3 error bases at a 10,000-sample
// site, chi 0.0136, |chi_a - chi_b| 5e-11 -- the argmin is then a coin toss between summation orders).
__device__ __forceinline__ bool near_tie(double chi, double cmin, double lr_alt, double lr)
{
    return fabs(chi - cmin) <= 1e-9 * fmax(fabs(chi), fabs(cmin)) + 1e-12 * (fabs(lr_alt) + fabs(lr));
}

__device__ __forceinline__ double dm_phred(double p)
{
    double v = -10.0 * log10(p);
    if (isinf(v)) v = 10000.0;
    else if (v == 0) v = 0.0;
    return v;
}

// kfunc.c:37-50 (AS245)
__device__ inline double dm_lgamma_as245(double z)
{
    double x = 0;
    x += 0.1659470187408462e-06 / (z + 7);
    x += 0.9934937113930748e-05 / (z + 6);
    x -= 0.1385710331296526 / (z + 5);
    x += 12.50734324009056 / (z + 4);
    x -= 176.6150291498386 / (z + 3);
    x += 771.3234287757674 / (z + 2);
    x -= 1259.139216722289 / (z + 1);
    x += 676.5203681218835 / z;
    x += 0.9999999999995183;
    return log(x) - 5.58106146679532777 - z + (z - 0.5) * log(z + 6.5);
}

// kfunc.c:56-82 (AS66)
__device__ __noinline__ double dm_erfc(double x)
{
    const double SQRT2 = 1.41421356237309504880;
    double z = fabs(x) * SQRT2;
    if (z > 37.) return x > 0. ? 0. : 2.;
    double e = exp(z * z * -.5), p;
    if (z < 10. / SQRT2) {
        double num = .03526249659989109;
        num = num * z + .7003830644436881;
        num = num * z + 6.37396220353165;
        num = num * z + 33.912866078383;
        num = num * z + 112.0792914978709;
        num = num * z + 221.2135961699311;
        num = num * z + 220.2068679123761;
        double den = .08838834764831844;
        den = den * z + 1.755667163182642;
        den = den * z + 16.06417757920695;
        den = den * z + 86.78073220294608;
        den = den * z + 296.5642487796737;
        den = den * z + 637.3336333788311;
        den = den * z + 793.8265125199484;
        den = den * z + 440.4137358247522;
        p = e * num / den;
    } else {
        p = e / 2.506628274631001 / (z + 1. / (z + 2. / (z + 3. / (z + 4. / (z + .65)))));
    }
    return x > 0. ? 2. * p : 2. * (1. - p);
}

__device__ inline double dm_normsf(double x) { return dm_erfc(x / sqrt(2.0)) / 2.0; }

// kfunc.c:101-131,138-141
__device__ __noinline__ double dm_gammaq(double s, double z)
{
    const double EPS = 1e-14, TINY = 1e-290;
    if (z <= 1. || z < s) {
        double sum = 1., x = 1.;
        for (int k = 1; k < 100; ++k) {
            x *= z / (s + k);
            sum += x;
            if (x / sum < EPS) break;
        }
        return 1. - exp(s * log(z) - z - dm_lgamma_as245(s + 1.) + log(sum));
    }
    double f = 1. + z - s, C = f, D = 0.;
    for (int j = 1; j < 100; ++j) {
        double a = j * (s - j), b = (j << 1) + 1 + z - s, d;
        D = b + a * D;
        if (D < TINY) D = TINY;
        C = b + a / C;
        if (C < TINY) C = TINY;
        D = 1. / D;
        d = C * D;
        f *= d;
        if (fabs(d - 1.) < EPS) break;
    }
    return exp(s * log(z) - z - dm_lgamma_as245(s) - log(f));
}

__device__ inline double dm_chisf(double x, double k) { return dm_gammaq(k / 2.0, x / 2.0); }

// kfunc.c:195-210 with lgam[k] = lgamma(k) from the host (glibc), k <= n+1
__device__ __forceinline__ double dm_lbinom(const double* __restrict__ lgam, int n, int k)
{
    if (k == 0 || n == k) return 0;
    return lgam[n + 1] - lgam[k + 1] - lgam[n - k + 1];
}
__device__ __forceinline__ double dm_hypergeo(const double* __restrict__ lgam, int n11, int n1_, int n_1, int n)
{
    return exp(dm_lbinom(lgam, n1_, n11) + dm_lbinom(lgam, n - n1_, n_1 - n11) - dm_lbinom(lgam, n, n_1));
}

struct HgAcc { int n11, n1_, n_1, n; double p; };

// kfunc.c:218-241, the "only n11 changed" call
__device__ __forceinline__ double dm_hg_move(const double* __restrict__ lgam, HgAcc& a, int n11)
{
    int n22 = n11 + a.n - a.n1_ - a.n_1;
    if ((n11 % 11) && n22) {
        if (n11 == a.n11 + 1) {
            a.p *= (double)(a.n1_ - a.n11) / n11 * (a.n_1 - a.n11) / n22;
            a.n11 = n11;
            return a.p;
        }
        if (n11 == a.n11 - 1) {
            a.p *= (double)a.n11 / (a.n1_ - n11) * (a.n11 + a.n - a.n1_ - a.n_1) / (a.n_1 - n11);
            a.n11 = n11;
            return a.p;
        }
    }
    a.n11 = n11;
    a.p = dm_hypergeo(lgam, a.n11, a.n1_, a.n_1, a.n);
    return a.p;
}

// kfunc.c:243-279 (two-sided p) + Algorithm.cpp:16-25.  Serial recurrence, warp-uniform.
__device__ __noinline__ double dm_fisher(const double* __restrict__ lgam, int n11, int n12, int n21, int n22)
{
    int n1_ = n11 + n12, n_1 = n11 + n21, n = n11 + n12 + n21 + n22;
    int hi = n_1 < n1_ ? n_1 : n1_;
    int lo = n1_ + n_1 - n;
    if (lo < 0) lo = 0;
    double two = 1.;
    if (lo != hi) {
        HgAcc a;
        a.n11 = n11; a.n1_ = n1_; a.n_1 = n_1; a.n = n;
        a.p = dm_hypergeo(lgam, n11, n1_, n_1, n);
        double q = a.p, left = 0., right = 0.;
        double p = dm_hg_move(lgam, a, lo);
        int i, j;
        for (i = lo + 1; p < 0.99999999 * q && i <= hi; ++i) { left += p; p = dm_hg_move(lgam, a, i); }
        if (p < 1.00000001 * q) left += p;
        p = dm_hg_move(lgam, a, hi);
        for (j = hi - 1; p < 0.99999999 * q && j >= 0; --j) { right += p; p = dm_hg_move(lgam, a, j); }
        if (p < 1.00000001 * q) right += p;
        two = left + right;
        if (two > 1.) two = 1.;
    }
    return dm_phred(two);
}

// src/BaseVarC.cpp:604-608, src/BaseType.cpp:193-197 (int products, as the reference)
__device__ __forceinline__ double dm_sor(int ref_fwd, int ref_rev, int alt_fwd, int alt_rev)
{
    if (alt_fwd * ref_rev > 0) return (double)(ref_fwd * alt_rev) / (ref_rev * alt_fwd);
    return 10000.0;
}

// Butterfly all-reduce: every lane ends with the same bits (fp add is commutative), fixed order.
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ long long warp_sum_ll(long long v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- warp-cooperative Fisher exact test -------------------------------------------------------------------------
// kt_fisher_exact's recurrence (kfunc.c:218-241) is re-anchored with lgamma whenever n11 % 11 == 0, so the chain of
// table probabilities falls into independent SEGMENTS of <= 11 terms: a direct (lgamma) value followed by incremental
// products.  Each lane takes one segment per round and produces bit-for-bit the p values of the serial walk; only the
// order in which the tail sums are added differs (per-lane partial sums + a butterfly), i.e. rounding in the last
// place.  The serial loop of kfunc.c:258-270 adds terms until the first one with p >= 0.99999999 q (index T, or the
// end of the range) and then adds p(T) iff p(T) < 1.00000001 q; that is what tail() returns.
struct FisherTail {
    const double* lgam; int n1_, n_1, n, lo, hi; double q;

    // the terms of segment `seg` of the walk starting at `lo` going up (dir = +1) or at `hi` going down (dir = -1)
    __device__ __forceinline__ void segment(int dir, int seg, double& sum, bool& stop, double& pstop) const
    {
        sum = 0.0; stop = false; pstop = 0.0;
        int first, last;                                          // inclusive, in walk order
        if (dir > 0) {
            first = seg == 0 ? lo : (lo - lo % 11) + 11 * seg;
            last = (first - first % 11) + 10;
            if (last > hi) last = hi;
            if (first > hi) return;
        } else {
            const int top = hi % 11 == 0 ? hi : hi - hi % 11 + 11;   // the first direct point is hi itself
            first = seg == 0 ? hi : top - 11 * seg;
            if (first < lo) return;
            last = first % 11 == 0 ? first - 10 : first - first % 11 + 1;
            if (last < lo) last = lo;
        }
        double p = dm_hypergeo(lgam, first, n1_, n_1, n);
        const double c = 0.99999999 * q;
        for (int i = first;; i += dir) {
            if (!(p < c)) { stop = true; pstop = p; return; }
            sum += p;
            if (i == last) return;
            const int nx = i + dir;
            const int n22 = nx + n - n1_ - n_1;
            if (dir > 0) p *= (double)(n1_ - i) / nx * (n_1 - i) / n22;                    // kfunc.c:224-225
            else         p *= (double)i / (n1_ - nx) * (i + n - n1_ - n_1) / (n_1 - nx);   // kfunc.c:230-231
        }
    }

    __device__ __forceinline__ double tail(int dir, int lane) const
    {
        double total = 0.0;
        const int n_seg = (hi - lo) / 11 + 2;
        for (int s0 = 0; s0 < n_seg; s0 += 32) {
            double sum, pstop; bool stop;
            segment(dir, s0 + lane, sum, stop, pstop);
            const unsigned m = __ballot_sync(0xffffffffu, stop);
            const int firststop = m ? __ffs(m) - 1 : 32;
            total += warp_sum(lane <= firststop ? sum : 0.0);
            if (m) {
                const double pT = __shfl_sync(0xffffffffu, pstop, firststop);
                if (pT < 1.00000001 * q) total += pT;
                return total;
            }
        }
        return total;                                             // ran off the range: every term was below q
    }
};

// same value on every lane; all 32 lanes must call it with the same table
__device__ __noinline__ double dm_fisher_warp(const double* __restrict__ lgam, int n11, int n12, int n21, int n22, int lane)
{
    FisherTail t;
    t.lgam = lgam; t.n1_ = n11 + n12; t.n_1 = n11 + n21; t.n = n11 + n12 + n21 + n22;
    t.hi = t.n_1 < t.n1_ ? t.n_1 : t.n1_;
    t.lo = t.n1_ + t.n_1 - t.n;
    if (t.lo < 0) t.lo = 0;
    double two = 1.;
    if (t.lo != t.hi) {
        t.q = dm_hypergeo(lgam, n11, t.n1_, t.n_1, t.n);
        two = t.tail(+1, lane) + t.tail(-1, lane);
        if (two > 1.) two = 1.;
    }
    return dm_phred(two);
}

}  // namespace bvc
