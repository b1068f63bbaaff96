/* oracle/basetype_oracle.c -- TEST INFRASTRUCTURE ONLY.  Never linked into the product path.
 *
 * A plain-C, single-threaded, per-READ restatement of the reference's per-site `basetype` calling
 * core, written from the reference's behaviour (paths under /root/reference):
 *   BaseType ctor / SetAlleleFreq / UpdateF / LRT / combs_   src/BaseType.cpp:5-139,237-255
 *   EM / singleEM / delta_bylog / chisf / normsf /
 *   bt_fisher_exact / RankSumTest                           src/Algorithm.cpp:3-130
 *   kf_lgamma / kf_erfc / kf_gammaq / kt_fisher_exact        SeqLib/htslib/kfunc.c:37-141,195-279 (htslib 1.5)
 *   bt_f counting, CVG strand table, SOR                     src/BaseVarC.cpp:536-669
 *   WriteVcf numeric part                                    src/BaseType.cpp:149-197
 * It keeps the reference's operation ORDER (sequential sums over reads, `((l0+l1)+l2)+l3`, etc.) so
 * that, compiled with -ffp-contract=off against the same glibc libm, its doubles are bit-identical
 * to the reference objects'.  PARITY IS PINNED: tests/test_oracle.py checks every function
 * here against oracle/_ref/libref_oracle.so (the unmodified reference objects) whenever that
 * library is present, and tests/golden/ holds vectors generated from it (the reference itself has no
 * golden vectors or known-answer tests for this path).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "basetype_oracle.h"

#define NT 4                                /* NTYPE, src/BaseType.h:13 */
static const double kMLN10TO10 = -0.23025850929940458;   /* src/BaseType.h:10 */

/* ------------------------------------------------------------------------------------------- */
/* htslib kfunc.c restated                                                                     */
/* ------------------------------------------------------------------------------------------- */

/* kfunc.c:37-50 (AS245): eight partial fractions added from z+7 down to z, then the constant. */
static double o_lgamma_as245(double z)
{
    static const double c[8] = { 0.1659470187408462e-06, 0.9934937113930748e-05, -0.1385710331296526,
                                 12.50734324009056, -176.6150291498386, 771.3234287757674,
                                 -1259.139216722289, 676.5203681218835 };
    double x = 0;
    for (int i = 0; i < 8; ++i) x += c[i] / (z + (7 - i));
    x += 0.9999999999995183;
    return log(x) - 5.58106146679532777 - z + (z - 0.5) * log(z + 6.5);
}

/* kfunc.c:56-82 (AS66) */
static double o_erfc(double x)
{
    static const double P[7] = { 220.2068679123761, 221.2135961699311, 112.0792914978709, 33.912866078383,
                                 6.37396220353165, .7003830644436881, .03526249659989109 };
    static const double Q[8] = { 440.4137358247522, 793.8265125199484, 637.3336333788311, 296.5642487796737,
                                 86.78073220294608, 16.06417757920695, 1.755667163182642, .08838834764831844 };
    double z = fabs(x) * M_SQRT2, e, p;
    if (z > 37.) return x > 0. ? 0. : 2.;
    e = exp(z * z * -.5);
    if (z < 10. / M_SQRT2) {
        double num = P[6], den = Q[7];
        for (int i = 5; i >= 0; --i) num = num * z + P[i];
        for (int i = 6; i >= 0; --i) den = den * z + Q[i];
        p = e * num / den;
    } else {
        p = e / 2.506628274631001 / (z + 1. / (z + 2. / (z + 3. / (z + 4. / (z + .65)))));
    }
    return x > 0. ? 2. * p : 2. * (1. - p);
}

#define O_GAMMA_EPS 1e-14
#define O_TINY 1e-290

/* kfunc.c:101-110: series for the regularised lower incomplete gamma */
static double o_gammap_series(double s, double z)
{
    double sum = 1., x = 1.;
    for (int k = 1; k < 100; ++k) {
        x *= z / (s + k);
        sum += x;
        if (x / sum < O_GAMMA_EPS) break;
    }
    return exp(s * log(z) - z - o_lgamma_as245(s + 1.) + log(sum));
}

/* kfunc.c:112-131: modified Lentz continued fraction for the upper incomplete gamma */
static double o_gammaq_cf(double s, double z)
{
    double f = 1. + z - s, C = f, D = 0.;
    for (int j = 1; j < 100; ++j) {
        double a = j * (s - j), b = (j << 1) + 1 + z - s, d;
        D = b + a * D;
        if (D < O_TINY) D = O_TINY;
        C = b + a / C;
        if (C < O_TINY) C = O_TINY;
        D = 1. / D;
        d = C * D;
        f *= d;
        if (fabs(d - 1.) < O_GAMMA_EPS) break;
    }
    return exp(s * log(z) - z - o_lgamma_as245(s) - log(f));
}

/* kfunc.c:138-141 */
static double o_gammaq(double s, double z)
{
    return (z <= 1. || z < s) ? 1. - o_gammap_series(s, z) : o_gammaq_cf(s, z);
}

/* kfunc.c:195-199 -- libm lgamma, as the reference */
static double o_lbinom(int n, int k)
{
    if (k == 0 || n == k) return 0;
    return lgamma(n + 1) - lgamma(k + 1) - lgamma(n - k + 1);
}

/* kfunc.c:207-210 */
static double o_hypergeo(int n11, int n1_, int n_1, int n)
{
    return exp(o_lbinom(n1_, n11) + o_lbinom(n - n1_, n_1 - n11) - o_lbinom(n, n_1));
}

typedef struct { int n11, n1_, n_1, n; double p; } o_hg;

/* kfunc.c:218-241: step the table by one cell with the ratio recurrence, re-anchor via lgamma when
 * n11 is a multiple of 11 or the n22 cell is zero, or when the move is not a unit step. */
static double o_hg_first(o_hg* a, int n11, int n1_, int n_1, int n)
{
    a->n11 = n11; a->n1_ = n1_; a->n_1 = n_1; a->n = n;
    a->p = o_hypergeo(n11, n1_, n_1, n);
    return a->p;
}
static double o_hg_move(o_hg* a, int n11)
{
    int n22 = n11 + a->n - a->n1_ - a->n_1;
    if (n11 % 11 && n22) {
        if (n11 == a->n11 + 1) {
            a->p *= (double)(a->n1_ - a->n11) / n11 * (a->n_1 - a->n11) / n22;
            a->n11 = n11;
            return a->p;
        }
        if (n11 == a->n11 - 1) {
            a->p *= (double)a->n11 / (a->n1_ - n11) * (a->n11 + a->n - a->n1_ - a->n_1) / (a->n_1 - n11);
            a->n11 = n11;
            return a->p;
        }
    }
    a->n11 = n11;
    a->p = o_hypergeo(a->n11, a->n1_, a->n_1, a->n);
    return a->p;
}

/* kfunc.c:243-279, two-sided p only */
static double o_fisher_two_sided(int n11, int n12, int n21, int n22)
{
    int n1_ = n11 + n12, n_1 = n11 + n21, n = n11 + n12 + n21 + n22;
    int hi = n_1 < n1_ ? n_1 : n1_;
    int lo = n1_ + n_1 - n;
    int i, j;
    double p, q, left, right, two;
    o_hg a;
    if (lo < 0) lo = 0;
    if (lo == hi) return 1.;
    /* NB: the reference calls hypergeo_acc(n11, n1_, n_1, n): with all three margins zero that call
     * would take the "only n11 changed" branch; the tables reaching here never have n == 0. */
    q = o_hg_first(&a, n11, n1_, n_1, n);
    p = o_hg_move(&a, lo);
    for (left = 0., i = lo + 1; p < 0.99999999 * q && i <= hi; ++i) { left += p; p = o_hg_move(&a, i); }
    --i;
    if (p < 1.00000001 * q) left += p; else --i;
    p = o_hg_move(&a, hi);
    for (right = 0., j = hi - 1; p < 0.99999999 * q && j >= 0; --j) { right += p; p = o_hg_move(&a, j); }
    ++j;
    if (p < 1.00000001 * q) right += p; else ++j;
    two = left + right;
    if (two > 1.) two = 1.;
    (void)i; (void)j;
    return two;
}

/* ------------------------------------------------------------------------------------------- */
/* src/Algorithm.cpp restated                                                                  */
/* ------------------------------------------------------------------------------------------- */

double bvc_oracle_chisf(double x, double k) { return o_gammaq(k / 2.0, x / 2.0); }      /* :3-7  */
double bvc_oracle_normsf(double x) { return o_erfc(x / sqrt(2.0)) / 2.0; }               /* :9-14 */

static double o_phred(double p)     /* the shared tail of :20-23 and :62-65 */
{
    double v = -10 * log10(p);
    if (isinf(v)) v = 10000.0;
    else if (v == 0) v = 0.0;
    return v;
}

double bvc_oracle_fisher(int n11, int n12, int n21, int n22)                              /* :16-25 */
{
    return o_phred(o_fisher_two_sided(n11, n12, n21, n22));
}

/* Algorithm.cpp:27-67.  Values are sorted DESCENDING; tied runs share the mean of their 1-based
 * ranks; r1 = rank sum of the first n1 values.  Every term is a multiple of 1/2 far below 2^52, so
 * the sum is exact whatever the (unstable) sort order of the reference was. */
typedef struct { double v; int32_t from_x; } o_rk;
static int o_rk_desc(const void* a, const void* b)
{
    double x = ((const o_rk*)a)->v, y = ((const o_rk*)b)->v;
    return (x < y) - (x > y);
}
double bvc_oracle_ranksum(const double* x, int32_t n1, const double* y, int32_t n2)
{
    int32_t s = n1 + n2;
    o_rk* r = (o_rk*)malloc(sizeof(o_rk) * (size_t)(s > 0 ? s : 1));
    double r1 = 0.0;
    for (int32_t i = 0; i < n1; ++i) { r[i].v = x[i]; r[i].from_x = 1; }
    for (int32_t i = 0; i < n2; ++i) { r[n1 + i].v = y[i]; r[n1 + i].from_x = 0; }
    qsort(r, (size_t)s, sizeof(o_rk), o_rk_desc);
    for (int32_t i = 0; i < s;) {
        int32_t e = i, ksum = 0, nx = 0;
        while (e < s && r[e].v == r[i].v) { ksum += e + 1; nx += r[e].from_x; ++e; }   /* int32 sum as :32,38,42 */
        double avg = (e - i > 1) ? (double)ksum / (e - i) : (double)(i + 1);
        for (int32_t t = 0; t < nx; ++t) r1 += avg;
        i = e;
    }
    free(r);
    size_t a = (size_t)n1, b = (size_t)n2;
    double expected = (double)(a * (a + b + 1)) / 2.0;
    double z = (r1 - expected) / sqrt((double)(a * b * (a + b + 1)) / 12.0);
    return o_phred(2 * bvc_oracle_normsf(fabs(z)));
}

/* ------------------------------------------------------------------------------------------- */
/* EM (src/Algorithm.cpp:69-130), per read                                                     */
/* ------------------------------------------------------------------------------------------- */

/* One E+M pass (:69-93).  ml[] must be zero on entry, eap[] must be zero on entry. `post` is the
 * reference's column-major scratch (n*4). */
static void o_single_em(const double* f, const double* L, double* ml, double* eap, double* post, int32_t n)
{
    double lik[NT];
    for (int32_t i = 0; i < n; ++i) {
        for (int j = 0; j < NT; ++j) { lik[j] = f[j] * L[(size_t)i * NT + j]; ml[i] += lik[j]; }
        for (int j = 0; j < NT; ++j) post[(size_t)j * n + i] = lik[j] / ml[i];
    }
    for (int j = 0; j < NT; ++j) {
        for (int32_t i = 0; i < n; ++i) eap[j] += post[(size_t)j * n + i];
        eap[j] = eap[j] / n;
    }
}

/* :115-130.  Returns the number of E+M passes done (a diagnostic the reference does not keep). */
static int o_em(double* f, const double* L, double* ml, double* eap, double* ml2, double* post,
                int32_t n, int iter_num, double eps)
{
    int passes = 1;
    o_single_em(f, L, ml, eap, post, n);
    for (int it = 0; it < iter_num; ++it) {
        for (int j = 0; j < NT; ++j) { f[j] = eap[j]; eap[j] = 0.0; }                    /* :95-101 */
        o_single_em(f, L, ml2, eap, post, n);
        ++passes;
        double delta = 0.0;                                                                /* :103-113 */
        for (int32_t i = 0; i < n; ++i) {
            delta += fabs(log(ml2[i]) - log(ml[i]));
            ml[i] = ml2[i];
            ml2[i] = 0.0;
        }
        if (delta < eps) break;
    }
    return passes;
}

/* ------------------------------------------------------------------------------------------- */
/* BaseType (src/BaseType.cpp) per read                                                        */
/* ------------------------------------------------------------------------------------------- */

typedef struct {
    int32_t n;
    const int8_t* bases; const int8_t* quals;
    int32_t depth[NT];
    double depth_total;
    double* L;          /* n*4 likelihood table (:9-20) */
    double *ml, *ml2, *post;
    int em_fits, em_passes;
    double eps; int iters; double lrt_thr;
} o_bt;

/* k-subsets of items[0..n) in the order std::prev_permutation visits a k-ones bitmask (:237-255):
 * lexicographically increasing index tuples. */
static int o_combs(const int8_t* items, int n, int k, int8_t out[][NT])
{
    int idx[NT], cnt = 0;
    for (int i = 0; i < k; ++i) idx[i] = i;
    for (;;) {
        for (int i = 0; i < k; ++i) out[cnt][i] = items[idx[i]];
        ++cnt;
        int i = k - 1;
        while (i >= 0 && idx[i] == n - k + i) --i;
        if (i < 0) break;
        ++idx[i];
        for (int t = i + 1; t < k; ++t) idx[t] = idx[t - 1] + 1;
    }
    return cnt;
}

/* UpdateF (:41-71) for every k-subset of `cand`; fills lr[] and bp[][4]; returns the subset count.
 * freq_sum == 0 cannot happen for min_af > 0 and is rejected upstream. */
static int o_update_f(o_bt* bt, const int8_t* cand, int n_cand, int k, int8_t bc[][NT], double* lr, double bp[][NT])
{
    int nc = o_combs(cand, n_cand, k, bc);
    for (int c = 0; c < nc; ++c) {
        double f[NT] = {0, 0, 0, 0}, eap[NT] = {0, 0, 0, 0};
        int32_t dsum = 0;                                                                  /* :25-39 */
        for (int t = 0; t < k; ++t) dsum += bt->depth[bc[c][t]];
        if (dsum > 0) for (int t = 0; t < k; ++t) f[bc[c][t]] = (double)bt->depth[bc[c][t]] / dsum;
        memset(bt->ml, 0, sizeof(double) * (size_t)bt->n);
        memset(bt->ml2, 0, sizeof(double) * (size_t)bt->n);
        bt->em_passes += o_em(f, bt->L, bt->ml, eap, bt->ml2, bt->post, bt->n, bt->iters, bt->eps);
        bt->em_fits += 1;
        double s = 0;
        for (int32_t i = 0; i < bt->n; ++i) s += log(bt->ml[i]);
        lr[c] = s;
        for (int j = 0; j < NT; ++j) bp[c][j] = eap[j];
    }
    return nc;
}

static void o_lrt(o_bt* bt, int32_t ref, double min_af, const int8_t* base_comb, int32_t n_comb, bvc_oracle_lrt* out)
{
    memset(out, 0, sizeof(*out));
    for (int j = 0; j < NT; ++j) out->depth[j] = bt->depth[j];
    out->depth_total = bt->depth_total;
    if (bt->depth_total == 0) return;                                                      /* :75 */
    int8_t cand[NT]; int n = 0;
    for (int t = 0; t < n_comb; ++t)                                                       /* :77-82 */
        if ((bt->depth[base_comb[t]] / bt->depth_total) >= min_af) cand[n++] = base_comb[t];
    if (n == 0) return;
    int8_t bc[6][NT]; double lr[6]; double bp[6][NT];
    o_update_f(bt, cand, n, n, bc, lr, bp);                                                /* :88 */
    double frq[NT]; memcpy(frq, bp[0], sizeof(frq));
    double lr_alt = lr[0], chi_t = 0.0;
    for (int k = n - 1; k > 0; --k) {                                                      /* :93-110 */
        int nc = o_update_f(bt, cand, n, k, bc, lr, bp);
        int imin = 0; double cmin = 2.0 * (lr_alt - lr[0]);
        for (int c = 1; c < nc; ++c) { double chi = 2.0 * (lr_alt - lr[c]); if (chi < cmin) { cmin = chi; imin = c; } }
        lr_alt = lr[imin];
        chi_t = cmin;
        if (chi_t < bt->lrt_thr) {
            memcpy(cand, bc[imin], (size_t)k); n = k;
            memcpy(frq, bp[imin], sizeof(frq));
        } else break;
    }
    for (int t = 0; t < n; ++t)                                                            /* :111-116 */
        if (cand[t] != ref) { out->alt_bases[out->n_alt++] = cand[t]; out->af_lrt[cand[t]] = frq[cand[t]]; }
    out->chi = chi_t;
    out->em_fits = bt->em_fits; out->em_passes = bt->em_passes;
    if (out->n_alt == 0) return;
    out->ok = 1;
    double r = bt->depth[cand[0]] / bt->depth_total;                                       /* :118-134 */
    if (n == 1 && bt->depth_total > 10 && r > 0.5) { out->var_qual = 5000.0; return; }
    if (chi_t <= 0) { out->var_qual = 0.0; return; }
    double p = bvc_oracle_chisf(chi_t, 1.0);
    out->var_qual = (p != 0) ? -10 * log10(p) : 10000;       /* `if (chi_prob)`: NaN is true */
    if (out->var_qual == 0) out->var_qual = 0.0;
}

/* The reference's compile-time constants (epsilon 0.001 and 100 iterations, BaseType.cpp:45-46; LRT_THRESHOLD 24, BaseType.h:9)
 * as arguments: bvc_params carries them (bvc_default_params sets the reference's values) and the threshold-boundary tests move
 * them onto a site's own delta / chi. */
static int o_lrt_reads(int32_t n, const int8_t* bases, const int8_t* quals, int32_t ref, double min_af,
                       const int8_t* base_comb, int32_t n_comb, double eps, int iters, double lrt_thr, bvc_oracle_lrt* out);

int bvc_oracle_lrt_reads(int32_t n, const int8_t* bases, const int8_t* quals, int32_t ref, double min_af,
                         const int8_t* base_comb, int32_t n_comb, bvc_oracle_lrt* out)
{
    return o_lrt_reads(n, bases, quals, ref, min_af, base_comb, n_comb, 0.001, 100, 24.0, out);
}

static int o_lrt_reads(int32_t n, const int8_t* bases, const int8_t* quals, int32_t ref, double min_af,
                       const int8_t* base_comb, int32_t n_comb, double eps, int iters, double lrt_thr, bvc_oracle_lrt* out)
{
    static const int8_t all4[NT] = {0, 1, 2, 3};
    o_bt bt; memset(&bt, 0, sizeof(bt));
    bt.n = n; bt.bases = bases; bt.quals = quals;
    bt.eps = eps; bt.iters = iters; bt.lrt_thr = lrt_thr;
    size_t m = (size_t)(n > 0 ? n : 1);
    bt.L = (double*)malloc(sizeof(double) * m * NT);
    bt.ml = (double*)calloc(m, sizeof(double)); bt.ml2 = (double*)calloc(m, sizeof(double));
    bt.post = (double*)malloc(sizeof(double) * m * NT);
    for (int32_t i = 0; i < n; ++i) {                                                      /* :9-20 */
        double e = exp(kMLN10TO10 * quals[i]);
        for (int j = 0; j < NT; ++j) bt.L[(size_t)i * NT + j] = (bases[i] == j) ? 1.0 - e : e / 3.0;
        if (bases[i] < 0 || bases[i] >= NT) { free(bt.L); free(bt.ml); free(bt.ml2); free(bt.post); return BVC_ERR_ARG; }
        bt.depth[bases[i]] += 1;
    }
    for (int j = 0; j < NT; ++j) bt.depth_total += bt.depth[j];                           /* :21-23 */
    if (!base_comb) { base_comb = all4; n_comb = NT; }
    o_lrt(&bt, ref, min_af, base_comb, n_comb, out);
    free(bt.L); free(bt.ml); free(bt.ml2); free(bt.post);
    return out->ok;
}

/* ------------------------------------------------------------------------------------------- */
/* bt_f (src/BaseVarC.cpp:536-669) + WriteVcf numeric part (src/BaseType.cpp:149-197) per site */
/* ------------------------------------------------------------------------------------------- */

static double o_sor(int ref_fwd, int ref_rev, int alt_fwd, int alt_rev)   /* BaseVarC.cpp:604-608, BaseType.cpp:193-197 */
{
    if (alt_fwd * ref_rev > 0) return (double)(ref_fwd * alt_rev) / (ref_rev * alt_fwd);
    return 10000.0;
}

static int o_in(const int32_t* set, int n, int v) { for (int i = 0; i < n; ++i) if (set[i] == v) return 1; return 0; }

static int o_site(const bvc_params* prm, const bvc_tile* t, int32_t s, bvc_site_result* r,
                  int32_t* gdepth, double* gaf, int8_t* bases, int8_t* quals, double* fa, double* fb)
{
    const uint32_t b0 = t->entry_off[s], b1 = t->entry_off[s + 1];
    const int32_t ref = t->ref_base[s];
    int32_t n = 0;
    memset(r, 0, sizeof(*r));
    r->n_entries = (int32_t)(b1 - b0);
    if (ref > 3) return BVC_ERR_ARG;
    for (uint32_t e = b0; e < b1; ++e) {                                   /* BaseVarC.cpp:550-567 */
        uint8_t c = t->code[e];
        if (BVC_CODE_INDEL(c)) continue;
        if (BVC_CODE_BASE(c) > 3 || t->qual[e] > 127) return BVC_ERR_ARG;
        r->depth[BVC_CODE_BASE(c)] += 1;
        bases[n] = (int8_t)BVC_CODE_BASE(c); quals[n] = (int8_t)t->qual[e]; ++n;
    }
    /* :576-583 descending argsort, ties in index order; top non-ref base */
    int order[4] = {0, 1, 2, 3};
    for (int i = 1; i < 4; ++i)
        for (int j = i; j > 0 && r->depth[order[j]] > r->depth[order[j - 1]]; --j) { int x = order[j]; order[j] = order[j - 1]; order[j - 1] = x; }
    int alt = (order[0] != ref) ? order[0] : order[1];
    r->cvg_alt_base = (int8_t)alt;
    for (uint32_t e = b0; e < b1; ++e) {                                   /* :587-602, ALL entries (Q1) */
        uint8_t c = t->code[e];
        int base = (int)BVC_CODE_BASE(c), fwd = (int)BVC_CODE_STRAND(c);
        if (base == ref) r->cvg_strand[fwd ? 0 : 1] += 1;
        else if (base == alt) r->cvg_strand[fwd ? 2 : 3] += 1;
    }
    r->cvg_fs = bvc_oracle_fisher(r->cvg_strand[0], r->cvg_strand[1], r->cvg_strand[2], r->cvg_strand[3]);
    r->cvg_sor = o_sor(r->cvg_strand[0], r->cvg_strand[1], r->cvg_strand[2], r->cvg_strand[3]);

    bvc_oracle_lrt L;                                                       /* :612-613 */
    o_lrt_reads(n, bases, quals, ref, prm->min_af, NULL, 0, prm->em_epsilon, prm->em_max_iter, prm->lrt_threshold, &L);
    r->lrt_ok = (int8_t)L.ok; r->n_alt = (int8_t)L.n_alt;
    for (int i = 0; i < 3; ++i) r->alt_bases[i] = (int8_t)(i < L.n_alt ? L.alt_bases[i] : -1);
    for (int j = 0; j < 4; ++j) r->af_lrt[j] = L.af_lrt[j];
    r->chi = L.chi; r->var_qual = L.var_qual;
    r->em_fits = L.em_fits; r->em_passes = L.em_passes;

    if (prm->n_groups > 0 && gdepth && gaf) {                               /* :617-661 */
        const int G = prm->n_groups;
        int8_t comb[4]; int nc = 0;
        comb[nc++] = (int8_t)ref;
        for (int i = 0; i < L.n_alt; ++i) comb[nc++] = (int8_t)L.alt_bases[i];
        for (int g = 0; g < G; ++g) {
            int32_t* gd = gdepth + ((size_t)s * G + g) * 4;
            double* ga = gaf + ((size_t)s * G + g) * 4;
            int32_t m = 0;
            for (int j = 0; j < 4; ++j) { gd[j] = 0; ga[j] = 0.0; }
            for (uint32_t e = b0; e < b1; ++e) {
                uint8_t c = t->code[e];
                if (BVC_CODE_INDEL(c)) continue;
                if (prm->sample_group[t->sample[e]] != g) continue;
                gd[BVC_CODE_BASE(c)] += 1;
                bases[m] = (int8_t)BVC_CODE_BASE(c); quals[m] = (int8_t)t->qual[e]; ++m;
            }
            if (L.ok && m > 0) {
                bvc_oracle_lrt GL;
                o_lrt_reads(m, bases, quals, ref, prm->min_af, comb, nc, prm->em_epsilon, prm->em_max_iter, prm->lrt_threshold, &GL);
                for (int i = 0; i < L.n_alt; ++i) {
                    int b = L.alt_bases[i];
                    ga[b] = o_in(GL.alt_bases, GL.n_alt, b) ? GL.af_lrt[b] : -1.0;   /* -1 = printed as "0" (:651) */
                }
            }
        }
    }
    if (!L.ok) return 0;

    /* WriteVcf: BaseType.cpp:152-197, entries in ascending sample order */
    for (int f = 0; f < 3; ++f) {
        const uint8_t* feat = f == 0 ? t->mapq : (f == 1 ? t->qual : t->rpr);
        int32_t na = 0, nb = 0;
        for (uint32_t e = b0; e < b1; ++e) {
            uint8_t c = t->code[e];
            int base = (int)BVC_CODE_BASE(c);
            if (BVC_CODE_INDEL(c) || base == 4) continue;
            if (base == ref) fa[na++] = feat[e];
            else if (o_in(L.alt_bases, L.n_alt, base)) fb[nb++] = feat[e];
        }
        r->ranksum[f] = bvc_oracle_ranksum(fa, na, fb, nb);
    }
    for (uint32_t e = b0; e < b1; ++e) {
        uint8_t c = t->code[e];
        int base = (int)BVC_CODE_BASE(c), fwd = (int)BVC_CODE_STRAND(c);
        if (BVC_CODE_INDEL(c) || base == 4) continue;
        if (base == ref) r->vcf_strand[fwd ? 0 : 1] += 1;
        else if (o_in(L.alt_bases, L.n_alt, base)) r->vcf_strand[fwd ? 2 : 3] += 1;
    }
    r->vcf_fs = bvc_oracle_fisher(r->vcf_strand[0], r->vcf_strand[1], r->vcf_strand[2], r->vcf_strand[3]);
    r->vcf_sor = o_sor(r->vcf_strand[0], r->vcf_strand[1], r->vcf_strand[2], r->vcf_strand[3]);
    return 0;
}

int bvc_oracle_call_range(const bvc_params* prm, const bvc_tile* tile, int32_t s0, int32_t s1,
                          bvc_site_result* out, int32_t* group_depth, double* group_af)
{
    if (!prm || !tile || !out || s0 < 0 || s1 > tile->n_sites) return BVC_ERR_ARG;
    if (prm->n_groups > 0 && (!prm->sample_group || !tile->sample)) return BVC_ERR_ARG;
    uint32_t maxn = 1;
    for (int32_t s = s0; s < s1; ++s) { uint32_t m = tile->entry_off[s + 1] - tile->entry_off[s]; if (m > maxn) maxn = m; }
    int8_t* bases = (int8_t*)malloc(maxn); int8_t* quals = (int8_t*)malloc(maxn);
    double* fa = (double*)malloc(sizeof(double) * maxn); double* fb = (double*)malloc(sizeof(double) * maxn);
    int rc = 0;
    for (int32_t s = s0; s < s1 && rc == 0; ++s)
        rc = o_site(prm, tile, s, &out[s], group_depth, group_af, bases, quals, fa, fb);
    free(bases); free(quals); free(fa); free(fb);
    return rc;
}

int bvc_oracle_call_sites(const bvc_params* prm, const bvc_tile* tile, bvc_site_result* out,
                          int32_t* group_depth, double* group_af)
{
    if (!tile) return BVC_ERR_ARG;
    return bvc_oracle_call_range(prm, tile, 0, tile->n_sites, out, group_depth, group_af);
}
