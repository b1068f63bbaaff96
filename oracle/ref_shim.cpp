// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE ONLY.
//
// A thin extern "C" shim (ours) over the UNMODIFIED reference objects BaseType.o, Algorithm.o
// and htslib's kfunc.o, which oracle/Makefile compiles straight from /root/reference.  It lets
// the test-suite call the reference's own functions at full double precision through ctypes:
//   BaseType::BaseType + SetBase + LRT      (src/BaseType.cpp:5-23, BaseType.h:67, BaseType.cpp:73-139)
//   bt_fisher_exact / RankSumTest / chisf / normsf   (src/Algorithm.cpp:3-67)
//   EM                                       (src/Algorithm.cpp:115-130)
//   WriteVcf                                 (src/BaseType.cpp:141-234)
// No reference code is copied: this file only #includes the reference headers where they lie.
#include <cstring>
#include <string>
#include <vector>
#include "BaseType.h"

extern "C" {

struct ref_lrt_out {
    int32_t ok;
    int32_t n_alt;
    int32_t alt_bases[4];
    int32_t depth[4];
    double af_lrt[4];      // indexed by base code; 0 when absent
    double var_qual;
    double depth_total;
};

// base_comb == NULL -> the default {0,1,2,3}; otherwise SetBase(base_comb[0..n_comb)) (group path).
int ref_lrt(int32_t n, const int8_t* bases, const int8_t* quals, int32_t ref, double min_af,
            const int8_t* base_comb, int32_t n_comb, ref_lrt_out* out)
{
    BaseV b(bases, bases + n), q(quals, quals + n);
    BaseType bt(b, q, (int8_t)ref, min_af);
    if (base_comb) bt.SetBase(BaseV(base_comb, base_comb + n_comb));
    bool ok = bt.LRT();
    std::memset(out, 0, sizeof(*out));
    out->ok = ok ? 1 : 0;
    out->n_alt = (int32_t)bt.alt_bases.size();
    for (size_t i = 0; i < bt.alt_bases.size() && i < 4; ++i) out->alt_bases[i] = bt.alt_bases[i];
    for (int j = 0; j < 4; ++j) out->depth[j] = bt.depth.at(j);
    for (auto const& kv : bt.af_lrt) if (kv.first >= 0 && kv.first < 4) out->af_lrt[kv.first] = kv.second;
    out->var_qual = bt.var_qual;
    out->depth_total = bt.depth_total;
    return out->ok;
}

double ref_fisher(int n11, int n12, int n21, int n22) { return bt_fisher_exact(n11, n12, n21, n22); }
double ref_chisf(double x, double k) { return chisf(x, k); }
double ref_normsf(double x) { return normsf(x); }

double ref_ranksum(const double* x, int32_t n1, const double* y, int32_t n2)
{
    std::vector<double> vx(x, x + n1), vy(y, y + n2);
    vx.reserve(n1 + n2 + 1);   // keeps the reference's idx[i+1] over-read (Algorithm.cpp:36) inside the allocation
    return RankSumTest(vx, vy);
}

// One reference EM fit (all vectors in the reference's layout; ml/eap must be zero on entry).
void ref_em(double* f4, const double* lik /* n*4 */, double* ml /* n */, double* eap4,
            int32_t n, int32_t iter_num, double eps)
{
    std::vector<double> f(f4, f4 + 4), L(lik, lik + (size_t)n * 4), m(n, 0.0), e(4, 0.0);
    EM(f, L, m, e, n, 4, iter_num, eps);
    std::memcpy(f4, f.data(), 4 * sizeof(double));
    std::memcpy(ml, m.data(), (size_t)n * sizeof(double));
    std::memcpy(eap4, e.data(), 4 * sizeof(double));
}

// WriteVcf on a site given as parallel arrays (entry k of the site belongs to sample `sample[k]`).
// Returns the record length; the text is copied into buf (truncated to cap-1).
int ref_writevcf(int32_t n, const uint8_t* base, const uint8_t* mapq, const uint8_t* qual, const uint8_t* rpr,
                 const uint8_t* strand, const uint8_t* is_indel, const int32_t* sample,
                 int32_t N, int32_t ref, double min_af, const char* chr, int32_t pos, char* buf, int32_t cap)
{
    AlleleInfoVector aiv;
    DepM idx;
    BaseV b, q;
    for (int32_t k = 0; k < n; ++k) {
        AlleleInfo a;
        a.base = base[k]; a.mapq = mapq[k]; a.qual = qual[k]; a.rpr = rpr[k];
        a.strand = strand[k]; a.is_indel = is_indel[k];
        if (a.is_indel) a.indel = "+A";
        aiv.push_back(a);
        idx.insert({sample[k], k});
        if (!a.is_indel) { b.push_back((int8_t)a.base); q.push_back((int8_t)a.qual); }
    }
    BaseType bt(b, q, (int8_t)ref, min_af);
    if (!bt.LRT()) { if (cap > 0) buf[0] = 0; return 0; }
    InfoM info;
    String s = WriteVcf(bt, chr, pos, (int8_t)ref, aiv, idx, info, N);
    int32_t len = (int32_t)s.size();
    if (cap > 0) { int32_t m = len < cap - 1 ? len : cap - 1; std::memcpy(buf, s.data(), m); buf[m] = 0; }
    return len;
}

}  // extern "C"
