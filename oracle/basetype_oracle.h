/* oracle/basetype_oracle.h -- TEST INFRASTRUCTURE ONLY (see basetype_oracle.c). */
#ifndef BVC_ORACLE_H
#define BVC_ORACLE_H
#include "../include/bvc_gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Whole-tile CPU restatement: same inputs and outputs as bvc_enqueue+bvc_wait (+bvc_fetch_groups).
 * group_depth / group_af may be NULL.  Returns 0, or BVC_ERR_ARG on malformed input. */
int bvc_oracle_call_sites(const bvc_params* prm, const bvc_tile* tile, bvc_site_result* out,
                          int32_t* group_depth, double* group_af);

/* Same, over sites [s0, s1) only (used by bench.py to run bounded samples on several host threads). */
int bvc_oracle_call_range(const bvc_params* prm, const bvc_tile* tile, int32_t s0, int32_t s1,
                          bvc_site_result* out, int32_t* group_depth, double* group_af);

/* The statistical leaves, exposed so that each can be pinned against the reference objects. */
double bvc_oracle_chisf(double x, double k);
double bvc_oracle_normsf(double x);
double bvc_oracle_fisher(int n11, int n12, int n21, int n22);
double bvc_oracle_ranksum(const double* x, int32_t n1, const double* y, int32_t n2);

typedef struct bvc_oracle_lrt {
    int32_t ok, n_alt;
    int32_t alt_bases[4];
    int32_t depth[4];
    double  af_lrt[4];
    double  var_qual, depth_total, chi;
    int32_t em_fits, em_passes;
} bvc_oracle_lrt;

/* BaseType(bases, quals, ref, min_af) [+ SetBase(base_comb)] + LRT(). base_comb NULL = {0,1,2,3}. */
int bvc_oracle_lrt_reads(int32_t n, const int8_t* bases, const int8_t* quals, int32_t ref, double min_af,
                         const int8_t* base_comb, int32_t n_comb, bvc_oracle_lrt* out);

#ifdef __cplusplus
}
#endif
#endif
