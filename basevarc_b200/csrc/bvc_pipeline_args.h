// bvc_pipeline_args.h -- device-side intermediate records and the argument block of the staged pipeline
// (shared by bvc_pipeline.cuh and the host API, which only allocates them).
#pragma once
#include <stdint.h>
#include "bvc_kernels.h"

namespace bvc {

// ---- intermediate records (device only) ------------------------------------------------------------
struct __align__(16) SiteSum {      // 64 B, written by K1
    int32_t  depth[4];              // non-indel entries per base
    int32_t  fwd[4];                // of which forward strand
    int32_t  cvg[4];                // CVG strand table over ALL entries (ref_fwd, ref_rev, alt_fwd, alt_rev)
    uint32_t cls_off;               // first class of this site in the pool
    uint16_t ncls;
    uint8_t  alt;                   // top-count non-ref base
    uint8_t  has_q0;                // bit 0: a qual-0 class exists (IEEE-faithful legacy path); bit 1: a qual <= 1 class exists (exact-delta path)
    int32_t  n_entries;
    uint16_t n_major;               // classes of the major (top-depth) base: they come first in the pool
    uint8_t  major;                 // that base
    uint8_t  pad;
};

struct __align__(16) FitRes {       // 96 B, one per (site, slot)
    double  f[4];                   // allele frequencies used by the LAST E-step (its ml enters lr)
    double  eap[4];                 // M-step after it (what UpdateF reports, BaseType.cpp:61-69)
                                    // (an UNFINISHED fit parks its state here: f = next f, eap = previous f)
    double  lr;                     // sum_c cnt_c log ml_c(f)  (BaseType.cpp:57-60), written by fit_loglik_kernel
    int32_t passes;                 // singleEM passes done (unfinished: index of the pass to resume at)
    uint32_t flags;
    uint32_t mask;                  // allele subset (bit b = base b)
    uint32_t unfinished;            // 1: pass budget hit in em_fit_kernel, em_long_kernel finishes it
};

struct __align__(16) Ladder {       // 96 B per site
    double  lr_alt;
    double  frq[4];
    double  chi_t;
    double  cvg_fs, cvg_sor;        // K2's CVG Fisher / SOR, carried to the one full-record store in K4
    uint32_t cand4;                 // candidate bases, one per byte in ladder order (byte t = candidate t; 0xff = none)
    int8_t  n;                      // current number of candidate bases
    int8_t  k;                      // size of the subsets in flight (== n: the full model)
    int8_t  nfit;                   // fits in flight
    int8_t  pad0;
    int32_t em_fits, em_passes;
    uint32_t flags;
    int32_t pad1;
};

constexpr uint8_t kStActive = 0, kStDone = 1, kStLegacy = 2;

struct PipeArgs {
    int mean_entries;           // entries / sites of the tile (host side: picks K1's lanes-per-site)
    int packed_hist;            // 1: K1 in its packed four-sites-per-warp form (every site has < 65,536 entries)
    int pass_budget;            // passes a fit may spend in em_fit_kernel before it is parked for em_long_kernel
    int drain_min_pass;         // once the queue is empty, fits at or beyond this pass move to em_long_kernel ...
    int drain_lanes;            // ... provided at most this many lanes of the warp still hold a fit
    int refill_min;             // em_fit_kernel: lanes without a fit wait until this many of the warp's lanes want one
    int long_exact;             // 1: em_long_kernel evaluates the exact delta on every pass (A/B check of its bound-based decision)
    int em_min_fits;            // a stage with fewer fits than this skips em_fit_kernel (every fit starts in em_long_kernel)
    int cls_stride;             // site_hist8_kernel: a site's classes live at pool word site * cls_stride (no allocation atomic)
    SiteArgs s;                 // tile arrays, out, lut, lgam, thresholds
    SiteSum*  sum;              // [n_sites]
    uint32_t* cls;              // class pool: cnt << 9 | base << 7 | qual
    uint32_t* cls_cursor;       // pool allocation cursor
    Ladder*   lad;              // [n_sites]
    FitRes*   fit;              // [n_sites][4]
    uint8_t*  state;            // [n_sites]
    uint32_t* q_in;             // fit queue of this stage as K2/K4 filled it (site << 2 | slot), unsorted
    uint32_t* q_out;            // fit queue K4 fills for the next stage
    uint32_t* q_in_n;           // counts
    uint32_t* q_out_n;
    uint32_t* q_in_head;        // K3 refill cursor
    uint32_t* q_sorted;         // q_in counting-sorted by descending class count (lanes of a warp get equal-length fits)
    uint32_t* bins;             // [kNBins + 1] histogram of class counts, then scatter cursors
    uint32_t* q_long;           // fits that hit the pass budget
    uint32_t* q_long_n;
    uint32_t* q_long_head;      // K3b refill cursor
    unsigned long long* work;   // [4] diagnostics: class-passes done by em_fit / em_long, loglik class evaluations, fits
    uint32_t* vcf_list;         // LRT-positive sites
    uint32_t* vcf_n;
    uint32_t* legacy_list;      // sites for the IEEE-faithful warp-per-site kernel
    uint32_t* legacy_n;
    uint32_t* deep_list;        // sites whose CVG Fisher table is deep: K2 leaves them to the warp-cooperative kernel
    uint32_t* deep_n;           //   (aliases q_long: that queue is idle until the first EM stage)
};

// ---- sample-sharded mode (bvc_shard.cuh) ---------------------------------------------------------------
constexpr int kShardHdr = 16;              // words before the bins: ind[8], n_entries, 7 spare (64-byte header)
constexpr int kVcfHistWords = 6 * 256;

struct ShardArgs {
    PipeArgs p;                 // p.s: the LOCAL tile (entries of this rank's samples); intermediates of the owned sites
    uint32_t* dense;            // [world * Sr][B]
    int B, qlo, qhi;
    int rank, world, Sr;        // Sr = ceil(S / world) sites per rank
    int n_tile_sites;           // S
    int n_own;                  // owned sites: [rank * Sr, rank * Sr + n_own)
    const uint8_t* tile_ref;    // [S] ref_base of the whole tile (p.s.ref_base points at the owned slice)
    const uint32_t* tile_off;   // [S + 1] local CSR offsets of the whole tile
    uint32_t* err;              // bit 0: a qual outside [qlo, qhi]
    uint8_t*  mask_all;         // [world * Sr]
    uint32_t* pos_list;         // ascending tile-site indices of the LRT-positive sites
    uint32_t* pos_n;
    uint32_t* hv;               // [P][kVcfHistWords]
};

}  // namespace bvc
