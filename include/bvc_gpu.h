/* include/bvc_gpu.h -- the C-ABI drop-in boundary of the B200 `basetype` calling core.
 *
 * The reference (Zilong-Li/BaseVarC) has no FFI: its per-site caller is reached by direct C++
 * calls inside one translation unit.  This header is the boundary a maintainer binds instead of
 * those calls; each entry point names the reference code it replaces (paths under /root/reference):
 *
 *   bvc_create / bvc_destroy     the per-thread state of bt_s            src/BaseVarC.cpp:316-398
 *                                + the constants of BaseType/UpdateF     src/BaseType.h:9-13, src/BaseType.cpp:45-46
 *                                + min_af of bt_f                        src/BaseVarC.cpp:541-543
 *   bvc_enqueue / bvc_wait       the body of the site loop: one bt_f()   src/BaseVarC.cpp:442-443, 536-669
 *                                per site, i.e. BaseType ctor + LRT()    src/BaseType.cpp:5-139
 *                                + the numeric part of WriteVcf()        src/BaseType.cpp:149-197
 *                                + Algorithm.cpp / kfunc.c underneath    src/Algorithm.cpp:3-130
 *   bvc_tile                     AlleleInfoVector + DepM idx per site    src/BamProcess.h:30-39, src/BaseType.h:21
 *   bvc_site_result              what bt_f/WriteVcf format into text     src/BaseVarC.cpp:609-610, src/BaseType.cpp:198-233
 *
 * Plain C: no STL, no exceptions, no torch types.  All host memory is caller-owned.  A context is
 * bound to one GPU and is NOT thread-safe (one dispatcher thread per GPU).  There is no CPU
 * fallback: every call fails with BVC_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef BVC_GPU_H
#define BVC_GPU_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BVC_ABI_VERSION 3

/* error codes (negative); 0 = success */
#define BVC_OK             0
#define BVC_ERR_ARG       -1   /* NULL pointer, bad size, qual > 127, non-indel base > 3, ... */
#define BVC_ERR_CUDA      -2   /* CUDA runtime error or no usable device (see bvc_last_error) */
#define BVC_ERR_CAPACITY  -3   /* tile larger than the capacity given at bvc_create */
#define BVC_ERR_STATE     -4   /* unknown ticket / nothing in flight */

/* entry code byte: base:3 | strand:1 | is_indel:1 (AlleleInfo, src/BamProcess.h:30-39).
 * For an indel entry base/strand (and qual/mapq/rpr) carry the STALE values of the last numeric
 * token, exactly as the reference's reused `ai` does (src/BaseVarC.cpp:393,432-435). */
#define BVC_CODE_BASE(c)    ((c) & 7u)
#define BVC_CODE_STRAND(c)  (((c) >> 3) & 1u)
#define BVC_CODE_INDEL(c)   (((c) >> 4) & 1u)
#define BVC_MAKE_CODE(base, strand, indel) ((uint8_t)(((base) & 7u) | (((strand) & 1u) << 3) | (((indel) & 1u) << 4)))

typedef struct bvc_params {
    int32_t abi_version;        /* = BVC_ABI_VERSION */
    int32_t n_samples;          /* N = lines of bam.list (src/BaseVarC.cpp:189); upper bound of entries per site */
    double  min_af;             /* min(100.0/N, 0.001, --maf)   src/BaseVarC.cpp:541-543 ; must be > 0 */
    double  lrt_threshold;      /* LRT_THRESHOLD 24.0           src/BaseType.h:9  */
    double  em_epsilon;         /* 0.001                        src/BaseType.cpp:45 */
    int32_t em_max_iter;        /* 100                          src/BaseType.cpp:46 */
    int32_t n_groups;           /* 0 = no --group file          src/BaseVarC.cpp:340-370 */
    const int32_t* sample_group;/* [n_samples] group id in [0,n_groups) or -1; NULL iff n_groups == 0 (copied) */
    int32_t max_sites;          /* capacity of one tile (sites)   */
    int32_t reserved0;
    int64_t max_entries;        /* capacity of one tile (entries) */
} bvc_params;

/* Number of CUDA devices visible to this process (0 when there is no driver / device). */
int  bvc_device_count(void);

/* Fills *p with the reference's constants for a cohort of n_samples and `--maf maf`. */
void bvc_default_params(bvc_params* p, int32_t n_samples, double maf);

/* A tile = n_sites consecutive sites of one window, CSR over covered entries, entries of a site in
 * ascending sample order (the order bt_s pushes them, src/BaseVarC.cpp:407-441). */
typedef struct bvc_tile {
    int32_t n_sites;
    int32_t reserved0;
    const uint8_t*  ref_base;   /* [n_sites]   BASE_INT8_TABLE[refseq[p-rg_s]] in 0..3  (src/BaseVarC.cpp:548) */
    const uint32_t* entry_off;  /* [n_sites+1] entry_off[0] == 0 */
    const uint8_t*  code;       /* [R] BVC_MAKE_CODE(base, strand, is_indel) */
    const uint8_t*  qual;       /* [R] 0..127 (the reference keeps quals in int8_t, src/BaseType.h:18) */
    const uint8_t*  mapq;       /* [R] */
    const uint8_t*  rpr;        /* [R] read position + 1, truncated to 8 bits as the bitfield does */
    const uint32_t* sample;     /* [R] global sample index; may be NULL when n_groups == 0 */
    /* Optional compact wire form of (code, qual): ONE byte per entry plus a dictionary of the tile's distinct quals.
     * When packed8 != NULL the library copies it instead of code / qual (half the host->device bytes of the two arrays;
     * the path is PCIe-bound end to end) and expands it on the device; code and qual may then be NULL.
     * mapq / rpr are unchanged.  The host packer (bvc_packer_tile) fills it whenever a tile has <= BVC_PACK8_MAX_QUALS
     * distinct quals. */
    const uint8_t*  packed8;    /* [R] BVC_PACK8(qi, strand, base) */
    const uint8_t*  qual_dict;  /* [n_qual_dict] qual value of dictionary index qi */
    int32_t n_qual_dict;        /* 1..BVC_PACK8_MAX_QUALS */
    int32_t reserved1;
} bvc_tile;

/* packed8 byte: base:2 | qi:5 | strand:1 (ABI 3; the byte is the bin index of K1's shared-memory histogram, and with the
 * base in the top bits the entries of a site's dominant base spread over all 32 banks).
 * qi < 30: a counted entry of base 0..3 with qual qual_dict[qi] (the dictionary need not be sorted).
 * qi == 31: an indel entry with stale base 0..3 / strand (its stale qual is not used by the calling core).
 * qi == 30: an indel entry whose stale base is 4 (N): it counts nowhere; only strand is kept. */
#define BVC_PACK8_MAX_QUALS 30
#define BVC_PACK8_INDEL     31u
#define BVC_PACK8_INDEL_N   30u
#define BVC_PACK8(qi, strand, base) ((uint8_t)((((base) & 3u) << 6) | (((qi) & 31u) << 1) | ((strand) & 1u)))
#define BVC_PACK8_QI(b)     (((b) >> 1) & 31u)
#define BVC_PACK8_STRAND(b) ((b) & 1u)
#define BVC_PACK8_BASE(b)   (((b) >> 6) & 3u)

/* boundary-flag bits: a hard FP64 threshold of the reference was met within a 1e-9 relative margin
 * at this site, so a differently-ordered summation may legitimately land on the other side. */
#define BVC_FLAG_EM_DELTA   1u   /* delta vs epsilon            src/Algorithm.cpp:124 */
#define BVC_FLAG_LRT_CHI    2u   /* chi vs LRT_THRESHOLD        src/BaseType.cpp:102  */
#define BVC_FLAG_MIN_AF     4u   /* depth/total vs min_af       src/BaseType.cpp:79   */
#define BVC_FLAG_ARGMIN_TIE 8u   /* two null models within 1e-9 src/BaseType.cpp:99   */
#define BVC_FLAG_QUAL60     16u  /* var_qual vs QUAL_THRESHOLD  src/BaseType.cpp:221  */

/* Fixed 176-byte record per site (little-endian, natural alignment). */
typedef struct bvc_site_result {
    double  cvg_fs;          /* bt_fisher_exact on cvg_strand            src/BaseVarC.cpp:603 */
    double  cvg_sor;         /*                                          src/BaseVarC.cpp:604-608 */
    double  af_lrt[4];       /* by base code; 0 unless base in alt_bases src/BaseType.cpp:111-116 */
    double  chi;             /* chi_sqrt_t after the ladder              src/BaseType.cpp:93-110 */
    double  var_qual;        /*                                          src/BaseType.cpp:117-135 */
    double  vcf_fs;          /* valid iff lrt_ok                         src/BaseType.cpp:192 */
    double  vcf_sor;         /*                                          src/BaseType.cpp:193-197 */
    double  ranksum[3];      /* mapq, qual, rpr                          src/BaseType.cpp:189-191 */
    int32_t n_entries;       /* aiv.size() */
    int32_t depth[4];        /* na,nc,ng,nt over non-indel entries       src/BaseVarC.cpp:550-559 */
    int32_t cvg_strand[4];   /* ref_fwd,ref_rev,alt_fwd,alt_rev, ALL entries   src/BaseVarC.cpp:587-602 */
    int32_t vcf_strand[4];   /* same, non-indel entries, alt = any LRT alt     src/BaseType.cpp:174-187 */
    int32_t em_fits;         /* diagnostics: EM() calls the reference would make */
    int32_t em_passes;       /* diagnostics: singleEM passes over all fits */
    uint32_t flags;          /* BVC_FLAG_* */
    int8_t  cvg_alt_base;    /* top-count non-ref base                   src/BaseVarC.cpp:576-583 */
    int8_t  lrt_ok;          /* BaseType::LRT() return value */
    int8_t  n_alt;           /* alt_bases.size() */
    int8_t  alt_bases[3];    /* in surviving-base order */
    int8_t  pad[2];
} bvc_site_result;

typedef struct bvc_ctx bvc_ctx;

int  bvc_create(bvc_ctx** out, int device, const bvc_params* params);
void bvc_destroy(bvc_ctx* ctx);
const char* bvc_last_error(const bvc_ctx* ctx);   /* ctx may be NULL: returns the last create() error */

/* Asynchronous host-buffer path (the reference-facing call): copies the tile H2D, runs the kernels,
 * copies n_sites records D2H into `out` (caller-owned; pinned memory makes the copies async).
 * Up to two tickets may be in flight; a third bvc_enqueue waits for the oldest. */
int  bvc_enqueue(bvc_ctx* ctx, const bvc_tile* tile, bvc_site_result* out, uint64_t* ticket);
int  bvc_wait(bvc_ctx* ctx, uint64_t ticket);
/* Split result for the PCIe-bound host-buffer path: every site gets the 72-byte core of its record (all a CVG row
 * prints, src/BaseVarC.cpp:568-575,610); the full 176-byte record travels only for the LRT-positive sites (the VCF rows,
 * ~0.3 % of a low-pass window), compacted into vcf[0 .. n_vcf) in no particular order -- core[site].vcf_index says where.
 * For the other sites the VCF half of the record is all zero by definition.  40 % of the bytes of bvc_enqueue's output. */
typedef struct bvc_site_core {
    double  cvg_fs, cvg_sor;
    int32_t n_entries;
    int32_t depth[4];
    int32_t cvg_strand[4];
    int32_t em_fits, em_passes;
    uint32_t flags;
    int8_t  cvg_alt_base, lrt_ok, n_alt, reserved;
    int32_t vcf_index;          /* index into vcf[] of this call when lrt_ok, else -1 */
} bvc_site_core;
/* core: [n_sites]; vcf: [vcf_cap] (both caller-owned until the ticket is retired).  Shares the two-slot ticket queue with
 * bvc_enqueue: a third enqueue retires the oldest ticket (its records are in place by then); bvc_wait_split may be called
 * for a ticket up to the moment its slot has retired ANOTHER split ticket. */
int  bvc_enqueue_split(bvc_ctx* ctx, const bvc_tile* tile, bvc_site_core* core, bvc_site_result* vcf, int32_t vcf_cap, uint64_t* ticket);
/* *n_vcf = LRT-positive sites of the tile.  BVC_ERR_CAPACITY (core valid, vcf untouched) if that exceeds vcf_cap. */
int  bvc_wait_split(bvc_ctx* ctx, uint64_t ticket, int32_t* n_vcf);
/* Optional per-group outputs of the most recently enqueued tile (n_groups > 0), valid after bvc_wait:
 * group_depth [n_sites][n_groups][4] int32 (src/BaseVarC.cpp:622-640) and
 * group_af    [n_sites][n_groups][4] double by base code, NaN-free (src/BaseVarC.cpp:641-660). */
int  bvc_fetch_groups(bvc_ctx* ctx, uint64_t ticket, int32_t* group_depth, double* group_af);

/* Pinned host allocation helpers (cudaHostAlloc / cudaFreeHost). */
void* bvc_host_alloc(size_t bytes);
void  bvc_host_free(void* p);
/* Page-lock caller-owned memory in place (cudaHostRegister / cudaHostUnregister): what the dispatcher does with a packer's
 * arrays once, so bvc_enqueue's copies are asynchronous and mapq / rpr are read in place.  0, or BVC_ERR_CUDA. */
int   bvc_host_register(void* p, size_t bytes);
void  bvc_host_unregister(void* p);

/* Device-resident path, used by bench.py for the kernel-only figure: upload once, run many times.
 * bvc_resident_run launches the site kernels `iters` times on the context's stream, optionally
 * flushing L2 (a >126 MB memset) before each timed launch, and returns the mean device time per
 * launch in milliseconds measured with CUDA events on that stream.  per_kernel_ms (may be NULL)
 * receives [n_kernels] mean times; n_kernels is returned through *n_kernels_out. */
/* per_kernel_ms is indexed by kernel KIND (all launches of a kind summed per run); bvc_kernel_name(kind) names it. */
int  bvc_resident_upload(bvc_ctx* ctx, const bvc_tile* tile);
int  bvc_resident_run(bvc_ctx* ctx, int iters, int flush_l2, float* ms_per_launch,
                      float* per_kernel_ms, int* n_kernels_out);
int  bvc_resident_download(bvc_ctx* ctx, bvc_site_result* out);
const char* bvc_kernel_name(int kind);                 /* kind in [0, n_kernels) */
/* Work done by the last bvc_resident_run pass: out4 = { class-passes in em_fit_kernel, class-passes in
 * em_long_kernel, class evaluations in fit_loglik_kernel, EM fits }.  One class-pass = one (base, qual)
 * class of one singleEM pass (src/Algorithm.cpp:69-93 + its delta term :103-113). */
int  bvc_work_counters(bvc_ctx* ctx, uint64_t* out4);
int64_t bvc_kernel_launch_count(const bvc_ctx* ctx);   /* kernels of ours launched so far */
/* Cumulative bytes moved by bvc_enqueue's copies (H2D of tiles, D2H of records).  mapq / rpr arrays that live
 * in pinned host memory are NOT copied: the kernels read them in place for LRT-positive sites only. */
void bvc_transfer_bytes(const bvc_ctx* ctx, int64_t* h2d, int64_t* d2h);

/* ---- Sample-sharded mode (cohorts of >= 100k samples) ------------------------------------------
 * The reference merges the sample batches of a site on one host thread: the loop over the batch files
 * (src/BaseVarC.cpp:405-441) appends every batch's tokens to the site's one AlleleInfoVector before bt_f runs.
 * Here rank r of W (one rank = one context = one GPU) is given, for EVERY site of a tile, only the entries of its
 * own sample range (its range of batch files).  Per tile the ranks exchange integer sufficient statistics once:
 * reduce-scatter of the per-site (base, qual, strand) histograms, all-gather of one byte per site, all-reduce of
 * the rank-sum histograms of the LRT-positive sites (NCCL over NVLink; the library dlopens libnccl.so.2).
 * (The histograms cover the quals the tile's entries actually carry: the ranks first agree on that range with a 2-word MAX
 * all-reduce, inside the window given to bvc_shard_init; BVC_SHARD_FIXED_WINDOW=1 keeps the whole window.)
 * Rank r returns the records of the site slice [r*ceil(S/W), ...) of the tile.  bvc_create must have been called
 * with n_samples = the size of the WHOLE cohort (min_af and the lgamma table depend on it) and n_groups = 0.
 * Every collective call below must be made by all W ranks with tiles of the same n_sites / ref_base.
 * Not served in this mode: qual 0 (qual_lo >= 1; an entry whose qual is outside [qual_lo, qual_hi] makes the call
 * fail with BVC_ERR_ARG on every rank that holds one, after all collectives have completed) and --group. */
#define BVC_SHARD_ID_BYTES 128      /* sizeof(ncclUniqueId) */
#define BVC_SHARD_N_PHASES 9
int  bvc_shard_unique_id(void* id128);                     /* rank 0: ncclGetUniqueId; the caller broadcasts the bytes */
int  bvc_shard_init(bvc_ctx* ctx, int rank, int world, const void* id128, int qual_lo, int qual_hi);   /* collective */
/* host buffers in, the owned slice of records out (synchronous; collective) */
int  bvc_shard_call(bvc_ctx* ctx, const bvc_tile* local_tile, bvc_site_result* out, int32_t* first_site, int32_t* n_out);
/* device-resident variant for timing: upload once, run many times (collective) */
int  bvc_shard_upload(bvc_ctx* ctx, const bvc_tile* local_tile);
int  bvc_shard_run(bvc_ctx* ctx, bvc_site_result* out, int32_t* first_site, int32_t* n_out);
/* In-process exchange WITHOUT NCCL: W contexts of one process on one device, ctxs[r] = rank r.  Same kernels and
 * phases; the sums are plain kernels.  This is how the sharded path is parity-tested on a single GPU. */
int  bvc_shard_init_group(bvc_ctx** ctxs, int world, int qual_lo, int qual_hi);
int  bvc_shard_call_group(bvc_ctx** ctxs, int world, const bvc_tile* const* local_tiles, bvc_site_result* const* outs,
                          int32_t* first_site /* [world] */, int32_t* n_out /* [world] */);
/* Device time of each phase of the last call in ms [BVC_SHARD_N_PHASES] (bvc_shard_phase_name(i) names them), the
 * cumulative payload bytes this rank handed to the collectives, and the number of LRT-positive sites of the tile. */
int  bvc_shard_stats(const bvc_ctx* ctx, float* phase_ms, int64_t* exchange_bytes, int32_t* n_positive);
const char* bvc_shard_phase_name(int i);

/* FP64 pipe microbenchmark (dependent-free DFMA chains on all SMs): returns TFLOP/s (2 flop/FMA). */
int  bvc_measure_fp64_peak(int device, double* tflops);

/* ---- Coded wire form: fewer bytes over a saturated host link ---------------------------------------------------------
 * No counterpart in the reference.  bvc_enqueue's host->device copy of the entries is what bounds the host-buffer path (401 MB
 * per 1 M-site C2 tile at 52 of the link's 55 GB/s; with 8 GPUs the box's host link saturates at 226 GB/s, DESIGN.md section
 * 6), and a packed8 byte carries < 8 bits: a low-pass tile has ~42 frequent byte values (21 quals x 2 strands at the reference
 * base).  The coded form sends W = 4, 5 or 6 bits per entry: a byte is made relative to its site's reference base (base field ->
 * (base - ref) & 3), the tile's 2^W - 1 most frequent relative values become codes, every other entry is the escape code
 * 2^W - 1 plus its packed8 byte in a side array -- and a kernel turns it back into the resident
 * packed8 bytes (0.75 B + escapes in, 1 B out per entry), after which the path is the packed8 one, byte for byte.
 *   words          bit stream: entries 32u .. 32u+31 occupy W consecutive words, entry j at bits [jW, jW+W) of that little-endian
 *                  32W-bit string; ceil(n/32) * W words
 *   escapes        the packed8 bytes of the escaped entries, in entry order
 *   chunk_escapes  [ceil(n / BVC_CODED_CHUNK) + 1] number of escapes before each chunk of BVC_CODED_CHUNK entries
 * bvc_encode_codes builds it from a tile's packed8 / entry_off / ref_base into caller buffers (words: bvc_coded_words(n, 6) words, escapes: n/8 + 16
 * bytes, chunk_escapes: n / BVC_CODED_CHUNK + 2 words) and leaves out->bits = 0 when it would not pay (escapes > n/8 or
 * < 10 % saved: send packed8).  bvc_enqueue_split_coded is bvc_enqueue_split with (tile->packed8, code, qual) == NULL and the
 * coded arrays instead; tile->qual_dict / n_qual_dict as for packed8.  bvc_decode_codes_host decodes on the host (tests). */
#define BVC_CODED_CHUNK 8192
typedef struct bvc_coded_tile {
    int32_t bits;                 /* W: 4, 5 or 6; 0 = not coded */
    int32_t n_codes;              /* direct codes in use (<= 2^W - 1) */
    uint8_t code_byte[64];        /* code -> packed8 byte */
    const uint32_t* words;
    const uint8_t*  escapes;
    const uint32_t* chunk_escapes;
    int64_t n_escape;
} bvc_coded_tile;
size_t bvc_coded_words(int64_t n_entries, int bits);
int  bvc_encode_codes(const bvc_tile* tile /* packed8, entry_off, ref_base */, uint32_t* words, uint8_t* escapes, uint32_t* chunk_escapes, bvc_coded_tile* out);
int  bvc_decode_codes_host(const bvc_coded_tile* coded, const bvc_tile* tile /* entry_off, ref_base */, uint8_t* packed8_out);
int  bvc_enqueue_split_coded(bvc_ctx* ctx, const bvc_tile* tile, const bvc_coded_tile* coded, bvc_site_core* core, bvc_site_result* vcf,
                             int32_t vcf_cap, uint64_t* ticket);

/* ---- Allocation-free call -------------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same records as
 * bvc_enqueue + bvc_wait for one small tile, computed by the one-warp-per-site kernel (site_call_kernel,
 * csrc/bvc_kernels.cu), with NO context and no allocation of any kind: the caller passes the tile and the record
 * array as DEVICE pointers, and everything the kernel needs besides -- its dynamic site counter, the checked site
 * list, the error word, the per-qual likelihood table and the lgamma table -- lives in ONE file-scope
 * `static __device__` array of the library (~390 KB, held by every GPU that loads the library until the process
 * exits).  Limits, fixed at build time: n_sites <= BVC_NOALLOC_MAX_SITES, prm->n_samples <= BVC_NOALLOC_MAX_SAMPLES,
 * n_entries < 2^32; anything larger fails with BVC_ERR_CAPACITY and a message in bvc_last_error(NULL).
 *   d_tile     a bvc_tile whose array pointers are device pointers (ref_base, entry_off, code, qual, mapq, rpr);
 *              packed8 must be NULL (the plain two-array form only), sample is ignored
 *   n_entries  length of the entry arrays; entry_off[n_sites] must not exceed it (entry_off lives on the device, so the
 *              host side cannot read it)
 *   prm        as for bvc_create: n_samples, min_af, lrt_threshold, em_epsilon, em_max_iter are used; n_groups must be 0
 *   d_out      [n_sites] records, device memory
 *   d_status   one device word, may be NULL: 0, or a mask of BVC_NOALLOC_BAD_* found by a checking kernel that runs first
 *              (the checks bvc_wait reports as BVC_ERR_ARG).  When it is non-zero NO record is written.
 *   stream     a cudaStream_t of `device` (NULL = the default stream).  The call only enqueues (one 256-byte memset and
 *              three kernels) and returns; it never synchronises.
 * The first call on a device also does the set-up for that device (the address of the array, two synchronous table
 * copies): make that call outside stream capture -- a call with n_sites == 0 does the set-up and nothing else.
 * All calls on a device SHARE the array, which keeps the previous call's contents: the call clears the words that must
 * start at zero itself, but two calls that overlap on the GPU (different streams, or graphs replayed concurrently)
 * race on it.  Order them on one stream or with events. */
#define BVC_NOALLOC_MAX_SITES    65536
#define BVC_NOALLOC_MAX_SAMPLES  16384
#define BVC_NOALLOC_BAD_QUAL     1u    /* an entry with qual > 127 */
#define BVC_NOALLOC_BAD_BASE     2u    /* a non-indel entry with base > 3 */
#define BVC_NOALLOC_BAD_OFFSETS  4u    /* entry_off[0] != 0, not ascending, beyond n_entries, or a site with > n_samples entries */
int  bvc_call_sites_noalloc(int device, const bvc_params* prm, const bvc_tile* d_tile, int64_t n_entries,
                            bvc_site_result* d_out, uint32_t* d_status, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BVC_GPU_H */
