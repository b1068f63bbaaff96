// bvc_kernels.h -- kernel argument block + launcher declarations shared by bvc_kernels.cu and bvc_api.cu
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/bvc_gpu.h"

namespace bvc {

constexpr int kWarpsPerBlock = 4;
constexpr int kNQ = 128;                 // qual bins per base (int8 quals, src/BaseType.h:18)
constexpr int kNBins = 4 * kNQ;          // (base, qual) classes
// bits of the device error word (SiteArgs::err): what bvc_wait turns into BVC_ERR_ARG
constexpr uint32_t kErrQual = 1u;        // a qual > 127 (plain form) / a dictionary index >= n_qual_dict (packed8 form)
constexpr uint32_t kErrBase = 2u;        // a non-indel entry with base > 3
// per-warp shared memory: lml f64[512] | hist u32[512] | keys u16[512] | codehist u32[32]
constexpr int kWarpSmemBytes = kNBins * 8 + kNBins * 4 + kNBins * 2 + 32 * 4;

// packed8 -> (code, qual): include/bvc_gpu.h BVC_PACK8
struct QualDict { uint8_t q[32]; };

struct SiteArgs {
    const uint8_t*  ref_base;
    const uint32_t* entry_off;
    const uint8_t*  code;
    const uint8_t*  qual;
    const uint8_t*  mapq;
    const uint8_t*  rpr;
    const uint8_t*  pk8;         // != NULL: the tile is resident in its ONE-byte form (BVC_PACK8) and code / qual are not filled:
    QualDict        dict;        //   K1 (site_hist8_kernel) and K5 read pk8 + dict directly
    int             n_dict;
    uint32_t*       err;         // device error word (kErr* bits): input the ABI promises to reject (include/bvc_gpu.h BVC_ERR_ARG)
    bvc_site_result* out;
    unsigned int*   counter;     // dynamic site fetch
    const uint32_t* site_list;   // optional: process only these sites (legacy path of the pipeline) ...
    const uint32_t* list_n;      // ... their count (device memory)
    const double*   lut;         // [4][128]: e/3, 1-e, log(e/3), log(1-e) with e = exp(MLN10TO10*q)
    const double*   lgam;        // [n_samples+3] lgamma(k)
    int    n_sites;
    int    max_iter;
    double min_af, lrt_thr, eps;
};

struct PipeArgs;
cudaError_t launch_site_call(const SiteArgs& a, int n_sm, cudaStream_t stream);
// The staged pipeline (bvc_pipeline.cuh).  q_a/q_b + their counters ping-pong between stages.
// ev/kinds (optional, [kMaxPipeLaunches + 1] / [kMaxPipeLaunches]): an event before every launch and one after the
// last, plus the kernel kind of each launch, for the per-kernel breakdown of bvc_resident_run.
enum KernelKind { kKHist = 0, kKPrepare, kKSort, kKEm, kKEmLong, kKLoglik, kKLadder, kKVcf, kKLegacy, kNumKernelKinds };
constexpr int kMaxPipeLaunches = 32;
const char* kernel_kind_name(int kind);
// parts: kPartHist = K1, kPartCore = K2 + the EM stages + K4, kPartVcf = K5 + the qual-0 legacy kernel.
enum PipeParts { kPartHist = 1, kPartCore = 2, kPartVcf = 4, kPartAll = 7 };
cudaError_t launch_pipeline(const PipeArgs& a, int n_sm, cudaStream_t stream, int* n_launches,
                            cudaEvent_t* ev = nullptr, int* kinds = nullptr, int parts = kPartAll);
// sample-sharded mode (bvc_shard.cuh): one launcher per phase; each returns the number of kernels it launched
struct ShardArgs;
cudaError_t launch_shard_qual_range(const uint8_t* code, const uint8_t* qual, size_t n_entries, uint32_t* out, int n_sm, cudaStream_t stream);
cudaError_t launch_shard_hist(const ShardArgs& a, int n_sm, cudaStream_t stream, int* n_launches);
cudaError_t launch_shard_hist8(const ShardArgs& a, const uint8_t* pk8, int n_sm, cudaStream_t stream, int* n_launches);
cudaError_t launch_shard_merge_and_fit(const ShardArgs& a, int n_sm, cudaStream_t stream, int* n_launches);
cudaError_t launch_shard_poslist(const ShardArgs& a, cudaStream_t stream, int* n_launches);
cudaError_t launch_shard_vcf_hist(const ShardArgs& a, int n_sm, uint32_t n_pos, cudaStream_t stream, int* n_launches);
cudaError_t launch_shard_vcf_finish(const ShardArgs& a, int n_sm, uint32_t n_pos, cudaStream_t stream, int* n_launches);
cudaError_t launch_sum_u32(uint32_t* dst, const uint32_t* src, size_t n, int n_sm, cudaStream_t stream);
// --group path (bvc_groups.cuh): per-group depths of every site + the group-level LRT of the LRT-positive sites
cudaError_t launch_groups(const SiteArgs& s, const uint32_t* sample, const int32_t* sample_group, int n_groups,
                          int32_t* gdepth, double* gaf, int n_sm, cudaStream_t stream);
cudaError_t launch_fp64_peak(double* out, int blocks, int threads, int iters, cudaStream_t stream);
// records -> 72-byte cores + compacted full records of the LRT-positive sites (bvc_enqueue_split)
cudaError_t launch_split_result(const bvc_site_result* rec, int n_sites, bvc_site_core* core, bvc_site_result* vcf, uint32_t* vcf_n,
                                cudaStream_t stream);
// coded wire form (bvc_codes.h) -> packed8 bytes
cudaError_t launch_decode_codes(int bits, const uint32_t* words, const uint8_t* esc, const uint32_t* chunk_esc, const uint8_t* code_byte,
                                const uint32_t* entry_off, const uint8_t* ref_base, int n_sites, uint8_t* pk8, int64_t n, int n_sm, cudaStream_t stream);
// packed8 -> (code, qual) arrays; n is rounded up to whole 16-byte words (the buffers are padded)
cudaError_t launch_expand8(const uint8_t* pk8, uint8_t* code, uint8_t* qual, int64_t n, const QualDict& dict, int n_sm, cudaStream_t stream);
}  // namespace bvc
