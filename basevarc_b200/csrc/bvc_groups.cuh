// bvc_groups.cuh -- the --group path of bt_f (src/BaseVarC.cpp:617-661 under /root/reference).
//
//   group_depth_kernel  warp per site: A/C/G/T counts of the non-indel entries of every population group
//                       (the "na:nc:ng:nt" CVG columns, :622-640).
//   group_lrt_kernel    warp per (LRT-positive site, group): the second-level BaseType over the group's reads,
//                       candidates restricted to {ref} + the site's ALT alleles (SetBase, :641-645), LRT() and the
//                       group's af_lrt per ALT allele (:646-655).  Uses the warp-per-site EM/LRT code of
//                       bvc_site_kernel.cuh (true divisions, IEEE semantics for qual-0 reads).
// group_af[site][g][b], b an ALT allele of the site = af_lrt[b] of the group's model, or -1 when b is not in it
// (the reference prints "0"); 0 for every other base, for empty groups and for sites without a VCF record.
#pragma once
#include <stdint.h>
#include "bvc_site_kernel.cuh"

namespace bvc {

struct GroupArgs {
    SiteArgs s;
    const uint32_t* sample;        // [R] global sample index per entry
    const int32_t*  sample_group;  // [n_samples] group id or -1
    int32_t* gdepth;               // [n_sites][n_groups][4], zeroed before the launch
    double*  gaf;                  // [n_sites][n_groups][4], zeroed before the launch
    int n_groups;
};

constexpr int kGroupWarps = 4;

__global__ void __launch_bounds__(kGroupWarps * 32)
group_depth_kernel(GroupArgs a)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int n_warps = gridDim.x * kGroupWarps;
    for (int site = blockIdx.x * kGroupWarps + wid; site < a.s.n_sites; site += n_warps) {
        const uint32_t b0 = a.s.entry_off[site], b1 = a.s.entry_off[site + 1];
        int32_t* out = a.gdepth + (size_t)site * a.n_groups * 4;
        for (uint32_t e = b0 + lane; e < b1; e += 32) {
            const unsigned c = a.s.code[e];
            if (c & 0x14u) continue;                               // indel entry or not A/C/G/T
            const int g = a.sample_group[a.sample[e]];
            if (g >= 0) atomicAdd(&out[g * 4 + (c & 3)], 1);
        }
    }
}

__global__ void __launch_bounds__(kGroupWarps * 32)
group_lrt_kernel(GroupArgs a)
{
    __shared__ __align__(16) unsigned char smem_w[kGroupWarps][kWarpSmemBytes];
    __shared__ double lut_s[4 * kNQ];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 4 * kNQ; i += blockDim.x) lut_s[i] = a.s.lut[i];
    WarpSmem w;
    w.lml = reinterpret_cast<double*>(smem_w[wid]);
    w.hist = reinterpret_cast<uint32_t*>(smem_w[wid] + kNBins * 8);
    w.keys = reinterpret_cast<uint16_t*>(smem_w[wid] + kNBins * 12);
    w.codehist = reinterpret_cast<uint32_t*>(smem_w[wid] + kNBins * 14);
    for (int i = lane; i < kNBins; i += 32) w.hist[i] = 0;
    __syncthreads();

    const long long n_items = (long long)a.s.n_sites * a.n_groups;
    const long long n_warps = (long long)gridDim.x * kGroupWarps;
    for (long long it = (long long)blockIdx.x * kGroupWarps + wid; it < n_items; it += n_warps) {
        const int site = (int)(it / a.n_groups), g = (int)(it % a.n_groups);
        const bvc_site_result* r = a.s.out + site;
        double* gaf = a.gaf + ((size_t)site * a.n_groups + g) * 4;
        if (!r->lrt_ok) continue;                                  // warp-uniform: the group model exists only for VCF sites
        const int32_t* gd = a.gdepth + ((size_t)site * a.n_groups + g) * 4;
        SiteState st;
        st.em_fits = 0; st.em_passes = 0; st.flags = 0; st.n_reads = 0;
        for (int b = 0; b < 4; ++b) { st.depth[b] = gd[b]; st.n_reads += gd[b]; }
        if (st.n_reads == 0) continue;                             // gr_bases.empty(): the reference prints "0" (gaf stays 0)
        const int ref = a.s.ref_base[site] & 3;
        // the group's reads -> (base, qual) histogram -> classes, as phases A/B of site_call_kernel
        const uint32_t b0 = a.s.entry_off[site], b1 = a.s.entry_off[site + 1];
        unsigned qmin = 127, qmax = 0;
        for (uint32_t e = b0 + lane; e < b1; e += 32) {
            const unsigned c = a.s.code[e];
            if (c & 0x14u) continue;
            if (a.sample_group[a.sample[e]] != g) continue;
            const unsigned q = a.s.qual[e] & (kNQ - 1);
            atomicAdd(&w.hist[((c & 3) << 7) | q], 1u);
            qmin = min(qmin, q); qmax = max(qmax, q);
        }
        __syncwarp();
        qmin = __reduce_min_sync(0xffffffffu, qmin);
        qmax = __reduce_max_sync(0xffffffffu, qmax);
        st.has_q0 = (qmin == 0);
        int ncls = 0;
        for (int b = 0; b < 4; ++b) {
            if (st.depth[b] == 0) continue;
            for (unsigned q0 = qmin & ~31u; q0 <= qmax; q0 += 32) {
                const unsigned bin = (b << 7) + q0 + lane;
                const uint32_t v = w.hist[bin];
                const unsigned m = __ballot_sync(0xffffffffu, v != 0);
                const int pos = ncls + __popc(m & ((1u << lane) - 1));
                __syncwarp();
                if (v) w.hist[bin] = 0;
                __syncwarp();
                if (v) { w.hist[pos] = v; w.keys[pos] = (uint16_t)bin; }
                ncls += __popc(m);
                __syncwarp();
            }
        }
        st.ncls = ncls;
        // base_comb = {ref} + alt_bases of the site-level model, in that order (src/BaseVarC.cpp:614-615,643)
        int comb[4]; int n_comb = 1;
        comb[0] = ref;
        for (int i = 0; i < r->n_alt && i < 3; ++i) comb[n_comb++] = r->alt_bases[i];
        LrtOut L;
        lrt_ladder(a.s, w, lut_s, st, ref, comb, n_comb, lane, L);
        __syncwarp();
        if (lane == 0) {
            for (int i = 0; i < r->n_alt && i < 3; ++i) gaf[r->alt_bases[i]] = -1.0;
            for (int i = 0; i < L.n_alt; ++i) gaf[L.alt_bases[i]] = L.af[L.alt_bases[i]];
        }
        for (int i = lane; i < ncls; i += 32) w.hist[i] = 0;
        __syncwarp();
    }
}

}  // namespace bvc
