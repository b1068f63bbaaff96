// bvc_kernels.h -- launcher declarations shared by bvc_kernels.cu and bvc_api.cu
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bvc {
struct SiteArgs;
cudaError_t launch_site_call(const SiteArgs& a, int n_sm, cudaStream_t stream);
cudaError_t launch_fp64_peak(double* out, int blocks, int threads, int iters, cudaStream_t stream);
}  // namespace bvc
