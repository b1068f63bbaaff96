// Microbenchmark: throughput of red.shared.add.u32 (no return) on sm_100a, in cycles per warp instruction per SM,
// for spread addresses (conflict-free), k-way same-address multiplicity, and for the LDS+STS private-counter
// alternative.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_red tools/ubench/red_shared.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(uint32_t* out, int iters, int mult, long long* cyc)
{
    __shared__ uint32_t h[8192];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) h[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // lanes of a warp: groups of `mult` lanes share one address; different warps use different rows
    uint32_t idx = wid * 64 + (lane / mult);
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(h);
    long long t0 = clock64();
    uint32_t x = lane;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const uint32_t a = sa + 4u * ((idx + (uint32_t)u * 32u + ((x >> 3) & 0u)) & 8191u);
            if (MODE == 0) asm volatile("red.shared.add.u32 [%0], %1;" :: "r"(a), "r"(1u) : "memory");
            if (MODE == 1) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(v + 1) : "memory"); }
            if (MODE == 2) { uint32_t v; asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(v) : "r"(a), "r"(1u) : "memory"); x += v; }
        }
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = h[threadIdx.x] + x;
}

int main()
{
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    const int iters = 2000;
    for (int mode = 0; mode < 3; ++mode)
        for (int warps : {4, 8, 16, 32})
            for (int mult : {1, 2, 4, 8, 32}) {
                if (mode == 1 && mult > 1) continue;
                if (mode == 0) k<0><<<148, warps * 32>>>(out, iters, mult, cyc);
                if (mode == 1) k<1><<<148, warps * 32>>>(out, iters, mult, cyc);
                if (mode == 2) k<2><<<148, warps * 32>>>(out, iters, mult, cyc);
                cudaDeviceSynchronize();
                long long c[148]; cudaMemcpy(c, cyc, sizeof c, cudaMemcpyDeviceToHost);
                double avg = 0; for (int i = 0; i < 148; ++i) avg += c[i]; avg /= 148;
                const double ops = (double)iters * 8 * warps;       // warp instructions per SM
                printf("%s warps/SM=%2d same-addr multiplicity=%2d : %.2f cycles per warp-op per SM (%.3f cyc/lane)\n",
                       mode == 0 ? "RED.shared" : mode == 1 ? "LDS+STS  " : "ATOMS.ret ", warps, mult, avg / ops, avg / ops / 32);
            }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
