// bvc_codes.h -- the CODED wire form of a tile's packed8 bytes (include/bvc_gpu.h "coded wire form"): W = 4, 5 or 6 bits per
// entry.  A byte is first made RELATIVE to its site's reference base (base field -> (base - ref) & 3: at a low-pass site ~99 % of
// the entries then share the ~2 x n_quals values with relative base 0, whatever the site's reference base is).  The tile's
// 2^W - 1 most frequent relative byte values get the codes 0 .. 2^W - 2; every other byte is the ESCAPE code
// 2^W - 1 and travels in a side array, in entry order.  Entries 32u .. 32u + 31 ("unit" u) occupy W consecutive 32-bit words:
// entry j of the unit sits at bits [jW, jW + W) of that little-endian 32W-bit string.  The unit decoder below is shared by
// the device kernel (bvc_kernels.cu) and by the host-side checker (bvc_decode_codes_host).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define BVC_HD __host__ __device__ __forceinline__
#else
#define BVC_HD inline
#endif

namespace bvc {

constexpr int kCodedChunk = 8192;        // entries per chunk (= per thread block of the decoder: 256 threads x 32 entries)

// 32 codes of one unit -> 32 bytes through `lut` (code -> packed8 byte; lut[escape] is ignored) + the mask of escaped entries
template <int W>
BVC_HD void decode_unit(const uint32_t* w, const uint8_t* lut, uint8_t* out, uint32_t* esc_mask)
{
    constexpr uint32_t mask = (1u << W) - 1u;
    uint32_t em = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int j = 0; j < 32; ++j) {
        const int bit = j * W, wi = bit >> 5, sh = bit & 31;
        uint32_t v = w[wi] >> sh;
        if (sh + W > 32) v |= w[wi + 1] << (32 - sh);
        v &= mask;
        if (v == mask) em |= 1u << j;
        out[j] = lut[v];
    }
    *esc_mask = em;
}

// relative byte <-> packed8 byte of an entry at a site whose reference base is `ref` (0..3)
BVC_HD uint8_t rel_to_abs(uint8_t r, unsigned ref) { return (uint8_t)((r & 0x3fu) | ((((unsigned)(r >> 6) + ref) & 3u) << 6)); }
BVC_HD uint8_t abs_to_rel(uint8_t b, unsigned ref) { return (uint8_t)((b & 0x3fu) | ((((unsigned)(b >> 6) + 4u - ref) & 3u) << 6)); }

}  // namespace bvc
