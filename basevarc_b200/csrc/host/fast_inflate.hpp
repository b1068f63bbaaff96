// fast_inflate.hpp -- a raw DEFLATE (RFC 1951) decoder for ONE BGZF block held in memory.
//
// The reference reads its batch files through htslib's bgzf_getline (src/BaseVarC.cpp:407), i.e. zlib's streaming inflate;
// with the calling core on the GPU that inflate is what bounds a `--rerun` from the text batch files (22 of 35 thread-seconds
// per 1 M sites x 1,000 samples).  A BGZF block is a complete deflate stream of <= 64 KiB whose compressed and inflated sizes
// are known up front, so the decoder can be a whole-buffer one: a 64-bit bit buffer refilled by one unaligned 8-byte load per
// symbol, 10-bit / 8-bit root tables with sub-tables, word-wise match copies into an output buffer with slack, no per-byte
// bounds checks in the main loop.  Written from the RFC; it is NOT the verdict on a block: BgzfReader keeps the CRC-32 check
// of every block, and a block this decoder declines (returns false: malformed, incomplete code, anything unexpected) is
// inflated again by zlib, whose answer stands.
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace bvc_host {

constexpr size_t kFastInflateInPad = 16;     // readable bytes required BEHIND in[in_len) (content irrelevant)
constexpr size_t kFastInflateOutSlack = 320; // writable bytes required BEHIND out[out_cap) (clobbered)

// true: the stream ended with its final block after exactly *out_len <= out_cap bytes.  false: see above.
bool fast_inflate(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_cap, size_t* out_len);

// CRC-32 (the gzip polynomial) of buf[0, len): carry-less-multiply folding where the CPU has PCLMULQDQ (4 x 128-bit lanes, the
// published fold constants for this polynomial), else zlib's crc32.  Same value as zlib's crc32(0, buf, len).
uint32_t fast_crc32(const uint8_t* buf, size_t len);

}  // namespace bvc_host
