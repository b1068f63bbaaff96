// fast_inflate.cc -- see fast_inflate.hpp.  RFC 1951 section numbers in the comments.
#include "fast_inflate.hpp"

#include <string.h>
#include <zlib.h>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace bvc_host {
namespace {

constexpr int kLitBits = 10, kDistBits = 8, kPreBits = 7, kMaxLen = 15;
constexpr int kLitCap = (1 << kLitBits) + 288 * 32, kDistCap = (1 << kDistBits) + 32 * 128;

// table entry: nbits:8 (bits this lookup consumes: code length, plus the extra bits of a length / distance) | extra:5 | kind:3 | value:16
enum Kind : uint32_t { kInvalid = 0, kLiteral = 1, kBase = 2 /* a length or a distance: base value + extra bits */, kEob = 3, kSub = 4 };
inline uint32_t entry(uint32_t kind, uint32_t value, uint32_t extra, uint32_t nbits) { return nbits | (extra << 8) | (kind << 13) | (value << 16); }
inline uint32_t e_nbits(uint32_t e) { return e & 0xffu; }
inline uint32_t e_extra(uint32_t e) { return (e >> 8) & 0x1fu; }
inline uint32_t e_kind(uint32_t e) { return (e >> 13) & 7u; }
inline uint32_t e_value(uint32_t e) { return e >> 16; }

// 3.2.5: length codes 257..285 and distance codes 0..29
const uint16_t kLenBase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
const uint8_t kLenExtra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
const uint16_t kDistBase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
const uint8_t kDistExtra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
const uint8_t kPreOrder[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};   // 3.2.7

enum Alphabet { kAlphaLit, kAlphaDist, kAlphaPre };

inline uint32_t symbol_entry(Alphabet a, int s, uint32_t nbits)
{
    if (a == kAlphaPre) return entry(kLiteral, (uint32_t)s, 0, nbits);
    if (a == kAlphaDist) return s < 30 ? entry(kBase, kDistBase[s], kDistExtra[s], nbits + kDistExtra[s]) : entry(kInvalid, 0, 0, nbits);
    if (s < 256) return entry(kLiteral, (uint32_t)s, 0, nbits);
    if (s == 256) return entry(kEob, 0, 0, nbits);
    return s < 286 ? entry(kBase, kLenBase[s - 257], kLenExtra[s - 257], nbits + kLenExtra[s - 257]) : entry(kInvalid, 0, 0, nbits);
}

inline uint32_t reverse_bits(uint32_t code, int len)
{
    uint32_t r = 0;
    for (int i = 0; i < len; ++i) { r = (r << 1) | (code & 1u); code >>= 1; }
    return r;
}

// Canonical Huffman code of 3.2.2 as a root table of 2^root entries indexed by the next `root` stream bits (codes are packed
// most-significant bit first into an LSB-first bit stream, hence the reversal) plus one sub-table per root slot shared by longer
// codes.  Slots no code reaches stay kInvalid (an incomplete code is legal for a lone distance code; hitting such a slot while
// decoding declines the block).  false: over-subscribed lengths or the table does not fit.
bool build_table(const uint8_t* lens, int n, Alphabet alpha, int root, uint32_t* table, int cap)
{
    int count[kMaxLen + 1] = {0};
    for (int s = 0; s < n; ++s) count[lens[s]]++;
    count[0] = 0;
    int left = 1;
    for (int l = 1; l <= kMaxLen; ++l) { left = (left << 1) - count[l]; if (left < 0) return false; }
    uint32_t next[kMaxLen + 2];
    next[1] = 0;
    for (int l = 1; l <= kMaxLen; ++l) next[l + 1] = (next[l] + (uint32_t)count[l]) << 1;
    const int root_size = 1 << root;
    for (int i = 0; i < root_size; ++i) table[i] = 0;
    uint8_t sub_max[1 << kLitBits];
    bool any_long = false;
    for (int l = root + 1; l <= kMaxLen; ++l) if (count[l]) any_long = true;
    uint32_t code_of[320];
    for (int s = 0; s < n; ++s) {
        const int l = lens[s];
        if (!l) continue;
        const uint32_t rev = reverse_bits(next[l]++, l);
        code_of[s] = rev;
        if (l <= root) for (uint32_t i = rev; i < (uint32_t)root_size; i += 1u << l) table[i] = symbol_entry(alpha, s, (uint32_t)l);
    }
    if (!any_long) return true;
    memset(sub_max, 0, (size_t)root_size);
    for (int s = 0; s < n; ++s) {
        const int l = lens[s];
        if (l > root) { uint8_t& m = sub_max[code_of[s] & (uint32_t)(root_size - 1)]; if (l > m) m = (uint8_t)l; }
    }
    int free_at = root_size;
    for (int p = 0; p < root_size; ++p) {
        if (!sub_max[p]) continue;
        const int sub_bits = sub_max[p] - root, size = 1 << sub_bits;
        if (free_at + size > cap || free_at > 0xffff) return false;
        for (int i = 0; i < size; ++i) table[free_at + i] = 0;
        table[p] = entry(kSub, (uint32_t)free_at, (uint32_t)sub_bits, (uint32_t)root);
        free_at += size;
    }
    for (int s = 0; s < n; ++s) {
        const int l = lens[s];
        if (l <= root) continue;
        const uint32_t rev = code_of[s], e = table[rev & (uint32_t)(root_size - 1)];
        const uint32_t start = e_value(e), size = 1u << e_extra(e);
        for (uint32_t i = rev >> root; i < size; i += 1u << (l - root)) table[start + i] = symbol_entry(alpha, s, (uint32_t)(l - root));
    }
    return true;
}

struct Work {
    uint32_t lit[kLitCap], dist[kDistCap], pre[1 << kPreBits];
    uint32_t fixed_lit[1 << kLitBits], fixed_dist[1 << kDistBits];
    bool fixed_ready = false;
};

inline uint64_t load64(const uint8_t* p) { uint64_t v; memcpy(&v, p, 8); return v; }   // little-endian hosts (x86-64, aarch64)
inline void store64(uint8_t* p, uint64_t v) { memcpy(p, &v, 8); }

}  // namespace

bool fast_inflate(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_cap, size_t* out_len)
{
    static thread_local Work w;
    const uint8_t* ip = in;
    const uint8_t* const in_end = in + in_len;
    uint8_t* op = out;
    uint8_t* const out_end = out + out_cap;
    uint64_t bb = 0;        // bit buffer: stream bits, next bit = bit 0
    int bc = 0;             // valid bits in bb (the bits above them are either zero or the stream's own next bits)

    // One unaligned load tops the buffer up to >= 56 bits.  Loads stop once ip has passed the input: every real bit is in bb by
    // then, a valid stream needs no more, and a stream that does runs bc negative and is declined.
#define BVC_REFILL()                                                                          \
    do {                                                                                      \
        if (bc < 0) return false;                                                             \
        if (ip <= in_end) { bb |= load64(ip) << bc; ip += (63 - bc) >> 3; bc |= 56; }          \
    } while (0)
#define BVC_TAKE(n) (bb >>= (n), bc -= (int)(n))
#define BVC_BITS(n) ((uint32_t)bb & ((1u << (n)) - 1u))

    for (bool last = false; !last;) {
        BVC_REFILL();
        last = BVC_BITS(1) != 0;
        const uint32_t type = ((uint32_t)bb >> 1) & 3u;
        BVC_TAKE(3);
        const uint32_t *lit, *dist;
        if (type == 0) {                                               // 3.2.4 stored
            BVC_TAKE((unsigned)bc & 7u);
            if (bc < 0) return false;
            ip -= bc >> 3;                                             // hand the whole bytes still in the buffer back
            bb = 0; bc = 0;
            if (ip + 4 > in_end) return false;
            const uint32_t len = (uint32_t)ip[0] | ((uint32_t)ip[1] << 8), nlen = (uint32_t)ip[2] | ((uint32_t)ip[3] << 8);
            ip += 4;
            if ((len ^ nlen) != 0xffffu || len > (size_t)(in_end - ip) || op > out_end || len > (size_t)(out_end - op)) return false;
            memcpy(op, ip, len);
            op += len; ip += len;
            continue;
        } else if (type == 1) {                                        // 3.2.6 fixed codes
            if (!w.fixed_ready) {
                uint8_t l[288];
                for (int i = 0; i < 144; ++i) l[i] = 8;
                for (int i = 144; i < 256; ++i) l[i] = 9;
                for (int i = 256; i < 280; ++i) l[i] = 7;
                for (int i = 280; i < 288; ++i) l[i] = 8;
                if (!build_table(l, 288, kAlphaLit, kLitBits, w.fixed_lit, 1 << kLitBits)) return false;
                for (int i = 0; i < 32; ++i) l[i] = 5;
                if (!build_table(l, 32, kAlphaDist, kDistBits, w.fixed_dist, 1 << kDistBits)) return false;
                w.fixed_ready = true;
            }
            lit = w.fixed_lit; dist = w.fixed_dist;
        } else if (type == 2) {                                        // 3.2.7 dynamic codes
            const uint32_t hlit = BVC_BITS(5) + 257; BVC_TAKE(5);
            const uint32_t hdist = BVC_BITS(5) + 1; BVC_TAKE(5);
            const uint32_t hclen = BVC_BITS(4) + 4; BVC_TAKE(4);
            if (hlit > 286 || hdist > 30) return false;
            uint8_t pl[19] = {0};
            BVC_REFILL();                                              // 19 x 3 = 57 bits: two refills
            for (uint32_t i = 0; i < hclen; ++i) {
                if (i == 12) BVC_REFILL();
                pl[kPreOrder[i]] = (uint8_t)BVC_BITS(3); BVC_TAKE(3);
            }
            if (!build_table(pl, 19, kAlphaPre, kPreBits, w.pre, 1 << kPreBits)) return false;
            uint8_t lens[286 + 30 + 138];
            uint32_t n = 0;
            const uint32_t total = hlit + hdist;
            while (n < total) {
                BVC_REFILL();
                const uint32_t e = w.pre[BVC_BITS(kPreBits)];
                if (e_kind(e) != kLiteral) return false;
                BVC_TAKE(e_nbits(e));
                const uint32_t sym = e_value(e);
                if (sym < 16) { lens[n++] = (uint8_t)sym; continue; }
                uint32_t rep; uint8_t v = 0;
                if (sym == 16) { if (n == 0) return false; v = lens[n - 1]; rep = 3 + BVC_BITS(2); BVC_TAKE(2); }
                else if (sym == 17) { rep = 3 + BVC_BITS(3); BVC_TAKE(3); }
                else { rep = 11 + BVC_BITS(7); BVC_TAKE(7); }
                if (n + rep > total) return false;
                memset(lens + n, v, rep);
                n += rep;
            }
            if (lens[256] == 0) return false;                          // no end-of-block code
            if (!build_table(lens, (int)hlit, kAlphaLit, kLitBits, w.lit, kLitCap)) return false;
            if (!build_table(lens + hlit, (int)hdist, kAlphaDist, kDistBits, w.dist, kDistCap)) return false;
            lit = w.lit; dist = w.dist;
        } else return false;

        // ---- the symbols of the block: one refill per literal run / per (length, distance) pair ---------------------------
        for (;;) {
            if (op > out_end) return false;                            // (writes may run into the slack behind out_cap, never past it)
            BVC_REFILL();
            uint32_t e = lit[BVC_BITS(kLitBits)];
            if (__builtin_expect(e_kind(e) == kSub, 0)) { BVC_TAKE(kLitBits); e = lit[e_value(e) + BVC_BITS(e_extra(e))]; }
            uint64_t saved = bb;                                       // the extra bits are cut out of this copy, off the critical path
            BVC_TAKE(e_nbits(e));
            if (__builtin_expect(e_kind(e) != kBase, 0)) {
                if (e_kind(e) == kEob) break;
                if (e_kind(e) != kLiteral) return false;
                *op++ = (uint8_t)e_value(e);
                // up to two more literals from the root table on the bits already here (<= 15 + 10 + 10 of >= 56)
                e = lit[BVC_BITS(kLitBits)];
                if (e_kind(e) != kLiteral) continue;
                BVC_TAKE(e_nbits(e));
                *op++ = (uint8_t)e_value(e);
                e = lit[BVC_BITS(kLitBits)];
                if (e_kind(e) != kLiteral) continue;
                BVC_TAKE(e_nbits(e));
                *op++ = (uint8_t)e_value(e);
                continue;
            }
            // <= 20 bits gone, >= 36 left: a distance needs <= 15 + 13
            uint32_t d = dist[BVC_BITS(kDistBits)];
            const uint32_t length = e_value(e) + ((uint32_t)(saved >> (e_nbits(e) - e_extra(e))) & ((1u << e_extra(e)) - 1u));
            if (__builtin_expect(e_kind(d) == kSub, 0)) { BVC_TAKE(kDistBits); d = dist[e_value(d) + BVC_BITS(e_extra(d))]; }
            saved = bb;
            BVC_TAKE(e_nbits(d));
            if (__builtin_expect(e_kind(d) != kBase, 0)) return false;
            const uint32_t offset = e_value(d) + ((uint32_t)(saved >> (e_nbits(d) - e_extra(d))) & ((1u << e_extra(d)) - 1u));
            if (offset > (size_t)(op - out)) return false;
            const uint8_t* src = op - offset;
            uint8_t* dst = op;
            op += length;                                              // <= 258 bytes: stays inside the slack even from out_end
            if (offset >= 8) {
                do { store64(dst, load64(src)); dst += 8; src += 8; } while (dst < op);
            } else if (offset == 1) {
                const uint64_t v = 0x0101010101010101ull * src[0];
                do { store64(dst, v); dst += 8; } while (dst < op);
            } else {
                do { *dst++ = *src++; } while (dst < op);
            }
        }
    }
#undef BVC_REFILL
#undef BVC_TAKE
#undef BVC_BITS
    if (bc < 0 || op > out_end) return false;
    // bits consumed = 8 * (bytes taken from the input) - (bits still buffered): must lie inside the input
    const long long used_bits = (long long)(ip - in) * 8 - bc;
    if (used_bits > (long long)in_len * 8) return false;
    *out_len = (size_t)(op - out);
    return true;
}


// ---- CRC-32 by carry-less multiplication -------------------------------------------------------------------------------
#if defined(__x86_64__)
namespace {
// x^(64*k) mod P folding constants of the reflected gzip polynomial P = 0x1db710641 (Gopal et al., "Fast CRC computation for
// generic polynomials using PCLMULQDQ"): fold by 512 bits, by 128 bits, 96 -> 64 bits, and the Barrett pair (P, floor(x^64 / P))
__attribute__((target("pclmul,sse4.1"))) uint32_t crc32_clmul(const uint8_t* buf, size_t len, uint32_t crc /* pre-inverted */)
{
    const __m128i k1k2 = _mm_set_epi64x(0x01c6e41596ll, 0x0154442bd4ll), k3k4 = _mm_set_epi64x(0x00ccaa009ell, 0x01751997d0ll);
    const __m128i k5 = _mm_set_epi64x(0, 0x0163cd6124ll), poly = _mm_set_epi64x(0x01f7011641ll, 0x01db710641ll);
    const __m128i* p = reinterpret_cast<const __m128i*>(buf);
    __m128i x1 = _mm_loadu_si128(p), x2 = _mm_loadu_si128(p + 1), x3 = _mm_loadu_si128(p + 2), x4 = _mm_loadu_si128(p + 3);
    x1 = _mm_xor_si128(x1, _mm_cvtsi32_si128((int)crc));
    p += 4; len -= 64;
    while (len >= 64) {
        const __m128i a1 = _mm_clmulepi64_si128(x1, k1k2, 0x00), a2 = _mm_clmulepi64_si128(x2, k1k2, 0x00);
        const __m128i a3 = _mm_clmulepi64_si128(x3, k1k2, 0x00), a4 = _mm_clmulepi64_si128(x4, k1k2, 0x00);
        x1 = _mm_xor_si128(_mm_xor_si128(_mm_clmulepi64_si128(x1, k1k2, 0x11), a1), _mm_loadu_si128(p));
        x2 = _mm_xor_si128(_mm_xor_si128(_mm_clmulepi64_si128(x2, k1k2, 0x11), a2), _mm_loadu_si128(p + 1));
        x3 = _mm_xor_si128(_mm_xor_si128(_mm_clmulepi64_si128(x3, k1k2, 0x11), a3), _mm_loadu_si128(p + 2));
        x4 = _mm_xor_si128(_mm_xor_si128(_mm_clmulepi64_si128(x4, k1k2, 0x11), a4), _mm_loadu_si128(p + 3));
        p += 4; len -= 64;
    }
#define BVC_FOLD128(acc, next) _mm_xor_si128(_mm_xor_si128(_mm_clmulepi64_si128(acc, k3k4, 0x11), _mm_clmulepi64_si128(acc, k3k4, 0x00)), next)
    x1 = BVC_FOLD128(x1, x2); x1 = BVC_FOLD128(x1, x3); x1 = BVC_FOLD128(x1, x4);
    while (len >= 16) { x1 = BVC_FOLD128(x1, _mm_loadu_si128(p)); ++p; len -= 16; }
#undef BVC_FOLD128
    const __m128i mask32 = _mm_setr_epi32(~0, 0, ~0, 0);
    __m128i t = _mm_clmulepi64_si128(x1, k3k4, 0x10);
    x1 = _mm_xor_si128(_mm_srli_si128(x1, 8), t);
    t = _mm_srli_si128(x1, 4);
    x1 = _mm_xor_si128(_mm_clmulepi64_si128(_mm_and_si128(x1, mask32), k5, 0x00), t);
    t = _mm_and_si128(_mm_clmulepi64_si128(_mm_and_si128(x1, mask32), poly, 0x10), mask32);
    x1 = _mm_xor_si128(x1, _mm_clmulepi64_si128(t, poly, 0x00));
    return (uint32_t)_mm_extract_epi32(x1, 1);
}
}  // namespace
#endif

uint32_t fast_crc32(const uint8_t* buf, size_t len)
{
#if defined(__x86_64__)
    static const bool have = __builtin_cpu_supports("pclmul") && __builtin_cpu_supports("sse4.1");
    if (have && len >= 64) {
        const size_t body = len & ~(size_t)15;
        const uint32_t c = ~crc32_clmul(buf, body, 0xffffffffu);
        return body == len ? c : (uint32_t)crc32(c, buf + body, (uInt)(len - body));
    }
#endif
    return (uint32_t)crc32(crc32(0L, Z_NULL, 0), buf, (uInt)len);
}

}  // namespace bvc_host
