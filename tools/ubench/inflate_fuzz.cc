// inflate_fuzz: fast_inflate on valid, corrupted and random streams with EXACT-size heap buffers, for AddressSanitizer / UBSan:
//   g++ -O1 -g -fsanitize=address,undefined -std=c++17 -Ibasevarc_b200/csrc/host tools/ubench/inflate_fuzz.cc basevarc_b200/csrc/host/fast_inflate.cc -lz -o gpurun_out/scratch/inflate_fuzz
#include <zlib.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>
#include "fast_inflate.hpp"
static std::vector<unsigned char> deflate_raw(const std::vector<unsigned char>& d, int level, int strategy)
{
    z_stream z; memset(&z, 0, sizeof z);
    deflateInit2(&z, level, Z_DEFLATED, -15, 8, strategy);
    std::vector<unsigned char> o(deflateBound(&z, d.size()) + 64);
    z.next_in = const_cast<unsigned char*>(d.data()); z.avail_in = (uInt)d.size(); z.next_out = o.data(); z.avail_out = (uInt)o.size();
    deflate(&z, Z_FINISH); o.resize(z.total_out); deflateEnd(&z);
    return o;
}
int main(int argc, char** argv)
{
    const int iters = argc > 1 ? atoi(argv[1]) : 20000;
    srand(12345);
    long ok = 0, declined = 0, wrong = 0;
    for (int it = 0; it < iters; ++it) {
        const size_t n = (size_t)rand() % 65537;
        std::vector<unsigned char> d(n);
        const int kind = rand() % 4;
        for (size_t i = 0; i < n; ++i) d[i] = kind == 0 ? (unsigned char)rand() : kind == 1 ? (unsigned char)("0123,. \n"[rand() % 8]) : kind == 2 ? (unsigned char)(i % (1 + it % 9) + 'a') : (unsigned char)(rand() % 3);
        std::vector<unsigned char> c = deflate_raw(d, rand() % 10, rand() % 5);
        const int mode = rand() % 3;            // 0: intact, 1: bit flips, 2: truncated
        if (mode == 1 && !c.empty()) for (int k = 0; k < 1 + rand() % 3; ++k) c[(size_t)rand() % c.size()] ^= (unsigned char)(1 << (rand() % 8));
        if (mode == 2 && !c.empty()) c.resize((size_t)rand() % c.size());
        const size_t cap = mode == 0 && rand() % 4 == 0 && n > 0 ? n - 1 - (size_t)rand() % n % 7 : 65536;
        std::unique_ptr<unsigned char[]> in(new unsigned char[c.size() + bvc_host::kFastInflateInPad]);
        if (!c.empty()) memcpy(in.get(), c.data(), c.size());
        memset(in.get() + c.size(), 0xEE, bvc_host::kFastInflateInPad);
        std::unique_ptr<unsigned char[]> out(new unsigned char[cap + bvc_host::kFastInflateOutSlack]);
        size_t got = 0;
        const bool r = bvc_host::fast_inflate(in.get(), c.size(), out.get(), cap, &got);
        if (!r) { ++declined; if (mode == 0 && cap >= n) { ++wrong; printf("declined an intact stream: it %d n %zu\n", it, n); } continue; }
        ++ok;
        if (got > cap) { ++wrong; printf("overlong: it %d\n", it); }
        if (mode == 0 && (got != n || memcmp(out.get(), d.data(), n) != 0)) { ++wrong; printf("wrong output: it %d n %zu got %zu\n", it, n, got); }
    }
    printf("%d streams: %ld decoded, %ld declined, %ld wrong\n", iters, ok, declined, wrong);
    return wrong != 0;
}
