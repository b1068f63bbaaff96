// inflate_bench: time BgzfReader over a set of batch files (thread-seconds of the "inflate+getline" phase in isolation)
//   g++ -O2 -std=c++17 -Ibasevarc_b200/csrc/host tools/ubench/inflate_bench.cc basevarc_b200/csrc/host/bgzf_io.cc basevarc_b200/csrc/host/fast_inflate.cc -lz -o gpurun_out/scratch/inflate_bench
#include <chrono>
#include <cstdio>
#include <string>
#include "bgzf_io.hpp"
int main(int argc, char** argv)
{
    using clk = std::chrono::steady_clock;
    size_t bytes = 0, lines = 0;
    unsigned long long h = 0;
    size_t zl = 0;
    const auto t0 = clk::now();
    for (int i = 1; i < argc; ++i) {
        bvc_host::BgzfReader r; std::string e;
        if (!r.open(argv[i], &e)) { fprintf(stderr, "%s\n", e.c_str()); return 1; }
        const char* p; size_t n;
        while (r.getline_view(&p, &n)) { bytes += n + 1; ++lines; h = h * 1315423911ull + (n ? (unsigned char)p[n - 1] : 7); }
        zl += r.zlib_blocks();
        if (r.failed()) { fprintf(stderr, "failed: %s\n", r.error().c_str()); return 1; }
    }
    const double s = std::chrono::duration<double>(clk::now() - t0).count();
    printf("%zu lines, %.1f MB inflated in %.3f s = %.1f MB/s (hash %llx, %zu blocks left to zlib)\n", lines, bytes / 1e6, s, bytes / 1e6 / s, h, zl);
    return 0;
}
