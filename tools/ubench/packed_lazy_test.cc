// packed_lazy_test: every tile of the given packed tile files read fully and lazily (PackedIo::next_tile); after
// Packer::expand_site on every site the lazily read arrays must equal the fully read ones, and before it the indel entries'
// codes / quals must already be there (all the CVG line reads).
//   g++ -O2 -std=c++17 -Iinclude -Ibasevarc_b200/csrc/host tools/ubench/packed_lazy_test.cc basevarc_b200/csrc/host/packed_io.cc basevarc_b200/csrc/host/packer.cc -o gpurun_out/scratch/packed_lazy_test
#include <cstdio>
#include <cstring>
#include <string>
#include "packed_io.hpp"
using namespace bvc_host;
int main(int argc, char** argv)
{
    long tiles = 0, sites = 0, entries = 0, indels = 0;
    for (int i = 1; i < argc; ++i) {
        PackedIo a, b; std::string e;
        if (!a.open(argv[i], &e) || !b.open(argv[i], &e)) { fprintf(stderr, "%s\n", e.c_str()); return 1; }
        Packer pa((int32_t)a.header().n_samples), pb((int32_t)a.header().n_samples);
        for (;;) {
            const int ra = a.next_tile(pa, &e, false), rb = b.next_tile(pb, &e, true);
            if (ra != rb) { fprintf(stderr, "return codes differ: %d %d %s\n", ra, rb, e.c_str()); return 1; }
            if (ra <= 0) { if (ra < 0) { fprintf(stderr, "%s\n", e.c_str()); return 1; } break; }
            bvc_tile ta, tb; pa.tile(&ta); pb.tile(&tb);
            const size_t S = (size_t)ta.n_sites, R = ta.entry_off[S];
            if (!pb.lazy() && ta.packed8) { fprintf(stderr, "a one-byte tile was not read lazily\n"); return 1; }
            for (size_t x = 0; x < R; ++x)
                if (BVC_CODE_INDEL(ta.code[x]) != BVC_CODE_INDEL(tb.code[x]) || (BVC_CODE_INDEL(ta.code[x]) && (ta.code[x] != tb.code[x] || ta.qual[x] != tb.qual[x]))) {
                    fprintf(stderr, "indel entry %zu differs before expansion\n", x); return 1; }
            for (size_t s = 0; s < S; ++s) { pb.expand_site((int32_t)s); pb.expand_site((int32_t)s); }
            if (memcmp(ta.code, tb.code, R) || memcmp(ta.qual, tb.qual, R) || memcmp(ta.sample, tb.sample, R * 4) || memcmp(ta.mapq, tb.mapq, R) ||
                memcmp(ta.rpr, tb.rpr, R) || memcmp(ta.entry_off, tb.entry_off, (S + 1) * 4)) { fprintf(stderr, "tile %ld differs after expansion\n", tiles); return 1; }
            for (size_t x = 0; x < R; ++x) {
                const char* ia = pa.indel((int64_t)x); const char* ib = pb.indel((int64_t)x);
                if ((ia == nullptr) != (ib == nullptr) || (ia && strcmp(ia, ib))) { fprintf(stderr, "indel text differs\n"); return 1; }
                if (ia) ++indels;
            }
            ++tiles; sites += (long)S; entries += (long)R;
        }
    }
    printf("packed_lazy_test: %ld tiles, %ld sites, %ld entries, %ld indel entries: lazy == full\n", tiles, sites, entries, indels);
    return 0;
}
