// bgzf_pool_test: BgzfWriter (memory sink) -> BgzfReader round trips with the block pool, from several threads at once; for
// ThreadSanitizer / AddressSanitizer:
//   g++ -O1 -g -fsanitize=thread -std=c++17 -Ibasevarc_b200/csrc/host tools/ubench/bgzf_pool_test.cc basevarc_b200/csrc/host/bgzf_io.cc basevarc_b200/csrc/host/fast_inflate.cc -lz -lpthread -o gpurun_out/scratch/bgzf_pool_test
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>
#include "bgzf_io.hpp"
using namespace bvc_host;
static int worker(int id, const std::string& dir)
{
    unsigned seed = 1000u + (unsigned)id;
    std::string text;
    for (int i = 0; i < 60000; ++i) {
        const int n = rand_r(&seed) % 40;
        for (int k = 0; k < n; ++k) text += (rand_r(&seed) % 3) ? ". " : "2,27,28,65,0 ";
        text += '\n';
    }
    BgzfWriter w;
    w.open_memory();
    for (size_t off = 0; off < text.size();) { const size_t n = 1 + rand_r(&seed) % 5000; w.write(text.data() + off, std::min(n, text.size() - off)); off += n; }
    w.close(true);
    const std::string path = dir + "/pool_test." + std::to_string(id) + ".gz";
    FILE* f = fopen(path.c_str(), "wb");
    fwrite(w.memory().data(), 1, w.memory().size(), f);
    fclose(f);
    BgzfReader r; std::string e, line, back;
    if (!r.open(path, &e)) return 1;
    while (r.getline(line)) { back += line; back += '\n'; }
    remove(path.c_str());
    if (r.failed() || back != text) { fprintf(stderr, "thread %d: round trip differs (%zu vs %zu) %s\n", id, back.size(), text.size(), r.error().c_str()); return 1; }
    // a reader abandoned with blocks still queued
    return 0;
}
int main(int argc, char** argv)
{
    const std::string dir = argc > 1 ? argv[1] : ".";
    const int pool = argc > 2 ? atoi(argv[2]) : 3;
    InflatePool::instance().start(pool);
    std::vector<std::thread> th; std::vector<int> rc(4, 0);
    for (int i = 0; i < 4; ++i) th.emplace_back([&, i] { rc[i] = worker(i, dir); });
    for (auto& t : th) t.join();
    InflatePool::instance().stop();
    int bad = 0; for (int x : rc) bad += x;
    printf("bgzf_pool_test: pool %d, %s\n", pool, bad ? "FAILED" : "ok");
    return bad;
}
