// batch_writer.cc -- a packed tile -> the reference's `--load` batch files (fixtures, benchmarks, the CPU-baseline
// run of the unmodified reference binary).  Format as written by bt_r (/root/reference/src/BaseVarC.cpp:467-534):
// PREFIX.tmp.thread.{t}/batch.{j}, BGZF; line 1 = sample names each followed by '\t'; then one line per site of
// window t with one token per sample of batch j, each followed by ' ':  "." | "b,mapq,qual,rpr,strand" | indel text.
// Window partition as the reference's (:399-403,497,523).
#include <stdio.h>
#include <string.h>
#include <sys/stat.h>

#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../../include/bvc_host.h"
#include "bgzf_io.hpp"

namespace {
inline char* put_u8(char* p, unsigned v)
{
    if (v >= 100) { *p++ = (char)('0' + v / 100); v %= 100; *p++ = (char)('0' + v / 10); *p++ = (char)('0' + v % 10); }
    else if (v >= 10) { *p++ = (char)('0' + v / 10); *p++ = (char)('0' + v % 10); }
    else *p++ = (char)('0' + v);
    return p;
}
}  // namespace

extern "C" int bvc_write_batch_files(const bvc_tile* t, int32_t n_samples, const char* prefix, int32_t threads, int32_t batch,
                                     int32_t level, const char* indel_token)
{
    if (!t || !prefix || n_samples <= 0 || threads <= 0 || batch <= 0 || !t->sample) return BVC_ERR_ARG;
    const size_t S = (size_t)t->n_sites;
    const size_t window = S % (size_t)threads + S / (size_t)threads;
    if (threads > 1 && (size_t)(threads - 1) * window > S) return BVC_ERR_ARG;
    const int nb = 1 + (n_samples - 1) / batch;
    const std::string itok = indel_token ? indel_token : "+A";
    std::vector<int> rcs((size_t)threads, 0);
    auto work = [&](int w) {
        const std::string dir = std::string(prefix) + ".tmp.thread." + std::to_string(w);
        mkdir(dir.c_str(), 0777);
        std::vector<std::unique_ptr<bvc_host::BgzfWriter>> out;
        for (int j = 0; j < nb; ++j) {
            out.emplace_back(new bvc_host::BgzfWriter(level));
            if (!out.back()->open(dir + "/batch." + std::to_string(j))) { rcs[(size_t)w] = BVC_ERR_ARG; return; }
            std::string h;
            for (int s = j * batch; s < n_samples && s < (j + 1) * batch; ++s) h += "S" + std::to_string(s) + "\t";
            h += "\n";
            out.back()->write(h);
        }
        const size_t s0 = std::min(S, (size_t)w * window), s1 = (w == threads - 1) ? S : std::min(S, (size_t)(w + 1) * window);
        std::string line;
        for (size_t s = s0; s < s1; ++s) {
            uint32_t e = t->entry_off[s];
            const uint32_t e1 = t->entry_off[s + 1];
            for (int j = 0; j < nb; ++j) {
                line.clear();
                const int a = j * batch, b = std::min(n_samples, (j + 1) * batch);
                for (int smp = a; smp < b; ++smp) {
                    if (e < e1 && t->sample[e] == (uint32_t)smp) {
                        const unsigned c = t->code[e];
                        if (BVC_CODE_INDEL(c)) line += itok;
                        else {
                            char tok[24]; char* p = tok;
                            *p++ = (char)('0' + BVC_CODE_BASE(c)); *p++ = ',';
                            p = put_u8(p, t->mapq[e]); *p++ = ',';
                            p = put_u8(p, t->qual[e]); *p++ = ',';
                            p = put_u8(p, t->rpr[e]); *p++ = ',';
                            *p++ = (char)('0' + BVC_CODE_STRAND(c));
                            line.append(tok, (size_t)(p - tok));
                        }
                        ++e;
                    } else line += '.';
                    line += ' ';
                }
                line += '\n';
                out[(size_t)j]->write(line);
            }
        }
        for (auto& o : out) if (!o->close()) rcs[(size_t)w] = BVC_ERR_ARG;
    };
    std::vector<std::thread> th;
    for (int w = 0; w < threads; ++w) th.emplace_back(work, w);
    for (auto& x : th) x.join();
    for (int rc : rcs) if (rc) return rc;
    return BVC_OK;
}
