// host_api.cc -- the plain-C surface of include/bvc_host.h over Packer / format_* (bvc_basetype_main is in rerun.cc)
#include <string.h>
#include <algorithm>
#include <string>

#include "../../../include/bvc_host.h"
#include "format.hpp"
#include "fast_inflate.hpp"
#include "packed_io.hpp"
#include "packer.hpp"
#include <sys/stat.h>

struct bvc_packer { bvc_host::Packer p; explicit bvc_packer(int32_t n) : p(n) {} };

namespace {
size_t emit(const std::string& s, char* buf, size_t cap)
{
    if (buf && cap) {
        const size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
        memcpy(buf, s.data(), n);
        buf[n] = 0;
    }
    return s.size();
}
}  // namespace

extern "C" {

bvc_packer* bvc_packer_create(int32_t n_samples) { return n_samples > 0 ? new bvc_packer(n_samples) : nullptr; }
void bvc_packer_destroy(bvc_packer* pk) { delete pk; }
void bvc_packer_clear(bvc_packer* pk) { if (pk) pk->p.clear(); }
void bvc_packer_begin_site(bvc_packer* pk, int32_t pos, uint8_t ref_base) { pk->p.begin_site(pos, ref_base); }
int bvc_packer_add_line(bvc_packer* pk, const char* line, size_t len) { return pk->p.add_line(line, len); }
int32_t bvc_packer_end_site(bvc_packer* pk) { return pk->p.end_site(); }
int32_t bvc_packer_n_sites(const bvc_packer* pk) { return pk->p.n_sites(); }
int64_t bvc_packer_n_entries(const bvc_packer* pk) { return pk->p.n_entries(); }
void bvc_packer_tile(const bvc_packer* pk, bvc_tile* out) { pk->p.tile(out); }
const int32_t* bvc_packer_positions(const bvc_packer* pk) { return pk->p.positions(); }
const char* bvc_packer_indel(const bvc_packer* pk, int64_t entry) { return pk->p.indel(entry); }
const char* bvc_packer_error(const bvc_packer* pk) { return pk->p.error().c_str(); }

size_t bvc_format_cvg(char* buf, size_t cap, const char* chr, int32_t pos, const bvc_tile* tile, int32_t site,
                      const bvc_site_result* rec, const bvc_packer* indel_src, int32_t n_groups, const int32_t* group_depth)
{
    std::string s;
    bvc_host::format_cvg(s, chr, pos, tile, site, rec, indel_src ? &indel_src->p : nullptr, n_groups, group_depth);
    return emit(s, buf, cap);
}

size_t bvc_format_vcf(char* buf, size_t cap, const char* chr, int32_t pos, const bvc_tile* tile, int32_t site,
                      const bvc_site_result* rec, int32_t n_samples, int32_t n_groups, const char* const* group_names,
                      const int32_t* group_depth, const double* group_af)
{
    std::string s;
    bvc_host::format_vcf(s, chr, pos, tile, site, rec, n_samples, n_groups, group_names, group_depth, group_af);
    return emit(s, buf, cap);
}

struct bvc_packed_reader { bvc_host::PackedIo io; std::string err; };

bvc_packed_reader* bvc_packed_open(const char* path, int32_t* n_samples, int32_t* n_tiles, int32_t* max_sites, int64_t* max_entries)
{
    if (!path) return nullptr;
    bvc_packed_reader* r = new bvc_packed_reader();
    if (!r->io.open(path, &r->err)) { delete r; return nullptr; }
    const bvc_host::PackedFileHeader& h = r->io.header();
    if (n_samples) *n_samples = (int32_t)h.n_samples;
    if (n_tiles) *n_tiles = (int32_t)h.n_tiles;
    if (max_sites) *max_sites = (int32_t)h.max_sites;
    if (max_entries) *max_entries = (int64_t)h.max_entries;
    return r;
}
int bvc_packed_next(bvc_packed_reader* r, bvc_packer* into) { return (r && into) ? r->io.next_tile(into->p, &r->err) : -1; }
const char* bvc_packed_names(const bvc_packed_reader* r) { return r ? r->io.names().c_str() : ""; }
const char* bvc_packed_error(const bvc_packed_reader* r) { return r ? r->err.c_str() : "no reader"; }
void bvc_packed_close(bvc_packed_reader* r) { delete r; }
int bvc_packed_is_complete(const char* path, int32_t n_samples) { return path && bvc_host::PackedIo::is_complete(path, n_samples) ? 1 : 0; }

int bvc_fast_inflate(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_cap, size_t* out_len)
{
    if ((!in && in_len) || (!out && out_cap) || !out_len) return 0;
    // the decoder wants readable padding behind the input and writable slack behind the output: private copies
    std::vector<uint8_t> ib(in_len + bvc_host::kFastInflateInPad, 0), ob(out_cap + bvc_host::kFastInflateOutSlack);
    if (in_len) memcpy(ib.data(), in, in_len);
    size_t n = 0;
    if (!bvc_host::fast_inflate(ib.data(), in_len, ob.data(), out_cap, &n) || n > out_cap) return 0;
    if (n) memcpy(out, ob.data(), n);
    *out_len = n;
    return 1;
}

uint32_t bvc_fast_crc32(const uint8_t* buf, size_t len) { return bvc_host::fast_crc32(buf, len); }

int bvc_write_packed_files(const bvc_tile* t, int32_t n_samples, const int32_t* pos, const char* prefix, int32_t threads,
                           int32_t tile_sites, const char* indel_token)
{
    if (!t || !prefix || n_samples <= 0 || threads <= 0 || tile_sites <= 0 || !t->sample || !t->code || !t->qual) return BVC_ERR_ARG;
    const size_t S = (size_t)t->n_sites;
    const size_t window = S % (size_t)threads + S / (size_t)threads;             // src/BaseVarC.cpp:399
    if (threads > 1 && (size_t)(threads - 1) * window > S) return BVC_ERR_ARG;
    std::string names;
    for (int s = 0; s < n_samples; ++s) names += "S" + std::to_string(s) + (s + 1 < n_samples ? "\t" : "");
    bvc_host::Packer pk(n_samples);
    for (int w = 0; w < threads; ++w) {
        const std::string dir = std::string(prefix) + ".tmp.thread." + std::to_string(w);
        mkdir(dir.c_str(), 0777);
        bvc_host::PackedIo out;
        std::string err;
        if (!out.create(dir + "/packed.bvc", n_samples, names, &err)) return BVC_ERR_ARG;
        const size_t s0 = std::min(S, (size_t)w * window), s1 = (w == threads - 1) ? S : std::min(S, (size_t)(w + 1) * window);
        for (size_t a = s0; a < s1; a += (size_t)tile_sites) {
            const size_t b = std::min(s1, a + (size_t)tile_sites);
            const int rc = bvc_host::PackedIo::adopt(pk, t, a, b, pos, indel_token ? indel_token : "+A");
            if (rc != BVC_OK) return rc;
            if (pk.n_sites() > 0 && !out.add_tile(pk, &err)) return BVC_ERR_ARG;
        }
        if (!out.finish(&err)) return BVC_ERR_ARG;
    }
    return BVC_OK;
}

size_t bvc_format_fixed(char* buf, size_t cap, double v, int32_t digits)
{
    std::string s;
    bvc_host::format_fixed(s, v, digits);
    return emit(s, buf, cap);
}

}  // extern "C"
