// host_api.cc -- the plain-C surface of include/bvc_host.h over Packer / format_* (bvc_basetype_main is in rerun.cc)
#include <string.h>
#include <string>

#include "../../../include/bvc_host.h"
#include "format.hpp"
#include "packer.hpp"

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

size_t bvc_format_fixed(char* buf, size_t cap, double v, int32_t digits)
{
    std::string s;
    bvc_host::format_fixed(s, v, digits);
    return emit(s, buf, cap);
}

}  // extern "C"
