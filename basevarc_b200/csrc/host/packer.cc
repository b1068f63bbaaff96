// packer.cc -- see packer.hpp
#include "packer.hpp"

#include <string.h>

namespace bvc_host {

namespace {
// std::atoi on a field that ends at `end` (the reference's fields are NUL-terminated by strtok_r)
inline int field_atoi(const char* p, const char* end)
{
    while (p < end && (*p == ' ' || (*p >= '\t' && *p <= '\r'))) ++p;
    bool neg = false;
    if (p < end && (*p == '+' || *p == '-')) { neg = (*p == '-'); ++p; }
    unsigned v = 0;
    while (p < end && *p >= '0' && *p <= '9') { v = v * 10u + (unsigned)(*p - '0'); ++p; }
    return neg ? -(int)v : (int)v;
}
}  // namespace

void Packer::clear()
{
    pos_.clear(); ref_base_.clear(); entry_off_.assign(1, 0u);
    code_.clear(); qual_.clear(); mapq_.clear(); rpr_.clear(); smp_.clear();
    indel_.clear(); err_.clear();
    sample_ = 0;
}

void Packer::begin_site(int32_t pos, uint8_t ref_base)
{
    pos_.push_back(pos);
    ref_base_.push_back(ref_base);
    sample_ = 0;
}

int Packer::add_line(const char* line, size_t len)
{
    const char* p = line;
    const char* const end = line + len;
    while (p < end) {
        while (p < end && *p == ' ') ++p;                    // strtok_r(buf, " "): runs of delimiters are skipped
        if (p >= end) break;
        const char* tok = p;
        while (p < end && *p != ' ') ++p;
        const char* tend = p;
        const char c0 = tok[0];
        if (c0 != '+' && c0 != '-' && c0 != 'N' && c0 != '.') {
            // "base,mapq,qual,rpr,strand": up to five non-empty comma fields, each through atoi into the bitfield
            const char* f = tok;
            for (int i = 0; i < 5; ++i) {
                while (f < tend && *f == ',') ++f;
                if (f >= tend) break;
                const char* fe = f;
                while (fe < tend && *fe != ',') ++fe;
                const int v = field_atoi(f, fe);
                switch (i) {
                case 0: st_base_ = (uint8_t)(v & 7); break;
                case 1: st_mapq_ = (uint8_t)(v & 255); break;
                case 2: st_qual_ = (uint8_t)(v & 255); break;
                case 3: st_rpr_ = (uint8_t)(v & 255); break;
                case 4: st_strand_ = (uint8_t)(v & 1); break;
                }
                f = fe;
            }
            if (st_base_ != 4) {                             // N base: skipped, the sample index still advances
                if (st_base_ > 4 || st_qual_ > 127) {
                    err_ = "token with base > 4 or qual > 127 at site " + std::to_string(pos_.empty() ? 0 : pos_.back());
                    return BVC_ERR_ARG;
                }
                code_.push_back(BVC_MAKE_CODE(st_base_, st_strand_, 0));
                qual_.push_back(st_qual_); mapq_.push_back(st_mapq_); rpr_.push_back(st_rpr_);
                smp_.push_back(sample_);
            }
        } else if (c0 != '.') {
            // indel token: only is_indel and the text change; every other field keeps its stale value (Q1)
            indel_.emplace((int64_t)code_.size(), std::string(tok, (size_t)(tend - tok)));
            code_.push_back(BVC_MAKE_CODE(st_base_, st_strand_, 1));
            qual_.push_back(st_qual_ > 127 ? (uint8_t)127 : st_qual_); mapq_.push_back(st_mapq_); rpr_.push_back(st_rpr_);
            smp_.push_back(sample_);
        }
        ++sample_;
    }
    return BVC_OK;
}

int32_t Packer::end_site()
{
    const uint32_t n = (uint32_t)code_.size() - entry_off_.back();
    entry_off_.push_back((uint32_t)code_.size());
    return (int32_t)n;
}

void Packer::drop_last_site()
{
    if (pos_.empty() || entry_off_.size() < 2 || entry_off_.back() != entry_off_[entry_off_.size() - 2]) return;
    pos_.pop_back(); ref_base_.pop_back(); entry_off_.pop_back();
}

void Packer::adopt_state(const Packer& o)
{
    st_base_ = o.st_base_; st_mapq_ = o.st_mapq_; st_qual_ = o.st_qual_; st_rpr_ = o.st_rpr_; st_strand_ = o.st_strand_;
}

void Packer::tile(bvc_tile* out) const
{
    memset(out, 0, sizeof(*out));
    out->n_sites = n_sites();
    out->ref_base = ref_base_.data();
    out->entry_off = entry_off_.data();
    out->code = code_.data(); out->qual = qual_.data(); out->mapq = mapq_.data(); out->rpr = rpr_.data();
    out->sample = smp_.data();
}

const char* Packer::indel(int64_t entry) const
{
    auto it = indel_.find(entry);
    return it == indel_.end() ? nullptr : it->second.c_str();
}

}  // namespace bvc_host
