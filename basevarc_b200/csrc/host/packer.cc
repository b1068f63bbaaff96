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
    n_ = 0;
    indel_.clear(); err_.clear();
    sample_ = 0;
    for (auto& q : qidx_) q = -1;
    ndict_ = 0; pk8_ok_ = true;
    trace_.clear(); early_.clear(); site_numeric_ = false;
    lazy_ = false;
}

void Packer::reserve_entries(size_t need)
{
    if (need <= cap_) return;
    size_t c = cap_ ? cap_ : (need > 4096 ? need : 4096);      // the first request is taken as given (Packer::reserve: fixed capacity)
    while (c < need) c *= 2;
    code_.resize(c); qual_.resize(c); mapq_.resize(c); rpr_.resize(c); smp_.resize(c); pk8_.resize(c);
    cap_ = c;
}

int Packer::buffers(const void** ptr, size_t* bytes, int max) const
{
    const void* p[6] = {code_.data(), qual_.data(), mapq_.data(), rpr_.data(), pk8_.data(), smp_.data()};
    const size_t b[6] = {cap_, cap_, cap_, cap_, cap_, cap_ * sizeof(uint32_t)};
    int n = 0;
    for (int i = 0; i < 6 && n < max; ++i) if (p[i] && b[i]) { ptr[n] = p[i]; bytes[n] = b[i]; ++n; }
    return n;
}

void Packer::begin_site(int32_t pos, uint8_t ref_base)
{
    pos_.push_back(pos);
    ref_base_.push_back(ref_base);
    sample_ = 0;
    site_numeric_ = false;
}

// One token through the reference's general grammar: strtok_r on ",", std::atoi per field (whitespace, sign, digits),
// empty fields skipped, fewer than five fields leave the rest stale (src/BaseVarC.cpp:410-438).
int Packer::add_token_slow(const char* tok, const char* tend)
{
    const char c0 = tok[0];
    if (c0 != '+' && c0 != '-' && c0 != 'N' && c0 != '.') {
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
        site_numeric_ = true;
        if (st_base_ != 4) {                             // N base: skipped, the sample index still advances
            if (st_base_ > 4 || st_qual_ > 127) {
                err_ = "token with base > 4 or qual > 127 at site " + std::to_string(pos_.empty() ? 0 : pos_.back());
                return BVC_ERR_ARG;
            }
            code_[n_] = BVC_MAKE_CODE(st_base_, st_strand_, 0);
            qual_[n_] = st_qual_; mapq_[n_] = st_mapq_; rpr_[n_] = st_rpr_; smp_[n_] = sample_;
            pk8_[n_] = pack_counted(st_base_, st_strand_, st_qual_);
            ++n_;
        }
    } else if (c0 != '.') {
        // indel token: only is_indel and the text change; every other field keeps its stale value (Q1)
        if (track_ && !site_numeric_) early_.emplace_back((uint32_t)(ref_base_.size() - 1), (uint32_t)n_);
        indel_.emplace((int64_t)n_, std::string(tok, (size_t)(tend - tok)));
        code_[n_] = BVC_MAKE_CODE(st_base_, st_strand_, 1);
        qual_[n_] = st_qual_ > 127 ? (uint8_t)127 : st_qual_; mapq_[n_] = st_mapq_; rpr_[n_] = st_rpr_; smp_[n_] = sample_;
        if (st_base_ > 4) pk8_ok_ = false;               // stale base 5..7: not representable (and nothing the writer emits)
        pk8_[n_] = BVC_PACK8(st_base_ == 4 ? BVC_PACK8_INDEL_N : BVC_PACK8_INDEL, st_strand_, st_base_);
        ++n_;
    }
    return BVC_OK;
}

// Two fast paths cover what the reference's writer emits (src/BaseVarC.cpp:513-521): ". " for an uncovered sample and
// "b,m..,q..,r..,s " with b, s one digit and 1-3 digit numbers.  Anything else takes add_token_slow.
int Packer::add_line(const char* line, size_t len)
{
    reserve_entries(n_ + len / 2 + 2);                       // a token and its delimiter are >= 2 bytes
    const char* p = line;
    const char* const end = line + len;
    uint8_t* const code = code_.data(); uint8_t* const qual = qual_.data();
    uint8_t* const mapq = mapq_.data(); uint8_t* const rpr = rpr_.data();
    uint32_t* const smp = smp_.data();
    uint8_t* const pk8 = pk8_.data();
    size_t n = n_;
    uint32_t sample = sample_;
    auto digit = [](char c) { return (unsigned)(c - '0') <= 9u; };
    while (p < end) {
        if (*p == ' ') { ++p; continue; }                    // strtok_r(buf, " "): runs of delimiters are skipped
        if (*p == '.' && (p + 1 == end || p[1] == ' ')) { p += 2; ++sample; continue; }
        if (end - p >= 16 && digit(p[0]) && p[1] == ',') {
            const char* q = p + 2;
            unsigned v[3];
            bool ok = true;
            for (int i = 0; i < 3 && ok; ++i) {
                if (!digit(*q)) { ok = false; break; }
                unsigned x = (unsigned)(*q++ - '0');
                if (digit(*q)) { x = x * 10u + (unsigned)(*q++ - '0'); if (digit(*q)) x = x * 10u + (unsigned)(*q++ - '0'); }
                if (*q != ',') { ok = false; break; }
                ++q;
                v[i] = x;
            }
            if (ok && digit(q[0]) && q[1] == ' ') {
                const unsigned b = (unsigned)(p[0] - '0') & 7u;
                const unsigned q8 = v[1] & 255u;
                if (b <= 4 && (b == 4 || q8 <= 127)) {       // errors are reported by the slow path
                    st_base_ = (uint8_t)b; st_mapq_ = (uint8_t)(v[0] & 255u); st_qual_ = (uint8_t)q8;
                    st_rpr_ = (uint8_t)(v[2] & 255u); st_strand_ = (uint8_t)((unsigned)(q[0] - '0') & 1u);
                    site_numeric_ = true;
                    if (b != 4) {
                        code[n] = BVC_MAKE_CODE(st_base_, st_strand_, 0);
                        qual[n] = st_qual_; mapq[n] = st_mapq_; rpr[n] = st_rpr_; smp[n] = sample;
                        pk8[n] = pack_counted(st_base_, st_strand_, st_qual_);
                        ++n;
                    }
                    p = q + 2; ++sample;
                    continue;
                }
            }
        }
        const char* tok = p;
        while (p < end && *p != ' ') ++p;
        n_ = n; sample_ = sample;
        const int rc = add_token_slow(tok, p);
        if (rc != BVC_OK) return rc;
        n = n_;
        ++sample;
    }
    n_ = n; sample_ = sample;
    return BVC_OK;
}

int32_t Packer::end_site()
{
    const uint32_t n = (uint32_t)n_ - entry_off_.back();
    entry_off_.push_back((uint32_t)n_);
    if (track_) trace_.push_back(SiteTrace{(uint8_t)site_numeric_, st_base_, st_mapq_, st_qual_, st_rpr_, st_strand_});
    return (int32_t)n;
}

void Packer::patch_entry(uint32_t e, const SiteTrace& st)
{
    code_[e] = BVC_MAKE_CODE(st.base, st.strand, 1);
    qual_[e] = st.qual > 127 ? (uint8_t)127 : st.qual; mapq_[e] = st.mapq; rpr_[e] = st.rpr;
    if (st.base > 4) pk8_ok_ = false;
    pk8_[e] = BVC_PACK8(st.base == 4 ? BVC_PACK8_INDEL_N : BVC_PACK8_INDEL, st.strand, st.base);
}

void Packer::append_raw(uint8_t code, uint8_t qual, uint8_t mapq, uint8_t rpr, uint32_t sample, const char* indel_text)
{
    reserve_entries(n_ + 1);
    code_[n_] = code; qual_[n_] = qual; mapq_[n_] = mapq; rpr_[n_] = rpr; smp_[n_] = sample; pk8_[n_] = 0;
    pk8_ok_ = false;
    if (indel_text) indel_.emplace((int64_t)n_, indel_text);
    ++n_;
}

void Packer::expand_site(int32_t site)
{
    if (!lazy_) return;
    const size_t nb = ((size_t)n_samples_ + 7) / 8;
    const uint32_t e0 = entry_off_[(size_t)site], e1 = entry_off_[(size_t)site + 1];
    for (uint32_t e = e0; e < e1; ++e) {
        const unsigned b = pk8_[e], qi = BVC_PACK8_QI(b), st = BVC_PACK8_STRAND(b), base = BVC_PACK8_BASE(b);
        if (qi < BVC_PACK8_MAX_QUALS) { code_[e] = BVC_MAKE_CODE(base, st, 0); qual_[e] = qi < (unsigned)ndict_ ? dict_[qi] : (uint8_t)0; }
        else code_[e] = qi == BVC_PACK8_INDEL ? BVC_MAKE_CODE(base, st, 1) : BVC_MAKE_CODE(4, st, 1);     // (stale qual: set with the indel table)
    }
    const uint8_t* bm = &lazy_bitmap_[(size_t)site * nb];
    uint32_t e = e0;
    for (size_t k = 0; k < nb && e < e1; ++k)
        for (unsigned w = bm[k]; w && e < e1; w &= w - 1) smp_[e++] = (uint32_t)(k * 8 + (unsigned)__builtin_ctz(w));
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
    if (pk8_ok_ && n_ > 0 && ndict_ > 0) { out->packed8 = pk8_.data(); out->qual_dict = dict_; out->n_qual_dict = ndict_; }
}

const char* Packer::indel(int64_t entry) const
{
    auto it = indel_.find(entry);
    return it == indel_.end() ? nullptr : it->second.c_str();
}

}  // namespace bvc_host
