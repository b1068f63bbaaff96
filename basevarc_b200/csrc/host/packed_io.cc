// packed_io.cc -- see packed_io.hpp and include/bvc_host.h ("packed tile files").
#include "packed_io.hpp"

#include <string.h>

#include <algorithm>

namespace bvc_host {

namespace {
const char kPad[16] = {0};
inline size_t pad16(size_t n) { return (16 - (n & 15)) & 15; }
}  // namespace

bool PackedIo::put(const void* p, size_t n)
{
    if (n && fwrite(p, 1, n, f_) != n) return false;
    const size_t z = pad16(n);
    return !z || fwrite(kPad, 1, z, f_) == z;
}

bool PackedIo::get(void* p, size_t n)
{
    if (n && fread(p, 1, n, f_) != n) return false;
    char skip[16];
    const size_t z = pad16(n);
    return !z || fread(skip, 1, z, f_) == z;
}

// ---- writer ------------------------------------------------------------------------------------------------
bool PackedIo::create(const std::string& path, int32_t n_samples, const std::string& names, std::string* err)
{
    f_ = fopen(path.c_str(), "wb");
    if (!f_) { if (err) *err = "cannot write " + path; return false; }
    memset(&hdr_, 0, sizeof hdr_);
    memcpy(hdr_.magic, kPackedMagic, 8);
    hdr_.version = kPackedVersion;
    hdr_.n_samples = (uint32_t)n_samples;
    hdr_.names_bytes = (uint32_t)names.size();
    // n_tiles stays 0 until finish(): a file cut short never validates (PackedIo::is_complete)
    if (fwrite(&hdr_, sizeof hdr_, 1, f_) != 1 || !put(names.data(), names.size())) { if (err) *err = "write error on " + path; return false; }
    return true;
}

bool PackedIo::add_tile(const Packer& pk, std::string* err)
{
    const size_t S = (size_t)pk.n_sites(), R = pk.n_, N = (size_t)pk.n_samples_, nb = (N + 7) / 8;
    PackedTileHeader th;
    memset(&th, 0, sizeof th);
    th.magic = kTileMagic; th.n_sites = (uint32_t)S; th.n_entries = (uint32_t)R;
    const bool packed = pk.pk8_ok_ && pk.ndict_ > 0;
    if (packed) { th.n_dict = (uint32_t)pk.ndict_; memcpy(th.dict, pk.dict_, 32); }
    // covered-sample bitmap per site: the entries of a site are in ascending sample order (src/BaseVarC.cpp:407-441),
    // so N bits per site replace a 4-byte sample index per entry
    bitmap_.assign(S * nb, 0);
    for (size_t s = 0; s < S; ++s)
        for (uint32_t e = pk.entry_off_[s]; e < pk.entry_off_[s + 1]; ++e) {
            const uint32_t j = pk.smp_[e];
            if (j >= N) { if (err) *err = "sample index out of range in a packed tile"; return false; }
            bitmap_[s * nb + (j >> 3)] |= (uint8_t)(1u << (j & 7));
        }
    // indel entries: index, stale qual (the one-byte form does not carry it; WriteVcf prints it, src/BaseType.cpp:156-163), token
    ie_.clear(); iq_.clear(); io_.assign(1, 0u); itext_.clear();
    for (auto& kv : pk.indel_) ie_.push_back((uint32_t)kv.first);
    std::sort(ie_.begin(), ie_.end());
    for (uint32_t e : ie_) {
        iq_.push_back(pk.qual_[e]);
        itext_ += pk.indel_.at((int64_t)e);
        io_.push_back((uint32_t)itext_.size());
    }
    th.n_indel = (uint32_t)ie_.size(); th.indel_bytes = (uint32_t)itext_.size();
    bool ok = fwrite(&th, sizeof th, 1, f_) == 1 && put(pk.pos_.data(), S * 4) && put(pk.ref_base_.data(), S) &&
              put(pk.entry_off_.data(), (S + 1) * 4) && put(bitmap_.data(), bitmap_.size());
    if (packed) ok = ok && put(pk.pk8_.data(), R);
    else ok = ok && put(pk.code_.data(), R) && put(pk.qual_.data(), R);
    ok = ok && put(pk.mapq_.data(), R) && put(pk.rpr_.data(), R) && put(ie_.data(), ie_.size() * 4) && put(iq_.data(), iq_.size()) &&
         put(io_.data(), io_.size() * 4) && put(itext_.data(), itext_.size());
    if (!ok) { if (err) *err = "write error on a packed tile file"; return false; }
    hdr_.n_tiles += 1;
    hdr_.max_sites = std::max(hdr_.max_sites, (uint32_t)S);
    hdr_.max_entries = std::max<uint64_t>(hdr_.max_entries, R);
    return true;
}

bool PackedIo::finish(std::string* err)
{
    bool ok = f_ && fseek(f_, 0, SEEK_SET) == 0 && fwrite(&hdr_, sizeof hdr_, 1, f_) == 1;
    if (f_) { ok = (fclose(f_) == 0) && ok; f_ = nullptr; }
    if (!ok && err) *err = "cannot finish a packed tile file";
    return ok;
}

// ---- reader ------------------------------------------------------------------------------------------------
bool PackedIo::open(const std::string& path, std::string* err)
{
    f_ = fopen(path.c_str(), "rb");
    if (!f_) { if (err) *err = "cannot open " + path; return false; }
    if (fread(&hdr_, sizeof hdr_, 1, f_) != 1 || memcmp(hdr_.magic, kPackedMagic, 8) != 0 || hdr_.version != kPackedVersion) {
        if (err) *err = path + " is not a packed tile file of this version";
        return false;
    }
    names_.resize(hdr_.names_bytes);
    if (!get(&names_[0], names_.size())) { if (err) *err = path + ": truncated"; return false; }
    tiles_read_ = 0;
    return true;
}

bool PackedIo::is_complete(const std::string& path, int32_t n_samples)
{
    PackedIo r;
    std::string e;
    // n_tiles is written last (finish): 0 means the writer never got there -- unless the window is empty, which the caller
    // cannot tell apart and therefore re-packs
    return r.open(path, &e) && r.hdr_.n_samples == (uint32_t)n_samples && r.hdr_.n_tiles > 0;
}

int PackedIo::next_tile(Packer& pk, std::string* err, bool lazy)
{
    if (tiles_read_ >= hdr_.n_tiles) return 0;
    auto bad = [&](const char* what) { if (err) *err = std::string("corrupt packed tile file: ") + what; return -1; };
    PackedTileHeader th;
    if (fread(&th, sizeof th, 1, f_) != 1 || th.magic != kTileMagic) return bad("tile header");
    const size_t S = th.n_sites, R = th.n_entries, N = (size_t)pk.n_samples_, nb = (N + 7) / 8;
    if (S > hdr_.max_sites || R > hdr_.max_entries || th.n_dict > BVC_PACK8_MAX_QUALS || hdr_.n_samples != (uint32_t)N) return bad("tile sizes");
    pk.clear();
    pk.pos_.resize(S); pk.ref_base_.resize(S); pk.entry_off_.resize(S + 1);
    pk.reserve_entries(R + 16);
    bitmap_.resize(S * nb);
    if (!get(pk.pos_.data(), S * 4) || !get(pk.ref_base_.data(), S) || !get(pk.entry_off_.data(), (S + 1) * 4) || !get(bitmap_.data(), bitmap_.size()))
        return bad("site arrays");
    if (pk.entry_off_[0] != 0 || pk.entry_off_[S] != R) return bad("entry offsets");
    for (size_t s = 0; s < S; ++s) if (pk.entry_off_[s] > pk.entry_off_[s + 1]) return bad("entry offsets");
    const bool packed = th.n_dict > 0;
    if (packed) {
        if (!get(pk.pk8_.data(), R)) return bad("packed8 array");
        // (code, qual) for the host-side formatter from the one-byte form (the GPU library takes the byte itself)
        uint8_t code_lut[256], qual_lut[256];
        for (unsigned b = 0; b < 256; ++b) {
            const unsigned qi = BVC_PACK8_QI(b), st = BVC_PACK8_STRAND(b), base = BVC_PACK8_BASE(b);
            code_lut[b] = qi < BVC_PACK8_MAX_QUALS ? BVC_MAKE_CODE(base, st, 0) : qi == BVC_PACK8_INDEL ? BVC_MAKE_CODE(base, st, 1) : BVC_MAKE_CODE(4, st, 1);
            qual_lut[b] = qi < th.n_dict ? th.dict[qi] : (uint8_t)0;
        }
        const uint8_t* p = pk.pk8_.data();
        uint8_t* c = pk.code_.data(); uint8_t* q = pk.qual_.data();
        if (lazy) memset(c, 0, R);                               // the indel entries' codes follow with the indel table
        else for (size_t e = 0; e < R; ++e) { c[e] = code_lut[p[e]]; q[e] = qual_lut[p[e]]; }
        memcpy(pk.dict_, th.dict, 32);
        pk.ndict_ = (int)th.n_dict; pk.pk8_ok_ = true;
        for (int i = 0; i < pk.ndict_; ++i) pk.qidx_[pk.dict_[i] & 127] = (int8_t)i;
    } else {
        if (!get(pk.code_.data(), R) || !get(pk.qual_.data(), R)) return bad("code / qual arrays");
        pk.pk8_ok_ = false;
        lazy = false;                                            // two-array tiles are stored expanded
    }
    if (!get(pk.mapq_.data(), R) || !get(pk.rpr_.data(), R)) return bad("mapq / rpr arrays");
    if (lazy) {
        // covered samples stay in their bitmaps (Packer::expand_site); the rows must still add up to the offsets
        for (size_t s = 0; s < S; ++s) {
            const uint8_t* bm = &bitmap_[s * nb];
            size_t n = 0, k = 0;
            for (; k + 8 <= nb; k += 8) { uint64_t w; memcpy(&w, bm + k, 8); n += (size_t)__builtin_popcountll(w); }
            for (; k < nb; ++k) n += (size_t)__builtin_popcount(bm[k]);
            if (n != (size_t)(pk.entry_off_[s + 1] - pk.entry_off_[s])) return bad("bitmap / offsets disagree");
        }
        pk.lazy_bitmap_.swap(bitmap_);
    } else {
        // sample indices back from the bitmaps
        uint32_t* smp = pk.smp_.data();
        for (size_t s = 0; s < S; ++s) {
            size_t e = pk.entry_off_[s];
            const size_t e1 = pk.entry_off_[s + 1];
            const uint8_t* bm = &bitmap_[s * nb];
            for (size_t k = 0; k < nb; ++k)
                for (unsigned w = bm[k]; w; w &= w - 1) { if (e >= e1) return bad("bitmap / offsets disagree"); smp[e++] = (uint32_t)(k * 8 + (unsigned)__builtin_ctz(w)); }
            if (e != e1) return bad("bitmap / offsets disagree");
        }
    }
    ie_.resize(th.n_indel); iq_.resize(th.n_indel); io_.resize((size_t)th.n_indel + 1); itext_.resize(th.indel_bytes);
    if (!get(ie_.data(), ie_.size() * 4) || !get(iq_.data(), iq_.size()) || !get(io_.data(), io_.size() * 4) || !get(&itext_[0], itext_.size()))
        return bad("indel table");
    for (size_t i = 0; i < ie_.size(); ++i) {
        if (ie_[i] >= R || io_[i] > io_[i + 1] || io_[i + 1] > itext_.size()) return bad("indel table");
        pk.qual_[ie_[i]] = iq_[i];
        if (lazy) {
            const unsigned b = pk.pk8_[ie_[i]], qi = BVC_PACK8_QI(b);
            if (qi < BVC_PACK8_MAX_QUALS) return bad("indel table points at a counted entry");
            pk.code_[ie_[i]] = qi == BVC_PACK8_INDEL ? BVC_MAKE_CODE(BVC_PACK8_BASE(b), BVC_PACK8_STRAND(b), 1) : BVC_MAKE_CODE(4, BVC_PACK8_STRAND(b), 1);
        }
        pk.indel_.emplace((int64_t)ie_[i], itext_.substr(io_[i], io_[i + 1] - io_[i]));
    }
    pk.lazy_ = lazy;
    pk.n_ = R;
    ++tiles_read_;
    return 1;
}

int PackedIo::adopt(Packer& pk, const bvc_tile* t, size_t s0, size_t s1, const int32_t* pos, const std::string& itok)
{
    pk.clear();
    const uint32_t e0 = t->entry_off[s0], e1 = t->entry_off[s1];
    pk.reserve_entries((size_t)(e1 - e0) + 16);
    for (size_t s = s0; s < s1; ++s) {
        if (t->entry_off[s + 1] == t->entry_off[s]) continue;                 // a site without entries prints nothing: not stored
        pk.pos_.push_back(pos ? pos[s] : (int32_t)(s + 1));
        pk.ref_base_.push_back(t->ref_base[s]);
        for (uint32_t e = t->entry_off[s]; e < t->entry_off[s + 1]; ++e) {
            const uint8_t c = t->code[e], q = t->qual[e];
            const unsigned base = BVC_CODE_BASE(c), st = BVC_CODE_STRAND(c);
            const size_t n = pk.n_;
            if (BVC_CODE_INDEL(c)) {
                if (base > 4) pk.pk8_ok_ = false;
                pk.pk8_[n] = BVC_PACK8(base == 4 ? BVC_PACK8_INDEL_N : BVC_PACK8_INDEL, st, base);
                pk.indel_.emplace((int64_t)n, itok);
            } else {
                if (base > 3 || q > 127) return BVC_ERR_ARG;
                pk.pk8_[n] = pk.pack_counted((uint8_t)base, (uint8_t)st, q);
            }
            pk.code_[n] = c; pk.qual_[n] = q; pk.mapq_[n] = t->mapq[e]; pk.rpr_[n] = t->rpr[e]; pk.smp_[n] = t->sample[e];
            pk.n_ = n + 1;
        }
        pk.entry_off_.push_back((uint32_t)pk.n_);
    }
    return BVC_OK;
}

}  // namespace bvc_host
