// packer.hpp -- batch-file lines -> structure-of-arrays pileup tile (host side of include/bvc_host.h).
//
// Restates what the reference's window loop does with each line it reads (paths under /root/reference):
// strtok_r on " " then on ",", std::atoi into the AlleleInfo bitfield, N-bases skipped while the sample
// index still advances, indel tokens reusing the stale numeric fields (src/BaseVarC.cpp:393-441,
// src/BamProcess.h:30-39).  Output layout = bvc_tile (CSR over covered entries, one byte per field).
#pragma once
#include <stdint.h>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/bvc_gpu.h"

namespace bvc_host {

class Packer {
public:
    explicit Packer(int32_t n_samples) : n_samples_(n_samples) { entry_off_.push_back(0); for (auto& q : qidx_) q = -1; }
    void clear();
    void begin_site(int32_t pos, uint8_t ref_base);
    int add_line(const char* line, size_t len);
    int32_t end_site();
    void drop_last_site();                       // undo begin_site/end_site of a site that got no entries
    void adopt_state(const Packer& other);       // take over the stale-token state (next tile of the same window)
    int32_t n_sites() const { return (int32_t)ref_base_.size(); }
    int64_t n_entries() const { return (int64_t)n_; }
    void tile(bvc_tile* out) const;
    const int32_t* positions() const { return pos_.data(); }
    const char* indel(int64_t entry) const;
    const std::string& error() const { return err_; }
    int32_t n_samples() const { return n_samples_; }
    // Fixed capacity: the per-entry arrays are sized once and never move afterwards, so the dispatcher can page-lock them
    // (bvc_host_register) and the GPU library copies / reads them asynchronously.  buffers() lists them as (pointer, bytes).
    void reserve(size_t entries) { reserve_entries(entries); }
    size_t capacity() const { return cap_; }
    int buffers(const void** ptr, size_t* bytes, int max) const;
    // ---- packed tile files (packed_io.cc): a tile adopted from / dumped to the binary form of include/bvc_host.h
    friend class PackedIo;
    // ---- sample-sharded windows (shard_rerun.cc): this packer sees only ITS range of a site's batch files, but the reference's
    // reused `AlleleInfo ai` (Q1) runs through ALL of them in file order.  With tracking on, the packer records per site whether
    // its range held a numeric token and the stale fields it ended the site with, and which indel entries came BEFORE the first
    // numeric token of the range: those carry the fields of the PRECEDING range / site and are patched once every rank has packed.
    struct SiteTrace { uint8_t has_numeric, base, mapq, qual, rpr, strand; };
    void track_ranges(bool on) { track_ = on; }
    const std::vector<SiteTrace>& traces() const { return trace_; }
    const std::vector<std::pair<uint32_t, uint32_t>>& early_indels() const { return early_; }     // (site of the tile, entry)
    void patch_entry(uint32_t entry, const SiteTrace& st);
    void set_state(const SiteTrace& st) { st_base_ = st.base; st_mapq_ = st.mapq; st_qual_ = st.qual; st_rpr_ = st.rpr; st_strand_ = st.strand; }
    // one raw entry into the current site (merged per-site views of the sharded formatter); indel_text != NULL for an indel entry
    void append_raw(uint8_t code, uint8_t qual, uint8_t mapq, uint8_t rpr, uint32_t sample, const char* indel_text);
    // code / qual / sample of the entries of one site of a lazily read tile (no-op otherwise; idempotent)
    void expand_site(int32_t site);
    bool lazy() const { return lazy_; }
    const uint8_t* codes() const { return code_.data(); }
    const uint8_t* quals() const { return qual_.data(); }
    const uint8_t* mapqs() const { return mapq_.data(); }
    const uint8_t* rprs() const { return rpr_.data(); }
    const uint32_t* samples() const { return smp_.data(); }
    const uint32_t* offsets() const { return entry_off_.data(); }

private:
    int32_t n_samples_;
    // the reference's reused `AlleleInfo ai` (Q1): survives tokens, lines, sites and clear()
    uint8_t st_base_ = 0, st_mapq_ = 0, st_qual_ = 0, st_rpr_ = 0, st_strand_ = 0;
    uint32_t sample_ = 0;          // j of src/BaseVarC.cpp:404: advances on every token of the current site
    std::vector<int32_t> pos_;
    std::vector<uint8_t> ref_base_;
    std::vector<uint32_t> entry_off_;
    // per-entry arrays: raw buffers of capacity cap_ with n_ entries in use (add_line writes through pointers)
    std::vector<uint8_t> code_, qual_, mapq_, rpr_;
    std::vector<uint32_t> smp_;
    size_t n_ = 0, cap_ = 0;
    // compact wire form of (code, qual) built alongside (include/bvc_gpu.h BVC_PACK8): valid while the tile has <= 30 quals
    std::vector<uint8_t> pk8_;
    // A tile read LAZILY from a packed tile file (PackedIo::next_tile with lazy = true: the dispatcher): code_ carries only the
    // indel entries (all the CVG line looks at), qual_ / smp_ nothing; expand_site fills the three arrays for one site -- the VCF
    // columns of an LRT-positive site are the only reader of them (src/BaseType.cpp:152-163).
    std::vector<uint8_t> lazy_bitmap_;     // [n_sites][ceil(N/8)] covered samples
    bool lazy_ = false;
    uint8_t dict_[32] = {0};
    int8_t qidx_[128];
    int ndict_ = 0;
    bool pk8_ok_ = true;
    inline uint8_t pack_counted(uint8_t base, uint8_t strand, uint8_t qual)
    {
        int qi = qidx_[qual];
        if (qi < 0) {
            if (ndict_ == BVC_PACK8_MAX_QUALS) { pk8_ok_ = false; return 0; }
            qi = ndict_; qidx_[qual] = (int8_t)qi; dict_[ndict_++] = qual;
        }
        return BVC_PACK8((unsigned)qi, strand, base);
    }
    void reserve_entries(size_t need);
    int add_token_slow(const char* tok, const char* tend);
    std::unordered_map<int64_t, std::string> indel_;
    std::string err_;
    bool track_ = false, site_numeric_ = false;
    std::vector<SiteTrace> trace_;
    std::vector<std::pair<uint32_t, uint32_t>> early_;
};

}  // namespace bvc_host
