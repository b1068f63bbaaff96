// packed_io.hpp -- the binary packed batch format (layout: include/bvc_host.h "packed tile files").
//
// The reference's `--load` phase writes one text token per (site, sample) into BGZF batch files
// (src/BaseVarC.cpp:509-527) and `--rerun` inflates and tokenises them again on every run (:403-441).  A packed
// file holds the SAME information as ready-made tiles -- the structure-of-arrays form the GPU library takes
// (include/bvc_gpu.h bvc_tile, packed8 wire form) -- so a rerun reads its arrays straight into page-locked
// buffers: no inflate, no tokeniser.  One file per reference thread directory: PREFIX.tmp.thread.{t}/packed.bvc.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "packer.hpp"

namespace bvc_host {

constexpr char kPackedMagic[8] = {'B', 'V', 'C', 'P', 'A', 'C', 'K', '1'};
constexpr uint32_t kPackedVersion = 1;
constexpr uint32_t kTileMagic = 0x454c4954u;      // "TILE"

struct PackedFileHeader {          // 64 bytes
    char     magic[8];
    uint32_t version;
    uint32_t n_samples;
    uint32_t n_tiles;              // patched at close
    uint32_t max_sites;            // largest tile of the file (sizes the reader's buffers), patched at close
    uint64_t max_entries;
    uint32_t names_bytes;          // the sample-name line(s) of the window's batch files, concatenated (src/BaseVarC.cpp:333-338)
    uint32_t reserved[7];
};
static_assert(sizeof(PackedFileHeader) == 64, "file header is 64 bytes");

struct PackedTileHeader {          // 64 bytes
    uint32_t magic;
    uint32_t n_sites;
    uint32_t n_entries;
    uint32_t n_dict;               // > 0: the (code, qual) pair is stored as packed8 + this dictionary
    uint32_t n_indel;              // indel entries: their index, stale qual and token text follow the arrays
    uint32_t indel_bytes;
    uint32_t reserved[2];
    uint8_t  dict[32];
};
static_assert(sizeof(PackedTileHeader) == 64, "tile header is 64 bytes");

class PackedIo {
public:
    ~PackedIo() { if (f_) fclose(f_); }
    // ---- writer -------------------------------------------------------------------------------------------------
    bool create(const std::string& path, int32_t n_samples, const std::string& names, std::string* err);
    bool add_tile(const Packer& pk, std::string* err);           // the packer's current sites (n_sites > 0)
    bool finish(std::string* err);                               // patches the header, closes the file
    // ---- reader -------------------------------------------------------------------------------------------------
    bool open(const std::string& path, std::string* err);
    const PackedFileHeader& header() const { return hdr_; }
    const std::string& names() const { return names_; }
    // Replaces the packer's sites with the next tile of the file.  Returns 1, 0 at the end of the file, -1 on a corrupt file.
    // lazy (the dispatcher; one-byte tiles only): the per-entry (code, qual, sample) arrays are left to Packer::expand_site
    int next_tile(Packer& pk, std::string* err, bool lazy = false);
    static bool is_complete(const std::string& path, int32_t n_samples);
    // sites [s0, s1) of a tile as the packer's content (fixtures / benchmarks: bvc_write_packed_files); tile->sample is required,
    // indel entries get the token text `itok`, pos == NULL numbers the sites 1, 2, ...
    static int adopt(Packer& pk, const bvc_tile* t, size_t s0, size_t s1, const int32_t* pos, const std::string& itok);

private:
    bool put(const void* p, size_t n);                           // + padding to 16 bytes
    bool get(void* p, size_t n);
    FILE* f_ = nullptr;
    PackedFileHeader hdr_{};
    std::string names_;
    uint32_t tiles_read_ = 0;
    std::vector<uint8_t> bitmap_;
    std::vector<uint32_t> ie_, io_;
    std::vector<uint8_t> iq_;
    std::string itext_;
};

}  // namespace bvc_host
