// bgzf_io.hpp -- minimal BGZF (blocked gzip, "BC" extra sub-field) line reader and writer on zlib.
//
// The reference reads its batch files and writes its .vcf.gz/.cvg.gz through htslib's bgzf_*
// (paths under /root/reference: src/BaseVarC.cpp:223-231 validity probe, :327-338,:407 bgzf_getline,
// :268-295 sub-file merge).  This is a from-scratch restatement of the container format only:
// every block is an independent gzip member of <= 64 KiB with its total size in the BC field; a
// well-formed file ends with the 28-byte empty block.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <condition_variable>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

namespace bvc_host {

// true iff `path` starts with a BGZF block header and ends with the EOF marker block
// (bgzf_compression(fp) == 2 && bgzf_check_EOF(fp) == 1, src/BaseVarC.cpp:224-227).
bool bgzf_is_complete(const std::string& path);

// Optional process-wide pool of threads for BGZF blocks (blocks are independent gzip members).  Readers
// opened while it runs keep their next blocks queued in it, and memory-sink writers hand it their full blocks to deflate;
// without it (the default) both work on their caller's thread.
class BgzfReader;
class InflatePool {
public:
    static InflatePool& instance();
    void start(int n_threads);          // once, before the readers are opened; n <= 0: no pool
    void stop();                        // after the readers are closed
    int threads() const;
    void submit(BgzfReader* r, int slot);
    void submit(std::function<void()> task);
private:
    InflatePool() = default;
    ~InflatePool();
    class Impl;
    Impl* impl_ = nullptr;
};

class BgzfReader {
public:
    BgzfReader();
    ~BgzfReader();
    BgzfReader(const BgzfReader&) = delete;
    BgzfReader& operator=(const BgzfReader&) = delete;
    bool open(const std::string& path, std::string* err);
    void close();
    // One line without its '\n' (bgzf_getline semantics); false at end of file.  On a corrupt block
    // returns false and sets failed().
    bool getline(std::string& line);
    // Zero-copy variant: the line as a [ptr, ptr+len) view, valid until the next call.
    bool getline_view(const char** ptr, size_t* len);
    bool failed() const { return failed_; }
    const std::string& error() const { return err_; }
    // blocks the whole-buffer decoder declined and zlib inflated instead (0 on well-formed files; BVC_ZLIB_INFLATE=1: zlib only)
    size_t zlib_blocks() const { return zlib_blocks_; }

private:
    friend class InflatePool;
    struct Block;
    static constexpr int kSlots = 3;    // with a pool: the block being consumed + two read ahead
    bool fill();                        // the next non-empty block, inflated, becomes the current one
    bool read_block(Block& b);
    static bool decode_block(Block& b, bool use_fast);
    void decode_slot(int slot);         // called by a pool thread
    FILE* f_ = nullptr;
    std::unique_ptr<Block[]> blk_;
    int head_ = 0, queued_ = 0;         // ring of slots handed to the pool, oldest first
    const char* cur_ = nullptr;         // inflated bytes of the current block
    size_t pos_ = 0, len_ = 0;
    std::string carry_;                 // a line that spans blocks
    bool failed_ = false, eof_ = false, use_fast_ = true, async_ = false;
    size_t zlib_blocks_ = 0;
    std::string err_;
    std::mutex m_;                      // slot states (pool threads <-> the reader's thread)
    std::condition_variable cv_;
};

// Writes BGZF either to a file or to an in-memory byte string (take()).  close() appends the EOF block
// unless `eof_marker` is false (used for window sub-streams that are concatenated later: BGZF blocks are
// independent gzip members, so concatenating compressed windows equals compressing the concatenation).
class BgzfWriter {
public:
    explicit BgzfWriter(int level = 6);
    ~BgzfWriter();
    BgzfWriter(const BgzfWriter&) = delete;
    BgzfWriter& operator=(const BgzfWriter&) = delete;
    bool open(const std::string& path);        // file sink
    void open_memory();                        // memory sink; with an InflatePool running, full blocks are deflated by the pool
    bool write(const char* data, size_t n);
    bool write(const std::string& s) { return write(s.data(), s.size()); }
    bool close(bool eof_marker = true);
    std::string& memory() { return mem_; }
    static const unsigned char kEofBlock[28];

private:
    struct Pending;
    static constexpr size_t kMaxFly = 8;       // blocks in the pool per writer before write() waits for the oldest
    bool flush_block(const char* data, size_t n);
    bool collect(bool all);                    // finished blocks, oldest first, into mem_
    bool sink(const void* p, size_t n);
    FILE* f_ = nullptr;
    bool mem_mode_ = false, open_ = false;
    std::string mem_;
    std::vector<char> pend_;
    std::deque<std::unique_ptr<Pending>> fly_;
    std::mutex m_;
    std::condition_variable cv_;
    int level_;
};

}  // namespace bvc_host
