// bgzf_io.cc -- see bgzf_io.hpp
#include "bgzf_io.hpp"
#include "fast_inflate.hpp"

#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#include <utility>

namespace bvc_host {

namespace {
constexpr size_t kBlockData = 0xff00;      // uncompressed bytes per block (htslib's BGZF_BLOCK_SIZE)
constexpr size_t kMaxBlock = 0x10000;

inline uint16_t rd16(const unsigned char* p) { return (uint16_t)(p[0] | (p[1] << 8)); }
inline uint32_t rd32(const unsigned char* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
inline void wr16(unsigned char* p, uint16_t v) { p[0] = (unsigned char)(v & 0xff); p[1] = (unsigned char)(v >> 8); }
inline void wr32(unsigned char* p, uint32_t v) { wr16(p, (uint16_t)(v & 0xffff)); wr16(p + 2, (uint16_t)(v >> 16)); }
}  // namespace

const unsigned char BgzfWriter::kEofBlock[28] = {0x1f, 0x8b, 0x08, 0x04, 0, 0, 0, 0, 0, 0xff, 0x06, 0x00, 0x42, 0x43,
                                                 0x02, 0x00, 0x1b, 0x00, 0x03, 0x00, 0, 0, 0, 0, 0, 0, 0, 0};

bool bgzf_is_complete(const std::string& path)
{
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    unsigned char h[18], t[28];
    bool ok = fread(h, 1, 18, f) == 18 && h[0] == 0x1f && h[1] == 0x8b && h[2] == 8 && (h[3] & 4) && h[12] == 'B' && h[13] == 'C';
    if (ok) ok = fseek(f, -28, SEEK_END) == 0 && fread(t, 1, 28, f) == 28 && memcmp(t, BgzfWriter::kEofBlock, 28) == 0;
    fclose(f);
    return ok;
}

// ---- reader ------------------------------------------------------------------------------------------
// A block goes through two steps: read_block (the file's next non-empty block into a slot: sequential, cheap) and decode_block
// (inflate + CRC: where the time goes).  Without an inflate pool both run on the caller's thread, one slot.  With a pool
// (InflatePool::start; the dispatcher starts one when it has fewer window threads than cores) the reader keeps kSlots blocks
// read ahead and queued for the pool, so the window thread finds its next block already inflated.
namespace {
struct ZStream {                     // zlib's verdict on blocks the whole-buffer decoder declines: one stream per thread
    z_stream z; bool init = false;
    ~ZStream() { if (init) inflateEnd(&z); }
};
thread_local ZStream t_zs;
}  // namespace

struct BgzfReader::Block {
    std::vector<unsigned char> comp;
    std::vector<char> out;
    size_t head = 0, bsize = 0, isize = 0;
    int state = 0;                   // 0 free, 1 queued / being decoded, 2 decoded, 3 failed
    bool used_zlib = false;
    std::string err;
};

class InflatePool::Impl {
public:
    std::mutex m;
    std::condition_variable cv;
    std::deque<std::function<void()>> q;
    std::vector<std::thread> th;
    bool stop = false;
    void run()
    {
        for (;;) {
            std::function<void()> t;
            {
                std::unique_lock<std::mutex> lk(m);
                cv.wait(lk, [&] { return stop || !q.empty(); });
                if (q.empty()) return;
                t = std::move(q.front()); q.pop_front();
            }
            t();
        }
    }
};

InflatePool& InflatePool::instance() { static InflatePool p; return p; }
InflatePool::~InflatePool() { stop(); }
int InflatePool::threads() const { return impl_ ? (int)impl_->th.size() : 0; }
void InflatePool::start(int n)
{
    if (impl_ || n <= 0) return;
    impl_ = new Impl;
    for (int i = 0; i < n; ++i) impl_->th.emplace_back([this] { impl_->run(); });
}
void InflatePool::stop()
{
    if (!impl_) return;
    { std::lock_guard<std::mutex> lk(impl_->m); impl_->stop = true; }
    impl_->cv.notify_all();
    for (auto& t : impl_->th) t.join();
    delete impl_;
    impl_ = nullptr;
}
void InflatePool::submit(std::function<void()> task)
{
    { std::lock_guard<std::mutex> lk(impl_->m); impl_->q.emplace_back(std::move(task)); }
    impl_->cv.notify_one();
}
void InflatePool::submit(BgzfReader* r, int slot) { submit([r, slot] { r->decode_slot(slot); }); }

BgzfReader::BgzfReader() = default;
BgzfReader::~BgzfReader() { close(); }

bool BgzfReader::open(const std::string& path, std::string* err)
{
    close();
    f_ = fopen(path.c_str(), "rb");
    if (!f_) { if (err) *err = "cannot open " + path; return false; }
    setvbuf(f_, nullptr, _IOFBF, 1 << 20);
    async_ = InflatePool::instance().threads() > 0;
    blk_.reset(new Block[kSlots]);
    for (int i = 0; i < (async_ ? kSlots : 1); ++i) { blk_[i].comp.resize(kMaxBlock + kFastInflateInPad); blk_[i].out.resize(kMaxBlock + kFastInflateOutSlack); }
    { const char* e = getenv("BVC_ZLIB_INFLATE"); use_fast_ = !(e && e[0] == '1'); }
    head_ = 0; queued_ = 0; cur_ = nullptr;
    pos_ = len_ = 0; failed_ = eof_ = false; carry_.clear(); err_.clear(); zlib_blocks_ = 0;
    return true;
}

void BgzfReader::close()
{
    if (blk_ && async_) {                                        // nothing of this reader may still be in the pool
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [&] { for (int i = 0; i < kSlots; ++i) if (blk_[i].state == 1) return false; return true; });
    }
    if (f_) fclose(f_);
    f_ = nullptr;
    blk_.reset();
    cur_ = nullptr;
}

// The file's next non-empty block into `b`.  false: end of file, or a malformed container (failed_).
bool BgzfReader::read_block(Block& b)
{
    while (!eof_) {
        unsigned char* h = b.comp.data();
        size_t got = fread(h, 1, 12, f_);
        if (got == 0) { eof_ = true; return false; }
        if (got != 12 || h[0] != 0x1f || h[1] != 0x8b || h[2] != 8 || !(h[3] & 4)) { failed_ = true; err_ = "not a BGZF block"; return false; }
        const size_t xlen = rd16(h + 10);
        if (12 + xlen > kMaxBlock || fread(h + 12, 1, xlen, f_) != xlen) { failed_ = true; err_ = "truncated BGZF header"; return false; }
        size_t bsize = 0;
        for (size_t o = 0; o + 4 <= xlen;) {
            const unsigned char* s = h + 12 + o;
            const size_t sl = rd16(s + 2);
            if (s[0] == 'B' && s[1] == 'C' && sl == 2) bsize = (size_t)rd16(s + 4) + 1;
            o += 4 + sl;
        }
        const size_t head = 12 + xlen;
        if (bsize < head + 8 || bsize > kMaxBlock) { failed_ = true; err_ = "bad BGZF block size"; return false; }
        const size_t rest = bsize - head;
        if (fread(h + head, 1, rest, f_) != rest) { failed_ = true; err_ = "truncated BGZF block"; return false; }
        const size_t isize = rd32(h + bsize - 4);
        if (isize > kMaxBlock) { failed_ = true; err_ = "bad BGZF ISIZE"; return false; }
        if (isize == 0) continue;              // empty block (EOF marker or flush point)
        b.head = head; b.bsize = bsize; b.isize = isize;
        return true;
    }
    return false;
}

// Inflate + CRC of a block that read_block filled (any thread).  false: b.err says why.
bool BgzfReader::decode_block(Block& b, bool use_fast)
{
    unsigned char* h = b.comp.data();
    const size_t rest = b.bsize - b.head, isize = b.isize;
    const uint32_t want_crc = rd32(h + b.bsize - 8);
    uint8_t* out = reinterpret_cast<uint8_t*>(b.out.data());
    // the whole-buffer decoder first (fast_inflate.hpp); whatever it declines -- or gets wrong by the CRC -- goes to zlib
    size_t fast_len = 0;
    b.used_zlib = false;
    if (use_fast && fast_inflate(h + b.head, rest - 8, out, kMaxBlock, &fast_len) && fast_len == isize && fast_crc32(out, isize) == want_crc) return true;
    b.used_zlib = use_fast;
    ZStream& zs = t_zs;
    if (!zs.init) {
        memset(&zs.z, 0, sizeof zs.z);
        if (inflateInit2(&zs.z, -15) != Z_OK) { b.err = "inflateInit2 failed"; return false; }
        zs.init = true;
    } else if (inflateReset(&zs.z) != Z_OK) { b.err = "inflateReset failed"; return false; }
    zs.z.next_in = h + b.head; zs.z.avail_in = (uInt)(rest - 8);
    zs.z.next_out = out; zs.z.avail_out = (uInt)kMaxBlock;
    const int rc = inflate(&zs.z, Z_FINISH);
    if (rc != Z_STREAM_END || zs.z.total_out != isize) { b.err = "inflate failed"; return false; }
    if (fast_crc32(out, isize) != want_crc) { b.err = "BGZF CRC mismatch"; return false; }
    return true;
}

void BgzfReader::decode_slot(int slot)                           // pool thread
{
    Block& b = blk_[slot];
    const bool ok = decode_block(b, use_fast_);
    // notified under the lock: the reader may be destroyed as soon as its thread sees the slot done
    std::lock_guard<std::mutex> lk(m_);
    b.state = ok ? 2 : 3;
    cv_.notify_all();
}

bool BgzfReader::fill()
{
    if (failed_) return false;
    if (!async_) {
        Block& b = blk_[0];
        if (!read_block(b)) return false;
        if (!decode_block(b, use_fast_)) { failed_ = true; err_ = b.err; return false; }
        if (b.used_zlib) ++zlib_blocks_;
        cur_ = b.out.data(); pos_ = 0; len_ = b.isize;
        return true;
    }
    if (cur_) {                                                  // the block just consumed: its slot is free again
        blk_[head_].state = 0;
        head_ = (head_ + 1) % kSlots; --queued_;
        cur_ = nullptr;
    }
    while (queued_ < kSlots && !eof_ && !failed_) {              // keep the pool kSlots blocks ahead of the caller
        const int slot = (head_ + queued_) % kSlots;
        if (!read_block(blk_[slot])) break;
        { std::lock_guard<std::mutex> lk(m_); blk_[slot].state = 1; }
        ++queued_;
        InflatePool::instance().submit(this, slot);
    }
    if (queued_ == 0) return false;                              // end of file (or failed_ by read_block with nothing decoded pending)
    Block& b = blk_[head_];
    { std::unique_lock<std::mutex> lk(m_); cv_.wait(lk, [&] { return b.state != 1; }); }
    if (b.state == 3) { failed_ = true; err_ = b.err; return false; }
    if (b.used_zlib) ++zlib_blocks_;
    cur_ = b.out.data(); pos_ = 0; len_ = b.isize;
    return true;
}

bool BgzfReader::getline_view(const char** ptr, size_t* len)
{
    carry_.clear();
    bool have_carry = false;
    for (;;) {
        if (pos_ >= len_ && !fill()) {
            if (have_carry && !failed_) { *ptr = carry_.data(); *len = carry_.size(); return true; }   // last line without '\n'
            return false;
        }
        const char* b = cur_ + pos_;
        const char* nl = static_cast<const char*>(memchr(b, '\n', len_ - pos_));
        if (nl) {
            const size_t n = (size_t)(nl - b);
            pos_ += n + 1;
            if (!have_carry) { *ptr = b; *len = n; return true; }
            carry_.append(b, n);
            *ptr = carry_.data(); *len = carry_.size();
            return true;
        }
        carry_.append(b, len_ - pos_);
        have_carry = true;
        pos_ = len_;
    }
}

bool BgzfReader::getline(std::string& line)
{
    const char* p; size_t n;
    if (!getline_view(&p, &n)) return false;
    line.assign(p, n);
    return true;
}

// ---- writer ------------------------------------------------------------------------------------------
bool BgzfWriter::open(const std::string& path)
{
    close();
    f_ = fopen(path.c_str(), "wb");
    if (!f_) return false;
    setvbuf(f_, nullptr, _IOFBF, 1 << 20);
    mem_mode_ = false; open_ = true; pend_.clear();
    return true;
}

void BgzfWriter::open_memory()
{
    close();
    mem_mode_ = true; open_ = true; mem_.clear(); pend_.clear();
}

bool BgzfWriter::sink(const void* p, size_t n)
{
    if (mem_mode_) { mem_.append(static_cast<const char*>(p), n); return true; }
    return fwrite(p, 1, n, f_) == n;
}

// One BGZF block from data[0, n), n <= kBlockData, appended to `out` (any thread).
static bool compress_block(const char* data, size_t n, int level, std::string& dst)
{
    unsigned char out[kMaxBlock + 64];
    static const unsigned char head[16] = {0x1f, 0x8b, 0x08, 0x04, 0, 0, 0, 0, 0, 0xff, 0x06, 0x00, 'B', 'C', 0x02, 0x00};
    memcpy(out, head, 16);
    z_stream zs;
    memset(&zs, 0, sizeof zs);
    for (;;) {
        if (deflateInit2(&zs, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) return false;
        zs.next_in = reinterpret_cast<Bytef*>(const_cast<char*>(data)); zs.avail_in = (uInt)n;
        zs.next_out = out + 18; zs.avail_out = (uInt)(kMaxBlock - 18 - 8);
        const int rc = deflate(&zs, Z_FINISH);
        deflateEnd(&zs);
        if (rc == Z_STREAM_END) break;
        if (level == 0) return false;          // cannot happen for n <= 0xff00 with stored blocks
        level = 0;                             // incompressible data: fall back to stored
    }
    const size_t clen = zs.total_out, bsize = 18 + clen + 8;
    wr16(out + 16, (uint16_t)(bsize - 1));
    wr32(out + 18 + clen, fast_crc32(reinterpret_cast<const uint8_t*>(data), n));
    wr32(out + 18 + clen + 4, (uint32_t)n);
    dst.append(reinterpret_cast<const char*>(out), bsize);
    return true;
}

// A block handed to the pool (memory sink only): compressed by a pool thread, appended to mem_ in order by the writer's thread.
struct BgzfWriter::Pending {
    std::vector<char> in;
    std::string out;
    bool done = false, ok = false;
};

BgzfWriter::BgzfWriter(int level) : level_(level) {}
BgzfWriter::~BgzfWriter() { close(); }

bool BgzfWriter::collect(bool all)
{
    bool ok = true;
    while (!fly_.empty()) {
        Pending& p = *fly_.front();
        {
            std::unique_lock<std::mutex> lk(m_);
            if (!p.done) {
                if (!all && fly_.size() < kMaxFly) break;
                cv_.wait(lk, [&] { return p.done; });
            }
        }
        ok = p.ok && ok;
        mem_.append(p.out);
        fly_.pop_front();
    }
    return ok;
}

bool BgzfWriter::flush_block(const char* data, size_t n)
{
    if (mem_mode_ && InflatePool::instance().threads() > 0) {
        fly_.emplace_back(new Pending);
        Pending* p = fly_.back().get();
        p->in.assign(data, data + n);
        const int level = level_;
        InflatePool::instance().submit([this, p, level] {
            const bool ok = compress_block(p->in.data(), p->in.size(), level, p->out);
            std::lock_guard<std::mutex> lk(m_);               // notified under the lock: the writer may go away once it sees `done`
            p->ok = ok; p->done = true;
            cv_.notify_all();
        });
        return collect(false);
    }
    if (mem_mode_) return compress_block(data, n, level_, mem_);
    std::string tmp;
    return compress_block(data, n, level_, tmp) && sink(tmp.data(), tmp.size());
}

bool BgzfWriter::write(const char* data, size_t n)
{
    if (!open_) return false;
    while (n) {
        const size_t room = kBlockData - pend_.size();
        const size_t take = n < room ? n : room;
        pend_.insert(pend_.end(), data, data + take);
        data += take; n -= take;
        if (pend_.size() == kBlockData) {
            if (!flush_block(pend_.data(), pend_.size())) return false;
            pend_.clear();
        }
    }
    return true;
}

bool BgzfWriter::close(bool eof_marker)
{
    if (!open_) return true;
    bool ok = true;
    if (!pend_.empty()) { ok = flush_block(pend_.data(), pend_.size()); pend_.clear(); }
    ok = collect(true) && ok;
    if (eof_marker) ok = sink(kEofBlock, 28) && ok;
    if (f_) { ok = (fclose(f_) == 0) && ok; f_ = nullptr; }
    open_ = false;
    return ok;
}

}  // namespace bvc_host
