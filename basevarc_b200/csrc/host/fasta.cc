// fasta.cc -- reference bases of a region from an indexed FASTA (plain or gzip/bgzip), upper-cased.
//
// Stands in for RefReader::GetTargetBase -> SeqLib::RefGenome::QueryRegion -> htslib faidx
// (/root/reference/src/RefReader.h:17-33): 0-based inclusive [beg, end], clipped to the contig, lower case
// folded to upper case.  The .fai index is read when present and built in memory otherwise.
#include "fasta.hpp"

#include <stdio.h>
#include <string.h>
#include <zlib.h>
#include <fstream>
#include <sstream>

namespace bvc_host {

namespace {
struct FaiEntry { std::string name; long long len = 0, offset = 0; int line_bases = 0, line_width = 0; };

bool read_fai(const std::string& path, std::vector<FaiEntry>& out)
{
    std::ifstream f(path);
    if (!f.is_open()) return false;
    std::string line;
    while (std::getline(f, line)) {
        std::istringstream ss(line);
        FaiEntry e;
        if (ss >> e.name >> e.len >> e.offset >> e.line_bases >> e.line_width) out.push_back(e);
    }
    return !out.empty();
}

// sequential reader over plain or gzip data (gzread passes plain files through)
struct Stream {
    gzFile g = nullptr;
    bool open(const std::string& p) { g = gzopen(p.c_str(), "rb"); if (g) gzbuffer(g, 1 << 20); return g != nullptr; }
    ~Stream() { if (g) gzclose(g); }
};

bool build_index(const std::string& path, std::vector<FaiEntry>& out)
{
    Stream s;
    if (!s.open(path)) return false;
    std::vector<char> buf(1 << 20);
    long long off = 0;
    FaiEntry cur; bool have = false, in_header = false, first_line = false;
    long long line_start = 0; int line_bases = 0;
    std::string header;
    int n;
    while ((n = gzread(s.g, buf.data(), (unsigned)buf.size())) > 0) {
        for (int i = 0; i < n; ++i, ++off) {
            const char c = buf[i];
            if (in_header) {
                if (c == '\n') {
                    in_header = false;
                    cur = FaiEntry();
                    cur.name = header.substr(0, header.find_first_of(" \t"));
                    cur.offset = off + 1; have = true; first_line = true; line_start = off + 1; line_bases = 0;
                } else header += c;
            } else if (c == '>') {
                if (have) out.push_back(cur);
                have = false; in_header = true; header.clear();
            } else if (have) {
                if (c == '\n') {
                    if (first_line && line_bases > 0) { cur.line_bases = line_bases; cur.line_width = (int)(off + 1 - line_start); first_line = false; }
                    line_start = off + 1; line_bases = 0;
                } else if (c != '\r') { ++cur.len; ++line_bases; }
            }
        }
    }
    if (have) {
        if (first_line) { cur.line_bases = line_bases > 0 ? line_bases : 1; cur.line_width = cur.line_bases + 1; }
        out.push_back(cur);
    }
    return !out.empty();
}
}  // namespace

bool Fasta::open(const std::string& path, std::string* err)
{
    path_ = path;
    contigs_.clear();
    std::vector<FaiEntry> idx;
    if (!read_fai(path + ".fai", idx) && !build_index(path, idx)) {
        if (err) *err = "cannot index " + path + " (reference must be indexed with samtools faidx)";
        return false;
    }
    for (auto& e : idx) contigs_.push_back({e.name, e.len, e.offset, e.line_bases, e.line_width});
    return true;
}

bool Fasta::fetch(const std::string& chr, long long beg, long long end, std::string& seq, std::string* err) const
{
    seq.clear();
    const Contig* c = nullptr;
    for (auto& k : contigs_) if (k.name == chr) { c = &k; break; }
    if (!c) { if (err) *err = "contig " + chr + " not in " + path_; return false; }
    if (beg < 0) beg = 0;
    if (end >= c->len) end = c->len - 1;
    if (end < beg) return true;
    const long long first = c->offset + beg / c->line_bases * c->line_width + beg % c->line_bases;
    const long long last = c->offset + end / c->line_bases * c->line_width + end % c->line_bases;
    Stream s;
    if (!s.open(path_)) { if (err) *err = "cannot open " + path_; return false; }
    if (gzseek(s.g, (z_off_t)first, SEEK_SET) < 0) { if (err) *err = "seek failed in " + path_; return false; }
    std::vector<char> buf((size_t)(last - first + 1));
    long long got = 0;
    while (got < (long long)buf.size()) {
        const int n = gzread(s.g, buf.data() + got, (unsigned)std::min<long long>((long long)buf.size() - got, 1 << 30));
        if (n <= 0) break;
        got += n;
    }
    seq.reserve((size_t)(end - beg + 1));
    for (long long i = 0; i < got; ++i) {
        char ch = buf[(size_t)i];
        if (ch == '\n' || ch == '\r') continue;
        if (!((ch >= 'A' && ch <= 'Z') || ch == '-')) ch = (char)(ch ^ 0x20);    // RefReader.h:28-31
        seq += ch;
    }
    return true;
}

}  // namespace bvc_host
