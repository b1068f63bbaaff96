// fasta.hpp -- indexed FASTA region fetch (see fasta.cc)
#pragma once
#include <string>
#include <vector>

namespace bvc_host {

class Fasta {
public:
    struct Contig { std::string name; long long len, offset; int line_bases, line_width; };
    bool open(const std::string& path, std::string* err);
    // 0-based inclusive [beg, end], clipped to the contig; upper-cased
    bool fetch(const std::string& chr, long long beg, long long end, std::string& seq, std::string* err) const;
    const std::vector<Contig>& contigs() const { return contigs_; }
    // true when PATH.fai exists on disk (the reference lists ##contig lines only then, src/BaseVarC.cpp:372-379)
    const std::string& path() const { return path_; }

private:
    std::string path_;
    std::vector<Contig> contigs_;
};

}  // namespace bvc_host
