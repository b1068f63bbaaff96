// format.cc -- result record + the site's entries -> the reference's CVG / VCF text.
//
// Restates the string building of bt_f (src/BaseVarC.cpp:568-575,609-610,640,662) and of WriteVcf
// (src/BaseType.cpp:143-163,198-233) under /root/reference.  No statistic is computed here: every number
// comes from the bvc_site_result; the only arithmetic is what the reference itself does while printing
// (1 - exp(MLN10TO10 * qual) per sample column, depth / depth_total, var_qual / sum of ALT depths).
// fmt's {:.Nf} is printf's %.Nf (correctly rounded, "nan"/"inf" with the sign bit's '-').
#include "format.hpp"

#include <math.h>
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <utility>
#include <vector>

namespace bvc_host {

namespace {
const char kBase2Char[8] = {'A', 'C', 'G', 'T', 'N', 'N', 'N', 'N'};   // the reference indexes a 4-char table: base > 3 is UB there
const char kStrand[2] = {'-', '+'};
const double kMLN10TO10 = -0.23025850929940458;                        // src/BaseType.h:10

inline void put_fixed(std::string& s, double v, int digits)
{
    char b[64];
    const int n = snprintf(b, sizeof b, "%.*f", digits, v);
    s.append(b, (size_t)n);
}
inline void put_int(std::string& s, long long v)
{
    char b[32];
    const int n = snprintf(b, sizeof b, "%lld", v);
    s.append(b, (size_t)n);
}
}  // namespace

void format_cvg(std::string& out, const char* chr, int32_t pos, const bvc_tile* t, int32_t site, const bvc_site_result* r,
                const Packer* indel_src, int32_t n_groups, const int32_t* group_depth)
{
    out.clear();
    const uint32_t e0 = t->entry_off[site], e1 = t->entry_off[site + 1];
    // Indels column: "token|count,..." (src/BaseVarC.cpp:568-575).  The reference iterates a robin_hood map,
    // so its order is unspecified; this prints first-occurrence order (compare as a multiset).
    std::vector<std::pair<std::string, int>> ind;
    if (indel_src) {
        for (uint32_t e = e0; e < e1; ++e) {
            if (!BVC_CODE_INDEL(t->code[e])) continue;
            const char* tok = indel_src->indel((int64_t)e);
            if (!tok) continue;
            bool found = false;
            for (auto& p : ind) if (p.first == tok) { p.second += 1; found = true; break; }
            if (!found) ind.emplace_back(tok, 1);
        }
    }
    const long long dep = (long long)r->depth[0] + r->depth[1] + r->depth[2] + r->depth[3];
    out += chr; out += '\t'; put_int(out, pos); out += '\t';
    out += kBase2Char[t->ref_base[site] & 7]; out += '\t';
    put_int(out, dep);
    for (int b = 0; b < 4; ++b) { out += '\t'; put_int(out, r->depth[b]); }
    out += '\t';
    if (ind.empty()) out += '.';
    else for (size_t i = 0; i < ind.size(); ++i) { if (i) out += ','; out += ind[i].first; out += '|'; put_int(out, ind[i].second); }
    out += '\t'; put_fixed(out, r->cvg_fs, 3);
    out += '\t'; put_fixed(out, r->cvg_sor, 3);
    out += '\t';
    for (int k = 0; k < 4; ++k) { if (k) out += ','; put_int(out, r->cvg_strand[k]); }
    for (int g = 0; g < n_groups; ++g) {                                 // src/BaseVarC.cpp:640
        out += '\t';
        for (int b = 0; b < 4; ++b) { if (b) out += ':'; put_int(out, group_depth[g * 4 + b]); }
    }
    out += '\n';
}

void format_vcf(std::string& out, const char* chr, int32_t pos, const bvc_tile* t, int32_t site, const bvc_site_result* r,
                int32_t n_samples, int32_t n_groups, const char* const* group_names, const int32_t* group_depth,
                const double* group_af)
{
    out.clear();
    if (!r->lrt_ok) return;
    const int n_alt = r->n_alt;
    const int ref = t->ref_base[site] & 7;
    const double depth_total = (double)r->depth[0] + (double)r->depth[1] + (double)r->depth[2] + (double)r->depth[3];

    std::vector<std::pair<std::string, std::string>> info;
    std::string alt, ac, af, caf;
    double ad_sum = 0;
    for (int i = 0; i < n_alt; ++i) {                                    // src/BaseType.cpp:199-206
        const int b = r->alt_bases[i];
        if (i) { alt += ','; ac += ','; af += ','; caf += ','; }
        ad_sum += r->depth[b];
        alt += kBase2Char[b];
        put_int(ac, r->depth[b]);
        put_fixed(af, r->af_lrt[b], 6);
        put_fixed(caf, r->depth[b] / depth_total, 6);
    }
    auto kv = [&](const char* k, std::string v) { info.emplace_back(k, std::move(v)); };
    auto fx = [&](double v, int d) { std::string s; put_fixed(s, v, d); return s; };
    kv("CM_AC", ac); kv("CM_AF", af); kv("CM_CAF", caf);
    kv("QD", fx(r->var_qual / ad_sum, 3));
    kv("CM_DP", fx(depth_total, 0));
    kv("MQRankSum", fx(r->ranksum[0], 3));
    kv("ReadPosRankSum", fx(r->ranksum[2], 3));
    kv("BaseQRankSum", fx(r->ranksum[1], 3));
    kv("FS", fx(r->vcf_fs, 3));
    kv("SOR", fx(r->vcf_sor, 3));
    { std::string s; put_int(s, r->vcf_strand[0]); s += ','; put_int(s, r->vcf_strand[1]); kv("SB_REF", s); }
    { std::string s; put_int(s, r->vcf_strand[2]); s += ','; put_int(s, r->vcf_strand[3]); kv("SB_ALT", s); }
    for (int g = 0; g < n_groups; ++g) {                                 // src/BaseVarC.cpp:641-660
        std::string key = std::string(group_names[g]) + "_AF", v;
        const int32_t* gd = group_depth + g * 4;
        if ((long long)gd[0] + gd[1] + gd[2] + gd[3] == 0) v = "0";
        else for (int i = 0; i < n_alt; ++i) {
            if (i) v += ',';
            const double x = group_af[g * 4 + r->alt_bases[i]];
            if (x < 0) v += '0'; else put_fixed(v, x, 6);               // negative = base absent from the group's af_lrt
        }
        info.emplace_back(std::move(key), std::move(v));
    }
    std::sort(info.begin(), info.end());                                 // InfoM is a std::map: keys in byte order

    out += chr; out += '\t'; put_int(out, pos); out += "\t.\t";
    out += kBase2Char[ref]; out += '\t'; out += alt; out += '\t';
    put_fixed(out, r->var_qual, 2); out += '\t';
    out += (r->var_qual > 60.0) ? "." : "LowQual";                       // QUAL_THRESHOLD, src/BaseType.cpp:221
    out += '\t';
    for (size_t i = 0; i < info.size(); ++i) { if (i) out += ';'; out += info[i].first; out += '='; out += info[i].second; }
    out += "\tGT:AB:SO:BP";

    // per-sample columns (src/BaseType.cpp:152-163): entries of a site are in ascending sample order
    char gt_of_base[8][4];
    for (int b = 0; b < 8; ++b) strcpy(gt_of_base[b], "./.");
    for (int i = 0; i < n_alt; ++i) { gt_of_base[r->alt_bases[i] & 7][2] = (char)('1' + i); }
    strcpy(gt_of_base[ref], "0/.");
    uint32_t e = t->entry_off[site];
    const uint32_t e1 = t->entry_off[site + 1];
    char col[48];
    for (int32_t i = 0; i < n_samples; ++i) {
        if (e < e1 && t->sample[e] == (uint32_t)i) {
            const unsigned c = t->code[e];
            const int b = BVC_CODE_BASE(c);
            const int n = snprintf(col, sizeof col, "\t%s:%c:%c:%.6f", gt_of_base[b], kBase2Char[b], kStrand[BVC_CODE_STRAND(c)],
                                   1 - exp(kMLN10TO10 * t->qual[e]));
            out.append(col, (size_t)n);
            ++e;
        } else out += "\t./.";
    }
    out += '\n';
}

}  // namespace bvc_host
