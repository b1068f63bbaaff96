// format.cc -- result record + the site's entries -> the reference's CVG / VCF text.
//
// Restates the string building of bt_f (src/BaseVarC.cpp:568-575,609-610,640,662) and of WriteVcf
// (src/BaseType.cpp:143-163,198-233) under /root/reference.  No statistic is computed here: every number
// comes from the bvc_site_result; the only arithmetic is what the reference itself does while printing
// (1 - exp(MLN10TO10 * qual) per sample column, depth / depth_total, var_qual / sum of ALT depths).
// fmt's {:.Nf} is printf's %.Nf ("nan"/"inf" with the sign bit's '-') except for tiny values: see put_fixed.
#include "format.hpp"

#include <math.h>
#include <stdint.h>
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

// ---- the reference's "{:.Nf}" ----------------------------------------------------------------------------------
// The reference prints through the fmt 5.3.1 it vendors (fmt/include/fmt/format-inl.h).  For ordinary values its fixed
// format equals printf's %.Nf.  For a TINY non-zero value -- all N requested decimals are leading zeros -- that version
// stops in fixed_handler::on_start (format-inl.h:591-605) with ONE digit, '1' or '0' by the rounding direction of the
// Grisu-scaled value against its own leading power of ten, and grisu_prettify (format.h:1092-1105) then prints that digit
// BEHIND the N zeros: 9.6e-5 -> "0.0001", 4.9e-5 -> "0.0000" (printf: "0.000").  FS / SOR / rank-sum values of that size
// occur at ~0.4 % of low-pass sites, so the drop-in restates the arithmetic: Grisu's normalised 64-bit product with the
// cached power of ten (computed exactly here and rounded half-up like fmt's table), its digit count, and the same three
// comparisons.  Anything else falls through to printf, as fmt does.
struct CachedPow10 { uint64_t f; int e; };

CachedPow10 pow10_normalized(int K)          // 10^K = f * 2^e with the top bit of f set; K >= 0
{
    uint32_t w[48] = {1};
    int n = 1;
    for (int i = 0; i < K; ++i) {
        uint64_t carry = 0;
        for (int j = 0; j < n; ++j) { const uint64_t x = (uint64_t)w[j] * 10u + carry; w[j] = (uint32_t)x; carry = x >> 32; }
        if (carry) w[n++] = (uint32_t)carry;
    }
    const int bl = 32 * (n - 1) + (32 - __builtin_clz(w[n - 1]));
    auto bit = [&](int i) -> uint64_t { return (i < 0 || i >= 32 * n) ? 0u : (w[i >> 5] >> (i & 31)) & 1u; };
    CachedPow10 c; c.f = 0; c.e = bl - 64;
    for (int i = 0; i < 64; ++i) c.f |= bit(bl - 64 + i) << i;
    if (bit(bl - 65)) { if (++c.f == 0) { c.f = 1ull << 63; ++c.e; } }
    return c;
}

// true: `v` (> 0, finite) takes fmt's one-digit path; *digit is the character printed behind *zeros zeros
bool fmt_tiny_fixed(double v, int precision, char* digit, int* zeros)
{
    uint64_t u; memcpy(&u, &v, 8);
    uint64_t f = u & ((1ull << 52) - 1);
    int biased = (int)((u >> 52) & 0x7ff);
    if (biased) f += 1ull << 52; else biased = 1;
    int e = biased - 1075;
    while (!(f & (1ull << 52))) { f <<= 1; --e; }               // fp::normalize
    f <<= 11; e -= 11;
    const int min_exponent = -60 - (e + 64);                    // get_cached_power (format-inl.h:452-463)
    int index = (int)ceil((min_exponent + 63) * 0.30102999566398114);
    index = (index + 348 - 1) / 8 + 1;
    const int K = -348 + index * 8;
    if (K <= 0 || K > 340) return false;                        // not a tiny value
    static CachedPow10 cache[44]; static bool have[44];
    const int ci = K / 8;                                       // K = 4, 12, ...
    CachedPow10 c;
    if (have[ci]) c = cache[ci]; else { c = pow10_normalized(K); cache[ci] = c; have[ci] = true; }   // idempotent: benign if raced
    const unsigned __int128 prod = (unsigned __int128)f * c.f;  // fp operator* (round half up on bit 63)
    uint64_t vf = (uint64_t)(prod >> 64);
    if ((uint64_t)prod & (1ull << 63)) ++vf;
    const int ve = e + c.e + 64;
    if (ve < -60 || ve > -32) return false;
    const uint32_t integral = (uint32_t)(vf >> -ve);
    if (integral == 0) return false;
    int kappa = 1;
    for (uint32_t x = integral; x >= 10; x /= 10) ++kappa;
    if (precision + kappa - K > 0) return false;                // digits are needed: the ordinary path (= printf)
    static const uint64_t p10[10] = {1ull, 10ull, 100ull, 1000ull, 10000ull, 100000ull, 1000000ull, 10000000ull, 100000000ull, 1000000000ull};
    const uint64_t divisor = p10[kappa - 1] << -ve, remainder = vf / 10, error = 10;
    bool up;                                                    // get_round_direction (format-inl.h:469-483)
    if (remainder <= divisor - remainder && error * 2 <= divisor - remainder * 2) up = false;
    else if (remainder >= error && remainder - error >= divisor - (remainder - error)) up = true;
    else return false;                                          // undecidable: fmt falls back to sprintf
    const int full_exp = 1 + kappa - K;
    if (full_exp > 0) return false;
    *digit = up ? '1' : '0';
    *zeros = -full_exp < precision ? -full_exp : precision;
    return true;
}

inline void put_fixed(std::string& s, double v, int digits)
{
    char b[64];
    if (v != 0 && v == v && fabs(v) < 1.0 && fabs(v) <= 1.7e308) {
        char d; int z;
        if (fmt_tiny_fixed(fabs(v), digits, &d, &z)) {
            if (signbit(v)) s += '-';
            if (digits == 0 && d == '0') { s += '0'; return; }     // precision 0: no trailing zeros (format.h:2787-2788)
            s += "0.";
            s.append((size_t)z, '0');
            s += d;
            return;
        }
    }
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

void format_fixed(std::string& out, double v, int digits) { put_fixed(out, v, digits); }

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
