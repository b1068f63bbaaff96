// format.hpp -- CVG / VCF text of one site (see format.cc)
#pragma once
#include <stdint.h>
#include <string>

#include "../../../include/bvc_gpu.h"
#include "packer.hpp"

namespace bvc_host {

void format_cvg(std::string& out, const char* chr, int32_t pos, const bvc_tile* t, int32_t site, const bvc_site_result* r,
                const Packer* indel_src, int32_t n_groups, const int32_t* group_depth);
void format_vcf(std::string& out, const char* chr, int32_t pos, const bvc_tile* t, int32_t site, const bvc_site_result* r,
                int32_t n_samples, int32_t n_groups, const char* const* group_names, const int32_t* group_depth,
                const double* group_af);

// the reference's "{:.Nf}" (fmt 5.3.1, including its one-digit output for tiny values)
void format_fixed(std::string& out, double v, int digits);

}  // namespace bvc_host
