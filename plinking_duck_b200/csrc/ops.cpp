// Operator layer (include/plinkgpu_ops.h): the per-row derivations of the reference's Scan callbacks
// on top of the C ABI.  Host-only C++; every genotype-touching step is a pd_* call.
#include "../../include/plinkgpu_ops.h"

#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstring>
#include <limits>
#include <numeric>
#include <string>
#include <vector>

namespace {

std::string Lower(const char* s) {
  std::string r(s ? s : "");
  for (auto& ch : r) ch = static_cast<char>(std::tolower(static_cast<unsigned char>(ch)));
  return r;
}

struct Par {
  int32_t par1_end = 0, par2_start = 0, par2_end = 0;
  bool active = false;
};

// ResolveParBounds (src/plink_common.cpp:1926-1958)
bool ResolvePar(const char* build, Par* pb) {
  std::string norm;
  for (char c : Lower(build)) {
    if (c != '-' && c != '_' && c != ' ' && c != '.') norm.push_back(c);
  }
  *pb = Par();
  if (norm.empty() || norm == "none") return true;
  if (norm == "grch38" || norm == "hg38" || norm == "b38" || norm == "38") {
    pb->par1_end = 2781479;
    pb->par2_start = 155701383;
    pb->par2_end = 156030895;
    pb->active = true;
    return true;
  }
  if (norm == "grch37" || norm == "hg19" || norm == "b37" || norm == "37") {
    pb->par1_end = 2699520;
    pb->par2_start = 154931044;
    pb->par2_end = 155260560;
    pb->active = true;
    return true;
  }
  return false;
}

// SexAwareCounts (src/include/plink_common.hpp:598-620) from stratum count quadruples instead of a
// per-sample switch (ComputeSexAwareCounts, src/plink_common.cpp:1996-2108): all = every effective
// sample, male / female = the sex strata.  Unknown-sex samples are what is left.
struct Sac {
  uint32_t alt_allele_ct = 0, obs_allele_ct = 0;
  uint32_t hwe_hom_ref = 0, hwe_het = 0, hwe_hom_alt = 0;
  bool hwe_defined = false;
  uint32_t geno_hom_ref = 0, geno_het = 0, geno_hom_alt = 0, geno_missing = 0;
  bool sex_unavailable = false;
};

Sac SexAware(uint8_t ploidy, bool have_sex, uint32_t n, const uint32_t* all, const uint32_t* male,
             const uint32_t* female) {
  Sac r;
  if ((ploidy == PD_PLOIDY_CHR_X || ploidy == PD_PLOIDY_CHR_Y) && !have_sex) {
    r.sex_unavailable = true;
    return r;
  }
  if (ploidy == PD_PLOIDY_CHR_MT) {
    r.obs_allele_ct = all[0] + all[2];
    r.alt_allele_ct = all[2];
    r.geno_hom_ref = all[0];
    r.geno_hom_alt = all[2];
    r.geno_missing = all[1] + all[3];  // missing, or an invalid het on a haploid genome
  } else if (ploidy == PD_PLOIDY_CHR_Y) {
    const uint32_t n_male = male[0] + male[1] + male[2] + male[3];
    r.obs_allele_ct = male[0] + male[2];
    r.alt_allele_ct = male[2];
    r.geno_hom_ref = male[0];
    r.geno_hom_alt = male[2];
    r.geno_missing = (n - n_male) + male[1] + male[3];  // females / unknown sex are excluded
  } else if (ploidy == PD_PLOIDY_CHR_X) {
    const uint32_t n_male = male[0] + male[1] + male[2] + male[3];
    const uint32_t n_female = female[0] + female[1] + female[2] + female[3];
    r.obs_allele_ct = 2 * (female[0] + female[1] + female[2]) + male[0] + male[2];
    r.alt_allele_ct = female[1] + 2 * female[2] + male[2];
    r.hwe_hom_ref = female[0];
    r.hwe_het = female[1];
    r.hwe_hom_alt = female[2];
    r.geno_hom_ref = female[0] + male[0];
    r.geno_het = female[1];
    r.geno_hom_alt = female[2] + male[2];
    r.geno_missing = female[3] + male[1] + male[3] + (n - n_male - n_female);
  } else {
    r.obs_allele_ct = 2 * (all[0] + all[1] + all[2]);
    r.alt_allele_ct = all[1] + 2 * all[2];
    r.hwe_hom_ref = r.geno_hom_ref = all[0];
    r.hwe_het = r.geno_het = all[1];
    r.hwe_hom_alt = r.geno_hom_alt = all[2];
    r.geno_missing = all[3];
  }
  r.hwe_defined = (ploidy == PD_PLOIDY_CHR_X || ploidy == PD_PLOIDY_AUTOSOMAL);
  return r;
}

// Male / female stratum counts over the smallest variant range that covers every sex-aware variant.
struct Strata {
  uint32_t lo = 0, hi = 0;  // [lo, hi) in file variant indices; empty when lo == hi
  std::vector<uint32_t> male, female;
  const uint32_t* Male(uint32_t v) const { return &male[4ull * (v - lo)]; }
  const uint32_t* Female(uint32_t v) const { return &female[4ull * (v - lo)]; }
};

int ComputeStrata(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, const uint8_t* ploidy, const uint8_t* sex, Strata* st) {
  if (!ploidy) return PD_OK;
  uint32_t lo = v_end, hi = v_begin;
  for (uint32_t v = v_begin; v < v_end; ++v) {
    if (ploidy[v] == PD_PLOIDY_CHR_X || ploidy[v] == PD_PLOIDY_CHR_Y) {
      lo = std::min(lo, v);
      hi = std::max(hi, v + 1);
    }
  }
  if (lo >= hi || !sex) return PD_OK;
  pd_info_t info;
  int rc = pd_info(ctx, &info);
  if (rc != PD_OK) return rc;
  const uint32_t N = info.sample_ct;
  const size_t words = (static_cast<size_t>(N) + 63) / 64 + 1;
  std::vector<uint64_t> m(words, 0), f(words, 0);
  for (uint32_t s = 0; s < N; ++s) {
    if (sex[s] == 1) m[s >> 6] |= 1ull << (s & 63);
    else if (sex[s] == 2) f[s >> 6] |= 1ull << (s & 63);
  }
  st->lo = lo;
  st->hi = hi;
  st->male.resize(4ull * (hi - lo));
  st->female.resize(4ull * (hi - lo));
  rc = pd_counts_masked(ctx, lo, hi, m.data(), st->male.data());
  if (rc == PD_OK) rc = pd_counts_masked(ctx, lo, hi, f.data(), st->female.data());
  return rc;
}

uint32_t EffectiveSampleCt(const uint32_t* counts) { return counts[0] + counts[1] + counts[2] + counts[3]; }

}  // namespace

extern "C" {

void pdop_ld_stats(const uint32_t* sums, uint32_t n_pairs, pd_ld_row* out) {
  if (!sums || !out) return;
  for (uint32_t i = 0; i < n_pairs; ++i) {
    const uint32_t* s = sums + 6ull * i;
    const uint32_t n = s[0];
    const double sum_a = s[1], sum_b = s[2], sum_ab = s[3], sum_a2 = s[4], sum_b2 = s[5];
    pd_ld_row r;
    r.obs_ct = n;
    r.is_valid = 0;
    r.r2 = 0;
    r.d_prime = 0;
    if (n >= 2) {
      const double dn = static_cast<double>(n);
      const double mean_a = sum_a / dn;
      const double mean_b = sum_b / dn;
      const double cov_ab = sum_ab / dn - mean_a * mean_b;
      const double var_a = sum_a2 / dn - mean_a * mean_a;
      const double var_b = sum_b2 / dn - mean_b * mean_b;
      if (!(var_a < 1e-15 || var_b < 1e-15)) {
        r.is_valid = 1;
        r.r2 = (cov_ab * cov_ab) / (var_a * var_b);
        const double D = cov_ab / 4.0;
        const double p_a = sum_a / (2.0 * dn);
        const double p_b = sum_b / (2.0 * dn);
        double D_max;
        if (D >= 0) {
          D_max = std::min(p_a * (1.0 - p_b), (1.0 - p_a) * p_b);
        } else {
          D_max = std::max(-p_a * p_b, -(1.0 - p_a) * (1.0 - p_b));
        }
        r.d_prime = std::abs(D_max) < 1e-15 ? 0.0 : D / D_max;
      }
    }
    out[i] = r;
  }
}

void pdop_list_entries(uint64_t base_offset, uint32_t rows, uint32_t n, uint64_t* out) {
  if (!out) return;
  for (uint32_t r = 0; r < rows; ++r) {
    out[2ull * r] = base_offset + static_cast<uint64_t>(r) * n;
    out[2ull * r + 1] = n;
  }
}

int pdop_classify_ploidy(const char* chrom, int32_t pos, const char* build) {
  Par par;
  if (!ResolvePar(build, &par)) return -1;
  std::string c = Lower(chrom);
  if (c.rfind("chr", 0) == 0) c = c.substr(3);
  if (c == "par1" || c == "par2" || c == "xy" || c == "25") return PD_PLOIDY_AUTOSOMAL;
  if (c == "y" || c == "24") return PD_PLOIDY_CHR_Y;
  if (c == "mt" || c == "m" || c == "26" || c == "chrm") return PD_PLOIDY_CHR_MT;
  if (c == "x" || c == "23") {
    if (par.active && ((pos > 0 && pos <= par.par1_end) || (pos >= par.par2_start && pos <= par.par2_end))) {
      return PD_PLOIDY_AUTOSOMAL;
    }
    return PD_PLOIDY_CHR_X;
  }
  return PD_PLOIDY_AUTOSOMAL;
}

int pdop_counts_pass(const uint32_t counts[4], double af_min, double af_max, int64_t ac_min, int64_t ac_max) {
  const uint32_t non_missing = counts[0] + counts[1] + counts[2];
  if (non_missing == 0) return 0;
  const uint32_t ac = counts[1] + 2 * counts[2];
  if (ac_min >= 0 && static_cast<double>(ac) < static_cast<double>(ac_min)) return 0;
  if (ac_max >= 0 && static_cast<double>(ac) > static_cast<double>(ac_max)) return 0;
  if (af_min >= 0.0 || af_max >= 0.0) {
    const double af = static_cast<double>(ac) / (2.0 * static_cast<double>(non_missing));
    if (af_min >= 0.0 && af < af_min) return 0;
    if (af_max >= 0.0 && af > af_max) return 0;
  }
  return 1;
}

void pdop_genotype_range(const uint32_t counts[4], uint32_t allowed_mask, uint32_t include_missing, uint32_t* any_pass,
                         uint32_t* all_pass) {
  uint32_t any = 0, all = 1;
  for (uint32_t g = 0; g < 3; ++g) {
    const bool selected = (allowed_mask >> g) & 1u;
    if (selected && counts[g] > 0) any = 1;
    if (!selected && counts[g] > 0) all = 0;
  }
  if (include_missing && counts[3] > 0) any = 1;
  if (any_pass) *any_pass = any;
  if (all_pass) *all_pass = all;
}

int pdop_predecomp_filter(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, const pd_predecomp_filter* f, uint8_t* skip,
                          uint8_t* all_pass) {
  if (!ctx || !f || v_begin > v_end) return PD_ERR_INVALID_ARG;
  const uint32_t n = v_end - v_begin;
  if (!n) return PD_OK;
  std::vector<uint32_t> counts(4ull * n);
  const int rc = pd_counts(ctx, v_begin, v_end, counts.data());
  if (rc != PD_OK) return rc;
  for (uint32_t i = 0; i < n; ++i) {
    const uint32_t* c = &counts[4ull * i];
    uint8_t sk = 0, ap = 1;
    if (f->use_count_filter && !pdop_counts_pass(c, f->af_min, f->af_max, f->ac_min, f->ac_max)) {
      sk = 1;
    } else if (f->use_genotype_filter) {
      uint32_t any = 0, all = 1;
      pdop_genotype_range(c, f->allowed_mask, f->include_missing, &any, &all);
      if (!any) sk = 1;
      else ap = static_cast<uint8_t>(all);
    }
    if (skip) skip[i] = sk;
    if (all_pass) all_pass[i] = ap;
  }
  return PD_OK;
}

int pdop_genotype_column(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, const pd_predecomp_filter* f, int8_t* out,
                         uint64_t* validity, uint32_t* kept, uint32_t* n_kept) {
  if (!ctx || v_begin > v_end || !kept || !n_kept) return PD_ERR_INVALID_ARG;
  const uint32_t n = v_end - v_begin;
  *n_kept = 0;
  if (!n) return PD_OK;
  const bool filtering = f && (f->use_count_filter || f->use_genotype_filter);
  if (!filtering) {
    for (uint32_t i = 0; i < n; ++i) kept[i] = v_begin + i;
    *n_kept = n;
    return pd_decode_i8_valid(ctx, v_begin, v_end, 7u, out, validity);
  }
  std::vector<uint8_t> skip(n);
  const int rc = pdop_predecomp_filter(ctx, v_begin, v_end, f, skip.data(), nullptr);
  if (rc != PD_OK) return rc;
  uint32_t m = 0;
  for (uint32_t i = 0; i < n; ++i) {
    if (!skip[i]) kept[m++] = v_begin + i;
  }
  *n_kept = m;
  // A variant whose calls all pass needs no null-out, and applying the mask to it changes nothing: one mask
  // serves every kept row (the reference branches per row on geno_range_all_pass only to save work).
  const uint32_t mask = f->use_genotype_filter ? (f->allowed_mask & 7u) : 7u;
  return pd_decode_i8_valid_list(ctx, kept, m, mask, out, validity);
}

int pdop_freq(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, const uint8_t* ploidy, const uint8_t* sex, uint32_t dosage,
              pd_freq_row* out) {
  if (!ctx || v_begin > v_end || (v_end > v_begin && !out)) return PD_ERR_INVALID_ARG;
  const uint32_t n = v_end - v_begin;
  if (!n) return PD_OK;
  std::vector<uint32_t> counts(4ull * n);
  std::vector<uint64_t> ad;
  std::vector<double> r2;
  int rc;
  if (dosage) {
    ad.resize(2ull * n);
    r2.resize(n);
    rc = pd_dosage_counts(ctx, v_begin, v_end, counts.data(), ad.data(), r2.data());
  } else {
    rc = pd_counts(ctx, v_begin, v_end, counts.data());
  }
  if (rc != PD_OK) return rc;
  Strata st;
  rc = ComputeStrata(ctx, v_begin, v_end, ploidy, sex, &st);
  if (rc != PD_OK) return rc;
  static const uint32_t kZero[4] = {0, 0, 0, 0};
  for (uint32_t i = 0; i < n; ++i) {
    const uint32_t v = v_begin + i;
    const uint32_t* gc = &counts[4ull * i];
    const uint8_t pl = ploidy ? ploidy[v] : PD_PLOIDY_AUTOSOMAL;
    pd_freq_row r;
    memset(&r, 0, sizeof(r));
    r.hom_ref_ct = static_cast<int32_t>(gc[0]);
    r.het_ct = static_cast<int32_t>(gc[1]);
    r.hom_alt_ct = static_cast<int32_t>(gc[2]);
    r.missing_ct = static_cast<int32_t>(gc[3]);
    if (pl != PD_PLOIDY_AUTOSOMAL) {
      const bool in_st = v >= st.lo && v < st.hi && !st.male.empty();
      const Sac sac = SexAware(pl, sex != nullptr, EffectiveSampleCt(gc), gc, in_st ? st.Male(v) : kZero,
                               in_st ? st.Female(v) : kZero);
      if (sac.sex_unavailable || sac.obs_allele_ct == 0) {
        r.freq_is_null = 1;
        r.counts_are_null = sac.sex_unavailable;
      } else {
        r.alt_freq = static_cast<double>(sac.alt_allele_ct) / static_cast<double>(sac.obs_allele_ct);
        r.obs_ct = sac.obs_allele_ct;
      }
      r.hom_ref_ct = static_cast<int32_t>(sac.geno_hom_ref);
      r.het_ct = static_cast<int32_t>(sac.geno_het);
      r.hom_alt_ct = static_cast<int32_t>(sac.geno_hom_alt);
      r.missing_ct = static_cast<int32_t>(sac.geno_missing);
    } else if (dosage) {
      const uint64_t total = ad[2ull * i] + ad[2ull * i + 1];
      r.imp_r2 = r2[i];
      if (total == 0) {
        r.freq_is_null = 1;
      } else {
        r.alt_freq = static_cast<double>(ad[2ull * i + 1]) / static_cast<double>(total);
        r.obs_ct = static_cast<uint32_t>(total / 16384);
      }
    } else {
      const uint32_t obs = gc[0] + gc[1] + gc[2];
      if (obs == 0) {
        r.freq_is_null = 1;
      } else {
        r.obs_ct = 2 * obs;
        r.alt_freq = (static_cast<double>(gc[1]) + 2.0 * static_cast<double>(gc[2])) / (2.0 * static_cast<double>(obs));
      }
    }
    out[i] = r;
  }
  return PD_OK;
}

int pdop_hardy(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, const uint8_t* ploidy, const uint8_t* sex, uint32_t midp,
               pd_hardy_row* out) {
  if (!ctx || v_begin > v_end || (v_end > v_begin && !out)) return PD_ERR_INVALID_ARG;
  const uint32_t n = v_end - v_begin;
  if (!n) return PD_OK;
  std::vector<uint32_t> counts(4ull * n);
  std::vector<double> p(n);
  int rc = pd_hardy(ctx, v_begin, v_end, midp, counts.data(), p.data());  // autosomal test for every row
  if (rc != PD_OK) return rc;
  Strata st;
  rc = ComputeStrata(ctx, v_begin, v_end, ploidy, sex, &st);
  if (rc != PD_OK) return rc;
  static const uint32_t kZero[4] = {0, 0, 0, 0};
  std::vector<uint32_t> x5;       // chrX rows that need the X test
  std::vector<uint32_t> x_rows;
  for (uint32_t i = 0; i < n; ++i) {
    const uint32_t v = v_begin + i;
    const uint32_t* gc = &counts[4ull * i];
    const uint8_t pl = ploidy ? ploidy[v] : PD_PLOIDY_AUTOSOMAL;
    pd_hardy_row r;
    memset(&r, 0, sizeof(r));
    r.p_hwe = 1.0;
    if (pl != PD_PLOIDY_AUTOSOMAL) {
      const bool in_st = v >= st.lo && v < st.hi && !st.male.empty();
      const Sac sac = SexAware(pl, sex != nullptr, EffectiveSampleCt(gc), gc, in_st ? st.Male(v) : kZero,
                               in_st ? st.Female(v) : kZero);
      if (sac.sex_unavailable) {
        r.counts_are_null = 1;
        r.stats_are_null = 1;
      } else if (sac.hwe_defined) {
        r.hom_ref_ct = static_cast<int32_t>(sac.hwe_hom_ref);
        r.het_ct = static_cast<int32_t>(sac.hwe_het);
        r.hom_alt_ct = static_cast<int32_t>(sac.hwe_hom_alt);
        const uint32_t fobs = sac.hwe_hom_ref + sac.hwe_het + sac.hwe_hom_alt;
        if (fobs == 0) {
          r.stats_are_null = 1;
        } else {
          r.o_het = static_cast<double>(sac.hwe_het) / static_cast<double>(fobs);
          const double pr = (2.0 * sac.hwe_hom_ref + sac.hwe_het) / (2.0 * fobs);
          r.e_het = 2.0 * pr * (1.0 - pr);
          x_rows.push_back(i);
          x5.push_back(sac.hwe_hom_ref);
          x5.push_back(sac.hwe_het);
          x5.push_back(sac.hwe_hom_alt);
          x5.push_back(sac.geno_hom_ref - sac.hwe_hom_ref);  // male REF hemizygotes
          x5.push_back(sac.geno_hom_alt - sac.hwe_hom_alt);  // male ALT hemizygotes
        }
      } else {
        r.hom_ref_ct = static_cast<int32_t>(sac.geno_hom_ref);
        r.het_ct = static_cast<int32_t>(sac.geno_het);
        r.hom_alt_ct = static_cast<int32_t>(sac.geno_hom_alt);
        r.stats_are_null = 1;
      }
    } else {
      const uint32_t obs = gc[0] + gc[1] + gc[2];
      r.hom_ref_ct = static_cast<int32_t>(gc[0]);
      r.het_ct = static_cast<int32_t>(gc[1]);
      r.hom_alt_ct = static_cast<int32_t>(gc[2]);
      r.stats_are_null = (obs == 0);
      if (obs) {
        r.o_het = static_cast<double>(gc[1]) / static_cast<double>(obs);
        const double pr = (2.0 * gc[0] + gc[1]) / (2.0 * obs);
        r.e_het = 2.0 * pr * (1.0 - pr);
        r.p_hwe = p[i];
      }
    }
    out[i] = r;
  }
  if (!x_rows.empty()) {
    std::vector<double> px(x_rows.size());
    rc = pd_hwe_xchr_from_counts(ctx, x5.data(), static_cast<uint32_t>(x_rows.size()), midp, px.data());
    if (rc != PD_OK) return rc;
    for (size_t k = 0; k < x_rows.size(); ++k) out[x_rows[k]].p_hwe = px[k];
  }
  return PD_OK;
}

int pdop_missing_variant(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, pd_missing_row* out) {
  if (!ctx || v_begin > v_end || (v_end > v_begin && !out)) return PD_ERR_INVALID_ARG;
  const uint32_t n = v_end - v_begin;
  if (!n) return PD_OK;
  std::vector<uint32_t> counts(4ull * n);
  int rc = pd_counts(ctx, v_begin, v_end, counts.data());
  if (rc != PD_OK) return rc;
  for (uint32_t i = 0; i < n; ++i) {
    const uint32_t* gc = &counts[4ull * i];
    const uint32_t sample_ct = EffectiveSampleCt(gc);
    out[i].missing_ct = gc[3];
    out[i].obs_ct = sample_ct - gc[3];
    out[i].f_miss = sample_ct ? static_cast<double>(gc[3]) / static_cast<double>(sample_ct) : 0.0;
  }
  return PD_OK;
}

int pdop_missing_sample(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, pd_missing_row* out) {
  if (!ctx || v_begin > v_end || !out) return PD_ERR_INVALID_ARG;
  pd_info_t info;
  int rc = pd_info(ctx, &info);
  if (rc != PD_OK) return rc;
  const uint32_t sample_ct = info.effective_sample_ct;  // the subset's size, else raw_sample_ct
  std::vector<uint32_t> miss(std::max<uint32_t>(sample_ct, 1));
  rc = pd_missing_by_sample(ctx, v_begin, v_end, miss.data());
  if (rc != PD_OK) return rc;
  const uint32_t total = v_end - v_begin;
  for (uint32_t s = 0; s < sample_ct; ++s) {
    out[s].missing_ct = miss[s];
    out[s].obs_ct = total - miss[s];
    out[s].f_miss = total ? static_cast<double>(miss[s]) / static_cast<double>(total) : 0.0;
  }
  return PD_OK;
}

int pdop_score(pd_ctx* ctx, const uint32_t* vidx, const uint8_t* flip, const double* weights, uint32_t n_var,
               uint32_t mode, pd_score_row* out) {
  if (!ctx || !out || (n_var && (!vidx || !weights))) return PD_ERR_INVALID_ARG;
  // bind rules: drop zero weights, sort by variant index (src/plink_score.cpp:401-407)
  std::vector<uint32_t> order;
  for (uint32_t i = 0; i < n_var; ++i) {
    if (weights[i] != 0.0) order.push_back(i);
  }
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return vidx[a] < vidx[b]; });
  std::vector<uint32_t> v(order.size());
  std::vector<uint8_t> f(order.size());
  std::vector<double> w(order.size());
  for (size_t k = 0; k < order.size(); ++k) {
    v[k] = vidx[order[k]];
    f[k] = flip ? flip[order[k]] : 0;
    w[k] = weights[order[k]];
  }
  pd_info_t info;
  int rc = pd_info(ctx, &info);
  if (rc != PD_OK) return rc;
  const uint32_t sample_ct = info.effective_sample_ct;  // the subset's size, else raw_sample_ct
  std::vector<double> s(std::max<uint32_t>(sample_ct, 1)), d(std::max<uint32_t>(sample_ct, 1));
  std::vector<uint32_t> a(std::max<uint32_t>(sample_ct, 1));
  rc = pd_score(ctx, v.data(), f.data(), w.data(), static_cast<uint32_t>(v.size()), 1, mode, s.data(), d.data(), a.data());
  if (rc != PD_OK) return rc;
  for (uint32_t i = 0; i < sample_ct; ++i) {
    out[i].allele_ct = a[i];
    out[i].denom = a[i];
    out[i].named_allele_dosage_sum = d[i];
    out[i].score_sum = s[i];
    out[i].score_avg = a[i] ? s[i] / static_cast<double>(a[i]) : 0.0;  // src/plink_score.cpp:686-690
  }
  return PD_OK;
}

namespace {
// the derived fields of MakeGenotypeStatsType (src/pgen_reader.cpp:875-899 == src/pfile_reader.cpp:3581-3605)
void FillGenoStats(uint32_t hom_ref, uint32_t het, uint32_t hom_alt, uint32_t missing, pd_geno_stats_row* r) {
  const double kNan = std::numeric_limits<double>::quiet_NaN();
  r->hom_ref = hom_ref;
  r->het = het;
  r->hom_alt = hom_alt;
  r->missing = missing;
  const uint32_t n = hom_ref + het + hom_alt;
  const uint32_t total = n + missing;
  r->n = n;
  r->carrier_count = het + hom_alt;
  if (n == 0) {
    r->af = r->maf = r->het_rate = kNan;
  } else {
    const double af = (static_cast<double>(het) + 2.0 * hom_alt) / (2.0 * n);
    r->af = af;
    r->maf = std::min(af, 1.0 - af);
    r->het_rate = static_cast<double>(het) / static_cast<double>(n);
  }
  r->missing_rate = total ? static_cast<double>(missing) / static_cast<double>(total) : kNan;
}
}  // namespace

int pdop_genotype_stats_variant(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, pd_geno_stats_row* out) {
  if (!ctx || v_begin > v_end || (v_end > v_begin && !out)) return PD_ERR_INVALID_ARG;
  const uint32_t n = v_end - v_begin;
  if (!n) return PD_OK;
  std::vector<uint32_t> counts(static_cast<size_t>(n) * 4);
  const int rc = pd_counts(ctx, v_begin, v_end, counts.data());
  if (rc != PD_OK) return rc;
  for (uint32_t i = 0; i < n; ++i) {
    const uint32_t* c = &counts[static_cast<size_t>(i) * 4];
    FillGenoStats(c[0], c[1], c[2], c[3], &out[i]);
  }
  return PD_OK;
}

int pdop_genotype_stats_sample(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, pd_geno_stats_row* out) {
  if (!ctx || v_begin > v_end || !out) return PD_ERR_INVALID_ARG;
  pd_info_t info;
  int rc = pd_info(ctx, &info);
  if (rc != PD_OK) return rc;
  const uint32_t sample_ct = info.effective_sample_ct;  // the subset's size, else raw_sample_ct
  const size_t cap = std::max<uint32_t>(sample_ct, 1);
  std::vector<uint32_t> het(cap), hom_alt(cap), missing(cap);
  rc = pd_sample_counts(ctx, v_begin, v_end, het.data(), hom_alt.data(), missing.data());
  if (rc != PD_OK) return rc;
  const uint32_t total_eff_variants = v_end - v_begin;  // smp_total_eff_variants, src/pfile_reader.cpp:3573
  for (uint32_t s = 0; s < sample_ct; ++s) {
    FillGenoStats(total_eff_variants - het[s] - hom_alt[s] - missing[s], het[s], hom_alt[s], missing[s], &out[s]);
  }
  return PD_OK;
}

}  // extern "C"
