/*
 * plinkgpu_ops.h -- operator layer above plinkgpu.h: what the reference's Scan callbacks compute per
 * output row from the hot-path results, written once in C++
 * so that the extension (or any other host) gets the reference's column values and NULL rules
 * without re-deriving them.  Each function names the reference code it mirrors; everything here
 * goes through the public C ABI of plinkgpu.h only.
 *
 * NULLs are reported through flag bytes next to the values (DuckDB validity is the caller's job).
 */
#ifndef PLINKGPU_OPS_H_
#define PLINKGPU_OPS_H_

#include "plinkgpu.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ChromPloidy (src/include/plink_common.hpp:568-573) */
#define PD_PLOIDY_AUTOSOMAL 0
#define PD_PLOIDY_CHR_X 1
#define PD_PLOIDY_CHR_Y 2
#define PD_PLOIDY_CHR_MT 3

/* ClassifyChromPloidy + ResolveParBounds (src/plink_common.cpp:1926-1979).  build: "GRCh38"/"hg38"/
 * "b38"/"38", "GRCh37"/"hg19"/"b37"/"37", or "none"/"" (PAR only through explicit chromosome codes).
 * Returns the ploidy class, or -1 for an unrecognised build string. */
int pdop_classify_ploidy(const char* chrom, int32_t pos, const char* build);

/* Sex strata for the sex-aware paths.  sex[raw_sample_ct]: 0 unknown, 1 male, 2 female
 * (psam_reader SampleInfo::sexes); NULL = no sex information (have_sex = false). */

/* ---- plink_freq (src/plink_freq.cpp:443-544) ---- */
typedef struct pd_freq_row {
  double alt_freq;      /* ALT_FREQ                                   */
  double imp_r2;        /* IMP_R2 (dosage mode only)                  */
  uint32_t obs_ct;      /* OBS_CT (allele observations)               */
  int32_t hom_ref_ct, het_ct, hom_alt_ct, missing_ct;
  uint8_t freq_is_null;    /* ALT_FREQ / OBS_CT are NULL              */
  uint8_t counts_are_null; /* the four count columns are NULL         */
  uint8_t pad[2];
} pd_freq_row;
/* ploidy: per variant of the FILE ([variant_ct]) or NULL (all autosomal).  dosage != 0 selects the
 * PgrGetDCounts path for autosomal variants (dosage := true). */
int pdop_freq(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, const uint8_t* ploidy, const uint8_t* sex, uint32_t dosage,
              pd_freq_row* out);

/* ---- plink_hardy (src/plink_hardy.cpp:481-589) ---- */
typedef struct pd_hardy_row {
  double o_het, e_het, p_hwe;
  int32_t hom_ref_ct, het_ct, hom_alt_ct;
  uint8_t stats_are_null;  /* O_HET / E_HET / P_HWE are NULL          */
  uint8_t counts_are_null;
  uint8_t pad[2];
} pd_hardy_row;
int pdop_hardy(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, const uint8_t* ploidy, const uint8_t* sex, uint32_t midp,
               pd_hardy_row* out);

/* ---- plink_missing (src/plink_missing.cpp:472-490 variant mode, :585-646 sample mode) ---- */
typedef struct pd_missing_row {
  uint32_t missing_ct, obs_ct;
  double f_miss;
} pd_missing_row;
int pdop_missing_variant(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, pd_missing_row* out);
/* out[effective sample_ct]; the denominator is the number of variants scanned. */
int pdop_missing_sample(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, pd_missing_row* out);

/* ---- plink_score (src/plink_score.cpp:309-427 bind rules, :562-751 scan) ----
 * One weight vector, like the reference.  Entries with weight == 0 are dropped before scoring
 * (src/plink_score.cpp:401,419); vidx need not be sorted (it is sorted here, :407). */
typedef struct pd_score_row {
  uint32_t allele_ct;   /* ALLELE_CT */
  uint32_t denom;       /* DENOM (= ALLELE_CT) */
  double named_allele_dosage_sum;
  double score_sum;
  double score_avg;     /* SCORE_SUM / ALLELE_CT, 0 when ALLELE_CT == 0 */
} pd_score_row;
int pdop_score(pd_ctx* ctx, const uint32_t* vidx, const uint8_t* flip, const double* weights, uint32_t n_var,
               uint32_t mode, pd_score_row* out);

/* ---- count-based pre-decompression filters (src/plink_common.cpp:1439-1515) ----
 * VariantPassesCountFilter: af / ac ranges evaluated from genotype counts.  Negative bound = unset. */
int pdop_counts_pass(const uint32_t counts[4], double af_min, double af_max, int64_t ac_min, int64_t ac_max);

/* CheckGenotypeRange (src/plink_common.cpp:1465-1493): allowed_mask bit g (g = 0 hom_ref, 1 het, 2 hom_alt)
 * = category g is selected.  *any_pass: some kept sample falls in a selected category (or is missing while
 * include_missing is set); *all_pass: no sample falls in an unselected hardcall category. */
void pdop_genotype_range(const uint32_t counts[4], uint32_t allowed_mask, uint32_t include_missing, uint32_t* any_pass,
                         uint32_t* all_pass);

/* CheckPreDecompFilters (src/plink_common.cpp:1495-1515) over a variant range, on GPU counts: one pd_counts
 * call (honouring the context's sample subset), then the two filters per variant.  skip[i] = 1: the variant
 * is dropped before any decode; all_pass[i] = 1: no per-element null-out is needed.  use_count_filter = 0
 * ignores the af / ac bounds, use_genotype_filter = 0 the category mask.  Either output may be NULL. */
typedef struct pd_predecomp_filter {
  uint32_t use_count_filter;
  double af_min, af_max;   /* negative = unset */
  int64_t ac_min, ac_max;  /* negative = unset */
  uint32_t use_genotype_filter;
  uint32_t allowed_mask;
  uint32_t include_missing;
} pd_predecomp_filter;
int pdop_predecomp_filter(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, const pd_predecomp_filter* filter, uint8_t* skip,
                          uint8_t* all_pass);

/* ---- plink_ld output columns from pd_ld_pairs sums: ComputeLdStats' arithmetic after its sample loop
 * (src/plink_ld.cpp:88-134), same operation order, so R2 / D_PRIME are bit-identical. ---- */
typedef struct pd_ld_row {
  double r2, d_prime;
  uint32_t obs_ct;
  uint32_t is_valid;   /* 0: fewer than 2 observations or a monomorphic variant -> R2 / D_PRIME are NULL */
} pd_ld_row;
void pdop_ld_stats(const uint32_t* sums, uint32_t n_pairs, pd_ld_row* out);

/* ---- LIST(TINYINT) genotype columns (genotypes := 'list', src/pgen_reader.cpp:1027-1046): the list_entry_t
 * {offset, length} pair of each of `rows` consecutive rows whose child values are pd_decode_i8_valid's output:
 * out[2 r] = base_offset + r * n, out[2 r + 1] = n. ---- */
void pdop_list_entries(uint64_t base_offset, uint32_t rows, uint32_t n, uint64_t* out);

/* ---- read_pgen / read_pfile genotype column for a range of variants, as the scan emits it
 * (src/pgen_reader.cpp:659-683 filter, :727-733 decode, :1008-1046 fill): counts on the GPU ->
 * CheckPreDecompFilters per variant -> the variants that are not skipped are decoded, in order, into
 * DuckDB's child-vector layout (pd_decode_i8_valid_list; call mask = the genotype filter's categories, 7 when it
 * is inactive).  kept[0 .. *n_kept) receives their variant indices; out / validity hold *n_kept rows and must
 * be sized for v_end - v_begin.  filter may be NULL (nothing skipped, nothing nulled but missing calls). */
int pdop_genotype_column(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, const pd_predecomp_filter* filter, int8_t* out,
                         uint64_t* validity, uint32_t* kept, uint32_t* n_kept);

/* ---- read_pgen / read_pfile with genotypes := 'counts' | 'stats' (MakeGenotypeCountsType /
 * MakeGenotypeStatsType, src/plink_common.cpp:43-68).  Variant orient: one PgrGetCounts per variant, then the
 * derived fields (src/pgen_reader.cpp:863-901, src/pfile_reader.cpp:2548-2600).  Sample orient: the per-sample
 * het / hom_alt / missing tallies of Phase 1, hom_ref = variants scanned - het - hom_alt - missing
 * (src/pfile_reader.cpp:3566-3607).  The counts mode is the first four fields of the same row.
 * af / maf / het_rate are NaN when n == 0 and missing_rate is NaN when n + missing == 0, as in the reference. */
typedef struct pd_geno_stats_row {
  uint32_t hom_ref, het, hom_alt, missing;
  uint32_t n;              /* hom_ref + het + hom_alt                  */
  uint32_t carrier_count;  /* het + hom_alt                            */
  double af, maf, missing_rate, het_rate;
} pd_geno_stats_row;
/* out[v_end - v_begin]; honours the context's sample subset. */
int pdop_genotype_stats_variant(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, pd_geno_stats_row* out);
/* out[effective sample_ct]; the variants scanned are [v_begin, v_end). */
int pdop_genotype_stats_sample(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, pd_geno_stats_row* out);

#ifdef __cplusplus
}
#endif
#endif /* PLINKGPU_OPS_H_ */
