/* TEST INFRASTRUCTURE ONLY -- never linked into, imported by or executed from the product path.
 *
 * pgen_oracle: a plain-C restatement of the reference's genotype path, small enough to audit by eye:
 * .pgen header/index parse, record decode (vrtype 0-7 incl. difflists, 1-bit and LD-compressed
 * records), genotype counts, missingness, 2-bit -> int8, the exact HWE test and the plink_score
 * accumulation loop.  Every function cites the reference lines it restates
 * (paths relative to the reference checkout; pgenlib = third_party/plink-ng/2.0/include).
 *
 * Pinning: tests/test_cpu_oracle.py checks this port against (1) the known answers of the reference's
 * own sqllogictests (tests/kat.py) and (2) tests/golden/golden.npz, which holds the outputs of the
 * reference's OWN pgenlib (oracle/_ref) over 19 fixtures.
 */
#ifndef PGEN_ORACLE_H_
#define PGEN_ORACLE_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct po_file po_file;

/* pgi_path: external index for storage modes 0x20/0x21, else NULL.  Returns 0 on success. */
int po_open(const char* pgen_path, const char* pgi_path, po_file** out, char* err, size_t errlen);
void po_close(po_file* f);
void po_info(const po_file* f, uint32_t* variant_ct, uint32_t* sample_ct, uint32_t* mode);
uint32_t po_vrtype(const po_file* f, uint32_t vidx);
uint64_t po_fpos(const po_file* f, uint32_t vidx); /* vidx in [0, V] */

/* One byte per sample: 0 hom-ref, 1 het, 2 hom-alt, 3 missing.  (PgrGet, main track.) */
int po_read_codes(po_file* f, uint32_t vidx, uint8_t* codes);
/* counts[v - v0][4] over the samples whose bit is set in include (N bits, LSB first; NULL = all).
 * (PgrGetCounts.) */
int po_counts(po_file* f, uint32_t v0, uint32_t v1, const uint8_t* include, uint32_t* counts);
/* out[(v - v0) * N + s] in {0, 1, 2, -9}.  (PgrGet + GenoarrToBytesMinus9.) */
int po_decode_i8(po_file* f, uint32_t v0, uint32_t v1, int8_t* out);
/* out[((v - v0) * N + s) * 2 + {0,1}] = allele pair: 0,0 / 0,1 or 1,0 / 1,1 / -9,-9.  (PgrGetP +
 * UnpackPhasedGenotypes.) */
int po_decode_phased(po_file* f, uint32_t v0, uint32_t v1, int8_t* out);
/* out[s] = number of variants of [v0, v1) where sample s is missing.  (plink_missing sample mode.) */
int po_missing_by_sample(po_file* f, uint32_t v0, uint32_t v1, uint32_t* out);
/* Exact Hardy-Weinberg test p-value (autosomal, optional mid-p), clamped to [0, 1]. */
double po_hwe_exact(uint32_t obs_hets, uint32_t obs_hom1, uint32_t obs_hom2, int midp);
/* ln p of the same test in x87 long double with a 64-bit exponent (no underflow): the parity yardstick. */
double po_hwe_lnp_exact(uint32_t obs_hets, uint32_t obs_hom1, uint32_t obs_hom2, int midp);
/* plink_score accumulation for one weight column.  mode 0 mean-impute, 1 no_mean_imputation,
 * 2 center.  score/dosage_sum/allele_ct: [N], accumulated into (caller zeroes them). */
int po_score(po_file* f, const uint32_t* vidx, const uint8_t* flip, const double* weights, uint32_t n, int mode,
             double* score, double* dosage_sum, uint32_t* allele_ct);

#ifdef __cplusplus
}
#endif
#endif
