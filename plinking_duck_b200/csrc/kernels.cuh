// Launch wrappers of the sm_100a scan kernels (definitions in kernels_*.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pdk {

// Where record v lives in device memory.
struct RecIndex {
  const uint8_t* data;      // device byte holding file offset data_fpos (>= 16 B slack both sides)
  uint64_t data_fpos;
  const uint64_t* fpos;     // [V+1] file offsets, nullptr for fixed-width files
  const uint8_t* vrtypes;   // [V], nullptr for fixed-width files
  const uint32_t* ldbase;   // [V], nullptr for fixed-width files
  const uint32_t* base_slot;// [V] scratch slot of the LD base when it is not a type-0 record, else ~0u
  uint64_t fixed_start;     // file offset of record 0 (fixed-width)
  uint32_t fixed_width;
  uint32_t const_vrtype;
  uint32_t N;               // raw sample count
  const uint8_t* allele_cts;// [V] allele count per variant, nullptr when the file stores none (all biallelic)
};

// Sample subset (and strata) in the layouts the kernels consume.
struct SubsetDev {
  const uint8_t* mask2_shifts;  // 16 shifted copies of the 2-bit-expanded include mask; nullptr = all samples
  uint32_t mask2_stride;
  const uint8_t* mask1_shifts;  // 16 shifted copies of the 1-bit include mask
  uint32_t mask1_stride;
  const uint32_t* include_words;// unshifted 1-bit mask as 32-bit words, padded
  const uint32_t* cum_popcount; // exclusive prefix popcount per 32-bit word (RawToSubsettedPos)
  const uint32_t* subset_idx;   // [subset_ct] raw sample index of every subset sample, ascending
  uint32_t subset_ct;
};

// Scratch for dense copies of non-type-0 LD bases.
struct LdScratch {
  uint8_t* dense = nullptr;  // slot s at dense + (s - slot_lo) * stride
  uint32_t stride = 0;       // bytes per slot, multiple of 16
  uint32_t slot_lo = 0;
  uint32_t slot_ct = 0;           // slots decoded (k_densify checks it; 0 = no scratch)
};

// Per-context launch state: every launch wrapper receives the owning context's copy, so one process can
// drive several devices (no process-wide caches of device properties).
struct LaunchStats {
  uint64_t launches = 0;
  int sm_count = 0;  // multiprocessors of the context's device (set at pd_open)
};
inline int sm_of(const LaunchStats* st) { return st && st->sm_count > 0 ? st->sm_count : 148; }
// Per-device function attributes (dynamic shared memory of k_score_mma); call with the device current.
void configure_device_kernels();

// ---- subset mask construction ----
void launch_build_masks(const uint32_t* include_words, uint32_t N, uint8_t* mask2_shifts, uint32_t mask2_stride,
                        uint8_t* mask1_shifts, uint32_t mask1_stride, cudaStream_t s, LaunchStats* st);
void launch_cum_popcount(const uint32_t* include_words, uint32_t nwords, uint32_t* cum, cudaStream_t s, LaunchStats* st);

// ---- K1/K2: genotype counts over fixed-width 2-bit records ----
// out[v - v_begin] = (hom_ref, het, hom_alt, missing).  main_bytes_only: records may carry
// trailing dosage tracks (modes 0x03/0x04); only the leading ceil(N/4) bytes are counted.
void launch_count_fixed(const RecIndex& ix, const SubsetDev& sub, uint32_t v_begin, uint32_t v_end, uint4* out,
                        cudaStream_t s, LaunchStats* st);

// ---- PLINK 1 .bed records (mode 0x01) -> .pgen 2-bit codes, in place in a staged tile / the resident copy:
// bed 3 / 2 / 0 / 1 = hom-ref / het / hom-alt / missing becomes 0 / 1 / 2 / 3 (what pgenlib does per record in
// PgrPlink1ToPlink2InplaceUnsafe, and to the counts at pgenlib_read.cc:3709-3719); the pad codes of a record's
// last byte are cleared.  data + first_off = first byte of record 0 of the range.
void launch_bed_to_pgen(uint8_t* data, uint64_t first_off, uint32_t width, uint32_t N, uint32_t nrec, cudaStream_t s,
                        LaunchStats* st);

// ---- K1h: plink_hardy over fixed-width records, counts and exact test in one kernel ----
// Returns false (nothing launched) when the records are too short for the fused form; the caller
// then runs launch_count_fixed + launch_hwe.  queue: as launch_hwe (kHweQueueWords + nrec words); ready: device uint32[nrec].
// out_p may be the device view of page-locked host memory (p-values are write-only for the kernel);
// host_out, when given, is such a view for the counts, written in addition to out.
bool hardy_fused_applicable(uint32_t N);
bool launch_hardy_fixed(const RecIndex& ix, const SubsetDev& sub, uint32_t v_begin, uint32_t v_end, uint4* out,
                        double* out_p, uint32_t midp, uint32_t* queue, uint32_t* ready, cudaStream_t s,
                        LaunchStats* st, uint4* host_out = nullptr);

// ---- K3: genotype counts over variable-width records (vrtype 0-7, LD resolved) ----
// Scans [v_begin, v_end); v_begin must not be LD-compressed.  out[v - v_begin].
// work_counter: device uint32 (zeroed by the wrapper).  err: device int[2] = {code, vidx}.
void launch_count_var(const RecIndex& ix, const SubsetDev& sub, const LdScratch& scratch, uint32_t v_begin,
                      uint32_t v_end, uint4* out, uint32_t* work_counter, int* err, cudaStream_t s, LaunchStats* st);

// Dense 2-bit copies of the LD bases slot_lo..slot_hi (inclusive) listed in slot_vidx.
void launch_decode_bases(const RecIndex& ix, const uint32_t* slot_vidx, const LdScratch& scratch, uint32_t slot_hi,
                         int* err, cudaStream_t s, LaunchStats* st);

// ---- dense 2-bit rows of a batch of variants (K3b) and their sample-wise consumers (K4, K5) ----
struct DenseRows {
  const uint8_t* const* ptrs;  // [nrows] byte pointers (any alignment), or nullptr:
  const uint8_t* base;         //   row r at base + r * stride (fixed-width records, used in place)
  uint64_t stride;
  uint32_t nrows;
  uint32_t N;                  // raw sample count; bits past N in a row are undefined
  const uint32_t* nrows_dev;   // when set: the row count lives on the device (<= nrows), see SparseAcc
};

// Sparse per-sample accumulation (plink_missing sample mode / read_pfile orient := 'sample' on
// difflist records): with this passed to launch_densify, records of type 4/5/6/7 are never expanded.
// Their difflist entries go straight into the per-sample plane counters as (entry - common) deltas,
// the common genotype's planes are added once per record to base[0..2], and only the remaining
// records (raw, 1-bit, LD) are decoded; their row pointers are appended to ptrs in arrival order
// and counted in *dense_count.  The sparse walk is the GPU form of the reference's Phase 1
// "sparse" branch (src/pfile_reader.cpp:3322-3428, PgrGetDifflistOrGenovec).
struct SparseAcc {
  uint32_t kinds;        // as launch_sample_acc
  uint32_t* acc_e;       // [N] plane counters, the same arrays launch_sample_acc adds to
  uint32_t* acc_o;
  uint32_t* acc_b;
  uint32_t* base;        // [3] E / O / B counts owed to every sample (zeroed by the caller per operation)
  uint32_t* dense_count; // [1] rows appended to ptrs by this launch (zeroed by launch_densify)
};

// bases: dense copies of the LD bases of the batch (launch_decode_bases); LD rows then copy their base instead of
// decoding it again.
// Row r = variant vlist[r] (vlist != nullptr) or v0 + r.  Type-0 records are pointed at in place,
// everything else (difflist, 1-bit, LD with its base resolved) is decoded into scratch slot r.
void launch_densify(const RecIndex& ix, const uint32_t* vlist, uint32_t v0, uint32_t nrows, uint8_t* scratch,
                    uint32_t slot_stride, const uint8_t** ptrs, uint32_t* work_counter, int* err, cudaStream_t s,
                    LaunchStats* st, const SparseAcc* sparse = nullptr, const LdScratch* bases = nullptr);
// out[(r * n_out + j) * 2 + {0,1}] = allele pair of sample subset_idx[j] (or j) -- PgrGetP +
// UnpackPhasedGenotypes.  raw_scratch: nrows x roundup8(N) x 2 bytes, only used with a subset.
void launch_expand_phased(const RecIndex& ix, const DenseRows& rows, const uint32_t* vlist, uint32_t v0,
                          const uint32_t* subset_idx, uint32_t n_out, int8_t* raw_scratch, int8_t* out, int* err,
                          cudaStream_t s, LaunchStats* st, uint32_t* mhet_scratch = nullptr);
// mhet_scratch: phased_mhet_scratch_bytes(N) of device memory, needed only for records that carry BOTH a multiallelic
// patch track and a phase track (one bit array of N bits per CTA of the launch); nullptr -> such records report
// PD_ERR_NOT_SUPPORTED.
constexpr uint32_t kPhasedMaxBlocks = 148u * 16u;
inline uint32_t phased_mhet_stride_words(uint32_t N) { return ((N + 31u) >> 5) + 2u; }
size_t phased_mhet_scratch_bytes(uint32_t N);
// K5v: int8 calls (0 where invalid) + validity bits in DuckDB's child-vector layout; vbits must be zero-filled
// except for bits below bitbase in its first word (the previous batch's calls).
void launch_expand_i8_valid(const DenseRows& rows, const uint32_t* subset_idx, uint32_t n_out, uint32_t call_mask,
                            int8_t* out, uint32_t* vbits, uint32_t bitbase, cudaStream_t s, LaunchStats* st = nullptr);
// out[r * n_out + j] = {0,1,2,-9} of sample subset_idx[j] (or j) -- PgrGet + GenoarrToBytesMinus9.
void launch_expand_i8(const DenseRows& rows, const uint32_t* subset_idx, uint32_t n_out, int8_t* out, cudaStream_t s,
                      LaunchStats* st);
// Per raw sample, number of rows whose E (low) bit / O (high) bit / both bits are set:
// acc_het += |E|, acc_hom += |O|, acc_mis += |E & O|; het = E - B, hom_alt = O - B, missing = B
// (launch_planes_to_counts).  kinds: 4 = missing (B) only, 7 = all three planes.
void launch_sample_acc(const DenseRows& rows, uint32_t kinds, uint32_t* acc_het, uint32_t* acc_hom, uint32_t* acc_mis,
                       cudaStream_t s, LaunchStats* st);
void launch_planes_to_counts(const uint32_t* pe, const uint32_t* po, const uint32_t* pb, const uint32_t* idx, uint32_t n,
                             uint32_t* out, cudaStream_t s, LaunchStats* st, const uint32_t* base = nullptr);
// dst[i] = src[idx[i]] (+ *add when given)
void launch_gather_u32(const uint32_t* src, const uint32_t* idx, uint32_t n, uint32_t* dst, cudaStream_t s,
                       LaunchStats* st, const uint32_t* add = nullptr);
void launch_subset_indices(const uint32_t* include_words, const uint32_t* cum, uint32_t nwords, uint32_t* out,
                           cudaStream_t s, LaunchStats* st);

// ---- K5t: genotype-orient ("tidy") rows: (variant, sample, call) per emitted call (kernels_tidy.cu) ----
// emit_mask bit g: calls of category g (0 hom-ref, 1 het, 2 hom-alt, 3 missing) produce a row.
void launch_tidy_rows(const uint4* counts, uint32_t n, uint32_t emit_mask, unsigned long long* rows_per, cudaStream_t s,
                      LaunchStats* st);
// row_off[r] - off0 = first output row of batch row r; variant = vlist ? vlist[r] : variant0 + r.
void launch_tidy(const DenseRows& rows, const uint32_t* subset_idx, uint32_t n_out, uint32_t emit_mask,
                 const unsigned long long* row_off, unsigned long long off0, uint32_t variant0, const uint32_t* vlist,
                 uint32_t* out_variant, uint32_t* out_sample, int8_t* out_geno, cudaStream_t s, LaunchStats* st);

// ---- plink_ld / plink_pca / plink_glm consumers of the dense rows (kernels_ld.cu) ----
// out[6 i + {0..5}] = n, sum gA, sum gB, sum gA gB, sum gA^2, sum gB^2 over the samples where both calls of pair i
// (rows ra[i], rb[i]) are present (and, with include_words, selected).
void launch_ld_pairs(const DenseRows& rows, const uint32_t* ra, const uint32_t* rb, uint32_t n_pairs,
                     const uint32_t* include_words, uint32_t* out, cudaStream_t s, LaunchStats* st);
// out[r * n_out + j] = (g - norm[r].x) * norm[r].y, 0 where the call is missing.
void launch_standardize(const DenseRows& rows, const uint32_t* subset_idx, uint32_t n_out, const double2* norm, double* out,
                        cudaStream_t s, LaunchStats* st);

// ---- K6: dosage tracks ----
struct DosageLoc {            // where a row's dosage tracks are (written by launch_dosage_locate)
  const uint8_t* track3;      // sample-id list (kind 1) or N-bit array (kind 3)
  const uint8_t* values;      // little-endian u16 dosages, any alignment
  uint32_t kind;              // 0 none, 1 list, 2 unconditional (65535 = missing), 3 bitarray
  uint32_t raw_ct;            // stored values
};
struct DosageStats {          // over the subset samples that carry a dosage
  unsigned long long sum;     // sum of u16 dosages
  unsigned long long sumsq;   // sum of squares
  uint32_t ct;
  uint32_t replaced[4];       // hardcall categories of those samples
  uint32_t kind;
  uint32_t pad;
};
// rank: [nrows x rank_stride] scratch for the bitarray form (exclusive prefix popcounts per 32 bits).
void launch_dosage_locate(const RecIndex& ix, const DenseRows& rows, const uint32_t* vlist, uint32_t v0, DosageLoc* loc,
                          uint32_t* rank, uint32_t rank_stride, int* err, cudaStream_t s, LaunchStats* st);
// out[r * n_out + j] = ALT dosage of subset sample j (hardcall where none is stored, -9 = missing).
void launch_dosage_fill(const DenseRows& rows, const DosageLoc* loc, const uint32_t* rank, uint32_t rank_stride,
                        const SubsetDev& sub, uint32_t n_out, double* out, int* err, cudaStream_t s, LaunchStats* st);
void launch_dosage_stats(const DenseRows& rows, const DosageLoc* loc, const uint32_t* rank, uint32_t rank_stride,
                         const SubsetDev& sub, uint32_t n_out, DosageStats* stats, cudaStream_t s, LaunchStats* st);
// Genotype counts of dense rows (any alignment), optionally under the subset mask.
void launch_count_rows(const DenseRows& rows, const SubsetDev& sub, uint4* out, cudaStream_t s, LaunchStats* st);

// ---- K7: plink_score ----
constexpr uint32_t kScoreActive = 1u;   // row contributes (non_missing_ct > 0, sd > 0)
constexpr uint32_t kScoreFlip = 2u;     // scored allele is REF: dosage -> 2 - dosage
constexpr uint32_t kScoreCenter = 4u;   // (scored - mean) / sd
constexpr uint32_t kScoreImpute = 8u;   // missing calls count with the mean dosage
struct ScoreRow {
  double T[4];          // contribution of hardcall 0 / 1 / 2 / missing
  double mean_scored;
  double sd;
  uint32_t flags;
  uint32_t kind;        // dosage storage kind of the row (DosageLoc::kind)
  uint32_t cmask;       // bit g set: hardcall g counts towards ALLELE_CT (0 for skipped rows)
  uint32_t pad;
};
// out[(r * n_out + j) * 2 + {0,1}] = allele codes of sample subset_idx[j] (or j) -- PgrGetM +
// PglMultiallelicSparseToDenseMiss: 0,0 / 0,1 / 1,1 / 255,255 from the main track, patched by the aux1a / aux1b sets of
// multiallelic records.  raw_scratch: nrows x ((N + 7) & ~7) x 2 bytes, only used with a subset.
void launch_expand_alleles(const RecIndex& ix, const DenseRows& rows, const uint32_t* vlist, uint32_t v0,
                           const uint32_t* subset_idx, uint32_t n_out, uint8_t* raw_scratch, uint8_t* out, int* err,
                           cudaStream_t s, LaunchStats* st);
// out[r, j] = raw[r, idx[j]] as 16-bit pairs (sample subset of a row of pairs).
void launch_gather_pairs(const uint16_t* raw, uint32_t raw_stride, const uint32_t* idx, uint32_t n_out, uint32_t nrows,
                         uint16_t* out, cudaStream_t s, LaunchStats* st);
// counts: (subset) hardcall counts per row; stats: dosage sums per row or nullptr; flip: per row or nullptr.
void launch_score_prep(const uint4* counts, const DosageStats* stats, const uint8_t* flip, uint32_t nrows, uint32_t mode,
                       ScoreRow* out, cudaStream_t s, LaunchStats* st);
// weights: nrows x K (row-major) for the rows of this batch.  score: n_out x K, dosage_sums / allele_cts: n_out;
// all three are accumulated into.  loc / rank may be nullptr when the file stores no dosages.
void launch_score_acc(const DenseRows& rows, const ScoreRow* srow, const DosageLoc* loc, const uint32_t* rank,
                      uint32_t rank_stride, const double* weights, uint32_t K, const SubsetDev& sub, uint32_t n_out,
                      uint32_t mode, double* score, double* dosage_sums, uint32_t* allele_cts, cudaStream_t s,
                      LaunchStats* st);

// ---- K7t: the same accumulation on the tensor cores (tcgen05 int8 limbs, exact integer sums) for
// hardcall rows without a sample subset.  workspace: score_tc_workspace_bytes(max_rows, n_samples)
// bytes of device memory kept across the batches of one call; first_batch clears its accumulators.
size_t score_tc_workspace_bytes(uint32_t max_rows, uint32_t n_samples);
bool score_tc_applicable(uint32_t K);
void launch_score_tc(const DenseRows& rows, const uint4* counts, const ScoreRow* srow, const double* weights, uint32_t K,
                     uint32_t mode, uint32_t n_samples, uint32_t max_rows, void* workspace, bool first_batch,
                     double* score, double* dosage_sums, uint32_t* allele_cts, cudaStream_t s, LaunchStats* st);

// ---- K8: exact Hardy-Weinberg test on the device ----
// counts[n] as produced by the count kernels; out_p[n] = p-value clamped to [0,1].
// queue: device uint32[kHweQueueWords + n], zeroed by the wrapper: [0] K1h record-chunk counter, [1] window
// counter of the exact-test warps, [2] number of tables set aside for k_hwe_scaled, [3] pad, [4..] those tables.
constexpr uint32_t kHweQueueWords = 4;
// Finishes the tables the walk kernels set aside (queue as above, filled by the preceding launch on s).
void launch_hwe_scaled(const uint4* counts, const uint32_t* queue, uint32_t midp, double* out_p, bool want_lnp,
                       cudaStream_t s, LaunchStats* st);
// want_lnp: store ln p (HweLnP's own return value; finite where p underflows) instead of p.
void launch_hwe(const uint4* counts, uint32_t n, uint32_t midp, double* out_p, uint32_t* queue, cudaStream_t s,
                LaunchStats* st, bool want_lnp = false);
// chrX exact test: c5[n][5] = female hom_ref, het, hom_alt, male ref, male alt.
void launch_hwe_xchr(const uint32_t* c5, uint32_t n, uint32_t midp, double* out_p, cudaStream_t s, LaunchStats* st);

}  // namespace pdk
