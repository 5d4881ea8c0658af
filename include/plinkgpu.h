/*
 * plinkgpu.h -- C ABI of libplinkgpu.so, the B200 (sm_100a) genotype hot path that
 * PlinkingDuck's table functions call instead of pgenlib.
 *
 * Conventions follow the reference's own C-ABI shim (src/plinking_pgen_vfs.hpp:25-47):
 * extern "C", POD arguments, caller-owned output arrays, no exceptions, no exit(), no
 * printing.  Every function returns 0 on success or one of PD_ERR_* (mirroring pgenlib's
 * PglErr classes the extension already maps to IOException / InvalidInputException); a
 * human-readable message is left in the context and readable with pd_last_error().
 *
 * There is NO CPU fallback: if no CUDA device is usable, pd_open* fails with PD_ERR_CUDA.
 *
 * Thread-safety: a pd_ctx may be shared by DuckDB's scan threads; blocking calls serialise on
 * the context's CUDA stream(s).  A multi-GPU context (pd_open_opts.n_devices) lives in the same single
 * process: it owns one sub-context, stream set and staging ring per GPU plus the NCCL communicator.  pd_counts_fetch may be called concurrently from any number of
 * threads on one job (src/plink_freq.cpp:434-441 keeps its fetch_add claim loop).
 *
 * Each entry point names the reference call site(s) it replaces.  Paths are relative to the
 * reference repository root.
 */
#ifndef PLINKGPU_H_
#define PLINKGPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PD_OK 0
#define PD_ERR_OPEN 1            /* kPglRetOpenFail */
#define PD_ERR_READ 2            /* kPglRetReadFail */
#define PD_ERR_MALFORMED 3       /* kPglRetMalformedInput */
#define PD_ERR_NOT_SUPPORTED 4   /* kPglRetNotYetSupported */
#define PD_ERR_NOMEM 5           /* kPglRetNomem */
#define PD_ERR_INVALID_ARG 6     /* kPglRetImproperFunctionCall */
#define PD_ERR_CUDA 7            /* CUDA runtime / device failure */

typedef struct pd_ctx pd_ctx;
typedef struct pd_job pd_job;

typedef struct pd_open_opts {
  int32_t device;           /* CUDA device ordinal; -1 = current device (unused when n_devices > 0) */
  uint64_t tile_bytes;      /* record bytes per staged tile; 0 = default (64 MiB)               */
  uint32_t n_staging;       /* pinned staging buffers (>= 2, double buffering); 0 = default (3) */
  uint32_t n_devices;       /* > 0: ONE context over the GPUs devices[0 .. n_devices) of this box:
                             * every call splits its variant range (or variant list) into contiguous
                             * pieces of equal record bytes, one per GPU; per-variant outputs land in
                             * disjoint slices of the caller's arrays; the per-sample accumulators of
                             * pd_missing_by_sample / pd_sample_counts / pd_score are summed with one
                             * ncclAllReduce over NVLink (the mutex merges of src/plink_missing.cpp:614-619,
                             * src/plink_score.cpp:657-664, src/pfile_reader.cpp:3432-3445).  NCCL is
                             * loaded at run time (libnccl.so.2) when n_devices > 1. */
  const int32_t* devices;   /* distinct CUDA device ordinals; read during pd_open* only        */
  uint32_t flags;           /* PD_OPEN_* bits                                                   */
  uint32_t raw_variant_ct;  /* PLINK 1 .bed (mode 0x01) only: the file stores neither count, the caller   */
  uint32_t raw_sample_ct;   /* supplies them from .bim / .fam as it does to PgfiInitPhase1                 */
  uint32_t reserved;        /* (src/pfile_reader.cpp:1319); 0 elsewhere                                    */
} pd_open_opts;

/* pd_open only: map the .pgen writable-shared and page-lock the mapping in place (cudaHostRegister), so that
 * every scan DMAs the records straight out of the page cache -- no pread into a staging buffer, which costs the
 * host three memory passes per byte where the reference's fread + popcount costs one.  Registration runs once at
 * open (~0.15 s per GiB per thread; pieces are registered in parallel) and holds the file's pages in RAM for the
 * life of the context: meant for files scanned more than once (a DuckDB session over a page-cache-hot fileset).
 * The file is never written.  Falls back to the pread pipeline, silently, when the file cannot be opened
 * read-write or the driver refuses the registration; pd_info_t.pinned_file tells which path is in use. */
#define PD_OPEN_PIN_FILE 1u  /* shared-memory (tmpfs / hugetlbfs) files: the kernel refuses to page-lock other file mappings */

typedef struct pd_info_t {
  uint32_t variant_ct;      /* raw_variant_ct                                                   */
  uint32_t sample_ct;       /* raw_sample_ct                                                    */
  uint32_t mode;            /* storage mode byte (0x02, 0x03, 0x10, 0x11, 0x20, 0x21, 0x01)     */
  uint32_t fixed_width;     /* record width in bytes for fixed-width modes, else 0             */
  uint32_t has_ld;          /* any vrtype 2/3 record (kfPgenGlobalLdCompressionPresent)         */
  uint32_t has_difflist;    /* any vrtype 1/4/6/7 record                                        */
  uint32_t has_dosage;      /* any record with dosage bits 5-6 (kfPgenGlobalDosagePresent)      */
  uint32_t has_phase;       /* any record with bit 4                                            */
  uint32_t has_multiallelic;/* any record with bit 3                                            */
  uint32_t resident;        /* 1 if a variant range is currently held in HBM                    */
  uint64_t data_bytes;      /* sum of record widths                                             */
  uint64_t resident_bytes;  /* bytes held in HBM (all devices)                                  */
  uint32_t effective_sample_ct; /* samples the outputs cover: the subset's size, else sample_ct  */
  uint32_t n_devices;       /* GPUs behind this context                                         */
  uint32_t pinned_file;     /* 1 if the .pgen mapping is page-locked (PD_OPEN_PIN_FILE took effect) */
  uint32_t reserved;
} pd_info_t;

typedef struct pd_stats_t {
  uint64_t launches;        /* kernels launched by this library since pd_open / last reset      */
  uint64_t h2d_bytes;       /* bytes copied host->device                                        */
  uint64_t d2h_bytes;       /* bytes copied device->host                                        */
  uint64_t record_bytes;    /* algorithmic record bytes scanned by the last call                */
  double kernel_ms;         /* CUDA-event time of the dominant scan kernel(s) of the last call  */
  double aux_kernel_ms;     /* CUDA-event time of follow-up kernels (LD fix-up, HWE, ...)       */
  double total_ms;          /* CUDA-event time of the last call's whole device section          */
  uint64_t scan_launches;   /* launches of the dominant scan kernel in the last call            */
  double call_ms;           /* CUDA-event time of the last call, first enqueue to last result copy */
  double host_stage_ms;     /* host wall time the last call spent filling staging buffers (reads, gathers) */
  double host_wait_ms;      /* host wall time the last call spent waiting for a free staging buffer        */
  double allreduce_ms;      /* CUDA-event time of the last call's NCCL all-reduce(s), first device          */
  uint64_t allreduce_bytes; /* bytes all-reduced by the last call (per device)                              */
  uint32_t n_devices;       /* GPUs behind this context; with several, counters are summed over them and
                             * the *_ms fields are the slowest device's (the devices run side by side)      */
  uint32_t reserved;
} pd_stats_t;

/* ---- open / close : replaces PgfiInitPhase1 + PgfiInitPhase2 + PgrInit per scan thread
 *      (src/plink_freq.cpp:342-390 and the same block in every *InitLocal); the header and
 *      per-variant index are parsed ONCE. ---- */
int pd_open(const char* pgen_path, const pd_open_opts* opts, pd_ctx** out, char* errbuf, size_t errbuf_len);
/* Same, over a .pgen image already in host memory (not copied; must outlive the context).
 * pgi/pgi_len: the external index for modes 0x20/0x21, else NULL/0.
 * pgen_bytes may also be DEVICE memory on the context's device (an image produced by an upstream
 * GPU stage): it is then used in place, permanently resident; the allocation must extend at least
 * 64 readable bytes past len. */
int pd_open_mem(const void* pgen_bytes, uint64_t len, const void* pgi_bytes, uint64_t pgi_len,
                const pd_open_opts* opts, pd_ctx** out, char* errbuf, size_t errbuf_len);
/* Header-only probe, no device needed: what the bind phase does with PgfiInitPhase1/2 to read V, N
 * and validate the file (src/plink_freq.cpp:175-208).  vrtypes (cap entries) / fpos (cap + 1
 * entries) are filled when non-NULL and cap >= variant_ct. */
int pd_probe(const char* pgen_path, pd_info_t* out, uint8_t* vrtypes, uint64_t* fpos, uint32_t cap, char* errbuf,
             size_t errbuf_len);
/* The plan a multi-GPU context uses for [v_begin, v_end) over k GPUs, without a device: cuts[0..k], GPU i takes
 * variants [cuts[i], cuts[i+1]).  Pieces hold (nearly) equal record BYTES and start on a variant whose main track
 * is not LD-compressed (what the 128-variant fetch_add claims of src/plink_freq.cpp:434-441 balance
 * dynamically on the CPU). */
int pd_plan_partition(const char* pgen_path, uint32_t v_begin, uint32_t v_end, uint32_t k, uint32_t* cuts, char* errbuf,
                      size_t errbuf_len);
void pd_close(pd_ctx* ctx);
const char* pd_last_error(const pd_ctx* ctx);
int pd_info(const pd_ctx* ctx, pd_info_t* out);
/* Per-variant index, as GetPgfiVrtype / GetPgfiFpos (pgenlib_read.h:195-214).
 * vrtypes[variant_ct], fpos[variant_ct + 1]; either may be NULL. */
int pd_index(const pd_ctx* ctx, uint8_t* vrtypes, uint64_t* fpos);

/* ---- HBM residency: stage the records of [v_begin, v_end) into device memory once and keep
 *      them for every later call (the whole 125 GB fileset fits one B200). ---- */
int pd_load_resident(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end);
int pd_unload(pd_ctx* ctx);

/* ---- sample subset: replaces BuildSampleSubset (src/plink_common.cpp:1222-1250).
 *      sample_include = raw_sample_ct bits, LSB-first in little-endian 64-bit words, or NULL
 *      for all samples.  Effective sample order = ascending raw index. ---- */
int pd_set_sample_subset(pd_ctx* ctx, const uint64_t* sample_include, uint32_t subset_sample_ct);

/* ---- per-variant genotype counts: replaces PgrGetCounts (src/plink_freq.cpp:482,
 *      src/plink_hardy.cpp:510, src/pgen_reader.cpp:673,864, src/pfile_reader.cpp:2374).
 *      out_counts[4*(v - v_begin) + {0,1,2,3}] = hom_ref, het, hom_alt, missing. ---- */
int pd_counts(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, uint32_t* out_counts);
/* Counts restricted to an extra stratum mask (male / female strata of ComputeSexAwareCounts,
 * src/plink_common.cpp:1996-2108); stratum is AND-ed with the context's subset. */
int pd_counts_masked(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, const uint64_t* stratum_include,
                     uint32_t* out_counts);
/* Asynchronous form for DuckDB's claim loop: start once, fetch row ranges from any thread. */
int pd_counts_start(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, pd_job** out);
int pd_counts_fetch(pd_job* job, uint32_t v_begin, uint32_t v_end, uint32_t* out_counts);
void pd_job_free(pd_job* job);

/* ---- plink_hardy: counts + exact-test p-value on the device: replaces PgrGetCounts +
 *      HweExactTestAutosomal/plink2::HweLnP (src/plink_hardy.cpp:67-79,510,572-589).
 *      out_p[v - v_begin] = P_HWE clamped to [0,1] (1.0 when no observations).
 *      out_counts may be NULL. ---- */
int pd_hardy(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, uint32_t midp, uint32_t* out_counts, double* out_p);
/* Asynchronous form, as pd_counts_start / pd_counts_fetch: PlinkHardyScan's threads claim 128-variant batches with
 * fetch_add (src/plink_hardy.cpp:481-510) and fetch exactly their rows; tiles are staged, counted and tested while
 * earlier rows are being emitted.  out_counts may be NULL.  Free with pd_job_free. */
int pd_hardy_start(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, uint32_t midp, pd_job** out);
int pd_hardy_fetch(pd_job* job, uint32_t v_begin, uint32_t v_end, uint32_t* out_counts, double* out_p);
/* Exact-test p-values for caller-supplied counts (n x 4, as pd_counts), computed on the device. */
int pd_hwe_from_counts(pd_ctx* ctx, const uint32_t* counts, uint32_t n, uint32_t midp, double* out_p);
/* Same test, but out_lnp[i] = ln p: what plink2::HweLnP itself returns (plink2_stats.cc:1784-2166) before
 * LnPToPvalue exponentiates and clamps it (src/plink_hardy.cpp:52-64).  Finite where p underflows to 0
 * (an observed table 10^-400 vs 10^-40000 under HWE), -inf only for a mid-p of exactly 0; 0.0 for tables
 * without observations. */
int pd_hwe_lnp_from_counts(pd_ctx* ctx, const uint32_t* counts, uint32_t n, uint32_t midp, double* out_lnp);

/* ---- Allocation-free form of the two calls above.  It has NO counterpart in the reference: it was added
 *      for callers that cannot call cudaMalloc (code under CUDA graph capture,
 *      tight serving loops, harnesses that forbid it).  No context: d_counts (n x 4 uint32, 16-byte aligned) and
 *      d_out (n doubles) are DEVICE buffers of the CURRENT device, owned by the caller; the kernels are enqueued
 *      on `stream` (a cudaStream_t, NULL = the default stream) and the call returns without synchronising.
 *      want_lnp = 0: p as pd_hwe_from_counts, != 0: ln p as pd_hwe_lnp_from_counts (bit-identical results).
 *      The kernels' scratch (work-queue counters + the list of tables that need the scaled walk) is a static
 *      device array of 16 + 4 * PD_NOALLOC_HWE_MAX_VARIANTS bytes (4 MiB) that every GPU using the library holds
 *      until exit; n beyond the limit returns PD_ERR_INVALID_ARG with a message in errbuf (split the call).
 *      The call allocates and frees nothing (no cudaMalloc / cudaFree / cudaHostAlloc, no library that does).
 *      Every call on a GPU shares that array: calls that overlap on the GPU (other streams, other threads) RACE
 *      and must be serialised by the caller; back-to-back calls on one stream are fine (each clears what must
 *      start at zero).  errbuf may be NULL. ---- */
#define PD_NOALLOC_HWE_MAX_VARIANTS (1u << 20)
uint32_t pd_hwe_noalloc_max_variants(void);
int pd_hwe_from_counts_noalloc(const uint32_t* d_counts, uint32_t n, uint32_t midp, uint32_t want_lnp, double* d_out,
                               void* stream, char* errbuf, size_t errbuf_len);

/* chrX non-PAR exact test on the device: replaces HweExactTestXchr -> plink2::HweXchrLnP
 * (src/plink_hardy.cpp:83-95).  c5[5*i + {0..4}] = female hom_ref, female het, female hom_alt,
 * male REF-hemizygous, male ALT-hemizygous. */
int pd_hwe_xchr_from_counts(pd_ctx* ctx, const uint32_t* c5, uint32_t n, uint32_t midp, double* out_p);

/* ---- plink_missing sample mode: replaces PgrGetMissingness + the ctz bit walk + mutex merge
 *      (src/plink_missing.cpp:585-619).  out_missing[effective sample_ct]. ---- */
int pd_missing_by_sample(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, uint32_t* out_missing);

/* ---- read_pfile(orient := 'sample', genotypes := 'counts'): replaces the dense and sparse
 *      Phase-1 loops (src/pfile_reader.cpp:3322-3445).  Each array is [effective sample_ct]. ---- */
int pd_sample_counts(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, uint32_t* out_het, uint32_t* out_hom_alt,
                     uint32_t* out_missing);

/* ---- genotype decode: replaces PgrGet + GenoarrToBytesMinus9 (src/pgen_reader.cpp:727-733,
 *      src/pfile_reader.cpp:2426-2432).  out is (v_end - v_begin) x effective sample_ct int8,
 *      row-major, values {0,1,2,-9}. ---- */
int pd_decode_i8(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, int8_t* out);
/* Asynchronous form for the scan callbacks of read_pgen / read_pfile, whose threads claim batches of variants from
 * an atomic counter (src/pgen_reader.cpp:635-736, src/pfile_reader.cpp:2790-2835): a job thread decodes the range
 * chunk by chunk (64 MiB of int8 per chunk, four page-locked chunk buffers) ahead of the fetches; pd_decode_fetch
 * blocks until rows [v_begin, v_end) are decoded and copies them to out ((v_end - v_begin) x effective sample_ct).
 * Thread-safe; every row may be fetched once, and fetches must keep up with the claim order (a chunk buffer is
 * reused when all of its rows were fetched).  The sample subset must not change while the job runs. */
int pd_decode_start(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, pd_job** out);
int pd_decode_fetch(pd_job* job, uint32_t v_begin, uint32_t v_end, int8_t* out);

/* ---- genotype column in DuckDB's child-vector layout: replaces PgrGet + GenoarrToBytesMinus9 AND the fill
 *      loop that follows it (src/pgen_reader.cpp:1008-1046: ARRAY(TINYINT, N) / LIST(TINYINT) child data +
 *      validity).  out[(v - v_begin) * n + s] = the call where it is valid, 0 where it is not; bit
 *      ((v - v_begin) * n + s) of validity (64-bit words, LSB first, as duckdb::ValidityMask) = 1 where
 *      valid.  A call is valid when it is not missing and call_mask has bit g set for its category g in
 *      {0,1,2} (GenotypeFilter::AllowsCall, src/include/plink_common.hpp:471-477; 7 = no genotype filter).
 *      validity holds ceil(rows * n / 64) words and is written whole (bits past the last call are 0). ---- */
int pd_decode_i8_valid(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, uint32_t call_mask, int8_t* out,
                       uint64_t* validity);
/* Same for a strictly ascending list of variants (read_pgen's effective_variant_indices, or the variants
 * that survive CheckPreDecompFilters, src/pgen_reader.cpp:659-683): row i is variant vidx[i]. */
int pd_decode_i8_valid_list(pd_ctx* ctx, const uint32_t* vidx, uint32_t n_var, uint32_t call_mask, int8_t* out,
                            uint64_t* validity);

/* ---- genotype-orient ("tidy") rows: replaces the load_variant / emit_sample_row double loop of PfileTidyScan
 *      (src/pfile_reader.cpp:2960-3281) for the columns that come from the genotype path: one row per (variant,
 *      sample) call -- out_variant[i] = variant index, out_sample[i] = sample index in FILE order (what
 *      sample_indices[] maps to, :3081-3082: the caller attaches the .psam columns), out_geno[i] = {0, 1, 2, -9}
 *      (-9: the caller sets the row's genotype NULL, :3163-3167).  emit_mask bit g (0 hom-ref, 1 het, 2 hom-alt,
 *      3 missing) selects the calls that produce a row: 15 = no genotype_range filter, otherwise
 *      GenotypeFilter::AllowsCall for bits 0-2 and include_missing for bit 3 (:3043-3075).  Rows are ordered by
 *      variant, then ascending sample -- the order one scan thread of the reference emits them in.
 *      *out_rows receives the number of rows; with out_geno == NULL nothing else is written (size query: the count
 *      follows from the genotype counts).  row_cap = capacity of the three arrays in rows. ---- */
int pd_tidy_i8(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, uint32_t emit_mask, uint64_t row_cap, uint32_t* out_variant,
               uint32_t* out_sample, int8_t* out_geno, uint64_t* out_rows);

/* ---- phased hardcalls : replaces PgrGetP + UnpackPhasedGenotypes (src/plink_common.cpp:1549-1584; call
 *      sites src/pgen_reader.cpp:705-733, src/pfile_reader.cpp:1722-1745, read_pgen / read_pfile with
 *      phased := true).  out: (v_end - v_begin) x n_out x 2 int8 allele pairs in subset order:
 *      hom-ref 0,0; hom-alt 1,1; missing -9,-9; het 1,0 when the record's hardcall-phase track marks the
 *      sample phased with the ALT allele first, else 0,1 (phased REF|ALT and unphased hets alike, as the
 *      reference).  Variants without a phase track decode every het as 0,1.  Records with BOTH a multiallelic
 *      patch track and a phase track are read once pd_set_allele_counts supplied the allele counts (the phase
 *      track counts the ALTx/ALTy calls of the patch set among its hets: pgenlib_read.cc:6874-6962; pgenlib needs
 *      allele_idx_offsets for the same reason) and return PD_ERR_NOT_SUPPORTED without them. ---- */
int pd_decode_phased(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, int8_t* out);
/* Multiallelic hardcalls: replaces PgrGetM + PglMultiallelicSparseToDenseMiss (pgenlib_read.cc:6668-6681,
 * pgenlib_misc.cc:3552-3578; the aux1a / aux1b patch sets of pgen_spec.tex:469-564).  out[(row * n + s) * 2 + {0,1}] =
 * the two allele codes of sample s (0 = REF, 1 = ALT1, ...; first <= second), 255,255 for a missing call.  Biallelic
 * records give 0,0 / 0,1 / 1,1.  The extension itself reads multiallelic files through the biallelic calls (pd_counts,
 * pd_decode_i8, ...: the main track); this entry point completes the codec. */
int pd_decode_alleles(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, uint8_t* out);
/* Allele count of every variant (n = variant count, each 2..255), as a caller derives them from the .pvar: the
 * drop-in for PgenFileInfo.allele_idx_offsets, which pgenlib expects preloaded the same way (pgenlib_read.cc:695-697,
 * 1113-1121; plink2's own writer does not store them in the .pgen).  Needed by pd_decode_alleles for records with a
 * patch track; files that do store counts in their header (header_ctrl bits 4-5) are read without this call.
 * NULL clears the table. */
int pd_set_allele_counts(pd_ctx* ctx, const uint32_t* allele_cts, uint32_t n);

/* ---- dosage decode: replaces PgrGetD + Dosage16ToDoublesMinus9 (src/plink_score.cpp:587-596,
 *      src/pgen_reader.cpp:694-703).  out is n_var x effective sample_ct double, -9.0 = missing. ---- */
int pd_dosage_f64(pd_ctx* ctx, const uint32_t* vidx, uint32_t n_var, double* out);

/* ---- plink_freq(dosage := true): replaces PgrGetDCounts (src/plink_freq.cpp:475).
 *      out_all_dosages[2*(v-v_begin)+{0,1}] in units of 1/16384 allele; out_imp_r2 may be NULL. ---- */
int pd_dosage_counts(pd_ctx* ctx, uint32_t v_begin, uint32_t v_end, uint32_t* out_counts,
                     uint64_t* out_all_dosages, double* out_imp_r2);

/* ---- plink_score: replaces the Phase-1 loop + merge (src/plink_score.cpp:562-664),
 *      generalised to K weight columns (the reference is K = 1).
 *      vidx[n_var] ascending scored variants, flip[n_var] (may be NULL), weights n_var x K
 *      row-major.  mode: 0 mean-impute, 1 no_mean_imputation, 2 center.
 *      score_sums: sample_ct x K row-major; dosage_sums[sample_ct] (NAMED_ALLELE_DOSAGE_SUM,
 *      weight-independent; may be NULL); allele_cts[sample_ct] (may be NULL). ---- */
#define PD_SCORE_MEAN_IMPUTE 0
#define PD_SCORE_NO_MEAN_IMPUTATION 1
#define PD_SCORE_CENTER 2
int pd_score(pd_ctx* ctx, const uint32_t* vidx, const uint8_t* flip, const double* weights, uint32_t n_var,
             uint32_t K, uint32_t mode, double* score_sums, double* dosage_sums, uint32_t* allele_cts);

/* ---- plink_ld: the sums behind ComputeLdStats (src/plink_ld.cpp:52-86) for a list of variant pairs, as popcount
 *      contractions of the two packed rows: out_sums[6 i + {0..5}] = n (samples with both calls present), sum gA,
 *      sum gB, sum gA gB, sum gA^2, sum gB^2 -- exact integers; pdop_ld_stats (plinkgpu_ops.h) turns them into R2 /
 *      D_PRIME / OBS_CT with the reference's arithmetic.  The caller builds the pair list from its window logic
 *      (src/plink_ld.cpp:620-690, which needs CHROM / POS).  The variant span of a call is staged into HBM for its
 *      duration unless it is already resident; the distinct variants of one call must fit one dense batch
 *      (65 535 rows; 1 GiB of decoded rows for variable-width files). ---- */
int pd_ld_pairs(pd_ctx* ctx, const uint32_t* va, const uint32_t* vb, uint32_t n_pairs, uint32_t* out_sums);

/* ---- plink_pca / plink_glm: standardised genotype rows, replaces PgrGet + GenoarrToBytesMinus9 +
 *      NormalizeGenotypes (src/plink_pca.cpp:895-905, src/plink_common.cpp:1521-1543): out[i * n + s] =
 *      (g - 2 f) / sqrt(2 f (1 - f)) with f the variant's ALT frequency over the (subset) samples
 *      (src/plink_pca.cpp:403-410), 0 where the call is missing.  Variants without observations or with f in {0, 1}
 *      (ComputeVariantNorm's skip) yield an all-zero row and out_skip[i] = 1 (out_skip may be NULL).  vidx strictly
 *      ascending. ---- */
int pd_standardized_f64(pd_ctx* ctx, const uint32_t* vidx, uint32_t n_var, double* out, uint8_t* out_skip);

/* ---- device-side results for multi-GPU reductions.  After pd_score / pd_missing_by_sample /
 *      pd_sample_counts (whose host output pointers may then be NULL), the accumulators of the last
 *      call stay in HBM; the caller all-reduces them in place over NVLink (NCCL) and reads them back
 *      once, instead of bouncing every rank's partial sums through host memory.
 *      which: PD_RESULT_SCORE (double[sample_ct x K]), PD_RESULT_DOSAGE_SUM (double[sample_ct]),
 *      PD_RESULT_ALLELE_CT (uint32[sample_ct]), PD_RESULT_SAMPLE_COUNTS (uint32[3 x sample_ct]:
 *      het, hom_alt, missing).  The pointers are valid until the next call on the context. ---- */
#define PD_RESULT_SCORE 0
#define PD_RESULT_DOSAGE_SUM 1
#define PD_RESULT_ALLELE_CT 2
#define PD_RESULT_SAMPLE_COUNTS 3
int pd_device_result(pd_ctx* ctx, uint32_t which, void** dev_ptr, uint64_t* bytes);

/* ---- instrumentation (the reference has only BindPhaseTimer, src/include/plink_profile.hpp) ---- */
int pd_get_stats(const pd_ctx* ctx, pd_stats_t* out);
int pd_reset_stats(pd_ctx* ctx);
/* Library / build identification: "plinkgpu <version> sm_100a". */
const char* pd_version(void);

/* ---- Multi-file filesets: variant-sharded, row-concatenated sources that share one sample set -- read_pfile's file
 *      lists (src/pfile_reader.cpp:325-400 sources, :1003-1012 the sample-count check, :2133-2151 scan batches that
 *      never span a file boundary).  One context per file (each may span GPUs through opts->n_devices); global variant
 *      index g = variant g - offsets[f] of file f.  Ranged per-variant calls return the concatenation over the files a
 *      range touches; per-sample calls return the sums over them (added in file order).  Every other entry point is
 *      reached per file through pd_fileset_ctx.  A fileset is used by one thread at a time. ---- */
typedef struct pd_fileset pd_fileset;
int pd_fileset_open(const char* const* pgen_paths, uint32_t n_files, const pd_open_opts* opts, pd_fileset** out,
                    char* errbuf, size_t errbuf_len);
void pd_fileset_close(pd_fileset* fs);
const char* pd_fileset_last_error(const pd_fileset* fs);
uint32_t pd_fileset_file_count(const pd_fileset* fs);
uint32_t pd_fileset_variant_ct(const pd_fileset* fs);   /* total over the files */
uint32_t pd_fileset_sample_ct(const pd_fileset* fs);
int pd_fileset_offsets(const pd_fileset* fs, uint32_t* out /* file_count + 1 */);
pd_ctx* pd_fileset_ctx(pd_fileset* fs, uint32_t file);
int pd_fileset_set_sample_subset(pd_fileset* fs, const uint64_t* sample_include, uint32_t subset_sample_ct);
int pd_fileset_counts(pd_fileset* fs, uint32_t g_begin, uint32_t g_end, uint32_t* out_counts);
int pd_fileset_hardy(pd_fileset* fs, uint32_t g_begin, uint32_t g_end, uint32_t midp, uint32_t* out_counts, double* out_p);
int pd_fileset_decode_i8(pd_fileset* fs, uint32_t g_begin, uint32_t g_end, int8_t* out);
int pd_fileset_missing_by_sample(pd_fileset* fs, uint32_t g_begin, uint32_t g_end, uint32_t* out_missing);
int pd_fileset_sample_counts(pd_fileset* fs, uint32_t g_begin, uint32_t g_end, uint32_t* out_het, uint32_t* out_hom_alt,
                             uint32_t* out_missing);
int pd_fileset_score(pd_fileset* fs, const uint32_t* g_vidx, const uint8_t* flip, const double* weights, uint32_t n_var,
                     uint32_t K, uint32_t mode, double* score_sums, double* dosage_sums, uint32_t* allele_cts);

#ifdef __cplusplus
}
#endif
#endif /* PLINKGPU_H_ */
