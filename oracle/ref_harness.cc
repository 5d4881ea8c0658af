// TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path.
//
// Thin C-ABI harness around the PRISTINE pgenlib sources that live under
// /root/reference/third_party/plink-ng/2.0/include (compiled where they lie by
// oracle/Makefile; nothing from the reference is copied into this repo).  It exposes the
// reference's own implementation of the genotype hot path so that tests/ and bench.py's
// cpu_baseline / --impl reference legs can use it as the checker and the timed CPU arm:
//
//   * open sequence             = src/plink_freq.cpp:342-396 (PgfiInitPhase1/2 + PgrInit per reader)
//   * sample subset             = src/plink_common.cpp:1222-1250 (BuildSampleSubset)
//   * counts                    = PgrGetCounts          (src/plink_freq.cpp:482)
//   * dosage counts             = PgrGetDCounts         (src/plink_freq.cpp:475)
//   * int8 decode               = PgrGet + GenoarrToBytesMinus9 (src/pfile_reader.cpp:2426-2432)
//   * missingness               = PgrGetMissingness     (src/plink_missing.cpp:479,593-609)
//   * dosage decode             = PgrGetD + Dosage16ToDoublesMinus9 (src/plink_score.cpp:587-596)
//   * score accumulation        = the loop of src/plink_score.cpp:598-653 around those calls
//   * sample-orient counts      = src/pfile_reader.cpp:3322-3428 (dense and sparse paths)
//   * HWE                       = plink2::HweLnP / HweXchrLnP (src/plink_hardy.cpp:67-95)
//   * multi-threaded scans      = fetch_add(128) claim loop of src/plink_freq.cpp:434-441, one
//                                 reader per thread (src/plink_freq.cpp:342-390)
//   * writer (fixture generator)= SpgwAppendBiallelicGenovec[Dosage16] (pgenlib_write.h:229,332)
//
// The extension-side arithmetic (score loop, sample-orient accumulation) is restated here
// because it lives in DuckDB scan callbacks that cannot be built without DuckDB.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "pgenlib_ffi_support.h"
#include "pgenlib_read.h"
#include "pgenlib_write.h"
#include "plink2_stats.h"

using namespace plink2;

namespace {

struct Reader {
  PgenFileInfo pgfi;
  PgenReader pgr;
  unsigned char* pgfi_alloc = nullptr;
  unsigned char* pgr_alloc = nullptr;
  uint32_t V = 0, N = 0;
  // subset
  uintptr_t* sample_include = nullptr;
  uintptr_t* interleaved = nullptr;
  uint32_t* cumpop = nullptr;
  uint32_t sample_ct = 0;
  PgrSampleSubsetIndex pssi;
  // work buffers
  uintptr_t* genovec = nullptr;
  uintptr_t* missingness = nullptr;
  uintptr_t* dosage_present = nullptr;
  uint16_t* dosage_main = nullptr;
  uintptr_t* raregeno = nullptr;
  uint32_t* difflist_ids = nullptr;
  // multiallelic view (ref_open_alleles): what a caller derives from the .pvar and hands to pgenlib
  std::vector<uintptr_t> allele_idx_offsets;
  bool open = false;
};

int OpenReader(const char* path, Reader* r, char* err, size_t errlen, uint32_t raw_v = UINT32_MAX, uint32_t raw_n = UINT32_MAX,
               const uint32_t* allele_cts = nullptr) {
  PreinitPgfi(&r->pgfi);
  PreinitPgr(&r->pgr);
  char errstr[kPglErrstrBufBlen];
  PgenHeaderCtrl hc;
  uintptr_t c1 = 0;
  PglErr e = PgfiInitPhase1(path, nullptr, raw_v, raw_n, &hc, &r->pgfi, &c1, errstr);
  if (e != kPglRetSuccess) {
    if (err) snprintf(err, errlen, "phase1: %s", errstr);
    PglErr ce = kPglRetSuccess;
    CleanupPgfi(&r->pgfi, &ce);
    return static_cast<int>(e);
  }
  if (c1) {
    if (cachealigned_malloc(c1 * kCacheline, &r->pgfi_alloc)) return static_cast<int>(kPglRetNomem);
  }
  uint32_t maxw = 0;
  uintptr_t c2 = 0;
  if (allele_cts) {
    // allele counts preloaded by the caller (they live in the .pvar, not in the .pgen): pgfi.allele_idx_offsets +
    // max_allele_ct before phase 2, allele_cts_already_loaded = 1 (pgenlib_read.cc:695-697,1113-1121)
    const uint32_t V = r->pgfi.raw_variant_ct;
    r->allele_idx_offsets.resize(static_cast<size_t>(V) + 1);
    uintptr_t cum = 0;
    uint32_t mx = 2;
    for (uint32_t v = 0; v < V; ++v) {
      r->allele_idx_offsets[v] = cum;
      cum += allele_cts[v];
      if (allele_cts[v] > mx) mx = allele_cts[v];
    }
    r->allele_idx_offsets[V] = cum;
    r->pgfi.allele_idx_offsets = r->allele_idx_offsets.data();
    r->pgfi.max_allele_ct = mx;
  }
  e = PgfiInitPhase2(hc, allele_cts ? 1 : 0, 0, 0, 0, r->pgfi.raw_variant_ct, &maxw, &r->pgfi, r->pgfi_alloc, &c2, errstr);
  if (e != kPglRetSuccess) {
    if (err) snprintf(err, errlen, "phase2: %s", errstr);
    PglErr ce = kPglRetSuccess;
    CleanupPgfi(&r->pgfi, &ce);
    return static_cast<int>(e);
  }
  if (c2) {
    if (cachealigned_malloc(c2 * kCacheline, &r->pgr_alloc)) return static_cast<int>(kPglRetNomem);
  }
  e = PgrInit(path, maxw, &r->pgfi, &r->pgr, r->pgr_alloc);
  if (e != kPglRetSuccess) {
    if (err) snprintf(err, errlen, "PgrInit failed");
    return static_cast<int>(e);
  }
  r->V = r->pgfi.raw_variant_ct;
  r->N = r->pgfi.raw_sample_ct;
  r->sample_ct = r->N;
  PgrClearSampleSubsetIndex(&r->pgr, &r->pssi);
  const uintptr_t gv_bytes = NypCtToAlignedWordCt(r->N) * sizeof(uintptr_t) + 64;
  const uintptr_t bit_bytes = BitCtToAlignedWordCt(r->N) * sizeof(uintptr_t) + 64;
  cachealigned_malloc(gv_bytes, &r->genovec);
  cachealigned_malloc(bit_bytes, &r->missingness);
  cachealigned_malloc(bit_bytes, &r->dosage_present);
  cachealigned_malloc(RoundUpPow2((r->N + 64) * sizeof(uint16_t), kCacheline), &r->dosage_main);
  const uint32_t max_dl = 2 * (r->N / kPglMaxDifflistLenDivisor);
  cachealigned_malloc(RoundUpPow2((max_dl + 3) / 4 + 64, kCacheline), &r->raregeno);
  cachealigned_malloc(RoundUpPow2((max_dl + 2) * sizeof(uint32_t) + 64, kCacheline), &r->difflist_ids);
  r->open = true;
  return 0;
}

void FreeSubset(Reader* r) {
  if (r->sample_include) aligned_free(r->sample_include);
  if (r->interleaved) aligned_free(r->interleaved);
  if (r->cumpop) aligned_free(r->cumpop);
  r->sample_include = nullptr;
  r->interleaved = nullptr;
  r->cumpop = nullptr;
}

// BuildSampleSubset (src/plink_common.cpp:1222-1250).  idx == nullptr clears the subset.
void SetSubset(Reader* r, const uint32_t* idx, uint32_t n) {
  FreeSubset(r);
  if (!idx) {
    r->sample_ct = r->N;
    PgrClearSampleSubsetIndex(&r->pgr, &r->pssi);
    return;
  }
  const uintptr_t aligned_word_ct = BitCtToAlignedWordCt(r->N);
  const uintptr_t bytes = aligned_word_ct * sizeof(uintptr_t);
  cachealigned_malloc(RoundUpPow2(bytes, kCacheline), &r->sample_include);
  cachealigned_malloc(RoundUpPow2(bytes, kCacheline), &r->interleaved);
  memset(r->sample_include, 0, RoundUpPow2(bytes, kCacheline));
  for (uint32_t i = 0; i < n; ++i) SetBit(idx[i], r->sample_include);
  FillInterleavedMaskVec(r->sample_include, aligned_word_ct / kWordsPerVec, r->interleaved);
  const uintptr_t word_ct = DivUp(r->N, static_cast<uint32_t>(kBitsPerWord));
  cachealigned_malloc(RoundUpPow2(word_ct * sizeof(uint32_t), kCacheline), &r->cumpop);
  FillCumulativePopcounts(r->sample_include, word_ct, r->cumpop);
  r->sample_ct = static_cast<uint32_t>(PopcountWords(r->sample_include, word_ct));
  PgrSetSampleSubsetIndex(r->cumpop, &r->pgr, &r->pssi);
}

void CloseReader(Reader* r) {
  if (!r->open) return;
  PglErr ce = kPglRetSuccess;
  CleanupPgr(&r->pgr, &ce);
  ce = kPglRetSuccess;
  CleanupPgfi(&r->pgfi, &ce);
  FreeSubset(r);
  aligned_free_cond(r->pgfi_alloc);
  aligned_free_cond(r->pgr_alloc);
  aligned_free_cond(r->genovec);
  aligned_free_cond(r->missingness);
  aligned_free_cond(r->dosage_present);
  aligned_free_cond(r->dosage_main);
  aligned_free_cond(r->raregeno);
  aligned_free_cond(r->difflist_ids);
  r->open = false;
}

struct Writer {
  STPgenWriter spgw;
  unsigned char* alloc = nullptr;
  uint32_t N = 0;
  // vector-aligned, zero-padded copies of the caller's arrays (pgenlib reads whole vectors)
  uintptr_t* genovec = nullptr;
  uintptr_t* dosage_present = nullptr;
  uint16_t* dosage_main = nullptr;
  std::vector<uintptr_t> allele_idx_offsets;  // multiallelic files: cumulative allele counts, V + 1 entries
};

}  // namespace

extern "C" {

int ref_open(const char* path, void** out, char* err, size_t errlen) {
  Reader* r = new Reader();
  int rc = OpenReader(path, r, err, errlen);
  if (rc) {
    delete r;
    *out = nullptr;
    return rc;
  }
  *out = r;
  return 0;
}

// Same with the variant / sample counts supplied by the caller (from .bim / .fam): PLINK 1 .bed files carry neither
// (PgfiInitPhase1's raw_variant_ct / raw_sample_ct arguments, as src/pfile_reader.cpp:1319 passes them).
int ref_open_counts(const char* path, uint32_t raw_v, uint32_t raw_n, void** out, char* err, size_t errlen) {
  Reader* r = new Reader();
  int rc = OpenReader(path, r, err, errlen, raw_v, raw_n);
  if (rc) {
    delete r;
    *out = nullptr;
    return rc;
  }
  *out = r;
  return 0;
}

// Reader with the caller's allele counts (V entries), for ref_decode_alleles.
int ref_open_alleles(const char* path, const uint32_t* allele_cts, void** out, char* err, size_t errlen) {
  Reader* r = new Reader();
  int rc = OpenReader(path, r, err, errlen, UINT32_MAX, UINT32_MAX, allele_cts);
  if (rc) {
    delete r;
    *out = nullptr;
    return rc;
  }
  *out = r;
  return 0;
}

// PgrGetM + PglMultiallelicSparseToDenseMiss for variants [v0, v1) under the reader's sample subset:
// out[(row * sample_ct + s) * 2 + {0, 1}] = allele codes, 255 = missing (pgenlib_read.cc:6668-6681,
// pgenlib_misc.cc:3552-3578).
int ref_decode_alleles(void* h, uint32_t v0, uint32_t v1, uint8_t* out) {
  Reader* r = static_cast<Reader*>(h);
  if (!r->pgfi.allele_idx_offsets) return static_cast<int>(kPglRetImproperFunctionCall);
  const uint32_t n = r->sample_ct;
  const uintptr_t bit_bytes = BitCtToAlignedWordCt(r->N) * sizeof(uintptr_t) + 64;
  PgenVariant pgv;
  memset(&pgv, 0, sizeof(pgv));
  unsigned char *s01 = nullptr, *s10 = nullptr, *v01 = nullptr, *v10 = nullptr, *wide = nullptr;
  cachealigned_malloc(RoundUpPow2(bit_bytes, kCacheline), &s01);
  cachealigned_malloc(RoundUpPow2(bit_bytes, kCacheline), &s10);
  cachealigned_malloc(RoundUpPow2(r->N + 64, kCacheline), &v01);
  cachealigned_malloc(RoundUpPow2(2 * static_cast<uintptr_t>(r->N) + 64, kCacheline), &v10);
  cachealigned_malloc(RoundUpPow2(2 * static_cast<uintptr_t>(r->N) + 256, kCacheline), &wide);
  pgv.genovec = r->genovec;
  pgv.patch_01_set = reinterpret_cast<uintptr_t*>(s01);
  pgv.patch_10_set = reinterpret_cast<uintptr_t*>(s10);
  pgv.patch_01_vals = reinterpret_cast<AlleleCode*>(v01);
  pgv.patch_10_vals = reinterpret_cast<AlleleCode*>(v10);
  int rc = 0;
  for (uint32_t v = v0; v < v1 && !rc; ++v) {
    PglErr e = PgrGetM(r->sample_include, r->pssi, n, v, &r->pgr, &pgv);
    if (e != kPglRetSuccess) {
      rc = static_cast<int>(e);
      break;
    }
    ZeroTrailingNyps(n, pgv.genovec);
    PglMultiallelicSparseToDenseMiss(&pgv, n, reinterpret_cast<AlleleCode*>(wide));
    memcpy(out + static_cast<size_t>(v - v0) * n * 2, wide, static_cast<size_t>(n) * 2);
  }
  aligned_free(s01);
  aligned_free(s10);
  aligned_free(v01);
  aligned_free(v10);
  aligned_free(wide);
  return rc;
}

void ref_close(void* h) {
  Reader* r = static_cast<Reader*>(h);
  if (!r) return;
  CloseReader(r);
  delete r;
}

void ref_info(void* h, uint32_t* V, uint32_t* N, uint32_t* sample_ct) {
  Reader* r = static_cast<Reader*>(h);
  *V = r->V;
  *N = r->N;
  *sample_ct = r->sample_ct;
}

uint32_t ref_vrtype(void* h, uint32_t vidx) {
  Reader* r = static_cast<Reader*>(h);
  return GetPgfiVrtype(&r->pgfi, vidx);
}

uint64_t ref_fpos(void* h, uint32_t vidx) {
  Reader* r = static_cast<Reader*>(h);
  return GetPgfiFpos(&r->pgfi, vidx);
}

void ref_set_subset(void* h, const uint32_t* idx, uint32_t n) { SetSubset(static_cast<Reader*>(h), idx, n); }

// PgrGetCounts over [v0, v1): out[4*(v-v0) + {0,1,2,3}] = hom_ref, het, hom_alt, missing.
int ref_counts(void* h, uint32_t v0, uint32_t v1, uint32_t* out) {
  Reader* r = static_cast<Reader*>(h);
  for (uint32_t v = v0; v < v1; ++v) {
    STD_ARRAY_DECL(uint32_t, 4, gc);
    PglErr e = PgrGetCounts(r->sample_include, r->interleaved, r->pssi, r->sample_ct, v, &r->pgr, gc);
    if (e != kPglRetSuccess) return static_cast<int>(e);
    memcpy(&out[4 * static_cast<size_t>(v - v0)], &gc[0], 16);
  }
  return 0;
}

// PgrGetDCounts (is_minimac3_r2 = 0, as src/plink_freq.cpp:475-477).
int ref_dcounts(void* h, uint32_t v0, uint32_t v1, uint32_t* counts, uint64_t* all_dosages, double* imp_r2) {
  Reader* r = static_cast<Reader*>(h);
  for (uint32_t v = v0; v < v1; ++v) {
    STD_ARRAY_DECL(uint32_t, 4, gc);
    uint64_t ad[2] = {0, 0};
    double r2 = 0.0;
    PglErr e = PgrGetDCounts(r->sample_include, r->interleaved, r->pssi, r->sample_ct, v, 0, &r->pgr, &r2, gc, ad);
    if (e != kPglRetSuccess) return static_cast<int>(e);
    const size_t k = v - v0;
    memcpy(&counts[4 * k], &gc[0], 16);
    all_dosages[2 * k] = ad[0];
    all_dosages[2 * k + 1] = ad[1];
    imp_r2[k] = r2;
  }
  return 0;
}

// PgrGet + GenoarrToBytesMinus9: out is (v1-v0) x sample_ct int8, -9 = missing.
int ref_decode_i8(void* h, uint32_t v0, uint32_t v1, int8_t* out) {
  Reader* r = static_cast<Reader*>(h);
  for (uint32_t v = v0; v < v1; ++v) {
    PglErr e = PgrGet(r->sample_include, r->pssi, r->sample_ct, v, &r->pgr, r->genovec);
    if (e != kPglRetSuccess) return static_cast<int>(e);
    GenoarrToBytesMinus9(r->genovec, r->sample_ct, &out[static_cast<size_t>(v - v0) * r->sample_ct]);
  }
  return 0;
}

// PgrGetP + UnpackPhasedGenotypes (src/plink_common.cpp:1549-1584; call sites src/pgen_reader.cpp:705-733,
// src/pfile_reader.cpp:1722-1745): out is (v1-v0) x sample_ct x 2 int8 allele pairs.  hom-ref 0,0;
// hom-alt 1,1; missing -9,-9; het 1,0 when phasepresent and phaseinfo are both set, else 0,1.  The phase
// buffers are zeroed per variant because PgrGetP leaves them untouched for variants without a phase track.
int ref_decode_phased(void* h, uint32_t v0, uint32_t v1, int8_t* out) {
  Reader* r = static_cast<Reader*>(h);
  const uintptr_t words = BitCtToAlignedWordCt(r->N) + 8;
  std::vector<uintptr_t> pp(words), pi(words);
  std::vector<int8_t> bytes(static_cast<size_t>(r->sample_ct) + 64);
  for (uint32_t v = v0; v < v1; ++v) {
    std::fill(pp.begin(), pp.end(), 0);
    std::fill(pi.begin(), pi.end(), 0);
    uint32_t phasepresent_ct = 0;
    PglErr e = PgrGetP(r->sample_include, r->pssi, r->sample_ct, v, &r->pgr, r->genovec, pp.data(), pi.data(),
                       &phasepresent_ct);
    if (e != kPglRetSuccess) return static_cast<int>(e);
    GenoarrToBytesMinus9(r->genovec, r->sample_ct, bytes.data());
    int8_t* o = &out[static_cast<size_t>(v - v0) * r->sample_ct * 2];
    for (uint32_t s = 0; s < r->sample_ct; ++s) {
      int8_t a = -9, b = -9;
      switch (bytes[s]) {
        case 0: a = 0; b = 0; break;
        case 2: a = 1; b = 1; break;
        case 1:
          if (IsSet(pp.data(), s) && IsSet(pi.data(), s)) { a = 1; b = 0; } else { a = 0; b = 1; }
          break;
        default: break;
      }
      o[2 * s] = a;
      o[2 * s + 1] = b;
    }
  }
  return 0;
}

// PgrGetMissingness + PopcountWords (variant mode) and the ctz bit walk (sample mode).
// Either output may be NULL.  per_sample must be zero-initialised by the caller.
int ref_missingness(void* h, uint32_t v0, uint32_t v1, uint32_t* per_variant, uint32_t* per_sample) {
  Reader* r = static_cast<Reader*>(h);
  const uintptr_t word_ct = BitCtToWordCt(r->sample_ct);
  for (uint32_t v = v0; v < v1; ++v) {
    PglErr e = PgrGetMissingness(r->sample_include, r->pssi, r->sample_ct, v, &r->pgr, r->missingness, r->genovec);
    if (e != kPglRetSuccess) return static_cast<int>(e);
    if (r->sample_ct % kBitsPerWord) {
      r->missingness[word_ct - 1] = bzhi(r->missingness[word_ct - 1], r->sample_ct % kBitsPerWord);
    }
    if (per_variant) per_variant[v - v0] = static_cast<uint32_t>(PopcountWords(r->missingness, word_ct));
    if (per_sample) {
      for (uintptr_t w = 0; w < word_ct; ++w) {
        uintptr_t bits = r->missingness[w];
        while (bits) {
          const uint32_t s = static_cast<uint32_t>(w * kBitsPerWord + ctzw(bits));
          if (s < r->sample_ct) per_sample[s]++;
          bits &= bits - 1;
        }
      }
    }
  }
  return 0;
}

// PgrGetD + Dosage16ToDoublesMinus9 for ONE variant: out[sample_ct], -9.0 = missing.
int ref_dosage(void* h, uint32_t vidx, double* out, uint32_t* dosage_ct_out) {
  Reader* r = static_cast<Reader*>(h);
  uint32_t dosage_ct = 0;
  PglErr e = PgrGetD(r->sample_include, r->pssi, r->sample_ct, vidx, &r->pgr, r->genovec, r->dosage_present,
                     r->dosage_main, &dosage_ct);
  if (e != kPglRetSuccess) return static_cast<int>(e);
  Dosage16ToDoublesMinus9(r->genovec, r->dosage_present, r->dosage_main, r->sample_ct, dosage_ct, out);
  if (dosage_ct_out) *dosage_ct_out = dosage_ct;
  return 0;
}

// The plink_score Phase-1 loop (src/plink_score.cpp:582-653), single thread, one weight
// column.  mode: 0 = mean-impute (default), 1 = no_mean_imputation, 2 = center.
// Accumulators (sample_ct each) are ADDED to, so K columns = K calls on zeroed arrays.
int ref_score(void* h, const uint32_t* vidx, const uint8_t* flip, const double* weight, uint32_t n_scored,
              uint32_t mode, double* score_sums, double* dosage_sums, uint32_t* allele_cts) {
  Reader* r = static_cast<Reader*>(h);
  const uint32_t sample_ct = r->sample_ct;
  std::vector<double> d(sample_ct);
  for (uint32_t si = 0; si < n_scored; ++si) {
    int rc = ref_dosage(h, vidx[si], d.data(), nullptr);
    if (rc) return rc;
    const bool fl = flip && flip[si];
    const double w = weight[si];
    double sum_alt = 0.0;
    uint32_t non_missing_ct = 0;
    for (uint32_t s = 0; s < sample_ct; ++s) {
      if (d[s] != -9.0) {
        sum_alt += d[s];
        non_missing_ct++;
      }
    }
    if (non_missing_ct == 0) continue;
    if (mode == 2) {
      const double mean_alt = sum_alt / static_cast<double>(non_missing_ct);
      const double freq = mean_alt / 2.0;
      const double sd = std::sqrt(2.0 * freq * (1.0 - freq));
      if (sd == 0.0) continue;
      const double mean_scored = fl ? (2.0 - mean_alt) : mean_alt;
      for (uint32_t s = 0; s < sample_ct; ++s) {
        if (d[s] == -9.0) continue;
        const double scored = fl ? (2.0 - d[s]) : d[s];
        const double standardized = (scored - mean_scored) / sd;
        score_sums[s] += w * standardized;
        allele_cts[s] += 2;
      }
    } else if (mode == 1) {
      for (uint32_t s = 0; s < sample_ct; ++s) {
        if (d[s] == -9.0) continue;
        const double scored = fl ? (2.0 - d[s]) : d[s];
        score_sums[s] += w * scored;
        dosage_sums[s] += scored;
        allele_cts[s] += 2;
      }
    } else {
      const double mean_alt = sum_alt / static_cast<double>(non_missing_ct);
      for (uint32_t s = 0; s < sample_ct; ++s) {
        const double alt = (d[s] == -9.0) ? mean_alt : d[s];
        const double scored = fl ? (2.0 - alt) : alt;
        score_sums[s] += w * scored;
        dosage_sums[s] += scored;
        allele_cts[s] += 2;
      }
    }
  }
  return 0;
}

// read_pfile(orient := 'sample', genotypes := 'counts') Phase-1 accumulation
// (src/pfile_reader.cpp:3322-3428).  use_sparse selects the PgrGetDifflistOrGenovec path.
int ref_sample_counts(void* h, uint32_t v0, uint32_t v1, uint32_t use_sparse, uint32_t* het, uint32_t* hom_alt,
                      uint32_t* missing) {
  Reader* r = static_cast<Reader*>(h);
  const uint32_t sample_ct = r->sample_ct;
  std::vector<int8_t> b(sample_ct + 64);
  const uint32_t max_dl = r->sample_ct / kPglMaxDifflistLenDivisor;
  for (uint32_t v = v0; v < v1; ++v) {
    bool dense = true;
    if (use_sparse) {
      uint32_t common_geno = 0, dl = 0;
      PglErr e = PgrGetDifflistOrGenovec(r->sample_include, r->pssi, sample_ct, max_dl, v, &r->pgr, r->genovec,
                                         &common_geno, r->raregeno, r->difflist_ids, &dl);
      if (e != kPglRetSuccess) return static_cast<int>(e);
      if (common_geno == 0) {
        dense = false;
        for (uint32_t j = 0; j < dl; ++j) {
          const uint32_t s = r->difflist_ids[j];
          const uint32_t g = (r->raregeno[j / kBitsPerWordD2] >> (2 * (j % kBitsPerWordD2))) & 3;
          if (g == 1) het[s]++;
          else if (g == 2) hom_alt[s]++;
          else if (g == 3) missing[s]++;
        }
      } else if (common_geno != UINT32_MAX) {
        e = PgrGet(r->sample_include, r->pssi, sample_ct, v, &r->pgr, r->genovec);
        if (e != kPglRetSuccess) return static_cast<int>(e);
      }
    } else {
      PglErr e = PgrGet(r->sample_include, r->pssi, sample_ct, v, &r->pgr, r->genovec);
      if (e != kPglRetSuccess) return static_cast<int>(e);
    }
    if (dense) {
      GenoarrToBytesMinus9(r->genovec, sample_ct, b.data());
      for (uint32_t s = 0; s < sample_ct; ++s) {
        const int8_t g = b[s];
        if (g == 1) het[s]++;
        else if (g == 2) hom_alt[s]++;
        else if (g != 0) missing[s]++;
      }
    }
  }
  return 0;
}

double ref_hwe_lnp(int32_t hets, int32_t hom1, int32_t hom2, uint32_t midp) { return HweLnP(hets, hom1, hom2, midp); }

double ref_hwe_xchr_lnp(int32_t fhets, int32_t fhom1, int32_t fhom2, int32_t male1, int32_t male2, uint32_t midp) {
  return HweXchrLnP(fhets, fhom1, fhom2, male1, male2, midp);
}

void ref_hwe_lnp_batch(const uint32_t* counts /* n x 4 */, uint32_t n, uint32_t midp, double* out_lnp) {
  for (uint32_t i = 0; i < n; ++i) {
    out_lnp[i] = HweLnP(static_cast<int32_t>(counts[4 * i + 1]), static_cast<int32_t>(counts[4 * i]),
                        static_cast<int32_t>(counts[4 * i + 2]), midp);
  }
}

// Multi-threaded scan = the reference's thread model without DuckDB: `threads` readers, each
// claiming 128-variant batches with fetch_add (src/plink_freq.cpp:434-441).
//   kind 0: PgrGetCounts                     (plink_freq / plink_hardy counts)
//   kind 1: PgrGetCounts + HweLnP            (plink_hardy)
//   kind 2: PgrGetMissingness + popcount     (plink_missing variant mode)
//   kind 3: PgrGetMissingness + ctz walk     (plink_missing sample mode, thread-local accumulators)
//   kind 4: PgrGet + GenoarrToBytesMinus9    (read_pfile decode)
// Returns wall seconds of the scan phase (readers opened before the clock starts);
// *checksum receives a sum over all outputs so the work cannot be optimised away.
double ref_bench_scan(const char* path, uint32_t threads, uint32_t v0, uint32_t v1, uint32_t kind,
                      uint64_t* checksum) {
  std::vector<Reader> readers(threads);
  for (uint32_t t = 0; t < threads; ++t) {
    char err[256];
    if (OpenReader(path, &readers[t], err, sizeof(err))) {
      for (uint32_t u = 0; u < t; ++u) CloseReader(&readers[u]);
      return -1.0;
    }
  }
  if (v1 > readers[0].V) v1 = readers[0].V;
  std::atomic<uint32_t> next(v0);
  std::atomic<uint64_t> total(0);
  auto worker = [&](uint32_t t) {
    Reader* r = &readers[t];
    uint64_t acc = 0;
    std::vector<uint32_t> per_sample;
    std::vector<int8_t> bytes;
    if (kind == 3) per_sample.assign(r->N, 0);
    if (kind == 4) bytes.resize(r->N + 64);
    const uintptr_t word_ct = BitCtToWordCt(r->N);
    while (true) {
      const uint32_t b0 = next.fetch_add(128);
      if (b0 >= v1) break;
      const uint32_t b1 = (b0 + 128 < v1) ? b0 + 128 : v1;
      for (uint32_t v = b0; v < b1; ++v) {
        if (kind <= 1) {
          STD_ARRAY_DECL(uint32_t, 4, gc);
          PgrGetCounts(nullptr, nullptr, r->pssi, r->N, v, &r->pgr, gc);
          acc += gc[0] + 3ull * gc[1] + 5ull * gc[2] + 7ull * gc[3];
          if (kind == 1) {
            const double lnp = HweLnP(static_cast<int32_t>(gc[1]), static_cast<int32_t>(gc[0]),
                                      static_cast<int32_t>(gc[2]), 0);
            acc += (lnp < -1.0);
          }
        } else if (kind <= 3) {
          PgrGetMissingness(nullptr, r->pssi, r->N, v, &r->pgr, r->missingness, r->genovec);
          if (kind == 2) {
            acc += PopcountWords(r->missingness, word_ct);
          } else {
            for (uintptr_t w = 0; w < word_ct; ++w) {
              uintptr_t bits = r->missingness[w];
              while (bits) {
                const uint32_t s = static_cast<uint32_t>(w * kBitsPerWord + ctzw(bits));
                if (s < r->N) per_sample[s]++;
                bits &= bits - 1;
              }
            }
          }
        } else {
          PgrGet(nullptr, r->pssi, r->N, v, &r->pgr, r->genovec);
          GenoarrToBytesMinus9(r->genovec, r->N, bytes.data());
          acc += static_cast<uint8_t>(bytes[v % r->N]);
        }
      }
    }
    if (kind == 3) {
      for (uint32_t s = 0; s < r->N; ++s) acc += per_sample[s];
    }
    total.fetch_add(acc);
  };
  const auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> pool;
  for (uint32_t t = 1; t < threads; ++t) pool.emplace_back(worker, t);
  worker(0);
  for (auto& th : pool) th.join();
  const auto t1 = std::chrono::steady_clock::now();
  for (uint32_t t = 0; t < threads; ++t) CloseReader(&readers[t]);
  if (checksum) *checksum = total.load();
  return std::chrono::duration<double>(t1 - t0).count();
}

// Multi-threaded plink_score over K weight columns.  The reference scores ONE weight vector per
// call (src/plink_score.cpp:329-427), so the K-column baseline shares the decode per variant and
// runs the K=1 accumulation loop K times (the generous reading of BASELINE.md).  Thread-local
// accumulators merged at the end as src/plink_score.cpp:657-664.  weights is n_scored x K.
double ref_bench_score(const char* path, uint32_t threads, const uint32_t* vidx, const uint8_t* flip,
                       const double* weights, uint32_t n_scored, uint32_t K, double* score_sums /* N x K */,
                       uint32_t* allele_cts /* N */) {
  std::vector<Reader> readers(threads);
  for (uint32_t t = 0; t < threads; ++t) {
    char err[256];
    if (OpenReader(path, &readers[t], err, sizeof(err))) return -1.0;
  }
  const uint32_t N = readers[0].N;
  std::atomic<uint32_t> next(0);
  std::vector<std::vector<double>> local(threads);
  std::vector<std::vector<uint32_t>> local_ct(threads);
  auto worker = [&](uint32_t t) {
    Reader* r = &readers[t];
    local[t].assign(static_cast<size_t>(N) * K, 0.0);
    local_ct[t].assign(N, 0);
    std::vector<double> d(N);
    while (true) {
      const uint32_t b0 = next.fetch_add(16);
      if (b0 >= n_scored) break;
      const uint32_t b1 = (b0 + 16 < n_scored) ? b0 + 16 : n_scored;
      for (uint32_t si = b0; si < b1; ++si) {
        uint32_t dosage_ct = 0;
        PgrGetD(nullptr, r->pssi, N, vidx[si], &r->pgr, r->genovec, r->dosage_present, r->dosage_main, &dosage_ct);
        Dosage16ToDoublesMinus9(r->genovec, r->dosage_present, r->dosage_main, N, dosage_ct, d.data());
        double sum_alt = 0.0;
        uint32_t nm = 0;
        for (uint32_t s = 0; s < N; ++s) {
          if (d[s] != -9.0) {
            sum_alt += d[s];
            nm++;
          }
        }
        if (!nm) continue;
        const double mean_alt = sum_alt / static_cast<double>(nm);
        const bool fl = flip && flip[si];
        for (uint32_t k = 0; k < K; ++k) {
          const double w = weights[static_cast<size_t>(si) * K + k];
          double* acc = &local[t][static_cast<size_t>(k) * N];
          for (uint32_t s = 0; s < N; ++s) {
            const double alt = (d[s] == -9.0) ? mean_alt : d[s];
            const double scored = fl ? (2.0 - alt) : alt;
            acc[s] += w * scored;
          }
        }
        for (uint32_t s = 0; s < N; ++s) local_ct[t][s] += 2;
      }
    }
  };
  const auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> pool;
  for (uint32_t t = 1; t < threads; ++t) pool.emplace_back(worker, t);
  worker(0);
  for (auto& th : pool) th.join();
  for (uint32_t t = 0; t < threads; ++t) {
    for (uint32_t k = 0; k < K; ++k)
      for (uint32_t s = 0; s < N; ++s) score_sums[static_cast<size_t>(s) * K + k] += local[t][static_cast<size_t>(k) * N + s];
    for (uint32_t s = 0; s < N; ++s) allele_cts[s] += local_ct[t][s];
  }
  const auto t1 = std::chrono::steady_clock::now();
  for (uint32_t t = 0; t < threads; ++t) CloseReader(&readers[t]);
  return std::chrono::duration<double>(t1 - t0).count();
}

// ---- plink_ld: ComputeLdStats restated (src/plink_ld.cpp:52-135; a static function of the extension, so it cannot
// be linked): PgrGet for both variants (ReadGenovec, :564-581), then the per-sample loop and the closed forms.
// out2 = {r2, d_prime}.
int ref_ld_pair(void* h, uint32_t va, uint32_t vb, double* out2, uint32_t* obs_ct, uint32_t* is_valid) {
  Reader* r = static_cast<Reader*>(h);
  const uint32_t sample_ct = r->sample_ct;
  const uintptr_t words = NypCtToAlignedWordCt(r->N) + 8;
  std::vector<uintptr_t> ga(words, 0), gb(words, 0);
  PglErr e = PgrGet(r->sample_include, r->pssi, sample_ct, va, &r->pgr, ga.data());
  if (e != kPglRetSuccess) return static_cast<int>(e);
  e = PgrGet(r->sample_include, r->pssi, sample_ct, vb, &r->pgr, gb.data());
  if (e != kPglRetSuccess) return static_cast<int>(e);
  double sum_a = 0, sum_b = 0, sum_ab = 0, sum_a2 = 0, sum_b2 = 0;
  uint32_t n = 0;
  const uint32_t word_ct = DivUp(sample_ct, kBitsPerWordD2);
  for (uint32_t widx = 0; widx < word_ct; widx++) {
    uintptr_t word_a = ga[widx];
    uintptr_t word_b = gb[widx];
    const uint32_t samples_remaining = sample_ct - widx * kBitsPerWordD2;
    const uint32_t samples_in_word = samples_remaining < kBitsPerWordD2 ? samples_remaining : kBitsPerWordD2;
    for (uint32_t sidx = 0; sidx < samples_in_word; sidx++) {
      const uint32_t geno_a = word_a & 3;
      const uint32_t geno_b = word_b & 3;
      word_a >>= 2;
      word_b >>= 2;
      if (geno_a == 3 || geno_b == 3) continue;
      const double a = static_cast<double>(geno_a), b = static_cast<double>(geno_b);
      sum_a += a;
      sum_b += b;
      sum_ab += a * b;
      sum_a2 += a * a;
      sum_b2 += b * b;
      n++;
    }
  }
  *obs_ct = n;
  *is_valid = 0;
  out2[0] = out2[1] = 0;
  if (n < 2) return 0;
  const double dn = static_cast<double>(n);
  const double mean_a = sum_a / dn, mean_b = sum_b / dn;
  const double cov_ab = sum_ab / dn - mean_a * mean_b;
  const double var_a = sum_a2 / dn - mean_a * mean_a;
  const double var_b = sum_b2 / dn - mean_b * mean_b;
  if (var_a < 1e-15 || var_b < 1e-15) return 0;
  *is_valid = 1;
  out2[0] = (cov_ab * cov_ab) / (var_a * var_b);
  const double D = cov_ab / 4.0;
  const double p_a = sum_a / (2.0 * dn), p_b = sum_b / (2.0 * dn);
  double D_max;
  if (D >= 0) {
    D_max = std::min(p_a * (1.0 - p_b), (1.0 - p_a) * p_b);
  } else {
    D_max = std::max(-p_a * p_b, -(1.0 - p_a) * (1.0 - p_b));
  }
  out2[1] = std::abs(D_max) < 1e-15 ? 0.0 : D / D_max;
  return 0;
}

// ---- writer (fixture / synthetic fileset generator; pgenlib picks the record type) ----
int ref_writer_open(const char* path, uint32_t V, uint32_t N, uint32_t dosage, void** out) {
  Writer* w = new Writer();
  PreinitSpgw(&w->spgw);
  uintptr_t alloc_ct = 0;
  uint32_t max_vrec_len = 0;
  PglErr e = SpgwInitPhase1(path, nullptr, nullptr, V, N, 0, kPgenWriteBackwardSeek,
                            ((dosage & 1) ? kfPgenGlobalDosagePresent : kfPgenGlobal0) |
                                ((dosage & 2) ? kfPgenGlobalHardcallPhasePresent : kfPgenGlobal0),
                            2, &w->spgw, &alloc_ct, &max_vrec_len);
  if (e != kPglRetSuccess) {
    delete w;
    return static_cast<int>(e);
  }
  if (cachealigned_malloc(alloc_ct * kCacheline, &w->alloc)) {
    delete w;
    return static_cast<int>(kPglRetNomem);
  }
  SpgwInitPhase2(max_vrec_len, &w->spgw, w->alloc);
  w->N = N;
  cachealigned_malloc(RoundUpPow2(NypCtToAlignedWordCt(N) * sizeof(uintptr_t) + 64, kCacheline), &w->genovec);
  cachealigned_malloc(RoundUpPow2(BitCtToAlignedWordCt(N) * sizeof(uintptr_t) + 64, kCacheline), &w->dosage_present);
  cachealigned_malloc(RoundUpPow2((static_cast<uintptr_t>(N) + 64) * sizeof(uint16_t), kCacheline), &w->dosage_main);
  *out = w;
  return 0;
}

static void StageGenovec(Writer* w, const uint64_t* genovec) {
  const uintptr_t aligned_words = NypCtToAlignedWordCt(w->N);
  memset(w->genovec, 0, aligned_words * sizeof(uintptr_t));
  memcpy(w->genovec, genovec, NypCtToWordCt(w->N) * sizeof(uintptr_t));
  ZeroTrailingNyps(w->N, w->genovec);
}

// genovec: ceil(N/32) 64-bit words, trailing bits zero.
int ref_writer_append(void* h, const uint64_t* genovec) {
  Writer* w = static_cast<Writer*>(h);
  StageGenovec(w, genovec);
  return static_cast<int>(SpgwAppendBiallelicGenovec(w->genovec, &w->spgw));
}

int ref_writer_append_dosage(void* h, const uint64_t* genovec, const uint64_t* dosage_present,
                             const uint16_t* dosage_main, uint32_t dosage_ct) {
  Writer* w = static_cast<Writer*>(h);
  StageGenovec(w, genovec);
  const uintptr_t aligned_words = BitCtToAlignedWordCt(w->N);
  memset(w->dosage_present, 0, aligned_words * sizeof(uintptr_t));
  memcpy(w->dosage_present, dosage_present, BitCtToWordCt(w->N) * sizeof(uintptr_t));
  memcpy(w->dosage_main, dosage_main, static_cast<size_t>(dosage_ct) * sizeof(uint16_t));
  return static_cast<int>(SpgwAppendBiallelicGenovecDosage16(w->genovec, w->dosage_present, w->dosage_main, dosage_ct,
                                                             &w->spgw));
}

// phasepresent / phaseinfo: ceil(N/64) 64-bit words each (writer opened with flag bit 1).
int ref_writer_append_phased(void* h, const uint64_t* genovec, const uint64_t* phasepresent, const uint64_t* phaseinfo) {
  Writer* w = static_cast<Writer*>(h);
  StageGenovec(w, genovec);
  const uintptr_t aligned_words = BitCtToAlignedWordCt(w->N);
  std::vector<uintptr_t> pp(aligned_words + 8, 0), pi(aligned_words + 8, 0);
  memcpy(pp.data(), phasepresent, BitCtToWordCt(w->N) * sizeof(uintptr_t));
  memcpy(pi.data(), phaseinfo, BitCtToWordCt(w->N) * sizeof(uintptr_t));
  return static_cast<int>(SpgwAppendBiallelicGenovecHphase(w->genovec, pp.data(), pi.data(), &w->spgw));
}

// Writer for files with multiallelic variants and / or dosage-phase tracks.  flags: 1 dosage, 2 hardcall phase,
// 4 dosage phase.  allele_cts: V allele counts (>= 2), or NULL for an all-biallelic file.
int ref_writer_open_ex(const char* path, uint32_t V, uint32_t N, uint32_t flags, const uint32_t* allele_cts, void** out) {
  Writer* w = new Writer();
  PreinitSpgw(&w->spgw);
  uint32_t max_allele_ct = 2;
  if (allele_cts) {
    w->allele_idx_offsets.resize(static_cast<size_t>(V) + 1);
    uintptr_t cum = 0;
    for (uint32_t v = 0; v < V; ++v) {
      w->allele_idx_offsets[v] = cum;
      cum += allele_cts[v];
      if (allele_cts[v] > max_allele_ct) max_allele_ct = allele_cts[v];
    }
    w->allele_idx_offsets[V] = cum;
  }
  uintptr_t alloc_ct = 0;
  uint32_t max_vrec_len = 0;
  PgenGlobalFlags gf = kfPgenGlobal0;
  if (flags & 1) gf |= kfPgenGlobalDosagePresent;
  if (flags & 2) gf |= kfPgenGlobalHardcallPhasePresent;
  if (flags & 4) gf |= kfPgenGlobalDosagePresent | kfPgenGlobalDosagePhasePresent | kfPgenGlobalHardcallPhasePresent;
  PglErr e = SpgwInitPhase1(path, allele_cts ? w->allele_idx_offsets.data() : nullptr, nullptr, V, N, max_allele_ct,
                            kPgenWriteBackwardSeek, gf, 2, &w->spgw, &alloc_ct, &max_vrec_len);
  if (e != kPglRetSuccess) {
    delete w;
    return static_cast<int>(e);
  }
  if (cachealigned_malloc(alloc_ct * kCacheline, &w->alloc)) {
    delete w;
    return static_cast<int>(kPglRetNomem);
  }
  SpgwInitPhase2(max_vrec_len, &w->spgw, w->alloc);
  w->N = N;
  cachealigned_malloc(RoundUpPow2(NypCtToAlignedWordCt(N) * sizeof(uintptr_t) + 64, kCacheline), &w->genovec);
  cachealigned_malloc(RoundUpPow2(BitCtToAlignedWordCt(N) * sizeof(uintptr_t) + 64, kCacheline), &w->dosage_present);
  cachealigned_malloc(RoundUpPow2((static_cast<uintptr_t>(N) + 64) * sizeof(uint16_t), kCacheline), &w->dosage_main);
  *out = w;
  return 0;
}

// Multiallelic hardcalls (SpgwAppendMultiallelicSparse): the main track holds 0 = REF/REF, 1 = REF/ALTx,
// 2 = ALTx/ALTy, 3 = missing; patch_01 names the code-1 samples whose ALT allele is not ALT1 (one allele code
// each), patch_10 the code-2 samples that are not ALT1/ALT1 (two allele codes each, low <= high).
int ref_writer_append_multiallelic(void* h, const uint64_t* genovec, const uint64_t* patch_01_set,
                                   const uint8_t* patch_01_vals, uint32_t patch_01_ct, const uint64_t* patch_10_set,
                                   const uint8_t* patch_10_vals, uint32_t patch_10_ct, uint32_t allele_ct) {
  Writer* w = static_cast<Writer*>(h);
  StageGenovec(w, genovec);
  const uintptr_t aligned_words = BitCtToAlignedWordCt(w->N);
  std::vector<uintptr_t> s01(aligned_words + 8, 0), s10(aligned_words + 8, 0);
  memcpy(s01.data(), patch_01_set, BitCtToWordCt(w->N) * sizeof(uintptr_t));
  memcpy(s10.data(), patch_10_set, BitCtToWordCt(w->N) * sizeof(uintptr_t));
  std::vector<AlleleCode> v01(patch_01_ct + 64, 0), v10(2 * static_cast<size_t>(patch_10_ct) + 64, 0);
  for (uint32_t i = 0; i < patch_01_ct; ++i) v01[i] = patch_01_vals[i];
  for (uint32_t i = 0; i < 2 * patch_10_ct; ++i) v10[i] = patch_10_vals[i];
  return static_cast<int>(SpgwAppendMultiallelicSparse(w->genovec, s01.data(), v01.data(), s10.data(), v10.data(), allele_ct,
                                                       patch_01_ct, patch_10_ct, &w->spgw));
}

// Multiallelic hardcalls with a hardcall-phase track (SpgwAppendMultiallelicGenovecHphase; writer opened with flag 2
// and allele counts).  phasepresent may only be set at hets of the full genotype (01 calls and ALTx/ALTy, x != y).
int ref_writer_append_multiallelic_hphase(void* h, const uint64_t* genovec, const uint64_t* patch_01_set,
                                          const uint8_t* patch_01_vals, uint32_t patch_01_ct, const uint64_t* patch_10_set,
                                          const uint8_t* patch_10_vals, uint32_t patch_10_ct, const uint64_t* phasepresent,
                                          const uint64_t* phaseinfo, uint32_t allele_ct) {
  Writer* w = static_cast<Writer*>(h);
  StageGenovec(w, genovec);
  const uintptr_t aligned_words = BitCtToAlignedWordCt(w->N);
  const size_t nw = BitCtToWordCt(w->N) * sizeof(uintptr_t);
  std::vector<uintptr_t> s01(aligned_words + 8, 0), s10(aligned_words + 8, 0), pp(aligned_words + 8, 0), pi(aligned_words + 8, 0);
  memcpy(s01.data(), patch_01_set, nw);
  memcpy(s10.data(), patch_10_set, nw);
  memcpy(pp.data(), phasepresent, nw);
  memcpy(pi.data(), phaseinfo, nw);
  std::vector<AlleleCode> v01(patch_01_ct + 64, 0), v10(2 * static_cast<size_t>(patch_10_ct) + 64, 0);
  for (uint32_t i = 0; i < patch_01_ct; ++i) v01[i] = patch_01_vals[i];
  for (uint32_t i = 0; i < 2 * patch_10_ct; ++i) v10[i] = patch_10_vals[i];
  return static_cast<int>(SpgwAppendMultiallelicGenovecHphase(w->genovec, s01.data(), v01.data(), s10.data(), v10.data(), pp.data(),
                                                              pi.data(), allele_ct, patch_01_ct, patch_10_ct, &w->spgw));
}

// Dosages with a dosage-phase track (SpgwAppendBiallelicGenovecDphase16; writer opened with flag 4).
int ref_writer_append_dphase(void* h, const uint64_t* genovec, const uint64_t* phasepresent, const uint64_t* phaseinfo,
                             const uint64_t* dosage_present, const uint64_t* dphase_present, const uint16_t* dosage_main,
                             const int16_t* dphase_delta, uint32_t dosage_ct, uint32_t dphase_ct) {
  Writer* w = static_cast<Writer*>(h);
  StageGenovec(w, genovec);
  const uintptr_t aligned_words = BitCtToAlignedWordCt(w->N);
  const size_t nw = BitCtToWordCt(w->N) * sizeof(uintptr_t);
  std::vector<uintptr_t> pp(aligned_words + 8, 0), pi(aligned_words + 8, 0), dp(aligned_words + 8, 0), dpp(aligned_words + 8, 0);
  memcpy(pp.data(), phasepresent, nw);
  memcpy(pi.data(), phaseinfo, nw);
  memcpy(dp.data(), dosage_present, nw);
  memcpy(dpp.data(), dphase_present, nw);
  std::vector<uint16_t> dm(dosage_ct + 64, 0);
  std::vector<int16_t> dd(dphase_ct + 64, 0);
  memcpy(dm.data(), dosage_main, static_cast<size_t>(dosage_ct) * 2);
  memcpy(dd.data(), dphase_delta, static_cast<size_t>(dphase_ct) * 2);
  return static_cast<int>(SpgwAppendBiallelicGenovecDphase16(w->genovec, pp.data(), pi.data(), dp.data(), dpp.data(), dm.data(),
                                                             dd.data(), dosage_ct, dphase_ct, &w->spgw));
}

int ref_writer_close(void* h) {
  Writer* w = static_cast<Writer*>(h);
  PglErr e = SpgwFinish(&w->spgw);
  PglErr ce = kPglRetSuccess;
  CleanupSpgw(&w->spgw, &ce);
  aligned_free_cond(w->alloc);
  aligned_free_cond(w->genovec);
  aligned_free_cond(w->dosage_present);
  aligned_free_cond(w->dosage_main);
  delete w;
  return static_cast<int>(e);
}

}  // extern "C"
