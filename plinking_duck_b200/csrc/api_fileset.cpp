// pd_fileset: variant-sharded, row-concatenated sources that share one sample set -- read_pfile's multi-file lists
// (src/pfile_reader.cpp:325-400 PfileSource / sources, :1003-1012 the sample-count check, :2133-2151 the
// source-bounded scan batches: "a batch never spans a file boundary").  One pd_ctx per file (each may itself span
// several GPUs, pd_open_opts.n_devices); a global variant index g addresses variant g - offset[f] of file f.
// Per-variant outputs of a range are the concatenation of the files' outputs; per-sample outputs (plink_missing sample
// mode, read_pfile orient := 'sample', plink_score) are the sums over the files, added on the host in file order
// (N or N x K numbers per file -- nothing next to the scans).
//
// Written on top of the public C ABI only (include/plinkgpu.h): what a caller could do by hand, done once.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/plinkgpu.h"

struct pd_fileset {
  std::vector<pd_ctx*> ctx;
  std::vector<uint32_t> off;  // n_files + 1 cumulative variant counts
  uint32_t sample_ct = 0;
  std::string last_error;

  int Fail(int code, const std::string& msg) {
    last_error = msg;
    return code;
  }
  int FailFrom(int code, uint32_t f) {
    last_error = std::string("file ") + std::to_string(f) + ": " + pd_last_error(ctx[f]);
    return code;
  }
  uint32_t EffectiveSamples() const {
    pd_info_t info;
    if (ctx.empty() || pd_info(ctx[0], &info) != PD_OK) return 0;
    return info.effective_sample_ct;
  }
  // Calls fn(file, local v0, local v1, rows before) for every file that [g0, g1) touches.
  template <class Fn>
  int ForRange(uint32_t g0, uint32_t g1, Fn fn) {
    if (g0 > g1 || g1 > off.back()) return Fail(PD_ERR_INVALID_ARG, "variant range out of bounds");
    for (uint32_t f = 0; f + 1 < off.size(); ++f) {
      const uint32_t a = std::max(g0, off[f]), b = std::min(g1, off[f + 1]);
      if (a >= b) continue;
      const int rc = fn(f, a - off[f], b - off[f], a - g0);
      if (rc != PD_OK) return FailFrom(rc, f);
    }
    return PD_OK;
  }
};

namespace {
void CopyErr(char* buf, size_t len, const std::string& s) {
  if (buf && len) snprintf(buf, len, "%s", s.c_str());
}
}  // namespace

extern "C" {

int pd_fileset_open(const char* const* pgen_paths, uint32_t n_files, const pd_open_opts* opts, pd_fileset** out,
                    char* errbuf, size_t errbuf_len) {
  if (!out) return PD_ERR_INVALID_ARG;
  *out = nullptr;
  if (!pgen_paths || !n_files) {
    CopyErr(errbuf, errbuf_len, "pd_fileset_open: an empty file list");
    return PD_ERR_INVALID_ARG;
  }
  pd_fileset* fs = new (std::nothrow) pd_fileset();
  if (!fs) return PD_ERR_NOMEM;
  fs->off.push_back(0);
  for (uint32_t f = 0; f < n_files; ++f) {
    pd_ctx* c = nullptr;
    char eb[512] = {0};
    int rc = pgen_paths[f] ? pd_open(pgen_paths[f], opts, &c, eb, sizeof(eb)) : PD_ERR_INVALID_ARG;
    pd_info_t info;
    if (rc == PD_OK) rc = pd_info(c, &info);
    std::string msg = std::string("pd_fileset_open: ") + (pgen_paths[f] ? pgen_paths[f] : "(null)") + ": " + eb;
    if (rc == PD_OK && f == 0) fs->sample_ct = info.sample_ct;
    if (rc == PD_OK && info.sample_ct != fs->sample_ct) {
      // every shard must carry the same samples (src/pfile_reader.cpp:1003-1012)
      rc = PD_ERR_INVALID_ARG;
      msg = std::string("pd_fileset_open: ") + pgen_paths[f] + " has " + std::to_string(info.sample_ct) +
            " samples, the first file has " + std::to_string(fs->sample_ct);
    }
    if (rc == PD_OK && static_cast<uint64_t>(fs->off.back()) + info.variant_ct > 0xffffffffull) {
      rc = PD_ERR_NOT_SUPPORTED;
      msg = "pd_fileset_open: more than 2^32 - 1 variants in total";
    }
    if (rc != PD_OK) {
      if (c) pd_close(c);
      for (pd_ctx* x : fs->ctx) pd_close(x);
      delete fs;
      CopyErr(errbuf, errbuf_len, msg);
      return rc;
    }
    fs->ctx.push_back(c);
    fs->off.push_back(fs->off.back() + info.variant_ct);
  }
  *out = fs;
  return PD_OK;
}

void pd_fileset_close(pd_fileset* fs) {
  if (!fs) return;
  for (pd_ctx* c : fs->ctx) pd_close(c);
  delete fs;
}

const char* pd_fileset_last_error(const pd_fileset* fs) { return fs ? fs->last_error.c_str() : ""; }
uint32_t pd_fileset_file_count(const pd_fileset* fs) { return fs ? static_cast<uint32_t>(fs->ctx.size()) : 0; }
uint32_t pd_fileset_variant_ct(const pd_fileset* fs) { return fs ? fs->off.back() : 0; }
uint32_t pd_fileset_sample_ct(const pd_fileset* fs) { return fs ? fs->sample_ct : 0; }

int pd_fileset_offsets(const pd_fileset* fs, uint32_t* out) {
  if (!fs || !out) return PD_ERR_INVALID_ARG;
  memcpy(out, fs->off.data(), fs->off.size() * sizeof(uint32_t));
  return PD_OK;
}

pd_ctx* pd_fileset_ctx(pd_fileset* fs, uint32_t file) { return (fs && file < fs->ctx.size()) ? fs->ctx[file] : nullptr; }

int pd_fileset_set_sample_subset(pd_fileset* fs, const uint64_t* sample_include, uint32_t subset_sample_ct) {
  if (!fs) return PD_ERR_INVALID_ARG;
  for (uint32_t f = 0; f < fs->ctx.size(); ++f) {
    const int rc = pd_set_sample_subset(fs->ctx[f], sample_include, subset_sample_ct);
    if (rc != PD_OK) return fs->FailFrom(rc, f);
  }
  return PD_OK;
}

int pd_fileset_counts(pd_fileset* fs, uint32_t g0, uint32_t g1, uint32_t* out_counts) {
  if (!fs || (!out_counts && g1 > g0)) return PD_ERR_INVALID_ARG;
  return fs->ForRange(g0, g1, [&](uint32_t f, uint32_t a, uint32_t b, uint32_t row) {
    return pd_counts(fs->ctx[f], a, b, out_counts + 4ull * row);
  });
}

int pd_fileset_hardy(pd_fileset* fs, uint32_t g0, uint32_t g1, uint32_t midp, uint32_t* out_counts, double* out_p) {
  if (!fs || (!out_p && g1 > g0)) return PD_ERR_INVALID_ARG;
  return fs->ForRange(g0, g1, [&](uint32_t f, uint32_t a, uint32_t b, uint32_t row) {
    return pd_hardy(fs->ctx[f], a, b, midp, out_counts ? out_counts + 4ull * row : nullptr, out_p + row);
  });
}

int pd_fileset_decode_i8(pd_fileset* fs, uint32_t g0, uint32_t g1, int8_t* out) {
  if (!fs || (!out && g1 > g0)) return PD_ERR_INVALID_ARG;
  const uint64_t n = fs->EffectiveSamples();
  return fs->ForRange(g0, g1, [&](uint32_t f, uint32_t a, uint32_t b, uint32_t row) {
    return pd_decode_i8(fs->ctx[f], a, b, out + row * n);
  });
}

int pd_fileset_sample_counts(pd_fileset* fs, uint32_t g0, uint32_t g1, uint32_t* out_het, uint32_t* out_hom_alt,
                             uint32_t* out_missing) {
  if (!fs || !out_het || !out_hom_alt || !out_missing) return PD_ERR_INVALID_ARG;
  const uint32_t n = fs->EffectiveSamples();
  std::fill(out_het, out_het + n, 0u);
  std::fill(out_hom_alt, out_hom_alt + n, 0u);
  std::fill(out_missing, out_missing + n, 0u);
  std::vector<uint32_t> h(n), a2(n), m(n);
  return fs->ForRange(g0, g1, [&](uint32_t f, uint32_t a, uint32_t b, uint32_t) {
    const int rc = pd_sample_counts(fs->ctx[f], a, b, h.data(), a2.data(), m.data());
    if (rc != PD_OK) return rc;
    for (uint32_t s = 0; s < n; ++s) {
      out_het[s] += h[s];
      out_hom_alt[s] += a2[s];
      out_missing[s] += m[s];
    }
    return static_cast<int>(PD_OK);
  });
}

int pd_fileset_missing_by_sample(pd_fileset* fs, uint32_t g0, uint32_t g1, uint32_t* out_missing) {
  if (!fs || !out_missing) return PD_ERR_INVALID_ARG;
  const uint32_t n = fs->EffectiveSamples();
  std::fill(out_missing, out_missing + n, 0u);
  std::vector<uint32_t> m(n);
  return fs->ForRange(g0, g1, [&](uint32_t f, uint32_t a, uint32_t b, uint32_t) {
    const int rc = pd_missing_by_sample(fs->ctx[f], a, b, m.data());
    if (rc != PD_OK) return rc;
    for (uint32_t s = 0; s < n; ++s) out_missing[s] += m[s];
    return static_cast<int>(PD_OK);
  });
}

int pd_fileset_score(pd_fileset* fs, const uint32_t* gvidx, const uint8_t* flip, const double* weights, uint32_t n_var,
                     uint32_t K, uint32_t mode, double* score_sums, double* dosage_sums, uint32_t* allele_cts) {
  if (!fs || !score_sums || !dosage_sums || !allele_cts || !K || (n_var && (!gvidx || !weights))) return PD_ERR_INVALID_ARG;
  for (uint32_t i = 0; i < n_var; ++i) {
    if (gvidx[i] >= fs->off.back() || (i && gvidx[i] <= gvidx[i - 1])) {
      return fs->Fail(PD_ERR_INVALID_ARG, "variant list must be strictly ascending and inside the fileset");
    }
  }
  const uint32_t n = fs->EffectiveSamples();
  std::fill(score_sums, score_sums + static_cast<size_t>(n) * K, 0.0);
  std::fill(dosage_sums, dosage_sums + n, 0.0);
  std::fill(allele_cts, allele_cts + n, 0u);
  std::vector<double> ss(static_cast<size_t>(n) * K), ds(n);
  std::vector<uint32_t> ac(n), local;
  uint32_t i0 = 0;
  for (uint32_t f = 0; f + 1 < fs->off.size() && i0 < n_var; ++f) {
    uint32_t i1 = i0;
    while (i1 < n_var && gvidx[i1] < fs->off[f + 1]) ++i1;
    if (i1 == i0) continue;
    local.resize(i1 - i0);
    for (uint32_t i = i0; i < i1; ++i) local[i - i0] = gvidx[i] - fs->off[f];
    const int rc = pd_score(fs->ctx[f], local.data(), flip ? flip + i0 : nullptr, weights + static_cast<size_t>(i0) * K, i1 - i0, K,
                            mode, ss.data(), ds.data(), ac.data());
    if (rc != PD_OK) return fs->FailFrom(rc, f);
    // files are added in list order: a fixed reduction order, whatever ran where
    for (size_t j = 0; j < ss.size(); ++j) score_sums[j] += ss[j];
    for (uint32_t s = 0; s < n; ++s) {
      dosage_sums[s] += ds[s];
      allele_cts[s] += ac[s];
    }
    i0 = i1;
  }
  return PD_OK;
}

}  // extern "C"
