// Dosage decode (read_pgen / read_pfile dosages), dosage-aware counts (plink_freq dosage := true)
// and plink_score.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>

#include "ctx.h"
#include "group.h"
#include "outpipe.h"

using namespace pdx;

namespace pdx {

int CountsToDevice(pd_ctx_impl* c, const SubsetState& sub, uint32_t v_begin, uint32_t v_end, uint32_t* v_ext);

namespace {

uint32_t RankStride(const pd_ctx_impl* c) { return ((c->idx.sample_ct + 31) / 32 + 3) & ~3u; }

// Scratch of a dosage batch: DosageLoc per row and, when the file has dosage tracks, the rank table.
int ReserveDosage(pd_ctx_impl* c, uint32_t rows) {
  cudaError_t e = c->d_loc.Reserve(static_cast<size_t>(rows) * sizeof(pdk::DosageLoc));
  if (e == cudaSuccess && c->idx.has_dosage) e = c->d_rank.Reserve(static_cast<size_t>(rows) * RankStride(c) * 4);
  if (e == cudaSuccess) e = c->d_dstats.Reserve(static_cast<size_t>(rows) * sizeof(pdk::DosageStats));
  return e == cudaSuccess ? PD_OK : CheckCuda(c, e, "dosage scratch allocation");
}

// Rows per batch such that the rank tables stay within ~64 MiB.
uint32_t DosageBatchRows(const pd_ctx_impl* c, uint64_t want) {
  uint64_t r = std::min<uint64_t>(want, 65535);
  if (c->idx.has_dosage) r = std::min<uint64_t>(r, std::max<uint64_t>(1, (64ull << 20) / (RankStride(c) * 4ull)));
  return static_cast<uint32_t>(std::max<uint64_t>(1, r));
}

void LocateBatch(pd_ctx_impl* c, const RowBatch& b, const pdk::RecIndex& ix, cudaStream_t s) {
  if (c->idx.has_dosage) {
    pdk::launch_dosage_locate(ix, b.rows, b.d_vlist, b.v_first, c->d_loc.as<pdk::DosageLoc>(), c->d_rank.as<uint32_t>(),
                              RankStride(c), DevErr(c), s, &c->launch_stats);
  } else {
    cudaMemsetAsync(c->d_loc.p, 0, static_cast<size_t>(b.n) * sizeof(pdk::DosageLoc), s);
  }
}

// BiallelicDiploidMinimac3R2 (pgenlib_read.cc:383-412), same integer / fp64 operation order.
double Minimac3R2(uint64_t alt1_dosage, uint64_t hap_alt1_ssq_x2, uint32_t nm_sample_ct) {
  if (!nm_sample_ct) return std::numeric_limits<double>::quiet_NaN();
  const uint64_t nm_x32768 = nm_sample_ct * 0x8000ull;
  if (nm_sample_ct < 131072) {
    const uint64_t sq = alt1_dosage * alt1_dosage;
    const uint64_t observed = hap_alt1_ssq_x2 * nm_sample_ct - alt1_dosage * alt1_dosage;
    const uint64_t expected = nm_x32768 * alt1_dosage - sq;
    return static_cast<double>(observed) / static_cast<double>(expected);
  }
  auto u63tod = [](uint64_t x) { return static_cast<double>(static_cast<int64_t>(x)); };
  const double alt1d = u63tod(alt1_dosage);
  const double expected = alt1d * u63tod(nm_x32768 - alt1_dosage);
  const uint64_t ssq_hi = hap_alt1_ssq_x2 >> 32;
  uint64_t left_lo = (hap_alt1_ssq_x2 & 0xffffffffull) * nm_sample_ct;
  const uint64_t left_hi = (left_lo >> 32) + ssq_hi * nm_sample_ct;
  left_lo &= 0xffffffffu;
  const uint64_t a_lo = alt1_dosage & 0xffffffffull;
  const uint64_t a_hi = alt1_dosage >> 32;
  uint64_t right_lo = a_lo * a_lo;
  const uint64_t right_hi = (right_lo >> 32) + (a_lo + alt1_dosage) * a_hi;
  right_lo &= 0xffffffffu;
  const double observed_hi = u63tod(left_hi - right_hi);
  const int64_t observed_lo = static_cast<int64_t>(left_lo) - static_cast<int64_t>(right_lo);
  const double observed = (observed_hi * 4294967296.0) + static_cast<double>(observed_lo);
  return observed / expected;
}

// GetBasicGenotypeCountsAndDosage16s with is_minimac3_r2 = 0 (pgenlib_read.cc:7946-7962 for
// hardcall-only records, :8062-8172 otherwise), from exact integer inputs.
void FinishDosageCounts(const uint32_t gc[4], const pdk::DosageStats* st, uint32_t sample_ct, uint64_t* all_dosages,
                        double* imp_r2) {
  if (!st || st->kind == 0) {
    all_dosages[0] = static_cast<uint64_t>(gc[0] * 2 + gc[1]) * 16384ull;
    all_dosages[1] = static_cast<uint64_t>(gc[2] * 2 + gc[1]) * 16384ull;
    if (imp_r2) {
      const uint32_t nm = sample_ct - gc[3];
      const uint64_t alt1 = gc[2] * 0x8000ull + gc[1] * 0x4000ull;
      const uint64_t ssq = gc[2] * 0x40000000ull + gc[1] * 0x10000000ull;
      *imp_r2 = Minimac3R2(alt1, ssq, nm) * 2;
    }
    return;
  }
  uint32_t rep[4];
  if (st->kind == 2) {
    // unconditional dosage: the reference takes every non-missing hardcall as replaced
    rep[0] = gc[0];
    rep[1] = gc[1];
    rep[2] = gc[2];
    rep[3] = gc[3] + st->ct - sample_ct;
  } else {
    for (int k = 0; k < 4; ++k) rep[k] = st->replaced[k];
  }
  const uint32_t replaced_ct = rep[0] + rep[1] + rep[2];
  const uint32_t remaining_het = gc[1] - rep[1];
  const uint32_t remaining_hom_alt = gc[2] - rep[2];
  const uint32_t alt1_ct = 2 * remaining_hom_alt + remaining_het;
  const uint64_t alt1 = st->sum + alt1_ct * 16384ull;
  all_dosages[1] = alt1;
  const uint32_t nondosage_nm = sample_ct - gc[3] - replaced_ct;
  const uint32_t new_nm = st->ct + nondosage_nm;
  all_dosages[0] = new_nm * 32768ull - alt1;
  if (imp_r2) {
    const uint64_t sq = st->sumsq + remaining_het * 0x10000000ull + remaining_hom_alt * 0x40000000ull;
    *imp_r2 = Minimac3R2(alt1, sq, new_nm) * 2;
  }
}

}  // namespace
}  // namespace pdx

extern "C" {

int pd_dosage_f64(pd_ctx* c, const uint32_t* vidx, uint32_t n_var, double* out) {
  if (!c || (n_var && (!vidx || !out))) return PD_ERR_INVALID_ARG;
  if (c->group) {
    const pd_ctx* s0 = c->group->shards[0];
    const size_t n = s0->subset.active ? s0->subset.count : s0->idx.sample_ct;
    return GroupRunList(c, vidx, n_var, 1, [&](pd_ctx* s, uint32_t i0, uint32_t i1) {
      return pd_dosage_f64(s, vidx + i0, i1 - i0, out + i0 * n);
    });
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  const uint32_t n_out = c->subset.active ? c->subset.count : c->idx.sample_ct;
  if (!n_var || !n_out) return PD_OK;
  const uint64_t kOutBytes = 128ull << 20;
  const uint32_t max_rows = DosageBatchRows(c, std::max<uint64_t>(1, kOutBytes / (8ull * n_out)));
  int rc = ReserveDosage(c, max_rows);
  if (rc != PD_OK) return rc;
  OutPipe pipe;
  rc = pipe.Init(c, out, static_cast<size_t>(max_rows) * n_out * 8);
  if (rc != PD_OK) return rc;
  const pdk::SubsetDev sd = c->subset.Dev();
  cudaEventRecord(c->ev_c0, c->stream);
  rc = RunRowsList(c, vidx, n_var, max_rows, [&](const RowBatch& b, const pdk::RecIndex& ix, cudaStream_t s) {
    void* dev = nullptr;
    int slot = 0;
    int r = pipe.Acquire(&dev, &slot);
    if (r != PD_OK) return r;
    LocateBatch(c, b, ix, s);
    pdk::launch_dosage_fill(b.rows, c->d_loc.as<pdk::DosageLoc>(), c->d_rank.as<uint32_t>(), RankStride(c), sd, n_out,
                            static_cast<double*>(dev), DevErr(c), s, &c->launch_stats);
    c->stats.scan_launches += 1;
    return pipe.Send(slot, s, out + static_cast<size_t>(b.row0) * n_out, static_cast<size_t>(b.n) * n_out * 8);
  });
  const int rc2 = pipe.Finish();
  cudaEventRecord(c->ev_c1, c->stream);
  cudaEventSynchronize(c->ev_c1);
  if (rc == PD_OK && rc2 == PD_OK) {
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
    c->stats.call_ms = ms;
    c->stats.kernel_ms = pipe.produce_ms;  // the decode kernels alone; call_ms includes the D2H stream
  }
  return rc != PD_OK ? rc : rc2;
}

int pd_dosage_counts(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t* out_counts, uint64_t* out_all_dosages,
                     double* out_imp_r2) {
  if (!c) return PD_ERR_INVALID_ARG;
  if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
  const uint32_t n = v_end - v_begin;
  if (!n) return PD_OK;
  if (!out_counts || !out_all_dosages) return PD_ERR_INVALID_ARG;
  if (c->group) {
    return GroupRunRanges(c, v_begin, v_end, 1, [&](pd_ctx* s, uint32_t a, uint32_t b, uint32_t row) {
      return pd_dosage_counts(s, a, b, out_counts + 4ull * row, out_all_dosages + 2ull * row,
                              out_imp_r2 ? out_imp_r2 + row : nullptr);
    });
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  const uint32_t sample_ct = c->subset.active ? c->subset.count : c->idx.sample_ct;
  cudaEventRecord(c->ev_c0, c->stream);
  if (!c->idx.has_dosage) {
    // hardcall-only file: the count kernels at full speed, then the closed forms
    uint32_t v_ext = 0;
    int rc = CountsToDevice(c, c->subset, v_begin, v_end, &v_ext);
    if (rc != PD_OK) return rc;
    rc = DownloadRows(c, c->d_counts.as<uint4>() + (v_begin - v_ext), static_cast<size_t>(n) * 16, out_counts);
    if (rc != PD_OK) return rc;
    for (uint32_t i = 0; i < n; ++i) {
      FinishDosageCounts(out_counts + 4ull * i, nullptr, sample_ct, out_all_dosages + 2ull * i,
                         out_imp_r2 ? out_imp_r2 + i : nullptr);
    }
  } else {
    const uint32_t max_rows = DosageBatchRows(c, 65535);
    int rc = ReserveDosage(c, max_rows);
    if (rc != PD_OK) return rc;
    cudaError_t e = c->d_counts.Reserve(static_cast<size_t>(n) * 16);
    if (e == cudaSuccess) e = c->d_dstats.Reserve(static_cast<size_t>(n) * sizeof(pdk::DosageStats));
    if (e != cudaSuccess) return CheckCuda(c, e, "allocation");
    const pdk::SubsetDev sd = c->subset.Dev();
    rc = RunRows(c, v_begin, v_end, max_rows, 0, [&](const RowBatch& b, const pdk::RecIndex& ix, cudaStream_t s) {
      pdk::launch_count_rows(b.rows, sd, c->d_counts.as<uint4>() + b.row0, s, &c->launch_stats);
      LocateBatch(c, b, ix, s);
      pdk::launch_dosage_stats(b.rows, c->d_loc.as<pdk::DosageLoc>(), c->d_rank.as<uint32_t>(), RankStride(c), sd,
                               sample_ct, c->d_dstats.as<pdk::DosageStats>() + b.row0, s, &c->launch_stats);
      c->stats.scan_launches += 1;
      return PD_OK;
    });
    if (rc != PD_OK) return rc;
    std::vector<pdk::DosageStats> st(n);
    rc = DownloadRows(c, c->d_counts.p, static_cast<size_t>(n) * 16, out_counts);
    if (rc == PD_OK) rc = DownloadRows(c, c->d_dstats.p, static_cast<size_t>(n) * sizeof(pdk::DosageStats), st.data());
    if (rc != PD_OK) return rc;
    for (uint32_t i = 0; i < n; ++i) {
      FinishDosageCounts(out_counts + 4ull * i, &st[i], sample_ct, out_all_dosages + 2ull * i,
                         out_imp_r2 ? out_imp_r2 + i : nullptr);
    }
  }
  cudaEventRecord(c->ev_c1, c->stream);
  cudaEventSynchronize(c->ev_c1);
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
  c->stats.call_ms = ms;
  return PD_OK;
}

int pd_score(pd_ctx* c, const uint32_t* vidx, const uint8_t* flip, const double* weights, uint32_t n_var, uint32_t K,
             uint32_t mode, double* score_sums, double* dosage_sums, uint32_t* allele_cts) {
  if (!c || !K || mode > 2 || (n_var && (!vidx || !weights))) return PD_ERR_INVALID_ARG;
  if (c->group) return GroupScore(c, vidx, flip, weights, n_var, K, mode, score_sums, dosage_sums, allele_cts);
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  const uint32_t n_out = c->subset.active ? c->subset.count : c->idx.sample_ct;
  if (!n_out) return PD_OK;
  cudaError_t e = c->d_score.Reserve(static_cast<size_t>(n_out) * K * 8);
  if (e == cudaSuccess) e = c->d_dsum.Reserve(static_cast<size_t>(n_out) * 8);
  if (e == cudaSuccess) e = c->d_act.Reserve(static_cast<size_t>(n_out) * 4);
  if (e == cudaSuccess) e = c->d_weights.Reserve(std::max<size_t>(8, static_cast<size_t>(n_var) * K * 8));
  if (e == cudaSuccess) e = c->d_flip.Reserve(std::max<size_t>(4, n_var));
  if (e != cudaSuccess) return CheckCuda(c, e, "score allocation");
  cudaEventRecord(c->ev_c0, c->stream);
  cudaMemsetAsync(c->d_score.p, 0, static_cast<size_t>(n_out) * K * 8, c->stream);
  cudaMemsetAsync(c->d_dsum.p, 0, static_cast<size_t>(n_out) * 8, c->stream);
  cudaMemsetAsync(c->d_act.p, 0, static_cast<size_t>(n_out) * 4, c->stream);
  int rc = PD_OK;
  if (n_var) {
    e = cudaMemcpyAsync(c->d_weights.p, weights, static_cast<size_t>(n_var) * K * 8, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess && flip) e = cudaMemcpyAsync(c->d_flip.p, flip, n_var, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return CheckCuda(c, e, "weights upload");
    c->stats.h2d_bytes += static_cast<size_t>(n_var) * (K * 8 + (flip ? 1 : 0));
    const pdk::SubsetDev sd = c->subset.Dev();
    const bool dos = c->idx.has_dosage;
    // K7t (tensor cores) takes hardcall-only files scanned over all samples; explicit dosages and sample
    // subsets stay on the fp64 CUDA-core kernel.  PD_SCORE_PATH=fp64 forces the latter (A/B timing).
    static const bool force_fp64 = [] {
      const char* v = getenv("PD_SCORE_PATH");
      return v && !strcmp(v, "fp64");
    }();
    const bool tc = !dos && !c->subset.active && pdk::score_tc_applicable(K) && !force_fp64;
    const uint32_t max_rows = DosageBatchRows(c, tc ? 65535 : 4096);
    rc = ReserveDosage(c, max_rows);
    if (rc != PD_OK) return rc;
    e = c->d_counts.Reserve(static_cast<size_t>(max_rows) * 16);
    if (e == cudaSuccess) e = c->d_srow.Reserve(static_cast<size_t>(max_rows) * sizeof(pdk::ScoreRow));
    if (e == cudaSuccess && tc) e = c->d_tc.Reserve(pdk::score_tc_workspace_bytes(max_rows, n_out));
    if (e != cudaSuccess) return CheckCuda(c, e, "score allocation");
    bool first = true;
    rc = RunRowsList(c, vidx, n_var, max_rows, [&](const RowBatch& b, const pdk::RecIndex& ix, cudaStream_t s) {
      pdk::launch_count_rows(b.rows, sd, c->d_counts.as<uint4>(), s, &c->launch_stats);
      if (dos) {
        LocateBatch(c, b, ix, s);
        pdk::launch_dosage_stats(b.rows, c->d_loc.as<pdk::DosageLoc>(), c->d_rank.as<uint32_t>(), RankStride(c), sd,
                                 n_out, c->d_dstats.as<pdk::DosageStats>(), s, &c->launch_stats);
      }
      pdk::launch_score_prep(c->d_counts.as<uint4>(), dos ? c->d_dstats.as<pdk::DosageStats>() : nullptr,
                             flip ? c->d_flip.as<uint8_t>() + b.row0 : nullptr, b.n, mode, c->d_srow.as<pdk::ScoreRow>(),
                             s, &c->launch_stats);
      const double* w = c->d_weights.as<double>() + static_cast<size_t>(b.row0) * K;
      if (tc) {
        pdk::launch_score_tc(b.rows, c->d_counts.as<uint4>(), c->d_srow.as<pdk::ScoreRow>(), w, K, mode, n_out, max_rows,
                             c->d_tc.p, first, c->d_score.as<double>(), c->d_dsum.as<double>(), c->d_act.as<uint32_t>(),
                             s, &c->launch_stats);
        first = false;
      } else {
        pdk::launch_score_acc(b.rows, c->d_srow.as<pdk::ScoreRow>(), dos ? c->d_loc.as<pdk::DosageLoc>() : nullptr,
                              c->d_rank.as<uint32_t>(), RankStride(c), w, K, sd, n_out, mode, c->d_score.as<double>(),
                              c->d_dsum.as<double>(), c->d_act.as<uint32_t>(), s, &c->launch_stats);
      }
      c->stats.scan_launches += 1;
      return PD_OK;
    });
    if (rc != PD_OK) return rc;
  }
  c->last_score_n = n_out;
  c->last_score_k = K;
  e = cudaSuccess;
  if (score_sums) {
    e = cudaMemcpyAsync(score_sums, c->d_score.p, static_cast<size_t>(n_out) * K * 8, cudaMemcpyDeviceToHost, c->stream);
  }
  if (e == cudaSuccess && dosage_sums) {
    e = cudaMemcpyAsync(dosage_sums, c->d_dsum.p, static_cast<size_t>(n_out) * 8, cudaMemcpyDeviceToHost, c->stream);
  }
  if (e == cudaSuccess && allele_cts) {
    e = cudaMemcpyAsync(allele_cts, c->d_act.p, static_cast<size_t>(n_out) * 4, cudaMemcpyDeviceToHost, c->stream);
  }
  cudaEventRecord(c->ev_c1, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "result download");
  c->stats.d2h_bytes += static_cast<size_t>(n_out) * (K * 8 + 12);
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
  c->stats.call_ms = ms;
  return PD_OK;
}

}  // extern "C"
