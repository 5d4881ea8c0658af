// pd_tidy_i8: read_pfile(orient := 'genotype') -- one row per (variant, sample) call, with the genotype_range row
// filter applied on the device (src/pfile_reader.cpp:2960-3281).  See kernels_tidy.cu for the row order.
#include <algorithm>
#include <cstring>
#include <vector>

#include "ctx.h"
#include "group.h"

using namespace pdx;

namespace pdx {
int CountsToDevice(pd_ctx_impl* c, const SubsetState& sub, uint32_t v_begin, uint32_t v_end, uint32_t* v_ext);
}

namespace {

uint64_t RowsOf(const uint32_t* cnt, uint32_t mask) {
  return ((mask & 1u) ? cnt[0] : 0u) + static_cast<uint64_t>((mask & 2u) ? cnt[1] : 0u) + ((mask & 4u) ? cnt[2] : 0u) +
         ((mask & 8u) ? cnt[3] : 0u);
}

int GroupTidy(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t emit_mask, uint64_t row_cap, uint32_t* out_variant,
              uint32_t* out_sample, int8_t* out_geno, uint64_t* out_rows) {
  // rows per variant from the (multi-GPU) counts, then every GPU expands its variant range at its row offset
  const uint32_t n = v_end - v_begin;
  std::vector<uint32_t> counts(static_cast<size_t>(n) * 4);
  int rc = pd_counts(c, v_begin, v_end, counts.data());
  if (rc != PD_OK) return rc;
  std::vector<uint64_t> off(static_cast<size_t>(n) + 1, 0);
  for (uint32_t i = 0; i < n; ++i) off[i + 1] = off[i] + RowsOf(&counts[4ull * i], emit_mask);
  if (out_rows) *out_rows = off[n];
  if (!out_geno) return PD_OK;
  if (off[n] > row_cap) return c->SetError(PD_ERR_INVALID_ARG, "pd_tidy_i8: the output arrays are too small (see *out_rows)");
  return GroupRunRanges(c, v_begin, v_end, 1, [&](pd_ctx* s, uint32_t a, uint32_t b, uint32_t row) {
    const uint64_t base = off[row];
    uint64_t got = 0;
    const int r = pd_tidy_i8(s, a, b, emit_mask, off[row + (b - a)] - base, out_variant + base, out_sample + base,
                             out_geno + base, &got);
    if (r == PD_OK && got != off[row + (b - a)] - base) return s->SetError(PD_ERR_READ, "pd_tidy_i8: row count changed between passes");
    return r;
  });
}

}  // namespace

extern "C" {

int pd_tidy_i8(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t emit_mask, uint64_t row_cap, uint32_t* out_variant,
               uint32_t* out_sample, int8_t* out_geno, uint64_t* out_rows) {
  if (!c || emit_mask > 15u) return PD_ERR_INVALID_ARG;
  const bool query = !out_geno;
  if (!query && (!out_variant || !out_sample)) return PD_ERR_INVALID_ARG;
  if (out_rows) *out_rows = 0;
  if (c->group) {
    if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
    return GroupTidy(c, v_begin, v_end, emit_mask, row_cap, out_variant, out_sample, out_geno, out_rows);
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
  const uint32_t n = v_end - v_begin;
  const uint32_t n_out = c->subset.active ? c->subset.count : c->idx.sample_ct;
  if (!n || !n_out || !emit_mask) return PD_OK;
  cudaEventRecord(c->ev_c0, c->stream);
  // pass 1: genotype counts -> rows per variant -> row offsets
  uint32_t v_ext = 0;
  int rc = CountsToDevice(c, c->subset, v_begin, v_end, &v_ext);
  if (rc != PD_OK) return rc;
  cudaError_t e = c->d_tidy_off.Reserve((static_cast<size_t>(n) + 1) * 8);
  if (e != cudaSuccess) return CheckCuda(c, e, "row offset allocation");
  unsigned long long* d_off = c->d_tidy_off.as<unsigned long long>();
  pdk::launch_tidy_rows(c->d_counts.as<uint4>() + (v_begin - v_ext), n, emit_mask, d_off + 1, c->stream, &c->launch_stats);
  std::vector<unsigned long long> off(static_cast<size_t>(n) + 1, 0);
  rc = DownloadRows(c, d_off + 1, static_cast<size_t>(n) * 8, off.data() + 1);
  if (rc != PD_OK) return rc;
  for (uint32_t i = 0; i < n; ++i) off[i + 1] += off[i];
  if (out_rows) *out_rows = off[n];
  if (query) return PD_OK;
  if (off[n] > row_cap) return c->SetError(PD_ERR_INVALID_ARG, "pd_tidy_i8: the output arrays are too small (see *out_rows)");
  e = cudaMemcpyAsync(d_off, off.data(), off.size() * 8, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "row offset upload");
  // pass 2: expansion, batch by batch; a batch's rows are three contiguous runs of the caller's arrays
  const uint64_t kBatchRows = 24ull << 20;  // 24 Mi rows = 216 MiB of output per buffer
  const uint32_t max_rows = static_cast<uint32_t>(std::max<uint64_t>(1, std::min<uint64_t>(65535, kBatchRows / n_out)));
  const size_t cap = static_cast<size_t>(max_rows) * n_out;
  const size_t off_s = (cap * 4 + 255) & ~static_cast<size_t>(255), off_g = 2 * off_s;
  for (int b = 0; b < 2; ++b) {
    e = c->d_out[b].Reserve(off_g + cap + 256);
    if (e != cudaSuccess) return CheckCuda(c, e, "output buffer allocation");
  }
  const uint32_t* idx = c->subset.active ? c->subset.d_idx.as<uint32_t>() : nullptr;
  uint32_t batch_no = 0;
  bool used[2] = {false, false};
  rc = RunRows(c, v_begin, v_end, max_rows, 0, [&](const RowBatch& b, const pdk::RecIndex&, cudaStream_t s) {
    const int slot = static_cast<int>(batch_no++ & 1u);
    if (used[slot]) {
      const cudaError_t ee = cudaEventSynchronize(c->ev_out_free[slot]);  // the slot's previous rows are on the host
      if (ee != cudaSuccess) return CheckCuda(c, ee, "result download");
    }
    used[slot] = true;
    char* dev = c->d_out[slot].as<char>();
    const unsigned long long o0 = off[b.row0], o1 = off[b.row0 + b.n];
    pdk::launch_tidy(b.rows, idx, n_out, emit_mask, d_off + b.row0, o0, b.v_first, b.d_vlist,
                     reinterpret_cast<uint32_t*>(dev), reinterpret_cast<uint32_t*>(dev + off_s),
                     reinterpret_cast<int8_t*>(dev + off_g), s, &c->launch_stats);
    c->stats.scan_launches += 1;
    cudaEventRecord(c->ev_out_ready[slot], s);
    cudaStreamWaitEvent(c->d2h_stream, c->ev_out_ready[slot], 0);
    const size_t rows = static_cast<size_t>(o1 - o0);
    if (rows) {
      cudaMemcpyAsync(out_variant + o0, dev, rows * 4, cudaMemcpyDeviceToHost, c->d2h_stream);
      cudaMemcpyAsync(out_sample + o0, dev + off_s, rows * 4, cudaMemcpyDeviceToHost, c->d2h_stream);
      cudaMemcpyAsync(out_geno + o0, dev + off_g, rows, cudaMemcpyDeviceToHost, c->d2h_stream);
      c->stats.d2h_bytes += rows * 9;
    }
    cudaEventRecord(c->ev_out_free[slot], c->d2h_stream);
    return static_cast<int>(PD_OK);
  });
  e = cudaStreamSynchronize(c->d2h_stream);
  if (rc == PD_OK && e != cudaSuccess) rc = CheckCuda(c, e, "result download");
  cudaEventRecord(c->ev_c1, c->stream);
  cudaEventSynchronize(c->ev_c1);
  if (rc == PD_OK) {
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
    c->stats.call_ms = ms;
  }
  return rc;
}

}  // extern "C"
