// Other consumers of the decode kernels: plink_ld's pairwise sums and the standardised genotype
// rows of plink_pca / plink_glm.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "ctx.h"
#include "group.h"
#include "outpipe.h"

using namespace pdx;

namespace {

// Makes [v_lo, v_hi) resident for the duration of a call that needs random access to its rows; restores whatever
// range the caller had loaded with pd_load_resident.
struct SpanGuard {
  pd_ctx_impl* c;
  bool swapped = false;
  bool resident;
  const uint8_t* res_ptr;
  uint32_t res_v0, res_v1;
  uint64_t res_f0, res_f1;
  explicit SpanGuard(pd_ctx_impl* ctx) : c(ctx) {}
  int Ensure(uint32_t v_lo, uint32_t v_hi) {
    if (c->dev_image) return PD_OK;
    if (c->resident && ExtendToBase(c, v_lo) >= c->res_v0 && v_hi <= c->res_v1) return PD_OK;
    resident = c->resident;
    res_ptr = c->res_ptr;
    res_v0 = c->res_v0;
    res_v1 = c->res_v1;
    res_f0 = c->res_f0;
    res_f1 = c->res_f1;
    swapped = true;
    return LoadRange(c, v_lo, v_hi, &c->d_span);
  }
  ~SpanGuard() {
    if (!swapped) return;
    c->resident = resident;
    c->res_ptr = res_ptr;
    c->res_v0 = res_v0;
    c->res_v1 = res_v1;
    c->res_f0 = res_f0;
    c->res_f1 = res_f1;
  }
};

}  // namespace

extern "C" {

int pd_ld_pairs(pd_ctx* c, const uint32_t* va, const uint32_t* vb, uint32_t n_pairs, uint32_t* out_sums) {
  if (!c || (n_pairs && (!va || !vb || !out_sums))) return PD_ERR_INVALID_ARG;
  if (!n_pairs) return PD_OK;
  if (c->group) {
    // pairs are independent: contiguous chunks of the list, one per GPU (each stages the span its chunk touches)
    GroupState* g = c->group;
    const uint32_t k = static_cast<uint32_t>(g->shards.size());
    std::vector<VRange> parts(k);
    for (uint32_t i = 0; i < k; ++i) {
      parts[i] = VRange{static_cast<uint32_t>(static_cast<uint64_t>(n_pairs) * i / k),
                        static_cast<uint32_t>(static_cast<uint64_t>(n_pairs) * (i + 1) / k)};
    }
    int rc = PD_OK;
    for (uint32_t i = 0; i < k; ++i) {
      g->workers[i]->Post([&, i] {
        const uint32_t a = parts[i].v0, n = parts[i].v1 - parts[i].v0;
        return n ? pd_ld_pairs(g->shards[i], va + a, vb + a, n, out_sums + 6ull * a) : static_cast<int>(PD_OK);
      });
    }
    for (uint32_t i = 0; i < k; ++i) {
      const int r = g->workers[i]->Wait();
      if (r != PD_OK && rc == PD_OK) {
        rc = r;
        c->last_error = g->shards[i]->last_error;
      }
    }
    return rc;
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  const uint32_t V = c->idx.variant_ct;
  std::vector<uint32_t> uniq;
  uniq.reserve(2ull * n_pairs);
  for (uint32_t i = 0; i < n_pairs; ++i) {
    if (va[i] >= V || vb[i] >= V) return c->SetError(PD_ERR_INVALID_ARG, "pd_ld_pairs: variant index out of range");
    uniq.push_back(va[i]);
    uniq.push_back(vb[i]);
  }
  std::sort(uniq.begin(), uniq.end());
  uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
  const uint32_t nu = static_cast<uint32_t>(uniq.size());
  // every row of the call must be addressable at once: one dense batch
  const uint64_t max_batch = c->idx.fixed ? 65535ull : std::min<uint64_t>(65535ull, (1024ull << 20) / DenseSlotStride(c));
  if (nu > max_batch) {
    char buf[160];
    snprintf(buf, sizeof(buf), "pd_ld_pairs: %u distinct variants in one call, at most %llu fit one batch; pass fewer anchors per call",
             nu, static_cast<unsigned long long>(max_batch));
    return c->SetError(PD_ERR_INVALID_ARG, buf);
  }
  std::vector<uint32_t> pos(2ull * n_pairs);
  for (uint32_t i = 0; i < n_pairs; ++i) {
    pos[i] = static_cast<uint32_t>(std::lower_bound(uniq.begin(), uniq.end(), va[i]) - uniq.begin());
    pos[n_pairs + i] = static_cast<uint32_t>(std::lower_bound(uniq.begin(), uniq.end(), vb[i]) - uniq.begin());
  }
  SpanGuard guard(c);
  int rc = guard.Ensure(uniq.front(), uniq.back() + 1);
  if (rc != PD_OK) return rc;
  cudaError_t e = c->d_pairs.Reserve(pos.size() * 4);
  if (e == cudaSuccess) e = c->d_out[0].Reserve(static_cast<size_t>(n_pairs) * 24 + 64);
  if (e != cudaSuccess) return CheckCuda(c, e, "pair list allocation");
  e = cudaMemcpyAsync(c->d_pairs.p, pos.data(), pos.size() * 4, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "pair list upload");
  cudaEventRecord(c->ev_c0, c->stream);
  uint32_t calls = 0;
  const uint32_t* inc = c->subset.active ? c->subset.d_words.as<uint32_t>() : nullptr;
  rc = RunRowsList(c, uniq.data(), nu, static_cast<uint32_t>(max_batch), [&](const RowBatch& b, const pdk::RecIndex&, cudaStream_t s) {
    if (b.n != nu || calls++) return c->SetError(PD_ERR_INVALID_ARG, "pd_ld_pairs: the variants of the call did not form one batch");
    pdk::launch_ld_pairs(b.rows, c->d_pairs.as<uint32_t>(), c->d_pairs.as<uint32_t>() + n_pairs, n_pairs, inc,
                         c->d_out[0].as<uint32_t>(), s, &c->launch_stats);
    c->stats.scan_launches += 1;
    return static_cast<int>(PD_OK);
  });
  if (rc != PD_OK) return rc;
  rc = DownloadRows(c, c->d_out[0].p, static_cast<size_t>(n_pairs) * 24, out_sums);
  cudaEventRecord(c->ev_c1, c->stream);
  cudaEventSynchronize(c->ev_c1);
  float ms = 0;
  if (cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1) == cudaSuccess) c->stats.call_ms = ms;
  return rc;
}

int pd_standardized_f64(pd_ctx* c, const uint32_t* vidx, uint32_t n_var, double* out, uint8_t* out_skip) {
  if (!c || (n_var && (!vidx || !out))) return PD_ERR_INVALID_ARG;
  if (!n_var) return PD_OK;
  if (c->group) {
    const pd_ctx* s0 = c->group->shards[0];
    const size_t n = s0->subset.active ? s0->subset.count : s0->idx.sample_ct;
    return GroupRunList(c, vidx, n_var, 1, [&](pd_ctx* s, uint32_t i0, uint32_t i1) {
      return pd_standardized_f64(s, vidx + i0, i1 - i0, out + i0 * n, out_skip ? out_skip + i0 : nullptr);
    });
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  const uint32_t n_out = c->subset.active ? c->subset.count : c->idx.sample_ct;
  if (!n_out) return PD_OK;
  const uint64_t kOutBytes = 128ull << 20;
  const uint32_t max_rows = static_cast<uint32_t>(std::max<uint64_t>(1, std::min<uint64_t>(65535, kOutBytes / (8ull * n_out))));
  cudaError_t e = c->d_counts.Reserve(static_cast<size_t>(max_rows) * 16);
  if (e == cudaSuccess) e = c->d_pairs.Reserve(static_cast<size_t>(max_rows) * 16);
  if (e == cudaSuccess) e = c->h_out.Reserve(static_cast<size_t>(max_rows) * 32);
  if (e != cudaSuccess) return CheckCuda(c, e, "allocation");
  OutPipe pipe;
  int rc = pipe.Init(c, out, static_cast<size_t>(max_rows) * n_out * 8);
  if (rc != PD_OK) return rc;
  const pdk::SubsetDev sd = c->subset.Dev();
  const uint32_t* idx = c->subset.active ? c->subset.d_idx.as<uint32_t>() : nullptr;
  uint32_t* h_counts = c->h_out.as<uint32_t>();
  double* h_norm = reinterpret_cast<double*>(c->h_out.as<char>() + static_cast<size_t>(max_rows) * 16);
  cudaEventRecord(c->ev_c0, c->stream);
  rc = RunRowsList(c, vidx, n_var, max_rows, [&](const RowBatch& b, const pdk::RecIndex&, cudaStream_t s) {
    // genotype counts of the batch -> ALT frequency -> centre and 1 / sd per variant, in the reference's own
    // operation order (src/plink_pca.cpp:403-410, src/plink_common.cpp:1521-1533), on the host
    pdk::launch_count_rows(b.rows, sd, c->d_counts.as<uint4>(), s, &c->launch_stats);
    cudaError_t ee = cudaMemcpyAsync(h_counts, c->d_counts.p, static_cast<size_t>(b.n) * 16, cudaMemcpyDeviceToHost, s);
    if (ee == cudaSuccess) ee = cudaStreamSynchronize(s);
    if (ee != cudaSuccess) return CheckCuda(c, ee, "counts download");
    for (uint32_t r = 0; r < b.n; ++r) {
      const uint32_t* g = h_counts + 4ull * r;
      const uint32_t obs = g[0] + g[1] + g[2];
      double center = 0.0, inv = 0.0;
      bool skip = true;
      if (obs) {
        const double alt_freq = (static_cast<double>(g[1]) + 2.0 * static_cast<double>(g[2])) / (2.0 * static_cast<double>(obs));
        if (!(alt_freq <= 0.0 || alt_freq >= 1.0)) {
          center = 2.0 * alt_freq;
          inv = 1.0 / std::sqrt(2.0 * alt_freq * (1.0 - alt_freq));
          skip = false;
        }
      }
      h_norm[2ull * r] = skip ? 0.0 : center;
      h_norm[2ull * r + 1] = inv;  // 0 for skipped variants: the whole row becomes +-0
      if (out_skip) out_skip[b.row0 + r] = skip ? 1 : 0;
    }
    ee = cudaMemcpyAsync(c->d_pairs.p, h_norm, static_cast<size_t>(b.n) * 16, cudaMemcpyHostToDevice, s);
    if (ee != cudaSuccess) return CheckCuda(c, ee, "norm upload");
    void* dev = nullptr;
    int slot = 0;
    int r2 = pipe.Acquire(&dev, &slot);
    if (r2 != PD_OK) return r2;
    pdk::launch_standardize(b.rows, idx, n_out, c->d_pairs.as<double2>(), static_cast<double*>(dev), s, &c->launch_stats);
    c->stats.scan_launches += 1;
    r2 = pipe.Send(slot, s, out + static_cast<size_t>(b.row0) * n_out, static_cast<size_t>(b.n) * n_out * 8);
    if (r2 != PD_OK) return r2;
    ee = cudaStreamSynchronize(s);  // h_norm / d_pairs are reused by the next batch
    return ee == cudaSuccess ? static_cast<int>(PD_OK) : CheckCuda(c, ee, "standardise");
  });
  const int rc2 = pipe.Finish();
  cudaEventRecord(c->ev_c1, c->stream);
  cudaEventSynchronize(c->ev_c1);
  float ms = 0;
  if (cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1) == cudaSuccess) c->stats.call_ms = ms;
  return rc != PD_OK ? rc : rc2;
}

}  // extern "C"
