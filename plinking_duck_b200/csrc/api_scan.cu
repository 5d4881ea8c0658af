// Per-variant reductions behind plink_freq / plink_hardy / plink_missing (variant mode).
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>

#include "ctx.h"
#include "group.h"

using namespace pdx;

namespace pdx {

// Enqueues the count kernels of one tile; out_dev indexes rows from v_ext.
int CountTile(pd_ctx_impl* c, const SubsetState& sub, const Tile& t, const pdk::RecIndex& ix, uint32_t v_ext,
              uint4* out_dev, cudaStream_t s) {
  const pdk::SubsetDev sd = sub.Dev();
  if (c->idx.fixed) {
    pdk::launch_count_fixed(ix, sd, t.v0, t.v1, out_dev + (t.v0 - v_ext), s, &c->launch_stats);
    c->stats.scan_launches += 1;
    return PD_OK;
  }
  pdk::LdScratch scratch{};
  // scratch slots used by LD records of this tile form a contiguous range
  uint32_t slot_lo = 0xffffffffu, slot_hi = 0;
  if (c->idx.has_ld && !c->slot_vidx.empty()) {
    auto lo = std::lower_bound(c->slot_vidx.begin(), c->slot_vidx.end(), t.v0);
    auto hi = std::lower_bound(c->slot_vidx.begin(), c->slot_vidx.end(), t.v1);
    if (lo != hi) {
      slot_lo = static_cast<uint32_t>(lo - c->slot_vidx.begin());
      slot_hi = static_cast<uint32_t>(hi - c->slot_vidx.begin()) - 1;
    }
  }
  if (slot_lo != 0xffffffffu) {
    scratch.stride = ((c->idx.sample_ct + 3) / 4 + 15) / 16 * 16 + 16;
    scratch.slot_lo = slot_lo;
    scratch.slot_ct = slot_hi - slot_lo + 1;
    const size_t need = static_cast<size_t>(slot_hi - slot_lo + 1) * scratch.stride;
    if (need > c->d_scratch.cap) {
      // growing frees the old buffer: make sure no earlier tile still reads it
      cudaStreamSynchronize(s);
      cudaError_t e = c->d_scratch.Reserve(need + (need >> 2));
      if (e != cudaSuccess) return CheckCuda(c, e, "LD scratch allocation");
    }
    scratch.dense = c->d_scratch.as<uint8_t>();
    pdk::launch_decode_bases(ix, c->d_slot_vidx.as<uint32_t>(), scratch, slot_hi, DevErr(c), s, &c->launch_stats);
  }
  pdk::launch_count_var(ix, sd, scratch, t.v0, t.v1, out_dev + (t.v0 - v_ext), DevWorkCounter(c), DevErr(c), s,
                        &c->launch_stats);
  c->stats.scan_launches += 1;
  return PD_OK;
}

// Counts of [v_begin, v_end) into d_counts (rows from *v_ext); leaves them on the device.
int CountsToDevice(pd_ctx_impl* c, const SubsetState& sub, uint32_t v_begin, uint32_t v_end, uint32_t* v_ext) {
  const uint32_t v0 = (v_begin < v_end) ? ExtendToBase(c, v_begin) : v_begin;
  cudaError_t e = c->d_counts.Reserve(std::max<size_t>(16, static_cast<size_t>(v_end - v0) * 16));
  if (e != cudaSuccess) return CheckCuda(c, e, "counts allocation");
  uint4* out_dev = c->d_counts.as<uint4>();
  // Unmasked counts never read the sample ids of a plain difflist record: stage those records trimmed.
  struct TrimScope {
    pd_ctx_impl* c;
    ~TrimScope() { c->trim_scan = false; }
  } scope{c};
  c->trim_scan = !sub.active && !c->idx.fixed;
  return RunScan(c, v_begin, v_end, v_ext, [&](const Tile& t, const pdk::RecIndex& ix, cudaStream_t s) {
    return CountTile(c, sub, t, ix, v0, out_dev, s);
  });
}

int DownloadRows(pd_ctx_impl* c, const void* dev, size_t bytes, void* host) {
  if (!bytes) return PD_OK;
  cudaError_t e = cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "result download");
  c->stats.d2h_bytes += bytes;
  return PD_OK;
}

}  // namespace pdx

extern "C" {

int pd_counts(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t* out_counts) {
  if (!c || (!out_counts && v_end > v_begin)) return PD_ERR_INVALID_ARG;
  if (c->group) {
    return GroupRunRanges(c, v_begin, v_end, 1, [&](pd_ctx* s, uint32_t a, uint32_t b, uint32_t row) {
      return pd_counts(s, a, b, out_counts + 4ull * row);
    });
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  cudaEventRecord(c->ev_c0, c->stream);
  uint32_t v_ext = 0;
  int rc = CountsToDevice(c, c->subset, v_begin, v_end, &v_ext);
  if (rc != PD_OK) return rc;
  if (v_end > v_begin) {
    cudaError_t e = cudaMemcpyAsync(out_counts, c->d_counts.as<uint4>() + (v_begin - v_ext),
                                    static_cast<size_t>(v_end - v_begin) * 16, cudaMemcpyDeviceToHost, c->stream);
    cudaEventRecord(c->ev_c1, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return CheckCuda(c, e, "result download");
    c->stats.d2h_bytes += static_cast<size_t>(v_end - v_begin) * 16;
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
    c->stats.call_ms = ms;
  }
  return PD_OK;
}

int pd_counts_masked(pd_ctx* c, uint32_t v_begin, uint32_t v_end, const uint64_t* stratum_include,
                     uint32_t* out_counts) {
  if (!c || !stratum_include || (!out_counts && v_end > v_begin)) return PD_ERR_INVALID_ARG;
  if (c->group) {
    return GroupRunRanges(c, v_begin, v_end, 1, [&](pd_ctx* s, uint32_t a, uint32_t b, uint32_t row) {
      return pd_counts_masked(s, a, b, stratum_include, out_counts + 4ull * row);
    });
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  SubsetState st;
  int rc = BuildSubset(c, stratum_include, &c->subset, &st);
  if (rc == PD_OK) {
    uint32_t v_ext = 0;
    rc = CountsToDevice(c, st, v_begin, v_end, &v_ext);
    if (rc == PD_OK) {
      rc = DownloadRows(c, c->d_counts.as<uint4>() + (v_begin - v_ext), static_cast<size_t>(v_end - v_begin) * 16,
                        out_counts);
    }
  }
  st.Free();
  return rc;
}

}  // extern "C"

// Starts the job thread behind pd_counts_start / pd_hardy_start: the scan runs tile by tile on the context's
// streams and every finished tile's rows (counts, and exact-test p-values for a hardy job) are published to the
// job's host arrays, where any number of scan threads fetch their claimed row ranges.
int pdx::StartScanJob(pd_ctx* c, uint32_t v_begin, uint32_t v_end, bool hardy, uint32_t midp, pd_job** out) {
  if (!c || !out) return PD_ERR_INVALID_ARG;
  *out = nullptr;
  if (c->group) return GroupCountsStart(c, v_begin, v_end, hardy, midp, out);
  if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
  pd_job* j = new (std::nothrow) pd_job();
  if (!j) return PD_ERR_NOMEM;
  j->ctx = c;
  j->v_begin = v_begin;
  j->v_end = v_end;
  j->hardy = hardy;
  j->midp = midp;
  j->counts.resize(static_cast<size_t>(v_end - v_begin) * 4);
  if (hardy) j->pvals.resize(v_end - v_begin);
  j->worker = std::thread([j]() {
    pd_ctx* c = j->ctx;
    std::lock_guard<std::mutex> lock(c->mu);
    {
      // pd_counts_start returns only now: the job owns the context (and the sample subset it was started
      // with) until it finishes; a pd_set_sample_subset issued right after the start waits for it
      std::lock_guard<std::mutex> l2(j->mu);
      j->started = true;
    }
    j->cv.notify_all();
    cudaSetDevice(c->device);
    const uint32_t v0 = (j->v_begin < j->v_end) ? ExtendToBase(c, j->v_begin) : j->v_begin;
    int rc = PD_OK;
    const size_t nrows_ext = std::max<size_t>(1, j->v_end - v0);
    cudaError_t e = c->d_counts.Reserve(nrows_ext * 16);
    if (e == cudaSuccess && j->hardy) e = c->d_pvals.Reserve(nrows_ext * 8);
    if (e == cudaSuccess && j->hardy) e = c->d_ready.Reserve(nrows_ext * 4);
    if (e == cudaSuccess && j->hardy) e = c->d_hweq.Reserve((nrows_ext + pdk::kHweQueueWords) * 4);
    if (e != cudaSuccess) rc = CheckCuda(c, e, "counts allocation");
    // Rows are published tile by tile: tile i's rows are copied out while tile i+1 is staged.
    struct Pending { Tile t; cudaEvent_t ev; };
    std::vector<Pending> pending;
    bool bad = false;  // a malformed record was flagged: nothing from that tile on is published
    auto publish = [&](const Pending& p) {
      if (bad) return;
      cudaEventSynchronize(p.ev);
      int flag[2] = {0, 0};
      cudaMemcpy(flag, DevErr(c), sizeof(flag), cudaMemcpyDeviceToHost);  // the tile's kernels are done: final for it
      if (flag[0]) {
        bad = true;
        return;
      }
      const uint32_t a = std::max(p.t.v0, j->v_begin), b = p.t.v1;
      if (b > a) {
        cudaMemcpy(j->counts.data() + static_cast<size_t>(a - j->v_begin) * 4, c->d_counts.as<uint4>() + (a - v0),
                   static_cast<size_t>(b - a) * 16, cudaMemcpyDeviceToHost);
        c->stats.d2h_bytes += static_cast<size_t>(b - a) * 16;
        if (j->hardy) {
          cudaMemcpy(j->pvals.data() + (a - j->v_begin), c->d_pvals.as<double>() + (a - v0), static_cast<size_t>(b - a) * 8,
                     cudaMemcpyDeviceToHost);
          c->stats.d2h_bytes += static_cast<size_t>(b - a) * 8;
        }
      }
      {
        std::lock_guard<std::mutex> l2(j->mu);
        j->rows_ready = b - j->v_begin;
      }
      j->cv.notify_all();
    };
    if (rc == PD_OK) {
      uint4* out_dev = c->d_counts.as<uint4>();
      rc = RunScan(
          c, j->v_begin, j->v_end, nullptr,
          [&](const Tile& t, const pdk::RecIndex& ix, cudaStream_t s) {
            if (!j->hardy) return CountTile(c, c->subset, t, ix, v0, out_dev, s);
            // plink_hardy: the tile's counts and exact tests (fused K1h for fixed-width records, else K3 + K8)
            const uint32_t off = t.v0 - v0;
            double* p_dev = c->d_pvals.as<double>() + off;
            if (c->idx.fixed && pdk::launch_hardy_fixed(ix, c->subset.Dev(), t.v0, t.v1, out_dev + off, p_dev, j->midp,
                                                       c->d_hweq.as<uint32_t>(), c->d_ready.as<uint32_t>() + off, s,
                                                       &c->launch_stats)) {
              c->stats.scan_launches += 1;
              return static_cast<int>(PD_OK);
            }
            const int r = CountTile(c, c->subset, t, ix, v0, out_dev, s);
            if (r != PD_OK) return r;
            pdk::launch_hwe(out_dev + off, t.v1 - t.v0, j->midp, p_dev, c->d_hweq.as<uint32_t>(), s, &c->launch_stats);
            return static_cast<int>(PD_OK);
          },
          [&](const Tile& t, cudaEvent_t ev) {
            if (!pending.empty()) {
              // malformed-input detection is only final at the end; rows are provisional until then
              publish(pending.back());
              pending.pop_back();
            }
            pending.push_back(Pending{t, ev});
          });
      if (rc == PD_OK && !pending.empty()) publish(pending.back());
    }
    {
      std::lock_guard<std::mutex> l2(j->mu);
      j->status = rc;
      j->finished = true;
      if (rc == PD_OK) j->rows_ready = j->v_end - j->v_begin;
    }
    j->cv.notify_all();
  });
  {
    std::unique_lock<std::mutex> lk(j->mu);
    j->cv.wait(lk, [&] { return j->started; });
  }
  *out = j;
  return PD_OK;
}

// Blocks until rows [v_begin, v_end) of the job are final, then copies them out (either output may be NULL).
static int FetchRows(pd_job* j, uint32_t v_begin, uint32_t v_end, uint32_t* out_counts, double* out_p) {
  if (!j || v_begin > v_end || v_begin < j->v_begin || v_end > j->v_end) return PD_ERR_INVALID_ARG;
  if (v_begin == v_end) return PD_OK;
  if (out_p && !j->hardy) return PD_ERR_INVALID_ARG;
  if (!j->subs.empty() || (j->ctx && j->ctx->group)) {
    for (pd_job* sub : j->subs) {
      const uint32_t a = std::max(v_begin, sub->v_begin), b = std::min(v_end, sub->v_end);
      if (a >= b) continue;
      const int rc = FetchRows(sub, a, b, out_counts ? out_counts + 4ull * (a - v_begin) : nullptr,
                               out_p ? out_p + (a - v_begin) : nullptr);
      if (rc != PD_OK) {
        j->ctx->last_error = sub->ctx->last_error;
        return rc;
      }
    }
    return PD_OK;
  }
  std::unique_lock<std::mutex> lock(j->mu);
  const uint32_t need = v_end - j->v_begin;
  j->cv.wait(lock, [&]() { return j->finished || j->rows_ready >= need; });
  if (j->status != PD_OK) return j->status;
  if (j->rows_ready < need) return PD_ERR_READ;
  if (out_counts) {
    memcpy(out_counts, j->counts.data() + static_cast<size_t>(v_begin - j->v_begin) * 4,
           static_cast<size_t>(v_end - v_begin) * 16);
  }
  if (out_p) memcpy(out_p, j->pvals.data() + (v_begin - j->v_begin), static_cast<size_t>(v_end - v_begin) * 8);
  return PD_OK;
}

extern "C" {

int pd_counts_start(pd_ctx* c, uint32_t v_begin, uint32_t v_end, pd_job** out) {
  return StartScanJob(c, v_begin, v_end, false, 0, out);
}

int pd_counts_fetch(pd_job* j, uint32_t v_begin, uint32_t v_end, uint32_t* out_counts) {
  if (!out_counts && v_end > v_begin) return PD_ERR_INVALID_ARG;
  return FetchRows(j, v_begin, v_end, out_counts, nullptr);
}

int pd_hardy_start(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t midp, pd_job** out) {
  return StartScanJob(c, v_begin, v_end, true, midp, out);
}

int pd_hardy_fetch(pd_job* j, uint32_t v_begin, uint32_t v_end, uint32_t* out_counts, double* out_p) {
  if (!out_p && v_end > v_begin) return PD_ERR_INVALID_ARG;
  return FetchRows(j, v_begin, v_end, out_counts, out_p);
}

void pd_job_free(pd_job* j) {
  if (!j) return;
  if (j->decode) FreeDecodeJob(j);
  for (pd_job* sub : j->subs) pd_job_free(sub);
  if (j->worker.joinable()) j->worker.join();
  delete j;
}

int pd_hardy(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t midp, uint32_t* out_counts, double* out_p) {
  if (!c || (!out_p && v_end > v_begin)) return PD_ERR_INVALID_ARG;
  if (c->group) {
    return GroupRunRanges(c, v_begin, v_end, 1, [&](pd_ctx* s, uint32_t a, uint32_t b, uint32_t row) {
      return pd_hardy(s, a, b, midp, out_counts ? out_counts + 4ull * row : nullptr, out_p + row);
    });
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  cudaEventRecord(c->ev_c0, c->stream);
  if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
  const uint32_t n = v_end - v_begin;
  cudaError_t e = c->d_pvals.Reserve(std::max<size_t>(8, static_cast<size_t>(n) * 8));
  if (e != cudaSuccess) return CheckCuda(c, e, "p-value allocation");
  uint32_t v_ext = v_begin;
  bool fused = false;
  bool direct = false;  // results already stored to the caller's arrays by the kernel
  int rc = PD_OK;
  if (c->idx.fixed && n) {
    // K1h: counts and exact tests of a tile in one kernel (fixed-width records have no LD bases,
    // so rows start at v_begin).
    e = c->d_counts.Reserve(static_cast<size_t>(n) * 16);
    if (e == cudaSuccess) e = c->d_ready.Reserve(static_cast<size_t>(n) * 4);
    if (e == cudaSuccess) e = c->d_hweq.Reserve((static_cast<size_t>(n) + pdk::kHweQueueWords) * 4);
    if (e != cudaSuccess) return CheckCuda(c, e, "counts allocation");
    const pdk::SubsetDev sd = c->subset.Dev();
    uint4* out_dev = c->d_counts.as<uint4>();
    double* p_dev = c->d_pvals.as<double>();
    uint32_t* ready = c->d_ready.as<uint32_t>();
    // Page-locked caller arrays are written by the kernel itself (mapped stores over PCIe while the scan
    // runs) instead of by two copies after it: 24 B per variant that no longer trail the last tile.
    uint4* host_counts = nullptr;
    if (pdk::hardy_fused_applicable(c->idx.sample_ct) && IsPinnedHost(out_p) && (!out_counts || IsPinnedHost(out_counts))) {
      void *dp = nullptr, *dc = nullptr;
      if (cudaHostGetDevicePointer(&dp, out_p, 0) == cudaSuccess &&
          (!out_counts || cudaHostGetDevicePointer(&dc, out_counts, 0) == cudaSuccess)) {
        p_dev = static_cast<double*>(dp);
        host_counts = static_cast<uint4*>(dc);
        direct = true;
      } else {
        cudaGetLastError();
      }
    }
    bool all_fused = true;
    rc = RunScan(c, v_begin, v_end, &v_ext, [&](const Tile& t, const pdk::RecIndex& ix, cudaStream_t s) {
      const uint32_t off = t.v0 - v_begin;
      if (!pdk::launch_hardy_fixed(ix, sd, t.v0, t.v1, out_dev + off, p_dev + off, midp, c->d_hweq.as<uint32_t>(),
                                   ready + off, s, &c->launch_stats, host_counts ? host_counts + off : nullptr)) {
        all_fused = false;
        pdk::launch_count_fixed(ix, sd, t.v0, t.v1, out_dev + off, s, &c->launch_stats);
      }
      c->stats.scan_launches += 1;
      return PD_OK;
    });
    if (rc != PD_OK) return rc;
    fused = all_fused;
  } else {
    rc = CountsToDevice(c, c->subset, v_begin, v_end, &v_ext);
    if (rc != PD_OK) return rc;
  }
  if (!n) return PD_OK;
  const uint4* counts = c->d_counts.as<uint4>() + (v_begin - v_ext);
  if (!fused) {
    e = c->d_hweq.Reserve((static_cast<size_t>(n) + pdk::kHweQueueWords) * 4);
    if (e != cudaSuccess) return CheckCuda(c, e, "exact-test queue allocation");
  }
  cudaEventRecord(c->ev_a, c->stream);
  if (!fused) pdk::launch_hwe(counts, n, midp, c->d_pvals.as<double>(), c->d_hweq.as<uint32_t>(), c->stream, &c->launch_stats);
  cudaEventRecord(c->ev_b, c->stream);
  e = cudaSuccess;
  if (!(direct && fused)) {
    e = cudaMemcpyAsync(out_p, c->d_pvals.p, static_cast<size_t>(n) * 8, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess && out_counts) {
      e = cudaMemcpyAsync(out_counts, counts, static_cast<size_t>(n) * 16, cudaMemcpyDeviceToHost, c->stream);
    }
  }
  cudaEventRecord(c->ev_c1, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "result download");
  c->stats.d2h_bytes += static_cast<size_t>(n) * (out_counts ? 24 : 8);
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev_a, c->ev_b);
  c->stats.aux_kernel_ms = fused ? 0.0 : ms;  // fused: the exact tests ran inside the scan kernel
  cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
  c->stats.call_ms = ms;
  return PD_OK;
}

static int HweFromCounts(pd_ctx* c, const uint32_t* counts, uint32_t n, uint32_t midp, double* out, bool want_lnp) {
  if (!c || (n && (!counts || !out))) return PD_ERR_INVALID_ARG;
  if (!n) return PD_OK;
  if (c->group) return HweFromCounts(c->group->shards[0], counts, n, midp, out, want_lnp);  // no file data involved
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  cudaError_t e = c->d_counts.Reserve(static_cast<size_t>(n) * 16);
  if (e == cudaSuccess) e = c->d_pvals.Reserve(static_cast<size_t>(n) * 8);
  if (e == cudaSuccess) e = c->d_hweq.Reserve((static_cast<size_t>(n) + pdk::kHweQueueWords) * 4);
  if (e != cudaSuccess) return CheckCuda(c, e, "allocation");
  e = cudaMemcpyAsync(c->d_counts.p, counts, static_cast<size_t>(n) * 16, cudaMemcpyHostToDevice, c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "counts upload");
  c->stats.h2d_bytes += static_cast<size_t>(n) * 16;
  cudaEventRecord(c->ev_a, c->stream);
  pdk::launch_hwe(c->d_counts.as<uint4>(), n, midp, c->d_pvals.as<double>(), c->d_hweq.as<uint32_t>(), c->stream,
                  &c->launch_stats, want_lnp);
  cudaEventRecord(c->ev_b, c->stream);
  int rc = DownloadRows(c, c->d_pvals.p, static_cast<size_t>(n) * 8, out);
  if (rc == PD_OK) {
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_a, c->ev_b);
    c->stats.aux_kernel_ms = ms;
  }
  return rc;
}

int pd_hwe_from_counts(pd_ctx* c, const uint32_t* counts, uint32_t n, uint32_t midp, double* out_p) {
  return HweFromCounts(c, counts, n, midp, out_p, false);
}

int pd_hwe_lnp_from_counts(pd_ctx* c, const uint32_t* counts, uint32_t n, uint32_t midp, double* out_lnp) {
  return HweFromCounts(c, counts, n, midp, out_lnp, true);
}

int pd_hwe_xchr_from_counts(pd_ctx* c, const uint32_t* c5, uint32_t n, uint32_t midp, double* out_p) {
  if (!c || (n && (!c5 || !out_p))) return PD_ERR_INVALID_ARG;
  if (!n) return PD_OK;
  if (c->group) return pd_hwe_xchr_from_counts(c->group->shards[0], c5, n, midp, out_p);
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  cudaError_t e = c->d_counts.Reserve(static_cast<size_t>(n) * 20);
  if (e == cudaSuccess) e = c->d_pvals.Reserve(static_cast<size_t>(n) * 8);
  if (e != cudaSuccess) return CheckCuda(c, e, "allocation");
  e = cudaMemcpyAsync(c->d_counts.p, c5, static_cast<size_t>(n) * 20, cudaMemcpyHostToDevice, c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "counts upload");
  c->stats.h2d_bytes += static_cast<size_t>(n) * 20;
  cudaEventRecord(c->ev_a, c->stream);
  pdk::launch_hwe_xchr(c->d_counts.as<uint32_t>(), n, midp, c->d_pvals.as<double>(), c->stream, &c->launch_stats);
  cudaEventRecord(c->ev_b, c->stream);
  int rc = DownloadRows(c, c->d_pvals.p, static_cast<size_t>(n) * 8, out_p);
  if (rc == PD_OK) {
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_a, c->ev_b);
    c->stats.aux_kernel_ms = ms;
  }
  return rc;
}

}  // extern "C"
