// Sample-wise consumers of the decoded genotype rows: int8 decode (read_pgen / read_pfile) and the
// per-sample accumulations of plink_missing (sample mode) and read_pfile(orient := 'sample').
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <thread>

#include "ctx.h"

using namespace pdx;

namespace pdx {

bool IsPinnedHost(const void* p) {
  cudaPointerAttributes attr{};
  if (cudaPointerGetAttributes(&attr, p) == cudaSuccess && attr.type == cudaMemoryTypeHost) return true;
  cudaGetLastError();
  return false;
}

// memcpy split over a few threads: one thread moves ~10 GB/s, PCIe delivers ~55 GB/s.
void ParallelMemcpy(void* dst, const void* src, size_t n) {
  const size_t kChunk = 8u << 20;
  const unsigned nthreads = static_cast<unsigned>(std::min<size_t>(8, (n + kChunk - 1) / kChunk));
  if (nthreads <= 1) {
    memcpy(dst, src, n);
    return;
  }
  std::atomic<size_t> next(0);
  auto worker = [&]() {
    for (;;) {
      const size_t o = next.fetch_add(kChunk);
      if (o >= n) break;
      memcpy(static_cast<char*>(dst) + o, static_cast<const char*>(src) + o, std::min(kChunk, n - o));
    }
  };
  std::vector<std::thread> pool;
  for (unsigned t = 1; t < nthreads; ++t) pool.emplace_back(worker);
  worker();
  for (auto& th : pool) th.join();
}

// Double-buffered device -> host streaming of per-batch outputs.  Batch i is produced into
// d_out[i & 1] on the compute stream and copied out on the D2H stream while batch i + 1 is being
// produced.  Pageable destinations go through a pinned bounce buffer.
struct OutPipe {
  pd_ctx_impl* c;
  bool direct;          // destination is page-locked: DMA straight into it
  size_t buf_bytes;
  struct Pending {
    void* dst = nullptr;
    size_t bytes = 0;
    bool live = false;
  } pend[2];
  uint32_t next = 0;

  int Init(pd_ctx_impl* ctx, const void* user_dst, size_t bytes_per_batch) {
    c = ctx;
    direct = IsPinnedHost(user_dst);
    buf_bytes = bytes_per_batch;
    for (int b = 0; b < 2; ++b) {
      cudaError_t e = c->d_out[b].Reserve(bytes_per_batch + 64);
      if (e == cudaSuccess && !direct) e = c->h_bounce[b].Reserve(bytes_per_batch);
      if (e != cudaSuccess) return CheckCuda(c, e, "output buffer allocation");
      pend[b] = Pending{};
    }
    return PD_OK;
  }
  // Finishes the host side of whatever was last sent from buffer b.
  int Retire(int b) {
    if (!pend[b].live) return PD_OK;
    cudaError_t e = cudaEventSynchronize(c->ev_out_free[b]);
    if (e != cudaSuccess) return CheckCuda(c, e, "result download");
    if (!direct) ParallelMemcpy(pend[b].dst, c->h_bounce[b].p, pend[b].bytes);
    pend[b].live = false;
    return PD_OK;
  }
  // Buffer the next batch should be produced into (its previous contents are safely out).
  int Acquire(void** dev, int* slot) {
    const int b = static_cast<int>(next++ & 1);
    const int rc = Retire(b);
    if (rc != PD_OK) return rc;
    *dev = c->d_out[b].p;
    *slot = b;
    return PD_OK;
  }
  // Call after the producing kernels of slot b were enqueued on stream s.
  int Send(int b, cudaStream_t s, void* host_dst, size_t bytes) {
    cudaEventRecord(c->ev_out_ready[b], s);
    cudaStreamWaitEvent(c->d2h_stream, c->ev_out_ready[b], 0);
    void* dst = direct ? host_dst : c->h_bounce[b].p;
    cudaError_t e = cudaMemcpyAsync(dst, c->d_out[b].p, bytes, cudaMemcpyDeviceToHost, c->d2h_stream);
    if (e != cudaSuccess) return CheckCuda(c, e, "result download");
    cudaEventRecord(c->ev_out_free[b], c->d2h_stream);  // Acquire() waits on it before slot b is reused
    pend[b].dst = host_dst;
    pend[b].bytes = bytes;
    pend[b].live = true;
    c->stats.d2h_bytes += bytes;
    return PD_OK;
  }
  int Finish() {
    int rc = Retire(0);
    const int rc2 = Retire(1);
    return rc != PD_OK ? rc : rc2;
  }
};

// Per-sample accumulation over [v_begin, v_end): kinds as pdk::launch_sample_acc.
// Results are left in d_out[0] as up to three arrays of n_out uint32 (het, hom_alt, missing).
int SampleAccumulate(pd_ctx_impl* c, uint32_t v_begin, uint32_t v_end, uint32_t kinds, uint32_t* n_out_p) {
  const uint32_t N = c->idx.sample_ct;
  const uint32_t n_out = c->subset.active ? c->subset.count : N;
  *n_out_p = n_out;
  const size_t npad = (static_cast<size_t>(N) + 63) & ~static_cast<size_t>(63);
  cudaError_t e = c->d_acc.Reserve(3 * npad * 4);
  if (e == cudaSuccess) e = c->d_out[0].Reserve(3 * static_cast<size_t>(std::max(n_out, 1u)) * 4 + 64);
  if (e != cudaSuccess) return CheckCuda(c, e, "accumulator allocation");
  uint32_t* acc = c->d_acc.as<uint32_t>();
  cudaMemsetAsync(acc, 0, 3 * npad * 4, c->stream);
  int rc = RunRows(c, v_begin, v_end, 0, 256ull << 20, [&](const RowBatch& b, const pdk::RecIndex&, cudaStream_t s) {
    pdk::launch_sample_acc(b.rows, kinds, acc, acc + npad, acc + 2 * npad, s, &c->launch_stats);
    c->stats.scan_launches += 1;
    return PD_OK;
  });
  if (rc != PD_OK) return rc;
  const uint32_t* idx = c->subset.active ? c->subset.d_idx.as<uint32_t>() : nullptr;
  uint32_t* out = c->d_out[0].as<uint32_t>();
  for (uint32_t k = 0; k < 3; ++k) {
    if (kinds & (1u << k)) {
      pdk::launch_gather_u32(acc + k * npad, idx, n_out, out + static_cast<size_t>(k) * n_out, c->stream, &c->launch_stats);
    }
  }
  return PD_OK;
}

}  // namespace pdx

extern "C" {

int pd_missing_by_sample(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t* out_missing) {
  if (!c || !out_missing) return PD_ERR_INVALID_ARG;
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  cudaEventRecord(c->ev_c0, c->stream);
  uint32_t n_out = 0;
  int rc = SampleAccumulate(c, v_begin, v_end, 4, &n_out);
  if (rc != PD_OK) return rc;
  cudaError_t e = cudaMemcpyAsync(out_missing, c->d_out[0].as<uint32_t>() + 2 * static_cast<size_t>(n_out),
                                  static_cast<size_t>(n_out) * 4, cudaMemcpyDeviceToHost, c->stream);
  cudaEventRecord(c->ev_c1, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "result download");
  c->stats.d2h_bytes += static_cast<size_t>(n_out) * 4;
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
  c->stats.call_ms = ms;
  return PD_OK;
}

int pd_sample_counts(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t* out_het, uint32_t* out_hom_alt,
                     uint32_t* out_missing) {
  if (!c || !out_het || !out_hom_alt || !out_missing) return PD_ERR_INVALID_ARG;
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  cudaEventRecord(c->ev_c0, c->stream);
  uint32_t n_out = 0;
  int rc = SampleAccumulate(c, v_begin, v_end, 7, &n_out);
  if (rc != PD_OK) return rc;
  const uint32_t* d = c->d_out[0].as<uint32_t>();
  const size_t bytes = static_cast<size_t>(n_out) * 4;
  cudaError_t e = cudaMemcpyAsync(out_het, d, bytes, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(out_hom_alt, d + n_out, bytes, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(out_missing, d + 2 * static_cast<size_t>(n_out), bytes, cudaMemcpyDeviceToHost, c->stream);
  cudaEventRecord(c->ev_c1, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "result download");
  c->stats.d2h_bytes += 3 * bytes;
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
  c->stats.call_ms = ms;
  return PD_OK;
}

int pd_decode_i8(pd_ctx* c, uint32_t v_begin, uint32_t v_end, int8_t* out) {
  if (!c || (!out && v_end > v_begin)) return PD_ERR_INVALID_ARG;
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  const uint32_t n_out = c->subset.active ? c->subset.count : c->idx.sample_ct;
  if (v_begin >= v_end || !n_out) {
    if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
    return PD_OK;
  }
  // rows per batch: 128 MiB of int8 output per buffer
  const uint64_t kOutBytes = 128ull << 20;
  const uint32_t max_rows = static_cast<uint32_t>(std::max<uint64_t>(1, std::min<uint64_t>(65536, kOutBytes / n_out)));
  OutPipe pipe;
  int rc = pipe.Init(c, out, static_cast<size_t>(max_rows) * n_out);
  if (rc != PD_OK) return rc;
  const uint32_t* idx = c->subset.active ? c->subset.d_idx.as<uint32_t>() : nullptr;
  cudaEventRecord(c->ev_c0, c->stream);
  rc = RunRows(c, v_begin, v_end, max_rows, 0, [&](const RowBatch& b, const pdk::RecIndex&, cudaStream_t s) {
    void* dev = nullptr;
    int slot = 0;
    int r = pipe.Acquire(&dev, &slot);
    if (r != PD_OK) return r;
    pdk::launch_expand_i8(b.rows, idx, n_out, static_cast<int8_t*>(dev), s, &c->launch_stats);
    c->stats.scan_launches += 1;
    return pipe.Send(slot, s, out + static_cast<size_t>(b.row0) * n_out, static_cast<size_t>(b.n) * n_out);
  });
  const int rc2 = pipe.Finish();
  cudaEventRecord(c->ev_c1, c->stream);
  cudaEventSynchronize(c->ev_c1);
  if (rc == PD_OK && rc2 == PD_OK) {
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
    c->stats.call_ms = ms;
  }
  return rc != PD_OK ? rc : rc2;
}

}  // extern "C"
