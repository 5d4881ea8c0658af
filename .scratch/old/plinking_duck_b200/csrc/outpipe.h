// Device -> host result streaming shared by the decode entry points.
#pragma once
#include "ctx.h"

namespace pdx {

bool IsPinnedHost(const void* p);
void ParallelMemcpy(void* dst, const void* src, size_t n);

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
    void* dst2 = nullptr;  // optional second region of the same buffer (validity words of pd_decode_i8_valid)
    size_t off2 = 0, bytes2 = 0;
    bool live = false;
  } pend[2];
  uint32_t next = 0;
  double produce_ms = 0;  // CUDA-event time of the producing kernels, summed over batches

  int Init(pd_ctx_impl* ctx, const void* user_dst, size_t bytes_per_batch, const void* user_dst2 = nullptr) {
    c = ctx;
    direct = IsPinnedHost(user_dst) && (!user_dst2 || IsPinnedHost(user_dst2));
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
    if (!direct) {
      ParallelMemcpy(pend[b].dst, c->h_bounce[b].p, pend[b].bytes);
      if (pend[b].bytes2) memcpy(pend[b].dst2, static_cast<char*>(c->h_bounce[b].p) + pend[b].off2, pend[b].bytes2);
    }
    float ms = 0;
    if (cudaEventElapsedTime(&ms, c->ev_prod0[b], c->ev_prod1[b]) == cudaSuccess) produce_ms += ms;
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
    cudaEventRecord(c->ev_prod0[b], c->stream);
    return PD_OK;
  }
  // Call after the producing kernels of slot b were enqueued on stream s.
  int Send(int b, cudaStream_t s, void* host_dst, size_t bytes, void* host_dst2 = nullptr, size_t off2 = 0,
           size_t bytes2 = 0) {
    cudaEventRecord(c->ev_prod1[b], s);
    cudaEventRecord(c->ev_out_ready[b], s);
    cudaStreamWaitEvent(c->d2h_stream, c->ev_out_ready[b], 0);
    void* dst = direct ? host_dst : c->h_bounce[b].p;
    cudaError_t e = cudaMemcpyAsync(dst, c->d_out[b].p, bytes, cudaMemcpyDeviceToHost, c->d2h_stream);
    if (e == cudaSuccess && bytes2) {
      void* d2 = direct ? host_dst2 : static_cast<char*>(c->h_bounce[b].p) + off2;
      e = cudaMemcpyAsync(d2, static_cast<char*>(c->d_out[b].p) + off2, bytes2, cudaMemcpyDeviceToHost, c->d2h_stream);
    }
    if (e != cudaSuccess) return CheckCuda(c, e, "result download");
    cudaEventRecord(c->ev_out_free[b], c->d2h_stream);  // Acquire() waits on it before slot b is reused
    pend[b].dst = host_dst;
    pend[b].bytes = bytes;
    pend[b].dst2 = host_dst2;
    pend[b].off2 = off2;
    pend[b].bytes2 = bytes2;
    pend[b].live = true;
    c->stats.d2h_bytes += bytes + bytes2;
    return PD_OK;
  }
  int Finish() {
    // oldest first: two batches of pd_decode_i8_valid may share a validity word, the later one holds its final value
    const int older = static_cast<int>(next & 1);
    int rc = Retire(older);
    const int rc2 = Retire(older ^ 1);
    return rc != PD_OK ? rc : rc2;
  }
};

}  // namespace pdx
