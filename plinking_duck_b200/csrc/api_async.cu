// Asynchronous genotype decode: pd_decode_start / pd_decode_fetch.
//
// read_pgen / read_pfile scan threads claim batches of variants from an atomic counter and emit them
// (src/pgen_reader.cpp:635-736, src/pfile_reader.cpp:2790-2835).  A decode job runs ahead of them: a job thread
// decodes the range chunk by chunk (the blocking pd_decode_i8 pipeline: staging, K3b, K5, D2H into page-locked
// chunk buffers) while the scan threads copy the rows they claimed out of finished chunks.  The ring of chunk
// buffers bounds the memory: a chunk's buffer is reused once every row of it has been fetched.
#include <algorithm>
#include <cstring>
#include <new>

#include "ctx.h"
#include "group.h"

using namespace pdx;

namespace pdx {

struct DecodeJob {
  static constexpr uint32_t kSlots = 4;
  uint32_t n_out = 0;          // samples per row
  uint32_t chunk_rows = 0;
  uint32_t n_chunks = 0;
  int8_t* buf[kSlots] = {nullptr, nullptr, nullptr, nullptr};  // page-locked, chunk_rows x n_out each
  // per slot: which chunk it holds (or ~0u), whether it is complete, rows not yet fetched
  uint32_t chunk_of[kSlots] = {~0u, ~0u, ~0u, ~0u};
  bool ready[kSlots] = {false, false, false, false};
  uint32_t unfetched[kSlots] = {0, 0, 0, 0};
  bool cancel = false;
};

}  // namespace pdx

namespace {

void DecodeWorker(pd_job* j) {
  DecodeJob* d = j->decode;
  pd_ctx* c = j->ctx;
  int rc = PD_OK;
  for (uint32_t k = 0; k < d->n_chunks && rc == PD_OK; ++k) {
    const uint32_t slot = k % DecodeJob::kSlots;
    const uint32_t a = j->v_begin + k * d->chunk_rows;
    const uint32_t b = std::min(j->v_end, a + d->chunk_rows);
    {
      std::unique_lock<std::mutex> lk(j->mu);
      j->cv.wait(lk, [&] { return d->cancel || d->chunk_of[slot] == ~0u; });  // every row of the old chunk was fetched
      if (d->cancel) break;
      d->chunk_of[slot] = k;
      d->ready[slot] = false;
      d->unfetched[slot] = b - a;
    }
    rc = pd_decode_i8(c, a, b, d->buf[slot]);
    {
      std::lock_guard<std::mutex> lk(j->mu);
      if (rc == PD_OK) d->ready[slot] = true;
    }
    j->cv.notify_all();
  }
  {
    std::lock_guard<std::mutex> lk(j->mu);
    j->status = rc;
    j->finished = true;
  }
  j->cv.notify_all();
}

}  // namespace

void pdx::FreeDecodeJob(pd_job* j) {
  DecodeJob* d = j->decode;
  if (!d) return;
  {
    std::lock_guard<std::mutex> lk(j->mu);
    d->cancel = true;
  }
  j->cv.notify_all();
  if (j->worker.joinable()) j->worker.join();
  for (int8_t* b : d->buf) {
    if (b) cudaFreeHost(b);
  }
  delete d;
  j->decode = nullptr;
}

extern "C" {

int pd_decode_start(pd_ctx* c, uint32_t v_begin, uint32_t v_end, pd_job** out) {
  if (!c || !out) return PD_ERR_INVALID_ARG;
  *out = nullptr;
  pd_info_t info;
  int rc = pd_info(c, &info);
  if (rc != PD_OK) return rc;
  if (v_begin > v_end || v_end > info.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
  pd_job* j = new (std::nothrow) pd_job();
  DecodeJob* d = new (std::nothrow) DecodeJob();
  if (!j || !d) {
    delete j;
    delete d;
    return PD_ERR_NOMEM;
  }
  j->ctx = c;
  j->v_begin = v_begin;
  j->v_end = v_end;
  j->decode = d;
  d->n_out = info.effective_sample_ct;
  // 64 MiB chunks, a multiple of 128 rows (the claim size of the reference's scans), at least 128
  const uint64_t per = std::max<uint64_t>(1, d->n_out);
  d->chunk_rows = static_cast<uint32_t>(std::max<uint64_t>(128, (64ull << 20) / per / 128 * 128));
  const uint32_t rows = v_end - v_begin;
  d->chunk_rows = std::min(d->chunk_rows, std::max(rows, 1u));
  d->n_chunks = rows ? (rows + d->chunk_rows - 1) / d->chunk_rows : 0;
  const uint32_t nslots = std::min<uint32_t>(DecodeJob::kSlots, std::max(d->n_chunks, 1u));
  for (uint32_t s = 0; s < nslots; ++s) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, std::max<size_t>(1, static_cast<size_t>(d->chunk_rows) * d->n_out), cudaHostAllocPortable) != cudaSuccess) {
      cudaGetLastError();
      FreeDecodeJob(j);
      delete j;
      return c->SetError(PD_ERR_NOMEM, "cannot page-lock the decode chunk buffers");
    }
    d->buf[s] = static_cast<int8_t*>(p);
  }
  j->worker = std::thread(DecodeWorker, j);
  *out = j;
  return PD_OK;
}

int pd_decode_fetch(pd_job* j, uint32_t v_begin, uint32_t v_end, int8_t* out) {
  if (!j || !j->decode || v_begin > v_end || v_begin < j->v_begin || v_end > j->v_end) return PD_ERR_INVALID_ARG;
  if (v_begin == v_end) return PD_OK;
  if (!out) return PD_ERR_INVALID_ARG;
  DecodeJob* d = j->decode;
  uint32_t v = v_begin;
  while (v < v_end) {
    const uint32_t k = (v - j->v_begin) / d->chunk_rows;
    const uint32_t slot = k % DecodeJob::kSlots;
    const uint32_t c0 = j->v_begin + k * d->chunk_rows;
    const uint32_t e = std::min(v_end, c0 + d->chunk_rows);
    {
      std::unique_lock<std::mutex> lk(j->mu);
      j->cv.wait(lk, [&] { return j->finished || (d->chunk_of[slot] == k && d->ready[slot]); });
      if (!(d->chunk_of[slot] == k && d->ready[slot])) {
        if (j->status != PD_OK) {
          return j->status;
        }
        return PD_ERR_INVALID_ARG;  // rows fetched twice, or a job that was cancelled
      }
    }
    memcpy(out + static_cast<size_t>(v - v_begin) * d->n_out, d->buf[slot] + static_cast<size_t>(v - c0) * d->n_out,
           static_cast<size_t>(e - v) * d->n_out);
    {
      std::lock_guard<std::mutex> lk(j->mu);
      d->unfetched[slot] -= e - v;
      if (d->unfetched[slot] == 0) {
        d->chunk_of[slot] = ~0u;
        d->ready[slot] = false;
      }
    }
    j->cv.notify_all();
    v = e;
  }
  return PD_OK;
}

}  // extern "C"
