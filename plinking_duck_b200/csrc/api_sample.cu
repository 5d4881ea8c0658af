// Sample-wise consumers of the decoded genotype rows: int8 decode (read_pgen / read_pfile) and the
// per-sample accumulations of plink_missing (sample mode) and read_pfile(orient := 'sample').
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <thread>

#include "ctx.h"
#include "group.h"
#include "host_pool.h"
#include "outpipe.h"

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
  const size_t kChunk = 4u << 20;
  const size_t nchunks = (n + kChunk - 1) / kChunk;
  if (nchunks <= 1) {
    memcpy(dst, src, n);
    return;
  }
  HostPool::Get().ParallelFor(nchunks, [&](uint64_t i) {
    const size_t o = static_cast<size_t>(i) * kChunk;
    memcpy(static_cast<char*>(dst) + o, static_cast<const char*>(src) + o, std::min(kChunk, n - o));
  });
}

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
  // Variable-width files: difflist records are walked sparsely inside the densify launch and only
  // the raw / 1-bit / LD records reach launch_sample_acc (pdk::SparseAcc).
  uint32_t* sw = DevSparseWords(c);
  cudaMemsetAsync(sw, 0, 4 * sizeof(uint32_t), c->stream);
  const pdk::SparseAcc sparse{kinds, acc, acc + npad, acc + 2 * npad, sw + 1, sw};
  struct Scope {
    pd_ctx_impl* c;
    ~Scope() { c->sparse_acc = nullptr; }
  } scope{c};
  c->sparse_acc = c->idx.fixed ? nullptr : &sparse;
  const uint32_t* base = c->idx.fixed ? nullptr : sw + 1;
  int rc = RunRows(c, v_begin, v_end, 0, 256ull << 20, [&](const RowBatch& b, const pdk::RecIndex&, cudaStream_t s) {
    pdk::launch_sample_acc(b.rows, kinds, acc, acc + npad, acc + 2 * npad, s, &c->launch_stats);
    c->stats.scan_launches += 1;
    return PD_OK;
  });
  if (rc != PD_OK) return rc;
  const uint32_t* idx = c->subset.active ? c->subset.d_idx.as<uint32_t>() : nullptr;
  uint32_t* out = c->d_out[0].as<uint32_t>();
  cudaMemsetAsync(out, 0, 3 * static_cast<size_t>(n_out) * 4, c->stream);
  if (kinds == 7) {
    pdk::launch_planes_to_counts(acc, acc + npad, acc + 2 * npad, idx, n_out, out, c->stream, &c->launch_stats, base);
  } else {
    pdk::launch_gather_u32(acc + 2 * npad, idx, n_out, out + 2 * static_cast<size_t>(n_out), c->stream, &c->launch_stats,
                           base ? base + 2 : nullptr);
  }
  return PD_OK;
}

}  // namespace pdx

extern "C" {

int pd_missing_by_sample(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t* out_missing) {
  if (!c) return PD_ERR_INVALID_ARG;
  if (c->group) return GroupMissingBySample(c, v_begin, v_end, out_missing);
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  cudaEventRecord(c->ev_c0, c->stream);
  uint32_t n_out = 0;
  int rc = SampleAccumulate(c, v_begin, v_end, 4, &n_out);
  if (rc != PD_OK) return rc;
  c->last_sample_n = n_out;
  cudaError_t e = cudaSuccess;
  if (out_missing) {
    e = cudaMemcpyAsync(out_missing, c->d_out[0].as<uint32_t>() + 2 * static_cast<size_t>(n_out),
                        static_cast<size_t>(n_out) * 4, cudaMemcpyDeviceToHost, c->stream);
  }
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
  if (!c) return PD_ERR_INVALID_ARG;
  const bool want = out_het || out_hom_alt || out_missing;
  if (want && !(out_het && out_hom_alt && out_missing)) return PD_ERR_INVALID_ARG;
  if (c->group) return GroupSampleCounts(c, v_begin, v_end, out_het, out_hom_alt, out_missing);
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  cudaEventRecord(c->ev_c0, c->stream);
  uint32_t n_out = 0;
  int rc = SampleAccumulate(c, v_begin, v_end, 7, &n_out);
  if (rc != PD_OK) return rc;
  const uint32_t* d = c->d_out[0].as<uint32_t>();
  const size_t bytes = static_cast<size_t>(n_out) * 4;
  c->last_sample_n = n_out;
  cudaError_t e = cudaSuccess;
  if (want) {
    e = cudaMemcpyAsync(out_het, d, bytes, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_hom_alt, d + n_out, bytes, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_missing, d + 2 * static_cast<size_t>(n_out), bytes, cudaMemcpyDeviceToHost, c->stream);
  }
  cudaEventRecord(c->ev_c1, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "result download");
  c->stats.d2h_bytes += 3 * bytes;
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
  c->stats.call_ms = ms;
  return PD_OK;
}

// Samples per output row: the subset's size, else the raw sample count (a group asks its first shard).
static uint32_t EffectiveSamples(const pd_ctx* c) {
  if (c->group) c = c->group->shards[0];
  return c->subset.active ? c->subset.count : c->idx.sample_ct;
}

int pd_decode_i8(pd_ctx* c, uint32_t v_begin, uint32_t v_end, int8_t* out) {
  if (!c || (!out && v_end > v_begin)) return PD_ERR_INVALID_ARG;
  if (c->group) {
    const size_t n = EffectiveSamples(c);
    return GroupRunRanges(c, v_begin, v_end, 1, [&](pd_ctx* s, uint32_t a, uint32_t b, uint32_t row) {
      return pd_decode_i8(s, a, b, out + row * n);
    });
  }
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
    c->stats.kernel_ms = pipe.produce_ms;  // the decode kernels alone; call_ms includes the D2H stream
  }
  return rc != PD_OK ? rc : rc2;
}

}  // extern "C"

// Shared body of pd_decode_i8_valid / pd_decode_i8_valid_list: run(max_rows, fn) feeds the row batches.
static int DecodeI8Valid(pd_ctx* c, uint32_t n_out, uint32_t call_mask, int8_t* out, uint64_t* validity,
                         const std::function<int(uint32_t, const RowFn&)>& run) {
  // Per batch and buffer: up to 128 MiB of calls, then (256-byte aligned) their validity words + one word of
  // slack.  A batch starts at any bit of the caller's bitmap: the word it shares with the batch before is
  // carried over on the device, so every download writes whole, final 64-bit words.
  const uint64_t kOutBytes = 128ull << 20;
  const uint32_t max_rows = static_cast<uint32_t>(std::max<uint64_t>(1, std::min<uint64_t>(65535, kOutBytes / n_out)));
  const size_t data_cap = (static_cast<size_t>(max_rows) * n_out + 255) & ~static_cast<size_t>(255);
  const size_t valid_cap = (data_cap / 64 + 2) * 8;
  OutPipe pipe;
  int rc = pipe.Init(c, out, data_cap + valid_cap, validity);
  if (rc != PD_OK) return rc;
  const uint32_t* idx = c->subset.active ? c->subset.d_idx.as<uint32_t>() : nullptr;
  const char* prev_last_word = nullptr;  // device address of the last validity word of the batch before
  cudaEventRecord(c->ev_c0, c->stream);
  rc = run(max_rows, [&](const RowBatch& b, const pdk::RecIndex&, cudaStream_t s) {
    void* dev = nullptr;
    int slot = 0;
    int r = pipe.Acquire(&dev, &slot);
    if (r != PD_OK) return r;
    const uint64_t bit0 = static_cast<uint64_t>(b.row0) * n_out, bit1 = bit0 + static_cast<uint64_t>(b.n) * n_out;
    const uint64_t w0 = bit0 >> 6, w1 = (bit1 + 63) >> 6;
    char* vdev = static_cast<char*>(dev) + data_cap;
    cudaMemsetAsync(vdev, 0, (w1 - w0 + 1) * 8, s);
    if ((bit0 & 63u) && prev_last_word) cudaMemcpyAsync(vdev, prev_last_word, 8, cudaMemcpyDeviceToDevice, s);
    pdk::launch_expand_i8_valid(b.rows, idx, n_out, call_mask, static_cast<int8_t*>(dev), reinterpret_cast<uint32_t*>(vdev),
                                static_cast<uint32_t>(bit0 & 63u), s, &c->launch_stats);
    prev_last_word = vdev + (w1 - w0 - 1) * 8;
    c->stats.scan_launches += 1;
    return pipe.Send(slot, s, out + static_cast<size_t>(b.row0) * n_out, static_cast<size_t>(b.n) * n_out, validity + w0,
                     data_cap, (w1 - w0) * 8);
  });
  const int rc2 = pipe.Finish();
  cudaEventRecord(c->ev_c1, c->stream);
  cudaEventSynchronize(c->ev_c1);
  if (rc == PD_OK && rc2 == PD_OK) {
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
    c->stats.call_ms = ms;
    c->stats.kernel_ms = pipe.produce_ms;
  }
  return rc != PD_OK ? rc : rc2;
}

extern "C" {

int pd_decode_i8_valid(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t call_mask, int8_t* out, uint64_t* validity) {
  if (!c || ((!out || !validity) && v_end > v_begin)) return PD_ERR_INVALID_ARG;
  if (c->group) {
    // pieces start at multiples of 64 rows, so no 64-bit validity word is shared between two GPUs
    const size_t n = EffectiveSamples(c);
    return GroupRunRanges(c, v_begin, v_end, 64, [&](pd_ctx* s, uint32_t a, uint32_t b, uint32_t row) {
      return pd_decode_i8_valid(s, a, b, call_mask, out + row * n, validity + row * n / 64);
    });
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  const uint32_t n_out = c->subset.active ? c->subset.count : c->idx.sample_ct;
  if (v_begin >= v_end || !n_out) {
    if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
    return PD_OK;
  }
  return DecodeI8Valid(c, n_out, call_mask, out, validity, [&](uint32_t max_rows, const RowFn& fn) {
    return RunRows(c, v_begin, v_end, max_rows, 0, fn);
  });
}

int pd_decode_i8_valid_list(pd_ctx* c, const uint32_t* vidx, uint32_t n_var, uint32_t call_mask, int8_t* out,
                            uint64_t* validity) {
  if (!c || ((!out || !validity || !vidx) && n_var)) return PD_ERR_INVALID_ARG;
  if (c->group) {
    const size_t n = EffectiveSamples(c);
    return GroupRunList(c, vidx, n_var, 64, [&](pd_ctx* s, uint32_t i0, uint32_t i1) {
      return pd_decode_i8_valid_list(s, vidx + i0, i1 - i0, call_mask, out + i0 * n, validity + i0 * n / 64);
    });
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  const uint32_t n_out = c->subset.active ? c->subset.count : c->idx.sample_ct;
  if (!n_var || !n_out) return PD_OK;
  return DecodeI8Valid(c, n_out, call_mask, out, validity, [&](uint32_t max_rows, const RowFn& fn) {
    return RunRowsList(c, vidx, n_var, max_rows, fn);
  });
}

int pd_decode_phased(pd_ctx* c, uint32_t v_begin, uint32_t v_end, int8_t* out) {
  if (!c || (!out && v_end > v_begin)) return PD_ERR_INVALID_ARG;
  if (c->group) {
    const size_t n = EffectiveSamples(c);
    return GroupRunRanges(c, v_begin, v_end, 1, [&](pd_ctx* s, uint32_t a, uint32_t b, uint32_t row) {
      return pd_decode_phased(s, a, b, out + 2 * row * n);
    });
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  const uint32_t N = c->idx.sample_ct;
  const uint32_t n_out = c->subset.active ? c->subset.count : N;
  if (v_begin >= v_end || !n_out) {
    if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
    return PD_OK;
  }
  // rows per batch: 128 MiB of pair output per buffer
  const uint64_t kOutBytes = 128ull << 20;
  const uint64_t row_bytes = 2ull * n_out;
  const uint32_t max_rows = static_cast<uint32_t>(std::max<uint64_t>(1, std::min<uint64_t>(65535, kOutBytes / row_bytes)));
  OutPipe pipe;
  int rc = pipe.Init(c, out, static_cast<size_t>(max_rows) * row_bytes);
  if (rc != PD_OK) return rc;
  const uint32_t* idx = c->subset.active ? c->subset.d_idx.as<uint32_t>() : nullptr;
  if (idx) {
    const cudaError_t e = c->d_phase_raw.Reserve(static_cast<size_t>(max_rows) * ((N + 7) & ~7u) * 2 + 64);
    if (e != cudaSuccess) return CheckCuda(c, e, "phase scratch allocation");
  }
  // records with a multiallelic patch track AND a phase track need the allele counts (pd_set_allele_counts) and a
  // bit array per CTA; without the counts they report PD_ERR_NOT_SUPPORTED
  uint32_t* mhet = nullptr;
  if (c->idx.has_multiallelic && c->idx.has_phase && !c->idx.allele_cts.empty()) {
    const cudaError_t e = c->d_mhet.Reserve(pdk::phased_mhet_scratch_bytes(N));
    if (e != cudaSuccess) return CheckCuda(c, e, "phase scratch allocation");
    mhet = c->d_mhet.as<uint32_t>();
  }
  cudaEventRecord(c->ev_c0, c->stream);
  rc = RunRows(c, v_begin, v_end, max_rows, 0, [&](const RowBatch& b, const pdk::RecIndex& ix, cudaStream_t s) {
    void* dev = nullptr;
    int slot = 0;
    int r = pipe.Acquire(&dev, &slot);
    if (r != PD_OK) return r;
    pdk::launch_expand_phased(ix, b.rows, b.d_vlist, b.v_first, idx, n_out, c->d_phase_raw.as<int8_t>(),
                              static_cast<int8_t*>(dev), DevErr(c), s, &c->launch_stats, mhet);
    c->stats.scan_launches += 1;
    return pipe.Send(slot, s, out + static_cast<size_t>(b.row0) * row_bytes, static_cast<size_t>(b.n) * row_bytes);
  });
  const int rc2 = pipe.Finish();
  cudaEventRecord(c->ev_c1, c->stream);
  cudaEventSynchronize(c->ev_c1);
  if (rc == PD_OK && rc2 == PD_OK) {
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
    c->stats.call_ms = ms;
    c->stats.kernel_ms = pipe.produce_ms;
  }
  return rc != PD_OK ? rc : rc2;
}

int pd_decode_alleles(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint8_t* out) {
  if (!c || (!out && v_end > v_begin)) return PD_ERR_INVALID_ARG;
  if (c->group) {
    const size_t n = EffectiveSamples(c);
    return GroupRunRanges(c, v_begin, v_end, 1, [&](pd_ctx* s, uint32_t a, uint32_t b, uint32_t row) {
      return pd_decode_alleles(s, a, b, out + 2 * row * n);
    });
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  const uint32_t N = c->idx.sample_ct;
  const uint32_t n_out = c->subset.active ? c->subset.count : N;
  if (v_begin >= v_end || !n_out) {
    if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
    return PD_OK;
  }
  if (c->idx.has_multiallelic && c->idx.allele_cts.empty()) {
    return c->SetError(PD_ERR_INVALID_ARG,
                       "pd_decode_alleles: the file has multiallelic records but stores no allele counts: "
                       "call pd_set_allele_counts (counts from the .pvar) first");
  }
  // rows per batch: 128 MiB of code pairs per buffer
  const uint64_t kOutBytes = 128ull << 20;
  const uint64_t row_bytes = 2ull * n_out;
  const uint32_t max_rows = static_cast<uint32_t>(std::max<uint64_t>(1, std::min<uint64_t>(65535, kOutBytes / row_bytes)));
  OutPipe pipe;
  int rc = pipe.Init(c, out, static_cast<size_t>(max_rows) * row_bytes);
  if (rc != PD_OK) return rc;
  const uint32_t* idx = c->subset.active ? c->subset.d_idx.as<uint32_t>() : nullptr;
  if (idx) {
    const cudaError_t e = c->d_phase_raw.Reserve(static_cast<size_t>(max_rows) * ((N + 7) & ~7u) * 2 + 64);
    if (e != cudaSuccess) return CheckCuda(c, e, "allele scratch allocation");
  }
  cudaEventRecord(c->ev_c0, c->stream);
  rc = RunRows(c, v_begin, v_end, max_rows, 0, [&](const RowBatch& b, const pdk::RecIndex& ix, cudaStream_t s) {
    void* dev = nullptr;
    int slot = 0;
    int r = pipe.Acquire(&dev, &slot);
    if (r != PD_OK) return r;
    pdk::launch_expand_alleles(ix, b.rows, b.d_vlist, b.v_first, idx, n_out, c->d_phase_raw.as<uint8_t>(),
                               static_cast<uint8_t*>(dev), DevErr(c), s, &c->launch_stats);
    c->stats.scan_launches += 1;
    return pipe.Send(slot, s, out + static_cast<size_t>(b.row0) * row_bytes, static_cast<size_t>(b.n) * row_bytes);
  });
  const int rc2 = pipe.Finish();
  cudaEventRecord(c->ev_c1, c->stream);
  cudaEventSynchronize(c->ev_c1);
  if (rc == PD_OK && rc2 == PD_OK) {
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_c0, c->ev_c1);
    c->stats.call_ms = ms;
    c->stats.kernel_ms = pipe.produce_ms;
  }
  return rc != PD_OK ? rc : rc2;
}

int pd_device_result(pd_ctx* c, uint32_t which, void** dev_ptr, uint64_t* bytes) {
  if (!c || !dev_ptr || !bytes) return PD_ERR_INVALID_ARG;
  if (c->group) {  // after the all-reduce every device holds the totals: hand out the first device's copy
    const int rc = pd_device_result(c->group->shards[0], which, dev_ptr, bytes);
    if (rc != PD_OK) c->last_error = c->group->shards[0]->last_error;
    return rc;
  }
  std::lock_guard<std::mutex> lock(c->mu);
  switch (which) {
    case PD_RESULT_SCORE:
      *dev_ptr = c->d_score.p;
      *bytes = static_cast<uint64_t>(c->last_score_n) * c->last_score_k * 8;
      break;
    case PD_RESULT_DOSAGE_SUM:
      *dev_ptr = c->d_dsum.p;
      *bytes = static_cast<uint64_t>(c->last_score_n) * 8;
      break;
    case PD_RESULT_ALLELE_CT:
      *dev_ptr = c->d_act.p;
      *bytes = static_cast<uint64_t>(c->last_score_n) * 4;
      break;
    case PD_RESULT_SAMPLE_COUNTS:
      *dev_ptr = c->d_out[0].p;
      *bytes = static_cast<uint64_t>(c->last_sample_n) * 12;
      break;
    default:
      return c->SetError(PD_ERR_INVALID_ARG, "unknown result id");
  }
  if (!*dev_ptr || !*bytes) return c->SetError(PD_ERR_INVALID_ARG, "no such result is held on the device");
  return PD_OK;
}

}  // extern "C"
