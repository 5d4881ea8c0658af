// Context management, HBM residency and the pinned double-buffered staging pipeline.
//
// What the reference does per scan thread -- fopen + fseeko + fread_unlocked per variant record
// into a private buffer (InitReadPtrs, pgenlib_read.cc:2763-2802) -- becomes: contiguous record
// ranges ("tiles", cut where no LD dependency crosses) read into page-locked buffers and pushed
// with cudaMemcpyAsync on a copy stream while the previous tile's kernels run.
#include "ctx.h"
#include "host_pool.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>

namespace pdx {

cudaError_t DevBuf::Reserve(size_t bytes) {
  if (bytes <= cap && p) return cudaSuccess;
  if (p) cudaFree(p);
  p = nullptr;
  cap = 0;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e == cudaSuccess) cap = bytes;
  return e;
}
void DevBuf::Free() {
  if (p) cudaFree(p);
  p = nullptr;
  cap = 0;
}
cudaError_t PinnedBuf::Reserve(size_t bytes) {
  if (bytes <= cap && p) return cudaSuccess;
  if (p) cudaFreeHost(p);
  p = nullptr;
  cap = 0;
  cudaError_t e = cudaHostAlloc(&p, bytes, cudaHostAllocDefault);
  if (e == cudaSuccess) cap = bytes;
  return e;
}
void PinnedBuf::Free() {
  if (p) cudaFreeHost(p);
  p = nullptr;
  cap = 0;
}

pdk::SubsetDev SubsetState::Dev() const {
  pdk::SubsetDev d{};
  if (!active) return d;
  d.mask2_shifts = d_mask2.as<uint8_t>();
  d.mask2_stride = stride2;
  d.mask1_shifts = d_mask1.as<uint8_t>();
  d.mask1_stride = stride1;
  d.include_words = d_words.as<uint32_t>();
  d.cum_popcount = d_cum.as<uint32_t>();
  d.subset_idx = d_idx.as<uint32_t>();
  d.subset_ct = count;
  return d;
}
void SubsetState::Free() {
  d_words.Free();
  d_cum.Free();
  d_mask2.Free();
  d_mask1.Free();
  d_idx.Free();
  active = false;
}

int CheckCuda(pd_ctx_impl* c, cudaError_t e, const char* what) {
  if (e == cudaSuccess) return PD_OK;
  char buf[256];
  snprintf(buf, sizeof(buf), "CUDA failure in %s: %s", what, cudaGetErrorString(e));
  return c->SetError(PD_ERR_CUDA, buf);
}

cudaError_t CopyFromPinnedSource(pd_ctx_impl* c, void* dst, uint64_t off, uint64_t len, cudaStream_t s) {
  // A page-locked file mapping is registered in pieces and one DMA may not span two registrations.  The shards of a
  // multi-GPU context see that mapping as a plain memory image, so the piece grid is looked up by address.
  const uint8_t* src = c->pgen.mem() + off;
  uint8_t* d = static_cast<uint8_t*>(dst);
  while (len) {
    const uint64_t room = ByteSource::PieceRoom(src);
    const uint64_t n = len < room ? len : room;
    const cudaError_t e = cudaMemcpyAsync(d, src, n, cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess) return e;
    d += n;
    src += n;
    len -= n;
  }
  return cudaSuccess;
}

int* DevErr(pd_ctx_impl* c) { return c->d_misc.as<int>(); }
uint32_t* DevWorkCounter(pd_ctx_impl* c) { return reinterpret_cast<uint32_t*>(c->d_misc.as<int>() + 4); }
uint32_t* DevSparseWords(pd_ctx_impl* c) { return reinterpret_cast<uint32_t*>(c->d_misc.as<int>() + 8); }

int CheckDeviceError(pd_ctx_impl* c) {
  int* h = c->h_misc.as<int>();
  cudaError_t e = cudaMemcpyAsync(h, DevErr(c), 2 * sizeof(int), cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "error readback");
  if (h[0]) {
    const int code = h[0];
    char buf[128];
    snprintf(buf, sizeof(buf), "malformed or unsupported variant record at variant %d", h[1]);
    cudaMemsetAsync(DevErr(c), 0, 2 * sizeof(int), c->stream);
    return c->SetError(code, buf);
  }
  return PD_OK;
}

uint32_t ExtendToBase(const pd_ctx_impl* c, uint32_t v) { return c->idx.Ldbase(v); }

// Subset mask upload + derived layouts (replaces BuildSampleSubset's interleaved vec and
// cumulative popcounts, src/plink_common.cpp:1222-1250).
int BuildSubset(pd_ctx_impl* c, const uint64_t* include, const SubsetState* and_with, SubsetState* out) {
  const uint32_t N = c->idx.sample_ct;
  const uint32_t nwords = (N + 31) / 32;
  const uint32_t padded = ((nwords + 3) & ~3u) + 8;  // 16-byte multiple plus slack
  out->words.assign(padded, 0u);
  const uint32_t* src32 = reinterpret_cast<const uint32_t*>(include);
  for (uint32_t w = 0; w < nwords; ++w) {
    uint32_t x = src32[w];
    if (and_with && and_with->active) x &= and_with->words[w];
    out->words[w] = x;
  }
  if (N & 31) out->words[nwords - 1] &= (1u << (N & 31)) - 1u;
  uint32_t cnt = 0;
  for (uint32_t w = 0; w < nwords; ++w) cnt += static_cast<uint32_t>(__builtin_popcount(out->words[w]));
  out->count = cnt;
  out->stride2 = ((N + 3) / 4 + 16 + 15) / 16 * 16 + 16;
  out->stride1 = ((N + 7) / 8 + 16 + 15) / 16 * 16 + 16;
  cudaError_t e = out->d_words.Reserve(padded * 4);
  if (e == cudaSuccess) e = out->d_cum.Reserve(padded * 4);
  if (e == cudaSuccess) e = out->d_mask2.Reserve(16ull * out->stride2);
  if (e == cudaSuccess) e = out->d_mask1.Reserve(16ull * out->stride1);
  if (e == cudaSuccess) e = out->d_idx.Reserve(std::max<size_t>(16, static_cast<size_t>(cnt) * 4));
  if (e != cudaSuccess) return CheckCuda(c, e, "subset allocation");
  e = cudaMemcpyAsync(out->d_words.p, out->words.data(), padded * 4, cudaMemcpyHostToDevice, c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "subset upload");
  c->stats.h2d_bytes += padded * 4;
  pdk::launch_build_masks(out->d_words.as<uint32_t>(), N, out->d_mask2.as<uint8_t>(), out->stride2,
                          out->d_mask1.as<uint8_t>(), out->stride1, c->stream, &c->launch_stats);
  pdk::launch_cum_popcount(out->d_words.as<uint32_t>(), padded, out->d_cum.as<uint32_t>(), c->stream,
                           &c->launch_stats);
  pdk::launch_subset_indices(out->d_words.as<uint32_t>(), out->d_cum.as<uint32_t>(), nwords, out->d_idx.as<uint32_t>(),
                             c->stream, &c->launch_stats);
  e = cudaStreamSynchronize(c->stream);  // out->words must stay valid until the copy is done
  if (e != cudaSuccess) return CheckCuda(c, e, "subset build");
  out->active = true;
  return PD_OK;
}

namespace {

// Parallel read of [off, off+len) into dst: page-cache pread (or memcpy from a pageable image)
// is a few GB/s per thread, PCIe Gen5 wants ~50 GB/s.
bool ParallelRead(const ByteSource& src, uint64_t off, uint8_t* dst, uint64_t len) {
  const uint64_t kChunk = 2ull << 20;
  const uint64_t nchunks = (len + kChunk - 1) / kChunk;
  if (nchunks <= 1) return src.ReadAt(off, dst, len);
  std::atomic<bool> ok(true);
  HostPool::Get().ParallelFor(nchunks, [&](uint64_t i) {
    const uint64_t o = i * kChunk;
    if (!src.ReadAt(off + o, dst + o, std::min(kChunk, len - o))) ok.store(false);
  });
  return ok.load();
}

// f(a, b) over [0, n) cut into pieces of about `per` items, on the host pool.
template <class F>
void ParallelRanges(uint64_t n, uint64_t per, const F& f) {
  if (!n) return;
  per = std::max<uint64_t>(1, per);
  const uint64_t pieces = (n + per - 1) / per;
  HostPool::Get().ParallelFor(pieces, [&](uint64_t i) { f(i * per, std::min(n, (i + 1) * per)); });
}

// Bytes of record v that genotype counting reads: everything for raw / 1-bit / LD records, LD bases and
// records with extra tracks; for a plain difflist record (vrtype 4 / 6 / 7) only the length, the group
// index and the 2-bit values -- not the sample-id deltas (layout: ParseDifflistHeader,
// pgenlib_read.cc:2184-2218).
uint64_t TrimmedLength(const PgenIndex& ix, const uint8_t* view, uint32_t v) {
  const uint64_t full = ix.fpos[v + 1] - ix.fpos[v];
  const uint32_t vt = ix.vrtypes[v];
  if ((vt & ~7u) || (vt & 7u) < 4 || (vt & 7u) == 5) return full;
  if (v + 1 < ix.variant_ct && ix.ldbase[v + 1] == v) return full;  // an LD record patches this one: ids needed
  const uint8_t* p = view + ix.fpos[v];
  uint32_t len = 0, shift = 0;
  uint64_t used = 0;
  for (;;) {  // varint
    if (used >= full || shift > 28) return full;
    const uint8_t b = p[used++];
    len |= static_cast<uint32_t>(b & 127u) << shift;
    if (!(b & 128u)) break;
    shift += 7;
  }
  if (!len) return full;
  const uint32_t n = ix.sample_ct;
  const uint32_t idbytes = 1u + ((31u - static_cast<uint32_t>(__builtin_clz(n))) >> 3);  // BytesToRepresentNzU32
  const uint64_t groups = (static_cast<uint64_t>(len) + 63) / 64;
  const uint64_t want = used + groups * (idbytes + 1) - 1 + (static_cast<uint64_t>(len) + 3) / 4;
  return want < full ? want : full;
}

// Builds c->trim_fpos once; usable only when the source is host-addressable and trimming saves >= 30 %.
bool EnsureTrimIndex(pd_ctx_impl* c) {
  if (c->trim_state) return c->trim_state > 0;
  c->trim_state = -1;
  const PgenIndex& ix = c->idx;
  const uint8_t* view = c->pgen.view();
  if (ix.fixed || !view || !ix.has_difflist || !ix.variant_ct) return false;
  const uint32_t V = ix.variant_ct;
  std::vector<uint64_t> pos(static_cast<size_t>(V) + 1);
  ParallelRanges(V, 4096, [&](uint64_t a, uint64_t b) {
    for (uint64_t v = a; v < b; ++v) pos[v + 1] = TrimmedLength(ix, view, static_cast<uint32_t>(v));
  });
  pos[0] = 0;
  for (uint32_t v = 0; v < V; ++v) pos[v + 1] += pos[v];
  if (pos[V] * 10 > ix.data_bytes * 7) return false;
  if (c->d_trim_fpos.Reserve(pos.size() * 8) != cudaSuccess) return false;
  if (cudaMemcpy(c->d_trim_fpos.p, pos.data(), pos.size() * 8, cudaMemcpyHostToDevice) != cudaSuccess) return false;
  c->trim_fpos.swap(pos);
  c->trim_state = 1;
  return true;
}

// Copies the counted bytes of records [v0, v1) into dst (their compacted stream).
void GatherTrimmed(const pd_ctx_impl* c, uint32_t v0, uint32_t v1, uint8_t* dst) {
  const PgenIndex& ix = c->idx;
  const uint8_t* view = c->pgen.view();
  const std::vector<uint64_t>& tp = c->trim_fpos;
  const uint64_t base = tp[v0];
  ParallelRanges(v1 - v0, 512, [&](uint64_t a, uint64_t b) {
    uint64_t v = v0 + a;
    const uint64_t e = v0 + b;
    while (v < e) {
      // runs of untrimmed records are contiguous in the file as well: one memcpy
      uint64_t w = v;
      while (w + 1 < e && (tp[w + 1] - tp[w]) == (ix.fpos[w + 1] - ix.fpos[w])) ++w;
      memcpy(dst + (tp[v] - base), view + ix.fpos[v], tp[w + 1] - tp[v]);
      v = w + 1;
    }
  });
}

void PlanTiles(const pd_ctx_impl* c, uint32_t v0, uint32_t v1, std::vector<Tile>* tiles, bool trim) {
  const PgenIndex& ix = c->idx;
  if (ix.fixed) {
    uint64_t per = c->tile_bytes / ix.fixed_width;
    if (per == 0) per = 1;
    for (uint64_t a = v0; a < v1; a += per) {
      const uint32_t b = static_cast<uint32_t>(std::min<uint64_t>(v1, a + per));
      tiles->push_back(Tile{static_cast<uint32_t>(a), b, ix.RecOff(static_cast<uint32_t>(a)),
                            ix.data_start + static_cast<uint64_t>(b) * ix.fixed_width});
    }
    return;
  }
  uint32_t a = v0;
  while (a < v1) {
    uint32_t b = a + 1;
    // extend while under budget; never cut in front of an LD-compressed record
    const std::vector<uint64_t>& pos = trim ? c->trim_fpos : ix.fpos;
    // trimmed tiles are gathered by the host before they can be copied: smaller tiles start the pipeline sooner
    const uint64_t budget = trim ? std::min<uint64_t>(c->tile_bytes, 16ull << 20) : c->tile_bytes;
    while (b < v1 && ((pos[b + 1] - pos[a] <= budget) || ix.ldbase[b] != b)) ++b;
    tiles->push_back(Tile{a, b, pos[a], pos[b]});
    a = b;
  }
}

cudaEvent_t PoolEvent(pd_ctx_impl* c, size_t i) {
  while (c->ev_pool.size() <= i) {
    cudaEvent_t e;
    cudaEventCreate(&e);
    c->ev_pool.push_back(e);
  }
  return c->ev_pool[i];
}

}  // namespace

int RunScan(pd_ctx_impl* c, uint32_t v_begin, uint32_t v_end, uint32_t* v_ext, const TileFn& fn,
            const std::function<void(const Tile&, cudaEvent_t)>& after_tile) {
  const PgenIndex& ixh = c->idx;
  if (v_begin > v_end || v_end > ixh.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
  const uint32_t v0 = (v_begin < v_end) ? ExtendToBase(c, v_begin) : v_begin;
  if (v_ext) *v_ext = v0;
  c->stats.record_bytes = 0;
  c->stats.kernel_ms = 0;
  c->stats.aux_kernel_ms = 0;
  c->stats.total_ms = 0;
  c->stats.scan_launches = 0;
  c->stats.host_stage_ms = 0;
  c->stats.host_wait_ms = 0;
  if (v_begin == v_end) return PD_OK;

  pdk::RecIndex ix{};
  ix.fpos = c->d_fpos.as<uint64_t>();
  ix.vrtypes = c->d_vrtypes.as<uint8_t>();
  ix.ldbase = c->d_ldbase.as<uint32_t>();
  ix.base_slot = c->d_base_slot.as<uint32_t>();
  ix.fixed_start = ixh.data_start;
  ix.fixed_width = ixh.fixed_width;
  ix.const_vrtype = ixh.const_vrtype;
  ix.N = ixh.sample_ct;
  ix.allele_cts = c->d_allele_cts.p ? c->d_allele_cts.as<uint8_t>() : nullptr;
  c->stats.record_bytes = ixh.RecOff(v_end - 1) + (ixh.fixed ? ixh.fixed_width : ixh.fpos[v_end] - ixh.fpos[v_end - 1]) -
                          ixh.RecOff(v_begin);

  // a failed earlier call may have left its malformed-record flag behind: every scan starts clean
  cudaMemsetAsync(DevErr(c), 0, 2 * sizeof(int), c->stream);
  memset(c->h_misc.as<int>() + 8, 0, 24 * sizeof(int));
  cudaEventRecord(c->ev_t0, c->stream);
  int rc = PD_OK;
  if (c->resident && v0 >= c->res_v0 && v_end <= c->res_v1) {
    Tile t{v0, v_end, ixh.RecOff(v0), ixh.fixed ? ixh.data_start + static_cast<uint64_t>(v_end) * ixh.fixed_width
                                               : ixh.fpos[v_end]};
    ix.data = c->res_ptr;
    ix.data_fpos = c->res_f0;
    cudaEvent_t k0 = PoolEvent(c, 0), k1 = PoolEvent(c, 1);
    cudaEventRecord(k0, c->stream);
    rc = fn(t, ix, c->stream);
    cudaEventRecord(k1, c->stream);
    if (after_tile) after_tile(t, k1);
    cudaError_t e = cudaEventSynchronize(k1);
    if (rc == PD_OK && e != cudaSuccess) rc = CheckCuda(c, e, "scan kernels");
    if (rc == PD_OK) {
      float ms = 0;
      cudaEventElapsedTime(&ms, k0, k1);
      c->stats.kernel_ms = ms;
    }
  } else {
    const bool trim = c->trim_scan && EnsureTrimIndex(c);
    if (trim) ix.fpos = c->d_trim_fpos.as<uint64_t>();
    std::vector<Tile> tiles;
    PlanTiles(c, v0, v_end, &tiles, trim);
    uint64_t max_len = 0;
    for (auto& t : tiles) max_len = std::max(max_len, t.f1 - t.f0);
    const uint32_t nbuf = c->n_staging;
    for (uint32_t b = 0; b < nbuf && rc == PD_OK; ++b) {
      cudaError_t e = c->d_stage[b].Reserve(max_len + 2 * kSlack);
      if (e == cudaSuccess && (trim || !c->src_pinned)) e = c->h_stage[b].Reserve(max_len);
      if (e != cudaSuccess) rc = CheckCuda(c, e, "staging allocation");
    }
    for (size_t i = 0; i < tiles.size() && rc == PD_OK; ++i) {
      const Tile& t = tiles[i];
      const uint32_t b = static_cast<uint32_t>(i % nbuf);
      const uint64_t len = t.f1 - t.f0;
      const auto tw0 = std::chrono::steady_clock::now();
      if (i >= nbuf) {
        cudaError_t e = cudaEventSynchronize(c->ev_done[b]);  // buffer b's previous tile fully consumed
        if (e != cudaSuccess) {
          rc = CheckCuda(c, e, "tile completion");
          break;
        }
        // that tile's copy of the device error word has landed too: a malformed record stops the scan here
        // (nbuf tiles after it was met) instead of at the end of a 125 GB file (pgenlib_read.cc:2195-2198
        // fails at the record)
        const int* flag = c->h_misc.as<int>() + 8 + 2 * (b < 12 ? b : 0);
        if (b < 12 && flag[0]) {
          char buf[128];
          snprintf(buf, sizeof(buf), "malformed or unsupported variant record at variant %d", flag[1]);
          rc = c->SetError(flag[0], buf);
          break;
        }
      }
      const auto tw1 = std::chrono::steady_clock::now();
      c->stats.host_wait_ms += std::chrono::duration<double, std::milli>(tw1 - tw0).count();
      const uint8_t* hsrc = nullptr;
      if (trim) {
        GatherTrimmed(c, t.v0, t.v1, c->h_stage[b].as<uint8_t>());
        hsrc = c->h_stage[b].as<uint8_t>();
      } else if (c->src_pinned) {
        // DMA straight from the caller's page-locked image / the pinned file mapping
      } else {
        if (!ParallelRead(c->pgen, t.f0, c->h_stage[b].as<uint8_t>(), len)) {
          rc = c->SetError(PD_ERR_READ, ".pgen read failure");
          break;
        }
        hsrc = c->h_stage[b].as<uint8_t>();
      }
      c->stats.host_stage_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tw1).count();
      cudaError_t e = hsrc ? cudaMemcpyAsync(c->d_stage[b].as<uint8_t>() + kSlack, hsrc, len, cudaMemcpyHostToDevice,
                                             c->copy_stream)
                           : CopyFromPinnedSource(c, c->d_stage[b].as<uint8_t>() + kSlack, t.f0, len, c->copy_stream);
      if (e != cudaSuccess) {
        rc = CheckCuda(c, e, "H2D staging copy");
        break;
      }
      c->stats.h2d_bytes += len;
      cudaEventRecord(c->ev_copied[b], c->copy_stream);
      cudaStreamWaitEvent(c->stream, c->ev_copied[b], 0);
      ix.data = c->d_stage[b].as<uint8_t>() + kSlack;
      ix.data_fpos = t.f0;
      if (ixh.bed) {  // PLINK 1 codes -> .pgen codes before any kernel reads the tile
        pdk::launch_bed_to_pgen(c->d_stage[b].as<uint8_t>() + kSlack, 0, ixh.fixed_width, ixh.sample_ct, t.v1 - t.v0,
                                c->stream, &c->launch_stats);
      }
      cudaEvent_t k0 = PoolEvent(c, 2 * i), k1 = PoolEvent(c, 2 * i + 1);
      cudaEventRecord(k0, c->stream);
      rc = fn(t, ix, c->stream);
      cudaEventRecord(k1, c->stream);
      if (b < 12) {  // h_misc holds 12 per-buffer copies of the error word behind the 8 ints CheckDeviceError uses
        cudaMemcpyAsync(c->h_misc.as<int>() + 8 + 2 * b, DevErr(c), 2 * sizeof(int), cudaMemcpyDeviceToHost, c->stream);
      }
      cudaEventRecord(c->ev_done[b], c->stream);
      if (after_tile) after_tile(t, k1);
    }
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (rc == PD_OK && e != cudaSuccess) rc = CheckCuda(c, e, "scan kernels");
    if (rc == PD_OK) {
      double tot = 0;
      for (size_t i = 0; i < tiles.size(); ++i) {
        float ms = 0;
        cudaEventElapsedTime(&ms, PoolEvent(c, 2 * i), PoolEvent(c, 2 * i + 1));
        tot += ms;
      }
      c->stats.kernel_ms = tot;
    }
  }
  cudaEventRecord(c->ev_t1, c->stream);
  cudaEventSynchronize(c->ev_t1);
  if (rc == PD_OK) {
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev_t0, c->ev_t1);
    c->stats.total_ms = ms;
    rc = CheckDeviceError(c);
  } else {
    cudaStreamSynchronize(c->stream);  // tiles still in flight must not outlive the buffers / the call
    cudaMemsetAsync(DevErr(c), 0, 2 * sizeof(int), c->stream);
  }
  return rc;
}


// ---- dense-row batches ---------------------------------------------------------------------------

uint32_t DenseSlotStride(const pd_ctx_impl* c) { return ((c->idx.sample_ct + 3) / 4 + 15) / 16 * 16 + 16; }

namespace {

// Decoded rows of one batch.  Measured on 2.5 KB difflist records (profiles/r01d_ncu_k_densify.txt): the
// expansion is latency-bound per row (one warp walks one record), so what matters is rows in flight -- with
// 64 MiB (L2-resident, 2.7k rows at 100k samples) the GPU ran 18 warps per SM; 1 GiB gives every SM a full
// complement and cuts the per-batch fixed costs of the consumers (plink_score finishes once per batch).
constexpr uint64_t kDenseScratchBytes = 1024ull << 20;
constexpr uint32_t kMaxBatchRows = 65535;  // also the gridDim.y limit

uint32_t ClampBatchRows(const pd_ctx_impl* c, uint32_t max_rows) {
  uint64_t r = max_rows ? max_rows : kMaxBatchRows;
  if (r > kMaxBatchRows) r = kMaxBatchRows;
  if (!c->idx.fixed) r = std::min<uint64_t>(r, std::max<uint64_t>(1, kDenseScratchBytes / DenseSlotStride(c)));
  return static_cast<uint32_t>(std::max<uint64_t>(1, r));
}

int ReserveDense(pd_ctx_impl* c, uint32_t rows, bool need_ptrs) {
  cudaError_t e = cudaSuccess;
  if (!c->idx.fixed) e = c->d_dense.Reserve(static_cast<size_t>(rows) * DenseSlotStride(c) + 64);
  if (e == cudaSuccess && need_ptrs) e = c->d_rowptrs.Reserve(static_cast<size_t>(rows) * sizeof(void*));
  return e == cudaSuccess ? PD_OK : CheckCuda(c, e, "dense row scratch allocation");
}

struct TileBytesGuard {
  pd_ctx_impl* c;
  uint64_t saved;
  TileBytesGuard(pd_ctx_impl* ctx, uint64_t min_bytes) : c(ctx), saved(ctx->tile_bytes) {
    if (min_bytes > c->tile_bytes) c->tile_bytes = min_bytes;
  }
  ~TileBytesGuard() { c->tile_bytes = saved; }
};

}  // namespace

// Dense copies of the LD bases that the rows of [v_lo, v_hi) refer to (those that are not plain 2-bit records, which
// are read in place), decoded once per batch by a CTA each; k_densify's LD rows copy them.  Without this every LD row
// decoded its base again, alone in its warp: with one-bit bases (a 12 k-entry difflist at 100k samples) that was most
// of the time of k_densify on LD-compressed files.  Leaves *out empty when the range has no such base.
int PrepareBases(pd_ctx_impl* c, const pdk::RecIndex& ix, uint32_t v_lo, uint32_t v_hi, cudaStream_t s,
                 pdk::LdScratch* out) {
  *out = pdk::LdScratch{};
  if (!c->idx.has_ld || c->slot_vidx.empty() || v_lo >= v_hi) return PD_OK;
  // bases of rows in [v_lo, v_hi) lie in [Ldbase(v_lo), v_hi); the slots of that range are contiguous
  const uint32_t b_lo = c->idx.Ldbase(v_lo);
  auto lo = std::lower_bound(c->slot_vidx.begin(), c->slot_vidx.end(), b_lo);
  auto hi = std::lower_bound(c->slot_vidx.begin(), c->slot_vidx.end(), v_hi);
  if (lo == hi) return PD_OK;
  const uint32_t slot_lo = static_cast<uint32_t>(lo - c->slot_vidx.begin());
  const uint32_t slot_ct = static_cast<uint32_t>(hi - lo);
  out->stride = ((c->idx.sample_ct + 3) / 4 + 15) / 16 * 16 + 16;
  out->slot_lo = slot_lo;
  const size_t need = static_cast<size_t>(slot_ct) * out->stride;
  if (need > c->d_scratch.cap) {
    cudaStreamSynchronize(s);  // growing frees the old buffer: no earlier batch may still read it
    const cudaError_t e = c->d_scratch.Reserve(need + (need >> 2));
    if (e != cudaSuccess) return CheckCuda(c, e, "LD scratch allocation");
  }
  out->dense = c->d_scratch.as<uint8_t>();
  out->slot_ct = slot_ct;
  pdk::launch_decode_bases(ix, c->d_slot_vidx.as<uint32_t>(), *out, slot_lo + slot_ct - 1, DevErr(c), s, &c->launch_stats);
  return PD_OK;
}

int RunRows(pd_ctx_impl* c, uint32_t v_begin, uint32_t v_end, uint32_t max_rows, uint64_t min_tile_bytes,
            const RowFn& fn) {
  // never reserve scratch for more rows than the range holds (a 1 GiB batch for a 100-variant file)
  const uint32_t cap = std::min(ClampBatchRows(c, max_rows), std::max(1u, v_end > v_begin ? v_end - v_begin : 1u));
  const bool fixed = c->idx.fixed;
  int rc = ReserveDense(c, cap, !fixed);
  if (rc != PD_OK) return rc;
  TileBytesGuard guard(c, min_tile_bytes);
  const uint32_t slot_stride = DenseSlotStride(c);
  const pdk::SparseAcc* sparse = fixed ? nullptr : c->sparse_acc;
  return RunScan(c, v_begin, v_end, nullptr, [&](const Tile& t, const pdk::RecIndex& ix, cudaStream_t s) {
    const uint32_t a = std::max(t.v0, v_begin);
    uint32_t n = 0;
    for (uint32_t v = a; v < t.v1; v += n) {
      n = std::min(cap, t.v1 - v);
      if (sparse) {
        // Difflist records need no scratch slot: a batch ends when cap records that do are in it.
        uint32_t dense = 0;
        n = 0;
        while (v + n < t.v1 && n < kMaxBatchRows) {
          const uint32_t main = c->idx.vrtypes[v + n] & 7u;
          if (main < 4 && ++dense > cap) break;
          ++n;
        }
      }
      RowBatch b{};
      b.row0 = v - v_begin;
      b.n = n;
      b.v_first = v;
      b.rows.nrows = b.n;
      b.rows.N = ix.N;
      if (fixed) {
        b.rows.base = ix.data + (ix.fixed_start + static_cast<uint64_t>(v) * ix.fixed_width - ix.data_fpos);
        b.rows.stride = ix.fixed_width;
      } else {
        pdk::LdScratch bases{};
        const int pr = PrepareBases(c, ix, v, v + b.n, s, &bases);
        if (pr != PD_OK) return pr;
        pdk::launch_densify(ix, nullptr, v, b.n, c->d_dense.as<uint8_t>(), slot_stride,
                            c->d_rowptrs.as<const uint8_t*>(), DevWorkCounter(c), DevErr(c), s, &c->launch_stats,
                            sparse, &bases);
        b.rows.ptrs = c->d_rowptrs.as<const uint8_t*>();
        if (sparse) {
          b.rows.nrows = std::min(b.n, cap);
          b.rows.nrows_dev = sparse->dense_count;
        }
      }
      const int r = fn(b, ix, s);
      if (r != PD_OK) return r;
    }
    return PD_OK;
  });
}

int RunRowsList(pd_ctx_impl* c, const uint32_t* vlist, uint32_t n, uint32_t max_rows, const RowFn& fn) {
  if (!n) return PD_OK;
  const PgenIndex& ixh = c->idx;
  for (uint32_t i = 0; i < n; ++i) {
    if (vlist[i] >= ixh.variant_ct || (i && vlist[i] <= vlist[i - 1])) {
      return c->SetError(PD_ERR_INVALID_ARG, "variant list must be strictly ascending and inside the file");
    }
  }
  const uint32_t cap = ClampBatchRows(c, max_rows);
  int rc = ReserveDense(c, cap, true);
  if (rc != PD_OK) return rc;
  cudaError_t e = c->d_vlist.Reserve(static_cast<size_t>(n) * 4);
  if (e == cudaSuccess) e = cudaMemcpyAsync(c->d_vlist.p, vlist, static_cast<size_t>(n) * 4, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);  // vlist may be pageable caller memory
  if (e != cudaSuccess) return CheckCuda(c, e, "variant list upload");
  c->stats.h2d_bytes += static_cast<size_t>(n) * 4;
  const uint32_t slot_stride = DenseSlotStride(c);
  // Segments: maximal runs of listed variants whose gaps are small enough that staging the gap
  // is cheaper than another pipeline start (always one segment when the range is resident).
  const uint64_t kGapBytes = 4ull << 20;
  uint32_t i0 = 0;
  double kernel_ms = 0, total_ms = 0;
  uint64_t rec_bytes = 0;
  while (i0 < n) {
    uint32_t i1 = i0 + 1;
    while (i1 < n) {
      const uint64_t gap = ixh.RecOff(ixh.Ldbase(vlist[i1])) - ixh.RecOff(vlist[i1 - 1]);
      const bool res = c->resident && vlist[i0] >= c->res_v0 && vlist[i1] < c->res_v1;
      if (!res && gap > kGapBytes) break;
      ++i1;
    }
    rc = RunScan(c, vlist[i0], vlist[i1 - 1] + 1, nullptr, [&](const Tile& t, const pdk::RecIndex& ix, cudaStream_t s) {
      const uint32_t* lo = std::lower_bound(vlist + i0, vlist + i1, t.v0);
      const uint32_t* hi = std::lower_bound(vlist + i0, vlist + i1, t.v1);
      for (uint32_t j = static_cast<uint32_t>(lo - vlist); j < static_cast<uint32_t>(hi - vlist); j += cap) {
        RowBatch b{};
        b.row0 = j;
        b.n = std::min<uint32_t>(cap, static_cast<uint32_t>(hi - vlist) - j);
        b.v_first = vlist[j];
        b.d_vlist = c->d_vlist.as<uint32_t>() + j;
        b.rows.nrows = b.n;
        b.rows.N = ix.N;
        // lists: pre-decode the bases only when the list is dense in its range (else most would not be used)
        pdk::LdScratch bases{};
        const uint32_t span = vlist[j + b.n - 1] + 1 - vlist[j];
        if (span <= 4 * b.n) {
          const int pr = PrepareBases(c, ix, vlist[j], vlist[j + b.n - 1] + 1, s, &bases);
          if (pr != PD_OK) return pr;
        }
        pdk::launch_densify(ix, b.d_vlist, 0, b.n, c->d_dense.as<uint8_t>(), slot_stride,
                            c->d_rowptrs.as<const uint8_t*>(), DevWorkCounter(c), DevErr(c), s, &c->launch_stats,
                            nullptr, &bases);
        b.rows.ptrs = c->d_rowptrs.as<const uint8_t*>();
        const int r = fn(b, ix, s);
        if (r != PD_OK) return r;
      }
      return PD_OK;
    });
    if (rc != PD_OK) return rc;
    kernel_ms += c->stats.kernel_ms;
    total_ms += c->stats.total_ms;
    rec_bytes += c->stats.record_bytes;
    i0 = i1;
  }
  c->stats.kernel_ms = kernel_ms;
  c->stats.total_ms = total_ms;
  c->stats.record_bytes = rec_bytes;
  return PD_OK;
}

}  // namespace pdx
