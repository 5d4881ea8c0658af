// extern "C" surface of libplinkgpu.so (see include/plinkgpu.h for the contract and the
// reference call sites each entry point replaces).
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>

#include "ctx.h"
#include "group.h"

using namespace pdx;

namespace {

void CopyErr(char* errbuf, size_t n, const std::string& msg) {
  if (errbuf && n) snprintf(errbuf, n, "%s", msg.c_str());
}

void DestroyCtx(pd_ctx* c) {
  if (!c) return;
  if (c->group) {  // multi-GPU context: the shards own every device resource
    GroupDestroy(c);
    delete c;
    return;
  }
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->copy_stream) cudaStreamSynchronize(c->copy_stream);
  if (c->d2h_stream) cudaStreamSynchronize(c->d2h_stream);
  c->d_fpos.Free(); c->d_vrtypes.Free(); c->d_ldbase.Free(); c->d_base_slot.Free(); c->d_slot_vidx.Free(); c->d_allele_cts.Free(); c->d_mhet.Free();
  c->d_resident.Free();
  for (auto& b : c->h_stage) b.Free();
  for (auto& b : c->d_stage) b.Free();
  for (auto e : c->ev_copied) cudaEventDestroy(e);
  for (auto e : c->ev_done) cudaEventDestroy(e);
  for (auto e : c->ev_pool) cudaEventDestroy(e);
  c->d_counts.Free(); c->d_scratch.Free(); c->d_misc.Free(); c->d_pvals.Free(); c->d_acc.Free(); c->d_ready.Free(); c->d_hweq.Free(); c->d_trim_fpos.Free(); c->d_phase_raw.Free(); c->d_tidy_off.Free(); c->d_span.Free(); c->d_pairs.Free();
  c->d_dense.Free(); c->d_rowptrs.Free(); c->d_vlist.Free();
  c->d_loc.Free(); c->d_rank.Free(); c->d_dstats.Free(); c->d_srow.Free(); c->d_weights.Free(); c->d_flip.Free();
  c->d_score.Free(); c->d_dsum.Free(); c->d_act.Free(); c->d_tc.Free();
  for (int b = 0; b < 2; ++b) {
    c->d_out[b].Free();
    c->h_bounce[b].Free();
    if (c->ev_out_ready[b]) cudaEventDestroy(c->ev_out_ready[b]);
    if (c->ev_out_free[b]) cudaEventDestroy(c->ev_out_free[b]);
    if (c->ev_prod0[b]) cudaEventDestroy(c->ev_prod0[b]);
    if (c->ev_prod1[b]) cudaEventDestroy(c->ev_prod1[b]);
  }
  if (c->d2h_stream) cudaStreamDestroy(c->d2h_stream);
  c->h_misc.Free(); c->h_out.Free();
  c->subset.Free();
  if (c->ev_a) cudaEventDestroy(c->ev_a);
  if (c->ev_b) cudaEventDestroy(c->ev_b);
  if (c->ev_t0) cudaEventDestroy(c->ev_t0);
  if (c->ev_t1) cudaEventDestroy(c->ev_t1);
  if (c->ev_c0) cudaEventDestroy(c->ev_c0);
  if (c->ev_c1) cudaEventDestroy(c->ev_c1);
  if (c->stream) cudaStreamDestroy(c->stream);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  delete c;
}

// Device setup shared by pd_open / pd_open_mem.  No CUDA device => hard failure (no CPU path).
int InitDevice(pd_ctx* c, const pd_open_opts* opts) {
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    return c->SetError(PD_ERR_CUDA, std::string("no usable CUDA device (libplinkgpu has no CPU fallback): ") +
                                        (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
  }
  int dev = opts ? opts->device : -1;
  if (dev < 0) {
    e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return CheckCuda(c, e, "cudaGetDevice");
  }
  if (dev >= ndev) return c->SetError(PD_ERR_INVALID_ARG, "device ordinal out of range");
  c->device = dev;
  if ((e = cudaSetDevice(dev)) != cudaSuccess) return CheckCuda(c, e, "cudaSetDevice");
  cudaDeviceGetAttribute(&c->launch_stats.sm_count, cudaDevAttrMultiProcessorCount, dev);
  pdk::configure_device_kernels();
  if (opts && opts->tile_bytes) c->tile_bytes = opts->tile_bytes;
  if (opts && opts->n_staging) c->n_staging = std::max<uint32_t>(2, opts->n_staging);
  if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) return CheckCuda(c, e, "stream");
  if ((e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking)) != cudaSuccess) return CheckCuda(c, e, "stream");
  cudaEventCreate(&c->ev_a); cudaEventCreate(&c->ev_b); cudaEventCreate(&c->ev_t0); cudaEventCreate(&c->ev_t1);
  cudaEventCreate(&c->ev_c0); cudaEventCreate(&c->ev_c1);
  if ((e = cudaStreamCreateWithFlags(&c->d2h_stream, cudaStreamNonBlocking)) != cudaSuccess) return CheckCuda(c, e, "stream");
  for (int b = 0; b < 2; ++b) {
    cudaEventCreateWithFlags(&c->ev_out_ready[b], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->ev_out_free[b], cudaEventDisableTiming);
    cudaEventCreate(&c->ev_prod0[b]);
    cudaEventCreate(&c->ev_prod1[b]);
  }
  c->h_stage.resize(c->n_staging);
  c->d_stage.resize(c->n_staging);
  c->ev_copied.resize(c->n_staging);
  c->ev_done.resize(c->n_staging);
  for (uint32_t b = 0; b < c->n_staging; ++b) {
    cudaEventCreateWithFlags(&c->ev_copied[b], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->ev_done[b], cudaEventDisableTiming);
  }
  if ((e = c->d_misc.Reserve(256)) != cudaSuccess) return CheckCuda(c, e, "cudaMalloc");
  if ((e = c->h_misc.Reserve(256)) != cudaSuccess) return CheckCuda(c, e, "cudaHostAlloc");
  cudaMemsetAsync(c->d_misc.p, 0, 256, c->stream);

  // Index upload (variable-width files): fpos, vrtypes, LD base table, scratch-slot table.
  const PgenIndex& ix = c->idx;
  if (!ix.fixed) {
    const uint32_t V = ix.variant_ct;
    c->base_slot.assign(V, 0xffffffffu);
    c->slot_vidx.clear();
    uint32_t last_base = 0xffffffffu, last_slot = 0xffffffffu;
    for (uint32_t v = 0; v < V; ++v) {
      const uint32_t b = ix.ldbase[v];
      if (b == v) continue;
      if ((ix.vrtypes[b] & 7) == 0) continue;  // dense base is used in place
      if (b != last_base) {
        last_base = b;
        last_slot = static_cast<uint32_t>(c->slot_vidx.size());
        c->slot_vidx.push_back(b);
      }
      c->base_slot[v] = last_slot;
    }
    std::vector<uint8_t> vt(ix.vrtypes);
    vt.resize(V + 16, 0);
    e = c->d_fpos.Reserve((static_cast<size_t>(V) + 1) * 8);
    if (e == cudaSuccess) e = c->d_vrtypes.Reserve(vt.size());
    if (e == cudaSuccess) e = c->d_ldbase.Reserve(static_cast<size_t>(V) * 4);
    if (e == cudaSuccess) e = c->d_base_slot.Reserve(static_cast<size_t>(V) * 4);
    if (e == cudaSuccess) e = c->d_slot_vidx.Reserve(std::max<size_t>(4, c->slot_vidx.size() * 4));
    if (e != cudaSuccess) return CheckCuda(c, e, "index allocation");
    cudaMemcpyAsync(c->d_fpos.p, ix.fpos.data(), (static_cast<size_t>(V) + 1) * 8, cudaMemcpyHostToDevice, c->stream);
    cudaMemcpyAsync(c->d_vrtypes.p, vt.data(), vt.size(), cudaMemcpyHostToDevice, c->stream);
    cudaMemcpyAsync(c->d_ldbase.p, ix.ldbase.data(), static_cast<size_t>(V) * 4, cudaMemcpyHostToDevice, c->stream);
    cudaMemcpyAsync(c->d_base_slot.p, c->base_slot.data(), static_cast<size_t>(V) * 4, cudaMemcpyHostToDevice, c->stream);
    if (!c->slot_vidx.empty()) {
      cudaMemcpyAsync(c->d_slot_vidx.p, c->slot_vidx.data(), c->slot_vidx.size() * 4, cudaMemcpyHostToDevice, c->stream);
    }
    c->stats.h2d_bytes += (static_cast<size_t>(V) + 1) * 8 + vt.size() + static_cast<size_t>(V) * 8;
    if (!ix.allele_cts.empty()) {
      e = c->d_allele_cts.Reserve(V);
      if (e != cudaSuccess) return CheckCuda(c, e, "index allocation");
      cudaMemcpyAsync(c->d_allele_cts.p, ix.allele_cts.data(), V, cudaMemcpyHostToDevice, c->stream);
      c->stats.h2d_bytes += V;
    }
  }
  e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "index upload");
  return PD_OK;
}

int FinishOpen(pd_ctx* c, const pd_open_opts* opts, pd_ctx** out, char* errbuf, size_t errbuf_len,
               const char* path = nullptr) {
  std::string err;
  int rc = ParsePgenIndex(c->pgen, c->have_pgi ? &c->pgi : nullptr, &c->idx, &err, opts ? opts->raw_variant_ct : 0,
                          opts ? opts->raw_sample_ct : 0);
  if (rc != PD_OK) {
    CopyErr(errbuf, errbuf_len, err);
    delete c;
    return rc;
  }
  if (path && opts && (opts->flags & PD_OPEN_PIN_FILE) && c->pgen.map_writable()) {
    // PD_OPEN_PIN_FILE: page-lock the mapping once; scans then DMA straight from the page cache
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) == cudaSuccess && ndev > 0) {
      int dev = opts->n_devices ? opts->devices[0] : opts->device;
      if (dev >= 0 && dev < ndev) cudaSetDevice(dev);
      if (c->pgen.PinMapping()) c->src_pinned = true;
    } else {
      cudaGetLastError();
    }
  }
  if (opts && opts->n_devices) {
    // one context over several GPUs: this object keeps the index (to partition calls), the shards do the work
    if (!opts->devices) {
      CopyErr(errbuf, errbuf_len, "pd_open_opts.n_devices is set but devices is NULL");
      delete c;
      return PD_ERR_INVALID_ARG;
    }
    // a page-locked mapping is shared by the shards as a memory image (registered once, portable)
    const bool as_mem = !path || c->src_pinned;
    std::vector<uint8_t> pgi_copy;
    const void* pgi_ptr = nullptr;
    uint64_t pgi_len = 0;
    if (c->have_pgi) {
      if (c->pgi.mem()) {
        pgi_ptr = c->pgi.mem();
        pgi_len = c->pgi.size();
      } else if (as_mem) {  // file-backed .pgi next to a pinned .pgen: the shards need it in memory
        c->pgi_image.resize(c->pgi.size());
        if (!c->pgi.ReadAt(0, c->pgi_image.data(), c->pgi.size())) {
          CopyErr(errbuf, errbuf_len, ".pgi read failure");
          delete c;
          return PD_ERR_READ;
        }
        pgi_ptr = c->pgi_image.data();
        pgi_len = c->pgi_image.size();
      }
    }
    rc = GroupOpen(c, opts, as_mem ? nullptr : path, as_mem ? c->pgen.mem() : nullptr, c->pgen.size(), pgi_ptr, pgi_len);
    if (rc != PD_OK) {
      CopyErr(errbuf, errbuf_len, c->last_error);
      DestroyCtx(c);
      return rc;
    }
    *out = c;
    return PD_OK;
  }
  rc = InitDevice(c, opts);
  if (rc != PD_OK) {
    CopyErr(errbuf, errbuf_len, c->last_error);
    DestroyCtx(c);
    return rc;
  }
  *out = c;
  return PD_OK;
}

}  // namespace

extern "C" {

const char* pd_version(void) { return "plinkgpu 0.2 sm_100a"; }

int pd_open(const char* pgen_path, const pd_open_opts* opts, pd_ctx** out, char* errbuf, size_t errbuf_len) {
  if (!out || !pgen_path) return PD_ERR_INVALID_ARG;
  *out = nullptr;
  pd_ctx* c = new (std::nothrow) pd_ctx();
  if (!c) return PD_ERR_NOMEM;
  std::string err;
  const bool want_pin = opts && (opts->flags & PD_OPEN_PIN_FILE);
  if (!c->pgen.OpenFile(pgen_path, &err, want_pin)) {
    CopyErr(errbuf, errbuf_len, err);
    delete c;
    return PD_ERR_OPEN;
  }
  // External index (modes 0x20/0x21): <pgen>.pgi next to the file (pgenlib_read.cc:811-829).
  uint8_t magic[3] = {0, 0, 0};
  if (c->pgen.size() >= 3 && c->pgen.ReadAt(0, magic, 3) && (magic[2] & 0xfe) == 0x20) {
    const std::string pgi_path = std::string(pgen_path) + ".pgi";
    if (!c->pgi.OpenFile(pgi_path.c_str(), &err)) {
      CopyErr(errbuf, errbuf_len, err);
      delete c;
      return PD_ERR_OPEN;
    }
    c->have_pgi = true;
  }
  return FinishOpen(c, opts, out, errbuf, errbuf_len, pgen_path);
}

int pd_open_mem(const void* pgen_bytes, uint64_t len, const void* pgi_bytes, uint64_t pgi_len,
                const pd_open_opts* opts, pd_ctx** out, char* errbuf, size_t errbuf_len) {
  if (!out || !pgen_bytes) return PD_ERR_INVALID_ARG;
  *out = nullptr;
  pd_ctx* c = new (std::nothrow) pd_ctx();
  if (!c) return PD_ERR_NOMEM;
  c->pgen.OpenMem(pgen_bytes, len);
  if (pgi_bytes) {
    c->pgi.OpenMem(pgi_bytes, pgi_len);
    c->have_pgi = true;
  }
  cudaPointerAttributes attr{};
  if (cudaPointerGetAttributes(&attr, pgen_bytes) == cudaSuccess && attr.type == cudaMemoryTypeHost) {
    c->src_pinned = true;
  } else if (attr.type == cudaMemoryTypeDevice) {
    // Image already in HBM (produced by an upstream GPU stage): used in place.  The allocation
    // must extend >= 64 readable bytes past `len` (aligned 16-byte loads on the last record).
    c->pgen.OpenDevice(pgen_bytes, len);
    c->dev_image = true;
  } else {
    cudaGetLastError();  // unregistered host memory reports an error on some drivers
  }
  if (opts && opts->n_devices && c->dev_image) {
    CopyErr(errbuf, errbuf_len, "a device-memory image cannot back a multi-GPU context");
    delete c;
    return PD_ERR_INVALID_ARG;
  }
  int rc = FinishOpen(c, opts, out, errbuf, errbuf_len);
  if (rc == PD_OK && c->dev_image && c->idx.bed) {
    CopyErr(errbuf, errbuf_len, "a PLINK 1 .bed image in device memory is not supported (its codes are rewritten while staging)");
    DestroyCtx(c);
    *out = nullptr;
    return PD_ERR_NOT_SUPPORTED;
  }
  if (rc == PD_OK && c->dev_image) {
    if (c->device != attr.device) {
      CopyErr(errbuf, errbuf_len, "device image lives on a different device than the context");
      DestroyCtx(c);
      *out = nullptr;
      return PD_ERR_INVALID_ARG;
    }
    c->resident = true;
    c->res_ptr = static_cast<const uint8_t*>(pgen_bytes);
    c->res_v0 = 0;
    c->res_v1 = c->idx.variant_ct;
    c->res_f0 = 0;
    c->res_f1 = len;
  }
  return rc;
}

void pd_close(pd_ctx* ctx) { DestroyCtx(ctx); }

int pd_probe(const char* pgen_path, pd_info_t* out, uint8_t* vrtypes, uint64_t* fpos, uint32_t cap, char* errbuf,
             size_t errbuf_len) {
  if (!pgen_path || !out) return PD_ERR_INVALID_ARG;
  ByteSource pgen, pgi;
  std::string err;
  if (!pgen.OpenFile(pgen_path, &err)) {
    CopyErr(errbuf, errbuf_len, err);
    return PD_ERR_OPEN;
  }
  bool have_pgi = false;
  uint8_t magic[3] = {0, 0, 0};
  if (pgen.size() >= 3 && pgen.ReadAt(0, magic, 3) && (magic[2] & 0xfe) == 0x20) {
    const std::string pgi_path = std::string(pgen_path) + ".pgi";
    if (!pgi.OpenFile(pgi_path.c_str(), &err)) {
      CopyErr(errbuf, errbuf_len, err);
      return PD_ERR_OPEN;
    }
    have_pgi = true;
  }
  PgenIndex ix;
  int rc = ParsePgenIndex(pgen, have_pgi ? &pgi : nullptr, &ix, &err);
  if (rc != PD_OK) {
    CopyErr(errbuf, errbuf_len, err);
    return rc;
  }
  memset(out, 0, sizeof(*out));
  out->variant_ct = ix.variant_ct;
  out->sample_ct = ix.sample_ct;
  out->mode = ix.mode;
  out->fixed_width = ix.fixed ? ix.fixed_width : 0;
  out->has_ld = ix.has_ld;
  out->has_difflist = ix.has_difflist;
  out->has_dosage = ix.has_dosage;
  out->has_phase = ix.has_phase;
  out->has_multiallelic = ix.has_multiallelic;
  out->data_bytes = ix.data_bytes;
  if (cap >= ix.variant_ct) {
    if (vrtypes) {
      for (uint32_t v = 0; v < ix.variant_ct; ++v) vrtypes[v] = static_cast<uint8_t>(ix.Vrtype(v));
    }
    if (fpos) {
      for (uint32_t v = 0; v < ix.variant_ct; ++v) fpos[v] = ix.RecOff(v);
      fpos[ix.variant_ct] = ix.data_start + ix.data_bytes;
    }
  }
  return PD_OK;
}

int pd_plan_partition(const char* pgen_path, uint32_t v_begin, uint32_t v_end, uint32_t k, uint32_t* cuts, char* errbuf,
                      size_t errbuf_len) {
  if (!pgen_path || !cuts || !k) return PD_ERR_INVALID_ARG;
  ByteSource pgen, pgi;
  std::string err;
  if (!pgen.OpenFile(pgen_path, &err)) {
    CopyErr(errbuf, errbuf_len, err);
    return PD_ERR_OPEN;
  }
  bool have_pgi = false;
  uint8_t magic[3] = {0, 0, 0};
  if (pgen.size() >= 3 && pgen.ReadAt(0, magic, 3) && (magic[2] & 0xfe) == 0x20) {
    if (!pgi.OpenFile((std::string(pgen_path) + ".pgi").c_str(), &err)) {
      CopyErr(errbuf, errbuf_len, err);
      return PD_ERR_OPEN;
    }
    have_pgi = true;
  }
  PgenIndex ix;
  const int rc = ParsePgenIndex(pgen, have_pgi ? &pgi : nullptr, &ix, &err);
  if (rc != PD_OK) {
    CopyErr(errbuf, errbuf_len, err);
    return rc;
  }
  if (v_begin > v_end || v_end > ix.variant_ct) {
    CopyErr(errbuf, errbuf_len, "variant range out of bounds");
    return PD_ERR_INVALID_ARG;
  }
  const std::vector<VRange> r = PartitionByBytes(ix, v_begin, v_end, k);
  for (uint32_t i = 0; i < k; ++i) cuts[i] = r[i].v0;
  cuts[k] = v_end;
  return PD_OK;
}

const char* pd_last_error(const pd_ctx* ctx) { return ctx ? ctx->last_error.c_str() : "null context"; }

int pd_info(const pd_ctx* c, pd_info_t* out) {
  if (!c || !out) return PD_ERR_INVALID_ARG;
  if (c->group) return GroupInfo(c, out);
  const PgenIndex& ix = c->idx;
  memset(out, 0, sizeof(*out));
  out->variant_ct = ix.variant_ct;
  out->sample_ct = ix.sample_ct;
  out->mode = ix.mode;
  out->fixed_width = ix.fixed ? ix.fixed_width : 0;
  out->has_ld = ix.has_ld;
  out->has_difflist = ix.has_difflist;
  out->has_dosage = ix.has_dosage;
  out->has_phase = ix.has_phase;
  out->has_multiallelic = ix.has_multiallelic;
  out->resident = c->resident;
  out->data_bytes = ix.data_bytes;
  out->resident_bytes = c->resident ? c->res_f1 - c->res_f0 : 0;
  out->effective_sample_ct = c->subset.active ? c->subset.count : ix.sample_ct;
  out->n_devices = 1;
  out->pinned_file = c->pgen.pinned() ? 1u : 0u;
  return PD_OK;
}

int pd_index(const pd_ctx* c, uint8_t* vrtypes, uint64_t* fpos) {
  if (!c) return PD_ERR_INVALID_ARG;
  const PgenIndex& ix = c->idx;
  const uint32_t V = ix.variant_ct;
  if (vrtypes) {
    for (uint32_t v = 0; v < V; ++v) vrtypes[v] = static_cast<uint8_t>(ix.Vrtype(v));
  }
  if (fpos) {
    for (uint32_t v = 0; v < V; ++v) fpos[v] = ix.RecOff(v);
    fpos[V] = ix.data_start + ix.data_bytes;
  }
  return PD_OK;
}

int pd_get_stats(const pd_ctx* c, pd_stats_t* out) {
  if (!c || !out) return PD_ERR_INVALID_ARG;
  if (c->group) return GroupGetStats(c, out);
  *out = c->stats;
  out->launches = c->launch_stats.launches;
  out->n_devices = 1;
  return PD_OK;
}

int pd_reset_stats(pd_ctx* c) {
  if (!c) return PD_ERR_INVALID_ARG;
  if (c->group) return GroupResetStats(c);
  std::lock_guard<std::mutex> lock(c->mu);
  memset(&c->stats, 0, sizeof(c->stats));
  c->launch_stats.launches = 0;
  return PD_OK;
}

int pd_set_sample_subset(pd_ctx* c, const uint64_t* sample_include, uint32_t subset_sample_ct) {
  if (!c) return PD_ERR_INVALID_ARG;
  if (c->group) return GroupSetSubset(c, sample_include, subset_sample_ct);
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  if (!sample_include) {
    c->subset.active = false;
    c->subset.count = c->idx.sample_ct;
    return PD_OK;
  }
  int rc = BuildSubset(c, sample_include, nullptr, &c->subset);
  if (rc != PD_OK) return rc;
  if (subset_sample_ct != c->subset.count) {
    c->subset.active = false;
    return c->SetError(PD_ERR_INVALID_ARG, "subset_sample_ct does not match the popcount of sample_include");
  }
  if (c->subset.count == c->idx.sample_ct) c->subset.active = false;  // full cohort: fast unmasked kernels
  return PD_OK;
}

int pd_set_allele_counts(pd_ctx* c, const uint32_t* allele_cts, uint32_t n) {
  if (!c) return PD_ERR_INVALID_ARG;
  if (allele_cts) {
    if (n != c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "pd_set_allele_counts: one count per variant is required");
    for (uint32_t v = 0; v < n; ++v) {
      if (allele_cts[v] < 2 || allele_cts[v] > 255) {
        return c->SetError(PD_ERR_NOT_SUPPORTED, "pd_set_allele_counts: allele counts must lie in 2..255 (pgenlib's own limit)");
      }
    }
  }
  if (c->group) {
    for (pd_ctx* s : c->group->shards) {
      const int rc = pd_set_allele_counts(s, allele_cts, n);
      if (rc != PD_OK) return c->SetError(rc, s->last_error);
    }
    return PD_OK;
  }
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);  // no kernel may be reading the old table
  if (!allele_cts) {
    c->idx.allele_cts.clear();
    c->d_allele_cts.Free();
    return PD_OK;
  }
  c->idx.allele_cts.resize(n);
  for (uint32_t v = 0; v < n; ++v) c->idx.allele_cts[v] = static_cast<uint8_t>(allele_cts[v]);
  if (!n) return PD_OK;
  cudaError_t e = c->d_allele_cts.Reserve(n);
  if (e == cudaSuccess) e = cudaMemcpy(c->d_allele_cts.p, c->idx.allele_cts.data(), n, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) return CheckCuda(c, e, "allele count upload");
  return PD_OK;
}

int pd_load_resident(pd_ctx* c, uint32_t v_begin, uint32_t v_end) {
  if (!c) return PD_ERR_INVALID_ARG;
  if (c->group) return GroupLoadResident(c, v_begin, v_end);
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  return LoadRange(c, v_begin, v_end, &c->d_resident);
}

}  // extern "C"

// Stages the records of [v_begin, v_end) (extended back to an LD base) into *buf and makes them the context's
// resident range.  Caller holds c->mu.
int pdx::LoadRange(pd_ctx_impl* c, uint32_t v_begin, uint32_t v_end, DevBuf* buf) {
  const PgenIndex& ix = c->idx;
  if (v_begin >= v_end || v_end > ix.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
  const uint32_t v0 = ExtendToBase(c, v_begin);
  const uint64_t f0 = ix.RecOff(v0);
  const uint64_t f1 = ix.fixed ? ix.data_start + static_cast<uint64_t>(v_end) * ix.fixed_width : ix.fpos[v_end];
  const uint64_t len = f1 - f0;
  if (c->dev_image) return PD_OK;  // already in HBM
  c->resident = false;
  cudaError_t e = buf->Reserve(len + 2 * kSlack);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return c->SetError(PD_ERR_NOMEM, std::string("cannot hold the requested range in HBM: ") + cudaGetErrorString(e));
  }
  uint8_t* dst = buf->as<uint8_t>() + kSlack;
  cudaMemsetAsync(buf->p, 0, kSlack, c->stream);
  cudaMemsetAsync(dst + len, 0, kSlack, c->stream);
  if (c->src_pinned) {
    e = CopyFromPinnedSource(c, dst, f0, len, c->stream);
    if (e != cudaSuccess) return CheckCuda(c, e, "resident copy");
  } else {
    // double-buffered chunks through the pinned staging ring
    const uint64_t chunk = std::max<uint64_t>(c->tile_bytes, 1ull << 20);
    for (uint32_t b = 0; b < 2; ++b) {
      if ((e = c->h_stage[b].Reserve(std::min(chunk, len))) != cudaSuccess) return CheckCuda(c, e, "cudaHostAlloc");
    }
    uint64_t off = 0;
    uint32_t i = 0;
    while (off < len) {
      const uint32_t b = i & 1;
      const uint64_t n = std::min(chunk, len - off);
      if (i >= 2) cudaEventSynchronize(c->ev_copied[b]);
      if (!c->pgen.ReadAt(f0 + off, c->h_stage[b].p, n)) return c->SetError(PD_ERR_READ, ".pgen read failure");
      e = cudaMemcpyAsync(dst + off, c->h_stage[b].p, n, cudaMemcpyHostToDevice, c->stream);
      if (e != cudaSuccess) return CheckCuda(c, e, "resident copy");
      cudaEventRecord(c->ev_copied[b], c->stream);
      off += n;
      ++i;
    }
  }
  if (ix.bed) pdk::launch_bed_to_pgen(dst, 0, ix.fixed_width, ix.sample_ct, v_end - v0, c->stream, &c->launch_stats);
  e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "resident copy");
  c->stats.h2d_bytes += len;
  c->resident = true;
  c->res_ptr = dst;
  c->res_v0 = v0;
  c->res_v1 = v_end;
  c->res_f0 = f0;
  c->res_f1 = f1;
  return PD_OK;
}

extern "C" {

int pd_unload(pd_ctx* c) {
  if (!c) return PD_ERR_INVALID_ARG;
  if (c->group) return GroupUnload(c);
  std::lock_guard<std::mutex> lock(c->mu);
  cudaSetDevice(c->device);
  if (c->dev_image) return PD_OK;
  c->resident = false;
  c->d_resident.Free();
  return PD_OK;
}

}  // extern "C"
