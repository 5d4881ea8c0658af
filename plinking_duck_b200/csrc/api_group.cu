// Multi-GPU contexts (see group.h): shard contexts, variant-range partition, NCCL all-reduce of the
// per-sample accumulators.  NCCL is bound at run time (dlopen of libnccl.so.2), so the library itself carries
// no link-time dependency on it and a single-GPU host does not need it installed.
#include "group.h"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <new>

namespace pdx {

// ---------------------------------------------------------------------------------------------------
// NCCL, bound lazily
// ---------------------------------------------------------------------------------------------------
namespace {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string error;

  bool Load() {
    if (handle) return true;
    if (!error.empty()) return false;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (handle) break;
    }
    if (!handle) {
      error = std::string("cannot load NCCL (libnccl.so.2): ") + dlerror();
      return false;
    }
    auto sym = [&](const char* n) { return dlsym(handle, n); };
    CommInitAll = reinterpret_cast<decltype(CommInitAll)>(sym("ncclCommInitAll"));
    CommDestroy = reinterpret_cast<decltype(CommDestroy)>(sym("ncclCommDestroy"));
    AllReduce = reinterpret_cast<decltype(AllReduce)>(sym("ncclAllReduce"));
    GroupStart = reinterpret_cast<decltype(GroupStart)>(sym("ncclGroupStart"));
    GroupEnd = reinterpret_cast<decltype(GroupEnd)>(sym("ncclGroupEnd"));
    GetErrorString = reinterpret_cast<decltype(GetErrorString)>(sym("ncclGetErrorString"));
    if (!CommInitAll || !CommDestroy || !AllReduce || !GroupStart || !GroupEnd || !GetErrorString) {
      error = "libnccl.so.2 lacks a required symbol";
      dlclose(handle);
      handle = nullptr;
      return false;
    }
    return true;
  }
};

NcclApi& Nccl() {
  static NcclApi api;  // function pointers of a shared library: process-wide by nature, holds no device state
  return api;
}
std::mutex g_nccl_mu;

}  // namespace

// ---------------------------------------------------------------------------------------------------
// ShardWorker
// ---------------------------------------------------------------------------------------------------
ShardWorker::ShardWorker(int device) : device_(device), th_([this] { Loop(); }) {}

ShardWorker::~ShardWorker() {
  {
    std::lock_guard<std::mutex> lk(mu_);
    stop_ = true;
  }
  cv_.notify_all();
  th_.join();
}

void ShardWorker::Post(std::function<int()> fn) {
  {
    std::lock_guard<std::mutex> lk(mu_);
    task_ = std::move(fn);
    has_task_ = true;
    done_ = false;
  }
  cv_.notify_all();
}

int ShardWorker::Wait() {
  std::unique_lock<std::mutex> lk(mu_);
  cv_.wait(lk, [&] { return done_; });
  return result_;
}

void ShardWorker::Loop() {
  cudaSetDevice(device_);
  for (;;) {
    std::function<int()> fn;
    {
      std::unique_lock<std::mutex> lk(mu_);
      cv_.wait(lk, [&] { return has_task_ || stop_; });
      if (stop_ && !has_task_) return;
      fn = std::move(task_);
      has_task_ = false;
    }
    const int r = fn();
    {
      std::lock_guard<std::mutex> lk(mu_);
      result_ = r;
      done_ = true;
    }
    cv_.notify_all();
  }
}

// ---------------------------------------------------------------------------------------------------
// Partition
// ---------------------------------------------------------------------------------------------------
std::vector<VRange> PartitionByBytes(const PgenIndex& ix, uint32_t v_begin, uint32_t v_end, uint32_t k) {
  std::vector<VRange> out(k, VRange{v_end, v_end});
  if (!k) return out;
  std::vector<uint32_t> cuts(k + 1, v_end);
  cuts[0] = v_begin;
  const uint64_t f0 = v_begin < v_end ? ix.RecOff(v_begin) : 0;
  const uint64_t f1 = v_begin < v_end ? ix.RecOff(v_end - 1) + (ix.fixed ? ix.fixed_width : ix.fpos[v_end] - ix.fpos[v_end - 1]) : 0;
  const uint64_t total = f1 - f0;
  for (uint32_t r = 1; r < k; ++r) {
    uint32_t v;
    if (v_begin >= v_end) {
      v = v_end;
    } else if (ix.fixed) {
      v = v_begin + static_cast<uint32_t>(static_cast<uint64_t>(v_end - v_begin) * r / k);
    } else {
      const uint64_t target = f0 + total / k * r + (total % k) * r / k;
      v = static_cast<uint32_t>(std::lower_bound(ix.fpos.begin() + v_begin, ix.fpos.begin() + v_end, target) - ix.fpos.begin());
      while (v > cuts[r - 1] && v < v_end && ix.ldbase[v] != v) --v;
    }
    cuts[r] = std::min(std::max(v, cuts[r - 1]), v_end);
  }
  for (uint32_t r = 0; r < k; ++r) out[r] = VRange{cuts[r], cuts[r + 1]};
  return out;
}

std::vector<VRange> PartitionListByBytes(const PgenIndex& ix, const uint32_t* vidx, uint32_t n, uint32_t k,
                                         uint32_t row_multiple) {
  std::vector<VRange> out(k, VRange{n, n});
  if (!k) return out;
  std::vector<uint64_t> cum(static_cast<size_t>(n) + 1, 0);
  for (uint32_t i = 0; i < n; ++i) {
    const uint32_t v = vidx[i];
    const uint64_t w = v < ix.variant_ct ? (ix.fixed ? ix.fixed_width : ix.fpos[v + 1] - ix.fpos[v]) : 0;
    cum[i + 1] = cum[i] + w + 1;  // + 1: empty (all-ref) records still cost a row
  }
  std::vector<uint32_t> cuts(k + 1, n);
  cuts[0] = 0;
  for (uint32_t r = 1; r < k; ++r) {
    const uint64_t target = cum[n] / k * r + (cum[n] % k) * r / k;
    uint32_t i = static_cast<uint32_t>(std::lower_bound(cum.begin(), cum.end(), target) - cum.begin());
    if (row_multiple > 1) i = i / row_multiple * row_multiple;
    cuts[r] = std::min(std::max(i, cuts[r - 1]), n);
  }
  for (uint32_t r = 0; r < k; ++r) out[r] = VRange{cuts[r], cuts[r + 1]};
  return out;
}

// ---------------------------------------------------------------------------------------------------
// Group plumbing
// ---------------------------------------------------------------------------------------------------
namespace {

// Runs fn(shard index) on every shard's worker; returns the first failure and copies its message.
int ForShards(pd_ctx* c, const std::function<int(uint32_t)>& fn) {
  GroupState* g = c->group;
  const uint32_t k = static_cast<uint32_t>(g->shards.size());
  for (uint32_t i = 0; i < k; ++i) g->workers[i]->Post([&fn, i] { return fn(i); });
  int rc = PD_OK;
  for (uint32_t i = 0; i < k; ++i) {
    const int r = g->workers[i]->Wait();
    if (r != PD_OK && rc == PD_OK) {
      rc = r;
      c->last_error = g->shards[i]->last_error + " (device " + std::to_string(g->devices[i]) + ")";
    }
  }
  return rc;
}

// Ranges of a call: the resident split when the range is held in HBM, else equal bytes.
std::vector<VRange> RangesFor(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t row_multiple) {
  GroupState* g = c->group;
  const uint32_t k = static_cast<uint32_t>(g->shards.size());
  std::vector<VRange> r;
  if (g->resident && v_begin >= g->res_v0 && v_end <= g->res_v1 && v_begin < v_end) {
    r.resize(k);
    for (uint32_t i = 0; i < k; ++i) {
      const uint32_t a = std::min(std::max(v_begin, g->res_ranges[i].v0), v_end);
      const uint32_t b = std::max(std::min(v_end, g->res_ranges[i].v1), a);
      r[i] = VRange{a, b};
    }
  } else {
    r = PartitionByBytes(c->idx, v_begin, v_end, k);
  }
  if (row_multiple > 1) {
    // cut points at multiples of row_multiple rows from v_begin (validity words must not straddle shards)
    for (uint32_t i = 1; i < k; ++i) {
      uint32_t cut = r[i].v0;
      cut = v_begin + (cut - v_begin) / row_multiple * row_multiple;
      cut = std::max(cut, r[i - 1].v0);
      r[i - 1].v1 = cut;
      r[i].v0 = cut;
    }
  }
  return r;
}

int NcclFail(pd_ctx* c, ncclResult_t r, const char* what) {
  return c->SetError(PD_ERR_CUDA, std::string("NCCL failure in ") + what + ": " + Nccl().GetErrorString(r));
}

// In-place sum of one device buffer per shard (same count everywhere), on the shards' compute streams.
int AllReduce(pd_ctx* c, const std::vector<void*>& bufs, size_t count, ncclDataType_t type) {
  GroupState* g = c->group;
  const uint32_t k = static_cast<uint32_t>(g->shards.size());
  if (k == 1 || !count) return PD_OK;
  NcclApi& api = Nccl();
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  ncclResult_t r = api.GroupStart();
  if (r != ncclSuccess) return NcclFail(c, r, "ncclGroupStart");
  for (uint32_t i = 0; i < k && r == ncclSuccess; ++i) {
    r = api.AllReduce(bufs[i], bufs[i], count, type, ncclSum, static_cast<ncclComm_t>(g->comms[i]), g->shards[i]->stream);
  }
  const ncclResult_t r2 = api.GroupEnd();
  if (r != ncclSuccess) return NcclFail(c, r, "ncclAllReduce");
  if (r2 != ncclSuccess) return NcclFail(c, r2, "ncclGroupEnd");
  g->allreduce_bytes += count * (type == ncclFloat64 ? 8 : 4);
  return PD_OK;
}

// Brackets the collectives of one call with events on the first shard's stream.
struct ReduceTimer {
  pd_ctx* c;
  explicit ReduceTimer(pd_ctx* ctx) : c(ctx) {
    GroupState* g = c->group;
    g->allreduce_ms = 0;
    g->allreduce_bytes = 0;
    cudaSetDevice(g->devices[0]);
    cudaEventRecord(g->ev_r0, g->shards[0]->stream);
  }
  void Stop() {
    GroupState* g = c->group;
    cudaSetDevice(g->devices[0]);
    cudaEventRecord(g->ev_r1, g->shards[0]->stream);
  }
  void Read() {
    GroupState* g = c->group;
    float ms = 0;
    if (cudaEventElapsedTime(&ms, g->ev_r0, g->ev_r1) == cudaSuccess) g->allreduce_ms = ms;
  }
};

int SyncAll(pd_ctx* c) {
  GroupState* g = c->group;
  for (size_t i = 0; i < g->shards.size(); ++i) {
    cudaSetDevice(g->devices[i]);
    const cudaError_t e = cudaStreamSynchronize(g->shards[i]->stream);
    if (e != cudaSuccess) return CheckCuda(c, e, "all-reduce");
  }
  return PD_OK;
}

int Download(pd_ctx* c, void* host, const void* dev, size_t bytes) {
  GroupState* g = c->group;
  pd_ctx* s0 = g->shards[0];
  cudaSetDevice(g->devices[0]);
  cudaError_t e = cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, s0->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s0->stream);
  if (e != cudaSuccess) return CheckCuda(c, e, "result download");
  s0->stats.d2h_bytes += bytes;
  return PD_OK;
}

}  // namespace

int GroupOpen(pd_ctx* parent, const pd_open_opts* opts, const char* path, const void* mem, uint64_t mem_len,
              const void* pgi, uint64_t pgi_len) {
  const uint32_t k = opts->n_devices;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    return parent->SetError(PD_ERR_CUDA, std::string("no usable CUDA device (libplinkgpu has no CPU fallback): ") +
                                             (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
  }
  for (uint32_t i = 0; i < k; ++i) {
    if (opts->devices[i] < 0 || opts->devices[i] >= ndev) return parent->SetError(PD_ERR_INVALID_ARG, "device ordinal out of range");
    for (uint32_t j = 0; j < i; ++j) {
      if (opts->devices[j] == opts->devices[i]) return parent->SetError(PD_ERR_INVALID_ARG, "device list holds a device twice");
    }
  }
  GroupState* g = new (std::nothrow) GroupState();
  if (!g) return PD_ERR_NOMEM;
  parent->group = g;
  g->devices.assign(opts->devices, opts->devices + k);
  char errbuf[512];
  for (uint32_t i = 0; i < k; ++i) {
    pd_open_opts so = *opts;
    so.device = g->devices[i];
    so.n_devices = 0;
    so.devices = nullptr;
    pd_ctx* s = nullptr;
    errbuf[0] = 0;
    const int rc = path ? pd_open(path, &so, &s, errbuf, sizeof(errbuf))
                        : pd_open_mem(mem, mem_len, pgi, pgi_len, &so, &s, errbuf, sizeof(errbuf));
    if (rc != PD_OK) return parent->SetError(rc, std::string(errbuf) + " (device " + std::to_string(g->devices[i]) + ")");
    if (s->dev_image) {
      pd_close(s);
      return parent->SetError(PD_ERR_INVALID_ARG, "a device-memory image cannot back a multi-GPU context");
    }
    g->shards.push_back(s);
    g->workers.push_back(new ShardWorker(g->devices[i]));
  }
  // peer access between every pair (NCCL uses it on its own; an enabled peer costs nothing)
  for (uint32_t i = 0; i < k; ++i) {
    cudaSetDevice(g->devices[i]);
    for (uint32_t j = 0; j < k; ++j) {
      if (i == j) continue;
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, g->devices[i], g->devices[j]) == cudaSuccess && can) {
        if (cudaDeviceEnablePeerAccess(g->devices[j], 0) != cudaSuccess) cudaGetLastError();  // already enabled
      }
    }
  }
  cudaSetDevice(g->devices[0]);
  cudaEventCreate(&g->ev_r0);
  cudaEventCreate(&g->ev_r1);
  if (k > 1) {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    NcclApi& api = Nccl();
    if (!api.Load()) return parent->SetError(PD_ERR_CUDA, api.error);
    std::vector<ncclComm_t> comms(k);
    const ncclResult_t r = api.CommInitAll(comms.data(), static_cast<int>(k), g->devices.data());
    if (r != ncclSuccess) return NcclFail(parent, r, "ncclCommInitAll");
    for (uint32_t i = 0; i < k; ++i) g->comms.push_back(comms[i]);
  }
  return PD_OK;
}

void GroupDestroy(pd_ctx* c) {
  GroupState* g = c->group;
  if (!g) return;
  for (auto* w : g->workers) delete w;
  if (!g->comms.empty()) {
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    for (void* cm : g->comms) Nccl().CommDestroy(static_cast<ncclComm_t>(cm));
  }
  if (g->ev_r0) {
    cudaSetDevice(g->devices[0]);
    cudaEventDestroy(g->ev_r0);
    cudaEventDestroy(g->ev_r1);
  }
  for (pd_ctx* s : g->shards) pd_close(s);
  delete g;
  c->group = nullptr;
}

int GroupInfo(const pd_ctx* c, pd_info_t* out) {
  const GroupState* g = c->group;
  int rc = pd_info(g->shards[0], out);
  if (rc != PD_OK) return rc;
  out->resident = g->resident;
  out->resident_bytes = 0;
  for (pd_ctx* s : g->shards) {
    pd_info_t si;
    pd_info(s, &si);
    out->resident_bytes += si.resident_bytes;
  }
  out->n_devices = static_cast<uint32_t>(g->shards.size());
  out->pinned_file = c->pgen.pinned() ? 1u : 0u;
  return PD_OK;
}

int GroupGetStats(const pd_ctx* c, pd_stats_t* out) {
  const GroupState* g = c->group;
  memset(out, 0, sizeof(*out));
  for (pd_ctx* s : g->shards) {
    pd_stats_t t;
    pd_get_stats(s, &t);
    out->launches += t.launches;
    out->h2d_bytes += t.h2d_bytes;
    out->d2h_bytes += t.d2h_bytes;
    out->record_bytes += t.record_bytes;
    out->scan_launches += t.scan_launches;
    // the devices run side by side: times are the slowest shard's (max over ranks)
    out->kernel_ms = std::max(out->kernel_ms, t.kernel_ms);
    out->aux_kernel_ms = std::max(out->aux_kernel_ms, t.aux_kernel_ms);
    out->total_ms = std::max(out->total_ms, t.total_ms);
    out->call_ms = std::max(out->call_ms, t.call_ms);
    out->host_stage_ms = std::max(out->host_stage_ms, t.host_stage_ms);
    out->host_wait_ms = std::max(out->host_wait_ms, t.host_wait_ms);
  }
  out->allreduce_ms = g->allreduce_ms;
  out->allreduce_bytes = g->allreduce_bytes;
  out->n_devices = static_cast<uint32_t>(g->shards.size());
  return PD_OK;
}

int GroupResetStats(pd_ctx* c) {
  for (pd_ctx* s : c->group->shards) pd_reset_stats(s);
  c->group->allreduce_ms = 0;
  c->group->allreduce_bytes = 0;
  return PD_OK;
}

int GroupSetSubset(pd_ctx* c, const uint64_t* include, uint32_t ct) {
  std::lock_guard<std::mutex> lock(c->mu);
  GroupState* g = c->group;
  return ForShards(c, [&](uint32_t i) { return pd_set_sample_subset(g->shards[i], include, ct); });
}

int GroupLoadResident(pd_ctx* c, uint32_t v_begin, uint32_t v_end) {
  std::lock_guard<std::mutex> lock(c->mu);
  GroupState* g = c->group;
  if (v_begin >= v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
  g->resident = false;
  const std::vector<VRange> r = PartitionByBytes(c->idx, v_begin, v_end, static_cast<uint32_t>(g->shards.size()));
  const int rc = ForShards(c, [&](uint32_t i) {
    if (r[i].v0 >= r[i].v1) return pd_unload(g->shards[i]);
    return pd_load_resident(g->shards[i], r[i].v0, r[i].v1);
  });
  if (rc != PD_OK) return rc;
  g->res_ranges = r;
  g->res_v0 = v_begin;
  g->res_v1 = v_end;
  g->resident = true;
  return PD_OK;
}

int GroupUnload(pd_ctx* c) {
  std::lock_guard<std::mutex> lock(c->mu);
  GroupState* g = c->group;
  g->resident = false;
  return ForShards(c, [&](uint32_t i) { return pd_unload(g->shards[i]); });
}

int GroupRunRanges(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t row_multiple,
                   const std::function<int(pd_ctx*, uint32_t, uint32_t, uint32_t)>& fn) {
  std::lock_guard<std::mutex> lock(c->mu);
  if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
  GroupState* g = c->group;
  const std::vector<VRange> r = RangesFor(c, v_begin, v_end, row_multiple);
  return ForShards(c, [&](uint32_t i) {
    if (r[i].v0 >= r[i].v1) return static_cast<int>(PD_OK);
    return fn(g->shards[i], r[i].v0, r[i].v1, r[i].v0 - v_begin);
  });
}

int GroupRunList(pd_ctx* c, const uint32_t* vidx, uint32_t n, uint32_t row_multiple,
                 const std::function<int(pd_ctx*, uint32_t, uint32_t)>& fn) {
  std::lock_guard<std::mutex> lock(c->mu);
  GroupState* g = c->group;
  for (uint32_t i = 0; i < n; ++i) {
    if (vidx[i] >= c->idx.variant_ct || (i && vidx[i] <= vidx[i - 1])) {
      return c->SetError(PD_ERR_INVALID_ARG, "variant list must be strictly ascending and inside the file");
    }
  }
  const std::vector<VRange> r = PartitionListByBytes(c->idx, vidx, n, static_cast<uint32_t>(g->shards.size()), row_multiple);
  return ForShards(c, [&](uint32_t i) {
    if (r[i].v0 >= r[i].v1) return static_cast<int>(PD_OK);
    return fn(g->shards[i], r[i].v0, r[i].v1);
  });
}

// plink_missing sample mode / read_pfile orient := 'sample': each shard accumulates its variant range into
// its own uint32[3 x n] (left in HBM), one all-reduce, one download (the mutex merges of
// src/plink_missing.cpp:614-619 and src/pfile_reader.cpp:3432-3445).
static int GroupSampleAcc(pd_ctx* c, uint32_t v_begin, uint32_t v_end, bool all_three, uint32_t* het, uint32_t* hom_alt,
                          uint32_t* missing) {
  std::lock_guard<std::mutex> lock(c->mu);
  if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
  GroupState* g = c->group;
  const uint32_t k = static_cast<uint32_t>(g->shards.size());
  const std::vector<VRange> r = RangesFor(c, v_begin, v_end, 1);
  int rc = ForShards(c, [&](uint32_t i) {
    const uint32_t a = r[i].v0, b = std::max(r[i].v0, r[i].v1);
    return all_three ? pd_sample_counts(g->shards[i], a, b, nullptr, nullptr, nullptr)
                     : pd_missing_by_sample(g->shards[i], a, b, nullptr);
  });
  if (rc != PD_OK) return rc;
  const uint32_t n = g->shards[0]->last_sample_n;
  std::vector<void*> bufs(k);
  for (uint32_t i = 0; i < k; ++i) {
    bufs[i] = all_three ? g->shards[i]->d_out[0].p : static_cast<void*>(g->shards[i]->d_out[0].as<uint32_t>() + 2 * static_cast<size_t>(n));
  }
  ReduceTimer timer(c);
  rc = AllReduce(c, bufs, static_cast<size_t>(n) * (all_three ? 3 : 1), ncclUint32);
  timer.Stop();
  if (rc == PD_OK) rc = SyncAll(c);
  if (rc != PD_OK) return rc;
  timer.Read();
  const uint32_t* d = g->shards[0]->d_out[0].as<uint32_t>();
  const size_t bytes = static_cast<size_t>(n) * 4;
  if (all_three && het) {
    rc = Download(c, het, d, bytes);
    if (rc == PD_OK) rc = Download(c, hom_alt, d + n, bytes);
  }
  if (rc == PD_OK && missing) rc = Download(c, missing, d + 2 * static_cast<size_t>(n), bytes);
  return rc;
}

int GroupMissingBySample(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t* out_missing) {
  return GroupSampleAcc(c, v_begin, v_end, false, nullptr, nullptr, out_missing);
}

int GroupSampleCounts(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t* het, uint32_t* hom_alt, uint32_t* missing) {
  const bool want = het || hom_alt || missing;
  if (want && !(het && hom_alt && missing)) return PD_ERR_INVALID_ARG;
  return GroupSampleAcc(c, v_begin, v_end, true, het, hom_alt, missing);
}

// plink_score: the scored-variant list (sorted by variant, src/plink_score.cpp:407) is cut into contiguous
// chunks of equal record bytes, each shard scores its chunk with its slice of the weights, and the per-sample
// partial sums are all-reduced -- the one collective of the path (merge of src/plink_score.cpp:657-664).
int GroupScore(pd_ctx* c, const uint32_t* vidx, const uint8_t* flip, const double* weights, uint32_t n_var, uint32_t K,
               uint32_t mode, double* score_sums, double* dosage_sums, uint32_t* allele_cts) {
  if (!K || mode > 2 || (n_var && (!vidx || !weights))) return PD_ERR_INVALID_ARG;
  std::lock_guard<std::mutex> lock(c->mu);
  GroupState* g = c->group;
  const uint32_t k = static_cast<uint32_t>(g->shards.size());
  for (uint32_t i = 0; i < n_var; ++i) {
    if (vidx[i] >= c->idx.variant_ct || (i && vidx[i] <= vidx[i - 1])) {
      return c->SetError(PD_ERR_INVALID_ARG, "variant list must be strictly ascending and inside the file");
    }
  }
  const std::vector<VRange> r = PartitionListByBytes(c->idx, vidx, n_var, k, 1);
  int rc = ForShards(c, [&](uint32_t i) {
    const uint32_t a = r[i].v0, n = r[i].v1 > r[i].v0 ? r[i].v1 - r[i].v0 : 0;
    return pd_score(g->shards[i], vidx + a, flip ? flip + a : nullptr, weights + static_cast<size_t>(a) * K, n, K, mode,
                    nullptr, nullptr, nullptr);
  });
  if (rc != PD_OK) return rc;
  const uint32_t n = g->shards[0]->last_score_n;
  if (!n) return PD_OK;
  std::vector<void*> s(k), d(k), a(k);
  for (uint32_t i = 0; i < k; ++i) {
    s[i] = g->shards[i]->d_score.p;
    d[i] = g->shards[i]->d_dsum.p;
    a[i] = g->shards[i]->d_act.p;
  }
  ReduceTimer timer(c);
  rc = AllReduce(c, s, static_cast<size_t>(n) * K, ncclFloat64);
  if (rc == PD_OK) rc = AllReduce(c, d, n, ncclFloat64);
  if (rc == PD_OK) rc = AllReduce(c, a, n, ncclUint32);
  timer.Stop();
  if (rc == PD_OK) rc = SyncAll(c);
  if (rc != PD_OK) return rc;
  timer.Read();
  if (score_sums) rc = Download(c, score_sums, s[0], static_cast<size_t>(n) * K * 8);
  if (rc == PD_OK && dosage_sums) rc = Download(c, dosage_sums, d[0], static_cast<size_t>(n) * 8);
  if (rc == PD_OK && allele_cts) rc = Download(c, allele_cts, a[0], static_cast<size_t>(n) * 4);
  return rc;
}

int GroupCountsStart(pd_ctx* c, uint32_t v_begin, uint32_t v_end, bool hardy, uint32_t midp, pd_job** out) {
  std::lock_guard<std::mutex> lock(c->mu);
  if (v_begin > v_end || v_end > c->idx.variant_ct) return c->SetError(PD_ERR_INVALID_ARG, "variant range out of bounds");
  GroupState* g = c->group;
  pd_job* j = new (std::nothrow) pd_job();
  if (!j) return PD_ERR_NOMEM;
  j->ctx = c;
  j->v_begin = v_begin;
  j->v_end = v_end;
  j->hardy = hardy;
  j->midp = midp;
  const std::vector<VRange> r = RangesFor(c, v_begin, v_end, 1);
  for (size_t i = 0; i < g->shards.size(); ++i) {
    if (r[i].v0 >= r[i].v1) continue;
    pd_job* sub = nullptr;
    const int rc = StartScanJob(g->shards[i], r[i].v0, r[i].v1, hardy, midp, &sub);
    if (rc != PD_OK) {
      c->last_error = g->shards[i]->last_error;
      pd_job_free(j);
      return rc;
    }
    j->subs.push_back(sub);
  }
  *out = j;
  return PD_OK;
}

}  // namespace pdx
