// Multi-GPU contexts: one pd_ctx that owns one single-device context ("shard") per GPU of the box.
//
// What the reference does inside one DuckDB process with up to MaxThreads() scan threads -- fetch_add batch
// claims (src/plink_freq.cpp:434-441) and mutex merges of per-thread accumulators (src/plink_missing.cpp:
// 614-619, src/plink_score.cpp:657-664, src/pfile_reader.cpp:3432-3445) -- becomes, still inside one process:
// contiguous variant ranges of equal record BYTES, one per GPU; per-variant outputs land in
// disjoint slices of the caller's arrays (no collective); per-sample accumulators (plink_missing sample mode,
// read_pfile orient := 'sample', plink_score) are summed with ONE ncclAllReduce over NVLink on the device
// buffers the shards leave in HBM, then read back once from the first device.
#pragma once
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "ctx.h"

namespace pdx {

// One persistent host thread per shard: keeps its device current, runs that shard's blocking calls.
class ShardWorker {
 public:
  explicit ShardWorker(int device);
  ~ShardWorker();
  ShardWorker(const ShardWorker&) = delete;
  ShardWorker& operator=(const ShardWorker&) = delete;
  void Post(std::function<int()> fn);  // one outstanding task at a time
  int Wait();                          // result of the posted task

 private:
  void Loop();
  int device_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::function<int()> task_;
  bool has_task_ = false, done_ = true, stop_ = false;
  int result_ = 0;
  std::thread th_;
};

struct VRange {
  uint32_t v0, v1;
};

struct GroupState {
  std::vector<int> devices;
  std::vector<pd_ctx*> shards;
  std::vector<ShardWorker*> workers;
  std::vector<void*> comms;          // ncclComm_t per shard (one communicator, created with ncclCommInitAll)
  bool resident = false;
  std::vector<VRange> res_ranges;    // what each shard holds in HBM after pd_load_resident
  uint32_t res_v0 = 0, res_v1 = 0;
  double allreduce_ms = 0;           // last call
  uint64_t allreduce_bytes = 0;
  cudaEvent_t ev_r0 = nullptr, ev_r1 = nullptr;  // on the first shard's device
};

// Cuts [v_begin, v_end) into k contiguous ranges of (nearly) equal record bytes; starts are moved back to
// the closest variant whose main track is not LD-compressed (any cut is correct -- a shard re-reads a missing
// LD base itself, ExtendToBase -- this only avoids the re-read).
std::vector<VRange> PartitionByBytes(const PgenIndex& ix, uint32_t v_begin, uint32_t v_end, uint32_t k);
// Same for an ascending variant list: index ranges [i0, i1) into vidx.
std::vector<VRange> PartitionListByBytes(const PgenIndex& ix, const uint32_t* vidx, uint32_t n, uint32_t k,
                                         uint32_t row_multiple = 1);

// Group forms of the entry points (include/plinkgpu.h); each extern "C" function dispatches here when
// ctx->group is set.
int GroupOpen(pd_ctx* parent, const pd_open_opts* opts, const char* path, const void* mem, uint64_t mem_len,
              const void* pgi, uint64_t pgi_len);
void GroupDestroy(pd_ctx* c);
int GroupInfo(const pd_ctx* c, pd_info_t* out);
int GroupGetStats(const pd_ctx* c, pd_stats_t* out);
int GroupResetStats(pd_ctx* c);
int GroupSetSubset(pd_ctx* c, const uint64_t* include, uint32_t ct);
int GroupLoadResident(pd_ctx* c, uint32_t v_begin, uint32_t v_end);
int GroupUnload(pd_ctx* c);
// Per-variant outputs: fn(shard, v0, v1, row_offset) runs on each shard's worker for its sub-range.
int GroupRunRanges(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t row_multiple,
                   const std::function<int(pd_ctx*, uint32_t, uint32_t, uint32_t)>& fn);
int GroupRunList(pd_ctx* c, const uint32_t* vidx, uint32_t n, uint32_t row_multiple,
                 const std::function<int(pd_ctx*, uint32_t, uint32_t)>& fn);
int GroupMissingBySample(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t* out_missing);
int GroupSampleCounts(pd_ctx* c, uint32_t v_begin, uint32_t v_end, uint32_t* het, uint32_t* hom_alt, uint32_t* missing);
int GroupScore(pd_ctx* c, const uint32_t* vidx, const uint8_t* flip, const double* weights, uint32_t n_var, uint32_t K,
               uint32_t mode, double* score_sums, double* dosage_sums, uint32_t* allele_cts);
int GroupCountsStart(pd_ctx* c, uint32_t v_begin, uint32_t v_end, bool hardy, uint32_t midp, pd_job** out);

}  // namespace pdx
