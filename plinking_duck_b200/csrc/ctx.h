// Host-side context of libplinkgpu: index, device buffers, staging ring, scan executor.
#pragma once
#include <cuda_runtime.h>

#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/plinkgpu.h"
#include "kernels.cuh"
#include "pgen_index.h"

struct pd_ctx;
struct pd_job;

namespace pdx {

struct GroupState;  // group.h
struct DecodeJob;   // api_async.cu
void FreeDecodeJob(pd_job* j);

constexpr uint32_t kSlack = 64;  // readable bytes before/after every record buffer (aligned 16 B loads)

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  // Grow-only device allocation.
  cudaError_t Reserve(size_t bytes);
  void Free();
  template <class T>
  T* as() const { return static_cast<T*>(p); }
};

struct PinnedBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t Reserve(size_t bytes);
  void Free();
  template <class T>
  T* as() const { return static_cast<T*>(p); }
};

// Device-side form of one sample mask (the context subset, or subset AND stratum).
struct SubsetState {
  bool active = false;
  uint32_t count = 0;
  std::vector<uint32_t> words;  // N bits as 32-bit words, padded
  DevBuf d_words, d_cum, d_mask2, d_mask1, d_idx;  // d_idx: raw index of every subset sample, ascending
  uint32_t stride2 = 0, stride1 = 0;
  pdk::SubsetDev Dev() const;
  void Free();
};

struct Tile {
  uint32_t v0, v1;      // variants staged (v0 is never LD-compressed)
  uint64_t f0, f1;      // file byte range
};

struct pd_ctx_impl {
  ByteSource pgen, pgi;
  std::vector<uint8_t> pgi_image;  // multi-GPU context over a pinned file: the .pgi bytes its shards parse
  bool have_pgi = false;
  PgenIndex idx;
  int device = 0;
  uint64_t tile_bytes = 64ull << 20;
  uint32_t n_staging = 3;
  bool src_pinned = false;  // memory image is page-locked: DMA straight from it
  // Trimmed staging (unmasked genotype counts over variable-width files): difflist records that are not an
  // LD base are staged without their sample-id deltas -- counting never reads them (K3), and they are
  // ~85 % of such a record.  trim_fpos[v] = offset of record v in that compacted stream.
  std::vector<uint64_t> trim_fpos;  // [V+1], built on first use
  DevBuf d_trim_fpos;
  DevBuf d_phase_raw;               // pd_decode_phased with a sample subset: unsubsetted pair rows
  DevBuf d_tidy_off;                // pd_tidy_i8: row offset of every variant of the call
  DevBuf d_pairs;                   // pd_ld_pairs: row positions of the pairs; pd_standardized_f64: per-row norms
  DevBuf d_span;                    // pd_ld_pairs: the variant span of a call, held in HBM for its duration
  int trim_state = 0;               // 0 not built, 1 usable, -1 not worth it / not possible
  bool trim_scan = false;           // set by the caller around RunScan

  cudaStream_t stream = nullptr;       // compute
  cudaStream_t copy_stream = nullptr;  // H2D staging
  std::mutex mu;
  std::string last_error;

  // device copies of the index (variable-width files)
  DevBuf d_fpos, d_vrtypes, d_ldbase, d_base_slot, d_slot_vidx;
  DevBuf d_allele_cts;  // [V] bytes, only for files that store allele counts
  DevBuf d_mhet;        // K5p: per-CTA bit arrays of ALTx/ALTy calls (multiallelic + phased records)
  std::vector<uint32_t> base_slot;  // [V]: scratch slot of the LD base (non-type-0 bases), else ~0u
  std::vector<uint32_t> slot_vidx;  // slot -> base variant

  // residency
  DevBuf d_resident;
  const uint8_t* res_ptr = nullptr;  // device byte holding file offset res_f0
  bool dev_image = false;            // the caller's image is device memory: always resident, never copied
  bool resident = false;
  uint32_t res_v0 = 0, res_v1 = 0;
  uint64_t res_f0 = 0, res_f1 = 0;

  // staging ring
  std::vector<PinnedBuf> h_stage;
  std::vector<DevBuf> d_stage;
  std::vector<cudaEvent_t> ev_copied, ev_done;

  // work buffers
  DevBuf d_counts, d_scratch, d_misc, d_pvals, d_acc, d_ready, d_hweq;  // d_hweq: exact-test queue words + overflow list
  const pdk::SparseAcc* sparse_acc = nullptr;  // set by SampleAccumulate around RunRows: difflist records skip densify
  DevBuf d_dense, d_rowptrs, d_vlist, d_out[2];  // dense-row batches and their outputs
  DevBuf d_loc, d_rank, d_dstats, d_srow, d_weights, d_flip, d_score, d_dsum, d_act, d_tc;  // dosage / score
  PinnedBuf h_bounce[2];
  cudaStream_t d2h_stream = nullptr;
  cudaEvent_t ev_out_ready[2] = {nullptr, nullptr}, ev_out_free[2] = {nullptr, nullptr};
  cudaEvent_t ev_prod0[2] = {nullptr, nullptr}, ev_prod1[2] = {nullptr, nullptr};  // producer kernel timing
  PinnedBuf h_misc;   // int err[2] + small readbacks
  PinnedBuf h_out;    // pinned bounce buffer for results
  cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_t0 = nullptr, ev_t1 = nullptr, ev_c0 = nullptr, ev_c1 = nullptr;
  std::vector<cudaEvent_t> ev_pool;

  SubsetState subset;
  uint32_t last_score_n = 0, last_score_k = 0, last_sample_n = 0;  // extents of the results left on the device

  pdk::LaunchStats launch_stats;
  pd_stats_t stats{};
  GroupState* group = nullptr;  // multi-GPU context: this object only holds the index; the shards do the work

  int SetError(int code, const std::string& msg) {
    last_error = msg;
    return code;
  }
};

// Per-tile callback: records of [t.v0, t.v1) are reachable through ix; enqueue kernels on s.
using TileFn = std::function<int(const Tile& t, const pdk::RecIndex& ix, cudaStream_t s)>;

// Plans tiles for [v_begin, v_end) (extended back to the LD base of v_begin), stages each one
// (or points at the resident copy) and calls fn.  *v_ext receives the first staged variant.
// after_tile (optional) runs on the host after tile i's kernels are enqueued, with the event that
// marks their completion.
int RunScan(pd_ctx_impl* c, uint32_t v_begin, uint32_t v_end, uint32_t* v_ext, const TileFn& fn,
            const std::function<void(const Tile&, cudaEvent_t)>& after_tile = nullptr);

int CheckCuda(pd_ctx_impl* c, cudaError_t e, const char* what);
// H2D copy of file bytes [off, off + len) of a page-locked source (memory image or pinned mapping); split where
// the mapping's separately registered pieces meet.
cudaError_t CopyFromPinnedSource(pd_ctx_impl* c, void* dst, uint64_t off, uint64_t len, cudaStream_t s);
int CheckDeviceError(pd_ctx_impl* c);  // reads the device error word after a synchronize
int* DevErr(pd_ctx_impl* c);
uint32_t* DevWorkCounter(pd_ctx_impl* c);
bool IsPinnedHost(const void* p);  // api_sample.cu
uint32_t* DevSparseWords(pd_ctx_impl* c);  // [0] dense row count of the current batch, [1..3] base planes
uint32_t ExtendToBase(const pd_ctx_impl* c, uint32_t v);
int BuildSubset(pd_ctx_impl* c, const uint64_t* include, const SubsetState* and_with, SubsetState* out);

// One batch of dense 2-bit rows handed to a sample-wise consumer.
struct RowBatch {
  uint32_t row0;           // index of the batch's first row in the operation's row space
  uint32_t n;              // rows in this batch
  uint32_t v_first;        // variant of row 0 when the batch is a contiguous range
  const uint32_t* d_vlist; // device list of the batch's variants (list mode), else nullptr
  pdk::DenseRows rows;
};
using RowFn = std::function<int(const RowBatch&, const pdk::RecIndex&, cudaStream_t)>;

// Streams [v_begin, v_end) through the staging pipeline (or the resident copy) and hands it to fn
// in batches of at most max_rows dense rows.  min_tile_bytes raises the staging tile size for
// consumers that amortise per-launch work over many rows.
int RunRows(pd_ctx_impl* c, uint32_t v_begin, uint32_t v_end, uint32_t max_rows, uint64_t min_tile_bytes,
            const RowFn& fn);
// Same for an ascending list of distinct variants (host array); rows follow the list order.
int RunRowsList(pd_ctx_impl* c, const uint32_t* vlist, uint32_t n, uint32_t max_rows, const RowFn& fn);
uint32_t DenseSlotStride(const pd_ctx_impl* c);
int DownloadRows(pd_ctx_impl* c, const void* dev, size_t bytes, void* host);
int LoadRange(pd_ctx_impl* c, uint32_t v_begin, uint32_t v_end, DevBuf* buf);
int StartScanJob(pd_ctx* c, uint32_t v_begin, uint32_t v_end, bool hardy, uint32_t midp, pd_job** out);

}  // namespace pdx

struct pd_ctx : pdx::pd_ctx_impl {};

// Asynchronous counts job (pd_counts_start / pd_counts_fetch).
struct pd_job {
  pd_ctx* ctx = nullptr;
  uint32_t v_begin = 0, v_end = 0;
  std::vector<uint32_t> counts;  // 4 x rows
  std::vector<double> pvals;     // rows (hardy jobs)
  bool hardy = false;            // also runs the exact test per tile
  uint32_t midp = 0;
  std::mutex mu;
  std::condition_variable cv;
  uint32_t rows_ready = 0;       // rows [0, rows_ready) of counts are final
  int status = PD_OK;
  bool finished = false;
  bool started = false;          // the worker holds the context
  std::thread worker;
  std::vector<pd_job*> subs;     // multi-GPU context: one job per shard with a non-empty range
  pdx::DecodeJob* decode = nullptr;  // pd_decode_start jobs
};
