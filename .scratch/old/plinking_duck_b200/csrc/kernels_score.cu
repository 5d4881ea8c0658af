// K7: plink_score -- per-sample sum over scored variants of weight x (imputed / flipped / centred)
// ALT dosage, for K weight columns at once.  Replaces the Phase-1 loop of
// src/plink_score.cpp:582-653 (PgrGetD + Dosage16ToDoublesMinus9 + two passes over double[N] per
// variant) without ever materialising the doubles: a hardcall contributes one of four per-variant
// table values T[g], explicit dosages are evaluated from their u16 in place.
//
// fp64 CUDA-core path (relative tolerance 1e-12 against the sequential fp64 oracle).  Work layout:
// a thread owns one sample and KT weight columns (KT accumulators in registers) and walks the rows
// of the batch; per-row tables and weights are staged in shared memory in chunks of kRowChunk
// rows and read as warp-wide broadcasts.
#include "device_common.cuh"
#include "kernels.cuh"

namespace pdk {

namespace {

constexpr int kThreads = 256;
constexpr int kRowChunk = 64;

__device__ __forceinline__ const uint8_t* row_ptr(const DenseRows& rows, uint32_t r) {
  return rows.ptrs ? rows.ptrs[r] : rows.base + static_cast<uint64_t>(r) * rows.stride;
}

__device__ __forceinline__ uint32_t ld_u16(const uint8_t* p, uint32_t i) {
  return static_cast<uint32_t>(p[2ull * i]) | (static_cast<uint32_t>(p[2ull * i + 1]) << 8);
}

// Contribution of an ALT dosage `alt` under the row's mode (src/plink_score.cpp:612-652).
__device__ __forceinline__ double score_value(const ScoreRow& R, double alt) {
  const double scored = (R.flags & kScoreFlip) ? (2.0 - alt) : alt;
  if (R.flags & kScoreCenter) return (scored - R.mean_scored) / R.sd;
  return scored;
}

// ---------------------------------------------------------------------------------------------
// k_score_prep: one thread per row.  From the (subset) hardcall counts and the dosage sums:
// mean ALT dosage, skip rules, and the table of the four hardcall contributions.
// ---------------------------------------------------------------------------------------------
__global__ void k_score_prep(const uint4* __restrict__ counts, const DosageStats* __restrict__ stats,
                             const uint8_t* __restrict__ flip, uint32_t nrows, uint32_t mode,
                             ScoreRow* __restrict__ out) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nrows) return;
  const uint4 c = counts[r];
  unsigned long long sum16 = 16384ull * c.y + 32768ull * c.z;
  uint32_t nm = c.x + c.y + c.z;
  uint32_t kind = 0;
  if (stats) {
    const DosageStats s = stats[r];
    kind = s.kind;
    if (kind) {
      sum16 = s.sum + 16384ull * (c.y - s.replaced[1]) + 32768ull * (c.z - s.replaced[2]);
      nm = s.ct + nm - (s.replaced[0] + s.replaced[1] + s.replaced[2]);
    }
  }
  ScoreRow R;
  R.flags = (flip && flip[r]) ? kScoreFlip : 0u;
  R.kind = kind;
  R.cmask = 0;
  R.mean_scored = 0.0;
  R.sd = 1.0;
#pragma unroll
  for (int g = 0; g < 4; ++g) R.T[g] = 0.0;
  if (nm) {
    const double sum_alt = static_cast<double>(sum16) * 0.00006103515625;
    const double mean_alt = sum_alt / static_cast<double>(nm);
    bool active = true;
    if (mode == 2) {
      const double freq = mean_alt / 2.0;
      const double sd = sqrt(2.0 * freq * (1.0 - freq));
      if (sd == 0.0) {
        active = false;
      } else {
        R.flags |= kScoreCenter;
        R.sd = sd;
        R.mean_scored = (R.flags & kScoreFlip) ? (2.0 - mean_alt) : mean_alt;
      }
    }
    if (active) {
      R.flags |= kScoreActive;
      R.cmask = (mode == 0) ? 0xFu : 0x7u;
      if (mode == 0) R.flags |= kScoreImpute;
      R.T[0] = score_value(R, 0.0);
      R.T[1] = score_value(R, 1.0);
      R.T[2] = score_value(R, 2.0);
      R.T[3] = (mode == 0) ? score_value(R, mean_alt) : 0.0;
    }
  }
  out[r] = R;
}

// ---------------------------------------------------------------------------------------------
// k_score_acc: grid (sample blocks, row chunks, K tiles); a thread owns kSpt = 4 consecutive samples
// (exactly one byte of every packed row) and KT weight columns: 4 x KT fp64 accumulators in
// registers.  Per row the fast path is 1 byte load, 4 table look-ups, KT/2 128-bit weight
// broadcasts and 4 x KT DFMA -- no branches: values that do not count are 0 in the table.
// Rows with dense dosage storage, and sample subsets, take the general per-sample path.
// Partial sums are added to the global accumulators with fp64 atomics once per launch.
// ---------------------------------------------------------------------------------------------
constexpr int kSpt = 4;
constexpr int kAccThreads = 128;
constexpr int kRowGroup = 8;

template <int KT>
__global__ void __launch_bounds__(kAccThreads)
k_score_acc(DenseRows rows, const ScoreRow* __restrict__ srow, const DosageLoc* __restrict__ loc,
            const uint32_t* __restrict__ rank, uint32_t rank_stride, const double* __restrict__ weights, uint32_t K,
            const uint32_t* __restrict__ subset_idx, uint32_t n_out, uint32_t mode, uint32_t rows_per_chunk,
            double* __restrict__ score, double* __restrict__ dosage_sums, uint32_t* __restrict__ allele_cts) {
  __shared__ ScoreRow s_row[kRowChunk];
  __shared__ __align__(16) double s_w[kRowChunk][KT];
  __shared__ const uint8_t* s_ptr[kRowChunk];
  __shared__ DosageLoc s_loc[kRowChunk];
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t j0 = t * kSpt;
  const uint32_t k0 = blockIdx.z * KT;
  const bool owner = blockIdx.z == 0;
  const bool live = j0 < n_out;
  const uint32_t row_lo = blockIdx.y * rows_per_chunk;
  const uint32_t row_hi = min(rows.nrows, row_lo + rows_per_chunk);
  double acc[kSpt][KT];
#pragma unroll
  for (int i = 0; i < kSpt; ++i) {
#pragma unroll
    for (int k = 0; k < KT; ++k) acc[i][k] = 0.0;
  }
  double dsum[kSpt] = {0.0, 0.0, 0.0, 0.0};
  uint32_t act[kSpt] = {0u, 0u, 0u, 0u};
  for (uint32_t rc = row_lo; rc < row_hi; rc += kRowChunk) {
    const uint32_t nr = min(static_cast<uint32_t>(kRowChunk), row_hi - rc);
    __syncthreads();
    for (uint32_t x = threadIdx.x; x < nr * KT; x += blockDim.x) {
      const uint32_t r = x / KT, k = x % KT;
      s_w[r][k] = (k0 + k < K) ? weights[static_cast<uint64_t>(rc + r) * K + k0 + k] : 0.0;
    }
    for (uint32_t x = threadIdx.x; x < nr; x += blockDim.x) {
      s_row[x] = srow[rc + x];
      s_ptr[x] = row_ptr(rows, rc + x);
      if (loc) s_loc[x] = loc[rc + x];
    }
    __syncthreads();
    if (!live) continue;
    for (uint32_t rg = 0; rg < nr; rg += kRowGroup) {
      // the genotype bytes of kRowGroup rows are requested together: one dependent load per row
      // would leave every warp waiting a full DRAM round trip per row
      uint32_t gb[kRowGroup];
#pragma unroll
      for (int q = 0; q < kRowGroup; ++q) {
        gb[q] = 0;
        if (rg + q < nr && !subset_idx && s_row[rg + q].kind < 2 && (s_row[rg + q].flags & kScoreActive)) {
          gb[q] = __ldg(s_ptr[rg + q] + t);
        }
      }
#pragma unroll
      for (int q = 0; q < kRowGroup; ++q) {
        const uint32_t r = rg + q;
        if (r >= nr) break;
        const ScoreRow& R = s_row[r];
        if (!(R.flags & kScoreActive)) continue;  // block-uniform
        double val[kSpt];
        uint32_t cnt[kSpt];
        if (!subset_idx && R.kind < 2) {
          const uint32_t b = gb[q];
#pragma unroll
          for (int i = 0; i < kSpt; ++i) {
            const uint32_t g = (b >> (2 * i)) & 3u;
            val[i] = R.T[g];
            cnt[i] = (R.cmask >> g) & 1u;
          }
        } else {
#pragma unroll
          for (int i = 0; i < kSpt; ++i) {
            val[i] = 0.0;
            cnt[i] = 0;
            const uint32_t j = j0 + i;
            if (j >= n_out) continue;
            const uint32_t raw = subset_idx ? subset_idx[j] : j;
            const uint32_t g = geno_at(s_ptr[r], raw);
            val[i] = R.T[g];
            cnt[i] = (R.cmask >> g) & 1u;
            if (R.kind == 2) {
              const uint32_t u = ld_u16(s_loc[r].values, raw);
              if (u != 65535u) {
                val[i] = score_value(R, static_cast<double>(u) * 0.00006103515625);
                cnt[i] = 1;
              }
            } else if (R.kind == 3) {
              const uint32_t bits = load_bits32(s_loc[r].track3, static_cast<uint64_t>(raw >> 5) << 5);
              if ((bits >> (raw & 31)) & 1u) {
                const uint32_t idx = rank[static_cast<uint64_t>(rc + r) * rank_stride + (raw >> 5)] +
                                     __popc(bits & ((1u << (raw & 31)) - 1u));
                val[i] = score_value(R, static_cast<double>(ld_u16(s_loc[r].values, idx)) * 0.00006103515625);
                cnt[i] = 1;
              }
            }
          }
        }
#pragma unroll
        for (int k = 0; k < KT; ++k) {
          const double w = s_w[r][k];
#pragma unroll
          for (int i = 0; i < kSpt; ++i) acc[i][k] = fma(w, val[i], acc[i][k]);
        }
        if (owner) {
#pragma unroll
          for (int i = 0; i < kSpt; ++i) {
            act[i] += cnt[i];
            dsum[i] += val[i];
          }
        }
      }
    }
  }
  if (!live) return;
#pragma unroll
  for (int i = 0; i < kSpt; ++i) {
    const uint32_t j = j0 + i;
    if (j >= n_out) break;
#pragma unroll
    for (int k = 0; k < KT; ++k) {
      if (k0 + k < K && acc[i][k] != 0.0) atomicAdd(score + static_cast<uint64_t>(j) * K + k0 + k, acc[i][k]);
    }
    if (owner) {
      if (mode != 2 && dsum[i] != 0.0) atomicAdd(dosage_sums + j, dsum[i]);
      if (act[i]) atomicAdd(allele_cts + j, 2u * act[i]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_score_list_fix: rows whose dosages are stored as a sample-id list (sparse by construction,
// <= N/8 entries).  The accumulate pass used the hardcall value for those samples; one warp per
// row walks the list and adds (dosage value - hardcall value) x weight with fp64 atomics.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
k_score_list_fix(DenseRows rows, const ScoreRow* __restrict__ srow, const DosageLoc* __restrict__ loc,
                 const double* __restrict__ weights, uint32_t K, SubsetDev sub, uint32_t mode,
                 double* __restrict__ score, double* __restrict__ dosage_sums, uint32_t* __restrict__ allele_cts) {
  const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= rows.nrows) return;
  const ScoreRow R = srow[r];
  const DosageLoc L = loc[r];
  if (R.kind != 1 || !L.raw_ct || !(R.flags & kScoreActive)) return;
  const uint8_t* row = row_ptr(rows, r);
  const uint8_t* q = L.track3;
  Difflist d;
  if (!parse_difflist(q, L.values, rows.N, false, d)) return;
  const double* w = weights + static_cast<uint64_t>(r) * K;
  difflist_walk(d, rows.N, lane, [&](uint32_t id, uint32_t, uint32_t k) {
    uint32_t pos = id;
    if (sub.include_words) {
      const uint32_t word = sub.include_words[id >> 5];
      if (!((word >> (id & 31)) & 1u)) return;
      pos = sub.cum_popcount[id >> 5] + __popc(word & ((1u << (id & 31)) - 1u));
    }
    const uint32_t g = geno_at(row, id);
    const bool was_counted = (g != 3) || (R.flags & kScoreImpute);
    const double val = score_value(R, static_cast<double>(ld_u16(L.values, k)) * 0.00006103515625);
    const double delta = val - (was_counted ? R.T[g] : 0.0);
    if (delta != 0.0) {
      for (uint32_t c = 0; c < K; ++c) atomicAdd(score + static_cast<uint64_t>(pos) * K + c, w[c] * delta);
      if (mode != 2) atomicAdd(dosage_sums + pos, delta);
    }
    if (!was_counted) atomicAdd(allele_cts + pos, 2u);
  });
}

}  // namespace

void launch_score_prep(const uint4* counts, const DosageStats* stats, const uint8_t* flip, uint32_t nrows, uint32_t mode,
                       ScoreRow* out, cudaStream_t s, LaunchStats* st) {
  if (!nrows) return;
  k_score_prep<<<(nrows + 127) / 128, 128, 0, s>>>(counts, stats, flip, nrows, mode, out);
  if (st) st->launches++;
}

void launch_score_acc(const DenseRows& rows, const ScoreRow* srow, const DosageLoc* loc, const uint32_t* rank,
                      uint32_t rank_stride, const double* weights, uint32_t K, const SubsetDev& sub, uint32_t n_out,
                      uint32_t mode, double* score, double* dosage_sums, uint32_t* allele_cts, cudaStream_t s,
                      LaunchStats* st) {
  if (!rows.nrows || !n_out || !K) return;
  const uint32_t bx = (n_out + kAccThreads * kSpt - 1) / (kAccThreads * kSpt);
  const uint32_t kt = K <= 4 ? 4 : 8;
  const uint32_t bz = (K + kt - 1) / kt;
  // split the rows over grid.y until the chip has ~6 CTAs per SM (results are merged with atomics)
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  uint32_t by = (static_cast<uint32_t>(sms) * 6 + bx * bz - 1) / (bx * bz);
  const uint32_t max_by = (rows.nrows + kRowChunk - 1) / kRowChunk;
  if (by > max_by) by = max_by;
  if (by < 1) by = 1;
  uint32_t rpc = (rows.nrows + by - 1) / by;
  rpc = (rpc + kRowChunk - 1) / kRowChunk * kRowChunk;
  by = (rows.nrows + rpc - 1) / rpc;
  dim3 grid(bx, by, bz);
  if (kt == 4) {
    k_score_acc<4><<<grid, kAccThreads, 0, s>>>(rows, srow, loc, rank, rank_stride, weights, K, sub.subset_idx, n_out,
                                               mode, rpc, score, dosage_sums, allele_cts);
  } else {
    k_score_acc<8><<<grid, kAccThreads, 0, s>>>(rows, srow, loc, rank, rank_stride, weights, K, sub.subset_idx, n_out,
                                               mode, rpc, score, dosage_sums, allele_cts);
  }
  if (st) st->launches++;
  if (loc) {
    const uint32_t blocks = (rows.nrows * 32 + kThreads - 1) / kThreads;
    k_score_list_fix<<<blocks, kThreads, 0, s>>>(rows, srow, loc, weights, K, sub, mode, score, dosage_sums, allele_cts);
    if (st) st->launches++;
  }
}

}  // namespace pdk
