// K5t: genotype-orient ("tidy") expansion -- one output row per (variant, sample) call.
//
// Replaces the per-thread double loop of PfileTidyScan (src/pfile_reader.cpp:2960-3281: load_variant = PgrGet +
// GenoarrToBytesMinus9, then emit_sample_row per sample with the genotype_range row filter of :3043-3075) for the
// three columns that come from the genotype path: which variant, which sample (file order index, so the caller can
// attach the .psam columns), and the call.  Rows of a variant are written in ascending sample order at the
// variant's row offset (exclusive prefix sum of the rows it emits, known from its genotype counts), so the output
// is exactly the reference's row sequence for a single scan thread.
#include "device_common.cuh"
#include "kernels.cuh"

namespace pdk {

namespace {

constexpr int kTidyThreads = 256;

// blockIdx.x = row of the batch.  A thread takes 16 consecutive output positions per step.
__global__ void __launch_bounds__(kTidyThreads)
k_tidy(DenseRows rows, const uint32_t* __restrict__ subset_idx, uint32_t n_out, uint32_t emit_mask,
       const unsigned long long* __restrict__ row_off, unsigned long long off0, uint32_t variant0,
       const uint32_t* __restrict__ vlist, uint32_t* __restrict__ out_variant, uint32_t* __restrict__ out_sample,
       int8_t* __restrict__ out_geno) {
  __shared__ uint32_t warp_tot[kTidyThreads / 32];
  __shared__ unsigned long long s_base;
  const uint32_t r = blockIdx.x;
  const uint8_t* row = rows.ptrs ? rows.ptrs[r] : rows.base + static_cast<uint64_t>(r) * rows.stride;
  const uint32_t variant = vlist ? vlist[r] : variant0 + r;
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_base = row_off[r] - off0;
  __syncthreads();
  for (uint32_t j0 = 0; j0 < n_out; j0 += kTidyThreads * 16) {
    const uint32_t j = j0 + threadIdx.x * 16;
    // codes of my 16 positions, 2 bits each
    uint32_t w = 0;
    uint32_t cnt = 0;
    if (j < n_out) {
      cnt = min(16u, n_out - j);
      if (!subset_idx) {
        w = load_bits32(row, static_cast<uint64_t>(j) * 2);
      } else {
        for (uint32_t k = 0; k < cnt; ++k) w |= geno_at(row, subset_idx[j + k]) << (2 * k);
      }
    }
    // positions that emit a row
    uint32_t emit = 0;
    for (uint32_t k = 0; k < cnt; ++k) emit |= ((emit_mask >> ((w >> (2 * k)) & 3u)) & 1u) << k;
    const uint32_t mine = __popc(emit);
    uint32_t incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += t;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int k = 0; k < kTidyThreads / 32; ++k) {
      const uint32_t c = warp_tot[k];
      before += k < static_cast<int>(wid) ? c : 0u;
      total += c;
    }
    unsigned long long o = s_base + before + (incl - mine);
    while (emit) {
      const uint32_t k = __ffs(emit) - 1u;
      emit &= emit - 1u;
      const uint32_t code = (w >> (2 * k)) & 3u;
      out_variant[o] = variant;
      out_sample[o] = subset_idx ? subset_idx[j + k] : j + k;
      out_geno[o] = code == 3u ? static_cast<int8_t>(-9) : static_cast<int8_t>(code);
      ++o;
    }
    __syncthreads();
    if (threadIdx.x == 0) s_base += total;
    __syncthreads();
  }
}

// rows emitted per variant from its genotype counts, as an inclusive running sum seeded by the previous batch
__global__ void k_tidy_rows(const uint4* __restrict__ counts, uint32_t n, uint32_t emit_mask,
                            unsigned long long* __restrict__ rows_per) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 c = counts[i];
  rows_per[i] = ((emit_mask & 1u) ? c.x : 0u) + ((emit_mask & 2u) ? c.y : 0u) + static_cast<unsigned long long>((emit_mask & 4u) ? c.z : 0u) +
                ((emit_mask & 8u) ? c.w : 0u);
}

}  // namespace

void launch_tidy_rows(const uint4* counts, uint32_t n, uint32_t emit_mask, unsigned long long* rows_per, cudaStream_t s,
                      LaunchStats* st) {
  if (!n) return;
  k_tidy_rows<<<(n + 255) / 256, 256, 0, s>>>(counts, n, emit_mask, rows_per);
  if (st) st->launches++;
}

void launch_tidy(const DenseRows& rows, const uint32_t* subset_idx, uint32_t n_out, uint32_t emit_mask,
                 const unsigned long long* row_off, unsigned long long off0, uint32_t variant0, const uint32_t* vlist,
                 uint32_t* out_variant, uint32_t* out_sample, int8_t* out_geno, cudaStream_t s, LaunchStats* st) {
  if (!rows.nrows || !n_out) return;
  k_tidy<<<rows.nrows, kTidyThreads, 0, s>>>(rows, subset_idx, n_out, emit_mask, row_off, off0, variant0, vlist, out_variant,
                                             out_sample, out_geno);
  if (st) st->launches++;
}

}  // namespace pdk
