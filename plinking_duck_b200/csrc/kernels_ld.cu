// Consumers of the decoded rows next to the counting path:
//   k_ld_pairs     -- the six integer sums behind plink_ld's genotype-covariance estimator, one warp per variant
//                     pair, as popcount contractions of the two 2-bit rows (ComputeLdStats, src/plink_ld.cpp:52-86);
//   k_standardize  -- (g - 2 f) / sqrt(2 f (1 - f)) with missing calls at 0, the rows plink_pca / plink_glm feed to
//                     their dense products (NormalizeGenotypes, src/plink_common.cpp:1535-1543).
#include <algorithm>

#include "device_common.cuh"
#include "kernels.cuh"

namespace pdk {

namespace {

constexpr uint32_t kEven = 0x55555555u;

// 16 include bits -> the even bits of a 32-bit word
__device__ __forceinline__ uint32_t spread_even(uint32_t x) {
  x = (x | (x << 8)) & 0x00FF00FFu;
  x = (x | (x << 4)) & 0x0F0F0F0Fu;
  x = (x | (x << 2)) & 0x33333333u;
  x = (x | (x << 1)) & 0x55555555u;
  return x;
}

__global__ void __launch_bounds__(256)
k_ld_pairs(DenseRows rows, const uint32_t* __restrict__ ra, const uint32_t* __restrict__ rb, uint32_t n_pairs,
           const uint32_t* __restrict__ include_words, uint32_t* __restrict__ out) {
  const uint32_t pair = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31u;
  if (pair >= n_pairs) return;
  const uint8_t* a = rows.ptrs ? rows.ptrs[ra[pair]] : rows.base + static_cast<uint64_t>(ra[pair]) * rows.stride;
  const uint8_t* b = rows.ptrs ? rows.ptrs[rb[pair]] : rows.base + static_cast<uint64_t>(rb[pair]) * rows.stride;
  const uint32_t N = rows.N;
  const uint32_t nwords = (N + 15) >> 4;
  uint32_t n = 0, sa1 = 0, sa2 = 0, sb1 = 0, sb2 = 0, p11 = 0, p12 = 0, p21 = 0, p22 = 0;
  for (uint32_t k = lane; k < nwords; k += 32) {
    const uint32_t wa = row_word(a, k), wb = row_word(b, k);
    const uint32_t al = wa & kEven, ah = (wa >> 1) & kEven, bl = wb & kEven, bh = (wb >> 1) & kEven;
    uint32_t v = ~(al & ah) & ~(bl & bh) & kEven;            // both calls present
    if (16 * k + 16 > N) v &= (1u << (2 * (N - 16 * k))) - 1u;  // ragged last word
    if (include_words) v &= spread_even((include_words[k >> 1] >> (16 * (k & 1))) & 0xffffu);
    n += __popc(v);
    const uint32_t a1 = al & v, a2 = ah & v, b1 = bl & v, b2 = bh & v;  // code 1 / code 2 of each side
    sa1 += __popc(a1);
    sa2 += __popc(a2);
    sb1 += __popc(b1);
    sb2 += __popc(b2);
    p11 += __popc(a1 & b1);
    p12 += __popc(a1 & b2);
    p21 += __popc(a2 & b1);
    p22 += __popc(a2 & b2);
  }
  n = group_sum<32>(n);
  sa1 = group_sum<32>(sa1);
  sa2 = group_sum<32>(sa2);
  sb1 = group_sum<32>(sb1);
  sb2 = group_sum<32>(sb2);
  p11 = group_sum<32>(p11);
  p12 = group_sum<32>(p12);
  p21 = group_sum<32>(p21);
  p22 = group_sum<32>(p22);
  if (lane == 0) {
    uint32_t* o = out + 6ull * pair;
    o[0] = n;
    o[1] = sa1 + 2 * sa2;                    // sum gA
    o[2] = sb1 + 2 * sb2;                    // sum gB
    o[3] = p11 + 2 * (p12 + p21) + 4 * p22;  // sum gA gB
    o[4] = sa1 + 4 * sa2;                    // sum gA^2
    o[5] = sb1 + 4 * sb2;                    // sum gB^2
  }
}

// blockIdx.y = row; thread = 16 output doubles per step.
__global__ void __launch_bounds__(256)
k_standardize(DenseRows rows, const uint32_t* __restrict__ subset_idx, uint32_t n_out, const double2* __restrict__ norm,
              double* __restrict__ out) {
  const uint32_t r = blockIdx.y;
  const uint8_t* row = rows.ptrs ? rows.ptrs[r] : rows.base + static_cast<uint64_t>(r) * rows.stride;
  const double center = norm[r].x, inv = norm[r].y;
  // the four possible values of the row (missing -> 0: mean imputation of a centred variable)
  const double t0 = (0.0 - center) * inv, t1 = (1.0 - center) * inv, t2 = (2.0 - center) * inv;
  double* o = out + static_cast<uint64_t>(r) * n_out;
  for (uint32_t j0 = (blockIdx.x * blockDim.x + threadIdx.x) * 16; j0 < n_out; j0 += gridDim.x * blockDim.x * 16) {
    const uint32_t cnt = min(16u, n_out - j0);
    uint32_t w = 0;
    if (!subset_idx) {
      w = load_bits32(row, static_cast<uint64_t>(j0) * 2);
    } else {
      for (uint32_t k = 0; k < cnt; ++k) w |= geno_at(row, subset_idx[j0 + k]) << (2 * k);
    }
    for (uint32_t k = 0; k < cnt; ++k) {
      const uint32_t code = (w >> (2 * k)) & 3u;
      o[j0 + k] = code == 0 ? t0 : (code == 1 ? t1 : (code == 2 ? t2 : 0.0));
    }
  }
}

}  // namespace

void launch_ld_pairs(const DenseRows& rows, const uint32_t* ra, const uint32_t* rb, uint32_t n_pairs,
                     const uint32_t* include_words, uint32_t* out, cudaStream_t s, LaunchStats* st) {
  if (!n_pairs) return;
  const uint64_t threads = static_cast<uint64_t>(n_pairs) * 32;
  k_ld_pairs<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, s>>>(rows, ra, rb, n_pairs, include_words, out);
  if (st) st->launches++;
}

void launch_standardize(const DenseRows& rows, const uint32_t* subset_idx, uint32_t n_out, const double2* norm, double* out,
                        cudaStream_t s, LaunchStats* st) {
  if (!rows.nrows || !n_out) return;
  const uint32_t bx = static_cast<uint32_t>(std::min<uint64_t>(64, (static_cast<uint64_t>(n_out) + 4095) / 4096));
  k_standardize<<<dim3(bx, rows.nrows), 256, 0, s>>>(rows, subset_idx, n_out, norm, out);
  if (st) st->launches++;
}

}  // namespace pdk
