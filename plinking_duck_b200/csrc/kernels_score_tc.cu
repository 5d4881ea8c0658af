// K7t: plink_score on the 5th-generation tensor cores (tcgen05, int8, accumulators in TMEM).
//
// The accumulation loop of src/plink_score.cpp:598-653 is, for hardcall records, a dense contraction
//     score[s, c] = sum_v  w[v, c] * T_v[g(v, s)]
// with T_v the per-variant table of the four hardcall contributions (k_score_prep).  T_v is affine
// in g for g in {0, 1, 2}, so with M(v, s) = [g == missing] and G(v, s) = g (0 where missing)
//     w T_v[g] = w (T1 - T0) * G  +  w (T3 - T0) * M  +  w T0 .
// That is one GEMM over "k-rows" (a G row per active variant, an M row per variant that has missing
// calls) between small-integer matrices G, M and real weights -- and the real weights are made
// exact for the tensor cores by fixed-point limbs (the int8 "Ozaki" splitting):
//     x ~ scale_c * q,   q = round(x / scale_c) a signed 55-bit integer = sum_l limb_l 256^l,
// seven signed int8 limbs per weight.  int8 x int8 products accumulate EXACTLY in the int32 TMEM
// accumulators (|sum| <= 2 * 128 * k-rows < 2^31 for up to 8M k-rows), the limbs are recombined as
// hi * 2^24 + lo, and a single rounding to fp64 happens per batch (<= 65535 scored variants) and column;
// batches are added in fp64.  Within a batch the result is independent of summation order
// (bit-reproducible across sample tiles, k splits and launches); its error is the quantisation
// max|x| 2^-55 per term, i.e. tighter than the fp64 FMA loop it replaces.
//
// GEMM shape per CTA:  two accumulators D_h[128 x 256] (int32, all 512 TMEM columns) += A[128 x 32] * B_h[32 x 256]
//   A (M side) = weight limbs: row r = column c * 7 + limb l, K-major, prepared once per batch in
//                global memory in the canonical no-swizzle shared-memory order (one 4 KB bulk copy);
//   B (N side) = 2 x 256 samples x 32 k-rows of unpacked genotypes, N-major (samples contiguous -- the
//                order the .pgen record already has), written to shared memory by 256 unpack threads
//                from the packed 2-bit rows (PRMT table look-up on nibble-spread codes).
// Two tcgen05.mma (M=128, N=256, K=32, kind::i8) per k-slab, issued by one thread; the second takes A from the
// tensor core's collector buffer (collector::a::fill / lastuse).  Stages are handed over with mbarriers
// (producers -> MMA) and tcgen05.commit (MMA -> producers).
//
// What bounds it is the shared-memory data pipe: one 128-byte wavefront per cycle shared by the LSU (st.shared of
// the unpacked bytes, ld.shared / cp.async of the packed ones) and the tensor core's operand reads -- ncu: LSU +
// tensor wavefronts = elapsed cycles.  Per MMA: 64 wavefronts of B written + 64 read are inherent to int8 operands
// in shared memory; everything else is overhead this layout keeps small: the 512-sample tile halves the A reads
// per MMA (32 -> 16) and the cache lines touched per staged byte (a row window is 144 B instead of 80 B: 2.1
// instead of 1.6 lines per row, for twice the samples), and the converters read the staged rows with
// conflict-free 128-bit loads.
//
// The unit of producer work is a STEP = 32 variants.  Variants with missing calls come first in the batch's
// k-row list ("pair" steps): their packed bytes are fetched and aligned ONCE and unpacked twice, into the G slab
// and the M slab of the same 32 variants (two MMAs per step); variants without missing calls follow ("single"
// steps, G slab only).  The order of k-rows inside a batch is free: the int32 sums are exact.
#include <algorithm>

#include "device_common.cuh"
#include "kernels.cuh"

namespace pdk {

namespace {

constexpr uint32_t kMmaN = 256;     // samples per MMA (N)
constexpr uint32_t kHalves = 2;     // MMAs per k-slab: the CTA's tile is two N = 256 halves that share the A slab
constexpr uint32_t kTileN = kMmaN * kHalves;  // samples per CTA tile
constexpr uint32_t kSlabK = 32;     // k-rows per MMA (K of kind::i8)
constexpr uint32_t kRowsM = 128;    // MMA M: weight-limb rows
constexpr uint32_t kLimbs = 7;
constexpr uint32_t kStages = 6;     // B stages (one k-slab of both halves, 16 KB each) between the unpackers and the MMA
constexpr uint32_t kAStages = 2 * kStages;  // A slabs (4 KB bulk copies from L2): their latency under load is ~2 us, so they run further ahead than the B slabs
constexpr uint32_t kBHalfBytes = kMmaN * kSlabK;      // 8 KB
constexpr uint32_t kBBytes = kHalves * kBHalfBytes;   // 16 KB
constexpr uint32_t kABytes = kRowsM * kSlabK;  // 4 KB
constexpr uint32_t kProducers = 256;  // threads that unpack one step: 32 rows x 8 chunks of 64 samples
constexpr uint32_t kGroups = 2;       // producer groups; group g unpacks steps g, g + kGroups, ... of the CTA
constexpr uint32_t kProdWarps = kGroups * kProducers / 32;
constexpr uint32_t kMmaThreads = kGroups * kProducers + 64 + 32 * kGroups;  // + MMA issue / TMEM warp, A-copy warp, one packed-row loader warp per group
constexpr uint32_t kPStages = 6;    // steps of packed rows in flight between the loader warps and the converters
constexpr uint32_t kRowWin = kTileN / 4;  // 128 wanted bytes of a row and step
constexpr uint32_t kPRow = kRowWin + 16;  // staged per row: the wanted bytes at any 16-byte misalignment; 9 chunks -- an ODD
                                          // count, so a quarter warp's 128-bit loads (8 rows, one chunk) never share a bank
constexpr uint32_t kPChunks = kPRow / 16;
constexpr uint32_t kPlaneM = 0x80000000u;

// ---- PTX wrappers ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void mbar_wait_addr(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_addr(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// 16-byte asynchronous copy global -> shared (LDGSTS); src_bytes = 0 zero-fills without reading.
__device__ __forceinline__ void cp_async16(void* dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(src_bytes) : "memory");
}
// Arrives on the mbarrier (without raising its pending count) once all cp.async of this thread so far are done.
__device__ __forceinline__ void cp_async_mbar_arrive(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], int8 operands, int32 accumulators.
__device__ __forceinline__ void umma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                        uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// The same with the A operand kept in the tensor core's collector buffer for the next MMA (fill) / taken from it
// for the last time (lastuse): the two halves of a tile multiply the same A slab, so it leaves shared memory once.
__device__ __forceinline__ void umma_i8_a_fill(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8.collector::a::fill [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_i8_a_lastuse(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                                  uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8.collector::a::lastuse [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrives on the mbarrier once every tcgen05.mma issued so far by this thread has completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void umma_commit_addr(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor, no swizzle (cute::UMMA::SmemDescriptor, mma_sm100_desc.hpp):
// bits [0,14) address >> 4, [16,30) leading byte offset >> 4, [32,46) stride byte offset >> 4,
// [46,48) version = 1 (sm_100), [61,64) layout type = 0.
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return static_cast<uint64_t>((addr & 0x3ffffu) >> 4) | (static_cast<uint64_t>(lbo_bytes >> 4) << 16) |
         (static_cast<uint64_t>(sbo_bytes >> 4) << 32) | (1ull << 46);
}

// Instruction descriptor (cute::UMMA::InstrDescriptor): c = S32 (2 << 4), a = b = signed int8
// (1 << 7, 1 << 10), A K-major (bit 15 = 0), B MN-major (bit 16 = 1), N >> 3 at [17,23), M >> 4 at [24,29).
constexpr uint32_t kIdesc = (2u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | ((kMmaN >> 3) << 17) | ((kRowsM >> 4) << 24);

// Byte offset of element (row r, k) of an A slab: canonical K-major no-swizzle order -- core matrix =
// 8 rows x 16 bytes of K (row stride 16 B); the two K halves 128 B apart (LBO), 8-row groups 256 B apart (SBO).
__device__ __forceinline__ uint32_t a_offset(uint32_t r, uint32_t k) {
  return (r & 7u) * 16u + (k >> 4) * 128u + (r >> 3) * 256u + (k & 15u);
}
constexpr uint32_t kALbo = 128, kASbo = 256;
// B slab (N-major, no swizzle): core matrix = 8 k-rows x 16 samples (k stride 16 B); 16-sample groups
// 128 B apart (SBO), 8-k groups 2048 B apart (LBO).
constexpr uint32_t kBSbo = 128, kBLbo = 2048;

// ---------------------------------------------------------------------------------------------
// k_tc_plan: (K + 2) x 32 + 1 CTAs per batch (32 row parts per column, one CTA for the k-row list).  Lists the k-rows (a G row per active variant, an M row when it
// has missing calls), and per column the fixed-point scale (a power of two >= max|x| 2^-54, so
// x / scale is exact up to the final rint) and the constant term sum_v w T0 (fixed reduction order).
// Columns: c < K score columns, c = K the dosage-sum column (weight 1), c = K + 1 counts the
// missing calls of a sample over the active rows (for ALLELE_CT when missing calls do not count).
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kPlanThreads = 1024;
constexpr uint32_t kMaxCols = 18;  // K + 2 <= 18  ->  7 * 18 = 126 limb rows <= 128

__device__ __forceinline__ double col_weight(const double* __restrict__ weights, uint32_t K, uint32_t mode, uint32_t r,
                                             uint32_t c) {
  if (c < K) return weights[static_cast<uint64_t>(r) * K + c];
  if (c == K) return mode == 2 ? 0.0 : 1.0;  // centred scores leave NAMED_ALLELE_DOSAGE_SUM alone
  return 0.0;
}

// Exclusive prefix sum of one value per thread over the block's 1024 threads (s_warp: 33 words); total = the sum.
__device__ __forceinline__ uint32_t plan_block_scan(uint32_t v, uint32_t* s_warp, uint32_t& total) {
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, inc, off);
    if (lane >= static_cast<uint32_t>(off)) inc += n;
  }
  if (lane == 31) s_warp[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    const uint32_t x = s_warp[lane];
    uint32_t xi = x;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t n = __shfl_up_sync(0xffffffffu, xi, off);
      if (lane >= static_cast<uint32_t>(off)) xi += n;
    }
    s_warp[lane] = xi - x;
    if (lane == 31) s_warp[32] = xi;
  }
  __syncthreads();
  total = s_warp[32];
  const uint32_t r = s_warp[wid] + inc - v;
  __syncthreads();
  return r;
}

constexpr uint32_t kPlanParts = 32;   // row parts per column in k_tc_plan
constexpr uint32_t kPlanChunks = 64;  // 1024-row chunks of a batch (<= 65 535 rows)

__global__ void __launch_bounds__(kPlanThreads)
k_tc_plan(const uint4* __restrict__ counts, const ScoreRow* __restrict__ srow, const double* __restrict__ weights,
          uint32_t nrows, uint32_t K, uint32_t mode, uint32_t* __restrict__ krow, double* __restrict__ col_scale,
          double* __restrict__ col_const, uint32_t* __restrict__ meta, double* __restrict__ part, uint32_t* __restrict__ part_ctr) {
  __shared__ uint32_t s_scan[40];
  __shared__ double s_red[64];
  __shared__ uint32_t s_cnt[2][kPlanChunks * 32];  // [list][chunk * 32 + warp]: entries of the warp in the chunk, then their offsets
  const uint32_t t = threadIdx.x;
  const uint32_t ncols = K + 2;
  if (blockIdx.x == ncols * kPlanParts) {
    // ---- last block: the k-row list: active variants with missing calls (pair steps), then those without, each
    // in row order.  Chunk i = rows [1024 i, 1024 i + 1024); every load is issued before the first barrier, the
    // offsets come from ONE scan over the (chunk, warp) counts. ----
    const uint32_t lane = t & 31u, wid = t >> 5;
    unsigned long long want0 = 0ull, want1 = 0ull;  // rows of this thread for the pair list / the single list
    for (uint32_t i0 = 0; i0 < kPlanChunks && i0 * kPlanThreads < nrows; i0 += 8) {
      // eight rows' loads in flight together (clamped, unconditional: behind a bounds branch they serialise)
      uint32_t fl[8], mis[8];
#pragma unroll
      for (uint32_t k = 0; k < 8; ++k) {
        const uint32_t r = min((i0 + k) * kPlanThreads + t, nrows - 1u);
        fl[k] = srow[r].flags;
        mis[k] = counts[r].w;
      }
#pragma unroll
      for (uint32_t k = 0; k < 8; ++k) {
        const uint32_t r = (i0 + k) * kPlanThreads + t;
        if (r < nrows && (fl[k] & kScoreActive)) {
          if (mis[k] != 0) {
            want0 |= 1ull << (i0 + k);
          } else {
            want1 |= 1ull << (i0 + k);
          }
        }
      }
    }
    for (uint32_t i = 0; i < kPlanChunks; ++i) {
#pragma unroll
      for (uint32_t pass = 0; pass < 2; ++pass) {
        const uint32_t bal = __ballot_sync(0xffffffffu, ((pass ? want1 : want0) >> i) & 1ull);
        if (lane == 0) s_cnt[pass][i * 32 + wid] = __popc(bal);
      }
    }
    __syncthreads();
    uint32_t n_real[2], n_steps[2], region[2];
    uint32_t region_base = 0;
#pragma unroll
    for (uint32_t pass = 0; pass < 2; ++pass) {
      const uint32_t a = s_cnt[pass][2 * t], b = s_cnt[pass][2 * t + 1];
      uint32_t total;
      const uint32_t ex = plan_block_scan(a + b, s_scan, total);
      s_cnt[pass][2 * t] = ex;
      s_cnt[pass][2 * t + 1] = ex + a;
      const uint32_t n_pad = (total + kSlabK - 1) / kSlabK * kSlabK;
      for (uint32_t i = total + t; i < n_pad; i += kPlanThreads) krow[region_base + i] = 0;  // zero weights; any valid row will do
      n_real[pass] = total;
      n_steps[pass] = n_pad / kSlabK;
      region[pass] = region_base;
      region_base += n_pad;
    }
    __syncthreads();
    for (uint32_t i = 0; i < kPlanChunks; ++i) {
#pragma unroll
      for (uint32_t pass = 0; pass < 2; ++pass) {
        const bool take = ((pass ? want1 : want0) >> i) & 1ull;
        const uint32_t bal = __ballot_sync(0xffffffffu, take);
        if (take) krow[region[pass] + s_cnt[pass][i * 32 + wid] + __popc(bal & ((1u << lane) - 1u))] = i * kPlanThreads + t;
      }
    }
    if (t == 0) {
      meta[0] = n_steps[0];            // pair steps (two slabs each)
      meta[1] = n_real[0] + n_real[1]; // active variants
      meta[2] = n_steps[1];            // single steps
      meta[3] = n_real[0];
      meta[4] = n_real[1];
    }
    return;
  }
  // ---- block (c, part): max |x| over both planes of column c and its constant term over one part of the rows;
  // the last block of a column to finish combines the parts in index order (a fixed reduction order whatever the
  // block schedule).  One block per column walked 64 rows per thread at one L2 round trip each: 190 us. ----
  const uint32_t c = blockIdx.x / kPlanParts, pt = blockIdx.x % kPlanParts;
  const uint32_t per = (nrows + kPlanParts - 1) / kPlanParts;
  const uint32_t r_end = min(nrows, (pt + 1) * per);
  double mx = 0.0, cs = 0.0;
  if (c == K + 1) {
    if (mode != 0) {
      for (uint32_t r = pt * per + t; r < r_end; r += kPlanThreads) {
        if (srow[r].flags & kScoreActive) mx = 1.0;
      }
    }
  } else {
#pragma unroll 2
    for (uint32_t r = pt * per + t; r < r_end; r += kPlanThreads) {
      const ScoreRow& R = srow[r];
      const uint32_t fl = R.flags;
      const double t0 = R.T[0], t2 = R.T[2], t3 = R.T[3];
      const double w = col_weight(weights, K, mode, r, c);
      const double xg = w * (0.5 * (t2 - t0));
      const double xm = w * (t3 - t0);
      if (fl & kScoreActive) {
        mx = fmax(mx, fmax(fabs(xg), fabs(xm)));
        cs += w * t0;
      }
    }
  }
  // block reduction in a fixed order: butterflies inside a warp, then the 32 warp results in warp order
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    cs += __shfl_xor_sync(0xffffffffu, cs, off);
  }
  if ((t & 31u) == 0) {
    s_red[t >> 5] = mx;
    s_red[32 + (t >> 5)] = cs;
  }
  __syncthreads();
  double partmax = 0.0, partsum = 0.0;
  if (t == 0) {
    for (uint32_t i = 0; i < kPlanThreads / 32; ++i) {
      partmax = fmax(partmax, s_red[i]);
      partsum += s_red[32 + i];
    }
  }
  if (t == 0) {
    double* pm = part + static_cast<size_t>(c) * 2 * kPlanParts;  // [max x kPlanParts][sum x kPlanParts]
    pm[pt] = partmax;
    pm[kPlanParts + pt] = partsum;
    __threadfence();
    if (atomicAdd(part_ctr + c, 1u) == kPlanParts - 1) {
      __threadfence();
      double colmax = 0.0, sum = 0.0;
      for (uint32_t i = 0; i < kPlanParts; ++i) {
        colmax = fmax(colmax, __ldcg(pm + i));
        sum += __ldcg(pm + kPlanParts + i);
      }
      int e = 0;
      double scale = 1.0;
      if (c <= K && colmax > 0.0) {
        frexp(colmax, &e);          // colmax < 2^e
        scale = ldexp(1.0, e - 54);  // |x / scale| < 2^54
      }
      col_scale[c] = scale;
      col_const[c] = sum;
      part_ctr[c] = 0;  // ready for the next batch
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_tc_quant: CTA per step, thread per (k-row, column): fixed-point value -> 7 signed int8 limbs, assembled in
// the shared-memory order of the A slabs (a_offset) and stored as whole slabs.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_tc_quant(DenseRows rows, const ScoreRow* __restrict__ srow, const double* __restrict__ weights, uint32_t K,
           uint32_t mode, const uint32_t* __restrict__ krow, const double* __restrict__ col_scale,
           const uint32_t* __restrict__ meta, int8_t* __restrict__ a_tiles, unsigned long long* __restrict__ krow_ptr) {
  // one CTA per step: its one (single step) or two (pair step: G plane, M plane) slabs are built in shared memory
  // -- padding slots and unused limb rows stay zero -- and leave as whole 4 KB images
  __shared__ __align__(16) int8_t s_slab[2][kABytes];
  const uint32_t ncols = K + 2;
  const uint32_t t = threadIdx.x;
  const uint32_t P = meta[0], S = meta[2];
  const uint32_t st = blockIdx.x;
  if (st >= P + S) return;
  const bool pair = st < P;
  const uint32_t planes = pair ? 2u : 1u;
  const uint32_t slot0 = st * kSlabK;                                  // first slot of the step in the k-row list
  const uint32_t j0 = (pair ? st : st - P) * kSlabK;                   // ... inside its region
  const uint32_t n_real = pair ? meta[3] : meta[4];
  const uint32_t slab_g = pair ? 2u * st : 2u * P + (st - P);
  for (uint32_t i = t; i < planes * kABytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(&s_slab[0][0])[i] = make_uint4(0u, 0u, 0u, 0u);
  if (t < kSlabK) {  // where the variant's packed genotypes start
    const uint32_t r = krow[slot0 + t];
    const uint8_t* row = rows.ptrs ? rows.ptrs[r] : rows.base + static_cast<uint64_t>(r) * rows.stride;
    krow_ptr[slot0 + t] = reinterpret_cast<unsigned long long>(row);
  }
  __syncthreads();
  for (uint32_t item = t; item < kSlabK * ncols; item += blockDim.x) {
    const uint32_t k = item / ncols, c = item - k * ncols;
    if (j0 + k >= n_real) continue;  // padding slots keep their zero limbs
    const uint32_t r = krow[slot0 + k];
    const ScoreRow& R = srow[r];
    const double t0 = R.T[0], t2 = R.T[2], t3 = R.T[3];
    const double w = c <= K ? col_weight(weights, K, mode, r, c) : 0.0;
    const double scale = c <= K ? col_scale[c] : 1.0;
    for (uint32_t plane = 0; plane < planes; ++plane) {
      long long q;
      if (c == K + 1) {
        q = (plane && mode != 0) ? 1 : 0;
      } else {
        const double x = plane ? w * (t3 - t0) : w * (0.5 * (t2 - t0));
        q = __double2ll_rn(x / scale);  // scale is a power of two: the division is exact
      }
#pragma unroll
      for (uint32_t l = 0; l < kLimbs; ++l) {
        const long long b = static_cast<long long>(static_cast<int8_t>(q & 0xff));
        s_slab[plane][a_offset(c * kLimbs + l, k)] = static_cast<int8_t>(b);
        q = (q - b) >> 8;
      }
    }
  }
  __syncthreads();
  uint4* dst = reinterpret_cast<uint4*>(a_tiles + static_cast<uint64_t>(slab_g) * kABytes);
  for (uint32_t i = t; i < planes * kABytes / 16; i += blockDim.x) dst[i] = reinterpret_cast<const uint4*>(&s_slab[0][0])[i];
}

// ---------------------------------------------------------------------------------------------
// k_score_mma: grid = (512-sample tiles, k splits), one CTA per SM.  Warps 0-15 unpack (two groups of 8 taking
// alternate steps), warp 16 issues the MMAs and owns the TMEM allocation, warp 17 copies the A slabs, warps 18-19
// stage the packed rows; after the last slab warps 0-15 drain the accumulators (TMEM lane = limb row, column =
// sample) into acc32[sample][row] with integer reductions, which is what makes k splits order-independent.
// ---------------------------------------------------------------------------------------------
struct MmaSmem {
  alignas(128) uint8_t b[kStages][kBBytes];
  alignas(128) uint8_t a[kAStages][kABytes];
  // packed 2-bit rows of a step (32 variants x 64 samples' bytes, in aligned 16-byte chunks) + each row's misalignment
  alignas(16) uint8_t pk[kPStages][kSlabK][kPRow];
  uint32_t pk_mis[kPStages][kSlabK];
  alignas(8) uint64_t p_full[kPStages];   // loader: 32 cp.async completion arrivals + 1 plain arrival
  uint64_t p_empty[kPStages];             // the 4 converter warps of the step's group have the bytes in registers
  uint64_t full[kStages];              // 4 producer-warp arrivals
  uint64_t empty[kStages];             // tcgen05.commit: the MMA has read B stage s
  uint64_t a_full[kAStages];           // the A copy's expect_tx arrival and its bytes
  uint64_t a_empty[kAStages];          // tcgen05.commit: the MMA has read A stage s
  uint64_t acc_ready;
  uint32_t tmem_base;
};

// 16 packed 2-bit codes -> 16 bytes through the 4-entry byte table `tab`.  A code sitting in the low
// bits of a nibble IS a PRMT selector, so the even codes (w & 0x3333..) and the odd codes
// ((w >> 2) & 0x3333..) are looked up four at a time.  The two byte streams are NOT interleaved back: the
// MMA does not care in which order the 16 samples of a group sit along N, so column j of a group holds sample
// kGroupPerm(j) = 0,2,4,6, 1,3,5,7, 8,10,12,14, 9,11,13,15 and the epilogue undoes it (half the PRMTs).
__device__ __forceinline__ uint4 unpack16(uint32_t w, uint32_t tab) {
  const uint32_t ev = w & 0x33333333u, od = (w >> 2) & 0x33333333u;
  uint4 o;
  o.x = __byte_perm(tab, 0u, ev);        // samples 0, 2, 4, 6
  o.y = __byte_perm(tab, 0u, od);        // samples 1, 3, 5, 7
  o.z = __byte_perm(tab, 0u, ev >> 16);  // samples 8, 10, 12, 14
  o.w = __byte_perm(tab, 0u, od >> 16);  // samples 9, 11, 13, 15
  return o;
}
__device__ __forceinline__ uint32_t group_perm(uint32_t j) { return (j & 8u) | ((j & 3u) << 1) | ((j >> 2) & 1u); }

__device__ __forceinline__ uint4 ldg128_stream(const uint8_t* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// The steps a CTA (k split y of ksplit) works through: its share of the pair steps, then its share of the
// single steps.  Virtual step j -> (first variant slot, first slab, planes).
struct StepRange {
  uint32_t p0, np, s0, ns;       // pair steps [p0, p0 + np), single steps [s0, s0 + ns)
  uint32_t pair_total;           // pair steps of the whole batch
  __device__ __forceinline__ uint32_t steps() const { return np + ns; }
  __device__ __forceinline__ uint32_t planes(uint32_t j) const { return j < np ? 2u : 1u; }
  __device__ __forceinline__ uint32_t slot0(uint32_t j) const {
    return (j < np ? p0 + j : pair_total + s0 + (j - np)) * kSlabK;
  }
  __device__ __forceinline__ uint32_t slab0(uint32_t j) const {   // index into a_tiles
    return j < np ? 2u * (p0 + j) : 2u * pair_total + s0 + (j - np);
  }
  __device__ __forceinline__ uint32_t seq0(uint32_t j) const {    // the CTA's own slab counter at step j
    return j < np ? 2u * j : 2u * np + (j - np);
  }
  __device__ __forceinline__ uint32_t slabs() const { return 2u * np + ns; }
};

__device__ __forceinline__ StepRange step_range(const uint32_t* __restrict__ meta, uint32_t y, uint32_t ksplit) {
  StepRange r;
  const uint32_t P = meta[0], S = meta[2];
  const uint32_t pp = (P + ksplit - 1) / ksplit, ps = (S + ksplit - 1) / ksplit;
  r.pair_total = P;
  r.p0 = min(P, y * pp);
  r.np = min(P, r.p0 + pp) - r.p0;
  r.s0 = min(S, y * ps);
  r.ns = min(S, r.s0 + ps) - r.s0;
  return r;
}

__global__ void __launch_bounds__(kMmaThreads, 1)
k_score_mma(const unsigned long long* __restrict__ krow_ptr, uint32_t n_raw, const int8_t* __restrict__ a_tiles,
            const uint32_t* __restrict__ meta, uint32_t ksplit, int32_t* __restrict__ acc32, uint32_t n_samples,
            uint32_t acc_stride) {
  extern __shared__ uint8_t smem_raw[];
  MmaSmem& sm = *reinterpret_cast<MmaSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t tile = blockIdx.x;
  const StepRange sr = step_range(meta, blockIdx.y, ksplit);
  const uint32_t nsteps = sr.steps();
  if (!nsteps) return;  // uniform for the CTA

  if (warp == kProdWarps) {
    if (lane == 0) {
      for (uint32_t i = 0; i < kStages; ++i) {
        mbar_init(&sm.full[i], kProducers / 32);
        mbar_init(&sm.empty[i], 1);
      }
      for (uint32_t i = 0; i < kAStages; ++i) {
        mbar_init(&sm.a_full[i], 1);
        mbar_init(&sm.a_empty[i], 1);
      }
      for (uint32_t i = 0; i < kPStages; ++i) {
        mbar_init(&sm.p_full[i], 33);
        mbar_init(&sm.p_empty[i], kProducers / 32);
      }
      mbar_init(&sm.acc_ready, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(&sm.tmem_base, kTileN);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sm.tmem_base;

  if (warp < kProdWarps) {
    // ================= producers: packed 2-bit rows -> int8 B slabs =================
    const uint32_t grp = warp >> 3, wq = warp & 7u;
    const uint32_t kk = (wq & 3u) * 8 + (lane & 7u);   // variant within the step = k-row within its slab(s)
    const uint32_t ch = (wq >> 2) * 4 + (lane >> 3);   // which 64 samples of the 512: half = ch >> 2, quarter = ch & 3
    const uint32_t row_bytes = (n_raw + 3) >> 2;
    const uint32_t off = tile * kRowWin + ch * 16;     // byte offset of this chunk inside a row
    // The packed rows of a step arrive through the loader warp (cp.async into sm.pk, below): these warps never
    // have a global load in flight, so the proxy fence in front of every slab hand-over (a MEMBAR that waits for
    // all of the thread's outstanding memory operations) costs shared-memory latency, not HBM latency -- with the
    // loads prefetched by the unpacking threads themselves every fence waited for them and a step took ~2500
    // cycles (profiles/r02_ncu_k_score_mma_*).  Step i of group grp is the CTA's step j = grp + i * kGroups.
    const uint32_t b_thread = smem_u32(sm.b[0]) + (ch >> 2) * kBHalfBytes + (kk & 7u) * 16u + (kk >> 3) * kBLbo +
                              ((ch & 3u) * 4u) * kBSbo;
    const bool in_row = off < row_bytes;  // chunks past the end of the row: zeros
    const uint32_t pk_row = smem_u32(&sm.pk[0][kk][0]) + ch * 16u;
    const uint32_t pk_mis = smem_u32(&sm.pk_mis[0][kk]);
    const uint32_t p_full0 = smem_u32(&sm.p_full[0]), p_empty0 = smem_u32(&sm.p_empty[0]);
    const uint32_t full0 = smem_u32(&sm.full[0]), empty0 = smem_u32(&sm.empty[0]);
    // ring positions advance by addition (the stage counts are not powers of two: no divisions in the loop)
    uint32_t pst = grp, pphase = 0;                 // packed stage of step j = j % kPStages
    uint32_t stage = sr.seq0(grp) % kStages;        // B stage of the step's first slab
    uint32_t sphase = (sr.seq0(grp) / kStages) & 1u;
    for (uint32_t j = grp; j < nsteps; j += kGroups) {
      mbar_wait_addr(p_full0 + pst * 8u, pphase);
      const uint32_t mis = lds32(pk_mis + pst * (kSlabK * 4u));
      // the 16 wanted bytes start `mis` bytes into this thread's two aligned chunks.  Two 128-bit loads: a quarter
      // warp reads one chunk of 8 consecutive rows, and with the 80-byte row pitch those sit in 8 different bank
      // groups (5 r mod 8) -- conflict-free, where word loads at a 20-word pitch were 4-way conflicts and the shared
      // memory data pipe is what bounds this kernel (profiles/r02_ncu_k_score_mma_*: LSU + tensor wavefronts = cycles)
      const uint32_t src = pk_row + pst * (kSlabK * kPRow);
      uint4 c0 = make_uint4(0u, 0u, 0u, 0u), c1 = c0;
      if (in_row) {
        c0 = lds128(src);
        c1 = lds128(src + 16u);
      }
      uint32_t w[5];
      {
        const bool q2 = (mis & 8u) != 0u, q1 = (mis & 4u) != 0u;
        const uint32_t x0 = q2 ? c0.z : c0.x, x1 = q2 ? c0.w : c0.y, x2 = q2 ? c1.x : c0.z, x3 = q2 ? c1.y : c0.w,
                       x4 = q2 ? c1.z : c1.x, x5 = q2 ? c1.w : c1.y;
        w[0] = q1 ? x1 : x0;
        w[1] = q1 ? x2 : x1;
        w[2] = q1 ? x3 : x2;
        w[3] = q1 ? x4 : x3;
        w[4] = q1 ? x5 : x4;
      }
      const uint32_t sh = (mis & 3u) * 8u;
#pragma unroll
      for (int u = 0; u < 4; ++u) w[u] = __funnelshift_r(w[u], w[u + 1], sh);
      __syncwarp();
      if (lane == 0) mbar_arrive_addr(p_empty0 + pst * 8u);  // the bytes are in registers
      const uint32_t planes = sr.planes(j);
      // G plane: 0, 1, 2, missing -> 0.  M plane (pair steps): missing -> 1.  Same bytes, two tables.
      uint32_t st = stage, ph = sphase;
      for (uint32_t pl = 0; pl < planes; ++pl) {
        const uint32_t tab = pl ? 0x01000000u : 0x00020100u;
        mbar_wait_addr(empty0 + st * 8u, ph ^ 1u);
        const uint32_t dst = b_thread + st * kBBytes;
#pragma unroll
        for (int u = 0; u < 4; ++u) sts128(dst + u * kBSbo, unpack16(w[u], tab));
        fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
        __syncwarp();
        if (lane == 0) mbar_arrive_addr(full0 + st * 8u);  // one arrival per warp: 32 would serialise on the barrier word
        if (++st == kStages) {
          st = 0;
          ph ^= 1u;
        }
      }
      // this group's next step is j + kGroups: skip the other group's step in between
      const uint32_t adv = planes + (j + 1 < nsteps ? sr.planes(j + 1) : 0u);  // <= 4 < kStages
      stage += adv;
      if (stage >= kStages) {
        stage -= kStages;
        sphase ^= 1u;
      }
      pst += kGroups;
      if (pst >= kPStages) {
        pst -= kPStages;
        pphase ^= 1u;
      }
    }
    // ================= epilogue: TMEM -> acc32 =================
    mbar_wait(&sm.acc_ready, 0);
    tc_fence_after();
    // 16 warps: TMEM lane quarter = warp % 4 (the only lanes a warp can reach), 128 of the 512 columns each
    const uint32_t lq = warp & 3u;
    const uint32_t r = lq * 32 + lane;  // limb row = TMEM lane
    const uint32_t cbeg = (warp >> 2) * (kTileN / 4);
    for (uint32_t c0 = cbeg; c0 < cbeg + kTileN / 4; c0 += 16) {
      if (tile * kTileN + c0 >= n_samples) break;  // whole groups past the last sample (their half may never have been multiplied)
      uint32_t v[16];
      tmem_ld16(tmem + ((lq * 32u) << 16) + c0, v);
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        const uint32_t s = tile * kTileN + c0 + group_perm(u);
        if (s < n_samples && v[u] != 0u) {
          atomicAdd(acc32 + static_cast<uint64_t>(s) * kRowsM + r, static_cast<int32_t>(v[u]));
        }
      }
    }
    tc_fence_before();
  } else if (warp == kProdWarps) {
    // ================= MMA issuer =================
    // The whole warp walks the slabs (uniform control flow), one elected lane issues.  Stage = slab % kStages for
    // both operands, so the ring is unrolled with compile-time descriptor offsets: per slab one barrier wait, one
    // MMA, one commit.  (A leaner loop matters: with ~65 dependent instructions per slab in one thread the issuer,
    // not the tensor pipe, set the pace -- profiles/r02_ncu_k_score_mma_*.)
    const uint32_t nslabs = sr.slabs();
    const bool two_halves = tile * kTileN + kMmaN < n_samples;  // a ragged last tile may end inside its first half
    const uint64_t a_desc0 = smem_desc(smem_u32(sm.a[0]), kALbo, kASbo);
    const uint64_t b_desc0 = smem_desc(smem_u32(sm.b[0]), kBLbo, kBSbo);
    const uint32_t full0 = smem_u32(&sm.full[0]), empty0 = smem_u32(&sm.empty[0]);
    const uint32_t a_full0 = smem_u32(&sm.a_full[0]), a_empty0 = smem_u32(&sm.a_empty[0]);
    uint32_t aphase = 0;
    for (uint32_t base = 0; base < nslabs; base += kAStages) {
#pragma unroll
      for (uint32_t st = 0; st < kAStages; ++st) {
        if (base + st < nslabs) {  // warp-uniform
          constexpr uint32_t kRounds = kAStages / kStages;
          const uint32_t bst = st % kStages;
          const uint32_t bphase = (aphase * kRounds + st / kStages) & 1u;
          mbar_wait_addr(a_full0 + st * 8u, aphase);
          mbar_wait_addr(full0 + bst * 8u, bphase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t ad = a_desc0 + ((st * kABytes) >> 4), bd = b_desc0 + ((bst * kBBytes) >> 4);
            const uint32_t accumulate = base + st ? 1u : 0u;
            if (two_halves) {
              umma_i8_a_fill(tmem, ad, bd, kIdesc, accumulate);
              umma_i8_a_lastuse(tmem + kMmaN, ad, bd + (kBHalfBytes >> 4), kIdesc, accumulate);
            } else {
              umma_i8(tmem, ad, bd, kIdesc, accumulate);
            }
            umma_commit_addr(empty0 + bst * 8u);    // frees the stages once the MMAs have read them
            umma_commit_addr(a_empty0 + st * 8u);
          }
          __syncwarp();
        }
      }
      aphase ^= 1u;
    }
    if (elect_one()) umma_commit(&sm.acc_ready);
    __syncwarp();
  } else if (warp >= kProdWarps + 2) {
    // ================= loaders: packed rows of every step, global -> shared with cp.async =================
    // (Loading through registers instead -- 9 x LDG.128 per lane, then 128-bit st.shared, six loader warps -- was
    // measured 4 % slower: the LDG return path costs the same pipe more than LDGSTS's sector-wise writes save.)
    // One loader warp per converter group (steps j = g, g + kGroups, ...).  Lane l owns row l of the step (its
    // pointer, its misalignment); the 32 x 9 aligned 16-byte chunks of a step are then copied by consecutive lanes
    // (9 lanes per row: contiguous global reads).  Row pointers are requested two of the warp's steps ahead.
    const uint32_t g = warp - (kProdWarps + 2);
    const uint32_t row_bytes = (n_raw + 3) >> 2;
    const uint32_t tile_off = tile * kRowWin;
    const uint32_t len = min(kRowWin, row_bytes - tile_off);  // wanted bytes of a row (the last tile is ragged)
    uint32_t row_q[kPChunks], ch16_q[kPChunks], dst_q[kPChunks];
#pragma unroll
    for (uint32_t q = 0; q < kPChunks; ++q) {
      const uint32_t idx = q * 32u + lane;
      row_q[q] = idx / kPChunks;
      ch16_q[q] = (idx - row_q[q] * kPChunks) * 16u;
      dst_q[q] = smem_u32(&sm.pk[0][row_q[q]][0]) + ch16_q[q];
    }
    auto ptr_of = [&](uint32_t j) { return j < nsteps ? __ldg(krow_ptr + sr.slot0(j) + lane) : 0ull; };
    unsigned long long kp0 = ptr_of(g), kp1 = ptr_of(g + kGroups);
    const uint32_t p_full0 = smem_u32(&sm.p_full[0]), p_empty0 = smem_u32(&sm.p_empty[0]);
    const uint32_t mis0 = smem_u32(&sm.pk_mis[0][lane]);
    uint32_t pst = g, pphase = 0;
    for (uint32_t j = g; j < nsteps; j += kGroups) {
      const unsigned long long kp = kp0;
      kp0 = kp1;
      kp1 = ptr_of(j + 2 * kGroups);
      mbar_wait_addr(p_empty0 + pst * 8u, pphase ^ 1u);
      const uintptr_t p = static_cast<uintptr_t>(kp) + tile_off;
      const uint32_t mis = static_cast<uint32_t>(p & 15);
      const unsigned long long base = static_cast<unsigned long long>(p & ~uintptr_t(15));
      asm volatile("st.shared.u32 [%0], %1;" ::"r"(mis0 + pst * (kSlabK * 4u)), "r"(mis) : "memory");
      __syncwarp();
      const uint32_t pk_off = pst * (kSlabK * kPRow);
#pragma unroll
      for (uint32_t q = 0; q < kPChunks; ++q) {
        const unsigned long long rb = __shfl_sync(0xffffffffu, base, row_q[q]);
        const uint32_t rmis = __shfl_sync(0xffffffffu, mis, row_q[q]);
        const uint32_t nb = ch16_q[q] < rmis + len ? 16u : 0u;  // chunks past the wanted bytes: zero-filled, not read
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst_q[q] + pk_off),
                     "l"(rb + ch16_q[q]), "r"(nb)
                     : "memory");
      }
      asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(p_full0 + pst * 8u) : "memory");
      if (lane == 0) mbar_arrive_addr(p_full0 + pst * 8u);  // publishes pk_mis (ordered behind the __syncwarp)
      pst += kGroups;
      if (pst >= kPStages) {
        pst -= kPStages;
        pphase ^= 1u;
      }
    }
    cp_async_wait<0>();
  } else {
    // ================= A slabs: one 4 KB bulk copy per slab, into the stage its B slab uses =================
    if (lane == 0) {
      uint32_t sq = 0;
      for (uint32_t j = 0; j < nsteps; ++j) {
        const uint32_t planes = sr.planes(j);
        for (uint32_t pl = 0; pl < planes; ++pl, ++sq) {
          const uint32_t stage = sq % kAStages;
          mbar_wait(&sm.a_empty[stage], ((sq / kAStages) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&sm.a_full[stage], kABytes);
          bulk_g2s(sm.a[stage], a_tiles + static_cast<uint64_t>(sr.slab0(j) + pl) * kABytes, kABytes, &sm.a_full[stage]);
        }
      }
    }
    __syncwarp();
  }
  __syncthreads();
  if (warp == kProdWarps) {
    tc_fence_after();
    tmem_dealloc(tmem, kTileN);
  }
}

// ---------------------------------------------------------------------------------------------
// k_tc_finish: warp per sample, lane per column.  Recombines the seven limb sums exactly
// (value = hi * 2^24 + lo with |hi|, |lo| < 2^53), applies the column scale, adds the constant
// term, accumulates into the fp64 results and clears acc32 for the next batch.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_tc_finish(int32_t* __restrict__ acc32, const double* __restrict__ col_scale, const double* __restrict__ col_const,
            const uint32_t* __restrict__ meta, uint32_t n_samples, uint32_t K, uint32_t mode, double* __restrict__ score,
            double* __restrict__ dosage_sums, uint32_t* __restrict__ allele_cts) {
  // A warp per sample, walking the samples with the grid's stride (a few resident blocks per SM): its 128 limb sums
  // are one 512-byte row -- a 128-bit load and a 128-bit clearing store per lane; the row is turned through shared
  // memory so that lane c gets the 7 limbs of column c (word pitch 7: no bank conflicts).  The old outputs are
  // requested BEFORE the clearing store: with the store directly behind the row load, and one short-lived block per
  // 8 samples, the same pass took 80-210 us instead of 20 (scripts/finish_probe.cu, profiles/r02_finish_probe.json).
  __shared__ __align__(16) int32_t s_row[8][kRowsM];
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  const uint32_t nwarps = gridDim.x * 8u;
  const uint32_t c = lane;
  const uint32_t n_active = meta[1];
  const double scale = c <= K ? col_scale[c] : 0.0, cst = c <= K ? col_const[c] : 0.0;
  for (uint32_t s = blockIdx.x * 8u + wid; s < n_samples; s += nwarps) {
    int4* row = reinterpret_cast<int4*>(acc32 + static_cast<uint64_t>(s) * kRowsM);
    const int4 v = row[lane];
    double old = 0.0;
    uint32_t old_ct = 0;
    if (c < K) {
      old = score[static_cast<uint64_t>(s) * K + c];
    } else if (c == K) {
      if (mode != 2) old = dosage_sums[s];
    } else if (c == K + 1) {
      old_ct = allele_cts[s];
    }
    row[lane] = make_int4(0, 0, 0, 0);
    __syncwarp();  // the previous sample's limbs have been read
    reinterpret_cast<int4*>(s_row[wid])[lane] = v;
    __syncwarp();
    if (c >= K + 2) continue;
    long long limb[kLimbs];
#pragma unroll
    for (uint32_t l = 0; l < kLimbs; ++l) limb[l] = s_row[wid][c * kLimbs + l];
    if (c == K + 1) {
      if (n_active) allele_cts[s] = old_ct + 2u * (n_active - static_cast<uint32_t>(limb[0]));
      continue;
    }
    const long long lo = limb[0] + (limb[1] << 8) + (limb[2] << 16);
    const long long hi = limb[3] + (limb[4] << 8) + (limb[5] << 16) + (limb[6] << 24);
    const double val = fma(static_cast<double>(hi), 16777216.0, static_cast<double>(lo)) * scale + cst;
    if (c < K) {
      score[static_cast<uint64_t>(s) * K + c] = old + val;
    } else if (mode != 2) {
      dosage_sums[s] = old + val;
    }
  }
}

}  // namespace

namespace {
struct TcWorkspace {
  uint32_t* krow;
  unsigned long long* krow_ptr;
  int8_t* a_tiles;
  double* col_scale;
  double* col_const;
  uint32_t* meta;
  double* plan_part;     // [kMaxCols][2][kPlanParts] partial column maxima / constant terms
  uint32_t* plan_ctr;    // [kMaxCols] finished parts per column (zero between launches)
  int32_t* acc32;
  uint32_t acc_stride;  // samples per limb row of acc32, padded
  size_t bytes;
};
// Slots of the k-row list: both regions are padded to whole steps.
size_t slot_cap_of(uint32_t max_rows) { return (static_cast<size_t>(max_rows) + kSlabK - 1) / kSlabK * kSlabK + kSlabK; }
// A slabs: two per pair step, one per single step.
size_t slab_cap_of(uint32_t max_rows) { return 2 * (slot_cap_of(max_rows) / kSlabK); }

TcWorkspace carve(void* base, uint32_t max_rows, uint32_t n_samples) {
  const size_t nk_cap = slot_cap_of(max_rows);
  uint8_t* p = static_cast<uint8_t*>(base);
  size_t off = 0;
  auto take = [&](size_t n) {
    uint8_t* r = p + off;
    off = (off + n + 255) / 256 * 256;
    return r;
  };
  TcWorkspace w;
  w.krow = reinterpret_cast<uint32_t*>(take(nk_cap * 4));
  w.krow_ptr = reinterpret_cast<unsigned long long*>(take(nk_cap * 8));
  w.a_tiles = reinterpret_cast<int8_t*>(take(slab_cap_of(max_rows) * kABytes));
  w.col_scale = reinterpret_cast<double*>(take(kMaxCols * 8));
  w.col_const = reinterpret_cast<double*>(take(kMaxCols * 8));
  w.meta = reinterpret_cast<uint32_t*>(take(32));
  w.plan_part = reinterpret_cast<double*>(take(static_cast<size_t>(kMaxCols) * 2 * kPlanParts * 8));
  w.plan_ctr = reinterpret_cast<uint32_t*>(take(kMaxCols * 4));
  w.acc_stride = (n_samples + 63u) & ~63u;
  w.acc32 = reinterpret_cast<int32_t*>(take(static_cast<size_t>(w.acc_stride) * kRowsM * 4));
  w.bytes = off;
  return w;
}
}  // namespace

size_t score_tc_workspace_bytes(uint32_t max_rows, uint32_t n_samples) {
  return carve(nullptr, max_rows, n_samples).bytes;
}

bool score_tc_applicable(uint32_t K) { return K + 2 <= kMaxCols; }

void launch_score_tc(const DenseRows& rows, const uint4* counts, const ScoreRow* srow, const double* weights, uint32_t K,
                     uint32_t mode, uint32_t n_samples, uint32_t max_rows, void* workspace, bool first_batch,
                     double* score, double* dosage_sums, uint32_t* allele_cts, cudaStream_t s, LaunchStats* st) {
  if (!rows.nrows || !n_samples) return;
  const TcWorkspace w = carve(workspace, max_rows, n_samples);
  if (first_batch) {
    cudaMemsetAsync(w.acc32, 0, static_cast<size_t>(w.acc_stride) * kRowsM * 4, s);
    cudaMemsetAsync(w.plan_ctr, 0, kMaxCols * 4, s);
  }
  const uint32_t slot_cap = static_cast<uint32_t>(slot_cap_of(rows.nrows));
  const uint32_t nk_batch = 2u * slot_cap;  // upper bound of the batch's k-rows (every variant a pair)
  k_tc_plan<<<(K + 2) * kPlanParts + 1, kPlanThreads, 0, s>>>(counts, srow, weights, rows.nrows, K, mode, w.krow,
                                                          w.col_scale, w.col_const, w.meta, w.plan_part, w.plan_ctr);
  k_tc_quant<<<slot_cap / kSlabK, 256, 0, s>>>(rows, srow, weights, K, mode, w.krow, w.col_scale, w.meta, w.a_tiles,
                                               w.krow_ptr);
  const uint32_t tiles = (n_samples + kTileN - 1) / kTileN;
  const int sms = sm_of(st);
  // k splits: the smallest count that fills whole waves of one CTA per SM to >= 90 % (each split pays one
  // accumulator drain per tile), keeping at least 32 slabs per CTA.
  const uint32_t slots = static_cast<uint32_t>(sms);
  const uint32_t max_split = nk_batch / kSlabK / 32u > 1u ? nk_batch / kSlabK / 32u : 1u;
  uint32_t ksplit = 1;
  double best = 0.0;
  for (uint32_t ks = 1; ks <= max_split && ks <= 64; ++ks) {
    const uint64_t ctas = static_cast<uint64_t>(tiles) * ks;
    const uint64_t waves = (ctas + slots - 1) / slots;
    const double eff = static_cast<double>(ctas) / static_cast<double>(waves * slots);
    if (eff > best + 1e-9) {
      best = eff;
      ksplit = ks;
    }
    if (eff >= 0.9) break;
  }
  k_score_mma<<<dim3(tiles, ksplit), kMmaThreads, sizeof(MmaSmem) + 128, s>>>(w.krow_ptr, rows.N, w.a_tiles, w.meta,
                                                                            ksplit, w.acc32, n_samples, w.acc_stride);
  k_tc_finish<<<std::min<uint32_t>((n_samples + 7) / 8, static_cast<uint32_t>(sms) * 8u), 256, 0, s>>>(w.acc32, w.col_scale, w.col_const, w.meta,
                                                                           n_samples, K, mode, score, dosage_sums,
                                                                           allele_cts);
  if (st) st->launches += 4;
}

void configure_device_kernels() {
  cudaFuncSetAttribute(k_score_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sizeof(MmaSmem) + 128));
}

}  // namespace pdk
