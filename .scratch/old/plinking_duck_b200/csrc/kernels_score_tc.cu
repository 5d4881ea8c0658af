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
// GEMM shape per CTA:  D[128 x 256] (int32, 256 TMEM columns) += A[128 x 32] * B[32 x 256]  per slab
//   A (M side) = weight limbs: row r = column c * 7 + limb l, K-major, prepared once per batch in
//                global memory in the canonical no-swizzle shared-memory order (one 4 KB bulk copy);
//   B (N side) = 256 samples x 32 k-rows of unpacked genotypes, N-major (samples contiguous -- the
//                order the .pgen record already has), written to shared memory by 128 unpack threads
//                from the packed 2-bit rows (PRMT table look-up on nibble-spread codes).
// One tcgen05.mma (M=128, N=256, K=32, kind::i8) per slab, issued by one thread; stages are handed
// over with mbarriers (producers -> MMA) and tcgen05.commit (MMA -> producers).
#include "device_common.cuh"
#include "kernels.cuh"

namespace pdk {

namespace {

constexpr uint32_t kTileN = 256;    // samples per CTA tile (MMA N)
constexpr uint32_t kSlabK = 32;     // k-rows per MMA (K of kind::i8)
constexpr uint32_t kRowsM = 128;    // MMA M: weight-limb rows
constexpr uint32_t kLimbs = 7;
constexpr uint32_t kStages = 4;     // B slabs (8 KB each) between the unpackers and the MMA
constexpr uint32_t kAStages = 6;    // A slabs (4 KB bulk copies): deep enough to cover their L2 latency
constexpr uint32_t kBBytes = kTileN * kSlabK;  // 8 KB
constexpr uint32_t kABytes = kRowsM * kSlabK;  // 4 KB
constexpr uint32_t kProducers = 128;  // threads that unpack one slab
constexpr uint32_t kGroups = 2;       // producer groups; group g unpacks slabs g, g + kGroups, ... of the CTA
constexpr uint32_t kProdWarps = kGroups * kProducers / 32;
constexpr uint32_t kMmaThreads = kGroups * kProducers + 64;  // + one warp (MMA issue, TMEM) + one warp (A copies)
constexpr uint32_t kPrefetch = 3;   // steps of packed bytes a producer group keeps in flight (cp.async ring)
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
// Arrives on the mbarrier once every tcgen05.mma issued so far by this thread has completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
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
constexpr uint32_t kIdesc = (2u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | ((kTileN >> 3) << 17) | ((kRowsM >> 4) << 24);

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
// k_tc_plan: K + 3 CTAs per batch (one per column, one for the k-row list).  Lists the k-rows (a G row per active variant, an M row when it
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

__global__ void __launch_bounds__(kPlanThreads)
k_tc_plan(const uint4* __restrict__ counts, const ScoreRow* __restrict__ srow, const double* __restrict__ weights,
          uint32_t nrows, uint32_t K, uint32_t mode, uint32_t* __restrict__ krow, double* __restrict__ col_scale,
          double* __restrict__ col_const, uint32_t* __restrict__ meta) {
  __shared__ uint32_t s_scan[kPlanThreads];
  __shared__ double s_red[kPlanThreads];
  __shared__ uint32_t s_base, s_active;
  const uint32_t t = threadIdx.x;
  const uint32_t ncols = K + 2;
  if (blockIdx.x == ncols) {
    // ---- block K+2: the k-row list ----
    if (t == 0) {
      s_base = 0;
      s_active = 0;
    }
    __syncthreads();
    for (uint32_t r0 = 0; r0 < nrows; r0 += kPlanThreads) {
      const uint32_t r = r0 + t;
      uint32_t n = 0;
      bool act = false, mis = false;
      if (r < nrows) {
        act = (srow[r].flags & kScoreActive) != 0;
        mis = act && counts[r].w != 0;
        n = (act ? 1u : 0u) + (mis ? 1u : 0u);
      }
      s_scan[t] = n;
      __syncthreads();
      for (uint32_t off = 1; off < kPlanThreads; off <<= 1) {
        const uint32_t add = t >= off ? s_scan[t - off] : 0u;
        __syncthreads();
        s_scan[t] += add;
        __syncthreads();
      }
      const uint32_t pos = s_base + s_scan[t] - n;
      if (act) krow[pos] = r;
      if (mis) krow[pos + 1] = r | kPlaneM;
      const uint32_t nact = __syncthreads_count(act);
      if (t == 0) {
        s_base += s_scan[kPlanThreads - 1];
        s_active += nact;
      }
      __syncthreads();
    }
    const uint32_t nk = s_base;
    const uint32_t nk_pad = (nk + kSlabK - 1) / kSlabK * kSlabK;
    for (uint32_t i = nk + t; i < nk_pad; i += kPlanThreads) krow[i] = 0;  // zero weights; any valid row will do
    if (t == 0) {
      meta[0] = nk_pad;
      meta[1] = s_active;
      meta[2] = nk;
    }
    return;
  }
  // ---- block c: max |x| over both planes of column c, and its constant term ----
  const uint32_t c = blockIdx.x;
  double mx = 0.0, cs = 0.0;
  for (uint32_t r = t; r < nrows; r += kPlanThreads) {
    const ScoreRow& R = srow[r];
    if (!(R.flags & kScoreActive)) continue;
    if (c == K + 1) {
      if (mode != 0) mx = 1.0;
      continue;
    }
    const double w = col_weight(weights, K, mode, r, c);
    const double xg = w * (0.5 * (R.T[2] - R.T[0]));
    const double xm = w * (R.T[3] - R.T[0]);
    mx = fmax(mx, fmax(fabs(xg), fabs(xm)));
    cs += w * R.T[0];
  }
  s_red[t] = mx;
  __syncthreads();
  for (uint32_t off = kPlanThreads / 2; off > 0; off >>= 1) {
    if (t < off) s_red[t] = fmax(s_red[t], s_red[t + off]);
    __syncthreads();
  }
  const double colmax = s_red[0];
  __syncthreads();
  s_red[t] = cs;
  __syncthreads();
  for (uint32_t off = kPlanThreads / 2; off > 0; off >>= 1) {
    if (t < off) s_red[t] += s_red[t + off];
    __syncthreads();
  }
  if (t == 0) {
    int e = 0;
    double scale = 1.0;
    if (c <= K && colmax > 0.0) {
      frexp(colmax, &e);          // colmax < 2^e
      scale = ldexp(1.0, e - 54);  // |x / scale| < 2^54
    }
    col_scale[c] = scale;
    col_const[c] = s_red[0];
  }
}

// ---------------------------------------------------------------------------------------------
// k_tc_quant: thread per (k-row, column): fixed-point value -> 7 signed int8 limbs, written straight
// into the shared-memory order of the A slabs (a_offset).  Slabs are zeroed beforehand.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_tc_quant(DenseRows rows, const ScoreRow* __restrict__ srow, const double* __restrict__ weights, uint32_t K,
           uint32_t mode, const uint32_t* __restrict__ krow, const double* __restrict__ col_scale,
           const uint32_t* __restrict__ meta, uint32_t nk_cap, int8_t* __restrict__ a_tiles,
           unsigned long long* __restrict__ krow_ptr) {
  const uint32_t ncols = K + 2;
  const uint64_t tid = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const uint32_t i = static_cast<uint32_t>(tid / ncols);
  const uint32_t c = static_cast<uint32_t>(tid % ncols);
  if (i >= nk_cap || i >= meta[0]) return;
  const uint32_t kr = krow[i];
  if (c == 0) {  // where the k-row's packed genotypes start, plane flag in bit 63
    const uint32_t rr = kr & ~kPlaneM;
    const uint8_t* row = rows.ptrs ? rows.ptrs[rr] : rows.base + static_cast<uint64_t>(rr) * rows.stride;
    krow_ptr[i] = reinterpret_cast<unsigned long long>(row) | ((kr & kPlaneM) ? (1ull << 63) : 0ull);
  }
  if (i >= meta[2]) return;  // padding k-rows keep their zero limbs
  const uint32_t r = kr & ~kPlaneM;
  const bool plane_m = (kr & kPlaneM) != 0;
  const ScoreRow& R = srow[r];
  long long q;
  if (c == K + 1) {
    q = (plane_m && mode != 0) ? 1 : 0;
  } else {
    const double w = col_weight(weights, K, mode, r, c);
    const double x = plane_m ? w * (R.T[3] - R.T[0]) : w * (0.5 * (R.T[2] - R.T[0]));
    q = __double2ll_rn(x / col_scale[c]);  // scale is a power of two: the division is exact
  }
  int8_t* slab = a_tiles + static_cast<uint64_t>(i / kSlabK) * kABytes;
  const uint32_t k = i % kSlabK;
#pragma unroll
  for (uint32_t l = 0; l < kLimbs; ++l) {
    const long long b = static_cast<long long>(static_cast<int8_t>(q & 0xff));
    slab[a_offset(c * kLimbs + l, k)] = static_cast<int8_t>(b);
    q = (q - b) >> 8;
  }
}

// ---------------------------------------------------------------------------------------------
// k_score_mma: grid = (sample tiles, k splits).  Warps 0-3 unpack, warp 4 issues the MMAs and owns
// the TMEM allocation, warp 5 copies the A slabs; after the last slab warps 0-3 drain the
// accumulators (TMEM lane = limb row, column = sample) into acc32[sample][row] with integer
// reductions, which is what makes k splits order-independent.
// ---------------------------------------------------------------------------------------------
struct MmaSmem {
  alignas(128) uint8_t b[kStages][kBBytes];
  alignas(128) uint8_t a[kAStages][kABytes];
  // packed 2-bit bytes in flight: per group and step, two 16-byte chunks per producer thread
  alignas(128) uint4 p[kGroups][kPrefetch][2][kProducers];
  alignas(8) uint64_t full[kStages];
  uint64_t empty[kStages];
  uint64_t a_full[kAStages];
  uint64_t a_empty[kAStages];
  uint64_t acc_ready;
  uint32_t tmem_base;
};

// 16 packed 2-bit codes -> 16 bytes through the 4-entry byte table `tab`.  A code sitting in the low
// bits of a nibble IS a PRMT selector, so the even codes (w & 0x3333..) and the odd codes
// ((w >> 2) & 0x3333..) are looked up four at a time and the two byte streams interleaved by PRMT.
__device__ __forceinline__ uint4 unpack16(uint32_t w, uint32_t tab) {
  const uint32_t ev = w & 0x33333333u, od = (w >> 2) & 0x33333333u;
  const uint32_t e0 = __byte_perm(tab, 0u, ev), o0 = __byte_perm(tab, 0u, od);              // codes 0,2,4,6 / 1,3,5,7
  const uint32_t e1 = __byte_perm(tab, 0u, ev >> 16), o1 = __byte_perm(tab, 0u, od >> 16);  // codes 8,10,.. / 9,11,..
  uint4 o;
  o.x = __byte_perm(e0, o0, 0x5140u);
  o.y = __byte_perm(e0, o0, 0x7362u);
  o.z = __byte_perm(e1, o1, 0x5140u);
  o.w = __byte_perm(e1, o1, 0x7362u);
  return o;
}

__global__ void __launch_bounds__(kMmaThreads, 2)
k_score_mma(const unsigned long long* __restrict__ krow_ptr, uint32_t n_raw, const int8_t* __restrict__ a_tiles,
            const uint32_t* __restrict__ meta, uint32_t ksplit, int32_t* __restrict__ acc32, uint32_t n_samples) {
  extern __shared__ uint8_t smem_raw[];
  MmaSmem& sm = *reinterpret_cast<MmaSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t tile = blockIdx.x;
  const uint32_t nslab_all = meta[0] / kSlabK;
  const uint32_t per = (nslab_all + ksplit - 1) / ksplit;
  const uint32_t s0 = blockIdx.y * per;
  const uint32_t s1 = min(nslab_all, s0 + per);
  if (s0 >= s1) return;  // uniform for the CTA

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
    const uint32_t grp = warp >> 2, wq = warp & 3u;
    const uint32_t kk = wq * 8 + (lane & 7u);     // k-row within the slab
    const uint32_t gq = lane >> 3;                // which 64 samples of the 256
    const uint32_t row_bytes = (n_raw + 3) >> 2;
    const uint32_t off = tile * (kTileN / 4) + gq * 16;  // byte offset of this quarter inside a row
    const uint32_t nit = s1 - s0;
    const uint32_t tg = threadIdx.x & (kProducers - 1);
    // Pipeline: the k-row's pointer is requested 2 x kPrefetch steps ahead (one register load), its
    // packed bytes kPrefetch steps ahead with cp.async into this thread's own ring slots -- rows start
    // at any byte, so the 32 aligned bytes around the 16 wanted ones are copied.  Asynchronous copies
    // hold no registers and no scoreboard, which is what lets ~20 KB per SM stay in flight.
    // Step j of group grp is the CTA's slab it = grp + j * kGroups.
    auto it_of = [&](uint32_t j) { return grp + j * kGroups; };
    auto fetch_ptr = [&](uint32_t j) {
      return it_of(j) < nit ? __ldg(krow_ptr + static_cast<uint64_t>(s0 + it_of(j)) * kSlabK + kk) : 0ull;
    };
    auto issue = [&](uint32_t j, unsigned long long kp) -> uint32_t {
      uint32_t sp = 0;
      if (it_of(j) < nit) {
        const uintptr_t p = static_cast<uintptr_t>(kp & ~(1ull << 63)) + off;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(p & ~uintptr_t(15));
        const uint32_t nb = off < row_bytes ? 16u : 0u;  // quarters past the end of the row: zeros
        cp_async16(&sm.p[grp][j % kPrefetch][0][tg], src, nb);
        cp_async16(&sm.p[grp][j % kPrefetch][1][tg], src + 16, nb);
        sp = (static_cast<uint32_t>(p & 15) << 3) | ((kp >> 63) ? kPlaneM : 0u);
      }
      cp_async_commit();  // one group per step, empty past the end, so that wait_group counts stay aligned
      return sp;
    };
    auto emit = [&](uint32_t j, uint32_t sp) {
      const uint32_t it = it_of(j);
      const uint32_t stage = it % kStages;
      const uint4 x0 = sm.p[grp][j % kPrefetch][0][tg], x1 = sm.p[grp][j % kPrefetch][1][tg];
      // the 16 wanted bytes start (sp >> 3) & 15 bytes into the 32: whole words first, then a funnel shift
      uint32_t w[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
      if (sp & 64u) {
#pragma unroll
        for (int i = 0; i < 6; ++i) w[i] = w[i + 2];
      }
      if (sp & 32u) {
#pragma unroll
        for (int i = 0; i < 5; ++i) w[i] = w[i + 1];
      }
      const uint32_t sh = sp & 31u;
      const uint32_t tab = (sp & kPlaneM) ? 0x01000000u : 0x00020100u;  // M: [3] -> 1;  G: 0, 1, 2, missing -> 0
      mbar_wait(&sm.empty[stage], ((it / kStages) & 1u) ^ 1u);
      uint8_t* dst = sm.b[stage] + (kk & 7u) * 16u + (kk >> 3) * kBLbo + (gq * 4u) * kBSbo;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        *reinterpret_cast<uint4*>(dst + u * kBSbo) = unpack16(__funnelshift_r(w[u], w[u + 1], sh), tab);
      }
      fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.full[stage]);  // one arrival per warp: 32 would serialise on the barrier word
    };
    uint32_t sps[kPrefetch];
    unsigned long long ptrs[kPrefetch];
#pragma unroll
    for (uint32_t d = 0; d < kPrefetch; ++d) ptrs[d] = fetch_ptr(d);
#pragma unroll
    for (uint32_t d = 0; d < kPrefetch; ++d) {
      sps[d] = issue(d, ptrs[d]);
      ptrs[d] = fetch_ptr(kPrefetch + d);
    }
    for (uint32_t jb = 0; it_of(jb) < nit; jb += kPrefetch) {
#pragma unroll
      for (uint32_t d = 0; d < kPrefetch; ++d) {
        const uint32_t j = jb + d;
        if (it_of(j) < nit) {  // uniform for the group
          cp_async_wait<kPrefetch - 1>();  // step j's copies have landed (this thread reads only its own)
          emit(j, sps[d]);
          sps[d] = issue(j + kPrefetch, ptrs[d]);
          ptrs[d] = fetch_ptr(j + 2 * kPrefetch);
        }
      }
    }
    cp_async_wait<0>();
    // ================= epilogue: TMEM -> acc32 =================
    mbar_wait(&sm.acc_ready, 0);
    tc_fence_after();
    const uint32_t r = wq * 32 + lane;  // limb row = TMEM lane (a warp reaches lanes 32 * (warp % 4) ..)
    for (uint32_t c0 = grp * 16; c0 < kTileN; c0 += 16 * kGroups) {
      uint32_t v[16];
      tmem_ld16(tmem + ((wq * 32u) << 16) + c0, v);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const uint32_t s = tile * kTileN + c0 + j;
        if (s < n_samples && v[j] != 0u) {
          atomicAdd(acc32 + static_cast<uint64_t>(s) * kRowsM + r, static_cast<int32_t>(v[j]));
        }
      }
    }
    tc_fence_before();
  } else if (warp == kProdWarps) {
    // ================= MMA issuer =================
    if (lane == 0) {
      for (uint32_t it = 0; it < s1 - s0; ++it) {
        const uint32_t stage = it % kStages, astage = it % kAStages;
        mbar_wait(&sm.a_full[astage], (it / kAStages) & 1u);
        mbar_wait(&sm.full[stage], (it / kStages) & 1u);
        tc_fence_after();
        const uint64_t a_desc = smem_desc(smem_u32(sm.a[astage]), kALbo, kASbo);
        const uint64_t b_desc = smem_desc(smem_u32(sm.b[stage]), kBLbo, kBSbo);
        umma_i8(tmem, a_desc, b_desc, kIdesc, it ? 1u : 0u);
        umma_commit(&sm.empty[stage]);  // frees the stages once the MMA has read them
        umma_commit(&sm.a_empty[astage]);
      }
      umma_commit(&sm.acc_ready);
    }
    __syncwarp();
  } else {
    // ================= A slabs: one 4 KB bulk copy per stage =================
    if (lane == 0) {
      for (uint32_t it = 0; it < s1 - s0; ++it) {
        const uint32_t astage = it % kAStages;
        mbar_wait(&sm.a_empty[astage], ((it / kAStages) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&sm.a_full[astage], kABytes);
        bulk_g2s(sm.a[astage], a_tiles + static_cast<uint64_t>(s0 + it) * kABytes, kABytes, &sm.a_full[astage]);
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
// k_tc_finish: thread per (sample, column).  Recombines the seven limb sums exactly
// (value = hi * 2^24 + lo with |hi|, |lo| < 2^53), applies the column scale, adds the constant
// term, accumulates into the fp64 results and clears acc32 for the next batch.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_tc_finish(int32_t* __restrict__ acc32, const double* __restrict__ col_scale, const double* __restrict__ col_const,
            const uint32_t* __restrict__ meta, uint32_t n_samples, uint32_t K, uint32_t mode, double* __restrict__ score,
            double* __restrict__ dosage_sums, uint32_t* __restrict__ allele_cts) {
  const uint32_t ncols = K + 2;
  const uint64_t tid = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const uint32_t s = static_cast<uint32_t>(tid / ncols);
  const uint32_t c = static_cast<uint32_t>(tid % ncols);
  if (s >= n_samples) return;
  int32_t* a = acc32 + static_cast<uint64_t>(s) * kRowsM + c * kLimbs;
  long long limb[kLimbs];
#pragma unroll
  for (uint32_t l = 0; l < kLimbs; ++l) {
    limb[l] = a[l];
    a[l] = 0;
  }
  if (c == K + 1) {
    const uint32_t n_active = meta[1];
    if (n_active) allele_cts[s] += 2u * (n_active - static_cast<uint32_t>(limb[0]));
    return;
  }
  const long long lo = limb[0] + (limb[1] << 8) + (limb[2] << 16);
  const long long hi = limb[3] + (limb[4] << 8) + (limb[5] << 16) + (limb[6] << 24);
  const double val = fma(static_cast<double>(hi), 16777216.0, static_cast<double>(lo)) * col_scale[c] + col_const[c];
  if (c < K) {
    score[static_cast<uint64_t>(s) * K + c] += val;
  } else if (mode != 2) {
    dosage_sums[s] += val;
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
  int32_t* acc32;
  size_t bytes;
};
TcWorkspace carve(void* base, uint32_t max_rows, uint32_t n_samples) {
  const size_t nk_cap = (2ull * max_rows + kSlabK - 1) / kSlabK * kSlabK;
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
  w.a_tiles = reinterpret_cast<int8_t*>(take(nk_cap / kSlabK * kABytes));
  w.col_scale = reinterpret_cast<double*>(take(kMaxCols * 8));
  w.col_const = reinterpret_cast<double*>(take(kMaxCols * 8));
  w.meta = reinterpret_cast<uint32_t*>(take(16));
  w.acc32 = reinterpret_cast<int32_t*>(take(static_cast<size_t>(n_samples) * kRowsM * 4));
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
  if (first_batch) cudaMemsetAsync(w.acc32, 0, static_cast<size_t>(n_samples) * kRowsM * 4, s);
  const uint32_t nk_batch = (2u * rows.nrows + kSlabK - 1) / kSlabK * kSlabK;
  cudaMemsetAsync(w.a_tiles, 0, static_cast<size_t>(nk_batch) / kSlabK * kABytes, s);
  k_tc_plan<<<K + 3, kPlanThreads, 0, s>>>(counts, srow, weights, rows.nrows, K, mode, w.krow, w.col_scale, w.col_const,
                                       w.meta);
  const uint64_t qthreads = static_cast<uint64_t>(nk_batch) * (K + 2);
  k_tc_quant<<<static_cast<unsigned>((qthreads + 255) / 256), 256, 0, s>>>(rows, srow, weights, K, mode, w.krow,
                                                                          w.col_scale, w.meta, nk_batch, w.a_tiles,
                                                                          w.krow_ptr);
  const uint32_t tiles = (n_samples + kTileN - 1) / kTileN;
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
    cudaFuncSetAttribute(k_score_mma, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         static_cast<int>(sizeof(MmaSmem) + 128));
  }
  // k splits: the smallest count that fills whole waves of 2 CTAs per SM to >= 90 % (each split pays one
  // accumulator drain per tile), keeping at least 32 slabs per CTA.
  const uint32_t slots = static_cast<uint32_t>(sms) * 2;
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
                                                                            ksplit, w.acc32, n_samples);
  const uint64_t fthreads = static_cast<uint64_t>(n_samples) * (K + 2);
  k_tc_finish<<<static_cast<unsigned>((fthreads + 255) / 256), 256, 0, s>>>(w.acc32, w.col_scale, w.col_const, w.meta,
                                                                           n_samples, K, mode, score, dosage_sums,
                                                                           allele_cts);
  if (st) st->launches += 4;
}

}  // namespace pdk
