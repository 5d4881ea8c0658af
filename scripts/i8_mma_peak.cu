// i8_mma_peak: what the tensor cores of this GPU deliver for tcgen05.mma kind::i8 (M128 x N256 x K32, int32
// accumulators in TMEM) when nothing else is in the way -- operands sit in shared memory, one thread per CTA
// issues back-to-back MMAs on one accumulator.  The denominator for K7t's tensor-pipe fraction (plink_score).
// The same shape is timed with kind::f8f6f4 (e4m3, K32) and kind::f16 (bf16, K16) for comparison.
// Build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a scripts/i8_mma_peak.cu -o scripts/build/i8_mma_peak
// Run:   scripts/build/i8_mma_peak        (prints one JSON object)
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

template <int KIND>  // 0 = i8, 1 = f8f6f4 (e4m3), 2 = f16 (bf16, K = 16), 3 = i8 with two tcgen05.commit after every MMA
__global__ void __launch_bounds__(128) k_peak(uint32_t iters, uint32_t* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint64_t bar2[2];
  __shared__ uint32_t tmem_slot;
  for (uint32_t i = threadIdx.x; i < 12288 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar2[0])));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar2[1])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    auto desc = [](uint32_t addr, uint32_t lbo, uint32_t sbo) {
      return static_cast<uint64_t>((addr & 0x3ffffu) >> 4) | (static_cast<uint64_t>(lbo >> 4) << 16) |
             (static_cast<uint64_t>(sbo >> 4) << 32) | (1ull << 46);
    };
    const uint64_t a_desc = desc(smem_u32(smem), 128, 256);           // K-major, no swizzle
    const uint64_t b_desc = desc(smem_u32(smem + 4096), 2048, 128);   // MN-major, no swizzle
    // c format (S32 = 2, F32 = 1) at [4,6); a / b format at [7,10) / [10,13); B MN-major: bit 16; N >> 3 at [17,23);
    // M >> 4 at [24,29)
    const uint32_t shape = (1u << 16) | ((256u >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t idesc = (KIND == 0 || KIND == 3) ? ((2u << 4) | (1u << 7) | (1u << 10) | shape)
                                     : (KIND == 1 ? ((1u << 4) | shape) : ((1u << 4) | (1u << 7) | (1u << 10) | shape));
    for (uint32_t it = 0; it < iters; ++it) {
      if (KIND == 0 || KIND == 3) {
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem),
                     "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(it)
                     : "memory");
        if (KIND == 3) {
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar2[0])) : "memory");
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar2[1])) : "memory");
        }
      } else if (KIND == 1) {
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem),
                     "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(it)
                     : "memory");
      } else {
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem),
                     "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(it)
                     : "memory");
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(
            smem_u32(&bar))
        : "memory");
    sink[blockIdx.x] = iters;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
  }
}

template <int KIND>
double run(int ctas_per_sm, int sms, uint32_t iters, uint32_t* sink, int k_per_mma) {
  cudaFuncSetAttribute(k_peak<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
  const int grid = sms * ctas_per_sm;
  k_peak<KIND><<<grid, 128, 16384>>>(64, sink);
  cudaDeviceSynchronize();
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  double best = 0;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(a);
    k_peak<KIND><<<grid, 128, 16384>>>(iters, sink);
    cudaEventRecord(b);
    if (cudaEventSynchronize(b) != cudaSuccess) return -1;
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    const double ops = 2.0 * 128 * 256 * k_per_mma * static_cast<double>(iters) * grid;
    const double pops = ops / (ms * 1e-3) / 1e15;
    if (pops > best) best = pops;
  }
  return best;
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  uint32_t* sink;
  cudaMalloc(&sink, 4096 * 4);
  printf("{\"sms\": %d, \"shape\": \"M128 x N256, cta_group::1, operands in shared memory (no swizzle), one accumulator per CTA\"", sms);
  for (int c = 1; c <= 2; ++c) {
    printf(", \"i8_K32_%dcta_per_sm_pops\": %.3f", c, run<0>(c, sms, 200000, sink, 32));
    printf(", \"e4m3_K32_%dcta_per_sm_pflops\": %.3f", c, run<1>(c, sms, 200000, sink, 32));
    printf(", \"bf16_K16_%dcta_per_sm_pflops\": %.3f", c, run<2>(c, sms, 200000, sink, 16));
    printf(", \"i8_K32_two_commits_per_mma_%dcta_per_sm_pops\": %.3f", c, run<3>(c, sms, 200000, sink, 32));
  }
  cudaError_t e = cudaGetLastError();
  printf(", \"cuda\": \"%s\"}\n", cudaGetErrorString(e));
  return 0;
}
