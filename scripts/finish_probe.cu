// finish_probe: why does K7t's k_tc_finish (read int32[N][128], clear it, add 18 doubles per sample) take 130-210 us
// for 51 MB?  Times a few formulations of the same pass after the accumulator was filled with RED atomics (as the
// MMA epilogue does) or with plain stores.
// Build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a scripts/finish_probe.cu -o scripts/build/finish_probe
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

constexpr uint32_t kRows = 128, kLimbs = 7;

__global__ void k_fill_atomic(int32_t* acc, uint32_t n) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < static_cast<uint64_t>(n) * kRows) atomicAdd(acc + i, static_cast<int32_t>(i & 63u) - 31);
}
__global__ void k_fill_plain(int32_t* acc, uint32_t n) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < static_cast<uint64_t>(n) * kRows) acc[i] = static_cast<int32_t>(i & 63u) - 31;
}

__device__ __forceinline__ double combine(const long long* limb, double scale, double cst) {
  const long long lo = limb[0] + (limb[1] << 8) + (limb[2] << 16);
  const long long hi = limb[3] + (limb[4] << 8) + (limb[5] << 16) + (limb[6] << 24);
  return fma(static_cast<double>(hi), 16777216.0, static_cast<double>(lo)) * scale + cst;
}

// V0: thread per (sample, column), scalar loads (round-1 kernel)
__global__ void __launch_bounds__(256) k_v0(int32_t* acc, uint32_t n, uint32_t K, double* score, double* dsum, uint32_t* act) {
  const uint32_t ncols = K + 2;
  const uint64_t tid = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const uint32_t s = static_cast<uint32_t>(tid / ncols), c = static_cast<uint32_t>(tid % ncols);
  if (s >= n) return;
  int32_t* a = acc + static_cast<uint64_t>(s) * kRows + c * kLimbs;
  long long limb[kLimbs];
  for (uint32_t l = 0; l < kLimbs; ++l) {
    limb[l] = a[l];
    a[l] = 0;
  }
  if (c == K + 1) {
    act[s] += 2u * (1000u - static_cast<uint32_t>(limb[0]));
    return;
  }
  const double v = combine(limb, 0.5, 1.0);
  if (c < K) score[static_cast<uint64_t>(s) * K + c] += v;
  else dsum[s] += v;
}

// V1: warp per sample through shared memory (FLAGS bit 0: skip the clearing store, bit 1: skip the output RMW)
template <int FLAGS>
__global__ void __launch_bounds__(256) k_v1(int32_t* acc, uint32_t n, uint32_t K, double* score, double* dsum, uint32_t* act) {
  __shared__ __align__(16) int32_t s_row[8][kRows];
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  const uint32_t s = blockIdx.x * 8u + wid;
  if (s >= n) return;
  int4* row = reinterpret_cast<int4*>(acc + static_cast<uint64_t>(s) * kRows);
  const int4 v4 = row[lane];
  if (!(FLAGS & 1)) row[lane] = make_int4(0, 0, 0, 0);
  reinterpret_cast<int4*>(s_row[wid])[lane] = v4;
  __syncwarp();
  const uint32_t c = lane;
  if (c >= K + 2) return;
  long long limb[kLimbs];
  for (uint32_t l = 0; l < kLimbs; ++l) limb[l] = s_row[wid][c * kLimbs + l];
  if (FLAGS & 2) {
    if (limb[0] == 123456789) act[s] = 1;
    return;
  }
  if (c == K + 1) {
    act[s] += 2u * (1000u - static_cast<uint32_t>(limb[0]));
    return;
  }
  const double v = combine(limb, 0.5, 1.0);
  if (c < K) score[static_cast<uint64_t>(s) * K + c] += v;
  else dsum[s] += v;
}

// V2: persistent-ish grid (fewer, longer blocks): a warp walks samples with a stride, loads issued two ahead
__global__ void __launch_bounds__(256) k_v2(int32_t* acc, uint32_t n, uint32_t K, double* score, double* dsum, uint32_t* act) {
  __shared__ __align__(16) int32_t s_row[8][kRows];
  const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  const uint32_t nwarps = gridDim.x * 8u;
  const uint32_t c = lane;
  for (uint32_t s = blockIdx.x * 8u + wid; s < n; s += nwarps) {
    int4* row = reinterpret_cast<int4*>(acc + static_cast<uint64_t>(s) * kRows);
    const int4 v4 = row[lane];
    double old = 0.0;
    if (c < K) old = score[static_cast<uint64_t>(s) * K + c];
    else if (c == K) old = dsum[s];
    uint32_t olda = c == K + 1 ? act[s] : 0u;
    row[lane] = make_int4(0, 0, 0, 0);
    __syncwarp();
    reinterpret_cast<int4*>(s_row[wid])[lane] = v4;
    __syncwarp();
    if (c < K + 2) {
      long long limb[kLimbs];
      for (uint32_t l = 0; l < kLimbs; ++l) limb[l] = s_row[wid][c * kLimbs + l];
      if (c == K + 1) act[s] = olda + 2u * (1000u - static_cast<uint32_t>(limb[0]));
      else {
        const double v = old + combine(limb, 0.5, 1.0);
        if (c < K) score[static_cast<uint64_t>(s) * K + c] = v;
        else dsum[s] = v;
      }
    }
  }
}

int main() {
  const uint32_t n = 100000, K = 16;
  int32_t* acc;
  double *score, *dsum;
  uint32_t* act;
  cudaMalloc(&acc, static_cast<size_t>(n) * kRows * 4);
  cudaMalloc(&score, static_cast<size_t>(n) * K * 8);
  cudaMalloc(&dsum, n * 8);
  cudaMalloc(&act, n * 4);
  cudaMemset(score, 0, static_cast<size_t>(n) * K * 8);
  cudaMemset(dsum, 0, n * 8);
  cudaMemset(act, 0, n * 4);
  cudaMemset(acc, 0, static_cast<size_t>(n) * kRows * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const unsigned fill_blocks = static_cast<unsigned>((static_cast<uint64_t>(n) * kRows + 255) / 256);
  printf("{");
  for (int fill = 0; fill < 2; ++fill) {
    for (int v = 0; v < 6; ++v) {
      float best = 1e9f;
      for (int rep = 0; rep < 5; ++rep) {
        if (fill == 0) k_fill_atomic<<<fill_blocks, 256>>>(acc, n);
        else k_fill_plain<<<fill_blocks, 256>>>(acc, n);
        cudaEventRecord(e0);
        switch (v) {
          case 0: k_v0<<<(n * (K + 2) + 255) / 256, 256>>>(acc, n, K, score, dsum, act); break;
          case 1: k_v1<0><<<(n + 7) / 8, 256>>>(acc, n, K, score, dsum, act); break;
          case 2: k_v1<1><<<(n + 7) / 8, 256>>>(acc, n, K, score, dsum, act); break;
          case 3: k_v1<2><<<(n + 7) / 8, 256>>>(acc, n, K, score, dsum, act); break;
          case 4: k_v2<<<148 * 8, 256>>>(acc, n, K, score, dsum, act); break;
          case 5: k_v2<<<148 * 2, 256>>>(acc, n, K, score, dsum, act); break;
        }
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
      }
      printf("%s\"%s_v%d_us\": %.1f", (fill || v) ? ", " : "", fill ? "plain" : "atomic", v, best * 1e3f);
    }
  }
  printf(", \"err\": \"%s\"}\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
