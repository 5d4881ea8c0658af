// pd_hwe_from_counts_noalloc: the exact Hardy-Weinberg test for caller-supplied counts WITHOUT any allocation.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the K8 launch of
// pd_hwe_from_counts / pd_hwe_lnp_from_counts (kernels_hwe.cu: k_hwe + k_hwe_scaled), whose scratch -- the work
// queue: two claim counters, the count and the list of the tables whose centre leaves the fp64 range -- lives here
// in a file-scope device array instead of a context-owned cudaMalloc buffer.  Inputs and outputs are the caller's
// DEVICE buffers; nothing is copied, allocated or freed, and the call does not synchronise.
//
// The array is per GPU (every device that loads the library holds its 4 MiB until exit) and shared by every call on
// that GPU: calls that overlap on the GPU (different streams, or different host threads) race on it.  Serialise them.
#include <cstdio>
#include <mutex>

#include "../../include/plinkgpu.h"
#include "kernels.cuh"

namespace {

constexpr uint32_t kMaxVariants = PD_NOALLOC_HWE_MAX_VARIANTS;
constexpr int kMaxDevices = 64;

// [0] K1h chunk counter (unused here), [1] window counter, [2] overflow count, [3] pad, [4..] overflow list.
// In host code this NAME is a host-side stand-in, not a device address: only cudaGetSymbolAddress yields the pointer
// a kernel may be given.
static __device__ __align__(256) uint32_t g_hwe_scratch[pdk::kHweQueueWords + kMaxVariants];

struct DeviceSlot {
  bool ready = false;
  uint32_t* scratch = nullptr;  // device address of g_hwe_scratch on this device
  int sm_count = 0;
};
std::mutex g_mu;
DeviceSlot g_slot[kMaxDevices];

int Fail(int code, char* errbuf, size_t errbuf_len, const char* what, const char* detail = nullptr) {
  if (errbuf && errbuf_len) {
    if (detail) {
      snprintf(errbuf, errbuf_len, "pd_hwe_from_counts_noalloc: %s: %s", what, detail);
    } else {
      snprintf(errbuf, errbuf_len, "pd_hwe_from_counts_noalloc: %s", what);
    }
  }
  return code;
}

// once per device, with that device current
int SlotOfCurrentDevice(DeviceSlot* out, char* errbuf, size_t errbuf_len) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return Fail(PD_ERR_CUDA, errbuf, errbuf_len, "no CUDA device", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return Fail(PD_ERR_NOT_SUPPORTED, errbuf, errbuf_len, "device ordinal beyond 63");
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceSlot& s = g_slot[dev];
  if (!s.ready) {
    uint32_t* ptr = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_hwe_scratch);
    if (e != cudaSuccess || !ptr) {
      cudaGetLastError();
      return Fail(PD_ERR_CUDA, errbuf, errbuf_len, "cudaGetSymbolAddress(scratch)", cudaGetErrorString(e));
    }
    int sms = 0;
    e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess || sms <= 0) {
      cudaGetLastError();
      return Fail(PD_ERR_CUDA, errbuf, errbuf_len, "cudaDeviceGetAttribute(SM count)", cudaGetErrorString(e));
    }
    s.scratch = ptr;
    s.sm_count = sms;
    s.ready = true;
  }
  *out = s;
  return PD_OK;
}

bool IsDevicePointer(const void* p) {
  cudaPointerAttributes attr{};
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
}

}  // namespace

extern "C" {

uint32_t pd_hwe_noalloc_max_variants(void) { return kMaxVariants; }

int pd_hwe_from_counts_noalloc(const uint32_t* d_counts, uint32_t n, uint32_t midp, uint32_t want_lnp, double* d_out,
                               void* stream, char* errbuf, size_t errbuf_len) {
  if (errbuf && errbuf_len) errbuf[0] = 0;
  if (n > kMaxVariants) {
    char detail[96];
    snprintf(detail, sizeof(detail), "n = %u, limit PD_NOALLOC_HWE_MAX_VARIANTS = %u (split the call)", n, kMaxVariants);
    return Fail(PD_ERR_INVALID_ARG, errbuf, errbuf_len, "too many variants for the static scratch", detail);
  }
  if (!n) return PD_OK;
  if (!d_counts || !d_out) return Fail(PD_ERR_INVALID_ARG, errbuf, errbuf_len, "NULL counts or output pointer");
  if ((reinterpret_cast<uintptr_t>(d_counts) & 15u) || (reinterpret_cast<uintptr_t>(d_out) & 7u)) {
    return Fail(PD_ERR_INVALID_ARG, errbuf, errbuf_len, "d_counts must be 16-byte and d_out 8-byte aligned");
  }
  if (!IsDevicePointer(d_counts) || !IsDevicePointer(d_out)) {
    return Fail(PD_ERR_INVALID_ARG, errbuf, errbuf_len, "d_counts and d_out must be device memory of the current device");
  }
  DeviceSlot slot;
  const int rc = SlotOfCurrentDevice(&slot, errbuf, errbuf_len);
  if (rc != PD_OK) return rc;
  pdk::LaunchStats st{};
  st.sm_count = slot.sm_count;
  // launch_hwe clears the queue header (what must start at zero) on the stream before the kernels; the list behind it
  // is only read up to the count the same launch wrote
  pdk::launch_hwe(reinterpret_cast<const uint4*>(d_counts), n, midp ? 1u : 0u, d_out, slot.scratch,
                  static_cast<cudaStream_t>(stream), &st, want_lnp != 0);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return Fail(PD_ERR_CUDA, errbuf, errbuf_len, "kernel launch", cudaGetErrorString(e));
  return PD_OK;
}

}  // extern "C"
