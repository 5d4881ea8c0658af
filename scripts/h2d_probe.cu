// h2d_probe: what the box gives the staging layer, measured without the library.
//   (1) raw concurrent page-locked H2D bandwidth at 1..n devices (cudaMemcpyAsync of 1 GiB blocks, one stream
//       per device, all devices at once) -- the ceiling the e2e numbers are quoted against;
//   (2) a tmpfs file mapped read-only and page-locked in place with cudaHostRegister: registration cost per
//       GiB and the H2D bandwidth straight from the page cache (no host copy);
//   (3) pread() of the same file into a page-locked buffer with T threads (the copy the default path pays).
// Build: nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a scripts/h2d_probe.cu -o scripts/build/h2d_probe
// Run:   scripts/build/h2d_probe [GiB per device = 4] [file GiB = 4]      (prints one JSON object)
#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

static double now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      fprintf(stderr, "%s failed: %s (line %d)\n", #x, cudaGetErrorString(e_), __LINE__); \
      exit(2);                                                                         \
    }                                                                                  \
  } while (0)

int main(int argc, char** argv) {
  const size_t gib = 1ull << 30;
  const size_t per_dev = (argc > 1 ? atoi(argv[1]) : 4) * gib;
  const size_t file_bytes = (argc > 2 ? atoi(argv[2]) : 4) * gib;
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  printf("{\"devices\": %d, \"host_threads\": %u", ndev, std::thread::hardware_concurrency());

  // ---- (1) raw pinned H2D, k devices at once ----
  std::vector<void*> h(ndev), d(ndev);
  std::vector<cudaStream_t> st(ndev);
  for (int i = 0; i < ndev; ++i) {
    CK(cudaSetDevice(i));
    CK(cudaHostAlloc(&h[i], gib, cudaHostAllocPortable));
    memset(h[i], i + 1, gib);
    CK(cudaMalloc(&d[i], gib));
    CK(cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking));
  }
  printf(", \"pinned_h2d_gbs\": {");
  for (int k = 1; k <= ndev; k *= 2) {
    double best = 0;
    for (int rep = 0; rep < 3; ++rep) {
      for (int i = 0; i < k; ++i) {
        CK(cudaSetDevice(i));
        CK(cudaStreamSynchronize(st[i]));
      }
      const double t0 = now();
      for (size_t off = 0; off < per_dev; off += gib) {
        for (int i = 0; i < k; ++i) {
          CK(cudaSetDevice(i));
          CK(cudaMemcpyAsync(d[i], h[i], gib, cudaMemcpyHostToDevice, st[i]));
        }
      }
      for (int i = 0; i < k; ++i) {
        CK(cudaSetDevice(i));
        CK(cudaStreamSynchronize(st[i]));
      }
      const double gbs = static_cast<double>(per_dev) * k / (now() - t0) / 1e9;
      if (gbs > best) best = gbs;
    }
    printf("%s\"%d\": %.1f", k > 1 ? ", " : "", k, best);
  }
  printf("}");

  // ---- (2) + (3): a tmpfs file ----
  const char* dir = access("/dev/shm", W_OK) == 0 ? "/dev/shm" : "/tmp";
  const std::string path = std::string(dir) + "/h2d_probe_" + std::to_string(getpid()) + ".bin";
  int fd = open(path.c_str(), O_CREAT | O_RDWR | O_TRUNC, 0600);
  if (fd < 0) {
    printf(", \"file\": \"cannot create %s\"}\n", path.c_str());
    return 0;
  }
  for (size_t off = 0; off < file_bytes; off += gib) {
    if (write(fd, h[0], gib) != static_cast<ssize_t>(gib)) {
      printf(", \"file\": \"short write\"}\n");
      unlink(path.c_str());
      return 0;
    }
  }
  printf(", \"file\": \"%s\", \"file_gib\": %zu", dir, file_bytes / gib);
  // (3) pread into the pinned buffer, T threads, 2 MiB chunks
  printf(", \"pread_into_pinned_gbs\": {");
  bool first = true;
  for (unsigned T : {4u, 8u, 16u, 32u}) {
    if (T > 2 * std::thread::hardware_concurrency()) break;
    const size_t chunk = 2u << 20;
    std::atomic<size_t> next{0};
    const size_t nchunks = file_bytes / chunk;
    const double t0 = now();
    std::vector<std::thread> th;
    for (unsigned t = 0; t < T; ++t) {
      th.emplace_back([&] {
        for (;;) {
          const size_t i = next.fetch_add(1);
          if (i >= nchunks) break;
          const size_t off = i * chunk;
          if (pread(fd, static_cast<char*>(h[0]) + (off % gib), chunk, static_cast<off_t>(off)) != static_cast<ssize_t>(chunk)) break;
        }
      });
    }
    for (auto& x : th) x.join();
    printf("%s\"%u\": %.1f", first ? "" : ", ", T, static_cast<double>(file_bytes) / (now() - t0) / 1e9);
    first = false;
  }
  printf("}");
  // (4) small page-locked chunks that stay cache-resident between the host copy and the DMA: T threads, each
  //     with its own stream and a ring of 3 chunks; pread (or memcpy from a mapping) -> cudaMemcpyAsync.
  {
    void* view = mmap(nullptr, file_bytes, PROT_READ, MAP_SHARED | MAP_POPULATE, fd, 0);
    printf(", \"chunked_file_to_device_gbs\": {");
    bool first_c = true;
    for (int use_memcpy = 0; use_memcpy < 2; ++use_memcpy) {
      if (use_memcpy && view == MAP_FAILED) break;
      for (size_t chunk_mib : {1, 2, 4, 8}) {
        for (unsigned T : {8u, 16u, 24u}) {
          if (T > std::thread::hardware_concurrency()) continue;
          const size_t chunk = chunk_mib << 20;
          const size_t nchunks = file_bytes / chunk;
          std::atomic<size_t> next{0};
          std::vector<std::thread> th;
          const double t0 = now();
          for (unsigned t = 0; t < T; ++t) {
            th.emplace_back([&, t] {
              const int dev = static_cast<int>(t % ndev);
              cudaSetDevice(dev);
              cudaStream_t s;
              cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
              char* ring;
              cudaHostAlloc(reinterpret_cast<void**>(&ring), 3 * chunk, cudaHostAllocPortable);
              cudaEvent_t ev[3];
              for (auto& e : ev) cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
              size_t k = 0;
              for (;;) {
                const size_t i = next.fetch_add(1);
                if (i >= nchunks) break;
                const int slot = static_cast<int>(k % 3);
                if (k >= 3) cudaEventSynchronize(ev[slot]);
                char* buf = ring + slot * chunk;
                if (use_memcpy) memcpy(buf, static_cast<char*>(view) + i * chunk, chunk);
                else if (pread(fd, buf, chunk, static_cast<off_t>(i * chunk)) != static_cast<ssize_t>(chunk)) break;
                cudaMemcpyAsync(static_cast<char*>(d[dev]) + (i * chunk) % gib, buf, chunk, cudaMemcpyHostToDevice, s);
                cudaEventRecord(ev[slot], s);
                ++k;
              }
              cudaStreamSynchronize(s);
              for (auto& e : ev) cudaEventDestroy(e);
              cudaFreeHost(ring);
              cudaStreamDestroy(s);
            });
          }
          for (auto& x : th) x.join();
          printf("%s\"%s_%zuMiB_T%u\": %.1f", first_c ? "" : ", ", use_memcpy ? "memcpy" : "pread", chunk_mib, T,
                 static_cast<double>(file_bytes) / (now() - t0) / 1e9);
          first_c = false;
        }
      }
    }
    printf("}");
    if (view != MAP_FAILED) munmap(view, file_bytes);
  }
  // (5) a writable shared mapping, page-locked in place without the read-only flag
  {
    void* mw = mmap(nullptr, file_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (mw != MAP_FAILED) {
      cudaSetDevice(0);
      const double t0 = now();
      cudaError_t e = cudaHostRegister(mw, file_bytes, cudaHostRegisterPortable);
      const double reg_s = now() - t0;
      if (e == cudaSuccess) {
        printf(", \"register_rw_s_per_gib\": %.4f", reg_s / (static_cast<double>(file_bytes) / gib));
        double best = 0;
        for (int rep = 0; rep < 3; ++rep) {
          const double t1 = now();
          for (size_t off = 0; off < file_bytes; off += gib) {
            cudaMemcpyAsync(d[0], static_cast<char*>(mw) + off, gib, cudaMemcpyHostToDevice, st[0]);
          }
          cudaStreamSynchronize(st[0]);
          const double gbs = static_cast<double>(file_bytes) / (now() - t1) / 1e9;
          if (gbs > best) best = gbs;
        }
        printf(", \"registered_rw_file_h2d_gbs\": %.1f", best);
        cudaHostUnregister(mw);
      } else {
        cudaGetLastError();
        printf(", \"register_rw\": \"%s\"", cudaGetErrorString(e));
      }
      munmap(mw, file_bytes);
    }
  }
  // (2) map + register in place
  void* m = mmap(nullptr, file_bytes, PROT_READ, MAP_SHARED, fd, 0);
  if (m == MAP_FAILED) {
    printf(", \"register\": \"mmap failed\"}\n");
    unlink(path.c_str());
    return 0;
  }
  CK(cudaSetDevice(0));
  double t0 = now();
  cudaError_t e = cudaHostRegister(m, file_bytes, cudaHostRegisterReadOnly | cudaHostRegisterPortable);
  double reg_s = now() - t0;
  if (e != cudaSuccess) {
    cudaGetLastError();
    printf(", \"register_readonly\": \"%s\"", cudaGetErrorString(e));
    t0 = now();
    e = cudaHostRegister(m, file_bytes, cudaHostRegisterPortable);
    reg_s = now() - t0;
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    printf(", \"register\": \"%s\"}\n", cudaGetErrorString(e));
    unlink(path.c_str());
    return 0;
  }
  printf(", \"register_s_per_gib\": %.4f", reg_s / (static_cast<double>(file_bytes) / gib));
  double best = 0;
  for (int rep = 0; rep < 3; ++rep) {
    t0 = now();
    for (size_t off = 0; off < file_bytes; off += gib) {
      CK(cudaMemcpyAsync(d[0], static_cast<char*>(m) + off, gib, cudaMemcpyHostToDevice, st[0]));
    }
    CK(cudaStreamSynchronize(st[0]));
    const double gbs = static_cast<double>(file_bytes) / (now() - t0) / 1e9;
    if (gbs > best) best = gbs;
  }
  printf(", \"registered_file_h2d_gbs\": %.1f", best);
  // registration in 256 MiB pieces from 4 threads (what an open could overlap with the first scan)
  CK(cudaHostUnregister(m));
  {
    const size_t piece = 256u << 20;
    std::atomic<size_t> next{0};
    const size_t npieces = file_bytes / piece;
    t0 = now();
    std::vector<std::thread> th;
    for (int t = 0; t < 4; ++t) {
      th.emplace_back([&] {
        cudaSetDevice(0);
        for (;;) {
          const size_t i = next.fetch_add(1);
          if (i >= npieces) break;
          cudaHostRegister(static_cast<char*>(m) + i * piece, piece, cudaHostRegisterReadOnly | cudaHostRegisterPortable);
        }
      });
    }
    for (auto& x : th) x.join();
    printf(", \"register_4threads_s_per_gib\": %.4f", (now() - t0) / (static_cast<double>(file_bytes) / gib));
  }
  printf("}\n");
  munmap(m, file_bytes);
  close(fd);
  unlink(path.c_str());
  return 0;
}
