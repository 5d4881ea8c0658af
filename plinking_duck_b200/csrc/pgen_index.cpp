// See pgen_index.h.  Format facts cite third_party/plink-ng/pgen_spec/pgen_spec.tex and, where
// pgenlib is stricter or differs from the spec text, pgenlib_read.cc.
#include "pgen_index.h"

#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>

#include "../../include/plinkgpu.h"

namespace pdx {

namespace {

struct PinnedRegion {
  const uint8_t* base;
  uint64_t size, piece;
};
std::mutex g_pin_mu;
std::vector<PinnedRegion> g_pinned;

uint64_t PinPieceBytes() {
  const char* e = getenv("PD_PIN_PIECE_MB");
  const long mb = e ? atol(e) : 0;
  return mb >= 2 ? static_cast<uint64_t>(mb) << 20 : 1ull << 30;
}

}  // namespace

ByteSource::~ByteSource() {
  if (!pinned_pieces_.empty()) {
    std::lock_guard<std::mutex> lk(g_pin_mu);
    for (size_t i = 0; i < g_pinned.size(); ++i) {
      if (g_pinned[i].base == map_) {
        g_pinned.erase(g_pinned.begin() + static_cast<long>(i));
        break;
      }
    }
  }
  for (auto& pc : pinned_pieces_) cudaHostUnregister(pc.first);
  if (map_) munmap(const_cast<uint8_t*>(map_), size_);
  if (fd_ >= 0) close(fd_);
}

uint64_t ByteSource::PieceRoom(const void* p) {
  const uint8_t* q = static_cast<const uint8_t*>(p);
  std::lock_guard<std::mutex> lk(g_pin_mu);
  for (const PinnedRegion& r : g_pinned) {
    if (q >= r.base && q < r.base + r.size) return r.piece - static_cast<uint64_t>(q - r.base) % r.piece;
  }
  return UINT64_MAX;
}

bool ByteSource::PinMapping() {
  if (!map_ || !map_writable_ || !size_) return false;
  // 1 GiB pieces on 2 MiB boundaries, registered from a few threads at once (the driver pins ~6.6 GiB/s per call)
  const uint64_t kPiece = PinPieceBytes();
  uint8_t* base = const_cast<uint8_t*>(map_);
  std::vector<std::pair<uint8_t*, uint64_t>> pieces;
  for (uint64_t off = 0; off < size_; off += kPiece) pieces.emplace_back(base + off, std::min(kPiece, size_ - off));
  std::atomic<size_t> next(0);
  std::atomic<bool> ok(true);
  std::vector<char> done(pieces.size(), 0);
  auto work = [&] {
    for (;;) {
      const size_t i = next.fetch_add(1);
      if (i >= pieces.size() || !ok.load()) break;
      if (cudaHostRegister(pieces[i].first, pieces[i].second, cudaHostRegisterPortable) == cudaSuccess) {
        done[i] = 1;
      } else {
        cudaGetLastError();
        ok.store(false);
      }
    }
  };
  const unsigned nt = static_cast<unsigned>(std::min<size_t>(8, pieces.size()));
  std::vector<std::thread> th;
  for (unsigned t = 1; t < nt; ++t) th.emplace_back(work);
  work();
  for (auto& x : th) x.join();
  if (!ok.load()) {
    for (size_t i = 0; i < pieces.size(); ++i) {
      if (done[i]) cudaHostUnregister(pieces[i].first);
    }
    return false;
  }
  pinned_pieces_ = pieces;
  {
    std::lock_guard<std::mutex> lk(g_pin_mu);
    g_pinned.push_back({map_, size_, kPiece});
  }
  mem_ = map_;  // from here on: a page-locked memory image
  return true;
}

bool ByteSource::OpenFile(const char* path, std::string* err, bool want_writable_map) {
  if (want_writable_map) {
    fd_ = open(path, O_RDWR | O_CLOEXEC);
    map_writable_ = fd_ >= 0;
  }
  if (fd_ < 0) fd_ = open(path, O_RDONLY | O_CLOEXEC);
  if (fd_ < 0) {
    if (err) *err = std::string("Failed to open ") + path + " : " + strerror(errno);
    return false;
  }
  struct stat st;
  if (fstat(fd_, &st) != 0) {
    if (err) *err = std::string(path) + " stat failure: " + strerror(errno);
    return false;
  }
  size_ = static_cast<uint64_t>(st.st_size);
  if (size_) {
    void* m = MAP_FAILED;
    if (map_writable_) m = mmap(nullptr, size_, PROT_READ | PROT_WRITE, MAP_SHARED, fd_, 0);
    if (m == MAP_FAILED) {
      map_writable_ = false;
      m = mmap(nullptr, size_, PROT_READ, MAP_SHARED, fd_, 0);
    }
    if (m != MAP_FAILED) map_ = static_cast<const uint8_t*>(m);  // optional: the trimmed scan and pinning need it
  }
  return true;
}

void ByteSource::OpenMem(const void* p, uint64_t len) {
  mem_ = static_cast<const uint8_t*>(p);
  size_ = len;
}

void ByteSource::OpenDevice(const void* p, uint64_t len) {
  dev_ = static_cast<const uint8_t*>(p);
  mem_ = nullptr;
  size_ = len;
}

bool ByteSource::ReadAt(uint64_t off, void* dst, uint64_t len) const {
  if (off > size_ || len > size_ - off) return false;
  if (dev_) return cudaMemcpy(dst, dev_ + off, len, cudaMemcpyDeviceToHost) == cudaSuccess;
  if (mem_) {
    memcpy(dst, mem_ + off, len);
    return true;
  }
  uint8_t* d = static_cast<uint8_t*>(dst);
  while (len) {
    ssize_t got = pread(fd_, d, len > (1u << 30) ? (1u << 30) : len, static_cast<off_t>(off));
    if (got < 0) {
      if (errno == EINTR) continue;
      return false;
    }
    if (got == 0) return false;
    d += got;
    off += static_cast<uint64_t>(got);
    len -= static_cast<uint64_t>(got);
  }
  return true;
}

namespace {

constexpr uint32_t kVblock = 65536;            // pgen_spec.tex:213-217
constexpr uint32_t kMaxVariantCt = 0x7ffffffd; // pgenlib_misc.h:1110-1112
constexpr uint32_t kMaxSampleCt = 0x7ffffffe;

int Fail(std::string* err, int code, const std::string& msg) {
  if (err) *err = msg;
  return code;
}

}  // namespace

int ParsePgenIndex(const ByteSource& pgen, const ByteSource* pgi, PgenIndex* out, std::string* err,
                   uint32_t raw_variant_ct, uint32_t raw_sample_ct) {
  PgenIndex& ix = *out;
  if (pgen.size() < 3) return Fail(err, PD_ERR_MALFORMED, "file is too small to be a valid .pgen file");
  uint8_t magic[3];
  if (!pgen.ReadAt(0, magic, 3)) return Fail(err, PD_ERR_READ, ".pgen read failure");
  if (magic[0] != 0x6c || magic[1] != 0x1b) {
    return Fail(err, PD_ERR_MALFORMED, "not a .pgen file (first two bytes don't match the magic number)");
  }
  ix.mode = magic[2];
  if (ix.mode == 0x00) return Fail(err, PD_ERR_NOT_SUPPORTED, "sample-major PLINK 1 .bed files are not supported");
  if (ix.mode == 0x01) {
    // PLINK 1 variant-major .bed: three header bytes, then V records of ceil(N/4) bytes; V and N come from the
    // caller (pgen_spec.tex:136-138; pgenlib_read.cc:744-790: both counts are required and the size must match).
    if (!raw_variant_ct || !raw_sample_ct) {
      return Fail(err, PD_ERR_INVALID_ARG, "PLINK 1 .bed file: pd_open_opts.raw_variant_ct / raw_sample_ct must be given");
    }
    if (raw_variant_ct > kMaxVariantCt || raw_sample_ct > kMaxSampleCt) return Fail(err, PD_ERR_MALFORMED, "invalid counts for a .bed file");
    const uint64_t width = (static_cast<uint64_t>(raw_sample_ct) + 3) / 4;
    if (3 + width * raw_variant_ct != pgen.size()) {
      char buf[128];
      snprintf(buf, sizeof(buf), "unexpected PLINK 1 .bed file size (expected %llu bytes)",
               static_cast<unsigned long long>(3 + width * raw_variant_ct));
      return Fail(err, PD_ERR_MALFORMED, buf);
    }
    ix.variant_ct = raw_variant_ct;
    ix.sample_ct = raw_sample_ct;
    ix.fixed = true;
    ix.bed = true;
    ix.fixed_width = static_cast<uint32_t>(width);
    ix.max_width = ix.fixed_width;
    ix.const_vrtype = 0;
    ix.data_start = 3;
    ix.data_bytes = width * raw_variant_ct;
    return PD_OK;
  }
  if ((ix.mode & 0xfe) == 0x30) return Fail(err, PD_ERR_MALFORMED, "this is a .pgen.pgi index file, not a .pgen file");

  const bool external = ((ix.mode & 0xfe) == 0x20);
  const ByteSource* hdr = &pgen;
  if (external) {
    if (!pgi) return Fail(err, PD_ERR_OPEN, "storage mode 0x20/0x21 needs the .pgen.pgi index file");
    hdr = pgi;
    uint8_t m2[3];
    if (hdr->size() < 12 || !hdr->ReadAt(0, m2, 3)) return Fail(err, PD_ERR_MALFORMED, ".pgen.pgi file is too small");
    if (m2[0] != 0x6c || m2[1] != 0x1b || (m2[2] & 0xfe) != 0x30) {
      return Fail(err, PD_ERR_MALFORMED, "not a .pgen.pgi file (first three bytes don't match the magic number)");
    }
  } else if (pgen.size() < 12) {
    return Fail(err, PD_ERR_MALFORMED, "file is too small to be a valid .pgen file");
  }
  uint8_t dims[9];
  if (!hdr->ReadAt(3, dims, 9)) return Fail(err, PD_ERR_READ, "header read failure");
  uint32_t V, N;
  memcpy(&V, dims, 4);
  memcpy(&N, dims + 4, 4);
  const uint32_t ctrl = dims[8];
  if (V == 0 || V > kMaxVariantCt) return Fail(err, PD_ERR_MALFORMED, "invalid variant count in .pgen file");
  if (N == 0 || N > kMaxSampleCt) return Fail(err, PD_ERR_MALFORMED, "invalid sample count in .pgen file");
  ix.variant_ct = V;
  ix.sample_ct = N;
  ix.header_ctrl = ctrl;
  const uint32_t nonref_storage = ctrl >> 6;
  const uint64_t geno_bytes = (static_cast<uint64_t>(N) + 3) / 4;

  if (ix.mode < 0x10) {
    // Fixed-width modes (pgen_spec.tex:139-145; pgenlib_read.cc:877-911).
    if (ix.mode > 4) return Fail(err, PD_ERR_NOT_SUPPORTED, "third byte does not correspond to a supported storage mode");
    if (ctrl & 63) {
      return Fail(err, PD_ERR_MALFORMED, "fixed-width storage mode but twelfth byte is only consistent with a variable-width mode");
    }
    uint64_t width = geno_bytes;
    ix.const_vrtype = 0;
    if (ix.mode == 3) {
      ix.const_vrtype = 0x40;
      width += 2ull * N;
      ix.has_dosage = true;
    } else if (ix.mode == 4) {
      ix.const_vrtype = 0xc0;
      width += 4ull * N;
      ix.has_dosage = true;
    }
    if (width > 0xfff00000ull) return Fail(err, PD_ERR_NOT_SUPPORTED, "variant record too wide");
    ix.fixed = true;
    ix.fixed_width = static_cast<uint32_t>(width);
    ix.max_width = ix.fixed_width;
    ix.data_start = 12;
    if (nonref_storage == 3) ix.data_start += (static_cast<uint64_t>(V) + 7) / 8;
    ix.data_bytes = static_cast<uint64_t>(V) * width;
    if (ix.data_start + ix.data_bytes != pgen.size()) {
      char buf[128];
      snprintf(buf, sizeof(buf), "unexpected .pgen file size (expected %llu bytes)",
               static_cast<unsigned long long>(ix.data_start + ix.data_bytes));
      return Fail(err, PD_ERR_MALFORMED, buf);
    }
    return PD_OK;
  }
  if (ix.mode >= 0x12 && !external) {
    return Fail(err, PD_ERR_NOT_SUPPORTED, "third byte does not correspond to a supported storage mode");
  }
  const uint32_t storage = ctrl & 15;
  const uint32_t allele_ct_bytes = (ctrl >> 4) & 3;
  if (allele_ct_bytes > 1) return Fail(err, PD_ERR_NOT_SUPPORTED, ">254 alternate alleles are not supported");
  // Compact header_ctrl modes 8-14 (pgenlib_read.cc:928-957 what they mean, :1260-1355 how a variant block is
  // read): one 1-, 2- or 4-bit code per variant that fixes both the record type and its length.
  // kNoCode marks the codes pgenlib leaves undefined.
  constexpr uint32_t kNoCode = 0xffffffffu;
  uint32_t compact_bits = 0;
  uint8_t compact_type[16];
  uint32_t compact_len[16];
  if (storage >= 8) {
    if (storage == 13 || storage == 15) {
      return Fail(err, PD_ERR_NOT_SUPPORTED, "twelfth byte does not correspond to a supported format (header_ctrl mode 13 / 15)");
    }
    if ((storage == 12 || storage == 14) && N != 1) {
      return Fail(err, PD_ERR_MALFORMED, "single-sample header_ctrl mode on a multi-sample file");
    }
    const uint32_t n4 = (N + 3) / 4;
    for (int i = 0; i < 16; ++i) {
      compact_type[i] = 0;
      compact_len[i] = kNoCode;
    }
    compact_type[0] = 5;
    compact_len[0] = 0;
    compact_type[1] = 0;
    compact_len[1] = n4;
    compact_bits = 2;
    if (storage == 8) {
      compact_bits = 1;
    } else if (storage == 9) {
      compact_type[2] = compact_type[3] = 8;
      compact_len[2] = n4 + 2;
      compact_len[3] = n4 + 3;
    } else if (storage == 10) {
      compact_type[2] = compact_type[3] = 16;
      compact_len[2] = n4 + 1;
      compact_len[3] = n4 + 2;
    } else if (storage == 11) {
      compact_bits = 4;
      for (uint32_t i = 2; i < 8; ++i) {
        compact_type[i] = 8;
        compact_len[i] = n4 + i;
        compact_type[i + 8] = 24;
        compact_len[i + 8] = n4 + 1 + i;
      }
      compact_type[9] = 16;
      compact_len[9] = n4 + 1;
    } else if (storage == 12) {
      compact_type[2] = 0x45;
      compact_len[2] = 2;
      compact_type[3] = 0x40;
      compact_len[3] = 3;
    } else {  // 14
      compact_bits = 4;
      compact_type[3] = 0x10;
      compact_len[3] = 2;
      compact_type[4] = 0x45;
      compact_len[4] = 2;
      compact_type[5] = 0x40;
      compact_len[5] = 3;
      compact_type[12] = 0xc5;
      compact_len[12] = 4;
      compact_type[13] = 0xc0;
      compact_len[13] = 5;
      compact_type[15] = 0xe0;
      compact_len[15] = 6;
    }
  }
  const bool wide_types = (storage & 4) != 0;       // 8-bit vs 4-bit vrtypes
  const uint32_t len_bytes = 1 + (storage & 3);

  const uint32_t vblock_ct = (V + kVblock - 1) / kVblock;
  uint64_t pos = 12;
  uint64_t first_fpos;
  if (!hdr->ReadAt(pos, &first_fpos, 8)) return Fail(err, PD_ERR_READ, "header read failure (variant block offsets)");
  pos += 8ull * vblock_ct;

  ix.vrtypes.resize(V);
  ix.fpos.resize(static_cast<size_t>(V) + 1);
  ix.ldbase.resize(V);
  std::vector<uint8_t> buf;
  uint64_t fpos = first_fpos;
  uint32_t or_types = 0;
  uint32_t max_w = 0;
  for (uint32_t b = 0; b < vblock_ct; ++b) {
    const uint32_t v0 = b * kVblock;
    const uint32_t n = (V - v0 < kVblock) ? V - v0 : kVblock;
    const uint64_t type_bytes = compact_bits ? (static_cast<uint64_t>(n) * compact_bits + 7) / 8 : (wide_types ? n : (n + 1) / 2);
    const uint64_t lens_bytes = compact_bits ? 0 : static_cast<uint64_t>(n) * len_bytes;
    const uint64_t total = type_bytes + lens_bytes;
    buf.resize(total + 4);
    if (!hdr->ReadAt(pos, buf.data(), total)) return Fail(err, PD_ERR_READ, "header read failure (record types/lengths)");
    memset(buf.data() + total, 0, 4);
    const uint8_t* tp = buf.data();
    const uint8_t* lp = buf.data() + type_bytes;
    for (uint32_t i = 0; i < n; ++i) {
      uint32_t vt;
      uint32_t len = 0;
      if (compact_bits) {
        const uint64_t bit = static_cast<uint64_t>(i) * compact_bits;
        const uint32_t code = (tp[bit >> 3] >> (bit & 7)) & ((1u << compact_bits) - 1);
        if (compact_len[code] == kNoCode) return Fail(err, PD_ERR_MALFORMED, "undefined record code in a compact .pgen header");
        vt = compact_type[code];
        len = compact_len[code];
      } else {
        if (wide_types) {
          vt = tp[i];
        } else {
          vt = (tp[i >> 1] >> (4 * (i & 1))) & 15;
        }
        memcpy(&len, lp + static_cast<size_t>(i) * len_bytes, len_bytes);
      }
      ix.vrtypes[v0 + i] = static_cast<uint8_t>(vt);
      ix.fpos[v0 + i] = fpos;
      fpos += len;
      or_types |= vt;
      if (len > max_w) max_w = len;
      const uint32_t main = vt & 7;
      if (main == 2 || main == 3) ix.has_ld = true;
      else if (main == 1 || main >= 4) ix.has_difflist = true;
    }
    pos += total;
    if (allele_ct_bytes) {  // one byte per variant (pgenlib_read.cc:1436-1471)
      if (ix.allele_cts.empty()) ix.allele_cts.assign(V, 2);
      if (!hdr->ReadAt(pos, ix.allele_cts.data() + v0, n)) return Fail(err, PD_ERR_READ, "header read failure (allele counts)");
      pos += static_cast<uint64_t>(n) * allele_ct_bytes;
    }
    if (nonref_storage == 3) pos += (n + 7) / 8;
  }
  ix.fpos[V] = fpos;
  ix.max_width = max_w;
  ix.has_multiallelic = (or_types & 8) != 0;
  ix.has_phase = (or_types & 0x10) != 0;
  ix.has_dosage = (or_types & 0x60) != 0;
  ix.data_start = first_fpos;
  ix.data_bytes = fpos - first_fpos;
  // pgenlib_read.cc:1531-1538: the header may not run past the first record; records may not
  // run past the end of the file.
  if (!external && pos > first_fpos) return Fail(err, PD_ERR_MALFORMED, "invalid .pgen header");
  if (fpos > pgen.size() || first_fpos > fpos) return Fail(err, PD_ERR_MALFORMED, "variant records run past the end of the .pgen file");
  // LD base resolution.  The first record of a variant block may not be LD-compressed
  // (pgen_spec.tex:446-451).
  uint32_t last_base = 0;
  for (uint32_t v = 0; v < V; ++v) {
    const uint32_t main = ix.vrtypes[v] & 7;
    if (main == 2 || main == 3) {
      if ((v % kVblock) == 0) return Fail(err, PD_ERR_MALFORMED, "LD-compressed record at the start of a variant block");
      ix.ldbase[v] = last_base;
    } else {
      last_base = v;
      ix.ldbase[v] = v;
    }
  }
  return PD_OK;
}

}  // namespace pdx
