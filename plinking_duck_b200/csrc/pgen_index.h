// Host-side .pgen header / per-variant index (parsed once per context).
//
// Replaces the per-thread PgfiInitPhase1 + PgfiInitPhase2Ex of the reference
// (third_party/plink-ng/2.0/include/pgenlib_read.cc:689-961, 1099-1680) with one parse whose
// result (vrtypes[V], fpos[V+1], ldbase[V]) is uploaded to the GPU.  Written from the format
// specification (third_party/plink-ng/pgen_spec/pgen_spec.tex:122-320); not a port of pgenlib.
#pragma once
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

namespace pdx {

// Random-access byte source over either a file descriptor or a caller-owned memory image.
class ByteSource {
 public:
  ByteSource() = default;
  ~ByteSource();
  ByteSource(const ByteSource&) = delete;
  ByteSource& operator=(const ByteSource&) = delete;
  // want_writable_map: map the file PROT_READ|PROT_WRITE, MAP_SHARED (needed to page-lock it in place; it is
  // never written); falls back to a read-only mapping.
  bool OpenFile(const char* path, std::string* err, bool want_writable_map = false);
  bool map_writable() const { return map_writable_; }
  // Page-locks the whole mapping for DMA (cudaHostRegister, in parallel pieces); afterwards the source behaves
  // like a page-locked memory image (mem() == the mapping).  False (nothing left registered) on any failure.
  bool PinMapping();
  bool pinned() const { return !pinned_pieces_.empty(); }
  // Page-locked pieces are separate registrations and one DMA must not span two of them.  PieceRoom(p): how many
  // bytes from host address p stay inside one registration of a mapping pinned by ANY ByteSource of this process
  // (the shards of a multi-GPU context read the parent's mapping as a plain memory image); UINT64_MAX when p is
  // not inside such a mapping.  Piece size: 1 GiB (PD_PIN_PIECE_MB overrides it so tests cross pieces with small files).
  static uint64_t PieceRoom(const void* p);
  void OpenMem(const void* p, uint64_t len);
  // A .pgen image that already lives in device memory (reads go through cudaMemcpy).
  void OpenDevice(const void* p, uint64_t len);
  const uint8_t* dev() const { return dev_; }
  uint64_t size() const { return size_; }
  bool is_mem() const { return mem_ != nullptr; }
  const uint8_t* mem() const { return mem_; }
  // Host-addressable bytes of the whole source: the caller's memory, or a read-only mapping of the
  // file (nullptr if the mapping failed or the image lives on the device).
  const uint8_t* view() const { return mem_ ? mem_ : map_; }
  // Reads exactly len bytes at off into dst; false on short read / error.
  bool ReadAt(uint64_t off, void* dst, uint64_t len) const;

 private:
  int fd_ = -1;
  bool map_writable_ = false;
  std::vector<std::pair<uint8_t*, uint64_t>> pinned_pieces_;
  const uint8_t* mem_ = nullptr;
  const uint8_t* map_ = nullptr;
  const uint8_t* dev_ = nullptr;
  uint64_t size_ = 0;
};

struct PgenIndex {
  uint32_t variant_ct = 0;
  uint32_t sample_ct = 0;
  uint32_t mode = 0;
  uint32_t header_ctrl = 0;
  bool fixed = false;          // storage modes 0x01-0x04
  bool bed = false;            // mode 0x01: PLINK 1 codes on disk (3 hom-ref, 2 het, 0 hom-alt, 1 missing); staged bytes
                               // are rewritten to .pgen codes on the device before any kernel sees them
  uint32_t fixed_width = 0;    // bytes per record when fixed
  uint32_t const_vrtype = 0;   // vrtype of every record when fixed
  uint64_t data_start = 0;     // file offset of record 0
  uint64_t data_bytes = 0;     // sum of record widths
  // Variable-width modes only (empty when fixed):
  std::vector<uint8_t> vrtypes;   // [V]
  std::vector<uint64_t> fpos;     // [V+1]
  // ldbase[v] = index of the last variant <= v whose main track is not LD-compressed
  // (semantics of GetLdbaseVidx, pgenlib_read.cc:1846-1894); ldbase[v] == v for non-LD records.
  std::vector<uint32_t> ldbase;   // [V]
  // allele count of every variant (header_ctrl bits 4-5 = 1: one byte each); empty when the file stores none
  // (every variant biallelic as far as the reader is concerned)
  std::vector<uint8_t> allele_cts;  // [V]
  bool has_ld = false, has_difflist = false, has_dosage = false, has_phase = false, has_multiallelic = false;
  uint32_t max_width = 0;

  uint64_t RecOff(uint32_t v) const { return fixed ? data_start + static_cast<uint64_t>(v) * fixed_width : fpos[v]; }
  uint32_t Vrtype(uint32_t v) const { return fixed ? const_vrtype : vrtypes[v]; }
  uint32_t Ldbase(uint32_t v) const { return fixed ? v : ldbase[v]; }
  uint32_t AlleleCt(uint32_t v) const { return allele_cts.empty() ? 2u : allele_cts[v]; }
};

// Error codes match include/plinkgpu.h PD_ERR_*.
// raw_variant_ct / raw_sample_ct: required for PLINK 1 .bed files (mode 0x01), ignored otherwise.
int ParsePgenIndex(const ByteSource& pgen, const ByteSource* pgi, PgenIndex* out, std::string* err,
                   uint32_t raw_variant_ct = 0, uint32_t raw_sample_ct = 0);

}  // namespace pdx
