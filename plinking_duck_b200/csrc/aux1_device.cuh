// Device helpers for the multiallelic patch track (aux1) of a .pgen record: pgen_spec.tex:469-564,
// pgenlib_read.cc:4091-4102 (GetAux1aWidth), :6173-6211 (GetAux1aCodes), :6377-6414 (GetAux1bCodes).
// Shared by K5m (kernels_alleles.cu) and K5p (kernels_dosage.cu: the phase track lies behind aux1 and counts the
// ALTx/ALTy hets among its heterozygous calls).
#pragma once
#include "device_common.cuh"

namespace pdk {

__device__ __forceinline__ uint32_t aux1a_width(uint32_t allele_ct) {  // GetAux1aWidth, pgenlib_read.cc:4091-4102
  return allele_ct < 5 ? allele_ct - 3 : (allele_ct < 7 ? 2u : (allele_ct < 19 ? 4u : 8u));
}
__device__ __forceinline__ uint32_t aux1b_width(uint32_t allele_ct) {  // bits per (x, y) entry
  return allele_ct == 3 ? 1u : (allele_ct < 6 ? 4u : (allele_ct < 18 ? 8u : 16u));
}
// k-th aux1a value -> allele code of the second allele
__device__ __forceinline__ uint32_t aux1a_code(const uint8_t* vals, uint32_t width, uint32_t k) {
  if (!width) return 2u;
  return (load_bits32(vals, static_cast<uint64_t>(k) * width) & ((1u << width) - 1u)) + 2u;
}
// k-th aux1b value -> low byte = first code, high byte = second code
__device__ __forceinline__ uint32_t aux1b_pair(const uint8_t* vals, uint32_t width, uint32_t k) {
  const uint32_t x = load_bits32(vals, static_cast<uint64_t>(k) * width);
  if (width == 1) return (x & 1u) ? 0x0202u : 0x0201u;
  const uint32_t h = width >> 1, m = (1u << h) - 1u;
  return ((x & m) + 1u) | ((((x >> h) & m) + 1u) << 8);
}

// One patch set as the kernel uses it.
struct PatchSet {
  uint32_t mode;        // 0 bit array, 1 id list, 15 empty
  uint32_t count;       // entries
  const uint8_t* fset;  // mode 0
  const uint8_t* vals;
  Difflist list;        // mode 1
};

// Parses one of the two sets at q (warp-uniform; every lane takes part).  raw_ct = calls of the class in the main
// track, width = bits per value.  Leaves q behind the values.
__device__ __forceinline__ bool parse_patch_set(uint32_t mode, const uint8_t*& q, const uint8_t* end, uint32_t N,
                                                uint32_t raw_ct, uint32_t width, int lane, PatchSet& ps) {
  ps.mode = mode;
  ps.count = 0;
  ps.fset = ps.vals = q;
  if (mode == 15) return true;
  if (mode == 0) {
    const uint32_t fset_bytes = (raw_ct + 7) >> 3;
    if (static_cast<uint64_t>(end - q) < fset_bytes) return false;
    ps.fset = q;
    ps.count = group_sum<32>(popcount_bits<32>(q, raw_ct, nullptr, 0, lane));
    q += fset_bytes;
  } else if (mode == 1) {
    if (!parse_difflist(q, end, N, false, ps.list)) return false;
    ps.count = ps.list.len;
    if (ps.list.len) {
      const uint8_t* e = q;
      if (!difflist_walk(ps.list, N, lane, [](uint32_t, uint32_t, uint32_t) {}, &e)) return false;
      q = e;
    }
  } else {
    return false;
  }
  const uint64_t val_bytes = (static_cast<uint64_t>(ps.count) * width + 7) >> 3;
  if (static_cast<uint64_t>(end - q) < val_bytes) return false;
  ps.vals = q;
  q += val_bytes;
  return true;
}

// End of the main track (warp-uniform).
__device__ __forceinline__ bool main_track_end(uint32_t main, const uint8_t* p, const uint8_t* end, uint32_t N, int lane,
                                               const uint8_t** out) {
  const uint8_t* q = p;
  if (main == 0) {
    q = p + ((N + 3) >> 2);
  } else if (main == 5) {
    q = p;
  } else {
    if (main == 1) q = p + 1 + ((N + 7) >> 3);
    if (q > end) return false;
    Difflist d;
    if (!parse_difflist(q, end, N, true, d)) return false;
    if (d.len) {
      const uint8_t* e = q;
      if (!difflist_walk(d, N, lane, [](uint32_t, uint32_t, uint32_t) {}, &e)) return false;
      q = e;
    }
  }
  if (q > end) return false;
  *out = q;
  return true;
}

}  // namespace pdk
