// Device-side building blocks shared by the scan kernels (sm_100a).
//
// Genotype encoding (pgen_spec.tex:428-436): 2 bits per sample, LSB first,
// 00 hom-ref, 01 het, 10 hom-alt, 11 missing.  "E" / "O" below are the even / odd bit
// planes of a packed genovec; B = E & O marks missing calls.  Counting identity
// (same algebra pgenlib uses, pgenlib_read.cc:141-144):
//   c1 = |E| - |B|, c2 = |O| - |B|, c3 = |B|, c0 = n - c1 - c2 - c3.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pdk {

constexpr uint32_t kM5 = 0x55555555u;
constexpr uint32_t kMA = 0xAAAAAAAAu;
constexpr int kErrMalformed = 3;      // PD_ERR_MALFORMED
constexpr int kErrNotSupported = 4;   // PD_ERR_NOT_SUPPORTED

// 128-bit streaming load: read-only path, do not allocate in L1 (each byte is used once).
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ uint32_t bitselect(uint32_t mask, uint32_t a, uint32_t b) {
  // (a & mask) | (b & ~mask): one LOP3.
  return (a & mask) | (b & ~mask);
}

struct Planes {
  uint32_t e, o, b;  // running popcounts of the E, O and B planes
};

// Two 32-bit genotype words -> the three plane popcounts with 3 POPC instead of 6:
// word a keeps its E plane on even bit positions, word b's E plane is moved to the odd
// positions of the same register (and likewise for O), so one POPC covers both words.
__device__ __forceinline__ void acc_pair(uint32_t a, uint32_t b, Planes& c) {
  const uint32_t x = bitselect(kM5, a, b << 1);   // E(a) | E(b) << 1
  const uint32_t y = bitselect(kM5, a >> 1, b);   // O(a) | O(b) << 1
  c.e += __popc(x);
  c.o += __popc(y);
  c.b += __popc(x & y);
}

__device__ __forceinline__ void acc_chunk(const uint4& v, Planes& c) {
  acc_pair(v.x, v.y, c);
  acc_pair(v.z, v.w, c);
}

// Masked flavour: m* hold 01 per included sample (2-bit expanded sample_include).
__device__ __forceinline__ void acc_pair_masked(uint32_t a, uint32_t b, uint32_t ma, uint32_t mb, Planes& c) {
  const uint32_t x = (a & ma) | ((b & mb) << 1);
  const uint32_t y = ((a >> 1) & ma) | (b & (mb << 1));
  c.e += __popc(x);
  c.o += __popc(y);
  c.b += __popc(x & y);
}

// Bits b of a 32-bit word whose stream position (word_bit0 + b) lies in [lo, hi).
__device__ __forceinline__ uint32_t valid_bits(int64_t word_bit0, int64_t lo, int64_t hi) {
  int64_t l = lo - word_bit0;
  int64_t h = hi - word_bit0;
  if (l < 0) l = 0;
  if (h > 32) h = 32;
  if (h <= l) return 0u;
  const uint32_t upto_h = (h == 32) ? 0xffffffffu : ((1u << h) - 1u);
  const uint32_t upto_l = (1u << l) - 1u;  // l < 32 here
  return upto_h & ~upto_l;
}

__device__ __forceinline__ void clip_chunk(uint4& v, int64_t chunk_bit0, int64_t lo, int64_t hi) {
  v.x &= valid_bits(chunk_bit0, lo, hi);
  v.y &= valid_bits(chunk_bit0 + 32, lo, hi);
  v.z &= valid_bits(chunk_bit0 + 64, lo, hi);
  v.w &= valid_bits(chunk_bit0 + 96, lo, hi);
}

// Plane popcounts of the packed 2-bit array of `nsamp` entries starting at byte pointer p
// (any alignment), cooperatively by LANES consecutive lanes; `lane` in [0, LANES).
// Every load is an aligned 16-byte load; bytes outside the array are clipped, so the buffer
// must have >= 16 bytes of readable slack on both sides.  Result is this lane's partial sum.
template <int LANES, int UNROLL, bool STREAM>
__device__ __forceinline__ void planes_2bit(const uint8_t* p, uint32_t nsamp, int lane, Planes& c) {
  const uintptr_t addr = reinterpret_cast<uintptr_t>(p);
  const uint32_t shift = static_cast<uint32_t>(addr & 15);
  const uint4* base = reinterpret_cast<const uint4*>(addr - shift);
  const int64_t lo = 8ll * shift;
  const int64_t hi = lo + 2ll * nsamp;
  const uint32_t nchunks = static_cast<uint32_t>((hi + 127) >> 7);
  for (uint32_t c0 = lane; c0 < nchunks; c0 += LANES * UNROLL) {
    uint4 v[UNROLL];
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      const uint32_t cc = c0 + j * LANES;
      if (cc < nchunks) {
        v[j] = STREAM ? ldg_stream(base + cc) : __ldg(base + cc);
      } else {
        v[j] = make_uint4(0, 0, 0, 0);
      }
    }
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      const uint32_t cc = c0 + j * LANES;
      if (cc == 0 || cc + 1 >= nchunks) {
        clip_chunk(v[j], 128ll * cc, lo, hi);
      }
      acc_chunk(v[j], c);
    }
  }
}

// Software-pipelined form for warps that share their SM with compute-heavy warps (K1h): the loads
// of block i+1 are issued before block i is reduced, so a warp always has 32 x UNROLL x 16 bytes
// in flight instead of alternating between waiting and counting.  Whole warp (LANES = 32).
template <int UNROLL>
__device__ __forceinline__ void planes_2bit_pipelined(const uint8_t* p, uint32_t nsamp, int lane, Planes& c) {
  const uintptr_t addr = reinterpret_cast<uintptr_t>(p);
  const uint32_t shift = static_cast<uint32_t>(addr & 15);
  const uint4* base = reinterpret_cast<const uint4*>(addr - shift);
  const int64_t lo = 8ll * shift;
  const int64_t hi = lo + 2ll * nsamp;
  const uint32_t nchunks = static_cast<uint32_t>((hi + 127) >> 7);
  constexpr uint32_t kStep = 32 * UNROLL;
  uint4 cur[UNROLL], nxt[UNROLL];
  auto load = [&](uint4 (&v)[UNROLL], uint32_t c0) {
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      const uint32_t cc = c0 + j * 32;
      v[j] = cc < nchunks ? ldg_stream(base + cc) : make_uint4(0, 0, 0, 0);
    }
  };
  load(cur, lane);
  for (uint32_t c0 = lane; c0 < nchunks; c0 += kStep) {
    load(nxt, c0 + kStep);  // past the end: zeros, no loads
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      const uint32_t cc = c0 + j * 32;
      if (cc == 0 || cc + 1 >= nchunks) clip_chunk(cur[j], 128ll * cc, lo, hi);
      acc_chunk(cur[j], c);
    }
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) cur[j] = nxt[j];
  }
}

// Same with a 2-bit-expanded include mask.  `mask` is the shifted copy whose byte 0 lines up
// with the aligned-down record address (see build_shifted_masks), so clipping is implicit.
template <int LANES, int UNROLL, bool STREAM>
__device__ __forceinline__ void planes_2bit_masked(const uint8_t* p, uint32_t nsamp, const uint8_t* mask_shifts,
                                                   uint32_t mask_stride, int lane, Planes& c) {
  const uintptr_t addr = reinterpret_cast<uintptr_t>(p);
  const uint32_t shift = static_cast<uint32_t>(addr & 15);
  const uint4* base = reinterpret_cast<const uint4*>(addr - shift);
  const uint4* mbase = reinterpret_cast<const uint4*>(mask_shifts + static_cast<size_t>(shift) * mask_stride);
  const uint32_t nchunks = static_cast<uint32_t>((8ll * shift + 2ll * nsamp + 127) >> 7);
  for (uint32_t c0 = lane; c0 < nchunks; c0 += LANES * UNROLL) {
    uint4 v[UNROLL], m[UNROLL];
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      const uint32_t cc = c0 + j * LANES;
      if (cc < nchunks) {
        v[j] = STREAM ? ldg_stream(base + cc) : __ldg(base + cc);
        m[j] = __ldg(mbase + cc);
      } else {
        v[j] = make_uint4(0, 0, 0, 0);
        m[j] = make_uint4(0, 0, 0, 0);
      }
    }
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      acc_pair_masked(v[j].x, v[j].y, m[j].x, m[j].y, c);
      acc_pair_masked(v[j].z, v[j].w, m[j].z, m[j].w, c);
    }
  }
}

// Popcount of an nbits-long bitarray at byte pointer p (any alignment), optionally AND-ed
// with a shifted 1-bit mask copy.  Partial sum for this lane.
template <int LANES>
__device__ __forceinline__ uint32_t popcount_bits(const uint8_t* p, uint32_t nbits, const uint8_t* mask1_shifts,
                                                  uint32_t mask1_stride, int lane) {
  const uintptr_t addr = reinterpret_cast<uintptr_t>(p);
  const uint32_t shift = static_cast<uint32_t>(addr & 15);
  const uint4* base = reinterpret_cast<const uint4*>(addr - shift);
  const uint4* mbase =
      mask1_shifts ? reinterpret_cast<const uint4*>(mask1_shifts + static_cast<size_t>(shift) * mask1_stride) : nullptr;
  const int64_t lo = 8ll * shift;
  const int64_t hi = lo + nbits;
  const uint32_t nchunks = static_cast<uint32_t>((hi + 127) >> 7);
  uint32_t acc = 0;
  for (uint32_t cc = lane; cc < nchunks; cc += LANES) {
    uint4 v = __ldg(base + cc);
    if (mbase) {
      const uint4 m = __ldg(mbase + cc);
      v.x &= m.x;
      v.y &= m.y;
      v.z &= m.z;
      v.w &= m.w;
    } else if (cc == 0 || cc + 1 >= nchunks) {
      clip_chunk(v, 128ll * cc, lo, hi);
    }
    acc += __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);
  }
  return acc;
}

template <int LANES>
__device__ __forceinline__ uint32_t group_sum(uint32_t v) {
#pragma unroll
  for (int off = LANES / 2; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

__device__ __forceinline__ uint4 planes_to_counts(const Planes& c, uint32_t n) {
  uint4 r;
  r.y = c.e - c.b;
  r.z = c.o - c.b;
  r.w = c.b;
  r.x = n - r.y - r.z - r.w;
  return r;
}

// ---- varints and difflists (pgen_spec.tex:354-390) -------------------------------------------

// Base-128 little-endian varint < 2^31 (GetVint31 semantics, plink2_base.h:4259-4280).
__device__ __forceinline__ bool get_vint31(const uint8_t*& p, const uint8_t* end, uint32_t& out) {
  uint32_t val = 0;
#pragma unroll 1
  for (int shift = 0; shift < 35; shift += 7) {
    if (p >= end) return false;
    const uint32_t b = *p++;
    if (shift == 28 && (b & 0xf8u)) return false;  // would reach 2^31 or continue past 5 bytes
    val |= (b & 127u) << shift;
    if (!(b & 128u)) {
      out = val;
      return true;
    }
  }
  return false;
}

__device__ __forceinline__ uint32_t bytes_to_represent(uint32_t n) {
  // BytesToRepresentNzU32 (plink2_base.h:4098): 1 + floor(log2(n)) / 8.  (For N == 256 this is 2,
  // where the spec text says 1; pgenlib's writer and reader agree with each other, so follow it.)
  return 1u + ((31u - __clz(n)) >> 3);
}

__device__ __forceinline__ uint32_t load_le(const uint8_t* p, uint32_t nbytes) {
  uint32_t v = p[0];
  if (nbytes > 1) v |= static_cast<uint32_t>(p[1]) << 8;
  if (nbytes > 2) v |= static_cast<uint32_t>(p[2]) << 16;
  if (nbytes > 3) v |= static_cast<uint32_t>(p[3]) << 24;
  return v;
}

struct Difflist {
  uint32_t len;          // L
  uint32_t ngroups;      // ceil(L / 64)
  uint32_t idbytes;
  const uint8_t* first_ids;
  const uint8_t* group_sizes;  // ngroups - 1 bytes, each (byte length of the group's deltas) - 63
  const uint8_t* raregeno;     // ceil(L/4) bytes, or nullptr for plain ID lists
  const uint8_t* deltas;
  const uint8_t* end;          // end of the enclosing record
};

// Parses the header at p; on success p is left at the start of the delta stream when len > 0,
// or just past the length varint when len == 0.  Mirrors the checks of ParseDifflistHeader
// (pgenlib_read.cc:2184-2218): L <= N / 8 and every component inside the record.
__device__ __forceinline__ bool parse_difflist(const uint8_t*& p, const uint8_t* end, uint32_t N, bool has_geno,
                                               Difflist& d) {
  uint32_t len;
  if (!get_vint31(p, end, len)) return false;
  d.len = len;
  d.ngroups = 0;
  d.end = end;
  d.raregeno = nullptr;
  d.first_ids = d.group_sizes = d.deltas = p;
  d.idbytes = bytes_to_represent(N);
  if (!len) return true;
  if (len > N / 8) return false;
  const uint32_t g = (len + 63) >> 6;
  d.ngroups = g;
  const uint64_t index_bytes = static_cast<uint64_t>(g) * (d.idbytes + 1) - 1;
  if (static_cast<uint64_t>(end - p) < index_bytes) return false;
  d.first_ids = p;
  d.group_sizes = p + static_cast<uint64_t>(g) * d.idbytes;
  p += index_bytes;
  if (has_geno) {
    const uint32_t rg = (len + 3) >> 2;
    if (static_cast<uint64_t>(end - p) < rg) return false;
    d.raregeno = p;
    p += rg;
  }
  d.deltas = p;
  return true;
}

// 4 bytes at any alignment (the two aligned words that cover them; up to 7 bytes past the window are touched, every
// record buffer carries >= 16 B of slack).
__device__ __forceinline__ uint32_t ldu32(const uint8_t* p) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~static_cast<uintptr_t>(3));
  const uint32_t sh = static_cast<uint32_t>(a & 3) << 3;
  const uint32_t lo = __ldg(w);
  return sh ? __funnelshift_r(lo, __ldg(w + 1), sh) : lo;
}

// Warp-cooperative walk: lane g of each batch of 32 groups decodes group g serially and calls
// f(sample_id, geno2bit, entry_ordinal) for every entry.  Group start offsets come from a warp prefix sum of
// the stored group byte lengths, which is what makes the varint stream parallel at group
// granularity.  Returns false (on all lanes) if any lane saw malformed input.
template <class F>
__device__ __forceinline__ bool difflist_walk_lanes(const Difflist& d, uint32_t N, int lane, F f,
                                                    const uint8_t** track_end = nullptr) {
  bool ok = true;
  uint64_t batch_off = 0;
  const uint8_t* my_end = d.deltas;  // meaningful on the lane that walked the last group
  for (uint32_t g0 = 0; g0 < d.ngroups; g0 += 32) {
    const uint32_t g = g0 + lane;
    uint32_t sz = 0;
    if (g + 1 < d.ngroups) sz = static_cast<uint32_t>(d.group_sizes[g]) + 63u;
    uint32_t incl = sz;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += t;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    if (g < d.ngroups) {
      const uint64_t start = batch_off + (incl - sz);
      const uint64_t avail = static_cast<uint64_t>(d.end - d.deltas);
      if (start > avail) {
        ok = false;
      } else {
        const uint8_t* ptr = d.deltas + start;
        const uint8_t* gend = d.end;
        if (g + 1 < d.ngroups && start + sz <= avail) gend = ptr + sz;
        uint32_t id = load_le(d.first_ids + static_cast<uint64_t>(g) * d.idbytes, d.idbytes);
        const uint32_t cnt = min(64u, d.len - (g << 6));
        // Buffered reader: up to 8 bytes of the group's deltas sit in a register and are refilled a word at a time,
        // the following word already requested -- with a byte load per varint byte the dependent load latency was
        // the largest stall of the sparse per-sample kernel (profiles/r02p_ncu_k_densify_sparse.txt: 29 % of samples
        // on the instruction behind that load).
        uint64_t buf = 0;
        uint32_t have = 0, consumed = 0;
        uint32_t remaining = static_cast<uint32_t>(gend - ptr);  // bytes of the group not yet in buf
        const uint8_t* rp = ptr;
        uint32_t nextw = remaining ? ldu32(rp) : 0u;
        // the group's 2-bit codes (16 bytes for a full group), requested up front as well
        uint32_t rg0 = 0, rg1 = 0, rg2 = 0, rg3 = 0;
        if (d.raregeno) {
          const uint8_t* rgp = d.raregeno + (g << 4);
          const uint32_t nbytes = (cnt + 3u) >> 2;
          rg0 = ldu32(rgp);
          if (nbytes > 4u) rg1 = ldu32(rgp + 4);
          if (nbytes > 8u) rg2 = ldu32(rgp + 8);
          if (nbytes > 12u) rg3 = ldu32(rgp + 12);
        }
#pragma unroll 1
        for (uint32_t k = 0; k < cnt; ++k) {
          if (id >= N) {
            ok = false;
            break;
          }
          const uint32_t wsel = k >> 4;
          const uint32_t rw = wsel == 0 ? rg0 : (wsel == 1 ? rg1 : (wsel == 2 ? rg2 : rg3));
          f(id, (rw >> (2 * (k & 15))) & 3u, (g << 6) + k);
          if (k + 1 < cnt) {
            while (have < 5u && remaining) {  // a whole varint (<= 5 bytes) or the rest of the stream
              const uint32_t take = min(4u, remaining);
              const uint32_t w = take < 4u ? (nextw & ((1u << (8u * take)) - 1u)) : nextw;
              buf |= static_cast<uint64_t>(w) << (8u * have);
              have += take;
              rp += take;
              remaining -= take;
              nextw = remaining ? ldu32(rp) : 0u;
            }
            const uint32_t lo = static_cast<uint32_t>(buf);
            uint32_t delta = lo & 127u, used = 1;
            if (lo & 128u) {  // GetVint31: up to 5 bytes, the fifth may carry 3 value bits
              bool done = false;
#pragma unroll
              for (uint32_t i = 1; i < 5; ++i) {
                if (!done) {
                  const uint32_t b = static_cast<uint32_t>(buf >> (8u * i)) & 0xffu;
                  if (i == 4 && (b & 0xf8u)) break;
                  delta |= (b & 127u) << (7u * i);
                  used = i + 1;
                  done = !(b & 128u);
                }
              }
              if (!done) {
                ok = false;
                break;
              }
            }
            if (used > have) {  // the stream ended inside the varint
              ok = false;
              break;
            }
            buf >>= 8u * used;
            have -= used;
            consumed += used;
            id += delta;
          }
        }
        ptr += consumed;
        my_end = ptr;
      }
    }
    batch_off += total;
  }
  if (track_end) {
    // byte after the last group's deltas = end of the whole difflist
    const uint32_t last_lane = d.ngroups ? ((d.ngroups - 1) & 31u) : 0u;
    const unsigned long long e =
        __shfl_sync(0xffffffffu, static_cast<unsigned long long>(reinterpret_cast<uintptr_t>(my_end)), last_lane);
    *track_end = reinterpret_cast<const uint8_t*>(static_cast<uintptr_t>(e));
  }
  return __all_sync(0xffffffffu, ok);
}

__device__ __forceinline__ uint32_t geno_at(const uint8_t* dense, uint32_t id) {
  return (static_cast<uint32_t>(dense[id >> 2]) >> (2 * (id & 3))) & 3u;
}

__device__ __forceinline__ bool bit_at(const uint8_t* bits, uint32_t id) { return (bits[id >> 3] >> (id & 7)) & 1u; }


// ---- unaligned bit-window loads ------------------------------------------------------------------
// 32 bits starting `bitoff` bits after byte pointer `row` (any alignment); reads the two aligned
// 32-bit words that cover the window, so up to 7 bytes past the window may be touched (every
// record buffer carries >= 16 B of slack).
__device__ __forceinline__ uint32_t load_bits32(const uint8_t* row, uint64_t bitoff) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(row) + (bitoff >> 3);
  const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~static_cast<uintptr_t>(3));
  const uint32_t sh = (static_cast<uint32_t>(a & 3) << 3) + static_cast<uint32_t>(bitoff & 7);
  const uint32_t lo = __ldg(w);
  if (sh == 0) return lo;
  return __funnelshift_r(lo, __ldg(w + 1), sh);
}

// Word k (16 genotypes) of a packed 2-bit row at any byte alignment.
__device__ __forceinline__ uint32_t row_word(const uint8_t* row, uint32_t k) {
  return load_bits32(row, static_cast<uint64_t>(k) << 5);
}

// The same walk with the WHOLE WARP on one 64-entry group at a time (bit-parallel varint decode): lane l takes bytes
// 4 l .. 4 l + 3 of the group's delta stream.  A byte with its high bit clear ends a varint; byte i of a varint
// contributes (b & 127) << 7 i, so the id behind the j-th terminator is first_id + (the sum of the contributions of
// all bytes up to it) -- two warp prefix sums (contributions, terminators) replace the 63 serial steps, and the
// position of a byte inside its varint needs only the run of continuation bytes in front of it (from the lanes
// before: two shuffles).  ~130 warp instructions per group against ~340 for lane-per-group on 20-group records
// (profiles/r01f_sparse_walk_experiment.txt).  Same callback, same error behaviour as difflist_walk.
// First 128 stream bytes of a group as one 32-bit word per lane (bytes 4 l .. 4 l + 3 from `ptr`; lanes past `room`
// read nothing and get 0), plus -- lanes 0-3 -- the four 32-bit words of the group's 64 two-bit raregeno codes.
struct SwarChunk {
  uint32_t w;   // stream bytes of this lane, bytes past the group's room cleared
  uint32_t rg;  // raregeno word `lane` of the group (lanes 0-3), else 0
};
__device__ __forceinline__ uint32_t swar_load(const uint8_t* ptr, uint32_t a, uint32_t room) {
  uint32_t w = 0;
  if (a < room) {
    w = load_bits32(ptr + a, 0);
    if (room - a < 4u) w &= (1u << (8u * (room - a))) - 1u;
  }
  return w;
}

// team_rank / team_size: the warps of a CTA share one (long) list -- warp r takes groups r, r + size, ... of every
// 32-group batch (size must divide 32; track_end is only meaningful with size 1).
template <class F>
__device__ __forceinline__ bool difflist_walk_swar(const Difflist& d, uint32_t N, int lane, F f,
                                                   const uint8_t** track_end = nullptr, uint32_t team_rank = 0,
                                                   uint32_t team_size = 1) {
  bool ok = true;
  uint32_t batch_off = 0;  // a record is far below 4 GiB: 32-bit offsets into its delta stream
  const uint8_t* list_end = d.deltas;
  const uint32_t avail = static_cast<uint32_t>(d.end - d.deltas);
  for (uint32_t g0 = 0; g0 < d.ngroups; g0 += 32) {
    const uint32_t gl = g0 + lane;
    uint32_t sz = 0;
    if (gl + 1 < d.ngroups) sz = static_cast<uint32_t>(d.group_sizes[gl]) + 63u;
    uint32_t incl = sz;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += t;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    const uint32_t my_first = gl < d.ngroups ? load_le(d.first_ids + static_cast<uint64_t>(gl) * d.idbytes, d.idbytes) : 0u;
    const uint32_t nb = min(32u, d.ngroups - g0);
    // where group k of the batch starts and how many bytes it may use (warp-uniform)
    auto start_of = [&](uint32_t k) { return batch_off + (__shfl_sync(0xffffffffu, incl, k) - __shfl_sync(0xffffffffu, sz, k)); };
    auto room_of = [&](uint32_t k, uint32_t start) -> uint32_t {
      const uint32_t gsz = __shfl_sync(0xffffffffu, sz, k);
      if (start > avail) return 0u;
      return (g0 + k + 1 != d.ngroups && start + gsz <= avail) ? gsz : avail - start;
    };
    auto load_group = [&](uint32_t k, uint32_t start, uint32_t room) {
      SwarChunk c;
      c.w = swar_load(d.deltas + start, 4u * static_cast<uint32_t>(lane), room);
      c.rg = 0;
      if (d.raregeno && lane < 4) {
        // 16 raregeno bytes per full group, fewer at the end of the list (what is read past them lies inside the
        // record or its slack and is never looked at)
        const uint32_t e0 = (g0 + k) << 6;
        const uint32_t nbytes = (min(64u, d.len - e0) + 3u) >> 2;
        if (4u * lane < nbytes) c.rg = load_bits32(d.raregeno + (e0 >> 2) + 4u * lane, 0);
      }
      return c;
    };
    if (team_rank >= nb) {  // nothing of this batch is mine (warp-uniform)
      batch_off += total;
      continue;
    }
    uint32_t start = start_of(team_rank);
    uint32_t room = room_of(team_rank, start);
    SwarChunk cur = load_group(team_rank, start, room);
#pragma unroll 1
    for (uint32_t k = team_rank; k < nb; k += team_size) {
      const uint32_t g = g0 + k;
      const uint32_t first_id = __shfl_sync(0xffffffffu, my_first, k);
      const bool last = g + 1 == d.ngroups;
      const uint32_t need = min(64u, d.len - (g << 6)) - 1u;  // varints of this group
      const uint32_t e0 = g << 6;
      if (start > avail) {
        ok = false;
        break;
      }
      const uint8_t* ptr = d.deltas + start;
      const uint32_t cur_room = room;
      // the next group's bytes are requested before this group's arithmetic starts
      SwarChunk nxt{0u, 0u};
      uint32_t nstart = 0, nroom = 0;
      if (k + team_size < nb) {
        nstart = start_of(k + team_size);
        nroom = room_of(k + team_size, nstart);
        nxt = load_group(k + team_size, nstart, nroom);
      }
      if (lane == 0) {
        if (first_id >= N) {
          ok = false;
        } else {
          f(first_id, cur.rg & 3u, e0);
        }
      }
      uint32_t c_sum = 0, c_terms = 0, c_run = 0;
      const uint8_t* my_end = ptr;  // (last group) byte behind the terminator that completes the list
      bool found_end = need == 0;
      uint32_t w = cur.w;
#pragma unroll 1
      for (uint32_t c = 0; c_terms < need; c += 128) {
        if (c >= cur_room) {
          ok = false;  // the stream ends before the group's entries do
          break;
        }
        const uint32_t a = c + 4u * static_cast<uint32_t>(lane);
        if (c) w = swar_load(ptr, a, cur_room);
        const uint32_t nvalid = a < cur_room ? min(cur_room - a, 4u) : 0u;
        const uint32_t vm = nvalid >= 4 ? 0x80808080u : (0x80808080u & ((1u << (8u * nvalid)) - 1u));
        const uint32_t term = ~w & vm;
        const uint32_t c0 = (w >> 7) & 1u, c1 = (w >> 15) & 1u, c2 = (w >> 23) & 1u, c3 = w >> 31;
        const uint32_t tr = c3 ? (c2 ? (c1 ? (c0 ? 4u : 3u) : 2u) : 1u) : 0u;
        uint32_t p1 = __shfl_up_sync(0xffffffffu, tr, 1), p2 = __shfl_up_sync(0xffffffffu, tr, 2);
        if (lane == 0) {
          p1 = c_run;
          p2 = 0;
        } else if (lane == 1) {
          p2 = c_run;
        }
        const uint32_t run_in = (lane >= 1 && p1 == 4u) ? p1 + p2 : p1;
        const uint32_t k0 = run_in, k1 = c0 ? k0 + 1u : 0u, k2 = c1 ? k1 + 1u : 0u, k3 = c2 ? k2 + 1u : 0u;
        const uint32_t run_out = c3 ? k3 + 1u : 0u;
        const uint32_t b0 = w & 0x7fu, b1 = (w >> 8) & 0x7fu, b2 = (w >> 16) & 0x7fu, b3 = (w >> 24) & 0x7fu;
        // GetVint31: at most 5 bytes, the fifth may carry 3 value bits.  Fifth bytes are rare: the test is one
        // comparison per lane unless one shows up.
        uint32_t bad = 0;
        uint32_t s0, s1, s2, s3;
        if (max(max(k0, k1), max(k2, k3)) < 4u) {
          s0 = b0 << (7u * k0);
          s1 = s0 + (b1 << (7u * k1));
          s2 = s1 + (b2 << (7u * k2));
          s3 = s2 + (b3 << (7u * k3));
        } else {
          bad = ((k0 > 4u) | (k0 == 4u && ((w & 0xf8u) != 0u))) | (((k1 > 4u) | (k1 == 4u && ((w & 0xf800u) != 0u))) << 1) |
                (((k2 > 4u) | (k2 == 4u && ((w & 0xf80000u) != 0u))) << 2) |
                (((k3 > 4u) | (k3 == 4u && ((w & 0xf8000000u) != 0u))) << 3);
          s0 = b0 << (7u * min(k0, 4u));
          s1 = s0 + (b1 << (7u * min(k1, 4u)));
          s2 = s1 + (b2 << (7u * min(k2, 4u)));
          s3 = s2 + (b3 << (7u * min(k3, 4u)));
        }
        const uint32_t tcount = __popc(term);
        uint32_t isum = s3, icnt = tcount;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
          const uint32_t ts = __shfl_up_sync(0xffffffffu, isum, off);
          const uint32_t tc = __shfl_up_sync(0xffffffffu, icnt, off);
          if (lane >= off) {
            isum += ts;
            icnt += tc;
          }
        }
        const uint32_t sbase = first_id + c_sum + isum - s3, cbase = c_terms + icnt - tcount;
        const uint32_t sums[4] = {s0, s1, s2, s3};
#pragma unroll
        for (uint32_t i = 0; i < 4; ++i) {
          const uint32_t before = cbase + __popc(term & ((1u << (8u * i)) - 1u));  // varints completed before byte i
          const uint32_t ord = before + 1u;
          // the entry's 2-bit code: word ord >> 4 of the group's raregeno sits in lane ord >> 4 (every lane shuffles)
          const uint32_t rgw = __shfl_sync(0xffffffffu, cur.rg, (ord >> 4) & 3u);
          if (i < nvalid && before < need) {
            if ((bad >> i) & 1u) ok = false;
            if ((term >> (8u * i + 7u)) & 1u) {
              const uint32_t id = sbase + sums[i];
              if (id >= N || id < first_id) {
                ok = false;
              } else {
                f(id, (rgw >> (2u * (ord & 15u))) & 3u, e0 + ord);
              }
              if (ord == need) {
                my_end = ptr + a + i + 1u;
                found_end = true;
              }
            }
          }
        }
        c_sum += __shfl_sync(0xffffffffu, isum, 31);
        c_terms += __shfl_sync(0xffffffffu, icnt, 31);
        c_run = __shfl_sync(0xffffffffu, run_out, 31);
      }
      if (last && track_end) {
        const uint32_t who = __ballot_sync(0xffffffffu, found_end && (need == 0 ? lane == 0 : true));
        if (who) {
          const unsigned long long e = __shfl_sync(
              0xffffffffu, static_cast<unsigned long long>(reinterpret_cast<uintptr_t>(my_end)), __ffs(who) - 1);
          list_end = reinterpret_cast<const uint8_t*>(static_cast<uintptr_t>(e));
        }
      }
      if (!__all_sync(0xffffffffu, ok)) break;
      cur = nxt;
      start = nstart;
      room = nroom;
    }
    batch_off += total;
    if (!__all_sync(0xffffffffu, ok)) break;
  }
  if (track_end) *track_end = list_end;
  return __all_sync(0xffffffffu, ok);
}

// The walker every kernel calls.  PD_WALK_LANES (compile-time) selects the round-1 lane-per-group walk for A/B runs.
template <class F>
__device__ __forceinline__ bool difflist_walk(const Difflist& d, uint32_t N, int lane, F f,
                                              const uint8_t** track_end = nullptr) {
#ifdef PD_WALK_LANES
  return difflist_walk_lanes(d, N, lane, f, track_end);
#else
  return difflist_walk_swar(d, N, lane, f, track_end);
#endif
}

// Same through L2 only (rows that other lanes patched with atomics moments ago).
__device__ __forceinline__ uint32_t geno_at_cg(const uint32_t* dense_words, uint32_t id) {
  return (__ldcg(dense_words + (id >> 4)) >> (2 * (id & 15))) & 3u;
}

// ---- full decode of one record into a dense, word-aligned 2-bit genovec --------------------------
// dst: ceil(N/16) 32-bit words (global), warp-cooperative.  Types 0, 1, 4, 5, 6, 7
// (ParseNonLdGenovecSubsetUnsafe, pgenlib_read.cc:2711-2741, without the subsetting).  Difflist
// entries are patched with atomicXor of (background ^ new) so that lanes touching the same word
// cannot lose updates.  Returns false on malformed input (uniform across the warp).
__device__ __forceinline__ uint32_t spread16(uint32_t x) {
  x = (x | (x << 8)) & 0x00FF00FFu;
  x = (x | (x << 4)) & 0x0F0F0F0Fu;
  x = (x | (x << 2)) & 0x33333333u;
  x = (x | (x << 1)) & 0x55555555u;
  return x;
}

// CTA = false: one warp decodes the record.  CTA = true: the `size` warps of a CTA decode it together (every thread of
// the CTA must call, with rank = its warp): the fill is strided over the warps and a long difflist is split by
// groups -- a one-bit record with a 12 k-entry list is ~200 us of serial latency for a single warp.
template <bool CTA>
__device__ __forceinline__ bool decode_nonld_dense_team(uint32_t main, const uint8_t* p, const uint8_t* end, uint32_t N,
                                                        uint32_t* dst, int lane, uint32_t rank, uint32_t size) {
  const uint32_t nwords = (N + 15) >> 4;
  const uint32_t geno_bytes = (N + 3) >> 2;
  const uint32_t first = CTA ? rank * 32u + static_cast<uint32_t>(lane) : static_cast<uint32_t>(lane);
  const uint32_t step = CTA ? size * 32u : 32u;
  auto sync = [] {
    if (CTA) {
      __syncthreads();
    } else {
      __syncwarp();
    }
  };
  if (main == 0) {
    if (static_cast<uint64_t>(end - p) < geno_bytes) return false;
    for (uint32_t w = first; w < nwords; w += step) dst[w] = row_word(p, w);
    sync();
    return true;
  }
  if (main == 5) {
    for (uint32_t w = first; w < nwords; w += step) dst[w] = 0u;
    sync();
    return true;
  }
  if (main == 1) {
    const uint32_t bit_bytes = (N + 7) >> 3;
    if (static_cast<uint64_t>(end - p) < static_cast<uint64_t>(bit_bytes) + 1) return false;
    const uint32_t common2 = p[0];
    const uint32_t low = common2 >> 2;
    const uint32_t delta = common2 & 3u;
    if (low + delta > 3u) return false;
    const uint8_t* bits = p + 1;
    for (uint32_t w = first; w < nwords; w += step) {
      const uint32_t b16 = load_bits32(bits, static_cast<uint64_t>(w) << 4) & 0xffffu;
      dst[w] = low * kM5 + spread16(b16) * delta;
    }
    sync();
    const uint8_t* q = bits + bit_bytes;
    Difflist d;
    if (!parse_difflist(q, end, N, true, d)) return false;
    bool ok = true;
    if (d.len) {
      auto patch = [&](uint32_t id, uint32_t g, uint32_t) {
        const uint32_t orig = low + (bit_at(bits, id) ? delta : 0u);
        atomicXor(&dst[id >> 4], (orig ^ g) << (2 * (id & 15)));
      };
      ok = CTA ? difflist_walk_swar(d, N, lane, patch, nullptr, rank, size) : difflist_walk(d, N, lane, patch);
    }
    sync();
    return ok;
  }
  if (main == 4 || main == 6 || main == 7) {
    const uint32_t common = main & 3u;
    for (uint32_t w = first; w < nwords; w += step) dst[w] = common * kM5;
    sync();
    const uint8_t* q = p;
    Difflist d;
    if (!parse_difflist(q, end, N, true, d)) return false;
    bool ok = true;
    if (d.len) {
      auto patch = [&](uint32_t id, uint32_t g, uint32_t) {
        atomicXor(&dst[id >> 4], (common ^ g) << (2 * (id & 15)));
      };
      ok = CTA ? difflist_walk_swar(d, N, lane, patch, nullptr, rank, size) : difflist_walk(d, N, lane, patch);
    }
    sync();
    return ok;
  }
  return false;
}

__device__ __forceinline__ bool decode_nonld_dense(uint32_t main, const uint8_t* p, const uint8_t* end, uint32_t N,
                                                   uint32_t* dst, int lane) {
  return decode_nonld_dense_team<false>(main, p, end, N, dst, lane, 0, 1);
}

// LD-compressed record (type 2 / 3) on top of an already decoded copy of its base in dst:
// patch the difflist entries in, then swap categories 0 and 2 for type 3
// (ReadGenovecSubsetUnsafe, pgenlib_read.cc:2863-2869; GenovecInvertUnsafe, pgenlib_misc.h:303).
__device__ __forceinline__ bool apply_ld_record(uint32_t main, const uint8_t* p, const uint8_t* end, uint32_t N,
                                                uint32_t* dst, int lane) {
  const uint32_t nwords = (N + 15) >> 4;
  const uint8_t* q = p;
  Difflist d;
  if (!parse_difflist(q, end, N, true, d)) return false;
  bool ok = true;
  if (d.len) {
    __threadfence_block();
    ok = difflist_walk(d, N, lane, [&](uint32_t id, uint32_t g, uint32_t) {
      const uint32_t old = geno_at_cg(dst, id);  // each sample id occurs once per difflist
      atomicXor(&dst[id >> 4], (old ^ g) << (2 * (id & 15)));
    });
  }
  __syncwarp();
  if (main == 3) {
    for (uint32_t w = lane; w < nwords; w += 32) {
      const uint32_t x = __ldcg(dst + w);
      dst[w] = x ^ ((~x & kM5) << 1);
    }
    __syncwarp();
  }
  return ok;
}

}  // namespace pdk
