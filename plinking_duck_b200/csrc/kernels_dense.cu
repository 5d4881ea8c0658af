// K3b/K4/K5: dense 2-bit rows for a batch of variants, and the consumers that read them
// sample-wise: int8 expansion (read_pgen / read_pfile) and per-sample het / hom-alt / missing
// accumulators (plink_missing sample mode, read_pfile orient := 'sample').
//
// Reference code these replace:
//   * PgrGet -> ReadGenovecSubsetUnsafe (pgenlib_read.cc:2847-2912) for every vrtype, incl. the
//     LD cache (LdLoadAndCopyGenovecSubset :2808-2838): k_densify
//   * GenoarrToBytesMinus9 (pgenlib_ffi_support.cc:25-43) + CopyNyparrNonemptySubset: k_expand_i8
//   * PgrGetMissingness + the ctz walk (src/plink_missing.cpp:593-609) and the dense Phase-1
//     loop of src/pfile_reader.cpp:3322-3428: k_sample_acc_v
#include <type_traits>

#include "device_common.cuh"
#include "kernels.cuh"

namespace pdk {

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ void report_error(int* err, int code, uint32_t vidx) {
  if (atomicCAS(&err[0], 0, code) == 0) err[1] = static_cast<int>(vidx);
}

__device__ __forceinline__ const uint8_t* row_ptr(const DenseRows& rows, uint32_t r) {
  return rows.ptrs ? rows.ptrs[r] : rows.base + static_cast<uint64_t>(r) * rows.stride;
}

// One difflist record (type 4 / 6 / 7, or the all-common type 5) added to the per-sample plane counters
// without expanding it: every listed sample gets planes(entry) - planes(common), the common planes
// go to base[] once.  Unsigned wrap-around makes the negative deltas exact: the final per-sample
// value base + acc is a true count >= 0.
__device__ __forceinline__ bool sparse_accumulate(uint32_t common, bool all_common, const uint8_t* p, const uint8_t* end,
                                                  uint32_t N, int lane, const SparseAcc& sa) {
  const uint32_t ce = common & 1u, co = common >> 1, cb = ce & co;
  if (lane == 0) {
    if ((sa.kinds & 1u) && ce) atomicAdd(sa.base + 0, 1u);
    if ((sa.kinds & 2u) && co) atomicAdd(sa.base + 1, 1u);
    if ((sa.kinds & 4u) && cb) atomicAdd(sa.base + 2, 1u);
  }
  if (all_common) return true;
  const uint8_t* q = p;
  Difflist d;
  if (!parse_difflist(q, end, N, true, d)) return false;
  if (!d.len) return true;
  const uint32_t kinds = sa.kinds;
  // Lane-per-group walk on purpose: every alternative measured this round was slower or equal here -- the
  // warp-per-group walker (+28 % time), a buffered stream reader (-2 %), the record staged in shared memory (+3 %),
  // per-CTA shared-memory counters for the missing plane (+15 %) -- while more resident warps helped (-10 %):
  // profiles/r02n_compressed_walk_ab.txt has the numbers and what they rule out.
  return difflist_walk_lanes(d, N, lane, [&](uint32_t id, uint32_t g, uint32_t) {
    const uint32_t ge = g & 1u, go = g >> 1, gb = ge & go;
    if ((kinds & 1u) && ge != ce) atomicAdd(sa.acc_e + id, ge - ce);
    if ((kinds & 2u) && go != co) atomicAdd(sa.acc_o + id, go - co);
    if ((kinds & 4u) && gb != cb) atomicAdd(sa.acc_b + id, gb - cb);
  });
}

// ---------------------------------------------------------------------------------------------
// k_densify: one warp per row (claimed from a counter: records differ in size by 1000x).
// Row r is variant vlist[r] (or v0 + r).  Type-0 records are used in place; every other type is
// decoded into scratch slot r.  LD records decode their base into their own slot first, so rows
// are independent of each other and of batch boundaries.
// ---------------------------------------------------------------------------------------------
template <bool SPARSE>
// 4 CTAs per SM (64 registers, a few spilled words): the kernel is latency-bound per row -- one warp per record --
// so resident warps are what it needs (3 CTAs: 4.26 ms, 4: 3.82 ms, 5 with 48 registers: no better;
// pd_missing_by_sample on the configs[3] fileset)
__global__ void __launch_bounds__(kThreads, 4)
k_densify(RecIndex ix, const uint32_t* __restrict__ vlist, uint32_t v0, uint32_t nrows, uint8_t* __restrict__ scratch,
          uint32_t slot_stride, const uint8_t** __restrict__ ptrs, uint32_t* __restrict__ work_counter,
          int* __restrict__ err, SparseAcc sa, LdScratch bases) {
  const int lane = threadIdx.x & 31;
  const uint32_t N = ix.N;
  const bool sparse = SPARSE && sa.dense_count != nullptr;
  for (;;) {
    uint32_t r = 0;
    if (lane == 0) r = atomicAdd(work_counter, 1u);
    r = __shfl_sync(0xffffffffu, r, 0);
    if (r >= nrows) return;
    const uint32_t v = vlist ? vlist[r] : v0 + r;
    if (!ix.fpos) {  // fixed-width file: every main track is a raw 2-bit array, used in place
      if (lane == 0) ptrs[r] = ix.data + (ix.fixed_start + static_cast<uint64_t>(v) * ix.fixed_width - ix.data_fpos);
      continue;
    }
    const uint32_t main = ix.vrtypes[v] & 7u;
    const uint8_t* p = ix.data + (ix.fpos[v] - ix.data_fpos);
    const uint8_t* end = ix.data + (ix.fpos[v + 1] - ix.data_fpos);
    if (SPARSE && sparse && main >= 4) {
      const bool ok = sparse_accumulate(main == 5 ? 0u : (main & 3u), main == 5, p, end, N, lane, sa);
      if (!ok && lane == 0) report_error(err, kErrMalformed, v);
      continue;
    }
    uint32_t slot_idx = r;
    if (SPARSE && sparse) {  // decoded rows are packed: slot and pointer index come from the counter
      if (lane == 0) slot_idx = atomicAdd(sa.dense_count, 1u);
      slot_idx = __shfl_sync(0xffffffffu, slot_idx, 0);
    }
    uint32_t* slot = reinterpret_cast<uint32_t*>(scratch + static_cast<uint64_t>(slot_idx) * slot_stride);
    bool ok = true;
    const uint8_t* out = reinterpret_cast<const uint8_t*>(slot);
    if (main == 0) {
      ok = static_cast<uint64_t>(end - p) >= ((N + 3) >> 2);
      out = p;
    } else if (main == 2 || main == 3) {
      const uint32_t b = ix.ldbase[v];
      const uint32_t bmain = ix.vrtypes[b] & 7u;
      const uint8_t* bp = ix.data + (ix.fpos[b] - ix.data_fpos);
      const uint8_t* bend = ix.data + (ix.fpos[b + 1] - ix.data_fpos);
      ok = (b < v) && bmain != 2 && bmain != 3;
      const uint32_t bs = ok && bases.slot_ct ? ix.base_slot[v] - bases.slot_lo : 0xffffffffu;
      if (ok && bs < bases.slot_ct) {
        // the base was decoded once for the whole batch (k_decode_bases): copy it
        const uint4* src = reinterpret_cast<const uint4*>(bases.dense + static_cast<uint64_t>(bs) * bases.stride);
        uint4* dst4 = reinterpret_cast<uint4*>(slot);
        const uint32_t n16 = (((N + 3) >> 2) + 15) >> 4;
#pragma unroll 4
        for (uint32_t i = lane; i < n16; i += 32) dst4[i] = __ldcs(src + i);  // four loads in flight per lane
        __syncwarp();
      } else if (ok) {
        ok = decode_nonld_dense(bmain, bp, bend, N, slot, lane);
      }
      if (ok) ok = apply_ld_record(main, p, end, N, slot, lane);
    } else {
      ok = decode_nonld_dense(main, p, end, N, slot, lane);
    }
    if (lane == 0) {
      if (!ok) report_error(err, kErrMalformed, v);
      ptrs[slot_idx] = out;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_densify_team: one CTA (8 warps) per row, the row built in SHARED memory and written out once.
// k_densify patches a row in global memory with one atomicXor per difflist entry (~1.4 k per 100k-sample rare-variant
// record: the atomics, not the walk, set its pace -- profiles/r02n_compressed_walk_ab.txt).  Here the warps of a CTA
// fill the row in shared memory (common value / expanded bit array / the LD base copied from the batch's base
// scratch), split the difflist by groups (difflist_walk_swar with team_rank / team_size) and patch with shared-memory
// atomics; the finished row leaves with coalesced 128-bit stores.  Used for dense decode (no SparseAcc) when a row
// fits the default dynamic shared memory.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kTeamMaxSmem = 48u << 10;

__global__ void __launch_bounds__(kThreads)
k_densify_team(RecIndex ix, const uint32_t* __restrict__ vlist, uint32_t v0, uint32_t nrows, uint8_t* __restrict__ scratch,
               uint32_t slot_stride, const uint8_t** __restrict__ ptrs, int* __restrict__ err, LdScratch bases) {
  extern __shared__ __align__(16) uint32_t s_row[];
  const int lane = threadIdx.x & 31;
  const uint32_t wid = threadIdx.x >> 5;
  constexpr uint32_t kTeam = kThreads / 32;
  const uint32_t N = ix.N;
  const uint32_t nwords = (N + 15) >> 4;
  const uint32_t n16 = (((N + 3) >> 2) + 15) >> 4;  // 16-byte chunks of a row
  const uint32_t r = blockIdx.x;
  if (r >= nrows) return;
  const uint32_t v = vlist ? vlist[r] : v0 + r;
  const uint32_t main = ix.vrtypes[v] & 7u;
  const uint8_t* p = ix.data + (ix.fpos[v] - ix.data_fpos);
  const uint8_t* end = ix.data + (ix.fpos[v + 1] - ix.data_fpos);
  if (main == 0) {  // used in place
    if (threadIdx.x == 0) {
      if (static_cast<uint64_t>(end - p) < ((N + 3) >> 2)) report_error(err, kErrMalformed, v);
      ptrs[r] = p;
    }
    return;
  }
  for (uint32_t w = nwords + threadIdx.x; w < n16 * 4; w += kThreads) s_row[w] = 0u;  // the padding of the last chunk
  bool ok = true;
  if (main == 2 || main == 3) {
    const uint32_t b = ix.ldbase[v];
    const uint32_t bmain = ix.vrtypes[b] & 7u;
    ok = (b < v) && bmain != 2 && bmain != 3;  // uniform for the CTA
    if (ok) {
      const uint32_t bs = bases.slot_ct ? ix.base_slot[v] - bases.slot_lo : 0xffffffffu;
      if (bs < bases.slot_ct) {
        const uint4* src = reinterpret_cast<const uint4*>(bases.dense + static_cast<uint64_t>(bs) * bases.stride);
        uint4* dst4 = reinterpret_cast<uint4*>(s_row);
        for (uint32_t i = threadIdx.x; i < n16; i += kThreads) dst4[i] = __ldcs(src + i);
        __syncthreads();
      } else {
        const uint8_t* bp = ix.data + (ix.fpos[b] - ix.data_fpos);
        const uint8_t* bend = ix.data + (ix.fpos[b + 1] - ix.data_fpos);
        ok = decode_nonld_dense_team<true>(bmain, bp, bend, N, s_row, lane, wid, kTeam);
        ok = __syncthreads_and(ok) != 0;  // a warp's walk may fail alone: every barrier below needs a uniform branch
      }
    }
    if (ok) {
      const uint8_t* q = p;
      Difflist d;
      ok = parse_difflist(q, end, N, true, d);
      if (ok && d.len) {
        ok = difflist_walk_swar(d, N, lane, [&](uint32_t id, uint32_t g, uint32_t) {
          const uint32_t old = (reinterpret_cast<volatile uint32_t*>(s_row)[id >> 4] >> (2 * (id & 15))) & 3u;
          atomicXor(&s_row[id >> 4], (old ^ g) << (2 * (id & 15)));  // each sample id occurs once per difflist
        }, nullptr, wid, kTeam);
      }
      ok = __syncthreads_and(ok) != 0;
      if (ok && main == 3) {  // GenovecInvertUnsafe: 0 <-> 2
        for (uint32_t w = threadIdx.x; w < nwords; w += kThreads) {
          const uint32_t x = s_row[w];
          s_row[w] = x ^ ((~x & 0x55555555u) << 1);
        }
      }
    }
  } else {
    ok = decode_nonld_dense_team<true>(main, p, end, N, s_row, lane, wid, kTeam);
  }
  ok = __syncthreads_and(ok) != 0;
  uint8_t* slot = scratch + static_cast<uint64_t>(r) * slot_stride;
  if (ok) {
    uint4* out4 = reinterpret_cast<uint4*>(slot);
    const uint4* row4 = reinterpret_cast<const uint4*>(s_row);
    for (uint32_t i = threadIdx.x; i < n16; i += kThreads) out4[i] = row4[i];
  }
  if (threadIdx.x == 0) {
    if (!ok) report_error(err, kErrMalformed, v);
    ptrs[r] = slot;
  }
}

// ---------------------------------------------------------------------------------------------
// k_expand_i8: blockIdx.y = row (no division), threads of a CTA stride over the row's 16-byte output
// chunks (one 128-bit streaming store each; consecutive lanes = consecutive chunks).  The output is a
// flat (nrows x n_out) byte matrix, so a row starts at any byte: its first chunk boundary is found
// once per row, the few bytes before it and after the last full chunk are written singly.  A chunk
// is 16 samples = one 32-bit window of the record at any 2-bit offset (two aligned loads + funnel
// shift); 2 bits -> {0, 1, 2, -9} is a PRMT byte look-up: the codes are spread to one per nibble and
// used as the selector into the constant 0xF7020100.  Subset gathers take the per-byte path.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t spread2to4(uint32_t x) {  // 16 bits -> 8 nibbles holding one 2-bit code each
  x = (x | (x << 8)) & 0x00FF00FFu;
  x = (x | (x << 4)) & 0x0F0F0F0Fu;
  x = (x | (x << 2)) & 0x33333333u;
  return x;
}

__device__ __forceinline__ uint4 expand16(uint32_t x, uint32_t kTable = 0xF7020100u) {  // bytes 0, 1, 2, -9
  const uint32_t lo = spread2to4(x & 0xffffu), hi = spread2to4(x >> 16);
  uint4 o;
  o.x = __byte_perm(kTable, 0u, lo);
  o.y = __byte_perm(kTable, 0u, lo >> 16);
  o.z = __byte_perm(kTable, 0u, hi);
  o.w = __byte_perm(kTable, 0u, hi >> 16);
  return o;
}

__global__ void __launch_bounds__(kThreads)
k_expand_i8(DenseRows rows, const uint32_t* __restrict__ subset_idx, uint32_t n_out, int8_t* __restrict__ out) {
  for (uint32_t r = blockIdx.y; r < rows.nrows; r += gridDim.y) {
    const uint8_t* row = row_ptr(rows, r);
    int8_t* orow = out + static_cast<uint64_t>(r) * n_out;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t nthr = gridDim.x * blockDim.x;
    if (subset_idx) {
      for (uint32_t j = tid; j < n_out; j += nthr) {
        const uint32_t g = geno_at(row, subset_idx[j]);
        orow[j] = static_cast<int8_t>(g == 3 ? -9 : static_cast<int>(g));
      }
      continue;
    }
    uint32_t head = static_cast<uint32_t>((16u - (reinterpret_cast<uintptr_t>(orow) & 15u)) & 15u);
    if (head > n_out) head = n_out;
    const uint32_t nchunks = (n_out - head) >> 4;
    const uint32_t tail0 = head + (nchunks << 4);
    for (uint32_t k = tid; k < nchunks; k += nthr) {
      const uint32_t col = head + (k << 4);
      const uint32_t x = load_bits32(row, static_cast<uint64_t>(col) << 1);
      __stcs(reinterpret_cast<uint4*>(orow + col), expand16(x));
    }
    // ragged ends: < 16 bytes each
    if (tid < head) {
      const uint32_t g = geno_at(row, tid);
      orow[tid] = static_cast<int8_t>(g == 3 ? -9 : static_cast<int>(g));
    }
    if (tid < n_out - tail0) {
      const uint32_t g = geno_at(row, tail0 + tid);
      orow[tail0 + tid] = static_cast<int8_t>(g == 3 ? -9 : static_cast<int>(g));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// k_expand_i8_valid: the genotype column in DuckDB's child-vector layout (the fill loops of
// src/pgen_reader.cpp:1008-1046 and src/pfile_reader.cpp's variant-orient emit): data byte = the call
// where it is valid and 0 where it is not, validity bit (row * n_out + sample) = 1 where valid.  A call is
// valid when it is not missing and GenotypeFilter::AllowsCall (src/include/plink_common.hpp:471-477)
// selects its category: call_mask bit g.  Same chunking as k_expand_i8; the 16 validity bits of a chunk
// are merged across the warp (512 consecutive bits) into 32-bit words: the 15 words that lie wholly inside
// the warp's bit range are stored, the two at its ends -- shared with the neighbouring warp or row -- are
// OR-ed atomically into the zero-filled bitmap.  bitbase = bit position of the batch's first call in vbits.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t valid16(uint32_t x, uint32_t call_mask) {
  const uint32_t lo = x & 0x55555555u, hi = (x >> 1) & 0x55555555u;
  uint32_t v = 0;
  if (call_mask & 1u) v |= ~(lo | hi) & 0x55555555u;
  if (call_mask & 2u) v |= lo & ~hi;
  if (call_mask & 4u) v |= hi & ~lo;
  v = (v | (v >> 1)) & 0x33333333u;
  v = (v | (v >> 2)) & 0x0F0F0F0Fu;
  v = (v | (v >> 4)) & 0x00FF00FFu;
  v = (v | (v >> 8)) & 0x0000FFFFu;
  return v;
}

__device__ __forceinline__ void or_bits(uint32_t* vbits, uint64_t bitpos, uint32_t bits) {
  if (!bits) return;
  const uint32_t sh = static_cast<uint32_t>(bitpos & 31u);
  uint32_t* w = vbits + (bitpos >> 5);
  atomicOr(w, bits << sh);
  if (sh && (bits >> (32u - sh))) atomicOr(w + 1, bits >> (32u - sh));
}

__global__ void __launch_bounds__(kThreads)
k_expand_i8_valid(DenseRows rows, const uint32_t* __restrict__ subset_idx, uint32_t n_out, uint32_t call_mask,
                  uint32_t table, int8_t* __restrict__ out, uint32_t* __restrict__ vbits, uint32_t bitbase) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t nthr = gridDim.x * blockDim.x;
  for (uint32_t r = blockIdx.y; r < rows.nrows; r += gridDim.y) {
    const uint8_t* row = row_ptr(rows, r);
    int8_t* orow = out + static_cast<uint64_t>(r) * n_out;
    const uint64_t p0 = bitbase + static_cast<uint64_t>(r) * n_out;  // bit of the row's first call
    if (subset_idx) {
      for (uint32_t j0 = tid - lane; j0 < n_out; j0 += nthr) {
        const uint32_t j = j0 + lane;
        bool ok = false;
        if (j < n_out) {
          const uint32_t g = geno_at(row, subset_idx[j]);
          ok = (call_mask >> g) & 1u;
          orow[j] = static_cast<int8_t>(ok ? g : 0u);
        }
        const uint32_t b = __ballot_sync(0xffffffffu, ok);
        if (lane == 0) or_bits(vbits, p0 + j0, b);
      }
      continue;
    }
    uint32_t head = static_cast<uint32_t>((16u - (reinterpret_cast<uintptr_t>(orow) & 15u)) & 15u);
    if (head > n_out) head = n_out;
    const uint32_t nchunks = (n_out - head) >> 4;
    const uint32_t tail0 = head + (nchunks << 4);
    for (uint32_t k0 = tid - lane; k0 < nchunks; k0 += nthr) {
      const uint32_t k = k0 + lane;
      uint32_t m = 0;
      if (k < nchunks) {
        const uint32_t col = head + (k << 4);
        const uint32_t x = load_bits32(row, static_cast<uint64_t>(col) << 1);
        __stcs(reinterpret_cast<uint4*>(orow + col), expand16(x, table));
        m = valid16(x, call_mask);
      }
      const uint32_t other = __shfl_xor_sync(0xffffffffu, m, 1);
      const uint32_t m32 = (lane & 1u) ? 0u : (m | (other << 16));  // chunks k and k + 1
      const uint32_t prev = __shfl_up_sync(0xffffffffu, m32, 2);
      const uint64_t P = p0 + head + (static_cast<uint64_t>(k0) << 4);  // bit of the warp's first chunk
      const uint32_t a = static_cast<uint32_t>(P & 31u);
      const bool full = k0 + 32u <= nchunks;
      if (!(lane & 1u)) {
        uint32_t* w = vbits + (P >> 5) + (lane >> 1);
        const uint32_t hi_part = a ? m32 >> (32u - a) : 0u;
        if (full && lane >= 2) {
          *w = (m32 << a) | (a ? prev >> (32u - a) : 0u);
          if (lane == 30 && hi_part) atomicOr(w + 1, hi_part);
        } else {
          if (m32 << a) atomicOr(w, m32 << a);
          if (!full && hi_part) atomicOr(w + 1, hi_part);
        }
      }
    }
    if (tid < head) {
      const uint32_t g = geno_at(row, tid);
      const bool ok = (call_mask >> g) & 1u;
      orow[tid] = static_cast<int8_t>(ok ? g : 0u);
      or_bits(vbits, p0 + tid, ok);
    }
    if (tid < n_out - tail0) {
      const uint32_t g = geno_at(row, tail0 + tid);
      const bool ok = (call_mask >> g) & 1u;
      orow[tail0 + tid] = static_cast<int8_t>(ok ? g : 0u);
      or_bits(vbits, p0 + tail0 + tid, ok);
    }
  }
}

// VEC consecutive 32-bit words (16 samples each) of a row starting at word w0.  ALIGNED (decided
// on the host for the whole launch): one vector load.  Otherwise VEC + 1 aligned words
// funnel-shifted into place.  Either way straight-line code, so the loads of several rows can be
// in flight together.
template <int VEC, bool ALIGNED>
__device__ __forceinline__ void load_words(const uint8_t* row, uint32_t w0, uint32_t (&x)[VEC]) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(row) + 4ull * w0;
  if (ALIGNED) {
    if (VEC == 4) {
      const uint4 v = ldg_stream(reinterpret_cast<const uint4*>(a));
      x[0] = v.x; x[1 % VEC] = v.y; x[2 % VEC] = v.z; x[3 % VEC] = v.w;
    } else if (VEC == 2) {
      const uint2 v = __ldg(reinterpret_cast<const uint2*>(a));
      x[0] = v.x; x[1 % VEC] = v.y;
    } else {
      x[0] = __ldg(reinterpret_cast<const uint32_t*>(a));
    }
  } else {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~static_cast<uintptr_t>(3));
    const uint32_t sh = static_cast<uint32_t>(a & 3) << 3;
    uint32_t t[VEC + 1];
#pragma unroll
    for (int k = 0; k <= VEC; ++k) t[k] = __ldg(w + k);
#pragma unroll
    for (int k = 0; k < VEC; ++k) x[k] = __funnelshift_r(t[k], t[k + 1], sh);  // sh == 0 -> t[k]
  }
}

// ---------------------------------------------------------------------------------------------
// k_sample_acc_v: transposed pass with VERTICAL counters and no global atomics in the steady
// state.  A thread owns 2 consecutive 32-bit words of a record (32 samples).  The E / O / B
// planes of the two words are interleaved into full 32-bit words (bit 2j = sample j of word 0,
// bit 2j+1 = sample j of word 1), so every bit of a register is a sample, and rows are added with
// carry-save adders (Harley-Seal, 16 rows per block): 15 CSAs (two LOP3 each: majority and 3-input
// xor) produce a weight-16 carry that ripples into four more bit planes -- 2.4 logic ops per row
// and plane.  Eight bit planes hold up to 255 rows; every 240 rows they are turned into per-sample
// bytes (nibble * 0x00204081 spreads 4 bits into 4 bytes) and added to thread-private 16-bit
// counters in shared memory; only at the end of the chunk (<= 65535 rows) each sample costs one
// atomicAdd.  (The first version kept widening 2/4/8-bit counters and flushed them to global memory every
// 255 rows: at 100k samples that is 0.4 atomics per record byte, which -- not the arithmetic -- bounded it
// at 0.63 / 0.37 of the HBM peak; profiles/r01c_kernel_table_200kx100k.jsonl.)
// The row pointer is carried, not recomputed, and full 16-row blocks run without bounds checks:
// the loop is bound by the integer pipe, so every instruction per row counts.
// MODE 0: rows 8-byte aligned (one 64-bit load); 1: 4-byte aligned (fixed-width files: records start
// 12 bytes into the file; two 32-bit loads); 3: any alignment (three loads + funnel shifts);
// 2: row pointer table, any alignment.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t lop3_maj(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ uint32_t lop3_xor(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ void csa(uint32_t& hi, uint32_t& lo, uint32_t a, uint32_t b, uint32_t c) {
  hi = lop3_maj(a, b, c);
  lo = lop3_xor(a, b, c);
}

struct VPlane {
  uint32_t ones, twos, fours, eights;
  uint32_t hi[4];  // weights 16 .. 128
  __device__ __forceinline__ void clear() {
    ones = twos = fours = eights = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) hi[i] = 0;
  }
  // 8 rows in, carry of weight 8 out
  __device__ __forceinline__ uint32_t add8(const uint32_t (&r)[8]) {
    uint32_t ta, tb, fa, fb, carry;
    csa(ta, ones, ones, r[0], r[1]);
    csa(tb, ones, ones, r[2], r[3]);
    csa(fa, twos, twos, ta, tb);
    csa(ta, ones, ones, r[4], r[5]);
    csa(tb, ones, ones, r[6], r[7]);
    csa(fb, twos, twos, ta, tb);
    csa(carry, fours, fours, fa, fb);
    return carry;
  }
  __device__ __forceinline__ void add_carries(uint32_t ea, uint32_t eb) {
    uint32_t carry;
    csa(carry, eights, eights, ea, eb);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t t = hi[i] & carry;
      hi[i] ^= carry;
      carry = t;
    }
  }
  // slot m of this thread (stride 128 words) holds bits 2m (low half) and 2m+1 (high half)
  __device__ __forceinline__ void to_shared(uint32_t* __restrict__ mine) {
    if ((ones | twos | fours | eights | hi[0] | hi[1] | hi[2] | hi[3]) != 0u) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        uint32_t v = (((ones >> (4 * k)) & 0xFu) * 0x00204081u) & 0x01010101u;
        v |= ((((twos >> (4 * k)) & 0xFu) * 0x00204081u) & 0x01010101u) << 1;
        v |= ((((fours >> (4 * k)) & 0xFu) * 0x00204081u) & 0x01010101u) << 2;
        v |= ((((eights >> (4 * k)) & 0xFu) * 0x00204081u) & 0x01010101u) << 3;
#pragma unroll
        for (int i = 0; i < 4; ++i) v |= ((((hi[i] >> (4 * k)) & 0xFu) * 0x00204081u) & 0x01010101u) << (4 + i);
        mine[(2 * k) * 128] += __byte_perm(v, 0u, 0x4140);
        mine[(2 * k + 1) * 128] += __byte_perm(v, 0u, 0x4342);
      }
    }
    clear();
  }
};

template <int MODE>
__device__ __forceinline__ void load_pair(const uint8_t* p, uint32_t (&x)[2]) {
  if (MODE == 0) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
    x[0] = v.x;
    x[1] = v.y;
  } else if (MODE == 1) {
    x[0] = __ldg(reinterpret_cast<const uint32_t*>(p));
    x[1] = __ldg(reinterpret_cast<const uint32_t*>(p) + 1);
  } else {
    load_words<2, false>(p, 0, x);
  }
}

template <int KINDS, int MODE>
__global__ void __launch_bounds__(128, 6)
k_sample_acc_v(DenseRows rows, uint32_t rows_per_chunk, uint32_t* __restrict__ acc_het, uint32_t* __restrict__ acc_hom,
               uint32_t* __restrict__ acc_mis) {
  constexpr int NP = ((KINDS & 1) ? 1 : 0) + ((KINDS & 2) ? 1 : 0) + ((KINDS & 4) ? 1 : 0);
  __shared__ uint32_t sm[NP * 16 * 128];
  const uint32_t N = rows.N;
  const uint32_t ncols = (N + 15) >> 4;
  const uint32_t w0 = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
  uint32_t nrows = rows.nrows;
  if (rows.nrows_dev) {  // packed rows of a sparse batch: spread what there is over the chunks of the grid
    nrows = min(nrows, __ldg(rows.nrows_dev));
    rows_per_chunk = ((nrows + gridDim.y - 1) / gridDim.y + 15) / 16 * 16;
    if (rows_per_chunk < 32) rows_per_chunk = 32;
  }
  const uint32_t r0 = blockIdx.y * rows_per_chunk;
  if (w0 >= ncols || r0 >= nrows) return;
  const uint32_t r1 = min(nrows, r0 + rows_per_chunk);
#pragma unroll
  for (int i = 0; i < NP * 16; ++i) sm[i * 128 + threadIdx.x] = 0u;
  VPlane vc[NP];
#pragma unroll
  for (int q = 0; q < NP; ++q) vc[q].clear();
  const uint64_t stride = rows.stride;
  // 8 rows starting at row rr; GUARD: rows >= r1 read as zero
  auto load8 = [&](uint32_t rr, uint32_t (&x)[8][2], auto guard) {
    constexpr bool GUARD = decltype(guard)::value;
    if (MODE == 2) {
      const uint8_t* rp[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) rp[j] = rows.ptrs[GUARD ? min(rr + j, r1 - 1) : rr + j] + 4ull * w0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (!GUARD || rr + j < r1) load_pair<MODE>(rp[j], x[j]);
        else x[j][0] = x[j][1] = 0u;
      }
    } else {
      const uint8_t* p = rows.base + static_cast<uint64_t>(rr) * stride + 4ull * w0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (!GUARD || rr + j < r1) load_pair<MODE>(p, x[j]);
        else x[j][0] = x[j][1] = 0u;
        p += stride;
      }
    }
  };
  auto process = [&](const uint32_t (&x)[8][2], uint32_t (&carry)[NP]) {
    uint32_t pl[NP][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t a = x[j][0], b = x[j][1];
      const uint32_t e = bitselect(kM5, a, b << 1);
      const uint32_t o = bitselect(kM5, a >> 1, b);
      int q = 0;
      if (KINDS & 1) pl[q++][j] = e;
      if (KINDS & 2) pl[q++][j] = o;
      if (KINDS & 4) pl[q++][j] = e & o;
    }
#pragma unroll
    for (int q = 0; q < NP; ++q) carry[q] = vc[q].add8(pl[q]);
  };
  auto to_shared_all = [&]() {
#pragma unroll
    for (int q = 0; q < NP; ++q) vc[q].to_shared(sm + q * 16 * 128 + threadIdx.x);
  };
  uint32_t xa[8][2], xb[8][2];
  uint32_t ca[NP], cb[NP];
  uint32_t blocks = 0;
  uint32_t r = r0;
  using No = std::integral_constant<bool, false>;
  using Yes = std::integral_constant<bool, true>;
  if (r + 16 <= r1) {
    load8(r, xa, No());
    for (; r + 32 <= r1; r += 16) {  // the 16 rows after this block are all there: no guards
      load8(r + 8, xb, No());
      process(xa, ca);
      load8(r + 16, xa, No());
      process(xb, cb);
#pragma unroll
      for (int q = 0; q < NP; ++q) vc[q].add_carries(ca[q], cb[q]);
      if (++blocks == 15) {
        blocks = 0;
        to_shared_all();
      }
    }
    // r + 16 <= r1 < r + 32: xa holds rows r .. r+7
    load8(r + 8, xb, No());
    process(xa, ca);
    process(xb, cb);
#pragma unroll
    for (int q = 0; q < NP; ++q) vc[q].add_carries(ca[q], cb[q]);
    r += 16;
    to_shared_all();
  }
  if (r < r1) {  // fewer than 16 rows left
    load8(r, xa, Yes());
    load8(r + 8, xb, Yes());
    process(xa, ca);
    process(xb, cb);
#pragma unroll
    for (int q = 0; q < NP; ++q) vc[q].add_carries(ca[q], cb[q]);
    to_shared_all();
  }
  // one atomicAdd per sample and plane for the whole chunk
  const uint32_t s0 = 16u * w0;
#pragma unroll
  for (int q = 0; q < NP; ++q) {
    int kind = 0, seen = 0;
#pragma unroll
    for (int b = 0; b < 3; ++b) {
      if (KINDS & (1 << b)) {
        if (seen == q) kind = b;
        ++seen;
      }
    }
    uint32_t* __restrict__ acc = kind == 0 ? acc_het : (kind == 1 ? acc_hom : acc_mis);
#pragma unroll
    for (int m = 0; m < 16; ++m) {
      const uint32_t v = sm[(q * 16 + m) * 128 + threadIdx.x];
      const uint32_t lo = v & 0xffffu, hi = v >> 16;
      if (lo && s0 + m < N) atomicAdd(acc + s0 + m, lo);
      if (hi && s0 + 16 + m < N) atomicAdd(acc + s0 + 16 + m, hi);
    }
  }
}

// acc arrays hold plane counts (E, O, B) per raw sample; out gets het / hom_alt / missing of the
// subset samples, three arrays of n each.
__global__ void k_planes_to_counts(const uint32_t* __restrict__ pe, const uint32_t* __restrict__ po,
                                   const uint32_t* __restrict__ pb, const uint32_t* __restrict__ idx, uint32_t n,
                                   uint32_t* __restrict__ out, const uint32_t* __restrict__ base) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t s = idx ? idx[i] : i;
  const uint32_t b = pb[s] + (base ? base[2] : 0u);
  out[i] = pe[s] + (base ? base[0] : 0u) - b;
  out[n + i] = po[s] + (base ? base[1] : 0u) - b;
  out[2ull * n + i] = b;
}

__global__ void k_gather_u32(const uint32_t* __restrict__ src, const uint32_t* __restrict__ idx, uint32_t n,
                             uint32_t* __restrict__ dst, const uint32_t* __restrict__ add) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[idx ? idx[i] : i] + (add ? *add : 0u);
}

// Raw index of every subset sample, ascending (the subset order of PgrSampleSubsetIndex).
__global__ void k_subset_indices(const uint32_t* __restrict__ include_words, const uint32_t* __restrict__ cum,
                                 uint32_t nwords, uint32_t* __restrict__ out) {
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= nwords) return;
  uint32_t bits = include_words[w];
  uint32_t pos = cum[w];
  while (bits) {
    out[pos++] = 32u * w + (__ffs(bits) - 1);
    bits &= bits - 1;
  }
}

// Genotype counts of dense rows: one warp per row, rows may sit at any byte alignment.
template <bool MASKED>
__global__ void __launch_bounds__(kThreads)
k_count_rows(DenseRows rows, SubsetDev sub, uint4* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t n_eff = MASKED ? sub.subset_ct : rows.N;
  for (uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < rows.nrows; r += nwarps) {
    const uint8_t* p = row_ptr(rows, r);
    Planes c = {0u, 0u, 0u};
    if (MASKED) {
      planes_2bit_masked<32, 4, false>(p, rows.N, sub.mask2_shifts, sub.mask2_stride, lane, c);
    } else {
      planes_2bit<32, 4, false>(p, rows.N, lane, c);
    }
    c.e = group_sum<32>(c.e);
    c.o = group_sum<32>(c.o);
    c.b = group_sum<32>(c.b);
    if (lane == 0) out[r] = planes_to_counts(c, n_eff);
  }
}

}  // namespace

void launch_densify(const RecIndex& ix, const uint32_t* vlist, uint32_t v0, uint32_t nrows, uint8_t* scratch,
                    uint32_t slot_stride, const uint8_t** ptrs, uint32_t* work_counter, int* err, cudaStream_t s,
                    LaunchStats* st, const SparseAcc* sparse, const LdScratch* bases) {
  const LdScratch bs = bases ? *bases : LdScratch{};
  if (sparse) cudaMemsetAsync(sparse->dense_count, 0, sizeof(uint32_t), s);
  if (!nrows) return;
  cudaMemsetAsync(work_counter, 0, sizeof(uint32_t), s);
  uint64_t blocks = (static_cast<uint64_t>(nrows) * 32 + kThreads - 1) / kThreads;
  const uint64_t max_blocks = static_cast<uint64_t>(sm_of(st)) * 8;
  if (blocks > max_blocks) blocks = max_blocks;
  if (sparse) {
    k_densify<true><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix, vlist, v0, nrows, scratch, slot_stride, ptrs,
                                                                      work_counter, err, *sparse, bs);
  } else {
    const uint32_t row_smem = (((ix.N + 3u) >> 2) + 15u) / 16u * 16u;
    if (ix.fpos && ix.N >= 200u && row_smem <= kTeamMaxSmem && row_smem <= slot_stride) {
      k_densify_team<<<nrows, kThreads, row_smem, s>>>(ix, vlist, v0, nrows, scratch, slot_stride, ptrs, err, bs);
    } else {
      k_densify<false><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix, vlist, v0, nrows, scratch, slot_stride,
                                                                         ptrs, work_counter, err, SparseAcc{}, bs);
    }
  }
  if (st) st->launches++;
}

void launch_expand_i8(const DenseRows& rows, const uint32_t* subset_idx, uint32_t n_out, int8_t* out, cudaStream_t s,
                      LaunchStats* st) {
  if (!rows.nrows || !n_out) return;
  // x: enough threads that one covers at most a few 16-byte chunks of its row; y: rows
  const uint32_t chunks = (n_out + 15) / 16;
  uint32_t bx = (chunks + 4 * kThreads - 1) / (4 * kThreads);
  if (bx < 1) bx = 1;
  const uint32_t by = rows.nrows < 65535u ? rows.nrows : 65535u;
  // short rows: no point in 256 threads for a handful of chunks (the ragged ends need <= 15 threads each)
  uint32_t threads = subset_idx ? (n_out + 31) / 32 * 32 : (chunks + 31) / 32 * 32;
  threads = threads < 32 ? 32 : (threads > kThreads ? kThreads : threads);
  k_expand_i8<<<dim3(bx, by), threads, 0, s>>>(rows, subset_idx, n_out, out);
  if (st) st->launches++;
}

void launch_expand_i8_valid(const DenseRows& rows, const uint32_t* subset_idx, uint32_t n_out, uint32_t call_mask,
                            int8_t* out, uint32_t* vbits, uint32_t bitbase, cudaStream_t s, LaunchStats* st) {
  if (!rows.nrows || !n_out) return;
  call_mask &= 7u;  // a missing call is never valid
  const uint32_t table = ((call_mask & 2u) ? 0x00000100u : 0u) | ((call_mask & 4u) ? 0x00020000u : 0u);
  const uint32_t chunks = (n_out + 15) / 16;
  uint32_t bx = (chunks + 4 * kThreads - 1) / (4 * kThreads);
  if (bx < 1) bx = 1;
  const uint32_t by = rows.nrows < 65535u ? rows.nrows : 65535u;
  uint32_t threads = subset_idx ? (n_out + 31) / 32 * 32 : (chunks + 31) / 32 * 32;
  threads = threads < 32 ? 32 : (threads > kThreads ? kThreads : threads);
  k_expand_i8_valid<<<dim3(bx, by), threads, 0, s>>>(rows, subset_idx, n_out, call_mask, table, out, vbits, bitbase);
  if (st) st->launches++;
}

void launch_sample_acc(const DenseRows& rows, uint32_t kinds, uint32_t* acc_het, uint32_t* acc_hom, uint32_t* acc_mis,
                       cudaStream_t s, LaunchStats* st) {
  if (!rows.nrows || !rows.N) return;
  const uint32_t ncols = (rows.N + 15) >> 4;
  const uint32_t nthreads = (ncols + 1) / 2;  // two 32-bit words of a record per thread
  const uint32_t bx = (nthreads + 127) / 128;
  // 6 CTAs of 128 threads fit an SM (80 registers, 24 KB of counters): the grid is exactly one resident
  // wave -- rounded DOWN, a 7th CTA on a few SMs would double the kernel time -- and every chunk ends with
  // one atomic per sample, so chunks are as long as that allows (>= 240 rows when rows allow).
  uint32_t want_y = static_cast<uint32_t>(sm_of(st)) * 6 / bx;
  if (want_y < 1) want_y = 1;
  uint32_t rpc = (rows.nrows + want_y - 1) / want_y;
  if (rpc < 240) rpc = 240;
  rpc = (rpc + 15) / 16 * 16;
  const uint32_t by = (rows.nrows + rpc - 1) / rpc;
  dim3 grid(bx, by);
  const uintptr_t addr_bits = reinterpret_cast<uintptr_t>(rows.base) | static_cast<uintptr_t>(rows.stride);
  const int mode = rows.ptrs ? 2 : (addr_bits % 8 == 0 ? 0 : (addr_bits % 4 == 0 ? 1 : 3));
#define PD_LAUNCH_V(K, M) k_sample_acc_v<K, M><<<grid, 128, 0, s>>>(rows, rpc, acc_het, acc_hom, acc_mis)
  if (kinds == 4) {
    if (mode == 0) PD_LAUNCH_V(4, 0);
    else if (mode == 1) PD_LAUNCH_V(4, 1);
    else if (mode == 2) PD_LAUNCH_V(4, 2);
    else PD_LAUNCH_V(4, 3);
  } else {
    if (mode == 0) PD_LAUNCH_V(7, 0);
    else if (mode == 1) PD_LAUNCH_V(7, 1);
    else if (mode == 2) PD_LAUNCH_V(7, 2);
    else PD_LAUNCH_V(7, 3);
  }
#undef PD_LAUNCH_V
  if (st) st->launches++;
}

void launch_count_rows(const DenseRows& rows, const SubsetDev& sub, uint4* out, cudaStream_t s, LaunchStats* st) {
  if (!rows.nrows) return;
  uint64_t blocks = (static_cast<uint64_t>(rows.nrows) * 32 + kThreads - 1) / kThreads;
  const uint64_t max_blocks = static_cast<uint64_t>(sm_of(st)) * 8;
  if (blocks > max_blocks) blocks = max_blocks;
  if (sub.mask2_shifts) {
    k_count_rows<true><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(rows, sub, out);
  } else {
    k_count_rows<false><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(rows, sub, out);
  }
  if (st) st->launches++;
}

void launch_planes_to_counts(const uint32_t* pe, const uint32_t* po, const uint32_t* pb, const uint32_t* idx, uint32_t n,
                             uint32_t* out, cudaStream_t s, LaunchStats* st, const uint32_t* base) {
  if (!n) return;
  k_planes_to_counts<<<(n + 255) / 256, 256, 0, s>>>(pe, po, pb, idx, n, out, base);
  if (st) st->launches++;
}

void launch_gather_u32(const uint32_t* src, const uint32_t* idx, uint32_t n, uint32_t* dst, cudaStream_t s,
                       LaunchStats* st, const uint32_t* add) {
  if (!n) return;
  k_gather_u32<<<(n + 255) / 256, 256, 0, s>>>(src, idx, n, dst, add);
  if (st) st->launches++;
}

void launch_subset_indices(const uint32_t* include_words, const uint32_t* cum, uint32_t nwords, uint32_t* out,
                           cudaStream_t s, LaunchStats* st) {
  if (!nwords) return;
  k_subset_indices<<<(nwords + 255) / 256, 256, 0, s>>>(include_words, cum, nwords, out);
  if (st) st->launches++;
}

}  // namespace pdk
