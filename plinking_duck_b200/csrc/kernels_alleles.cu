// K5m: multiallelic hardcalls -- PgrGetM + PglMultiallelicSparseToDenseMiss (pgenlib_read.cc:6668-6681,
// GetMultiallelicCodes :6617-6666; pgenlib_misc.cc:3552-3578): two allele codes per sample.
//
// The main track of a multiallelic record is the biallelic view (0 = REF/REF, 1 = REF/ALT1, 2 = ALT1/ALT1,
// 3 = missing); the aux1 track behind it (pgen_spec.tex:469-564) patches it:
//   aux1a  which of the "01" calls are REF/ALTx with x >= 2, and x;
//   aux1b  which of the "10" calls are ALTx/ALTy other than ALT1/ALT1, and (x, y).
// Each set is stored either as a bit array over the 01 (10) calls in sample order (mode 0) or as a list of sample
// ids in difflist form without genotypes (mode 1); mode 15 = empty.  The values follow, packed:
//   aux1a  0 / 1 / 2 / 4 / 8 bits per entry for 3 / 4 / 5-6 / 7-18 / >= 19 alleles, value + 2
//          (GetAux1aCodes, :6173-6211);
//   aux1b  1 bit (ALT1/ALT2 or ALT2/ALT2) for 3 alleles, else 2+2 / 4+4 / 8+8 bits for < 6 / < 18 / more alleles,
//          each + 1 (GetAux1bCodes, :6377-6414).
// Output row = N x 2 bytes: 0,0 / 0,1 / 1,1 / 255,255 (missing) from the main track, then the patches.
#include "aux1_device.cuh"
#include "device_common.cuh"
#include "kernels.cuh"

namespace pdk {

namespace {

constexpr int kThreads = 256;
constexpr uint32_t kM5 = 0x55555555u;

__device__ __forceinline__ void report_error(int* err, int code, uint32_t vidx) {
  if (atomicCAS(&err[0], 0, code) == 0) err[1] = static_cast<int>(vidx);
}

__device__ __forceinline__ const uint8_t* row_ptr(const DenseRows& rows, uint32_t r) {
  return rows.ptrs ? rows.ptrs[r] : rows.base + static_cast<uint64_t>(r) * rows.stride;
}

// One CTA (8 warps) per row, the row cut into 8 parts of whole 512-sample steps (as k_expand_phased): a warp counts
// the 01 and 10 calls of its part, warp 0 finds and measures the two patch sets, then every warp expands its part --
// with bit-array sets the patches are applied on the fly (the rank of a call among the 01 / 10 calls of the row is a
// prefix sum, the rank of a flagged call among the flagged ones a popcount over the set) -- and id-list sets are
// scattered over the finished row by warps 0 and 1 at the end.
__global__ void __launch_bounds__(kThreads)
k_expand_alleles(RecIndex ix, DenseRows rows, const uint32_t* __restrict__ vlist, uint32_t v0, uint32_t out_stride_pairs,
                 uint8_t* __restrict__ out, int* __restrict__ err) {
  constexpr int kWarps = kThreads / 32;
  __shared__ uint32_t sm_c01[kWarps], sm_c10[kWarps];
  __shared__ PatchSet sm_set[2];
  __shared__ uint32_t sm_allele_ct;
  __shared__ int sm_bad;
  const int lane = threadIdx.x & 31;
  const int wid = threadIdx.x >> 5;
  const uint32_t N = ix.N;
  const uint32_t nwords = (N + 15) >> 4;
  const uint32_t part = ((nwords + kWarps - 1) / kWarps + 31) & ~31u;  // words per warp, whole steps
  const uint32_t wbeg = min(nwords, wid * part), wend = min(nwords, wbeg + part);
  for (uint32_t r = blockIdx.x; r < rows.nrows; r += gridDim.x) {
    const uint32_t v = vlist ? vlist[r] : v0 + r;
    uint32_t vrtype;
    const uint8_t *p, *end;
    if (ix.fpos) {
      vrtype = ix.vrtypes[v];
      p = ix.data + (ix.fpos[v] - ix.data_fpos);
      end = ix.data + (ix.fpos[v + 1] - ix.data_fpos);
    } else {
      vrtype = ix.const_vrtype;
      p = ix.data + (ix.fixed_start + static_cast<uint64_t>(v) * ix.fixed_width - ix.data_fpos);
      end = p + ix.fixed_width;
    }
    const bool multi = (vrtype & 8u) != 0;
    const uint8_t* row = row_ptr(rows, r);
    auto word_at = [&](uint32_t w) -> uint32_t {
      uint32_t x = load_bits32(row, static_cast<uint64_t>(w) << 5);
      const uint32_t valid = N - (w << 4);
      if (valid < 16) x &= (1u << (2 * valid)) - 1u;
      return x;
    };
    // ---- 01 / 10 calls of my part ----
    uint32_t my01 = 0, my10 = 0;
    if (multi) {
      for (uint32_t w = wbeg + lane; w < wend; w += 32) {
        const uint32_t x = word_at(w);
        my01 += __popc(x & ~(x >> 1) & kM5);
        my10 += __popc((x >> 1) & ~x & kM5);
      }
      my01 = group_sum<32>(my01);
      my10 = group_sum<32>(my10);
    }
    if (lane == 0) {
      sm_c01[wid] = my01;
      sm_c10[wid] = my10;
    }
    __syncthreads();
    uint32_t tot01 = 0, tot10 = 0, base01 = 0, base10 = 0;
#pragma unroll
    for (int j = 0; j < kWarps; ++j) {
      tot01 += sm_c01[j];
      tot10 += sm_c10[j];
      if (j < wid) {
        base01 += sm_c01[j];
        base10 += sm_c10[j];
      }
    }
    // ---- warp 0: the two patch sets ----
    if (wid == 0) {
      int bad = 0;
      PatchSet a, b;
      a.mode = b.mode = 15;
      a.count = b.count = 0;
      a.fset = a.vals = b.fset = b.vals = p;
      uint32_t allele_ct = 2;
      if (multi) {
        allele_ct = ix.allele_cts ? ix.allele_cts[v] : 2u;
        const uint8_t* q = nullptr;
        if (allele_ct < 3) {
          bad = kErrMalformed;  // a patch track needs a third allele
        } else if (!main_track_end(vrtype & 7u, p, end, N, lane, &q) || q >= end) {
          bad = kErrMalformed;
        } else {
          const uint32_t modes = *q++;
          if (!parse_patch_set(modes & 15u, q, end, N, tot01, aux1a_width(allele_ct), lane, a) ||
              !parse_patch_set(modes >> 4, q, end, N, tot10, aux1b_width(allele_ct), lane, b)) {
            bad = kErrMalformed;
          }
        }
        if (bad) a.mode = b.mode = 15;
      }
      if (lane == 0) {
        sm_set[0] = a;
        sm_set[1] = b;
        sm_allele_ct = allele_ct;
        sm_bad = bad;
        if (bad) report_error(err, bad, v);
      }
    }
    __syncthreads();
    const PatchSet A = sm_set[0], B = sm_set[1];
    const uint32_t wa = aux1a_width(max(sm_allele_ct, 3u)), wb = aux1b_width(max(sm_allele_ct, 3u));
    // flagged calls in front of my part (bit-array sets)
    uint32_t flagA = 0, flagB = 0;
    if (A.mode == 0) flagA = group_sum<32>(popcount_bits<32>(A.fset, base01, nullptr, 0, lane));
    if (B.mode == 0) flagB = group_sum<32>(popcount_bits<32>(B.fset, base10, nullptr, 0, lane));
    // ---- expand my part ----
    uint8_t* orow = out + static_cast<uint64_t>(r) * out_stride_pairs * 2;
    const bool vec_ok = (reinterpret_cast<uintptr_t>(orow) & 15u) == 0;
    for (uint32_t w0 = wbeg; w0 < wend; w0 += 32) {
      const uint32_t w = w0 + lane;
      const uint32_t x = w < wend ? word_at(w) : 0u;
      const uint32_t m01 = x & ~(x >> 1) & kM5, m10 = (x >> 1) & ~x & kM5;
      uint32_t bitsA = 0, bitsB = 0, offA = 0, offB = 0;
      if (A.mode == 0 || B.mode == 0) {  // uniform for the CTA
        const uint32_t c01 = __popc(m01), c10 = __popc(m10);
        uint32_t i01 = c01, i10 = c10;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
          const uint32_t t1 = __shfl_up_sync(0xffffffffu, i01, off), t2 = __shfl_up_sync(0xffffffffu, i10, off);
          if (lane >= off) {
            i01 += t1;
            i10 += t2;
          }
        }
        const uint32_t h01 = base01 + i01 - c01, h10 = base10 + i10 - c10;  // rank of my first 01 / 10 call
        if (A.mode == 0 && c01) bitsA = load_bits32(A.fset, h01) & ((c01 >= 32 ? 0u : (1u << c01)) - 1u);
        if (B.mode == 0 && c10) bitsB = load_bits32(B.fset, h10) & ((c10 >= 32 ? 0u : (1u << c10)) - 1u);
        uint32_t fA = __popc(bitsA), fB = __popc(bitsB);
        uint32_t jA = fA, jB = fB;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
          const uint32_t t1 = __shfl_up_sync(0xffffffffu, jA, off), t2 = __shfl_up_sync(0xffffffffu, jB, off);
          if (lane >= off) {
            jA += t1;
            jB += t2;
          }
        }
        offA = flagA + jA - fA;
        offB = flagB + jB - fB;
        base01 += __shfl_sync(0xffffffffu, i01, 31);
        base10 += __shfl_sync(0xffffffffu, i10, 31);
        flagA += __shfl_sync(0xffffffffu, jA, 31);
        flagB += __shfl_sync(0xffffffffu, jB, 31);
      }
      if (w < wend) {
        uint32_t o[8];
        uint32_t n01 = 0, n10 = 0;  // 01 / 10 calls of this word seen so far
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          uint32_t word = 0;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint32_t g = (x >> (4 * i + 2 * h)) & 3u;
            uint32_t pair = g == 0 ? 0x0000u : (g == 1 ? 0x0100u : (g == 2 ? 0x0101u : 0xFFFFu));
            if (g == 1) {
              if ((bitsA >> n01) & 1u) pair = aux1a_code(A.vals, wa, offA++) << 8;
              ++n01;
            } else if (g == 2) {
              if ((bitsB >> n10) & 1u) pair = aux1b_pair(B.vals, wb, offB++);
              ++n10;
            }
            word |= pair << (16 * h);
          }
          o[i] = word;
        }
        const uint32_t s0 = w << 4;
        if (vec_ok && s0 + 16 <= N) {
          uint4* dst = reinterpret_cast<uint4*>(orow + 2ull * s0);
          __stcs(dst, make_uint4(o[0], o[1], o[2], o[3]));
          __stcs(dst + 1, make_uint4(o[4], o[5], o[6], o[7]));
        } else {
          for (uint32_t k = 0; k < 16 && s0 + k < N; ++k) {
            const uint32_t pair = (o[k >> 1] >> (16 * (k & 1))) & 0xffffu;
            orow[2ull * (s0 + k)] = static_cast<uint8_t>(pair & 0xffu);
            orow[2ull * (s0 + k) + 1] = static_cast<uint8_t>(pair >> 8);
          }
        }
      }
    }
    // ---- id-list sets: scattered over the finished row ----
    if (A.mode == 1 || B.mode == 1) {  // uniform for the CTA
      __syncthreads();
      if (wid == 0 && A.mode == 1 && A.count) {
        difflist_walk(A.list, N, lane, [&](uint32_t id, uint32_t, uint32_t k) {
          orow[2ull * id + 1] = static_cast<uint8_t>(aux1a_code(A.vals, wa, k));
        });
      }
      if (wid == 1 && B.mode == 1 && B.count) {
        difflist_walk(B.list, N, lane, [&](uint32_t id, uint32_t, uint32_t k) {
          const uint32_t pair = aux1b_pair(B.vals, wb, k);
          orow[2ull * id] = static_cast<uint8_t>(pair & 0xffu);
          orow[2ull * id + 1] = static_cast<uint8_t>(pair >> 8);
        });
      }
    }
    __syncthreads();  // the shared row state is rewritten for the next row
  }
}

}  // namespace

void launch_expand_alleles(const RecIndex& ix, const DenseRows& rows, const uint32_t* vlist, uint32_t v0,
                           const uint32_t* subset_idx, uint32_t n_out, uint8_t* raw_scratch, uint8_t* out, int* err,
                           cudaStream_t s, LaunchStats* st) {
  if (!rows.nrows || !n_out) return;
  uint64_t blocks = rows.nrows;  // one CTA per row
  const uint64_t cap = static_cast<uint64_t>(sm_of(st)) * 16;
  if (blocks > cap) blocks = cap;
  if (!subset_idx) {
    k_expand_alleles<<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix, rows, vlist, v0, ix.N, out, err);
    if (st) st->launches++;
    return;
  }
  // subset: expand all raw samples, then gather the kept pairs
  const uint32_t raw_stride = (ix.N + 7) & ~7u;  // keeps every scratch row 16-byte aligned
  k_expand_alleles<<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix, rows, vlist, v0, raw_stride, raw_scratch, err);
  launch_gather_pairs(reinterpret_cast<const uint16_t*>(raw_scratch), raw_stride, subset_idx, n_out, rows.nrows,
                      reinterpret_cast<uint16_t*>(out), s, st);
  if (st) st->launches++;
}

}  // namespace pdk
