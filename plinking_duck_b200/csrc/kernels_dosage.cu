// K6: dosage tracks.  Replaces PgrGetD -> ReadGenovecHphaseSubsetUnsafe + ParseDosage16
// (pgenlib_read.cc:7484-7505, 7210-7482) + Dosage16ToDoublesMinus9 (pgenlib_ffi_support.cc:268-279)
// and the dosage half of PgrGetDCounts -> GetBasicGenotypeCountsAndDosage16s (:7921-8172).
//
// Record layout after the main genotype track (pgen_spec.tex:469-610): [aux1 multiallelic patches,
// bit 3] [aux2 hardcall phase, bit 4] [aux3 dosage-present list / bitarray] [aux4 u16 dosages]
// [aux7/8 dosage phase, bit 7].  The dosage track can only be found by walking what precedes it:
// the main track's difflist ends where its last varint group ends, and the phase track's length
// depends on the number of heterozygous hardcalls (SkipAux2, pgenlib_read.cc:6844-6865).
#include "aux1_device.cuh"
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

__device__ __forceinline__ uint32_t ld_u16(const uint8_t* p, uint32_t i) {
  return static_cast<uint32_t>(p[2ull * i]) | (static_cast<uint32_t>(p[2ull * i + 1]) << 8);
}

// End of a difflist that starts at q (warp-uniform result).  has_geno: 2-bit values present.
__device__ __forceinline__ bool difflist_end(const uint8_t* q, const uint8_t* end, uint32_t N, bool has_geno, int lane,
                                             const uint8_t** out_end, Difflist* out_d) {
  Difflist d;
  if (!parse_difflist(q, end, N, has_geno, d)) return false;
  if (out_d) *out_d = d;
  if (!d.len) {
    *out_end = q;
    return true;
  }
  return difflist_walk_lanes(d, N, lane, [](uint32_t, uint32_t, uint32_t) {}, out_end);  // end-finding only: the small walker keeps K5p at 50 registers
}

// ---------------------------------------------------------------------------------------------
// k_dosage_locate: one warp per row.  Finds aux3 / aux4, validates their extent, and for the
// bitarray form writes the exclusive prefix popcount per 32-bit word (rank table) used to map a
// sample to its slot in the u16 value array.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
k_dosage_locate(RecIndex ix, DenseRows rows, const uint32_t* __restrict__ vlist, uint32_t v0,
                DosageLoc* __restrict__ loc, uint32_t* __restrict__ rank, uint32_t rank_stride,
                int* __restrict__ err) {
  const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= rows.nrows) return;
  const uint32_t N = ix.N;
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
  DosageLoc out;
  out.track3 = nullptr;
  out.values = nullptr;
  out.kind = 0;
  out.raw_ct = 0;
  const uint32_t dos = (vrtype >> 5) & 3u;
  int bad = 0;
  if (dos) {
    if (vrtype & 0x08u) {
      bad = kErrNotSupported;  // multiallelic dosage: pgenlib_read.cc:7350-7356 refuses it too
    } else {
      const uint32_t main = vrtype & 7u;
      const uint8_t* q = p;
      bool ok = true;
      if (main == 0) {
        q = p + ((N + 3) >> 2);
      } else if (main == 1) {
        q = p + 1 + ((N + 7) >> 3);
        ok = q <= end && difflist_end(q, end, N, true, lane, &q, nullptr);
      } else if (main != 5) {
        ok = difflist_end(q, end, N, true, lane, &q, nullptr);
      }
      ok = ok && q <= end;
      if (ok && (vrtype & 0x10u)) {
        // aux2: 1 + het_ct/8 bytes; if bit 0 is set they are an explicit phasepresent array and
        // ceil(popcount - 1, 8) bytes of phaseinfo follow.
        Planes c = {0u, 0u, 0u};
        planes_2bit<32, 2, false>(row_ptr(rows, r), N, lane, c);
        const uint32_t het = group_sum<32>(c.e) - group_sum<32>(c.b);
        const uint32_t first = 1 + (het >> 3);
        ok = static_cast<uint64_t>(end - q) >= first;
        if (ok) {
          const bool explicit_pp = q[0] & 1u;
          uint32_t pc = popcount_bits<32>(q, first * 8, nullptr, 0, lane);
          pc = group_sum<32>(pc);
          q += first;
          if (explicit_pp) {
            const uint32_t second = (pc - 1 + 7) >> 3;
            ok = static_cast<uint64_t>(end - q) >= second;
            q += second;
          }
        }
      }
      if (ok) {
        if (dos == 1) {
          Difflist d;
          const uint8_t* vals = q;
          ok = difflist_end(q, end, N, false, lane, &vals, &d);
          out.track3 = q;
          out.values = vals;
          out.raw_ct = ok ? d.len : 0;
          out.kind = 1;
          // same header rules as a genotype difflist, incl. the N/8 length cap
          // (ParseAndSaveDeltalistAsBitarr -> ParseDifflistHeader, pgenlib_read.cc:6136-6138)
        } else if (dos == 2) {
          out.values = q;
          out.raw_ct = N;
          out.kind = 2;
        } else {
          const uint32_t bit_bytes = (N + 7) >> 3;
          ok = static_cast<uint64_t>(end - q) >= bit_bytes;
          if (ok) {
            out.track3 = q;
            out.values = q + bit_bytes;
            out.kind = 3;
            // rank table + total
            const uint32_t nw = (N + 31) >> 5;
            uint32_t running = 0;
            uint32_t* rk = rank + static_cast<uint64_t>(r) * rank_stride;
            for (uint32_t w0 = 0; w0 < nw; w0 += 32) {
              const uint32_t w = w0 + lane;
              uint32_t bits = 0;
              if (w < nw) {
                bits = load_bits32(q, static_cast<uint64_t>(w) << 5);
                const uint32_t valid = min(32u, N - 32u * w);
                if (valid < 32) bits &= (1u << valid) - 1u;
              }
              const uint32_t pc = __popc(bits);
              uint32_t incl = pc;
#pragma unroll
              for (int off = 1; off < 32; off <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
                if (lane >= off) incl += t;
              }
              if (w < nw) rk[w] = running + incl - pc;
              running += __shfl_sync(0xffffffffu, incl, 31);
            }
            out.raw_ct = running;
          }
        }
      }
      if (ok) ok = static_cast<uint64_t>(end - out.values) >= 2ull * out.raw_ct;
      if (!ok) bad = kErrMalformed;
    }
  }
  if (lane == 0) {
    if (bad) {
      report_error(err, bad, v);
      out.kind = 0;
    }
    loc[r] = out;
  }
}

// ---------------------------------------------------------------------------------------------
// k_dosage_fill: out[r][j] = hardcall {0,1,2,-9} of subset sample j, overridden by its dosage
// (u16 * 2^-14) for the unconditional and bitarray forms.  List-form rows are patched afterwards
// by k_dosage_scatter.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
k_dosage_fill(DenseRows rows, const DosageLoc* __restrict__ loc, const uint32_t* __restrict__ rank,
              uint32_t rank_stride, const uint32_t* __restrict__ subset_idx, uint32_t n_out, double* __restrict__ out) {
  const uint32_t r = blockIdx.y;
  const uint8_t* row = row_ptr(rows, r);
  const DosageLoc L = loc[r];
  const uint32_t* rk = rank + static_cast<uint64_t>(r) * rank_stride;
  double* o = out + static_cast<uint64_t>(r) * n_out;
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n_out; j += gridDim.x * blockDim.x) {
    const uint32_t raw = subset_idx ? subset_idx[j] : j;
    const uint32_t g = geno_at(row, raw);
    double val = (g == 3) ? -9.0 : static_cast<double>(g);
    if (L.kind == 2) {
      const uint32_t u = ld_u16(L.values, raw);
      if (u != 65535u) val = static_cast<double>(u) * 0.00006103515625;
    } else if (L.kind == 3) {
      const uint32_t bits = load_bits32(L.track3, static_cast<uint64_t>(raw >> 5) << 5);
      if ((bits >> (raw & 31)) & 1u) {
        const uint32_t k = rk[raw >> 5] + __popc(bits & ((1u << (raw & 31)) - 1u));
        val = static_cast<double>(ld_u16(L.values, k)) * 0.00006103515625;
      }
    }
    o[j] = val;
  }
}

__device__ __forceinline__ bool subset_pos(const SubsetDev& sub, uint32_t id, uint32_t& pos) {
  if (!sub.include_words) {
    pos = id;
    return true;
  }
  const uint32_t w = sub.include_words[id >> 5];
  if (!((w >> (id & 31)) & 1u)) return false;
  pos = sub.cum_popcount[id >> 5] + __popc(w & ((1u << (id & 31)) - 1u));
  return true;
}

__global__ void __launch_bounds__(kThreads)
k_dosage_scatter(DenseRows rows, const DosageLoc* __restrict__ loc, SubsetDev sub, uint32_t n_out,
                 double* __restrict__ out, int* __restrict__ err) {
  const uint32_t r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= rows.nrows) return;
  const DosageLoc L = loc[r];
  if (L.kind != 1 || !L.raw_ct) return;
  const uint8_t* q = L.track3;
  Difflist d;
  if (!parse_difflist(q, L.values, rows.N, false, d)) return;  // validated by k_dosage_locate
  double* o = out + static_cast<uint64_t>(r) * n_out;
  difflist_walk(d, rows.N, lane, [&](uint32_t id, uint32_t, uint32_t k) {
    uint32_t pos;
    if (subset_pos(sub, id, pos)) o[pos] = static_cast<double>(ld_u16(L.values, k)) * 0.00006103515625;
  });
}

// ---------------------------------------------------------------------------------------------
// k_dosage_stats: one CTA (8 warps) per row -> integer sums over the (subset) samples that carry a
// dosage: count, sum u16, sum u16^2 and the hardcall categories they replace.  The sample range of
// the unconditional (one u16 per sample) and bitarray forms is split over the warps; a lane of the
// unconditional form takes 16 consecutive samples per step -- one word of the dense row and nine
// aligned words holding their 16 values (the value track starts at any byte).  The list form is
// short (<= N/8 entries) and stays on warp 0.  Warp partial sums meet in shared memory.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
k_dosage_stats(DenseRows rows, const DosageLoc* __restrict__ loc, const uint32_t* __restrict__ rank,
               uint32_t rank_stride, SubsetDev sub, uint32_t n_out, DosageStats* __restrict__ stats) {
  constexpr int kWarps = kThreads / 32;
  __shared__ unsigned long long sm_sum[kWarps], sm_sq[kWarps];
  __shared__ uint32_t sm_ct[kWarps][5];
  const int lane = threadIdx.x & 31;
  const int wid = threadIdx.x >> 5;
  const uint32_t N = rows.N;
  for (uint32_t r = blockIdx.x; r < rows.nrows; r += gridDim.x) {
    const DosageLoc L = loc[r];
    const uint8_t* row = row_ptr(rows, r);
    unsigned long long sum = 0, sumsq = 0;
    uint32_t ct = 0, r0 = 0, r1 = 0, r2 = 0, r3 = 0;
    auto take = [&](uint32_t g, uint32_t u) {
      sum += u;
      sumsq += static_cast<unsigned long long>(u) * u;
      ++ct;
      r0 += (g == 0);
      r1 += (g == 1);
      r2 += (g == 2);
      r3 += (g == 3);
    };
    if (L.kind == 1 && L.raw_ct) {
      if (wid == 0) {
        const uint8_t* q = L.track3;
        Difflist d;
        if (parse_difflist(q, L.values, N, false, d)) {
          difflist_walk(d, N, lane, [&](uint32_t id, uint32_t, uint32_t k) {
            uint32_t pos;
            if (subset_pos(sub, id, pos)) take(geno_at(row, id), ld_u16(L.values, k));
          });
        }
      }
    } else if (L.kind == 2) {
      if (!sub.subset_idx) {
        const uint32_t nblk = (N + 15) >> 4;  // blocks of 16 samples
        for (uint32_t blk = threadIdx.x; blk < nblk; blk += kThreads) {
          const uint32_t x = load_bits32(row, static_cast<uint64_t>(blk) << 5);
          const uintptr_t a = reinterpret_cast<uintptr_t>(L.values) + 32ull * blk;
          const uint32_t* wp = reinterpret_cast<const uint32_t*>(a & ~static_cast<uintptr_t>(3));
          const uint32_t sh = static_cast<uint32_t>(a & 3) << 3;
          const uint32_t nval = min(16u, N - (blk << 4));
          const uint32_t nwords = (2 * nval + (static_cast<uint32_t>(a & 3)) + 3) >> 2;  // words that hold them
          uint32_t t[9];
#pragma unroll
          for (int k = 0; k < 9; ++k) t[k] = static_cast<uint32_t>(k) < nwords ? __ldg(wp + k) : 0u;
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint32_t two = __funnelshift_r(t[k], t[k + 1], sh);
            if (2u * k < nval) {
              const uint32_t u = two & 0xffffu;
              if (u != 65535u) take((x >> (4 * k)) & 3u, u);
            }
            if (2u * k + 1 < nval) {
              const uint32_t u = two >> 16;
              if (u != 65535u) take((x >> (4 * k + 2)) & 3u, u);
            }
          }
        }
      } else {
        for (uint32_t j = threadIdx.x; j < n_out; j += kThreads) {
          const uint32_t raw = sub.subset_idx[j];
          const uint32_t u = ld_u16(L.values, raw);
          if (u != 65535u) take(geno_at(row, raw), u);
        }
      }
    } else if (L.kind == 3) {
      const uint32_t nw = (N + 31) >> 5;
      const uint32_t* rk = rank + static_cast<uint64_t>(r) * rank_stride;
      for (uint32_t w = threadIdx.x; w < nw; w += kThreads) {
        uint32_t all = load_bits32(L.track3, static_cast<uint64_t>(w) << 5);
        const uint32_t valid = min(32u, N - 32u * w);
        if (valid < 32) all &= (1u << valid) - 1u;
        uint32_t bits = sub.include_words ? (all & sub.include_words[w]) : all;
        const uint32_t base = rk[w];
        // the 32 samples of this word are two words of the dense row
        const uint32_t xlo = load_bits32(row, static_cast<uint64_t>(w) << 6);
        const uint32_t xhi = load_bits32(row, (static_cast<uint64_t>(w) << 6) + 32);
        while (bits) {
          const uint32_t b = __ffs(bits) - 1;
          const uint32_t g = b < 16 ? (xlo >> (2 * b)) & 3u : (xhi >> (2 * (b - 16))) & 3u;
          take(g, ld_u16(L.values, base + __popc(all & ((1u << b) - 1u))));
          bits &= bits - 1;
        }
      }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      sum += __shfl_xor_sync(0xffffffffu, sum, off);
      sumsq += __shfl_xor_sync(0xffffffffu, sumsq, off);
    }
    ct = group_sum<32>(ct);
    r0 = group_sum<32>(r0);
    r1 = group_sum<32>(r1);
    r2 = group_sum<32>(r2);
    r3 = group_sum<32>(r3);
    if (lane == 0) {
      sm_sum[wid] = sum;
      sm_sq[wid] = sumsq;
      sm_ct[wid][0] = ct;
      sm_ct[wid][1] = r0;
      sm_ct[wid][2] = r1;
      sm_ct[wid][3] = r2;
      sm_ct[wid][4] = r3;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      DosageStats st{};
#pragma unroll
      for (int j = 0; j < kWarps; ++j) {
        st.sum += sm_sum[j];
        st.sumsq += sm_sq[j];
        st.ct += sm_ct[j][0];
        st.replaced[0] += sm_ct[j][1];
        st.replaced[1] += sm_ct[j][2];
        st.replaced[2] += sm_ct[j][3];
        st.replaced[3] += sm_ct[j][4];
      }
      st.kind = L.kind;
      st.pad = 0;
      stats[r] = st;
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------
// k_expand_phased: PgrGetP + UnpackPhasedGenotypes (src/plink_common.cpp:1549-1584) -- one warp per
// row; output row = N x 2 int8 allele pairs: hom-ref 0,0  hom-alt 1,1  missing -9,-9  het 0,1 unless
// the aux2 track says "phased, ALT first" -> 1,0.
// aux2 (pgen_spec.tex "hardcall phase", ReadGenovecHphaseSubsetUnsafe pgenlib_read.cc:6874-6962):
// 1 + het_ct bits; bit 0 clear: every het is phased and bits 1.. are the phaseinfo bits; bit 0 set:
// bits 1.. are phasepresent flags and a second bit array (one bit per set flag) follows at the next
// byte.  A lane owns one 32-bit word of the dense row (16 samples) per step of 512 samples; the
// index of its first het among the row's hets, and of its first set flag among the flags, come from
// warp prefix sums carried across steps.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t warp_excl_scan(uint32_t x, int lane, uint32_t* total) {
  uint32_t incl = x;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += t;
  }
  *total = __shfl_sync(0xffffffffu, incl, 31);
  return incl - x;
}

// MULTI: compiled with the multiallelic + phased pre-pass (only launched for files that need it: it costs registers).
template <bool MULTI>
__global__ void __launch_bounds__(kThreads)
k_expand_phased(RecIndex ix, DenseRows rows, const uint32_t* __restrict__ vlist, uint32_t v0, uint32_t out_stride_pairs,
                int8_t* __restrict__ out, int* __restrict__ err, uint32_t* __restrict__ mhet_scratch,
                uint32_t mhet_stride_words) {
  // One CTA (8 warps) per row: the row is cut into 8 contiguous parts of whole 512-sample steps; a warp
  // first counts the hets of its part, the counts are exchanged through shared memory, and the flag
  // offset of a part is a popcount over the aux2 bits before it -- so the parts expand independently.
  constexpr int kWarps = kThreads / 32;
  __shared__ uint32_t sm_het[kWarps];
  __shared__ uint32_t sm_c10[kWarps];
  __shared__ PatchSet sm_setb;
  __shared__ uint32_t sm_wb;
  __shared__ const uint8_t* sm_q;
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
    const uint8_t* row = row_ptr(rows, r);
    auto word_at = [&](uint32_t w) -> uint32_t {
      uint32_t x = load_bits32(row, static_cast<uint64_t>(w) << 5);
      const uint32_t valid = N - (w << 4);
      if (valid < 16) x &= (1u << (2 * valid)) - 1u;
      return x;
    };
    // ---- multiallelic record with a phase track: the phase track counts the ALTx/ALTy (x != y) calls of the patch
    // set aux1b among its hets (PgrDetectGenoarrHetsMultiallelic / GetAux1bHets, pgenlib_read.cc:5270-5409,6874-6962),
    // although the biallelic view returns them as hom-alt and their phase bits are dropped.  They are marked in a
    // per-CTA bit array (halfword w = samples 16 w .. 16 w + 15) before the hets are counted. ----
    const bool mp = MULTI && (vrtype & 0x18u) == 0x18u && mhet_scratch != nullptr && ix.allele_cts != nullptr;
    uint16_t* mh = reinterpret_cast<uint16_t*>(mhet_scratch + static_cast<uint64_t>(blockIdx.x) * mhet_stride_words);
    if (mp) {  // uniform for the CTA
      uint32_t my01 = 0, my10 = 0;
      for (uint32_t w = wbeg + lane; w < wend; w += 32) {
        const uint32_t x = word_at(w);
        my01 += __popc(x & ~(x >> 1) & kM5);
        my10 += __popc((x >> 1) & ~x & kM5);
        mh[w] = 0;
      }
      my01 = group_sum<32>(my01);
      my10 = group_sum<32>(my10);
      if (lane == 0) {
        sm_het[wid] = my01;
        sm_c10[wid] = my10;
      }
      __syncthreads();
      uint32_t tot01 = 0, tot10 = 0, base10 = 0;
#pragma unroll
      for (int j = 0; j < kWarps; ++j) {
        tot01 += sm_het[j];
        tot10 += sm_c10[j];
        if (j < wid) base10 += sm_c10[j];
      }
      if (wid == 0) {
        int bad = 0;
        PatchSet a, b;
        b.mode = 15;
        b.count = 0;
        b.fset = b.vals = p;
        const uint32_t allele_ct = ix.allele_cts[v];
        const uint8_t* q = nullptr;
        if (allele_ct < 3) {
          bad = kErrMalformed;
        } else if (!main_track_end(vrtype & 7u, p, end, N, lane, &q) || q >= end) {
          bad = kErrMalformed;
        } else {
          const uint32_t modes = *q++;
          if (!parse_patch_set(modes & 15u, q, end, N, tot01, aux1a_width(allele_ct), lane, a) ||
              !parse_patch_set(modes >> 4, q, end, N, tot10, aux1b_width(allele_ct), lane, b)) {
            bad = kErrMalformed;
          }
        }
        if (bad) {
          b.mode = 15;
          q = nullptr;
        }
        if (lane == 0) {
          sm_setb = b;
          sm_wb = aux1b_width(max(allele_ct, 3u));
          sm_q = q;  // start of aux2
          sm_bad = bad;
        }
      }
      __syncthreads();
      const PatchSet B = sm_setb;
      const uint32_t wb = sm_wb;
      if (B.mode == 0) {
        uint32_t flag = group_sum<32>(popcount_bits<32>(B.fset, base10, nullptr, 0, lane));
        for (uint32_t w0 = wbeg; w0 < wend; w0 += 32) {
          const uint32_t w = w0 + lane;
          const uint32_t x = w < wend ? word_at(w) : 0u;
          const uint32_t m10 = (x >> 1) & ~x & kM5;
          const uint32_t c10 = __popc(m10);
          uint32_t t10 = 0, tf = 0;
          const uint32_t h10 = base10 + warp_excl_scan(c10, lane, &t10);
          const uint32_t bits = c10 ? (load_bits32(B.fset, h10) & ((1u << c10) - 1u)) : 0u;
          uint32_t off = flag + warp_excl_scan(__popc(bits), lane, &tf);
          uint32_t marks = 0, n10 = 0;
          for (uint32_t m = m10; m; m &= m - 1u, ++n10) {
            if ((bits >> n10) & 1u) {
              const uint32_t pair = aux1b_pair(B.vals, wb, off++);
              if ((pair & 0xffu) != (pair >> 8)) marks |= 1u << ((__ffs(m) - 1) >> 1);
            }
          }
          if (w < wend && marks) mh[w] = static_cast<uint16_t>(marks);
          base10 += t10;
          flag += tf;
        }
      } else if (B.mode == 1 && wid == 0 && B.count) {
        uint32_t* words = mhet_scratch + static_cast<uint64_t>(blockIdx.x) * mhet_stride_words;
        difflist_walk(B.list, N, lane, [&](uint32_t id, uint32_t, uint32_t k) {
          const uint32_t pair = aux1b_pair(B.vals, wb, k);
          if ((pair & 0xffu) != (pair >> 8)) atomicOr(words + (id >> 5), 1u << (id & 31u));
        });
      }
      __syncthreads();
    }
    // all hets of word w: the 01 calls plus (multiallelic + phased records) the marked ALTx/ALTy calls
    auto hets_of = [&](uint32_t w, uint32_t x) -> uint32_t {
      uint32_t h = x & ~(x >> 1) & kM5;
      if (mp) h |= spread16(__ldcg(reinterpret_cast<const unsigned short*>(mh) + w));
      return h;
    };
    // ---- hets of my part; warp 0 also finds the end of the main track ----
    uint32_t my_het = 0;
    for (uint32_t w = wbeg + lane; w < wend; w += 32) {
      const uint32_t x = word_at(w);
      my_het += __popc(hets_of(w, x));
    }
    my_het = group_sum<32>(my_het);
    if (lane == 0) sm_het[wid] = my_het;
    if (wid == 0) {
      int bad = 0;
      const uint8_t* q = nullptr;
      if (vrtype & 0x10u) {
        if (vrtype & 0x08u) {
          // the patch sets precede aux2: found above when the caller supplied allele counts (pd_set_allele_counts);
          // without them the track cannot be measured (pgenlib needs allele_idx_offsets for the same reason)
          if (mp) {
            q = sm_q;
            bad = sm_bad;
          } else {
            bad = kErrNotSupported;
          }
        } else {
          const uint32_t main = vrtype & 7u;
          q = p;
          bool ok = true;
          if (main == 0) {
            q = p + ((N + 3) >> 2);
          } else if (main == 1) {
            q = p + 1 + ((N + 7) >> 3);
            ok = q <= end && difflist_end(q, end, N, true, lane, &q, nullptr);
          } else if (main != 5) {
            ok = difflist_end(q, end, N, true, lane, &q, nullptr);
          }
          if (!ok || q > end) {
            bad = kErrMalformed;
            q = nullptr;
          }
        }
      }
      if (lane == 0) {
        sm_q = q;
        sm_bad = bad;
      }
    }
    __syncthreads();
    uint32_t het_total = 0, het_base = 0;
#pragma unroll
    for (int j = 0; j < kWarps; ++j) {
      het_total += sm_het[j];
      if (j < wid) het_base += sm_het[j];
    }
    const uint8_t* aux2 = sm_q;  // bit 0 = explicit flag, bits 1.. = flags or phaseinfo
    const uint8_t* second = nullptr;
    bool explicit_pp = false;
    int bad = sm_bad;
    uint32_t flag_base = 0;
    if (aux2) {
      const uint32_t first = 1 + (het_total >> 3);
      if (static_cast<uint64_t>(end - aux2) < first) {
        bad = kErrMalformed;
      } else {
        explicit_pp = aux2[0] & 1u;
        if (explicit_pp) {
          second = aux2 + first;
          uint32_t pc = popcount_bits<32>(aux2, het_total + 1, nullptr, 0, lane);
          pc = group_sum<32>(pc) - 1;  // flags in total (minus the form bit itself)
          if (static_cast<uint64_t>(end - second) < ((pc + 7) >> 3)) bad = kErrMalformed;
          uint32_t before = popcount_bits<32>(aux2, het_base + 1, nullptr, 0, lane);
          flag_base = group_sum<32>(before) - 1;
        }
      }
    }
    if (bad) {
      if (wid == 0 && lane == 0) report_error(err, bad, v);
      aux2 = nullptr;
      explicit_pp = false;
    }
    // ---- expand my part ----
    int8_t* orow = out + static_cast<uint64_t>(r) * out_stride_pairs * 2;
    const bool vec_ok = (reinterpret_cast<uintptr_t>(orow) & 15u) == 0;
    for (uint32_t w0 = wbeg; w0 < wend; w0 += 32) {
      const uint32_t w = w0 + lane;
      const uint32_t x = w < wend ? word_at(w) : 0u;
      const uint32_t hm = w < wend ? hets_of(w, x) : 0u;  // hets, one bit per sample at even positions
      const uint32_t cnt = __popc(hm);
      uint32_t tot_h = 0, tot_f = 0;
      const uint32_t hoff = het_base + warp_excl_scan(cnt, lane, &tot_h);
      uint32_t bits = 0;  // this lane's aux2 bits, bit j = het hoff + j
      uint32_t nflag = 0;
      if (aux2 && cnt) {
        bits = load_bits32(aux2, 1ull + hoff) & ((1u << cnt) - 1u);
        nflag = explicit_pp ? __popc(bits) : 0u;
      }
      uint32_t info = bits;  // all-phased form: the bits ARE the phaseinfo
      if (explicit_pp) {
        const uint32_t foff = flag_base + warp_excl_scan(nflag, lane, &tot_f);
        // spread the packed phaseinfo bits of the flagged hets back to het positions
        const uint32_t packed = nflag ? load_bits32(second, foff) : 0u;
        info = 0;
        uint32_t k = 0;
        for (uint32_t f = bits; f; f &= f - 1) {
          info |= ((packed >> k) & 1u) << (__ffs(f) - 1);
          ++k;
        }
      }
      if (w < wend) {
        // 16 samples -> 16 pairs (32 bytes)
        uint32_t o[8];
        uint32_t hj = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          uint32_t word = 0;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint32_t g = (x >> (4 * i + 2 * h)) & 3u;
            uint32_t pair = g == 0 ? 0x0000u : (g == 2 ? 0x0101u : (g == 3 ? 0xF7F7u : 0x0100u));
            // a het of the phase track (with MULTI also the marked ALTx/ALTy calls); only 01 calls show their phase
            const bool track_het = MULTI ? (((hm >> (4 * i + 2 * h)) & 1u) != 0u) : (g == 1);
            if (track_het) {
              if (g == 1 && ((info >> hj) & 1u)) pair = 0x0001u;
              ++hj;
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
            orow[2ull * (s0 + k)] = static_cast<int8_t>(pair & 0xffu);
            orow[2ull * (s0 + k) + 1] = static_cast<int8_t>(pair >> 8);
          }
        }
      }
      het_base += tot_h;
      flag_base += tot_f;
    }
    __syncthreads();  // sm_het / sm_q are rewritten for the next row
  }
}

// out[r, j] = raw[r, idx[j]] as 16-bit pairs (sample subset of a phased row).
__global__ void k_gather_pairs(const uint16_t* __restrict__ raw, uint32_t raw_stride, const uint32_t* __restrict__ idx,
                               uint32_t n_out, uint64_t total, uint16_t* __restrict__ out) {
  for (uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
    const uint64_t r = i / n_out;
    const uint32_t j = static_cast<uint32_t>(i - r * n_out);
    out[i] = raw[r * raw_stride + idx[j]];
  }
}

}  // namespace

void launch_dosage_locate(const RecIndex& ix, const DenseRows& rows, const uint32_t* vlist, uint32_t v0, DosageLoc* loc,
                          uint32_t* rank, uint32_t rank_stride, int* err, cudaStream_t s, LaunchStats* st) {
  if (!rows.nrows) return;
  const uint32_t blocks = (rows.nrows * 32 + kThreads - 1) / kThreads;
  k_dosage_locate<<<blocks, kThreads, 0, s>>>(ix, rows, vlist, v0, loc, rank, rank_stride, err);
  if (st) st->launches++;
}

void launch_dosage_fill(const DenseRows& rows, const DosageLoc* loc, const uint32_t* rank, uint32_t rank_stride,
                        const SubsetDev& sub, uint32_t n_out, double* out, int* err, cudaStream_t s, LaunchStats* st) {
  if (!rows.nrows || !n_out) return;
  uint32_t bx = (n_out + kThreads * 4 - 1) / (kThreads * 4);
  if (bx < 1) bx = 1;
  dim3 grid(bx, rows.nrows);
  k_dosage_fill<<<grid, kThreads, 0, s>>>(rows, loc, rank, rank_stride, sub.subset_idx, n_out, out);
  const uint32_t blocks = (rows.nrows * 32 + kThreads - 1) / kThreads;
  k_dosage_scatter<<<blocks, kThreads, 0, s>>>(rows, loc, sub, n_out, out, err);
  if (st) st->launches += 2;
}

void launch_dosage_stats(const DenseRows& rows, const DosageLoc* loc, const uint32_t* rank, uint32_t rank_stride,
                         const SubsetDev& sub, uint32_t n_out, DosageStats* stats, cudaStream_t s, LaunchStats* st) {
  if (!rows.nrows) return;
  const uint32_t blocks = rows.nrows < 148u * 16 ? rows.nrows : 148u * 16;  // one CTA per row
  k_dosage_stats<<<blocks, kThreads, 0, s>>>(rows, loc, rank, rank_stride, sub, n_out, stats);
  if (st) st->launches++;
}

void launch_expand_phased(const RecIndex& ix, const DenseRows& rows, const uint32_t* vlist, uint32_t v0,
                          const uint32_t* subset_idx, uint32_t n_out, int8_t* raw_scratch, int8_t* out, int* err,
                          cudaStream_t s, LaunchStats* st, uint32_t* mhet_scratch) {
  if (!rows.nrows || !n_out) return;
  uint64_t blocks = rows.nrows;  // one CTA per row
  if (blocks > kPhasedMaxBlocks) blocks = kPhasedMaxBlocks;
  const uint32_t mstride = phased_mhet_stride_words(ix.N);
  if (!subset_idx) {
    if (mhet_scratch) {
      k_expand_phased<true><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix, rows, vlist, v0, ix.N, out, err,
                                                                               mhet_scratch, mstride);
    } else {
      k_expand_phased<false><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix, rows, vlist, v0, ix.N, out, err, nullptr, 0);
    }
    if (st) st->launches++;
    return;
  }
  // subset: expand all raw samples, then gather the kept pairs (phase is a per-sample property)
  const uint32_t raw_stride = (ix.N + 7) & ~7u;  // keeps every scratch row 16-byte aligned
  if (mhet_scratch) {
    k_expand_phased<true><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix, rows, vlist, v0, raw_stride, raw_scratch,
                                                                             err, mhet_scratch, mstride);
  } else {
    k_expand_phased<false><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix, rows, vlist, v0, raw_stride, raw_scratch,
                                                                              err, nullptr, 0);
  }
  launch_gather_pairs(reinterpret_cast<const uint16_t*>(raw_scratch), raw_stride, subset_idx, n_out, rows.nrows,
                      reinterpret_cast<uint16_t*>(out), s, st);
  if (st) st->launches++;
}

size_t phased_mhet_scratch_bytes(uint32_t N) { return static_cast<size_t>(kPhasedMaxBlocks) * phased_mhet_stride_words(N) * 4; }

void launch_gather_pairs(const uint16_t* raw, uint32_t raw_stride, const uint32_t* idx, uint32_t n_out, uint32_t nrows,
                         uint16_t* out, cudaStream_t s, LaunchStats* st) {
  const uint64_t total = static_cast<uint64_t>(nrows) * n_out;
  if (!total) return;
  uint64_t gb = (total + 255) / 256;
  if (gb > 148ull * 32) gb = 148ull * 32;
  k_gather_pairs<<<static_cast<unsigned>(gb), 256, 0, s>>>(raw, raw_stride, idx, n_out, total, out);
  if (st) st->launches++;
}

}  // namespace pdk
