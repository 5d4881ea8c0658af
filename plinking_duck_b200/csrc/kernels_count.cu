// K1/K2/K3: per-variant genotype counts straight from packed .pgen records.
//
// Replaces, for the whole variant range at once, what the reference does one variant at a time
// in PgrGetCounts -> GetBasicGenotypeCounts (pgenlib_read.cc:3624-3739) and the helpers it
// dispatches to (GenoarrbCountFreqs :72-145, GenoarrbCountSubsetFreqs :147-294,
// CountparseDifflistSubset :3453-3511, CountparseOnebitSubset :3516-3593,
// LdSubsetAdjustGenocounts :3294-3384).  HBM-bound: 0.25 B per genotype call.
#include "device_common.cuh"
#include "hwe_device.cuh"
#include "kernels.cuh"

namespace pdk {

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ void report_error(int* err, int code, uint32_t vidx) {
  if (atomicCAS(&err[0], 0, code) == 0) err[1] = static_cast<int>(vidx);
}

// ---------------------------------------------------------------------------------------------
// K1 / K2: fixed-width records.  LANES consecutive lanes own one record; each lane issues
// UNROLL independent 128-bit loads per iteration, so a resident warp keeps UNROLL x 512 B in
// flight and 64 warps/SM cover the HBM latency-bandwidth product with margin.
// ---------------------------------------------------------------------------------------------
template <int LANES, bool MASKED>
__global__ void __launch_bounds__(kThreads)
k_count_fixed(const uint8_t* __restrict__ data, uint64_t first_off, uint32_t width, uint32_t N, uint32_t nrec,
              SubsetDev sub, uint4* __restrict__ out) {
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t ngroups = (gridDim.x * blockDim.x) / LANES;
  const int lane = threadIdx.x & (LANES - 1);
  const uint32_t n_eff = MASKED ? sub.subset_ct : N;
  // The loop condition is warp-uniform (first group of the warp) so that the shuffles always see
  // all 32 lanes; a group past the end in the last iteration just skips its record.
  for (uint64_t r_warp = (tid & ~31u) / LANES, r = tid / LANES; r_warp < nrec; r_warp += ngroups, r += ngroups) {
    Planes c = {0u, 0u, 0u};
    if (r < nrec) {
      const uint8_t* p = data + first_off + r * width;
      if (MASKED) {
        planes_2bit_masked<LANES, 4, true>(p, N, sub.mask2_shifts, sub.mask2_stride, lane, c);
      } else {
        planes_2bit<LANES, 4, true>(p, N, lane, c);
      }
    }
    c.e = group_sum<LANES>(c.e);
    c.o = group_sum<LANES>(c.o);
    c.b = group_sum<LANES>(c.b);
    if (lane == 0 && r < nrec) out[r] = planes_to_counts(c, n_eff);
  }
}

// ---------------------------------------------------------------------------------------------
// K1h: plink_hardy over fixed-width records in ONE kernel -- counts (HBM-bound) and the exact test
// (fp64-bound) run side by side on every SM instead of back to back.  Warps 0..CW-1 of a CTA claim
// chunks of records from a counter and count them exactly like K1 (a whole warp per record); each
// finished record is published with a release store to ready[r].  The other warps walk exact-test
// tables: they claim windows of 32 variants, and every lane that finishes a table takes the next
// variant of the window (hwe_consume), waiting only for that variant's ready flag.
// Counting warps never wait for anything, so the kernel cannot deadlock whatever the residency.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kHardyChunk = 8;   // records per claim of a counting warp
constexpr int kHardyCountWarps = 6;   // of the 8 warps of a CTA; measured best split on B200 (profiles/r01_hardy_split.txt)

__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <bool MASKED, int CW>
__global__ void __launch_bounds__(kThreads, 4)
k_hardy_fixed(const uint8_t* __restrict__ data, uint64_t first_off, uint32_t width, uint32_t N, uint32_t nrec,
              SubsetDev sub, uint4* out, double* __restrict__ out_p, uint32_t midp, uint32_t* __restrict__ queue,
              uint32_t* ready, uint4* __restrict__ host_out) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t warp = threadIdx.x >> 5;
  if (warp >= CW) {
    hwe_consume(queue + 1, nrec, ready, out, out_p, midp != 0);
    return;
  }
  const uint32_t n_eff = MASKED ? sub.subset_ct : N;
  for (;;) {
    uint32_t chunk = 0;
    if (lane == 0) chunk = atomicAdd(queue, 1u);
    chunk = __shfl_sync(0xffffffffu, chunk, 0);
    const uint64_t r0 = static_cast<uint64_t>(chunk) * kHardyChunk;
    if (r0 >= nrec) return;
    const uint32_t r1 = static_cast<uint32_t>(r0 + kHardyChunk < nrec ? r0 + kHardyChunk : nrec);
    for (uint32_t r = static_cast<uint32_t>(r0); r < r1; ++r) {
      const uint8_t* p = data + first_off + static_cast<uint64_t>(r) * width;
      Planes c = {0u, 0u, 0u};
      if (MASKED) {
        planes_2bit_masked<32, 4, true>(p, N, sub.mask2_shifts, sub.mask2_stride, lane, c);
      } else {
        planes_2bit_pipelined<4>(p, N, lane, c);
      }
      c.e = group_sum<32>(c.e);
      c.o = group_sum<32>(c.o);
      c.b = group_sum<32>(c.b);
      if (lane == 0) {
        const uint4 cnt = planes_to_counts(c, n_eff);
        out[r] = cnt;
        st_release_u32(ready + r, 1u);
        if (host_out) host_out[r] = cnt;  // caller's page-locked array, written over PCIe as results appear
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// K3: variable-width records, one warp per record, records claimed from a global counter so
// that 10-byte difflists and 125 KB dense records balance across the chip.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool included(const SubsetDev& sub, uint32_t id) {
  return (sub.include_words[id >> 5] >> (id & 31)) & 1u;
}

template <bool MASKED>
__global__ void __launch_bounds__(kThreads)
k_count_var(RecIndex ix, SubsetDev sub, LdScratch scratch, uint32_t v_begin, uint32_t v_end, uint4* __restrict__ out,
            uint32_t* __restrict__ work_counter, int* __restrict__ err) {
  const int lane = threadIdx.x & 31;
  const uint32_t N = ix.N;
  const uint32_t n_eff = MASKED ? sub.subset_ct : N;
  const uint32_t geno_bytes = (N + 3) >> 2;
  for (;;) {
    uint32_t v = 0;
    if (lane == 0) v = atomicAdd(work_counter, 1u);
    v = __shfl_sync(0xffffffffu, v, 0) + v_begin;
    if (v >= v_end) return;
    const uint64_t f0 = ix.fpos[v];
    const uint64_t f1 = ix.fpos[v + 1];
    const uint8_t* p = ix.data + (f0 - ix.data_fpos);
    const uint8_t* end = ix.data + (f1 - ix.data_fpos);
    const uint32_t main = ix.vrtypes[v] & 7u;
    uint4 res = make_uint4(0, 0, 0, 0);
    bool ok = true;

    if (main == 0) {
      // raw 2-bit track
      if (static_cast<uint64_t>(end - p) < geno_bytes) {
        ok = false;
      } else {
        Planes c = {0u, 0u, 0u};
        if (MASKED) {
          planes_2bit_masked<32, 4, false>(p, N, sub.mask2_shifts, sub.mask2_stride, lane, c);
        } else {
          planes_2bit<32, 4, false>(p, N, lane, c);
        }
        c.e = group_sum<32>(c.e);
        c.o = group_sum<32>(c.o);
        c.b = group_sum<32>(c.b);
        res = planes_to_counts(c, n_eff);
      }
    } else if (main == 5) {
      res.x = n_eff;  // all hom-ref, empty track (pgenlib_read.cc:3685-3690)
    } else if (main == 1) {
      // 1-bit track: common2 byte, ceil(N/8) bitarray, difflist of the two rarer categories.
      const uint32_t bit_bytes = (N + 7) >> 3;
      if (static_cast<uint64_t>(end - p) < static_cast<uint64_t>(bit_bytes) + 1) {
        ok = false;
      } else {
        const uint32_t common2 = p[0];
        const uint32_t code_low = common2 >> 2;
        const uint32_t code_high = (common2 & 3u) + code_low;
        const uint8_t* bits = p + 1;
        uint32_t high_ct = popcount_bits<32>(bits, N, MASKED ? sub.mask1_shifts : nullptr, sub.mask1_stride, lane);
        high_ct = group_sum<32>(high_ct);
        const uint8_t* q = bits + bit_bytes;
        Difflist d;
        ok = parse_difflist(q, end, N, true, d) && code_high <= 3;
        if (ok) {
          uint32_t cnt[4] = {0u, 0u, 0u, 0u};
          if (d.len) {
            if (!MASKED) {
              Planes c = {0u, 0u, 0u};
              planes_2bit<32, 1, false>(d.raregeno, d.len, lane, c);
              c.e = group_sum<32>(c.e);
              c.o = group_sum<32>(c.o);
              c.b = group_sum<32>(c.b);
              const uint4 rc = planes_to_counts(c, d.len);
              cnt[0] = rc.x; cnt[1] = rc.y; cnt[2] = rc.z; cnt[3] = rc.w;
              cnt[code_low] = n_eff - d.len - high_ct;
              cnt[code_high] = high_ct;
            } else {
              uint32_t t0 = 0, t1 = 0, t2 = 0, t3 = 0, hdec = 0;
              ok = difflist_walk(d, N, lane, [&](uint32_t id, uint32_t g, uint32_t) {
                if (included(sub, id)) {
                  t0 += (g == 0); t1 += (g == 1); t2 += (g == 2); t3 += (g == 3);
                  hdec += bit_at(bits, id);
                }
              });
              cnt[0] = group_sum<32>(t0); cnt[1] = group_sum<32>(t1);
              cnt[2] = group_sum<32>(t2); cnt[3] = group_sum<32>(t3);
              high_ct -= group_sum<32>(hdec);
              cnt[code_low] = n_eff - high_ct - cnt[0] - cnt[1] - cnt[2] - cnt[3];
              cnt[code_high] = high_ct;
            }
          } else {
            cnt[code_low] = n_eff - high_ct;
            cnt[code_high] = high_ct;
          }
          res = make_uint4(cnt[0], cnt[1], cnt[2], cnt[3]);
        }
      }
    } else if (main >= 4) {
      // difflist against an all-`common` background (vrtype 4 / 6 / 7)
      const uint32_t common = main & 3u;
      const uint8_t* q = p;
      Difflist d;
      ok = parse_difflist(q, end, N, true, d);
      if (ok) {
        uint32_t cnt[4] = {0u, 0u, 0u, 0u};
        if (d.len) {
          if (!MASKED) {
            Planes c = {0u, 0u, 0u};
            planes_2bit<32, 1, false>(d.raregeno, d.len, lane, c);
            c.e = group_sum<32>(c.e);
            c.o = group_sum<32>(c.o);
            c.b = group_sum<32>(c.b);
            const uint4 rc = planes_to_counts(c, d.len);
            cnt[0] = rc.x; cnt[1] = rc.y; cnt[2] = rc.z; cnt[3] = rc.w;
            cnt[common] = n_eff - d.len;
          } else {
            uint32_t t0 = 0, t1 = 0, t2 = 0, t3 = 0;
            ok = difflist_walk(d, N, lane, [&](uint32_t id, uint32_t g, uint32_t) {
              if (included(sub, id)) {
                t0 += (g == 0); t1 += (g == 1); t2 += (g == 2); t3 += (g == 3);
              }
            });
            cnt[0] = group_sum<32>(t0); cnt[1] = group_sum<32>(t1);
            cnt[2] = group_sum<32>(t2); cnt[3] = group_sum<32>(t3);
            cnt[common] = 0;
            cnt[common] = n_eff - cnt[0] - cnt[1] - cnt[2] - cnt[3];
          }
        } else {
          cnt[common] = n_eff;
        }
        res = make_uint4(cnt[0], cnt[1], cnt[2], cnt[3]);
      }
    } else {
      // LD-compressed (2 / 3): store the signed count delta against the base record; the
      // fix-up kernel adds the base's counts and applies the 0<->2 swap of type 3.
      const uint32_t b = ix.ldbase[v];
      const uint8_t* dense;
      if ((ix.vrtypes[b] & 7u) == 0) {
        dense = ix.data + (ix.fpos[b] - ix.data_fpos);
      } else {
        dense = scratch.dense + static_cast<uint64_t>(ix.base_slot[v] - scratch.slot_lo) * scratch.stride;
      }
      const uint8_t* q = p;
      Difflist d;
      ok = parse_difflist(q, end, N, true, d);
      if (ok && d.len) {
        int32_t i0 = 0, i1 = 0, i2 = 0, i3 = 0;
        ok = difflist_walk(d, N, lane, [&](uint32_t id, uint32_t g, uint32_t) {
          if (!MASKED || included(sub, id)) {
            const uint32_t old = geno_at(dense, id);
            i0 += static_cast<int32_t>(g == 0) - static_cast<int32_t>(old == 0);
            i1 += static_cast<int32_t>(g == 1) - static_cast<int32_t>(old == 1);
            i2 += static_cast<int32_t>(g == 2) - static_cast<int32_t>(old == 2);
            i3 += static_cast<int32_t>(g == 3) - static_cast<int32_t>(old == 3);
          }
        });
        res.x = group_sum<32>(static_cast<uint32_t>(i0));
        res.y = group_sum<32>(static_cast<uint32_t>(i1));
        res.z = group_sum<32>(static_cast<uint32_t>(i2));
        res.w = group_sum<32>(static_cast<uint32_t>(i3));
      }
    }
    if (!ok) {
      if (lane == 0) report_error(err, kErrMalformed, v);
      res = make_uint4(0, 0, 0, 0);
    }
    if (lane == 0) out[v - v_begin] = res;
  }
}

__global__ void k_ld_fixup(RecIndex ix, uint32_t v_begin, uint32_t v_end, uint4* __restrict__ out) {
  const uint32_t v = v_begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= v_end) return;
  const uint32_t main = ix.vrtypes[v] & 7u;
  if (main != 2 && main != 3) return;
  const uint32_t b = ix.ldbase[v];
  if (b < v_begin) return;  // caller guarantees v_begin is a base; defensive
  const uint4 base = out[b - v_begin];
  const uint4 d = out[v - v_begin];
  uint4 r = make_uint4(base.x + d.x, base.y + d.y, base.z + d.z, base.w + d.w);
  if (main == 3) {
    const uint32_t t = r.x;
    r.x = r.z;
    r.z = t;
  }
  out[v - v_begin] = r;
}

}  // namespace

namespace {

// One CTA per scratch slot: dense copy of a non-type-0 record that LD records refer to.  The warps of the CTA share
// the record (decode_nonld_dense_team): bases are often common variants stored as one-bit records with difflists of
// thousands of entries, which one warp walks in ~200 us (profiles/r02p_*: this kernel was 80 % of pd_counts on the
// LD-compressed fileset with a warp per slot).
__global__ void __launch_bounds__(kThreads)
k_decode_bases(RecIndex ix, const uint32_t* __restrict__ slot_vidx, LdScratch scratch, uint32_t slot_hi,
               int* __restrict__ err) {
  const uint32_t w = blockIdx.x;
  const int lane = threadIdx.x & 31;
  const uint32_t slot = scratch.slot_lo + w;
  if (slot > slot_hi) return;  // uniform for the CTA
  const uint32_t v = slot_vidx[slot];
  const uint8_t* p = ix.data + (ix.fpos[v] - ix.data_fpos);
  const uint8_t* end = ix.data + (ix.fpos[v + 1] - ix.data_fpos);
  uint32_t* dst = reinterpret_cast<uint32_t*>(scratch.dense + static_cast<uint64_t>(w) * scratch.stride);
  const bool ok = decode_nonld_dense_team<true>(ix.vrtypes[v] & 7u, p, end, ix.N, dst, lane, threadIdx.x >> 5, kThreads / 32);
  if (!ok && lane == 0) report_error(err, kErrMalformed, v);
}

// Shifted mask copies.  For shift s in [0,16): copy_s[j] = mask[j - s] (0 outside the mask), so
// that the 16-byte chunk c of a record whose first byte sits s bytes into an aligned chunk lines
// up with chunk c of copy_s.  Bytes past the last sample are zero, which is what clips the tail.
__global__ void k_build_masks(const uint32_t* __restrict__ include_words, uint32_t N, uint8_t* __restrict__ mask2,
                              uint32_t stride2, uint8_t* __restrict__ mask1, uint32_t stride1) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;  // byte index inside a copy
  const uint32_t s = blockIdx.y;                              // shift
  const uint8_t* inc = reinterpret_cast<const uint8_t*>(include_words);
  const uint32_t bytes1 = (N + 7) >> 3;
  const uint32_t bytes2 = (N + 3) >> 2;
  if (j < stride2) {
    uint8_t val = 0;
    if (j >= s && j - s < bytes2) {
      const uint32_t k = j - s;            // covers samples 4k .. 4k+3
      const uint32_t nib = (inc[k >> 1] >> (4 * (k & 1))) & 15u;
      val = static_cast<uint8_t>((nib & 1u) | ((nib & 2u) << 1) | ((nib & 4u) << 2) | ((nib & 8u) << 3));
    }
    mask2[static_cast<size_t>(s) * stride2 + j] = val;
  }
  if (j < stride1) {
    uint8_t val = 0;
    if (j >= s && j - s < bytes1) val = inc[j - s];
    mask1[static_cast<size_t>(s) * stride1 + j] = val;
  }
}

// Exclusive prefix popcount per 32-bit word; single block, sequential over chunks (N <= 2^31 bits
// is at most 2^26 words, but real subsets are built once per query so simplicity wins).
__global__ void k_cum_popcount(const uint32_t* __restrict__ words, uint32_t nwords, uint32_t* __restrict__ cum) {
  __shared__ uint32_t warp_tot[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (uint32_t base = 0; base < nwords; base += blockDim.x) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t pc = (i < nwords) ? __popc(words[i]) : 0u;
    uint32_t incl = pc;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += t;
    }
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    uint32_t before = carry;
    for (int k = 0; k < wid; ++k) before += warp_tot[k];
    if (i < nwords) cum[i] = before + incl - pc;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = before + incl;
    __syncthreads();
  }
}

int pick_lanes(uint32_t N) {
  const uint64_t chunks = (2ull * N + 127) / 128 + 1;
  if (chunks >= 96) return 32;
  if (chunks >= 48) return 16;
  if (chunks >= 24) return 8;
  return 4;
}

template <int LANES>
void launch_fixed_t(const RecIndex& ix, const SubsetDev& sub, uint32_t v_begin, uint32_t nrec, uint4* out,
                    cudaStream_t s, int sms) {
  const uint64_t first_off = ix.fixed_start + static_cast<uint64_t>(v_begin) * ix.fixed_width - ix.data_fpos;
  // Persistent grid: 8 CTAs of 256 threads per SM (64 resident warps), a multiple of the SM count.
  const uint64_t groups_needed = nrec;
  uint64_t blocks = (groups_needed * LANES + kThreads - 1) / kThreads;
  const uint64_t max_blocks = static_cast<uint64_t>(sms) * 8;
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks == 0) blocks = 1;
  if (sub.mask2_shifts) {
    k_count_fixed<LANES, true><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix.data, first_off, ix.fixed_width,
                                                                                 ix.N, nrec, sub, out);
  } else {
    k_count_fixed<LANES, false><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix.data, first_off, ix.fixed_width,
                                                                                  ix.N, nrec, sub, out);
  }
}


}  // namespace

namespace {
__global__ void __launch_bounds__(256) k_bed_to_pgen(uint8_t* __restrict__ data, uint32_t width, uint32_t N, uint64_t total) {
  const uint32_t tail = N & 3u;
  for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
    const uint32_t b = data[i];
    // per 2-bit code (h, l): h' = ~h, l' = l ^ h
    uint32_t o = ((~b) & 0xAAu) | ((b ^ (b >> 1)) & 0x55u);
    if (tail && (i % width) == width - 1) o &= (1u << (2 * tail)) - 1u;
    data[i] = static_cast<uint8_t>(o);
  }
}
}  // namespace

void launch_bed_to_pgen(uint8_t* data, uint64_t first_off, uint32_t width, uint32_t N, uint32_t nrec, cudaStream_t s,
                        LaunchStats* st) {
  if (!nrec) return;
  const uint64_t total = static_cast<uint64_t>(nrec) * width;
  uint64_t blocks = (total + 255) / 256;
  const uint64_t max_blocks = static_cast<uint64_t>(sm_of(st)) * 16;
  if (blocks > max_blocks) blocks = max_blocks;
  k_bed_to_pgen<<<static_cast<unsigned>(blocks), 256, 0, s>>>(data + first_off, width, N, total);
  if (st) st->launches++;
}

// Enough record bytes per variant that counting, not the exact test, is the long pole.
bool hardy_fused_applicable(uint32_t N) { return pick_lanes(N) == 32; }

void launch_build_masks(const uint32_t* include_words, uint32_t N, uint8_t* mask2_shifts, uint32_t mask2_stride,
                        uint8_t* mask1_shifts, uint32_t mask1_stride, cudaStream_t s, LaunchStats* st) {
  const uint32_t maxstride = mask2_stride > mask1_stride ? mask2_stride : mask1_stride;
  dim3 grid((maxstride + 255) / 256, 16);
  k_build_masks<<<grid, 256, 0, s>>>(include_words, N, mask2_shifts, mask2_stride, mask1_shifts, mask1_stride);
  if (st) st->launches++;
}

void launch_cum_popcount(const uint32_t* include_words, uint32_t nwords, uint32_t* cum, cudaStream_t s,
                         LaunchStats* st) {
  k_cum_popcount<<<1, 1024, 0, s>>>(include_words, nwords, cum);
  if (st) st->launches++;
}

void launch_count_fixed(const RecIndex& ix, const SubsetDev& sub, uint32_t v_begin, uint32_t v_end, uint4* out,
                        cudaStream_t s, LaunchStats* st) {
  const uint32_t nrec = v_end - v_begin;
  if (!nrec) return;
  switch (pick_lanes(ix.N)) {
    case 32: launch_fixed_t<32>(ix, sub, v_begin, nrec, out, s, sm_of(st)); break;
    case 16: launch_fixed_t<16>(ix, sub, v_begin, nrec, out, s, sm_of(st)); break;
    case 8: launch_fixed_t<8>(ix, sub, v_begin, nrec, out, s, sm_of(st)); break;
    default: launch_fixed_t<4>(ix, sub, v_begin, nrec, out, s, sm_of(st)); break;
  }
  if (st) st->launches++;
}

bool launch_hardy_fixed(const RecIndex& ix, const SubsetDev& sub, uint32_t v_begin, uint32_t v_end, uint4* out,
                        double* out_p, uint32_t midp, uint32_t* queue, uint32_t* ready, cudaStream_t s,
                        LaunchStats* st, uint4* host_out) {
  const uint32_t nrec = v_end - v_begin;
  if (!nrec) return true;
  if (!hardy_fused_applicable(ix.N)) return false;
  const uint64_t first_off = ix.fixed_start + static_cast<uint64_t>(v_begin) * ix.fixed_width - ix.data_fpos;
  cudaMemsetAsync(queue, 0, kHweQueueWords * sizeof(uint32_t), s);
  cudaMemsetAsync(ready, 0, static_cast<size_t>(nrec) * sizeof(uint32_t), s);
  // One resident wave: 4 CTAs per SM (launch bounds); 6 warps of a CTA count, 2 test.
  uint64_t blocks = (static_cast<uint64_t>(nrec) + kHardyChunk * kHardyCountWarps - 1) / (kHardyChunk * kHardyCountWarps);
  const uint64_t max_blocks = static_cast<uint64_t>(sm_of(st)) * 4;
  if (blocks > max_blocks) blocks = max_blocks;
  if (sub.mask2_shifts) {
    k_hardy_fixed<true, kHardyCountWarps><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(
        ix.data, first_off, ix.fixed_width, ix.N, nrec, sub, out, out_p, midp, queue, ready, host_out);
  } else {
    k_hardy_fixed<false, kHardyCountWarps><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(
        ix.data, first_off, ix.fixed_width, ix.N, nrec, sub, out, out_p, midp, queue, ready, host_out);
  }
  if (st) st->launches++;
  launch_hwe_scaled(out, queue, midp, out_p, false, s, st);
  return true;
}

void launch_decode_bases(const RecIndex& ix, const uint32_t* slot_vidx, const LdScratch& scratch, uint32_t slot_hi,
                         int* err, cudaStream_t s, LaunchStats* st) {
  if (slot_hi < scratch.slot_lo) return;
  const uint32_t nslots = slot_hi - scratch.slot_lo + 1;
  k_decode_bases<<<nslots, kThreads, 0, s>>>(ix, slot_vidx, scratch, slot_hi, err);
  if (st) st->launches++;
}

void launch_count_var(const RecIndex& ix, const SubsetDev& sub, const LdScratch& scratch, uint32_t v_begin,
                      uint32_t v_end, uint4* out, uint32_t* work_counter, int* err, cudaStream_t s, LaunchStats* st) {
  const uint32_t nrec = v_end - v_begin;
  if (!nrec) return;
  cudaMemsetAsync(work_counter, 0, sizeof(uint32_t), s);
  uint64_t blocks = (static_cast<uint64_t>(nrec) * 32 + kThreads - 1) / kThreads;
  const uint64_t max_blocks = static_cast<uint64_t>(sm_of(st)) * 8;
  if (blocks > max_blocks) blocks = max_blocks;
  if (sub.mask2_shifts) {
    k_count_var<true><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix, sub, scratch, v_begin, v_end, out,
                                                                        work_counter, err);
  } else {
    k_count_var<false><<<static_cast<unsigned>(blocks), kThreads, 0, s>>>(ix, sub, scratch, v_begin, v_end, out,
                                                                         work_counter, err);
  }
  if (st) st->launches++;
  const uint32_t fb = (nrec + 255) / 256;
  k_ld_fixup<<<fb, 256, 0, s>>>(ix, v_begin, v_end, out);
  if (st) st->launches++;
}

}  // namespace pdk
