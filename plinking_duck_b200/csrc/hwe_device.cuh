// Device-side exact Hardy-Weinberg test shared by K8 (k_hwe) and the fused count+HWE kernel K1h.
//
// Replaces HweExactTestAutosomal -> plink2::HweLnP + LnPToPvalue (src/plink_hardy.cpp:52-79;
// plink2_stats.cc:1784-2166).  Exact test of Wigginton, Cutler & Abecasis (2005): conditional on
// the allele counts, the probability of k heterozygotes relative to the observed table follows
//     P(k-2)/P(k) = k (k-1) / (4 (r+1)(c+1)),      P(k+2)/P(k) = 4 r c / ((k+2)(k+1)),
// (r, c = rarer / commoner homozygote counts at k).  The p-value is the mass of all tables no more
// likely than the observed one over the total mass; mid-p (Graffelman & Moreno 2013) removes half
// of the mass tied with the observed table.  The whole distribution is walked with the
// recurrences in units of the observed table's probability (no log-gamma), see DESIGN.md.
//
// How the walk is made cheap (it is ~1000 steps per variant at N = 1e5, fp64-pipe bound):
//  * Both ratios are A_j / B_j with A and B quadratic in the step index j, and in BOTH directions
//    their second differences are +8.  So a step is  A += dA, dA += 8, B += dB, dB += 8  (exact:
//    integers below 2^53) and the direction is data, not control flow.
//  * 1/B comes from the hardware seed (rcp.approx, MUFU.RCP64H) plus two Newton steps instead of
//    an IEEE division: 4 DFMA, relative error < 2^-52; the walk accumulates ~2 ulp per step, far
//    inside the stated tolerance (tests/test_gpu_counts.py::hwe_tolerance).
//  * The two directions are independent dependency chains; they are stepped together, which halves
//    the number of loop trips and gives every lane two fp64 chains to overlap.
//  * The walk length differs by ~4x between variants (it scales with sqrt(N p q)), so the lanes
//    of a warp pull variants from a window of 32 claimed by the warp (hwe_consume): a lane that finishes its
//    table takes the next variant while its neighbours keep walking theirs.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace pdk {

constexpr double kHweTieEps = 3.637978807091713e-12;  // 2^-38: relative window for "same probability"
constexpr int kHweCenterHiLimit = 0x7de00000;         // high word of 2^991

struct HweDir {
  double t;              // probability of the current table relative to the observed one (0 once finished)
  double A, dA, B, dB;   // next ratio A / B and its first differences
};

struct HweLane {
  HweDir dn, up;         // towards fewer / more heterozygotes: two independent chains, walked together
  double tail, center;   // mass <= observed (observed included) / mass > observed
  uint32_t ties;
};

__device__ __forceinline__ double hwe_rcp(double b) {
  double x;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(b));
  double e = fma(-b, x, 1.0);
  x = fma(x, e, x);
  e = fma(-b, x, 1.0);
  x = fma(x, e, x);
  return x;
}

// Starts a table.  Returns true when the p-value is already known (*p).
__device__ __forceinline__ bool hwe_begin(uint32_t obs_hets, uint32_t hom_a, uint32_t hom_b, bool midp, HweLane& s,
                                          double* p) {
  const uint32_t n = obs_hets + hom_a + hom_b;
  if (n == 0) {  // src/plink_hardy.cpp:73-75
    *p = 1.0;
    return true;
  }
  const uint32_t r0 = hom_a < hom_b ? hom_a : hom_b;
  const uint32_t c0 = hom_a < hom_b ? hom_b : hom_a;
  if (2ull * r0 + obs_hets < 2) {
    *p = midp ? 0.5 : 1.0;
    return true;
  }
  const double k = static_cast<double>(obs_hets), r = static_cast<double>(r0), c = static_cast<double>(c0);
  // fewer hets: step j uses k = k0 - 2j, r = r0 + j, c = c0 + j:  A = k (k-1),  B = 4 (r+1)(c+1).
  // A reaches exactly 0 at k = 1 or 0, which zeroes t for good: the end of the range needs no counter.
  s.dn.A = k * (k - 1.0);
  s.dn.dA = 6.0 - 4.0 * k;
  s.dn.B = 4.0 * (r + 1.0) * (c + 1.0);
  s.dn.dB = 4.0 * (r + c) + 12.0;
  s.dn.t = obs_hets >= 2 ? 1.0 : 0.0;
  // more hets: step j uses r = r0 - j, c = c0 - j, k = k0 + 2j:  A = 4 r c (0 at r = 0),  B = (k+2)(k+1)
  s.up.A = 4.0 * r * c;
  s.up.dA = 4.0 - 4.0 * (r + c);
  s.up.B = (k + 2.0) * (k + 1.0);
  s.up.dB = 4.0 * k + 10.0;
  s.up.t = r0 ? 1.0 : 0.0;
  s.tail = 1.0;
  s.center = 0.0;
  s.ties = 1;
  return false;
}

// One step of one direction.  A finished direction has t = 0: stepping it again adds exact zeros,
// so the two directions need no separate control flow.
template <bool MIDP>
__device__ __forceinline__ void hwe_dir_step(HweDir& d, double& tail, double& center, uint32_t& ties) {
  const double t = d.t * (d.A * hwe_rcp(d.B));
  d.A += d.dA;
  d.dA += 8.0;
  d.B += d.dB;
  d.dB += 8.0;
  const bool big = t > 1.0 + kHweTieEps;
  const double ts = big ? 0.0 : t;
  center += t - ts;  // t or 0, exactly
  const double nt = tail + ts;
  if (MIDP) ties += (!big && t >= 1.0 - kHweTieEps) ? 1u : 0u;
  // nt <= tail: this term (and every later one: they only shrink from here) cannot change the sum;
  // also true once A has hit zero at the end of the range.
  // (!(nt > tail) also ends a chain whose t overflowed to inf and then met A = 0: NaN.)
  d.t = (!big && !(nt > tail)) ? 0.0 : t;
  tail = nt;
}

// Steps every lane that has `on` set until one of them has finished both directions or max_steps
// have passed.  Whole-warp call; returns this lane's finished flag.  The loop body is the fp64 hot
// spot of plink_hardy: two independent chains of 14 fp64 instructions, no branches.
template <bool MIDP>
__device__ __forceinline__ bool hwe_walk_t(HweLane& s, bool on, int max_steps) {
  HweDir dn = s.dn, up = s.up;
  double tail = s.tail, center = s.center;
  uint32_t ties = s.ties;
  // t is 0 or a normal number >= 2^-53 relative to tail >= 1 while a direction is live
  bool fin = on && (__double2hiint(dn.t) | __double2hiint(up.t)) == 0;
#pragma unroll 1
  for (int it = 0; it < max_steps; ++it) {
    if (__any_sync(0xffffffffu, fin)) break;
    if (on) {
      hwe_dir_step<MIDP>(dn, tail, center, ties);
      hwe_dir_step<MIDP>(up, tail, center, ties);
      // also leave when the centre mass nears the top of the fp64 range (>= 2^991, inf and NaN included):
      // the table is finished by hwe_scaled_lnp instead
      fin = ((__double2hiint(dn.t) | __double2hiint(up.t)) == 0) | (__double2hiint(center) >= kHweCenterHiLimit);
    }
  }
  s.dn = dn;
  s.up = up;
  s.tail = tail;
  s.center = center;
  s.ties = ties;
  return fin;
}

__device__ __forceinline__ bool hwe_walk(HweLane& s, bool on, int max_steps, bool midp) {
  return midp ? hwe_walk_t<true>(s, on, max_steps) : hwe_walk_t<false>(s, on, max_steps);
}

// A table whose observed cell is more than ~2^990 times less likely than the mode (p below ~1e-298): the plain
// walk above overflows.  This one carries the running term and the centre mass as mantissa x 2^(512 e), so
// it returns ln p however small p is (HweLnP returns -40000 just the same, plink2_stats.cc:1845-1849
// "A p-value of 1e-400 may be worth distinguishing from 1e-40000").  One lane, both directions in turn; rare.
__device__ __forceinline__ double hwe_scaled_lnp(uint32_t obs_hets, uint32_t hom_a, uint32_t hom_b, bool midp) {
  const uint32_t r0 = hom_a < hom_b ? hom_a : hom_b;
  const uint32_t c0 = hom_a < hom_b ? hom_b : hom_a;
  const double k = static_cast<double>(obs_hets), r = static_cast<double>(r0), c = static_cast<double>(c0);
  constexpr double kUp = 1.3407807929942597e154, kDown = 7.458340731200207e-155;  // 2^512, 2^-512
  double tail = 1.0, cm = 0.0;
  int ce = 0;  // centre mass = cm * 2^(512 ce)
  uint32_t ties = 1;
#pragma unroll 1
  for (int dir = 0; dir < 2; ++dir) {
    double A, dA, B, dB, t;
    if (dir == 0) {
      A = k * (k - 1.0); dA = 6.0 - 4.0 * k; B = 4.0 * (r + 1.0) * (c + 1.0); dB = 4.0 * (r + c) + 12.0;
      t = obs_hets >= 2 ? 1.0 : 0.0;
    } else {
      A = 4.0 * r * c; dA = 4.0 - 4.0 * (r + c); B = (k + 2.0) * (k + 1.0); dB = 4.0 * k + 10.0;
      t = r0 ? 1.0 : 0.0;
    }
    int te = 0;  // term = t * 2^(512 te)
#pragma unroll 1
    while (t != 0.0) {
      t *= A * hwe_rcp(B);
      A += dA; dA += 8.0; B += dB; dB += 8.0;
      if (t == 0.0) break;  // end of the range (A reached 0)
      if (t >= kUp) {
        t *= kDown;
        ++te;
      } else if (te > 0 && t < kDown) {
        t *= kUp;
        --te;
      }
      if (te > 0 || t > 1.0 + kHweTieEps) {
        if (te == ce) {
          cm += t;
        } else if (te > ce) {
          const int d = te - ce;
          cm = d == 1 ? fma(cm, kDown, t) : (d == 2 ? fma(cm * kDown, kDown, t) : t);
          ce = te;
        } else {
          const int d = ce - te;
          if (d == 1) cm = fma(t, kDown, cm);
          else if (d == 2) cm = fma(t * kDown, kDown, cm);
        }
        if (cm >= kUp) {
          cm *= kDown;
          ++ce;
        }
      } else {
        ties += (t >= 1.0 - kHweTieEps) ? 1u : 0u;
        const double nt = tail + t;
        if (!(nt > tail)) break;  // unimodal: every later term is smaller still
        tail = nt;
      }
    }
  }
  double num = tail;
  if (midp) num -= 0.5 * static_cast<double>(ties);
  if (!(num > 0.0)) return -CUDART_INF;
  double total = cm;
  if (ce == 0) total += tail;
  else if (ce == 1) total = fma(tail, kDown, cm);
  else if (ce == 2) total = fma(tail * kDown, kDown, cm);
  const double lnp = log(num / total) - static_cast<double>(ce) * 354.89135644669199843;  // 512 ln 2
  return lnp < 0.0 ? lnp : 0.0;
}

// Stores the result of a table hwe_walk reported finished: p clamped to [0,1] (LnPToPvalue,
// src/plink_hardy.cpp:52-64) or ln p.  A table that left the walk because its centre mass neared the top of
// the fp64 range is appended to the overflow list instead (ovf[0] = count, ovf[2 + i] = item) and finished by
// k_hwe_scaled right after this kernel -- keeping that rare, long path out of the hot kernels' register budget.
template <bool LNP>
__device__ __forceinline__ void hwe_store(const HweLane& s, bool midp, uint32_t item, double* out_p, uint32_t* ovf) {
  if (__double2hiint(s.center) >= kHweCenterHiLimit) {
    ovf[2u + atomicAdd(ovf, 1u)] = item;
    return;
  }
  double tail = s.tail;
  if (midp) tail -= 0.5 * static_cast<double>(s.ties);
  double q = tail / (s.tail + s.center);
  if (q < 0.0) q = 0.0;
  if (q > 1.0) q = 1.0;
  out_p[item] = LNP ? log(q) : q;
}

__device__ __forceinline__ uint32_t hwe_ld_acquire(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Whole-warp call, any number of warps of a grid at once: exact tests of variants [0, n), claimed
// in windows of 32 from the device counter *next_window; next_window + 1 is the overflow list of hwe_store
// (count, pad, then up to n items).  ready (optional): variant v may only be
// read once ready[v] != 0 (release-published by the warp that counted it); lanes whose variant is
// not ready yet keep polling while the others walk.
template <bool LNP = false>
__device__ __forceinline__ void hwe_consume(uint32_t* next_window, uint32_t n, const uint32_t* ready,
                                            const uint4* counts, double* out_p, bool midp) {
  const uint32_t lane = threadIdx.x & 31u;
  uint32_t win_base = 0, win_n = 0, taken = 0;  // warp-uniform
  bool exhausted = false;
  uint32_t mine = ~0u;
  bool started = false;
  HweLane s;
  s.dn.t = s.up.t = 0.0;
  for (;;) {
    const uint32_t idle = __ballot_sync(0xffffffffu, mine == ~0u);
    if (idle && !exhausted) {
      if (taken == win_n) {
        uint32_t b = 0;
        if (lane == 0) b = atomicAdd(next_window, 32u);
        b = __shfl_sync(0xffffffffu, b, 0);
        win_base = b;
        win_n = b < n ? (n - b < 32u ? n - b : 32u) : 0u;
        taken = 0;
        exhausted = win_n == 0;
      }
      if (!exhausted) {
        const uint32_t avail = win_n - taken;
        const uint32_t rank = __popc(idle & ((1u << lane) - 1u));
        if (mine == ~0u && rank < avail) {
          mine = win_base + taken + rank;
          started = false;
        }
        const uint32_t want = __popc(idle);
        taken += want < avail ? want : avail;
      }
    }
    if (__ballot_sync(0xffffffffu, mine != ~0u) == 0u) {
      if (exhausted) return;
      continue;
    }
    if (mine != ~0u && !started) {
      if (!ready || hwe_ld_acquire(ready + mine)) {
        const uint4 cnt = __ldcg(counts + mine);
        double pv;
        if (hwe_begin(cnt.y, cnt.x, cnt.z, midp, s, &pv)) {
          out_p[mine] = LNP ? log(pv) : pv;
          mine = ~0u;
        } else {
          started = true;
        }
      }
    }
    if (__ballot_sync(0xffffffffu, mine != ~0u && started) == 0u) {
      if (ready) __nanosleep(256);  // every table of this warp still waits for its counts
      continue;
    }
    // walk until a lane needs attention (bounded when lanes still waiting for counts must be polled again)
    if (hwe_walk(s, mine != ~0u && started, ready ? 64 : 0x7fffffff, midp)) {
      hwe_store<LNP>(s, midp, mine, out_p, next_window + 1);
      mine = ~0u;
    }
  }
}

}  // namespace pdk
