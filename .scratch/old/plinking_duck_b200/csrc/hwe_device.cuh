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
//    of a warp pull variants from a warp-local queue (hwe_warp_queue): a lane that finishes its
//    table takes the next variant while its neighbours keep walking theirs.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pdk {

constexpr double kHweTieEps = 3.637978807091713e-12;  // 2^-38: relative window for "same probability"

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
      fin = (__double2hiint(dn.t) | __double2hiint(up.t)) == 0;
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

// p-value of a finished table.
__device__ __forceinline__ double hwe_finish(const HweLane& s, bool midp) {
  double q = 0.0;  // center overflowed: the observed table is > 1e308 times less likely than the mode
  if (s.center < 1.7e308) {
    double tail = s.tail;
    if (midp) tail -= 0.5 * static_cast<double>(s.ties);
    q = tail / (s.tail + s.center);
    if (q < 0.0) q = 0.0;
    if (q > 1.0) q = 1.0;
  }
  return q;
}

// Whole-warp call: p-values of items [0, n) -- get(i) returns the item's (hom_ref, het, hom_alt, *)
// counts, put(i, p) stores its result.  All 32 lanes must call it with the same n.
template <class Get, class Put>
__device__ __forceinline__ void hwe_warp_queue(uint32_t n, bool midp, Get get, Put put) {
  const uint32_t lane = threadIdx.x & 31u;
  uint32_t next = 0;      // warp-uniform: first item nobody has taken yet
  uint32_t mine = ~0u;    // item this lane is walking
  HweLane s;
  s.dn.t = s.up.t = 0.0;
  for (;;) {
    const uint32_t idle = __ballot_sync(0xffffffffu, mine == ~0u);
    if (idle && next < n) {
      if (mine == ~0u) {
        const uint32_t i = next + __popc(idle & ((1u << lane) - 1u));
        if (i < n) {
          const uint4 cnt = get(i);
          double p;
          if (hwe_begin(cnt.y, cnt.x, cnt.z, midp, s, &p)) {
            put(i, p);
          } else {
            mine = i;
          }
        }
      }
      next += __popc(idle);
      continue;  // lanes whose item was trivial try again
    }
    if (idle == 0xffffffffu) break;  // queue empty and nobody walking
    if (hwe_walk(s, mine != ~0u, 0x7fffffff, midp)) {
      put(mine, hwe_finish(s, midp));
      mine = ~0u;
    }
  }
}

__device__ __forceinline__ uint32_t hwe_ld_acquire(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Whole-warp call, any number of warps of a grid at once: exact tests of variants [0, n), claimed
// in windows of 32 from the device counter *next_window.  ready (optional): variant v may only be
// read once ready[v] != 0 (release-published by the warp that counted it); lanes whose variant is
// not ready yet keep polling while the others walk.
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
          out_p[mine] = pv;
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
      out_p[mine] = hwe_finish(s, midp);
      mine = ~0u;
    }
  }
}

}  // namespace pdk
