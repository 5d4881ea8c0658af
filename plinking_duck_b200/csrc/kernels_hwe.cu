// K8: exact Hardy-Weinberg equilibrium test, one thread per variant, fp64.
//
// Replaces HweExactTestAutosomal -> plink2::HweLnP + LnPToPvalue (src/plink_hardy.cpp:52-79;
// plink2_stats.cc:1784-2166) for the autosomal / diploid-stratum test.  This is the exact test of
// Wigginton, Cutler & Abecasis (2005): conditional on the allele counts, the probability of k
// heterozygotes relative to the observed table follows the two-term recurrences
//     P(k-2)/P(k) = k (k-1) / (4 (r+1)(c+1)),      P(k+2)/P(k) = 4 r c / ((k+2)(k+1)),
// (r, c = rarer / commoner homozygote counts at k).  The p-value is the mass of all tables no
// more likely than the observed one over the total mass; the mid-p variant (Graffelman & Moreno
// 2013) removes half of the mass tied with the observed table.
//
// Unlike plink2, which reaches the far tail through log-factorial (Lanczos) evaluations, the
// whole distribution is walked with the recurrences in units of the observed table's
// probability, so no log-gamma rounding enters: agreement with HweLnP is limited by the
// reference's own Lfact noise (ulp(Lfact(2N)) relative, ~5e-10 at N = 1e5), see DESIGN.md.
#include "kernels.cuh"

namespace pdk {

namespace {

constexpr double kTieEps = 3.637978807091713e-12;  // 2^-38: relative window for "same probability"

__device__ __forceinline__ double hwe_pvalue(uint32_t obs_hets, uint32_t hom_a, uint32_t hom_b, bool midp) {
  const uint32_t n = obs_hets + hom_a + hom_b;
  if (n == 0) return 1.0;  // src/plink_hardy.cpp:73-75
  const uint32_t r0 = hom_a < hom_b ? hom_a : hom_b;
  const uint32_t c0 = hom_a < hom_b ? hom_b : hom_a;
  const uint64_t rare = 2ull * r0 + obs_hets;
  if (rare < 2) return midp ? 0.5 : 1.0;

  double tail = 1.0;    // tables with probability <= observed (observed included), in observed units
  double center = 0.0;  // tables strictly more likely than the observed one
  uint32_t ties = 1;

  // Direction 1: fewer heterozygotes.
  {
    double k = static_cast<double>(obs_hets);
    double r = static_cast<double>(r0);
    double c = static_cast<double>(c0);
    double t = 1.0;
    while (k > 1.5) {
      r += 1.0;
      c += 1.0;
      t *= (k * (k - 1.0)) / (4.0 * r * c);
      k -= 2.0;
      if (t > 1.0 + kTieEps) {
        center += t;
        if (!(center < 1.7e308)) return 0.0;
      } else {
        if (t >= 1.0 - kTieEps) ++ties;
        const double before = tail;
        tail += t;
        if (tail <= before) break;  // further terms cannot change the sum
      }
    }
  }
  // Direction 2: more heterozygotes.
  {
    double k = static_cast<double>(obs_hets);
    double r = static_cast<double>(r0);
    double c = static_cast<double>(c0);
    double t = 1.0;
    while (r > 0.5) {
      k += 2.0;
      t *= (4.0 * r * c) / (k * (k - 1.0));
      r -= 1.0;
      c -= 1.0;
      if (t > 1.0 + kTieEps) {
        center += t;
        if (!(center < 1.7e308)) return 0.0;
      } else {
        if (t >= 1.0 - kTieEps) ++ties;
        const double before = tail;
        tail += t;
        if (tail <= before) break;
      }
    }
  }
  const double total = tail + center;
  if (midp) tail -= 0.5 * static_cast<double>(ties);
  double p = tail / total;
  if (p < 0.0) p = 0.0;
  if (p > 1.0) p = 1.0;
  return p;
}

__global__ void __launch_bounds__(128) k_hwe(const uint4* __restrict__ counts, uint32_t n, uint32_t midp,
                                             double* __restrict__ out_p) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 c = counts[i];
  out_p[i] = hwe_pvalue(c.y, c.x, c.z, midp != 0);
}

}  // namespace

void launch_hwe(const uint4* counts, uint32_t n, uint32_t midp, double* out_p, cudaStream_t s, LaunchStats* st) {
  if (!n) return;
  k_hwe<<<(n + 127) / 128, 128, 0, s>>>(counts, n, midp, out_p);
  if (st) st->launches++;
}

}  // namespace pdk
