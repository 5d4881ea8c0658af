// K8: exact Hardy-Weinberg equilibrium test in fp64; the test itself is in hwe_device.cuh.
//
// Replaces HweExactTestAutosomal -> plink2::HweLnP + LnPToPvalue (src/plink_hardy.cpp:52-79;
// plink2_stats.cc:1784-2166) for the autosomal / diploid-stratum test.  One warp owns a contiguous
// run of variants and its lanes pull tables from it as they finish (hwe_consume), so a warp is
// not held up by its slowest table.
#include "hwe_device.cuh"
#include "kernels.cuh"

namespace pdk {

namespace {

constexpr double kTieEps = kHweTieEps;
// Persistent grid; every warp pulls windows of 32 variants from the device counter *queue.
// LNP: store ln p (what HweLnP returns, unclamped by exp) instead of p.
template <bool LNP>
__global__ void __launch_bounds__(256, 4) k_hwe(const uint4* __restrict__ counts, uint32_t n, uint32_t midp,
                                                double* __restrict__ out_p, uint32_t* __restrict__ queue) {
  hwe_consume<LNP>(queue + 1, n, nullptr, counts, out_p, midp != 0);
}

// The tables hwe_store set aside (observed cell > 2^990 times less likely than the mode): one lane each.
template <bool LNP>
__global__ void __launch_bounds__(32) k_hwe_scaled(const uint4* __restrict__ counts, const uint32_t* __restrict__ ovf,
                                                   uint32_t midp, double* __restrict__ out_p) {
  const uint32_t n = ovf[0];
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t item = ovf[2u + i];
    const uint4 cnt = counts[item];
    const double lnp = hwe_scaled_lnp(cnt.y, cnt.x, cnt.z, midp != 0);
    out_p[item] = LNP ? lnp : exp(lnp);
  }
}


// ---------------------------------------------------------------------------------------------
// K8x: exact test for a biallelic X-chromosome marker (Graffelman & Weir 2016), replacing
// HweExactTestXchr -> plink2::HweXchrLnP (src/plink_hardy.cpp:83-95; plink2_stats.cc:3414-3592).
// Conditional on the allele counts and the numbers of males and females, a table is determined by
// (mb = ALT hemizygous males, fh = heterozygous females) and
//     P(mb, fh)  ~  2^fh / ( mb! (nm - mb)!  f0! fh! f2! ).
// The p-value is the mass of all tables no more likely than the observed one.  One block per
// marker; a thread owns rows (fixed mb), positions itself at the row's most likely fh through
// log-gamma, then walks the row in both directions with the autosomal two-term recurrences.
// All terms are in units of the observed table's probability.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double xchr_lnp(double mb, double nm, double f0, double fh, double f2) {
  return fh * 0.6931471805599453 - lgamma(mb + 1.0) - lgamma(nm - mb + 1.0) - lgamma(f0 + 1.0) - lgamma(fh + 1.0) -
         lgamma(f2 + 1.0);
}

__global__ void __launch_bounds__(256) k_hwe_xchr(const uint32_t* __restrict__ c5, uint32_t n, uint32_t midp,
                                                   double* __restrict__ out_p) {
  const uint32_t v = blockIdx.x;
  if (v >= n) return;
  const uint32_t of0 = c5[5 * v], ofh = c5[5 * v + 1], of2 = c5[5 * v + 2], oma = c5[5 * v + 3], omb = c5[5 * v + 4];
  const uint32_t nf = of0 + ofh + of2, nm = oma + omb;
  __shared__ double s_tail[256], s_center[256];
  __shared__ uint32_t s_ties[256];
  if (nf + nm == 0) {
    if (threadIdx.x == 0) out_p[v] = 1.0;  // src/plink_hardy.cpp:88-90
    return;
  }
  const uint64_t nB = 2ull * of2 + ofh + omb;  // ALT alleles
  const uint64_t nt = 2ull * nf + nm;
  const double l_obs = xchr_lnp(omb, nm, of0, ofh, of2);
  // rows: mb in [max(0, nB - 2 nf), min(nm, nB)].  The row maxima are unimodal in mb, so the rows are taken
  // from the most likely one outwards, 256 at a time, and a thread stops in a direction once its row's maximum
  // is below e^-64 of the observed table (every later row of that direction is smaller still).  Rows on the
  // far side of the observed one are part of the tail and must not be cut at the observed row.
  const int64_t lo = nB > 2ull * nf ? static_cast<int64_t>(nB - 2ull * nf) : 0;
  const int64_t hi = static_cast<int64_t>(nB < nm ? nB : nm);
  int64_t mid = static_cast<int64_t>(static_cast<double>(nm) * static_cast<double>(nB) / static_cast<double>(nt) + 0.5);
  mid = mid < lo ? lo : (mid > hi ? hi : mid);
  double tail = 0.0, center = 0.0;
  uint32_t ties = 0;
  bool overflow = false;
  // one row; returns false when the row (and every row further out) is negligible
  auto row = [&](int64_t mbi) -> bool {
    const uint64_t mb = static_cast<uint64_t>(mbi);
    const uint64_t fb = nB - mb;  // ALT alleles carried by females
    const uint64_t fa = 2ull * nf - fb;
    const uint64_t fh_max = fb < fa ? fb : fa;
    // most likely heterozygote count of the row ~ fa fb / (2 nf - 1), with the parity of fb
    double est = (2.0 * nf > 1.0) ? static_cast<double>(fa) * static_cast<double>(fb) / (2.0 * nf - 1.0) : 0.0;
    uint64_t fh0 = static_cast<uint64_t>(est);
    if ((fh0 ^ fb) & 1ull) fh0 += 1;
    if (fh0 > fh_max) fh0 = fh_max - ((fh_max ^ fb) & 1ull);
    if (mb == omb) fh0 = ofh;  // start the observed row exactly on the observed table
    const double f2s = static_cast<double>((fb - fh0) >> 1);
    const double f0s = static_cast<double>(nf) - static_cast<double>(fh0) - f2s;
    const double l0 = xchr_lnp(static_cast<double>(mb), nm, f0s, static_cast<double>(fh0), f2s) - l_obs;
    if (l0 > 700.0) {
      overflow = true;  // the observed table is > e^700 times less likely than this one: p = 0
      return false;
    }
    if (l0 < -64.0 && mb != omb) return false;
    const double t0 = (mb == omb) ? 1.0 : exp(l0);
    auto put = [&](double t) {
      if (t > 1.0 + kTieEps) {
        center += t;
      } else {
        if (t >= 1.0 - kTieEps) ++ties;
        tail += t;
      }
    };
    if (mb == omb) {
      tail += 1.0;
      ++ties;
    } else {
      put(t0);
    }
    // Both walks move away from the row's maximum, so the terms only shrink: below 1e-24 of the observed
    // table they can change neither the tail (>= 1) nor the total.
    {  // fewer heterozygotes
      double k = static_cast<double>(fh0), r = f0s, c = f2s, t = t0;
      while (k > 1.5) {
        r += 1.0;
        c += 1.0;
        t *= (k * (k - 1.0)) / (4.0 * r * c);
        k -= 2.0;
        put(t);
        if (t < 1e-24) break;
      }
    }
    {  // more heterozygotes
      double k = static_cast<double>(fh0), r = f0s, c = f2s, t = t0;
      while (r > 0.5 && c > 0.5) {
        k += 2.0;
        t *= (4.0 * r * c) / (k * (k - 1.0));
        r -= 1.0;
        c -= 1.0;
        put(t);
        if (t < 1e-24) break;
      }
    }
    if (!(center < 1.0e300)) {
      overflow = true;
      return false;
    }
    return true;
  };
  // Every row between the most likely one and the observed one has a maximum >= the observed table, so no
  // direction stops before it has passed the observed row.
  for (int64_t mb = mid + threadIdx.x; mb <= hi; mb += blockDim.x) {
    if (!row(mb)) break;
  }
  for (int64_t mb = mid - 1 - static_cast<int64_t>(threadIdx.x); mb >= lo; mb -= blockDim.x) {
    if (!row(mb)) break;
  }
  s_tail[threadIdx.x] = overflow ? -1.0 : tail;
  s_center[threadIdx.x] = center;
  s_ties[threadIdx.x] = ties;
  __syncthreads();
  if (threadIdx.x == 0) {
    double T = 0.0, C = 0.0;
    uint32_t ti = 0;
    bool ovf = false;
    for (int i = 0; i < 256; ++i) {
      if (s_tail[i] < 0.0) ovf = true;
      else T += s_tail[i];
      C += s_center[i];
      ti += s_ties[i];
    }
    double p = 0.0;
    if (!ovf && C < 1.0e300) {
      const double total = T + C;
      if (midp) T -= 0.5 * static_cast<double>(ti);
      p = T / total;
      if (p < 0.0) p = 0.0;
      if (p > 1.0) p = 1.0;
    }
    out_p[v] = p;
  }
}

}  // namespace

void launch_hwe(const uint4* counts, uint32_t n, uint32_t midp, double* out_p, uint32_t* queue, cudaStream_t s,
                LaunchStats* st, bool want_lnp) {
  if (!n) return;
  cudaMemsetAsync(queue, 0, kHweQueueWords * sizeof(uint32_t), s);
  const int sms = st && st->sm_count > 0 ? st->sm_count : 148;
  uint32_t blocks = (n + 255) / 256;  // 8 warps x 32 variants per CTA at least
  if (blocks > static_cast<uint32_t>(sms) * 4) blocks = sms * 4;
  if (want_lnp) {
    k_hwe<true><<<blocks, 256, 0, s>>>(counts, n, midp, out_p, queue);
  } else {
    k_hwe<false><<<blocks, 256, 0, s>>>(counts, n, midp, out_p, queue);
  }
  if (st) st->launches++;
  launch_hwe_scaled(counts, queue, midp, out_p, want_lnp, s, st);
}

void launch_hwe_scaled(const uint4* counts, const uint32_t* queue, uint32_t midp, double* out_p, bool want_lnp,
                       cudaStream_t s, LaunchStats* st) {
  const unsigned blocks = static_cast<unsigned>(sm_of(st)) * 4u;
  if (want_lnp) {
    k_hwe_scaled<true><<<blocks, 32, 0, s>>>(counts, queue + 2, midp, out_p);
  } else {
    k_hwe_scaled<false><<<blocks, 32, 0, s>>>(counts, queue + 2, midp, out_p);
  }
  if (st) st->launches++;
}

void launch_hwe_xchr(const uint32_t* c5, uint32_t n, uint32_t midp, double* out_p, cudaStream_t s, LaunchStats* st) {
  if (!n) return;
  k_hwe_xchr<<<n, 256, 0, s>>>(c5, n, midp, out_p);
  if (st) st->launches++;
}

}  // namespace pdk
