#!/usr/bin/env python
"""Regenerates tests/golden/hwe_grid.npz (run in the build container, where oracle/_ref exists).

For N in GRID_N: the far-from-equilibrium count grid of tests/hwe_check.py::departure_grid and, per table and
for midp in {0, 1}: ln p from the reference's own plink2::HweLnP (oracle/_ref) and from the extended-precision
restatement oracle/pgen_oracle.c::po_hwe_lnp_exact.  Also prints (and stores) the calibration the tolerances in
tests/hwe_check.py quote: HweLnP's own error against the extended-precision value, in ulp(2n ln 2n), and pins
po_hwe_lnp_exact to a 60-digit mpmath evaluation of the definition (log-gamma based, no recurrences).
"""
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import port, ref  # noqa: E402
import hwe_check  # noqa: E402

GRID_N = (1000, 100003, 500000)


def mp_lnp(hets, h1, h2, midp):
    """ln p straight from the definition in 60 significant digits (mpmath), O(n) log-gammas."""
    import mpmath as mp

    mp.mp.dps = 60
    r0 = min(h1, h2)
    n = hets + h1 + h2
    rare = 2 * r0 + hets
    if rare < 2:
        return math.log(0.5) if midp else 0.0

    def lp(k):
        r = (rare - k) // 2
        c = n - k - r
        return k * mp.log(2) - mp.loggamma(k + 1) - mp.loggamma(r + 1) - mp.loggamma(c + 1)

    l_obs = lp(hets)
    hi, lo = mp.log(1 + mp.mpf(2) ** -38), mp.log(1 - mp.mpf(2) ** -38)
    tail = tot = mp.mpf(0)
    ties = 0
    for k in range(rare & 1, rare + 1, 2):
        d = lp(k) - l_obs
        t = mp.e ** d
        tot += t
        if d <= hi:
            tail += t
            ties += d >= lo
    if midp:
        tail -= mp.mpf(ties) / 2
    return float(mp.log(tail / tot))


def main():
    out = {}
    # 1. pin the extended-precision restatement to the definition
    rng = np.random.default_rng(1)
    worst = 0.0
    for _ in range(60):
        n = int(rng.integers(10, 4000))
        h1 = int(rng.integers(0, n))
        hets = int(rng.integers(0, n - h1 + 1))
        h2 = n - h1 - hets
        for midp in (False, True):
            a, b = port.hwe_lnp_exact(hets, h1, h2, midp), mp_lnp(hets, h1, h2, midp)
            if math.isinf(a) and math.isinf(b):
                continue
            worst = max(worst, abs(a - b) / max(1.0, abs(b)))
    print(f"po_hwe_lnp_exact vs 60-digit mpmath, 120 tables: worst |d ln p| / max(1, |ln p|) = {worst:.3g}")
    assert worst < 1e-15
    out["exact_vs_mpmath_worst"] = np.array(worst)

    # 2. the grid + both yardsticks, and the calibration of the reference's own noise
    for N in GRID_N:
        counts = hwe_check.departure_grid(N, seed=N)
        out[f"n{N}/counts"] = counts
        n_obs = counts[:, :3].sum(axis=1)
        for midp in (0, 1):
            rf = ref.hwe_lnp_batch(counts, bool(midp))
            ex = np.array([port.hwe_lnp_exact(int(c[1]), int(c[0]), int(c[2]), bool(midp)) for c in counts])
            out[f"n{N}/ref_lnp_midp{midp}"] = rf
            out[f"n{N}/exact_lnp_midp{midp}"] = ex
            ulp = np.spacing(2.0 * np.maximum(n_obs, 1) * np.log(2.0 * np.maximum(n_obs, 1) + 2.0))
            with np.errstate(invalid="ignore"):
                tail = np.exp(ex) < hwe_check.CENTRE_P
                d_ln = np.where(np.isinf(ex) & np.isinf(rf), 0.0, np.abs(rf - ex)) / ulp
                d_p = np.abs(np.exp(rf) - np.exp(ex)) / ulp
            err = np.where(tail, d_ln, np.minimum(d_ln, d_p))
            print(f"N={N} midp={midp}: {counts.shape[0]} tables, ln p in [{ex[np.isfinite(ex)].min():.1f}, 0]; "
                  f"HweLnP vs extended precision: max {err.max():.2f} ulp(2n ln 2n)")
            assert err.max() < hwe_check.REF_ULPS / 2
            assert hwe_check.violations_lnp_vs_reference(ex, rf, n_obs).size == 0
    np.savez_compressed(hwe_check.GOLDEN, **out)
    print("wrote", hwe_check.GOLDEN, os.path.getsize(hwe_check.GOLDEN), "bytes")


if __name__ == "__main__":
    main()
