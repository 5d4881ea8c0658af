"""HWE parity: the stated tolerances and the checks that enforce them (test infrastructure).

Two yardsticks for the exact Hardy-Weinberg test (plink2::HweLnP, plink2_stats.cc:1784-2166;
wrappers src/plink_hardy.cpp:52-79):

* the reference itself -- `oracle.ref.hwe_lnp` (pgenlib's own HweLnP) or the committed golden
  `tests/golden/hwe_grid.npz` generated from it.  HweLnP builds ln P(observed) from log-factorials as
  large as Lfact(2N) ~ 2N ln 2N, so its OWN result is only good to a few ulp(2N ln 2N): in ln p on its tail
  path (ln p = starting_lnprob + log(tailp)), in p itself on its centre path (p = 1 - centerp * exp(...),
  taken when P(observed) > 2^-10, i.e. p >= ~2^-11 with mid-p).  Measured against the extended-precision
  yardstick below over 12 000 tables at N = 1e2..5e5 (tests/golden/make_hwe_golden.py prints the table):
  max 2.4 ulp (tail, ln p) and 1.9 ulp (centre, p).  REF_ULPS = 8 leaves a 3x margin:
      N = 1e3 -> 1.5e-11,   N = 1e5 -> 3.7e-9,   N = 5e5 -> 1.5e-8.
* `oracle.port.hwe_lnp_exact`: the same test in x87 long double with an unbounded exponent, itself pinned
  to a 60-digit mpmath evaluation (make_hwe_golden.py).  Against it the device result must hold
      |ln p_gpu - ln p_exact| <= 1e-12 + 4 ulp(|ln p|)          (relative 1e-12 in p)
  at every N -- this is the bound that catches a 1e-9 relative error at p = 1e-40 whatever N is.
"""
from __future__ import annotations

import math
import os

import numpy as np

REF_ULPS = 8.0
EXACT_ABS = 1e-12
CENTRE_P = 2.0 ** -12  # the reference's centre path cannot be taken below this p

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hwe_grid.npz")


def ref_bound(n_obs) -> np.ndarray:
    """REF_ULPS * ulp(2n ln(2n + 2)) per table, n = non-missing genotypes."""
    n = np.maximum(np.asarray(n_obs, dtype=np.float64), 1.0)
    return REF_ULPS * np.spacing(2.0 * n * np.log(2.0 * n + 2.0))


def exact_bound(lnp_exact) -> np.ndarray:
    return EXACT_ABS + 4.0 * np.spacing(np.abs(np.asarray(lnp_exact, dtype=np.float64)))


def violations_lnp_vs_reference(lnp_gpu, lnp_ref, n_obs) -> np.ndarray:
    """Indices where ln p disagrees with the reference beyond the reference's own noise."""
    lnp_gpu = np.asarray(lnp_gpu, dtype=np.float64)
    lnp_ref = np.asarray(lnp_ref, dtype=np.float64)
    b = ref_bound(n_obs)
    with np.errstate(invalid="ignore", over="ignore"):
        d_ln = np.abs(lnp_gpu - lnp_ref)
        same_inf = np.isinf(lnp_gpu) & np.isinf(lnp_ref) & (lnp_gpu == lnp_ref)
        p_ref = np.exp(lnp_ref)
        d_p = np.abs(np.exp(lnp_gpu) - p_ref)
    ok = same_inf | (d_ln <= b) | ((p_ref >= CENTRE_P) & (d_p <= b))
    return np.nonzero(~ok)[0]


def violations_lnp_vs_exact(lnp_gpu, lnp_exact) -> np.ndarray:
    lnp_gpu = np.asarray(lnp_gpu, dtype=np.float64)
    lnp_exact = np.asarray(lnp_exact, dtype=np.float64)
    with np.errstate(invalid="ignore"):
        d = np.abs(lnp_gpu - lnp_exact)
        same_inf = np.isinf(lnp_gpu) & np.isinf(lnp_exact) & (lnp_gpu == lnp_exact)
    ok = same_inf | (d <= exact_bound(lnp_exact))
    return np.nonzero(~ok)[0]


LN_1E300 = math.log(1e-300)
LN_ZERO = -745.2  # exp() of anything below is 0.0 in IEEE double


def _p_to_lnp(p_gpu, lnp_yard):
    """ln of a p-value array for comparison in ln space; entries the yardstick puts below 1e-300 are checked
    separately (p_small_ok) and masked here."""
    p_gpu = np.asarray(p_gpu, dtype=np.float64)
    with np.errstate(divide="ignore"):
        ln = np.log(p_gpu)
    return np.where(np.asarray(lnp_yard) < LN_1E300, lnp_yard, ln)


def p_small_ok(p_gpu, lnp_yard) -> np.ndarray:
    """Below 1e-300 (subnormal or zero p) relative error is not defined on the fp64 p-value the API returns:
    require p <= 2e-300 there, and exactly 0 where exp(ln p) underflows."""
    p_gpu = np.asarray(p_gpu, dtype=np.float64)
    lnp_yard = np.asarray(lnp_yard, dtype=np.float64)
    small = lnp_yard < LN_1E300
    ok = ~small | ((p_gpu >= 0.0) & (p_gpu <= 2e-300))
    ok &= ~(lnp_yard < LN_ZERO - 1.0) | (p_gpu == 0.0)
    return ok


def violations_p_vs_reference(p_gpu, lnp_ref, n_obs) -> np.ndarray:
    """P_HWE as the API returns it (LnPToPvalue: clamp(exp(ln p), 0, 1); no observations -> 1.0)."""
    lnp_ref = np.where(np.asarray(n_obs) == 0, 0.0, np.minimum(np.asarray(lnp_ref, dtype=np.float64), 0.0))
    bad = set(violations_lnp_vs_reference(_p_to_lnp(p_gpu, lnp_ref), lnp_ref, n_obs).tolist())
    bad |= set(np.nonzero(~p_small_ok(p_gpu, lnp_ref))[0].tolist())
    return np.array(sorted(bad), dtype=np.int64)


def violations_p_vs_exact(p_gpu, lnp_exact, n_obs) -> np.ndarray:
    lnp_exact = np.where(np.asarray(n_obs) == 0, 0.0, np.asarray(lnp_exact, dtype=np.float64))
    bad = set(violations_lnp_vs_exact(_p_to_lnp(p_gpu, lnp_exact), lnp_exact).tolist())
    bad |= set(np.nonzero(~p_small_ok(p_gpu, lnp_exact))[0].tolist())
    return np.array(sorted(bad), dtype=np.int64)


def describe(idx, counts, got, lnp_ref=None, lnp_exact=None, limit=5) -> str:
    rows = []
    for i in idx[:limit]:
        r = f"#{int(i)} counts={tuple(int(x) for x in counts[i][:3])} got={got[i]!r}"
        if lnp_ref is not None:
            r += f" ref_lnp={lnp_ref[i]!r}"
        if lnp_exact is not None:
            r += f" exact_lnp={lnp_exact[i]!r}"
        rows.append(r)
    return f"{len(idx)} violations: " + "; ".join(rows)


# ---- the far-from-equilibrium table grid ---------------------------------------------------------------------

def departure_grid(N: int, seed: int, n_random: int = 260) -> np.ndarray:
    """(n, 4) uint32 counts (hom_ref, het, hom_alt, missing) summing to N: MAF from 1e-4 to 0.5, inbreeding
    coefficient F from -1 (all possible hets) to +1 (no hets) -- so p spans 1 down to far below fp64 underflow
    -- plus the corner tables."""
    rng = np.random.default_rng(seed)
    rows = []
    for i in range(n_random):
        maf = 10.0 ** rng.uniform(-4.0, math.log10(0.5))
        kind = i % 6
        F = [0.0, rng.uniform(-1.0, 1.0), rng.uniform(-0.3, 0.3), rng.uniform(-0.03, 0.03), 1.0, -1.0][kind]
        miss = int(rng.integers(0, max(1, N // 50))) if i % 3 == 0 else 0
        n = N - miss
        p, q = maf, 1.0 - maf
        F = max(F, -p / q)  # smallest feasible F
        f = np.array([q * q + F * p * q, 2.0 * p * q * (1.0 - F), p * p + F * p * q])
        f = np.clip(f, 0.0, None)
        f /= f.sum()
        if kind >= 4:       # exact expectation: the most extreme tables
            c = np.floor(f * n).astype(np.int64)
            c[0] += n - c.sum()
        else:
            c = rng.multinomial(n, f)
        if rng.random() < 0.5:
            c = c[::-1]
        rows.append((int(c[0]), int(c[1]), int(c[2]), miss))
    n = N
    h = n // 2
    corners = [
        (n, 0, 0, 0), (0, n, 0, 0), (0, 0, n, 0), (0, 0, 0, n), (n - 1, 1, 0, 0), (n - 1, 0, 1, 0), (n - 2, 1, 1, 0),
        (h, 0, n - h, 0), (h, n - h, 0, 0), (1, n - 2, 1, 0), (n // 3, n // 3, n - 2 * (n // 3), 0),
        (n // 4, n // 2, n - n // 4 - n // 2, 0), (n - 3, 2, 1, 0), (n - 3, 1, 2, 0), (n - 4, 0, 2, 2),
    ]
    rows.extend(c for c in corners if min(c) >= 0)
    return np.array(rows, dtype=np.uint32)


def load_golden():
    return np.load(GOLDEN)
