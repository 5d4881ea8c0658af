"""HWE yardsticks and tolerances on the CPU: the committed golden grid (tests/golden/hwe_grid.npz) against
the live reference (oracle/_ref, plink2::HweLnP) and the extended-precision restatement, and the checker
itself (tests/hwe_check.py) -- it must reject a 1e-9 relative error at p = 1e-40."""
import numpy as np
import pytest

import hwe_check as hc
from oracle import port

GRID_N = (1000, 100003, 500000)


@pytest.mark.parametrize("N", GRID_N)
def test_golden_grid_matches_live_oracles(N, built):
    from oracle import ref

    g = hc.load_golden()
    counts = g[f"n{N}/counts"]
    assert np.array_equal(counts, hc.departure_grid(N, seed=N))
    assert np.all(counts.sum(axis=1) == N)
    sel = np.arange(0, counts.shape[0], 5)
    for midp in (0, 1):
        ex = np.array([port.hwe_lnp_exact(int(c[1]), int(c[0]), int(c[2]), bool(midp)) for c in counts[sel]])
        assert np.array_equal(ex, g[f"n{N}/exact_lnp_midp{midp}"][sel])
        if ref.available():
            assert np.array_equal(ref.hwe_lnp_batch(counts, bool(midp)), g[f"n{N}/ref_lnp_midp{midp}"])
    # the grid really is far from equilibrium: p from 1 down to far below fp64 underflow
    ex = g[f"n{N}/exact_lnp_midp0"]
    assert ex.max() == 0.0 and ex.min() < -600.0
    assert np.sum((ex < np.log(1e-10)) & (ex > np.log(1e-300))) >= 10


@pytest.mark.parametrize("N", GRID_N)
@pytest.mark.parametrize("midp", [0, 1])
def test_reference_and_f64_walk_inside_stated_bounds(N, midp):
    """HweLnP sits within REF_ULPS ulp(2n ln 2n) of the extended-precision value (that is the reference's own
    noise), and a plain fp64 restatement of the recurrences (po_hwe_exact, what the device kernel does) within
    1e-12 relative of it where p is representable."""
    g = hc.load_golden()
    counts = g[f"n{N}/counts"]
    n_obs = counts[:, :3].sum(axis=1)
    rf, ex = g[f"n{N}/ref_lnp_midp{midp}"], g[f"n{N}/exact_lnp_midp{midp}"]
    assert hc.violations_lnp_vs_reference(ex, rf, n_obs).size == 0
    p = np.array([port.hwe_exact(int(c[1]), int(c[0]), int(c[2]), bool(midp)) for c in counts])
    v = hc.violations_p_vs_exact(p, ex, n_obs)
    assert v.size == 0, hc.describe(v, counts, p, rf, ex)
    v = hc.violations_p_vs_reference(p, rf, n_obs)
    assert v.size == 0, hc.describe(v, counts, p, rf, ex)


@pytest.mark.parametrize("N", GRID_N)
def test_checker_rejects_1e9_relative_error_at_1e40(N):
    g = hc.load_golden()
    counts = g[f"n{N}/counts"]
    n_obs = counts[:, :3].sum(axis=1)
    ex, rf = g[f"n{N}/exact_lnp_midp0"], g[f"n{N}/ref_lnp_midp0"]
    i = int(np.argmin(np.abs(ex - np.log(1e-40))))
    assert np.log(1e-60) < ex[i] < np.log(1e-25)
    p = np.exp(ex)
    assert hc.violations_p_vs_exact(p, ex, n_obs).size == 0
    for rel in (1e-9, -1e-9, 3e-12):
        q = p.copy()
        q[i] *= 1.0 + rel
        assert i in hc.violations_p_vs_exact(q, ex, n_obs), (N, rel)
        ln = ex.copy()
        ln[i] += rel
        assert i in hc.violations_lnp_vs_exact(ln, ex), (N, rel)
    if N <= 1000:  # the reference itself is only good to ~4e-9 at N = 1e5: there the exact yardstick is the gate
        q = p.copy()
        q[i] *= 1.0 + 1e-9
        assert i in hc.violations_p_vs_reference(q, rf, n_obs)
    # an underflowed p-value must be exactly zero, a subnormal one small
    j = int(np.argmin(ex))
    q = p.copy()
    q[j] = 1e-200
    assert j in hc.violations_p_vs_exact(q, ex, n_obs)
