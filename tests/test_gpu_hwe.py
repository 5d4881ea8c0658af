"""GPU parity of the exact Hardy-Weinberg test far from equilibrium, in ln-p space.

Yardsticks and tolerances: tests/hwe_check.py.  Inputs: tests/golden/hwe_grid.npz -- count tables at
N = 1000 / 100003 / 500000 with MAF 1e-4..0.5 and inbreeding coefficient -1..+1 (het excess and deficit), so
p runs from 1 through 1e-10..1e-300 down to ln p = -346 000, each with the reference's own HweLnP value
(plink2_stats.cc:1784-2166 via oracle/_ref) and the extended-precision value.  Every kernel that computes the
test is driven through the C ABI: K8 `k_hwe` (+ `k_hwe_scaled`) via pd_hwe_from_counts /
pd_hwe_lnp_from_counts, the fused K1h `k_hardy_fixed` via pd_hardy on fixed-width files built to have exactly
those tables, and K3 + K8 via pd_hardy on a pgenlib-written variable-width file."""
import numpy as np
import pytest

import hwe_check as hc

pytestmark = pytest.mark.gpu

GRID_N = (1000, 100003, 500000)


def _any_ctx():
    from conftest import fixture_path
    from plinking_duck_b200.pgen import PgenContext

    return PgenContext(fixture_path("pgen_example"))


def _live_ref(counts, midp, golden_ref):
    """The reference's HweLnP, live when oracle/_ref travelled with the snapshot (it must equal the golden)."""
    from oracle import ref

    if ref.available():
        live = ref.hwe_lnp_batch(counts, bool(midp))
        assert np.array_equal(live, golden_ref)
        return live
    return golden_ref


@pytest.mark.parametrize("N", GRID_N)
@pytest.mark.parametrize("midp", [0, 1])
def test_k8_lnp_far_from_equilibrium(N, midp):
    g = hc.load_golden()
    counts = g[f"n{N}/counts"]
    n_obs = counts[:, :3].sum(axis=1)
    ex = g[f"n{N}/exact_lnp_midp{midp}"]
    rf = _live_ref(counts, midp, g[f"n{N}/ref_lnp_midp{midp}"])
    # tables without observations: the entry points follow HweExactTestAutosomal (p = 1, src/plink_hardy.cpp:73-75),
    # not HweLnP's own mid-p value of ln 0.5 for an empty table
    ex = np.where(n_obs == 0, 0.0, ex)
    rf = np.where(n_obs == 0, 0.0, rf)
    with _any_ctx() as c:
        lnp = c.hwe_lnp_from_counts(counts, midp=bool(midp))
        p = c.hwe_from_counts(counts, midp=bool(midp))
        assert c.stats().launches > 0
    v = hc.violations_lnp_vs_exact(lnp, ex)
    assert v.size == 0, hc.describe(v, counts, lnp, rf, ex)
    v = hc.violations_lnp_vs_reference(lnp, rf, n_obs)
    assert v.size == 0, hc.describe(v, counts, lnp, rf, ex)
    v = hc.violations_p_vs_exact(p, ex, n_obs)
    assert v.size == 0, hc.describe(v, counts, p, rf, ex)
    v = hc.violations_p_vs_reference(p, rf, n_obs)
    assert v.size == 0, hc.describe(v, counts, p, rf, ex)
    # plain relative check where p is a normal fp64 number, against exp(ln p) of both yardsticks
    ok = ex > hc.LN_1E300
    rel = np.abs(p[ok] - np.exp(ex[ok])) / np.exp(ex[ok])
    assert rel.max() <= 1e-12, rel.max()
    small = (ex < np.log(1e-10)) & ok
    assert small.sum() >= 10  # p-values an HWE filter acts on are really compared
    # p-values below fp64: the ln-p entry point still resolves them, the p entry point returns exactly 0
    deep = ex < -800.0
    assert np.all(p[deep] == 0.0) and np.all(np.isfinite(lnp[deep]))
    assert N < 100000 or deep.sum() >= 20


@pytest.mark.parametrize("N", [100003, 500000])
def test_k1h_fused_far_from_equilibrium(N, tmp_path):
    """The fused count + exact-test kernel on files that hold exactly the grid's tables: resident, streamed
    in small tiles, page-locked outputs (kernel stores p over PCIe), sub-range."""
    import torch

    import synth
    from plinking_duck_b200.pgen import PgenContext

    g = hc.load_golden()
    step = 2 if N == 100003 else 4
    counts = g[f"n{N}/counts"][::step]
    n_obs = counts[:, :3].sum(axis=1)
    geno = synth.genotypes_from_counts(counts, seed=N)
    img = synth.fixed_pgen_image(geno)
    V = counts.shape[0]
    for midp in (0, 1):
        ex = g[f"n{N}/exact_lnp_midp{midp}"][::step]
        rf = _live_ref(counts, midp, g[f"n{N}/ref_lnp_midp{midp}"][::step])
        for kw in (dict(), dict(tile_bytes=1 << 20, n_staging=2)):
            with PgenContext(image=img, **kw) as c:
                for resident in (False, True):
                    if resident:
                        c.load_resident()
                    got_counts, p = c.hardy(midp=bool(midp))
                    assert np.array_equal(got_counts, counts)
                    v = hc.violations_p_vs_exact(p, ex, n_obs)
                    assert v.size == 0, hc.describe(v, counts, p, rf, ex)
                    v = hc.violations_p_vs_reference(p, rf, n_obs)
                    assert v.size == 0, hc.describe(v, counts, p, rf, ex)
                pc = torch.zeros((V, 4), dtype=torch.int32).pin_memory().numpy().view(np.uint32)
                pp = torch.full((V,), -1.0, dtype=torch.float64).pin_memory().numpy()
                c.hardy(midp=bool(midp), out_counts=pc, out_p=pp)
                assert np.array_equal(pc, counts) and np.array_equal(pp, p)
                a, b = V // 4, V - 3
                _, ps = c.hardy(a, b, midp=bool(midp))
                assert np.array_equal(ps, p[a:b])
    ok = ex > hc.LN_1E300
    assert ((ex < np.log(1e-10)) & ok).sum() >= 5


def test_k3_k8_variable_width_far_from_equilibrium(tmp_path):
    """pd_hardy on a pgenlib-written (mode 0x10) file with the same tables: K3 counts + K8."""
    import synth
    from oracle import ref
    from plinking_duck_b200.pgen import PgenContext

    if not ref.available():
        pytest.skip("oracle/_ref not built: no pgenlib writer")
    N = 100003
    g = hc.load_golden()
    counts = g[f"n{N}/counts"][1::5]
    n_obs = counts[:, :3].sum(axis=1)
    geno = synth.genotypes_from_counts(counts, seed=7)
    path = str(tmp_path / "dep.pgen")
    ref.write_pgen(path, synth.to_genovec_words(geno), N)
    with ref.RefPgen(path) as r:
        assert np.array_equal(r.counts(), counts)
    for midp in (0, 1):
        ex = g[f"n{N}/exact_lnp_midp{midp}"][1::5]
        rf = ref.hwe_lnp_batch(counts, bool(midp))
        with PgenContext(path) as c:
            got_counts, p = c.hardy(midp=bool(midp))
        assert np.array_equal(got_counts, counts)
        v = hc.violations_p_vs_exact(p, ex, n_obs)
        assert v.size == 0, hc.describe(v, counts, p, rf, ex)
        v = hc.violations_p_vs_reference(p, rf, n_obs)
        assert v.size == 0, hc.describe(v, counts, p, rf, ex)
