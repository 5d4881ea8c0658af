"""the other consumers of the decode kernels.  plink_ld's R2 / D_PRIME / OBS_CT (pd_ld_pairs +
pdop_ld_stats) against the reference's ComputeLdStats restated over pgenlib's PgrGet (oracle/ref_harness.cc::
ref_ld_pair) -- bit-identical, the sums are exact integers; plink_pca / plink_glm standardised rows
(pd_standardized_f64) against NormalizeGenotypes restated in numpy over pgenlib's calls and counts -- bit-identical."""
import numpy as np
import pytest

from conftest import fixture_path

pytestmark = pytest.mark.gpu


def _ctx(name, **kw):
    from plinking_duck_b200.pgen import PgenContext

    return PgenContext(fixture_path(name), **kw)


@pytest.mark.parametrize("name", ["pgen_example", "pca_example", "syn_n1000", "streaming_example", "syn_n70000", "rare_small"])
@pytest.mark.parametrize("kw", [dict(), dict(tile_bytes=4096, n_staging=2), dict(devices=[0])])
def test_plink_ld_pairs_bit_identical(name, kw, golden):
    from oracle import ref
    from plinking_duck_b200 import ops

    if not ref.available():
        pytest.skip("oracle/_ref not built")
    V = golden[name + "/counts"].shape[0]
    rng = np.random.default_rng(3)
    # windowed-scan shaped pair list (anchor, anchor + 1 .. anchor + w) plus self pairs and far pairs
    va, vb = [], []
    for a in rng.choice(V, size=min(V, 12), replace=False):
        for j in range(int(a), min(V, int(a) + 6)):
            va.append(int(a))
            vb.append(j)
    va += [0, V - 1, V - 1]
    vb += [V - 1, 0, V - 1]
    with ref.RefPgen(fixture_path(name)) as r, _ctx(name, **kw) as c:
        for subset in (None, np.arange(0, r.N, 2) if r.N >= 6 else None):
            if subset is not None:
                r.set_subset(subset)
                c.set_subset(subset)
            rows = ops.plink_ld(c, va, vb)
            for i, (a, b) in enumerate(zip(va, vb)):
                r2, dp, obs, valid = r.ld_pair(a, b)
                assert (int(rows["OBS_CT"][i]), int(rows["is_valid"][i])) == (obs, valid), (name, a, b)
                assert rows["R2"][i] == r2 and rows["D_PRIME"][i] == dp, (name, a, b, rows[i], r2, dp)
            c.load_resident()   # already-resident range: no temporary staging
            assert np.array_equal(ops.plink_ld(c, va, vb), rows)
            c.unload()
            if subset is None:
                break


@pytest.mark.parametrize("name", ["pca_example", "syn_n1000", "streaming_example", "all_missing"])
def test_standardized_rows_bit_identical(name, golden):
    i8 = golden[name + "/i8"]
    V, N = i8.shape
    vidx = np.arange(0, V, 2, dtype=np.uint32)

    def expected(mat):
        out = np.zeros(mat.shape, dtype=np.float64)
        skip = np.zeros(mat.shape[0], dtype=np.uint8)
        for k, row in enumerate(mat):
            c1, c2 = int((row == 1).sum()), int((row == 2).sum())
            obs = int((row != -9).sum())
            f = (float(c1) + 2.0 * float(c2)) / (2.0 * float(obs)) if obs else 0.0   # src/plink_pca.cpp:403-409
            if obs == 0 or f <= 0.0 or f >= 1.0:                                      # ComputeVariantNorm skip
                skip[k] = 1
                continue
            center = 2.0 * f
            inv = 1.0 / np.sqrt(2.0 * f * (1.0 - f))
            out[k] = np.where(row == -9, 0.0, (row.astype(np.float64) - center) * inv)  # NormalizeGenotypes
        return out, skip

    for kw in (dict(), dict(tile_bytes=2048, n_staging=2), dict(devices=[0])):
        with _ctx(name, **kw) as c:
            got, skip = c.standardized(vidx)
            eo, es = expected(i8[vidx])
            assert np.array_equal(skip, es)
            assert np.array_equal(got, eo), name
            if N >= 8:
                sub = np.arange(1, N, 3)
                c.set_subset(sub)
                got, skip = c.standardized(vidx)
                eo, es = expected(i8[vidx][:, sub])
                assert np.array_equal(skip, es) and np.array_equal(got, eo)
