"""BASELINE.json configs[0] at its own size: plink_freq + plink_missing on the S1 fileset (10 000 variants x
1 000 samples, chr 1 / 2 / X / Y / MT with a SEX column; tests/s1.py), written as a fixed-width mode 0x02 file AND
through pgenlib's writer, full result arrays against the reference's own pgenlib (oracle/_ref: PgrGetCounts,
PgrGet + GenoarrToBytesMinus9, PgrGetMissingness) plus the numpy restatement of ComputeSexAwareCounts."""
import numpy as np
import pytest

import s1

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fileset(tmp_path_factory):
    from oracle import ref

    if not ref.available():
        pytest.skip("oracle/_ref not built")
    d = str(tmp_path_factory.mktemp("s1"))
    g = s1.genotypes()
    return g, s1.write(d, g, with_pgenlib=True)


@pytest.mark.parametrize("kind", ["fixed", "pgenlib"])
@pytest.mark.parametrize("devices", [None, [0]])
def test_s1_freq_and_missing_full_arrays(fileset, kind, devices):
    from oracle import ref
    from plinking_duck_b200 import ops
    from plinking_duck_b200.pgen import PgenContext

    g, paths = fileset
    chrom, pos, sex = s1.layout()
    ploidy = np.array([ops.classify_ploidy(c, int(p)) for c, p in zip(chrom, pos)], dtype=np.uint8)
    assert (ploidy == ops.PLOIDY_CHR_X).sum() == 320 and (ploidy == ops.PLOIDY_CHR_Y).sum() == 100
    assert (ploidy == ops.PLOIDY_CHR_MT).sum() == 100 and (ploidy[9400:9460] == ops.PLOIDY_AUTOSOMAL).all()
    with ref.RefPgen(paths[kind]) as r:
        assert (r.V, r.N) == (s1.V, s1.N)
        rc = r.counts()
        ri8 = r.decode_i8()
        miss_v, miss_s = r.missingness()
        vt = r.vrtypes()
    assert np.array_equal(rc, __import__("synth").counts_numpy(g))
    if kind == "pgenlib":
        assert len(set((vt & 7).tolist())) >= 2  # the writer really chose several record types
    exp = s1.expected_freq_rows(rc, ri8, ploidy, sex, ops.FREQ_DTYPE)
    with PgenContext(paths[kind], devices=devices) as c:
        f = ops.plink_freq(c, ploidy=ploidy, sex=sex)
        for col in ("ALT_FREQ", "OBS_CT", "HOM_REF_CT", "HET_CT", "HOM_ALT_CT", "MISSING_CT", "freq_is_null",
                    "counts_are_null"):
            assert np.array_equal(f[col], exp[col]), col
        assert f["freq_is_null"][17] == 1 and f["freq_is_null"][9405] == 1
        # without a SEX column chrX / chrY rows are NULL, MT and autosomes are not (src/plink_common.cpp:2000-2006)
        fn = ops.plink_freq(c, ploidy=ploidy, sex=None)
        xy = (ploidy == ops.PLOIDY_CHR_X) | (ploidy == ops.PLOIDY_CHR_Y)
        assert np.all(fn["freq_is_null"][xy] == 1) and np.all(fn["counts_are_null"][xy] == 1)
        assert np.array_equal(fn["ALT_FREQ"][~xy], exp["ALT_FREQ"][~xy])
        # plink_missing, variant mode: MISSING_CT / OBS_CT / F_MISS (src/plink_missing.cpp:472-490)
        mv = ops.plink_missing(c, mode="variant")
        assert np.array_equal(mv["MISSING_CT"], miss_v)
        assert np.array_equal(mv["OBS_CT"], s1.N - miss_v)
        assert np.array_equal(mv["F_MISS"], miss_v.astype(np.float64) / float(s1.N))
        # sample mode: the denominator is the number of variants scanned (:643-646)
        ms = ops.plink_missing(c, mode="sample")
        assert np.array_equal(ms["MISSING_CT"], miss_s)
        assert np.array_equal(ms["OBS_CT"], s1.V - miss_s)
        assert np.array_equal(ms["F_MISS"], miss_s.astype(np.float64) / float(s1.V))
        # region + subset, as test/sql/plink_missing_sample.test:107
        sub = np.arange(3, s1.N, 7)
        c.set_subset(sub)
        a, b = 1234, 8765
        ms = ops.plink_missing(c, a, b, mode="sample")
        assert np.array_equal(ms["MISSING_CT"], (ri8[a:b][:, sub] == -9).sum(axis=0).astype(np.uint32))
        f2 = ops.plink_freq(c, a, b)
        cs = np.stack([(ri8[a:b][:, sub] == k).sum(axis=1) for k in (0, 1, 2, -9)], axis=1)
        assert np.array_equal(np.stack([f2["HOM_REF_CT"], f2["HET_CT"], f2["HOM_ALT_CT"], f2["MISSING_CT"]], axis=1), cs)
        assert c.stats().launches > 0
