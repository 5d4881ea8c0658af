"""CPU: pins the oracle.  oracle/_ref is the reference's own pgenlib compiled in place; here it is
checked against (a) every known-answer value the reference's sqllogictests hold for the hot path
(tests/kat.py) and (b) the committed goldens (tests/golden/golden.npz must be what the oracle
produces today).  The GPU parity tests then compare the CUDA path with those same goldens."""
import numpy as np
import pytest

import kat
from conftest import fixture_names, fixture_path


def test_counts_kat(refmod):
    with refmod.RefPgen(fixture_path("pgen_example")) as r:
        c = r.counts()
    assert np.array_equal(c, kat.PGEN_EXAMPLE_COUNTS)
    obs = c[:, :3].sum(1)
    assert list(2 * obs) == kat.PGEN_EXAMPLE_OBS_CT
    assert np.allclose((c[:, 1] + 2.0 * c[:, 2]) / (2.0 * obs), kat.PGEN_EXAMPLE_ALT_FREQ)
    with refmod.RefPgen(fixture_path("all_missing")) as r:
        assert np.array_equal(r.counts(), kat.ALL_MISSING_COUNTS)


def test_hardy_kat(refmod):
    c = kat.PGEN_EXAMPLE_COUNTS
    p = refmod.lnp_to_pvalue(refmod.hwe_lnp_batch(c, False))
    pm = refmod.lnp_to_pvalue(refmod.hwe_lnp_batch(c, True))
    assert np.allclose(p, kat.PGEN_EXAMPLE_P_HWE, atol=5e-7)
    assert np.allclose(pm, kat.PGEN_EXAMPLE_P_HWE_MIDP, atol=5e-7)
    obs = c[:, :3].sum(1).astype(float)
    assert np.allclose(c[:, 1] / obs, kat.PGEN_EXAMPLE_O_HET, atol=5e-7)
    pr = (2.0 * c[:, 0] + c[:, 1]) / (2.0 * obs)
    assert np.allclose(2 * pr * (1 - pr), kat.PGEN_EXAMPLE_E_HET, atol=5e-7)
    with refmod.RefPgen(fixture_path("large_example")) as r:
        cl = r.counts()
    pl = refmod.lnp_to_pvalue(refmod.hwe_lnp_batch(cl, False))
    assert np.allclose(pl, kat.LARGE_EXAMPLE_P_HWE, atol=5e-7)


def test_xchr_kat(refmod):
    x = kat.XCHR_X1
    lnp = refmod.hwe_xchr_lnp(x["fhet"], x["fhom_ref"], x["fhom_alt"], x["male_ref"], x["male_alt"], False)
    lnpm = refmod.hwe_xchr_lnp(x["fhet"], x["fhom_ref"], x["fhom_alt"], x["male_ref"], x["male_alt"], True)
    assert abs(float(refmod.lnp_to_pvalue(lnp)) - x["p"]) < 5e-7
    assert abs(float(refmod.lnp_to_pvalue(lnpm)) - x["p_midp"]) < 5e-7


def test_missing_kat(refmod):
    with refmod.RefPgen(fixture_path("pgen_example")) as r:
        pv, ps = r.missingness()
        N = r.N
    assert [(int(m), int(N - m), float(m) / N) for m in pv] == kat.PGEN_EXAMPLE_MISSING
    assert list(ps) == kat.PGEN_EXAMPLE_MISSING_BY_SAMPLE
    with refmod.RefPgen(fixture_path("large_example")) as r:
        pv, ps = r.missingness()
        V = r.V
    m, o, f = kat.LARGE_EXAMPLE_SAMPLE_MISSING
    assert (ps == m).all() and V - m == o and m / V == f


def test_score_kat(refmod):
    with refmod.RefPgen(fixture_path("pgen_example")) as r:
        vid = np.arange(4)
        s, d, a = r.score(vid, np.zeros(4, np.uint8), kat.SCORE_W, 0)
        assert np.allclose(s, kat.SCORE_SUM) and np.allclose(d, kat.SCORE_DOSAGE_SUM) and (a == 8).all()
        assert np.allclose(s / a, kat.SCORE_AVG)
        s, d, a = r.score(vid, np.zeros(4, np.uint8), kat.SCORE_W, 1)
        for i, (act, ss, ds) in kat.SCORE_NOIMPUTE.items():
            assert (int(a[i]), float(s[i]), float(d[i])) == (act, ss, ds)
        s, d, a = r.score(np.array([1]), np.zeros(1, np.uint8), np.array([1.0]), 2)
        for i, v in kat.SCORE_CENTER_RS2.items():
            assert abs(s[i] - v) < 1e-15
    with refmod.RefPgen(fixture_path("all_missing")) as r:
        s, d, a = r.score(np.arange(2), np.zeros(2, np.uint8), np.ones(2), 0)
        assert (s == 0).all() and (a == 0).all()


def test_sparse_and_dosage_kat(refmod):
    with refmod.RefPgen(fixture_path("rare_small")) as r:
        c = r.counts()
        het, hom, mis = r.sample_counts(sparse=True)
    assert tuple(int(x) for x in c.sum(0)) == kat.RARE_SMALL_SUMS
    assert (int(het.sum()), int(hom.sum()), int(mis.sum())) == kat.RARE_SMALL_SUMS[1:]
    with refmod.RefPgen(fixture_path("dosage_example")) as r:
        _, ad, r2 = r.dcounts()
    ad = ad.astype(float)
    assert np.allclose(ad[:, 1] / ad.sum(1), kat.DOSAGE_ALT_FREQ, rtol=1e-15)
    assert list(ad.sum(1) / 16384) == kat.DOSAGE_OBS_CT
    assert np.allclose(r2, kat.DOSAGE_IMP_R2, rtol=1e-14)


@pytest.mark.parametrize("name", fixture_names())
def test_goldens_are_current(name, golden, refmod):
    """golden.npz (made by tests/golden/make_golden.py) is exactly what the oracle returns."""
    with refmod.RefPgen(fixture_path(name)) as r:
        assert np.array_equal(r.counts(), golden[name + "/counts"])
        pv, ps = r.missingness()
        assert np.array_equal(pv, golden[name + "/missing_by_variant"])
        assert np.array_equal(ps, golden[name + "/missing_by_sample"])
        if name + "/i8" in golden and r.V * r.N <= 400_000:
            assert np.array_equal(r.decode_i8(), golden[name + "/i8"])
        if name + "/dcounts" in golden:
            dc, ad, r2 = r.dcounts()
            assert np.array_equal(ad, golden[name + "/all_dosages"])


def test_genotype_counts_struct_kat(refmod):
    """genotypes := 'counts' / 'stats' known answers (test/sql/read_pfile_genotypes_counts.test,
    read_pfile_genotypes_stats.test) against the reference's own pgenlib."""
    with refmod.RefPgen(fixture_path("pgen_example")) as r:
        c = r.counts()
        het, hom, mis = r.sample_counts(sparse=False)
    assert [tuple(int(x) for x in row) for row in c] == [k[:4] for k in kat.GENO_STATS_VARIANT]
    hom_ref = c.shape[0] - het - hom - mis  # src/pfile_reader.cpp:3573
    assert list(zip(hom_ref.tolist(), het.tolist(), hom.tolist(), mis.tolist())) == kat.GENO_COUNTS_SAMPLE
    n = hom_ref + het + hom
    af = (het + 2.0 * hom) / (2.0 * n)
    assert [(int(a), round(float(b), 4), round(float(m) / 4, 4), int(cc)) for a, b, m, cc in zip(n, af, mis, het + hom)] \
        == kat.GENO_STATS_SAMPLE
    with refmod.RefPgen(fixture_path("rare_small")) as r:
        het, hom, mis = r.sample_counts(sparse=True)
        nv = r.counts().shape[0]
    n = nv * het.size - int(mis.sum())
    assert (n, int(het.sum() + hom.sum()), round(float(het.sum()) / n, 6)) == kat.RARE_SMALL_STATS


@pytest.mark.parametrize("storage,N", [(8, 5), (9, 7), (10, 4), (11, 13), (12, 1), (14, 1)])
def test_compact_header_modes_index_matches_pgenlib(storage, N, refmod, tmp_path):
    """header_ctrl modes 8-14 (pgenlib_read.cc:928-957, :1260-1355): the host index parser (pd_probe) and the
    reference's own reader agree on record types and offsets of hand-built files that cross a variant-block
    boundary, and pgenlib's counts on them are the numpy tally (the extra tracks are never parsed)."""
    import synth
    from plinking_duck_b200.pgen import probe

    V = 65536 + 777
    g = synth.gen_hwe_genotypes(V, N, seed=storage, missing=0.05)
    g[::3] = 0
    image, vrtypes, fpos = synth.compact_pgen_image(storage, g, seed=storage)
    path = str(tmp_path / f"compact{storage}.pgen")
    image.tofile(path)
    info, vt, fp = probe(path)
    assert (info.variant_ct, info.sample_ct) == (V, N)
    assert np.array_equal(vt, vrtypes) and np.array_equal(fp, fpos)
    with refmod.RefPgen(path) as r:
        assert np.array_equal(r.vrtypes(), vrtypes) and np.array_equal(r.fpos(), fpos)
        assert np.array_equal(r.counts(), synth.counts_numpy(g))
    # an undefined code is rejected, not read as a random record type
    if storage in (11, 14):
        bad = image.copy()
        bad[12 + 16] = 0x88 if storage == 11 else 0x22
        bad.tofile(path)
        with pytest.raises(Exception):
            probe(path)
