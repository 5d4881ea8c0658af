"""GPU: the operator layer (plink_freq / plink_hardy / plink_missing / plink_score rows) against the
reference's own sqllogictest known answers (tests/kat.py cites the lines) -- these read like
test/sql/plink_*.test with the SQL replaced by the ops call."""
import numpy as np
import pytest

import kat
from conftest import FIX_DIR, fixture_path

pytestmark = pytest.mark.gpu


def _ctx(name, **kw):
    from plinking_duck_b200.pgen import PgenContext

    return PgenContext(fixture_path(name), **kw)


def _psam_sex(name):
    import os

    rows = [l.split() for l in open(os.path.join(FIX_DIR, name + ".psam")) if l.strip()]
    col = rows[0].index("SEX")
    return np.array([int(r[col]) if r[col] in ("1", "2") else 0 for r in rows[1:]], dtype=np.uint8)


def _pvar(name):
    import os

    rows = [l.split() for l in open(os.path.join(FIX_DIR, name + ".pvar")) if l.strip() and not l.startswith("##")]
    return [(r[0], int(r[1]), r[2]) for r in rows[1:]]


def test_plink_freq_known_answers():
    from plinking_duck_b200 import ops

    with _ctx("pgen_example") as c:
        f = ops.plink_freq(c)
        assert np.allclose(f["ALT_FREQ"], kat.PGEN_EXAMPLE_ALT_FREQ) and list(f["OBS_CT"]) == kat.PGEN_EXAMPLE_OBS_CT
        got = np.stack([f["HOM_REF_CT"], f["HET_CT"], f["HOM_ALT_CT"], f["MISSING_CT"]], axis=1)
        assert np.array_equal(got, kat.PGEN_EXAMPLE_COUNTS)
        assert not f["freq_is_null"].any()
        # samples := [1, 3] style subset (test/sql/plink_freq.test:120-160): counts over 2 samples
        c.set_subset([0, 2])
        fs = ops.plink_freq(c)
        assert list(fs["OBS_CT"]) == [4, 4, 4, 4]
    with _ctx("all_missing") as c:
        f = ops.plink_freq(c)
        assert f["freq_is_null"].all() and (f["MISSING_CT"] == 2).all()  # test/sql/plink_freq.test:213-229
    with _ctx("large_example") as c:
        f = ops.plink_freq(c)
        assert (f["ALT_FREQ"] == 0.5).all() and (f["OBS_CT"] == 12).all()  # test/sql/plink_freq.test:247-255


def test_plink_freq_dosage_known_answers():
    from plinking_duck_b200 import ops

    with _ctx("dosage_example") as c:
        f = ops.plink_freq(c, dosage=True)
    assert np.allclose(f["ALT_FREQ"], kat.DOSAGE_ALT_FREQ, rtol=1e-15)
    assert list(f["OBS_CT"]) == kat.DOSAGE_OBS_CT
    assert np.allclose(f["IMP_R2"], kat.DOSAGE_IMP_R2, rtol=1e-14)


def test_plink_hardy_known_answers():
    from plinking_duck_b200 import ops

    with _ctx("pgen_example") as c:
        h = ops.plink_hardy(c)
        hm = ops.plink_hardy(c, midp=True)
    assert np.allclose(h["P_HWE"], kat.PGEN_EXAMPLE_P_HWE, atol=5e-7)
    assert np.allclose(hm["P_HWE"], kat.PGEN_EXAMPLE_P_HWE_MIDP, atol=5e-7)
    assert np.allclose(h["O_HET"], kat.PGEN_EXAMPLE_O_HET, atol=5e-7)
    assert np.allclose(h["E_HET"], kat.PGEN_EXAMPLE_E_HET, atol=5e-7)
    assert (hm["P_HWE"] <= h["P_HWE"] + 1e-15).all()
    with _ctx("large_example") as c:
        assert np.allclose(ops.plink_hardy(c)["P_HWE"], kat.LARGE_EXAMPLE_P_HWE, atol=5e-7)
    with _ctx("all_missing") as c:
        h = ops.plink_hardy(c)
        assert h["stats_are_null"].all()


def test_plink_missing_known_answers():
    from plinking_duck_b200 import ops

    with _ctx("pgen_example") as c:
        m = ops.plink_missing(c)
        assert [(int(r["MISSING_CT"]), int(r["OBS_CT"]), float(r["F_MISS"])) for r in m] == kat.PGEN_EXAMPLE_MISSING
        ms = ops.plink_missing(c, mode="sample")
        assert list(ms["MISSING_CT"]) == kat.PGEN_EXAMPLE_MISSING_BY_SAMPLE
        assert list(ms["OBS_CT"]) == [4, 3, 4, 3] and list(ms["F_MISS"]) == [0.0, 0.25, 0.0, 0.25]
        with pytest.raises(ValueError):
            ops.plink_missing(c, mode="bogus")
    with _ctx("large_example") as c:
        ms = ops.plink_missing(c, mode="sample")
        m_, o_, f_ = kat.LARGE_EXAMPLE_SAMPLE_MISSING
        assert (ms["MISSING_CT"] == m_).all() and (ms["OBS_CT"] == o_).all() and (ms["F_MISS"] == f_).all()
    with _ctx("all_missing") as c:  # test/sql/plink_missing.test:187-188
        m = ops.plink_missing(c)
        assert list(m["MISSING_CT"]) == [2, 2] and list(m["OBS_CT"]) == [0, 0] and list(m["F_MISS"]) == [1.0, 1.0]


def test_plink_score_known_answers():
    from plinking_duck_b200 import ops

    with _ctx("pgen_example") as c:
        s = ops.plink_score(c, np.arange(4), kat.SCORE_W)
        assert np.allclose(s["SCORE_SUM"], kat.SCORE_SUM) and np.allclose(s["SCORE_AVG"], kat.SCORE_AVG)
        assert np.allclose(s["NAMED_ALLELE_DOSAGE_SUM"], kat.SCORE_DOSAGE_SUM)
        assert (s["ALLELE_CT"] == 8).all() and (s["DENOM"] == 8).all()
        s = ops.plink_score(c, np.arange(4), kat.SCORE_W, no_mean_imputation=True)
        for i, (act, ss, ds) in kat.SCORE_NOIMPUTE.items():
            assert (int(s["ALLELE_CT"][i]), float(s["SCORE_SUM"][i]), float(s["NAMED_ALLELE_DOSAGE_SUM"][i])) == (act, ss, ds)
        s = ops.plink_score(c, [1], [1.0], center=True)
        for i, v in kat.SCORE_CENTER_RS2.items():
            assert abs(s["SCORE_SUM"][i] - v) < 1e-15
        # zero weights are dropped at bind (src/plink_score.cpp:401): ALLELE_CT counts 3 variants
        s = ops.plink_score(c, [3, 0, 2, 1], [2.0, 0.0, -0.5, 0.5])
        assert (s["ALLELE_CT"] == 6).all()
        # flip: weight on REF allele = 2 - dosage (test/sql/plink_score.test flip cases)
        a = ops.plink_score(c, [1], [1.0])
        b = ops.plink_score(c, [1], [1.0], flip=[1])
        assert np.allclose(a["SCORE_SUM"] + b["SCORE_SUM"], 2.0)
        with pytest.raises(ValueError):
            ops.plink_score(c, [0, 1], [1.0])
    with _ctx("all_missing") as c:
        s = ops.plink_score(c, [0, 1], [1.0, 1.0])
        assert (s["SCORE_SUM"] == 0).all() and (s["ALLELE_CT"] == 0).all() and (s["SCORE_AVG"] == 0).all()


def test_sex_chromosomes_known_answers():
    """test/sql/plink_sexchr.test: chrX / chrY / chrMT ploidy-aware freq and hardy."""
    from plinking_duck_b200 import ops

    name = "sexchr_example"
    sex = _psam_sex(name)
    pv = _pvar(name)
    ids = [r[2] for r in pv]
    ploidy = np.array([ops.classify_ploidy(ch, pos) for ch, pos, _ in pv], dtype=np.uint8)
    with _ctx(name) as c:
        f = ops.plink_freq(c, ploidy=ploidy, sex=sex)
        h = ops.plink_hardy(c, ploidy=ploidy, sex=sex)
        hm = ops.plink_hardy(c, ploidy=ploidy, sex=sex, midp=True)
        f_nosex = ops.plink_freq(c, ploidy=ploidy, sex=None)
    row = {i: k for k, i in enumerate(ids)}
    for vid, (af, obs) in kat.SEXCHR_FREQ.items():
        assert abs(f["ALT_FREQ"][row[vid]] - af) < 5e-7 and f["OBS_CT"][row[vid]] == obs
    assert abs(f["ALT_FREQ"][row["par1"]] - 0.333333) < 5e-7 and f["OBS_CT"][row["par1"]] == 12  # :30-36
    y = f[row["y1"]]
    assert (y["HOM_REF_CT"], y["HET_CT"], y["HOM_ALT_CT"], y["MISSING_CT"]) == (1, 0, 2, 3)  # :151
    a1 = h[row["a1"]]
    assert (a1["HOM_REF_CT"], a1["HET_CT"], a1["HOM_ALT_CT"]) == (2, 2, 2) and abs(a1["P_HWE"] - 0.480519) < 5e-7
    p1 = h[row["par1"]]
    assert (p1["HOM_REF_CT"], p1["HET_CT"], p1["HOM_ALT_CT"]) == (3, 2, 1) and abs(p1["E_HET"] - 0.444444) < 5e-7
    x1 = h[row["x1"]]
    assert (x1["HOM_REF_CT"], x1["HET_CT"], x1["HOM_ALT_CT"]) == (1, 1, 1)
    assert abs(x1["O_HET"] - 0.333333) < 5e-7 and abs(x1["E_HET"] - 0.5) < 5e-7 and abs(x1["P_HWE"] - 1.0) < 5e-7
    assert abs(hm["P_HWE"][row["x1"]] - kat.XCHR_X1["p_midp"]) < 5e-7
    y1 = h[row["y1"]]
    assert (y1["HOM_REF_CT"], y1["HET_CT"], y1["HOM_ALT_CT"], y1["stats_are_null"]) == (1, 0, 2, 1)
    m1 = h[row["mt1"]]
    assert (m1["HOM_REF_CT"], m1["HET_CT"], m1["HOM_ALT_CT"], m1["stats_are_null"]) == (2, 0, 4, 1)
    # no sex information: chrX / chrY rows are NULL, MT still works (sex-independent)
    assert f_nosex["freq_is_null"][row["x1"]] and f_nosex["counts_are_null"][row["y1"]]
    assert not f_nosex["freq_is_null"][row["mt1"]]


def _xchr_bruteforce(f0, fh, f2, ma, mb, midp):
    """Definition of the test (Graffelman & Weir 2016) by full enumeration, for small tables."""
    from math import exp, lgamma, log

    nf, nm, nB = f0 + fh + f2, ma + mb, 2 * f2 + fh + mb

    def lp(mbx, fhx):
        fb = nB - mbx
        if fb < 0 or fb > 2 * nf or (fb - fhx) % 2 or fhx > fb:
            return None
        f2x = (fb - fhx) // 2
        f0x = nf - fhx - f2x
        if f0x < 0:
            return None
        return fhx * log(2) - lgamma(mbx + 1) - lgamma(nm - mbx + 1) - lgamma(f0x + 1) - lgamma(fhx + 1) - lgamma(f2x + 1)

    lo = lp(mb, fh)
    terms = [exp(l - lo) for mbx in range(nm + 1) for fhx in range(nf + 1) for l in [lp(mbx, fhx)] if l is not None]
    tail = sum(t for t in terms if t <= 1 + 1e-9)
    ties = sum(1 for t in terms if abs(t - 1) < 1e-9)
    return (tail - (0.5 * ties if midp else 0.0)) / sum(terms)


def test_xchr_exact_test_matches_reference():
    """pd_hwe_xchr_from_counts vs plink2::HweXchrLnP over random tables.  Standard p-values agree to the
    stated tolerance everywhere.  For mid-p, HweXchrLnP subtracts half of every table tied with the observed
    one, but finds ties through accumulated log() sums and misses some exact ties at small counts (about 6 % of
    tied tables in a random sweep); the kernel counts every tie, so where the two differ the kernel must equal
    the full enumeration of the test's definition."""
    import hwe_check as hc
    from oracle import ref

    if not ref.available():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(11)
    tabs = []
    for _ in range(400):
        n = int(rng.choice([3, 6, 10, 40, 200, 3000, 40000]))
        p = rng.uniform(0.02, 0.5)
        nf, nm = int(rng.integers(0, n + 1)), int(rng.integers(0, n + 1))
        f = rng.multinomial(nf, [(1 - p) ** 2, 2 * p * (1 - p), p * p])
        if rng.random() < 0.3 and nf:  # some tables far from equilibrium
            f = rng.multinomial(nf, [0.45, 0.1, 0.45])
        mb = int(rng.binomial(nm, p))
        tabs.append([f[0], f[1], f[2], nm - mb, mb])
    tabs += [[1, 1, 1, 2, 1], [0, 0, 0, 0, 0], [5, 0, 0, 3, 0], [0, 0, 0, 4, 3], [2, 3, 1, 0, 0], [2, 4, 3, 6, 4]]
    tabs = np.array(tabs, dtype=np.uint32)
    tie_quirks = 0
    worst = 0.0
    with _ctx("pgen_example") as c:
        for midp in (False, True):
            got = c.hwe_xchr_from_counts(tabs, midp=midp)
            for k, t in enumerate(tabs):
                if t.sum() == 0:
                    assert got[k] == 1.0
                    continue
                lnp = ref.hwe_xchr_lnp(t[1], t[0], t[2], t[3], t[4], midp)
                exp = float(ref.lnp_to_pvalue(lnp))
                # RELATIVE bound (no absolute floor: small p-values are compared too).  Both sides anchor rows
                # with log-gamma sums of magnitude ~ 2n ln 2n, so the bound is stated in ulps of that:
                # 128 x the autosomal reference noise, floored at 1e-12; p below fp64 must come out 0.
                tol = max(1e-12, 128.0 * float(hc.ref_bound(int(t.sum()))))
                if lnp < hc.LN_1E300:
                    assert got[k] <= 2e-300, (t, midp, got[k], lnp)
                    continue
                if abs(got[k] - exp) <= tol * exp:
                    worst = max(worst, abs(got[k] - exp) / exp if exp > 0 else 0.0)
                    continue
                assert midp and int(t.sum()) <= 100, (t, midp, got[k], exp)
                bf = _xchr_bruteforce(*[int(x) for x in t], True)
                assert abs(got[k] - max(bf, 0.0)) <= 1e-12, (t, got[k], exp, bf)
                tie_quirks += 1
    assert tie_quirks <= 20
    print(f"xchr: worst relative error vs HweXchrLnP {worst:.3g}")



def test_genotype_counts_and_stats_struct_known_answers(golden):
    """read_pfile / read_pgen with genotypes := 'counts' | 'stats', both orients
    (test/sql/read_pfile_genotypes_stats.test, read_pfile_genotypes_counts.test,
    read_pfile_sample_counts_sparse.test:37-41)."""
    from plinking_duck_b200 import ops

    cnt = ("hom_ref", "het", "hom_alt", "missing")
    with _ctx("pgen_example") as c:
        v = ops.genotype_stats(c)
        for r, k in zip(v, kat.GENO_STATS_VARIANT):
            assert tuple(int(r[f]) for f in cnt) == k[:4] and int(r["n"]) == k[4] and int(r["carrier_count"]) == k[8]
            assert (round(float(r["af"]), 6), round(float(r["maf"]), 6), round(float(r["missing_rate"]), 6),
                    round(float(r["het_rate"]), 6)) == (k[5], k[6], k[7], k[9])
        # "Cross-check af against plink_freq -- values must match exactly" (:56-66)
        assert np.array_equal(v["af"], ops.plink_freq(c)["ALT_FREQ"])
        s = ops.genotype_stats(c, orient="sample")
        assert [tuple(int(r[f]) for f in cnt) for r in s] == kat.GENO_COUNTS_SAMPLE
        assert [(int(r["n"]), round(float(r["af"]), 4), round(float(r["missing_rate"]), 4), int(r["carrier_count"]))
                for r in s] == kat.GENO_STATS_SAMPLE
        # variant counts and sample counts tally the same carriers (read_pfile_genotypes_counts.test:105-113)
        assert int(v["carrier_count"].sum()) == int(s["carrier_count"].sum())
        c.set_subset([0, 2])
        v = ops.genotype_stats(c)
        assert [tuple(int(r[f]) for f in cnt) for r in v] == kat.GENO_COUNTS_SUBSET_1_3
        assert [(int(r["hom_ref"]), int(r["het"]), round(float(r["af"]), 4), round(float(r["missing_rate"]), 4))
                for r in v] == kat.GENO_STATS_SUBSET_1_3
        s = ops.genotype_stats(c, orient="sample")
        assert len(s) == 2 and [tuple(int(r[f]) for f in cnt) for r in s] == [kat.GENO_COUNTS_SAMPLE[0], kat.GENO_COUNTS_SAMPLE[2]]
        with pytest.raises(ValueError):
            ops.genotype_stats(c, orient="genotype")
    with _ctx("rare_small") as c:
        s = ops.genotype_stats(c, orient="sample")
        assert tuple(int(s[f].sum()) for f in cnt) == kat.RARE_SMALL_SUMS
        n_, cc_, hr_ = kat.RARE_SMALL_STATS
        assert (int(s["n"].sum()), int(s["carrier_count"].sum())) == (n_, cc_)
        assert round(float(s["het"].sum()) / float(s["n"].sum()), 6) == hr_
        # empty variant range: all-zero counts, NaN rates (read_pfile_sample_counts_sparse.test:49-53)
        e = ops.genotype_stats(c, 0, 0, orient="sample")
        assert len(e) == 256 and int(e["het"].sum()) == 0 and int(e["hom_ref"].sum()) == 0
        assert np.isnan(e["af"]).all() and np.isnan(e["missing_rate"]).all()
    with _ctx("all_missing") as c:  # n == 0: NaN af / maf / het_rate, missing_rate 1
        v = ops.genotype_stats(c)
        assert np.isnan(v["af"]).all() and np.isnan(v["maf"]).all() and np.isnan(v["het_rate"]).all()
        assert list(v["missing_rate"]) == [1.0, 1.0] and list(v["n"]) == [0, 0]
    with _ctx("large_example") as c:  # read_pfile_genotypes_stats.test:145-148 and the pgenlib golden counts
        v = ops.genotype_stats(c)
        assert len(v) == 3000
        assert np.array_equal(np.stack([v[f] for f in cnt], axis=1), golden["large_example/counts"])


def test_predecomp_filters_on_gpu_counts(golden):
    """CheckPreDecompFilters (src/plink_common.cpp:1495-1515) over GPU counts: af / ac ranges and the
    genotype-category filter vs the same rules evaluated with numpy on pgenlib's golden counts."""
    from plinking_duck_b200 import ops

    name = "syn_n1000"
    cnt = golden[name + "/counts"].astype(np.int64)
    nm = cnt[:, 0] + cnt[:, 1] + cnt[:, 2]
    ac = cnt[:, 1] + 2 * cnt[:, 2]
    af = np.where(nm > 0, ac / np.maximum(2.0 * nm, 1), 0.0)
    with _ctx(name) as c:
        skip, allp = ops.predecomp_filter(c, af=(0.05, 0.4))
        exp = (nm == 0) | (af < 0.05) | (af > 0.4)
        assert np.array_equal(skip.astype(bool), exp) and np.all(allp[~exp] == 1)
        skip, _ = ops.predecomp_filter(c, ac=(10, None))
        assert np.array_equal(skip.astype(bool), (nm == 0) | (ac < 10))
        skip, allp = ops.predecomp_filter(c, allowed=(1, 2))
        any_pass = (cnt[:, 1] > 0) | (cnt[:, 2] > 0)
        assert np.array_equal(skip.astype(bool), ~any_pass)
        assert np.array_equal(allp[any_pass].astype(bool), (cnt[:, 0] == 0)[any_pass])
        skip, _ = ops.predecomp_filter(c, allowed=(2,), include_missing=True)
        assert np.array_equal(skip.astype(bool), ~((cnt[:, 2] > 0) | (cnt[:, 3] > 0)))
        skip, _ = ops.predecomp_filter(c, 10, 50, af=(0.0, 1.0), allowed=(0, 1, 2))
        assert skip.shape == (40,) and np.array_equal(skip.astype(bool), (nm == 0)[10:50])


def _scan_genotype_column(i8, counts, af, ac, allowed, include_missing):
    """The per-variant flow of the reference's read_pgen scan restated on pgenlib's goldens:
    CheckPreDecompFilters (src/plink_common.cpp:1495-1515) on the counts, skip, then the child-vector fill
    (src/pgen_reader.cpp:1008-1025) with the per-row geno_range_all_pass short-cut."""
    from oracle import port

    kept, rows = [], []
    for v in range(i8.shape[0]):
        c = [int(x) for x in counts[v]]
        nm = c[0] + c[1] + c[2]
        all_pass = True
        if af is not None or ac is not None:
            a = c[1] + 2 * c[2]
            if nm == 0:
                continue
            f = a / (2.0 * nm)
            lo, hi = af or (None, None)
            if (lo is not None and f < lo) or (hi is not None and f > hi):
                continue
            lo, hi = ac or (None, None)
            if (lo is not None and a < lo) or (hi is not None and a > hi):
                continue
        if allowed is not None:
            anyp = any(c[g] > 0 for g in allowed) or (include_missing and c[3] > 0)
            if not anyp:
                continue
            all_pass = all(c[g] == 0 for g in (0, 1, 2) if g not in allowed) and (include_missing or c[3] == 0)
        kept.append(v)
        mask = (True, True, True) if (allowed is None or all_pass) else tuple(g in allowed for g in (0, 1, 2))
        rows.append(port.child_vector_layout(i8[v : v + 1], mask)[0][0])
    data = np.array(rows, dtype=np.int8).reshape(len(kept), i8.shape[1])
    valid = data != 0
    # validity from the same rule, flattened over the kept rows
    ok = np.zeros(data.shape, dtype=bool)
    for r, v in enumerate(kept):
        g = i8[v]
        m = (True, True, True) if allowed is None else tuple(gg in allowed for gg in (0, 1, 2))
        ok[r] = (g != -9) & np.isin(g, [gg for gg in (0, 1, 2) if m[gg]])
    bits = np.packbits(ok.reshape(-1), bitorder="little")
    words = np.zeros(((ok.size + 63) // 64) * 8, dtype=np.uint8)
    words[: bits.size] = bits
    del valid
    return np.array(kept, dtype=np.uint32), data, words.view(np.uint64)


@pytest.mark.parametrize("name", ["pgen_example", "syn_n1000", "rare_small", "streaming_example", "pca_example", "all_missing"])
def test_genotype_column_op_matches_scan_flow(name, golden):
    """pdop_genotype_column = counts -> CheckPreDecompFilters -> compacted decode into DuckDB's child layout."""
    from plinking_duck_b200 import ops

    if name + "/i8" not in golden:
        pytest.skip("no int8 golden")
    i8, cnt = golden[name + "/i8"], golden[name + "/counts"]
    cases = [
        dict(af=None, ac=None, allowed=None, include_missing=False),
        dict(af=(0.05, 0.45), ac=None, allowed=None, include_missing=False),
        dict(af=None, ac=(2, None), allowed=(1, 2), include_missing=False),
        dict(af=None, ac=None, allowed=(2,), include_missing=True),
        dict(af=None, ac=None, allowed=(0,), include_missing=False),
    ]
    with _ctx(name, tile_bytes=1200, n_staging=2) as c:
        for kw in cases:
            ek, ed, ev = _scan_genotype_column(i8, cnt, **kw)
            k, d, v = ops.genotype_column(c, **kw)
            assert np.array_equal(k, ek), kw
            assert np.array_equal(d, ed), kw
            assert np.array_equal(v, ev), kw
