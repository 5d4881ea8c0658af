"""CPU: pins oracle/pgen_oracle.c (the plain-C restatement of the reference path) against
(a) the reference's sqllogictest known answers (tests/kat.py) and (b) golden.npz, i.e. the outputs of the
reference's OWN pgenlib over every fixture: index, counts (all / subset), int8 decode, missingness, the
exact HWE test (vs HweLnP, within the restatement's stated fp64 noise) and plink_score in all three modes."""
import subprocess
import os

import numpy as np
import pytest

import kat
from conftest import ROOT, fixture_names, fixture_path


@pytest.fixture(scope="module")
def port():
    subprocess.run(["make", "-s", "port"], cwd=os.path.join(ROOT, "oracle"), check=True)
    from oracle import port as p

    return p


def test_port_kat(port):
    with port.PortPgen(fixture_path("pgen_example")) as r:
        c = r.counts()
        assert np.array_equal(c, kat.PGEN_EXAMPLE_COUNTS)
        assert list(r.missing_by_sample()) == kat.PGEN_EXAMPLE_MISSING_BY_SAMPLE
        s, d, a = r.score(np.arange(4), None, kat.SCORE_W, 0)
        assert np.allclose(s, kat.SCORE_SUM) and np.allclose(d, kat.SCORE_DOSAGE_SUM) and (a == 8).all()
        s, d, a = r.score(np.arange(4), None, kat.SCORE_W, 1)
        for i, (act, ss, ds) in kat.SCORE_NOIMPUTE.items():
            assert (int(a[i]), float(s[i]), float(d[i])) == (act, ss, ds)
        s, d, a = r.score(np.array([1]), None, np.array([1.0]), 2)
        for i, v in kat.SCORE_CENTER_RS2.items():
            assert abs(s[i] - v) < 1e-15
    p = [port.hwe_exact(int(x[1]), int(x[0]), int(x[2])) for x in kat.PGEN_EXAMPLE_COUNTS]
    pm = [port.hwe_exact(int(x[1]), int(x[0]), int(x[2]), True) for x in kat.PGEN_EXAMPLE_COUNTS]
    assert np.allclose(p, kat.PGEN_EXAMPLE_P_HWE, atol=5e-7)
    assert np.allclose(pm, kat.PGEN_EXAMPLE_P_HWE_MIDP, atol=5e-7)
    with port.PortPgen(fixture_path("all_missing")) as r:
        assert np.array_equal(r.counts(), kat.ALL_MISSING_COUNTS)
    with port.PortPgen(fixture_path("rare_small")) as r:
        assert tuple(int(x) for x in r.counts().sum(0)) == kat.RARE_SMALL_SUMS


@pytest.mark.parametrize("name", fixture_names())
def test_port_matches_reference_goldens(name, golden, port):
    with port.PortPgen(fixture_path(name)) as r:
        vt, fp = r.index()
        assert np.array_equal(vt, golden[name + "/vrtypes"])
        assert np.array_equal(fp, golden[name + "/fpos"])
        c = r.counts()
        assert np.array_equal(c, golden[name + "/counts"])
        assert np.array_equal(r.missing_by_sample(), golden[name + "/missing_by_sample"])
        assert np.array_equal(c[:, 3], golden[name + "/missing_by_variant"])
        if name + "/i8" in golden:
            assert np.array_equal(r.decode_i8(), golden[name + "/i8"])
        if name + "/subset" in golden:
            sub = golden[name + "/subset"]
            assert np.array_equal(r.counts(subset=sub), golden[name + "/subset_counts"])
        # exact test vs HweLnP: |p - exp(lnp)| within the noise of HweLnP's log-factorials
        lnp = golden[name + "/hwe_lnp"]
        n2 = 2.0 * c[:, :3].sum(1).astype(float)
        tol = 1e-13 + 16 * 2.2e-16 * n2 * np.log(n2 + 2.0)
        p = np.array([port.hwe_exact(int(x[1]), int(x[0]), int(x[2])) for x in c])
        assert np.all(np.abs(p - np.clip(np.exp(lnp), 0, 1)) <= tol)
        if name + "/score_m0_sum" in golden and not (vt & 0x60).any():
            w, fl = golden[name + "/score_w"], golden[name + "/score_flip"]
            for mode in (0, 1, 2):
                s, d, a = r.score(np.arange(r.variant_ct), fl, w, mode)
                es, ed, ea = (golden[name + f"/score_m{mode}_{k}"] for k in ("sum", "dsum", "ct"))
                assert np.array_equal(a, ea.reshape(-1))
                assert np.allclose(s, es.reshape(-1), rtol=1e-12, atol=1e-9)
                assert np.allclose(d, ed.reshape(-1), rtol=1e-12, atol=1e-9)


def test_all_ref_vrtype5_records(port, tmp_path):
    """vrtype 5 (all-ref, zero-length record) never leaves pgenlib's stock writer, so the fixtures lack it:
    rewrite a pgenlib-written file with those records (synth.rewrite_with_all_ref_records) and require the
    reference's reader and the port to agree with the genotype matrix on it."""
    import synth
    from oracle import ref

    if not ref.available():
        pytest.skip("oracle/_ref not built: no pgenlib writer")
    V, N = 300, 257
    g = synth.gen_rare_genotypes(V, N, 1)
    g[0] = 0
    g[1] = 2
    g[2] = 3
    g[10:14] = 0
    a, b = str(tmp_path / "a.pgen"), str(tmp_path / "b.pgen")
    ref.write_pgen(a, synth.to_genovec_words(g), N)
    assert synth.rewrite_with_all_ref_records(a, b) >= 5
    with ref.RefPgen(b) as r:
        assert 5 in set(int(x) & 7 for x in r.vrtypes())
        assert np.array_equal(r.counts(), synth.counts_numpy(g))
    with port.PortPgen(b) as p:
        assert np.array_equal(p.counts(), synth.counts_numpy(g))
        assert np.array_equal(p.missing_by_sample(), (g == 3).sum(0))
        assert np.array_equal(p.decode_i8(), np.where(g == 3, -9, g))


def test_phased_decode_kat_and_pgenlib(port, tmp_path):
    """PgrGetP + UnpackPhasedGenotypes: the reference's known answers for phased_example
    (test/sql/read_pgen_phased.test:45-63), then pgenlib-written files with implicit (all hets phased) and
    explicit (phasepresent) aux2 tracks on every main record type: pgenlib's reader, the port and the
    generator's own expectation must agree."""
    import synth
    from oracle import ref

    with port.PortPgen(fixture_path("phased_example")) as p:
        assert p.decode_phased().tolist() == kat.PHASED_EXAMPLE_PAIRS
    if not ref.available():
        pytest.skip("oracle/_ref not built: no pgenlib writer")
    with ref.RefPgen(fixture_path("phased_example")) as r:
        assert r.decode_phased().tolist() == kat.PHASED_EXAMPLE_PAIRS
    for (V, N) in [(90, 300), (60, 257), (30, 5000)]:
        g = synth.gen_mixed_genotypes(V, N, seed=V)
        phase, exp = synth.gen_phase(g, 7)
        path = str(tmp_path / f"ph_{V}_{N}.pgen")
        ref.write_pgen(path, synth.to_genovec_words(g), N, phase=phase)
        with ref.RefPgen(path) as r:
            assert np.array_equal(r.decode_phased(), exp)
            assert any(int(x) & 0x10 for x in r.vrtypes())
        with port.PortPgen(path) as p:
            assert np.array_equal(p.decode_phased(), exp)
            assert np.array_equal(p.decode_phased(5, 17), exp[5:17])


def test_child_vector_layout_restatement(port, golden):
    """The DuckDB child-vector fill loop (data + validity): vectorised restatement == the literal loop, on the
    reference's own fixture and for every genotype-filter mask."""
    g = golden["pgen_example/i8"]
    d, v = port.child_vector_layout(g)
    assert d.tolist() == np.where(g == -9, 0, g).tolist()
    assert int(sum(bin(int(x)).count("1") for x in v)) == int((g != -9).sum())
    rng = np.random.default_rng(3)
    for rows, n in [(1, 1), (3, 7), (5, 64), (4, 129), (9, 100)]:
        m = rng.choice(np.array([0, 1, 2, -9], dtype=np.int8), size=(rows, n))
        for mask in range(8):
            allowed = tuple(bool(mask >> k & 1) for k in range(3))
            d1, v1 = port.child_vector_layout(m, allowed)
            d2, v2 = port.child_vector_layout_loop(m, allowed)
            assert np.array_equal(d1, d2) and np.array_equal(v1, v2)
