"""GPU parity: dosage decode (PgrGetD + Dosage16ToDoublesMinus9), dosage-aware counts (PgrGetDCounts)
and plink_score through the C ABI vs pgenlib goldens.

Integer outputs (counts, all_dosages numerators, allele counts) are bit-exact.  Decoded dosages and
IMP_R2 are computed with the reference's own fp64 operation order and must be bit-identical too.
Score sums: relative tolerance 1e-12 of the sum's natural scale (sum of |w| x max dosage 2), the
north_star's fp64 bound -- the reference's own sums depend on thread arrival order
(src/plink_score.cpp:657-664), so tolerance parity is the only meaningful kind."""
import numpy as np
import pytest

import synth
from conftest import fixture_names, fixture_path

pytestmark = pytest.mark.gpu

SCORE_RTOL = 1e-12


def _ctx(name, **kw):
    from plinking_duck_b200.pgen import PgenContext

    return PgenContext(fixture_path(name), **kw)


def _dosage_fixtures(golden):
    return [n for n in fixture_names() if n + "/dcounts" in golden]


def pytest_generate_tests(metafunc):
    if "dname" in metafunc.fixturenames:
        import os

        g = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden.npz"))
        metafunc.parametrize("dname", [n for n in fixture_names() if n + "/dcounts" in g])


@pytest.mark.parametrize("mode", ["stream", "stream_small_tiles", "resident"])
def test_dosage_counts(dname, mode, golden):
    kw = dict(tile_bytes=600, n_staging=2) if mode == "stream_small_tiles" else {}
    with _ctx(dname, **kw) as c:
        if mode == "resident":
            c.load_resident()
        counts, ad, r2 = c.dosage_counts()
    assert np.array_equal(counts, golden[dname + "/dcounts"])
    assert np.array_equal(ad, golden[dname + "/all_dosages"])
    exp = golden[dname + "/imp_r2"]
    assert np.array_equal(np.isnan(r2), np.isnan(exp))
    ok = ~np.isnan(exp)
    assert np.array_equal(r2[ok], exp[ok])


def test_dosage_example_known_answers():
    """test/sql/plink_freq_dosage.test:103-118: ALT_FREQ / OBS_CT / IMP_R2 of dosage_example."""
    with _ctx("dosage_example") as c:
        counts, ad, r2 = c.dosage_counts()
    ad = ad.astype(np.float64)
    alt_freq = ad[:, 1] / (ad[:, 0] + ad[:, 1])
    obs_ct = (ad[:, 0] + ad[:, 1]) / 16384.0
    assert np.allclose(alt_freq, [0.4583333333333333, 0.4499969482421875, 0.5, 0.20833333333333334], rtol=1e-15)
    assert np.array_equal(obs_ct, [6, 8, 8, 6])
    assert np.allclose(r2, [1.2900887786091624, 0.7374153374407626, 1.0, 0.3857199844561125], rtol=1e-15)


def test_dosage_f64(dname, golden):
    exp = golden[dname + "/dosage_first"]
    nd = exp.shape[0]
    with _ctx(dname) as c:
        got = c.dosage_f64(np.arange(nd))
        assert np.array_equal(got, exp)
        # sparse list + resident
        pick = np.arange(0, nd, 3)
        c.load_resident()
        assert np.array_equal(c.dosage_f64(pick), exp[pick])


def test_dosage_subset(golden):
    from oracle import ref

    if not ref.available():
        pytest.skip("oracle/_ref not built")
    for name in ("syn_dosage", "dosage_example", "syn_n257"):
        with ref.RefPgen(fixture_path(name)) as r, _ctx(name) as c:
            N = r.N
            sub = np.arange(0, N, 2) if N > 4 else np.array([1, 3])
            r.set_subset(sub)
            c.set_subset(sub)
            exp = np.stack([r.dosage(v) for v in range(r.V)])
            assert np.array_equal(c.dosage_f64(np.arange(r.V)), exp), name
            ec, ead, er2 = r.dcounts()
            gc, gad, gr2 = c.dosage_counts()
            assert np.array_equal(gc, ec) and np.array_equal(gad, ead), name
            ok = ~np.isnan(er2)
            assert np.array_equal(gr2[ok], er2[ok]) and np.array_equal(np.isnan(gr2), np.isnan(er2)), name


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_score_matches_oracle(dname, mode, golden):
    w = golden[dname + "/score_w"]
    fl = golden[dname + "/score_flip"]
    V = w.size
    with _ctx(dname) as c:
        s, d, a = c.score(np.arange(V), fl, w, mode)
    scale = 2.0 * np.abs(w).sum() + 1.0
    es, ed, ea = (golden[dname + f"/score_m{mode}_{k}"] for k in ("sum", "dsum", "ct"))
    assert np.array_equal(a, ea)
    assert np.max(np.abs(s[:, 0] - es)) <= SCORE_RTOL * scale * (50.0 if mode == 2 else 1.0)
    assert np.max(np.abs(d - ed)) <= SCORE_RTOL * (2.0 * V + 1.0)


def test_score_known_answers():
    """test/sql/plink_score.test:48-75: weights [1.0, 0.5, -0.5, 2.0] on pgen_example ->
    SCORE_SUM -0.5 / 1.0 / 3.5 / 6.0, NAMED_ALLELE_DOSAGE_SUM 3 / 3 / 4 / 5."""
    w = np.array([1.0, 0.5, -0.5, 2.0])
    with _ctx("pgen_example") as c:
        s, d, a = c.score(np.arange(4), None, w, 0)
    assert np.allclose(s[:, 0], [-0.5, 1.0, 3.5, 6.0], rtol=1e-15, atol=1e-15)
    assert np.allclose(d, [3, 3, 4, 5])
    assert np.array_equal(a, [8, 8, 8, 8])
    assert np.allclose(s[:, 0] / a, [-0.0625, 0.125, 0.4375, 0.75])


def test_score_multi_column_and_subset(golden):
    """K = 16 columns equal 16 single-column runs; all-ones weights give SCORE_SUM == NAMED_ALLELE_DOSAGE_SUM
    (test/sql/streaming_threading.test:226-233)."""
    name = "syn_n1000"
    rng = np.random.default_rng(3)
    with _ctx(name) as c:
        V = c.variant_ct
        vid = np.sort(rng.choice(V, size=V // 2, replace=False)).astype(np.uint32)
        W = rng.standard_normal((vid.size, 16))
        fl = (rng.random(vid.size) < 0.2).astype(np.uint8)
        S, D, A = c.score(vid, fl, W, 0)
        for k in (0, 7, 15):
            s1, d1, a1 = c.score(vid, fl, W[:, k], 0)
            assert np.allclose(S[:, k], s1[:, 0], rtol=1e-13, atol=1e-10)
            assert np.array_equal(a1, A) and np.allclose(d1, D)
        ones, dn, _ = c.score(vid, None, np.ones(vid.size), 0)
        assert np.allclose(ones[:, 0], dn, rtol=1e-13)
        sub = np.arange(0, c.raw_sample_ct, 3)
        c.set_subset(sub)
        Ss, Ds, As = c.score(vid, fl, W[:, :3], 1)
        assert Ss.shape == (sub.size, 3)
    from oracle import ref

    if ref.available():
        with ref.RefPgen(fixture_path(name)) as r:
            r.set_subset(sub)
            for k in range(3):
                es, ed, ea = r.score(vid, fl, W[:, k], 1)
                assert np.allclose(Ss[:, k], es, rtol=1e-12, atol=1e-9)
                assert np.array_equal(As, ea)


def _score_numpy(g, w, flip, mode):
    """src/plink_score.cpp:598-653 restated with numpy in fp64 (hardcalls only): per scored variant
    mean = sum_alt / non_missing; skip if non_missing == 0; mode 0 missing -> mean, mode 1 missing skipped,
    mode 2 (d' - mean') / sqrt(2 f (1 - f)) with missing contributing nothing."""
    V, N = g.shape
    K = w.shape[1]
    score = np.zeros((N, K))
    dsum = np.zeros(N)
    act = np.zeros(N, dtype=np.uint32)
    for v in range(V):
        row = g[v]
        nm = row != 3
        n = int(nm.sum())
        if n == 0:
            continue
        mean = float(row[nm].sum()) / n
        d = row.astype(np.float64)
        if mode == 2:
            f = mean / 2.0
            sd = np.sqrt(2.0 * f * (1.0 - f))
            if sd == 0.0:
                continue
            dd = (2.0 - d) if flip[v] else d
            mm = (2.0 - mean) if flip[v] else mean
            val = np.where(nm, (dd - mm) / sd, 0.0)
            act += 2 * nm.astype(np.uint32)
        else:
            if mode == 0:
                d = np.where(nm, d, mean)
                cnt = np.ones(N, dtype=bool)
            else:
                cnt = nm
            dd = (2.0 - d) if flip[v] else d
            val = np.where(cnt, dd, 0.0)
            dsum += val
            act += 2 * cnt.astype(np.uint32)
        score += val[:, None] * w[v][None, :]
    return score, dsum, act


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("N,K", [(1000, 16), (257, 1), (2049, 5)])
def test_score_tensor_core_path_vs_fp64_loop(N, K, mode, tmp_path):
    """K7t (tcgen05 int8 limbs) on a fixed-width hardcall file: sample tiles that end inside a tile, k-row
    splits, variants without missing calls, an all-missing variant and a monomorphic one (sd = 0 in mode 2).
    Tolerance: the stated fp64 bound 1e-12 relative to the magnitude of the sum of |terms|."""
    V = 700
    g = synth.gen_hwe_genotypes(V, N, seed=11 + N, missing=0.02)
    g[5] = 3            # all missing: skipped
    g[6] = 0            # monomorphic
    g[7][g[7] == 3] = 1  # no missing calls: no M row
    rng = np.random.default_rng(N + K)
    w = rng.standard_normal((V, K)) * np.exp(rng.uniform(-6, 6, size=(1, K)))  # columns of very different scale
    fl = (rng.random(V) < 0.3).astype(np.uint8)
    from plinking_duck_b200.pgen import PgenContext

    with PgenContext(image=synth.fixed_pgen_image(g), device=0) as c:
        c.load_resident()
        S, D, A = c.score(np.arange(V, dtype=np.uint32), fl, w, mode)
        st = c.stats()
    es, ed, ea = _score_numpy(g, w, fl, mode)
    assert np.array_equal(A, ea)
    bound = 1e-12 * (np.abs(w).sum(axis=0) * 2.0 + 1.0)
    assert np.all(np.abs(S - es) <= bound[None, :]), float(np.max(np.abs(S - es) / bound[None, :]))
    if mode != 2:
        assert np.allclose(D, ed, rtol=1e-13, atol=1e-9)
    assert st.launches > 0
    from oracle import ref

    if ref.available():
        path = str(tmp_path / "tc.pgen")
        synth.write_fixed_pgen(path, g)
        with ref.RefPgen(path) as r:
            rs, rd, ra = r.score(np.arange(V, dtype=np.uint32), fl, w[:, 0], mode)
        assert np.array_equal(A, ra)
        assert np.all(np.abs(S[:, 0] - np.asarray(rs).reshape(-1)) <= bound[0])


def test_score_tensor_core_multibatch_ksplit_vs_reference_loop(tmp_path):
    """K7t at scale: 150 000 scored variants (three batches of <= 65 535, accumulated across launches) x 20 011
    samples (79 sample tiles -> the k-rows of a batch are split 7 ways over the grid, partial sums meet in the int32
    accumulators) x 16 weight columns, against the reference's own loop (PgrGetD + Dosage16ToDoublesMinus9 + the
    accumulation of src/plink_score.cpp:598-653 run per column on every host thread, oracle/ref_harness.cc) for the
    default mode, and against exact integer identities for no_mean_imputation.  Tolerance: 1e-12 (2 sum|w| + 1)."""
    import os
    import tempfile

    import torch

    from bench import generate_records_gpu
    from oracle import ref
    from plinking_duck_b200.pgen import PgenContext

    if not ref.available():
        pytest.skip("oracle/_ref not built")
    V, N, K = 150_000, 20_011, 16
    Wb = (N + 3) // 4
    dev = torch.device("cuda", 0)
    body = torch.empty(V * Wb, dtype=torch.uint8, device=dev)
    generate_records_gpu(V, N, 4242, dev, body)
    img = np.concatenate([np.frombuffer(synth.fixed_header(V, N), dtype=np.uint8), body.cpu().numpy()])
    del body
    rng = np.random.default_rng(99)
    vidx = np.arange(V, dtype=np.uint32)
    w = rng.standard_normal((V, K)) * np.exp(rng.uniform(-3, 3, size=(1, K)))
    fl = (rng.random(V) < 0.1).astype(np.uint8)
    shm = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    with tempfile.TemporaryDirectory(dir=shm or str(tmp_path)) as d:
        path = os.path.join(d, "score_scale.pgen")
        img.tofile(path)
        with PgenContext(path) as c:
            c.load_resident()
            S, D, A = c.score(vidx, fl, w, 0)
            st = c.stats()
            assert st.scan_launches >= 3  # > 65 535 scored variants: several batches
            S1, D1, A1 = c.score(vidx, None, w, 1)
            het, hom, mis = c.sample_counts()
        _t, rs, ra = ref.bench_score(path, os.cpu_count() or 1, vidx, fl, w)
    assert np.array_equal(A, ra)
    bound = 1e-12 * (np.abs(w).sum(axis=0) * 2.0 + 1.0)
    err = np.abs(S - rs) / bound[None, :]
    assert err.max() <= 1.0, float(err.max())
    # no_mean_imputation, no flips: dosage sums and allele counts are exact integers
    assert np.array_equal(D1, het.astype(np.float64) + 2.0 * hom.astype(np.float64))
    assert np.array_equal(A1, (2 * (V - mis.astype(np.int64))).astype(np.uint32))


def test_single_sample_compact_dosage_mode_vs_pgenlib(tmp_path):
    """header_ctrl mode 12 (pgenlib_read.cc:946-948): a single-sample file whose 2-bit codes select vrtype 5 / 0 /
    0x45 (no hardcall track, one unconditional dosage value) / 0x40 (hardcall byte + dosage value): doubles,
    genotype counts, dosage sums and IMP_R2 against PgrGetD / PgrGetDCounts."""
    import synth
    from oracle import ref
    from plinking_duck_b200.pgen import PgenContext

    if not ref.available():
        pytest.skip("oracle/_ref not built")
    V = 3000
    g = synth.gen_hwe_genotypes(V, 1, seed=12, missing=0.05)
    g[::2] = 0
    image, vrtypes, _ = synth.compact_pgen_image(12, g, seed=5)
    assert {0, 5, 0x40, 0x45} <= set(int(x) for x in vrtypes)
    path = str(tmp_path / "compact12.pgen")
    image.tofile(path)
    with ref.RefPgen(path) as r:
        exp_c, exp_ad, exp_r2 = r.dcounts()
        exp_d = np.stack([r.dosage(v) for v in range(V)])
    with PgenContext(path) as c:
        got_c, got_ad, got_r2 = c.dosage_counts()
        got_d = c.dosage_f64(np.arange(V))
    assert np.array_equal(got_c, exp_c) and np.array_equal(got_ad, exp_ad)
    assert np.array_equal(got_d, exp_d)
    assert np.array_equal(np.isnan(got_r2), np.isnan(exp_r2))
    ok = ~np.isnan(exp_r2)
    assert np.array_equal(got_r2[ok], exp_r2[ok])
