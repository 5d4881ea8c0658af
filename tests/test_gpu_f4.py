"""records with multiallelic patch sets (vrtype bit 3) and dosage-phase tracks (bit 7), written by
pgenlib's own writer (tests/f4.py).  The extension reads them through the biallelic entry points, so the bar is
what PgrGetCounts / PgrGet / PgrGetMissingness / PgrGetP / PgrGetD / PgrGetDCounts return on the same file."""
import numpy as np
import pytest

import f4

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def refmod_gpu():
    from oracle import ref

    if not ref.available():
        pytest.skip("oracle/_ref not built: no pgenlib writer")
    return ref


@pytest.mark.parametrize("kw", [dict(), dict(tile_bytes=2048, n_staging=2), dict(devices=[0]), dict(devices=[0, 1])])
def test_multiallelic_records_read_as_pgenlib_reads_them(tmp_path, refmod_gpu, kw):
    from conftest import need_devices

    need_devices(kw)
    from plinking_duck_b200 import _native as nat
    from plinking_duck_b200.pgen import PgenContext

    ref = refmod_gpu
    V, N = 400, 1501
    path = str(tmp_path / "ma.pgen")
    codes, allele_cts = f4.write_multiallelic(path, V, N, seed=8)
    with ref.RefPgen(path) as r:
        vt = r.vrtypes()
        exp_counts = r.counts()
        exp_i8 = r.decode_i8()
        _, exp_ms = r.missingness()
        eh, ea, em = r.sample_counts(sparse=True)
        exp_d = np.stack([r.dosage(v) for v in range(0, V, 7)])
    assert (vt & 8).any() and not (vt & 8).all()
    with PgenContext(path, **kw) as c:
        assert c.info().has_multiallelic == 1
        for resident in (False, True):
            if resident:
                c.load_resident()
            assert np.array_equal(c.counts(), exp_counts)
            assert np.array_equal(c.decode_i8(), exp_i8)
            assert np.array_equal(c.missing_by_sample(), exp_ms)
            h, a, m = c.sample_counts()
            assert np.array_equal(h, eh) and np.array_equal(a, ea) and np.array_equal(m, em)
            counts, p = c.hardy()
            assert np.array_equal(counts, exp_counts) and np.all((p >= 0) & (p <= 1))
        # PgrGetD on a multiallelic record without a dosage track: the hardcalls as doubles
        assert np.array_equal(c.dosage_f64(np.arange(0, V, 7, dtype=np.uint32)), exp_d)
        sub = np.arange(2, N, 3)
        c.set_subset(sub)
        assert np.array_equal(c.decode_i8(), exp_i8[:, sub])
        assert isinstance(nat.PD_ERR_NOT_SUPPORTED, int)


def _aux1_modes(path, vrtypes, fpos, N):
    """(aux1a mode, aux1b mode) of every multiallelic record whose main track is a plain 2-bit array (its patch track
    then starts ceil(N / 4) bytes into the record; pgen_spec.tex:469-479)."""
    raw = np.fromfile(path, dtype=np.uint8)
    out = []
    for v in np.nonzero(((vrtypes & 8) != 0) & ((vrtypes & 7) == 0))[0]:
        b = int(raw[int(fpos[v]) + (N + 3) // 4])
        out.append((b & 15, b >> 4))
    return out


@pytest.mark.parametrize("kw", [dict(), dict(tile_bytes=4096, n_staging=2), dict(devices=[0]), dict(devices=[0, 1])])
@pytest.mark.parametrize("N", [37, 1501, 20011])
def test_decode_alleles_matches_pgrgetm(tmp_path, refmod_gpu, kw, N):
    """pd_decode_alleles = PgrGetM + PglMultiallelicSparseToDenseMiss (pgenlib_read.cc:6668-6681,
    pgenlib_misc.cc:3552-3578): allele counts 2..40 (every aux1a / aux1b value width), patch sets stored as bit arrays
    and as sample-id lists, every main-track type the writer picks, streamed / tiny tiles / resident / a group context,
    sub-ranges and a sample subset."""
    from conftest import need_devices

    need_devices(kw)
    from plinking_duck_b200 import _native as nat
    from plinking_duck_b200.pgen import PgenContext

    ref = refmod_gpu
    V = 240 if N < 20000 else 90
    path = str(tmp_path / "ma.pgen")
    codes, allele_cts = f4.write_multiallelic(path, V, N, seed=21 + N, rare_every=2)
    with ref.RefPgen(path, allele_cts=allele_cts) as r:
        exp = r.decode_alleles()
        sub = np.arange(1, N, 3, dtype=np.uint32)
        r.set_subset(sub)
        exp_sub = r.decode_alleles()
    # the biallelic view of the codes is the main track that went in
    lo, hi = exp[..., 0].astype(int), exp[..., 1].astype(int)
    assert np.array_equal(np.where(lo == 255, 3, np.where(hi == 0, 0, np.where(lo == 0, 1, 2))), codes)
    with PgenContext(path, **kw) as c:
        vt, fpos = c.index()
        if N >= 1501:
            modes = _aux1_modes(path, vt, fpos, N)
            assert any(a == 0 for a, _ in modes) and any(b == 0 for _, b in modes), "no bit-array patch set in the fixture"
            assert any(a == 1 for a, _ in modes) and any(b == 1 for _, b in modes), "no id-list patch set in the fixture"
        with pytest.raises(nat.PlinkGpuError) as ei:
            c.decode_alleles()
        assert "pd_set_allele_counts" in str(ei.value)
        c.set_allele_counts(allele_cts)
        for resident in (False, True):
            if resident:
                c.load_resident()
            got = c.decode_alleles()
            bad = np.nonzero((got != exp).any(axis=(1, 2)))[0]
            assert bad.size == 0, f"variants {bad[:8]} differ (allele counts {allele_cts[bad[:8]]}, vrtypes {vt[bad[:8]]})"
            assert np.array_equal(c.decode_alleles(V // 3, V // 3 + 17), exp[V // 3:V // 3 + 17])
        c.set_subset(sub)
        assert np.array_equal(c.decode_alleles(), exp_sub)
        c.set_subset(None)
        # the biallelic entry points are unaffected by the table
        assert np.array_equal(c.decode_i8(), np.where(codes == 3, -9, codes).astype(np.int8))
        with pytest.raises(nat.PlinkGpuError):
            c.set_allele_counts(allele_cts[:-1])


@pytest.mark.parametrize("kw", [dict(), dict(tile_bytes=4096, n_staging=2), dict(devices=[0]), dict(devices=[0, 1])])
@pytest.mark.parametrize("N", [37, 1501, 20011])
def test_phased_decode_of_multiallelic_records(tmp_path, refmod_gpu, kw, N):
    """Records with BOTH a multiallelic patch track and a hardcall-phase track: aux2 sits behind aux1 and counts the
    ALTx/ALTy (x != y) calls of aux1b among its hets (ReadGenovecHphaseSubsetUnsafe, pgenlib_read.cc:6874-6962).
    pd_decode_phased must return what PgrGetP + UnpackPhasedGenotypes return once pgenlib knows the allele counts --
    and, like pgenlib, cannot read such records without them."""
    from conftest import need_devices

    need_devices(kw)
    from plinking_duck_b200 import _native as nat
    from plinking_duck_b200.pgen import PgenContext

    ref = refmod_gpu
    V = 240 if N < 20000 else 90
    path = str(tmp_path / "map.pgen")
    pairs, pps, pis, allele_cts = f4.write_multiallelic_phased(path, V, N, seed=3 + N)
    with ref.RefPgen(path, allele_cts=allele_cts) as r:
        vt = r.vrtypes()
        exp = r.decode_phased()
        exp_al = r.decode_alleles()
        sub = np.arange(1, N, 3, dtype=np.uint32)
        r.set_subset(sub)
        exp_sub = r.decode_phased()
    assert ((vt & 0x18) == 0x18).any() and ((vt & 0x18) == 0x10).any() and ((vt & 0x18) == 0x08).any()
    assert np.array_equal(exp_al, pairs)
    with PgenContext(path, **kw) as c:
        with pytest.raises(nat.PlinkGpuError) as ei:
            c.decode_phased()
        assert ei.value.code == nat.PD_ERR_NOT_SUPPORTED
        c.set_allele_counts(allele_cts)
        for resident in (False, True):
            if resident:
                c.load_resident()
            got = c.decode_phased()
            bad = np.nonzero((got != exp).any(axis=(1, 2)))[0]
            assert bad.size == 0, f"variants {bad[:8]} differ (allele counts {allele_cts[bad[:8]]}, vrtypes {vt[bad[:8]]})"
            assert np.array_equal(c.decode_phased(V // 3, V // 3 + 17), exp[V // 3:V // 3 + 17])
            assert np.array_equal(c.decode_alleles(), pairs)
        c.set_subset(sub)
        assert np.array_equal(c.decode_phased(), exp_sub)


def test_decode_alleles_on_biallelic_files(golden):
    """Files without patch tracks need no allele counts: 0,0 / 0,1 / 1,1 / 255,255 from the main track."""
    from conftest import fixture_path
    from plinking_duck_b200.pgen import PgenContext

    for name in ("pgen_example", "streaming_example", "syn_n1000"):
        with PgenContext(fixture_path(name)) as c:
            got = c.decode_alleles()
        g = golden[name + "/i8"].astype(int)
        exp = np.stack([np.where(g == -9, 255, np.where(g == 2, 1, 0)), np.where(g == -9, 255, np.where(g >= 1, 1, 0))], axis=-1)
        assert np.array_equal(got, exp.astype(np.uint8)), name


def test_dosage_phase_tracks_are_skipped_like_pgenlib(tmp_path, refmod_gpu):
    """vrtype bit 7: PgrGetD / PgrGetDCounts ignore the dphase values behind the dosage track; the hardcall phase
    track in front of it is read by PgrGetP."""
    from plinking_duck_b200.pgen import PgenContext

    ref = refmod_gpu
    V, N = 180, 777
    path = str(tmp_path / "dp.pgen")
    f4.write_dphase(path, V, N, seed=9)
    with ref.RefPgen(path) as r:
        vt = r.vrtypes()
        exp_counts, exp_ad, exp_r2 = r.dcounts()
        exp_d = np.stack([r.dosage(v) for v in range(V)])
        exp_i8 = r.decode_i8()
        exp_ph = r.decode_phased()
        rng = np.random.default_rng(1)
        w = rng.standard_normal(V)
        fl = (rng.random(V) < 0.3).astype(np.uint8)
        es, ed, ea = r.score(np.arange(V, dtype=np.uint32), fl, w, 0)
    assert (vt & 0x80).any() and (vt & 0x60).any() and (vt & 0x10).any()
    for kw in (dict(), dict(tile_bytes=4096, n_staging=2)):
        with PgenContext(path, **kw) as c:
            assert np.array_equal(c.dosage_f64(np.arange(V, dtype=np.uint32)), exp_d)
            counts, ad, r2 = c.dosage_counts()
            assert np.array_equal(counts, exp_counts) and np.array_equal(ad, exp_ad)
            assert np.array_equal(r2, exp_r2, equal_nan=True)
            assert np.array_equal(c.decode_i8(), exp_i8)
            assert np.array_equal(c.decode_phased(), exp_ph)
            s, d, a = c.score(np.arange(V, dtype=np.uint32), fl, w, 0)
            assert np.array_equal(a, ea)
            assert np.allclose(s[:, 0], es, rtol=1e-12, atol=1e-9) and np.allclose(d, ed, rtol=1e-12, atol=1e-9)
