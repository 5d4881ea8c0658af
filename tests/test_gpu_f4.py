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


@pytest.mark.parametrize("kw", [dict(), dict(tile_bytes=2048, n_staging=2), dict(devices=[0])])
def test_multiallelic_records_read_as_pgenlib_reads_them(tmp_path, refmod_gpu, kw):
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
