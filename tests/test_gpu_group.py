"""GPU parity of multi-GPU contexts (pd_open_opts.n_devices): ONE pd_ctx in ONE process over k GPUs
-- variant ranges of equal record bytes per GPU, per-variant outputs in disjoint slices, per-sample accumulators
summed with ncclAllReduce inside the library -- must return exactly what the reference's pgenlib returns
(tests/golden/golden.npz), as the single-GPU context does.  k = 1 exercises the partition / dispatch plumbing on
a one-GPU box; k = 2 and 8 run where that many devices are visible (gpurun --gpus N)."""
import numpy as np
import pytest

from conftest import fixture_names, fixture_path

pytestmark = pytest.mark.gpu


def _devices(k):
    import torch

    if torch.cuda.device_count() < k:
        pytest.skip(f"needs {k} GPUs")
    return list(range(k))


def _ctx(name, k, **kw):
    from plinking_duck_b200.pgen import PgenContext

    return PgenContext(fixture_path(name), devices=_devices(k), **kw)


K = [1, 2, 8]


@pytest.mark.parametrize("k", K)
@pytest.mark.parametrize("name", fixture_names())
def test_group_counts_hardy_bit_exact(name, k, golden):
    import hwe_check as hc

    exp = golden[name + "/counts"]
    V = exp.shape[0]
    for kw in (dict(), dict(tile_bytes=1024, n_staging=2)):
        with _ctx(name, k, **kw) as c:
            info = c.info()
            assert info.n_devices == k and info.variant_ct == V
            assert np.array_equal(c.counts(), exp)
            a, b = V // 3, V - V // 4
            assert np.array_equal(c.counts(a, b), exp[a:b])
            counts, p = c.hardy()
            assert np.array_equal(counts, exp)
            v = hc.violations_p_vs_reference(p, golden[name + "/hwe_lnp"], exp[:, :3].sum(axis=1))
            assert v.size == 0, hc.describe(v, exp, p, golden[name + "/hwe_lnp"])
            c.load_resident()
            assert c.info().resident == 1
            assert np.array_equal(c.counts(), exp)
            assert np.array_equal(c.counts(a, b), exp[a:b])
            st = c.stats()
            assert st.launches > 0 and st.n_devices == k
            job = c.counts_async(0, V)
            for r0 in range(0, V, 128):
                r1 = min(V, r0 + 128)
                assert np.array_equal(job.fetch(r0, r1), exp[r0:r1])
            job.close()
            if name + "/subset" in golden:
                c.set_subset(golden[name + "/subset"])
                assert np.array_equal(c.counts(), golden[name + "/subset_counts"])
                c.set_subset(None)


@pytest.mark.parametrize("k", K)
@pytest.mark.parametrize("name", fixture_names())
def test_group_per_sample_accumulators_all_reduced(name, k, golden):
    het, hom_alt, miss = golden[name + "/sample_counts"]
    with _ctx(name, k, tile_bytes=4096) as c:
        assert np.array_equal(c.missing_by_sample(), golden[name + "/missing_by_sample"])
        h, a, m = c.sample_counts()
        assert np.array_equal(h, het) and np.array_equal(a, hom_alt) and np.array_equal(m, miss)
        st = c.stats()
        if k > 1:
            assert st.allreduce_bytes == 12 * het.size
        V = c.variant_ct
        lo, hi = V // 4, V - V // 5
        m2 = c.missing_by_sample(lo, hi)
        if name + "/i8" in golden:
            assert np.array_equal(m2, (golden[name + "/i8"][lo:hi] == -9).sum(axis=0).astype(np.uint32))


@pytest.mark.parametrize("k", K)
@pytest.mark.parametrize("name", fixture_names())
def test_group_decode_bit_exact(name, k, golden):
    if name + "/i8" not in golden:
        pytest.skip("no int8 golden for this fixture")
    exp = golden[name + "/i8"]
    V, N = exp.shape
    with _ctx(name, k, tile_bytes=2048) as c:
        assert np.array_equal(c.decode_i8(), exp)
        a, b = V // 5, V - V // 3
        assert np.array_equal(c.decode_i8(a, b), exp[a:b])
        data, valid = c.decode_i8_valid(0, V, call_mask=7)
        bits = np.unpackbits(valid.view(np.uint8), bitorder="little")[: V * N].reshape(V, N).astype(bool)
        assert np.array_equal(bits, exp != -9)
        assert np.array_equal(data, np.where(exp == -9, 0, exp))


@pytest.mark.parametrize("k", K)
@pytest.mark.parametrize("name", ["pgen_example", "pca_example", "syn_n1000", "streaming_example", "syn_dosage"])
def test_group_score_all_reduced(name, k, golden):
    """plink_score across GPUs: the scored-variant list is cut by record bytes, partial sums all-reduced."""
    from plinking_duck_b200.pgen import PgenContext

    with PgenContext(fixture_path(name)) as c1:
        V = c1.variant_ct
        rng = np.random.default_rng(5)
        vidx = np.sort(rng.choice(V, size=max(1, (2 * V) // 3), replace=False)).astype(np.uint32)
        W = rng.standard_normal((vidx.size, 3))
        flip = (rng.random(vidx.size) < 0.3).astype(np.uint8)
        for mode in (0, 1, 2):
            s1, d1, a1 = c1.score(vidx, flip, W, mode)
            with _ctx(name, k) as c:
                s, d, a = c.score(vidx, flip, W, mode)
                if k > 1:
                    assert c.stats().allreduce_bytes == s.size * 8 + d.size * 8 + a.size * 4
            assert np.array_equal(a, a1)
            scale = 1e-12 * (2.0 * np.abs(W).sum(axis=0) + 1.0)
            assert np.all(np.abs(s - s1) <= scale), (name, mode)
            assert np.allclose(d, d1, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("k", K)
def test_group_dosage_and_ops(k, golden):
    from plinking_duck_b200 import ops
    from plinking_duck_b200.pgen import PgenContext

    name = "syn_dosage"
    with _ctx(name, k) as c:
        V = c.variant_ct
        counts, ad, r2 = c.dosage_counts()
        assert np.array_equal(counts, golden[name + "/dcounts"])
        assert np.array_equal(ad, golden[name + "/all_dosages"])
        assert np.array_equal(r2, golden[name + "/imp_r2"], equal_nan=True)
        with PgenContext(fixture_path(name)) as c1:
            d1 = c1.dosage_f64(np.arange(V, dtype=np.uint32))
            rows1 = ops.plink_freq(c1, dosage=True)
        assert np.array_equal(c.dosage_f64(np.arange(V, dtype=np.uint32)), d1)
        gd = golden[name + "/dosage_first"]
        assert np.array_equal(d1[: gd.shape[0]], gd)
        # the operator layer only sees the C ABI: a multi-GPU context is a drop-in
        rows = ops.plink_freq(c, dosage=True)
        assert rows.tobytes() == rows1.tobytes()
        m1 = ops.plink_missing(c, mode="sample")
        assert np.array_equal(m1["MISSING_CT"], golden[name + "/missing_by_sample"])


@pytest.mark.parametrize("k", K)
def test_group_over_pinned_file(k, tmp_path, golden):
    """A multi-GPU context over a PD_OPEN_PIN_FILE mapping: registered once, shared by every shard."""
    import shutil

    from plinking_duck_b200.pgen import PgenContext

    import os
    import tempfile

    name = "streaming_example"
    shm = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    with tempfile.TemporaryDirectory(dir=shm or str(tmp_path)) as d:
        dst = os.path.join(d, name + ".pgen")
        shutil.copyfile(fixture_path(name), dst)
        with PgenContext(dst, devices=_devices(k), pin_file=True, tile_bytes=4096) as c:
            info = c.info()
            assert (info.pinned_file == 1 or shm is None) and info.n_devices == k
            assert np.array_equal(c.counts(), golden[name + "/counts"])
            assert np.array_equal(c.missing_by_sample(), golden[name + "/missing_by_sample"])


@pytest.mark.parametrize("k", K)
def test_group_over_pinned_file_crossing_registration_pieces(k, tmp_path, monkeypatch):
    """The page-locked mapping is registered in pieces and one DMA may not span two of them: with 2 MiB pieces
    (PD_PIN_PIECE_MB) a 9 MB file is five registrations, tiles of 1.5 MiB straddle them, and the shards -- which
    see the parent's mapping as a plain memory image -- must cut their copies at the same grid (the N = 2 bench
    failed with 'invalid argument' before they did)."""
    import os
    import tempfile

    import synth
    from plinking_duck_b200.pgen import PgenContext

    shm = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    if shm is None:
        pytest.skip("needs a tmpfs to page-lock a file mapping")
    monkeypatch.setenv("PD_PIN_PIECE_MB", "2")
    g = synth.gen_hwe_genotypes(3000, 12011, seed=77)
    exp = synth.counts_numpy(g)
    with tempfile.TemporaryDirectory(dir=shm) as d:
        path = os.path.join(d, "pieces.pgen")
        synth.write_fixed_pgen(path, g)
        for devices in (None, _devices(k)):
            with PgenContext(path, devices=devices, pin_file=True, tile_bytes=3 << 19) as c:
                assert c.info().pinned_file == 1
                assert np.array_equal(c.counts(), exp)
                counts, p = c.hardy()
                assert np.array_equal(counts, exp)
                assert np.array_equal(c.missing_by_sample(), (g == 3).sum(0).astype(np.uint32))
                c.load_resident()
                assert np.array_equal(c.counts(), exp)


def test_group_rejects_bad_device_lists():
    from plinking_duck_b200 import _native as nat
    from plinking_duck_b200.pgen import PgenContext

    with pytest.raises(nat.PlinkGpuError) as e:
        PgenContext(fixture_path("pgen_example"), devices=[0, 0])
    assert e.value.code == nat.PD_ERR_INVALID_ARG
    with pytest.raises(nat.PlinkGpuError) as e:
        PgenContext(fixture_path("pgen_example"), devices=[99])
    assert e.value.code == nat.PD_ERR_INVALID_ARG
