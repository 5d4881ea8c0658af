"""GPU parity: sample-wise consumers of decoded genotype rows through the C ABI --
pd_decode_i8 (PgrGet + GenoarrToBytesMinus9), pd_missing_by_sample (plink_missing sample mode),
pd_sample_counts (read_pfile orient := 'sample') -- vs pgenlib goldens, bit-exact."""
import numpy as np
import pytest

from conftest import fixture_names, fixture_path

pytestmark = pytest.mark.gpu


def _ctx(name, **kw):
    from plinking_duck_b200.pgen import PgenContext

    return PgenContext(fixture_path(name), **kw)


@pytest.mark.parametrize("name", fixture_names())
@pytest.mark.parametrize("mode", ["stream", "stream_small_tiles", "resident"])
def test_decode_i8_bit_exact(name, mode, golden):
    key = name + "/i8"
    if key not in golden:
        pytest.skip("no int8 golden (matrix too large)")
    kw = dict(tile_bytes=700, n_staging=2) if mode == "stream_small_tiles" else {}
    with _ctx(name, **kw) as c:
        if mode == "resident":
            c.load_resident()
        got = c.decode_i8()
        assert c.stats().launches > 0
    assert np.array_equal(got, golden[key])


@pytest.mark.parametrize("name", fixture_names())
def test_decode_i8_subranges_and_subset(name, golden):
    key = name + "/i8"
    if key not in golden:
        pytest.skip("no int8 golden")
    exp = golden[key]
    V = exp.shape[0]
    rng = np.random.default_rng(9)
    with _ctx(name, tile_bytes=4096) as c:
        for _ in range(4):
            a = int(rng.integers(0, V))
            b = int(rng.integers(a, V + 1))
            assert np.array_equal(c.decode_i8(a, b), exp[a:b]), (a, b)
        if name + "/subset_i8" in golden:
            c.set_subset(golden[name + "/subset"])
            assert np.array_equal(c.decode_i8(), golden[name + "/subset_i8"])
            c.set_subset(None)
            assert np.array_equal(c.decode_i8(0, min(V, 7)), exp[: min(V, 7)])


def test_decode_i8_large_n_against_counts(golden):
    """syn_n70000 has no int8 golden: check the decoded matrix through its per-variant counts."""
    name = "syn_n70000"
    with _ctx(name) as c:
        g = c.decode_i8()
    exp = golden[name + "/counts"]
    got = np.stack([(g == 0).sum(1), (g == 1).sum(1), (g == 2).sum(1), (g == -9).sum(1)], axis=1)
    assert np.array_equal(got.astype(np.uint32), exp)


@pytest.mark.parametrize("name", fixture_names())
@pytest.mark.parametrize("mode", ["stream", "stream_small_tiles", "resident"])
def test_missing_by_sample_bit_exact(name, mode, golden):
    kw = dict(tile_bytes=900, n_staging=2) if mode == "stream_small_tiles" else {}
    with _ctx(name, **kw) as c:
        if mode == "resident":
            c.load_resident()
        got = c.missing_by_sample()
    assert np.array_equal(got, golden[name + "/missing_by_sample"])


@pytest.mark.parametrize("name", fixture_names())
def test_missing_by_sample_subset_and_range(name, golden):
    with _ctx(name) as c:
        if name + "/subset" in golden:
            c.set_subset(golden[name + "/subset"])
            assert np.array_equal(c.missing_by_sample(), golden[name + "/subset_missing_by_sample"])
            c.set_subset(None)
        if name + "/i8" in golden:
            i8 = golden[name + "/i8"]
            V = i8.shape[0]
            a, b = V // 3, max(V // 3, (2 * V) // 3)
            assert np.array_equal(c.missing_by_sample(a, b), (i8[a:b] == -9).sum(0).astype(np.uint32))
        # variant-mode and sample-mode totals agree (test/sql/streaming_threading.test:165-171)
        assert int(c.missing_by_sample().sum()) == int(c.counts()[:, 3].sum())


@pytest.mark.parametrize("name", fixture_names())
@pytest.mark.parametrize("mode", ["stream", "resident"])
def test_sample_counts_bit_exact(name, mode, golden):
    with _ctx(name) as c:
        if mode == "resident":
            c.load_resident()
        het, hom_alt, missing = c.sample_counts()
    exp = golden[name + "/sample_counts"]
    assert np.array_equal(het, exp[0])
    assert np.array_equal(hom_alt, exp[1])
    assert np.array_equal(missing, exp[2])


def test_rare_small_sparse_golden_sums():
    """test/sql/read_pfile_sample_counts_sparse.test:13 -- sums 99217 / 1947 / 224 / 1012."""
    with _ctx("rare_small") as c:
        het, hom_alt, missing = c.sample_counts()
        V, N = c.variant_ct, c.raw_sample_ct
    assert (int(het.sum()), int(hom_alt.sum()), int(missing.sum())) == (1947, 224, 1012)
    assert V * N - 1947 - 224 - 1012 == 99217


def test_fixed_width_sample_paths(tmp_path):
    import synth
    from plinking_duck_b200.pgen import PgenContext

    for (V, N) in [(700, 1), (300, 37), (1000, 1000), (2100, 4099), (40, 100003)]:
        g = synth.gen_hwe_genotypes(V, N, seed=3 * V + N, missing=0.05)
        img = synth.fixed_pgen_image(g)
        exp_i8 = np.where(g == 3, -9, g).astype(np.int8)
        with PgenContext(image=img, tile_bytes=1 << 16) as c:
            for resident in (False, True):
                if resident:
                    c.load_resident()
                assert np.array_equal(c.decode_i8(), exp_i8), (V, N, resident)
                het, hom, mis = c.sample_counts()
                assert np.array_equal(het, (g == 1).sum(0).astype(np.uint32)), (V, N, resident)
                assert np.array_equal(hom, (g == 2).sum(0).astype(np.uint32)), (V, N, resident)
                assert np.array_equal(mis, (g == 3).sum(0).astype(np.uint32)), (V, N, resident)
                assert np.array_equal(c.missing_by_sample(), mis)
            sub = np.arange(0, N, 3)
            c.set_subset(sub)
            assert np.array_equal(c.decode_i8(5, min(V, 50)), exp_i8[5:min(V, 50)][:, sub])
            assert np.array_equal(c.missing_by_sample(), (g[:, sub] == 3).sum(0).astype(np.uint32))


def test_sparse_sample_accumulation_on_pgenlib_written_records(tmp_path):
    """Difflist records (vrtype 4/5/6/7) are accumulated per sample without being expanded
    (pdk::SparseAcc, the GPU form of src/pfile_reader.cpp:3322-3428's sparse branch): a file written by
    pgenlib's own writer with all-ref (type 5), all-hom-alt (type 6, empty difflist), all-missing (type 7),
    rare-variant and common rows interleaved, so every batch mixes sparse and decoded rows."""
    import synth
    from oracle import ref
    from plinking_duck_b200.pgen import PgenContext

    if not ref.available():
        pytest.skip("oracle/_ref not built: no pgenlib writer")
    for (V, N) in [(900, 300), (260, 70001)]:
        g = synth.gen_rare_genotypes(V, N, seed=V + N)
        common = synth.gen_hwe_genotypes(V, N, seed=V, missing=0.03)
        pick = np.arange(V) % 7 == 3
        g[pick] = common[pick]
        g[0] = 0
        g[1] = 2
        g[2] = 3
        g[3] = 0
        g[3, N // 2] = 1
        g[V - 1] = 0
        g[V - 2] = 2
        g[V - 2, ::5] = 3
        g[10:14] = 0
        path = str(tmp_path / f"sparse_{V}_{N}.pgen")
        ref.write_pgen(path + ".tmp", synth.to_genovec_words(g), N)
        assert synth.rewrite_with_all_ref_records(path + ".tmp", path) >= 3
        with ref.RefPgen(path) as r:
            vt = set(int(x) & 7 for x in r.vrtypes())
            assert np.array_equal(r.counts(), synth.counts_numpy(g))
        assert {4, 5, 6, 7} <= vt, vt
        exp = [(g == k).sum(0).astype(np.uint32) for k in (1, 2, 3)]
        for kw in (dict(), dict(tile_bytes=3000, n_staging=2)):
            with PgenContext(path, **kw) as c:
                for resident in (False, True):
                    if resident:
                        c.load_resident()
                    het, hom, mis = c.sample_counts()
                    assert np.array_equal(het, exp[0]) and np.array_equal(hom, exp[1]) and np.array_equal(mis, exp[2])
                    assert np.array_equal(c.missing_by_sample(), exp[2])
                    assert np.array_equal(c.counts(), synth.counts_numpy(g))
                    a, b = V // 3, V - 1
                    assert np.array_equal(c.missing_by_sample(a, b), (g[a:b] == 3).sum(0).astype(np.uint32))
                sub = np.arange(1, N, 3)
                c.set_subset(sub)
                het, hom, mis = c.sample_counts()
                assert np.array_equal(het, exp[0][sub]) and np.array_equal(hom, exp[1][sub])
                assert np.array_equal(mis, exp[2][sub])
                assert np.array_equal(c.missing_by_sample(), exp[2][sub])


def test_decode_phased_known_answers():
    """test/sql/read_pgen_phased.test:45-63."""
    import kat

    for kw in (dict(), dict(tile_bytes=64, n_staging=2)):
        with _ctx("phased_example", **kw) as c:
            assert c.decode_phased().tolist() == kat.PHASED_EXAMPLE_PAIRS
            c.load_resident()
            assert c.decode_phased(1, 3).tolist() == kat.PHASED_EXAMPLE_PAIRS[1:3]
            c.set_subset(np.array([3, 1]))
            exp = np.asarray(kat.PHASED_EXAMPLE_PAIRS, dtype=np.int8)[:, [1, 3]]
            assert np.array_equal(c.decode_phased(), exp)


def test_decode_phased_vs_pgenlib(tmp_path):
    """pd_decode_phased (PgrGetP + UnpackPhasedGenotypes) on pgenlib-written files: implicit and explicit
    aux2 tracks behind every main record type (vrtypes 0x10-0x17), records without a phase track, ragged N,
    streamed / tiny tiles / resident, sub-range, sample subset; and a fixed-width file (no phase: hets 0,1)."""
    import synth
    from oracle import ref
    from plinking_duck_b200.pgen import PgenContext

    if not ref.available():
        pytest.skip("oracle/_ref not built: no pgenlib writer")
    for (V, N) in [(90, 300), (60, 257), (40, 5000), (12, 70001)]:
        g = synth.gen_mixed_genotypes(V, N, seed=V)
        phase, exp = synth.gen_phase(g, 7)
        path = str(tmp_path / f"ph_{V}_{N}.pgen")
        ref.write_pgen(path, synth.to_genovec_words(g), N, phase=phase)
        with ref.RefPgen(path) as r:
            assert np.array_equal(r.decode_phased(), exp)
        for kw in (dict(), dict(tile_bytes=900, n_staging=2)):
            with PgenContext(path, **kw) as c:
                assert np.array_equal(c.decode_phased(), exp), (V, N, kw)
                a, b = V // 4, V - 3
                assert np.array_equal(c.decode_phased(a, b), exp[a:b]), (V, N, kw)
                assert np.array_equal(c.decode_i8(), np.where(g == 3, -9, g))
                c.load_resident()
                assert np.array_equal(c.decode_phased(), exp), (V, N, kw, "resident")
                sub = np.arange(1, N, 3)
                c.set_subset(sub)
                assert np.array_equal(c.decode_phased(), exp[:, sub]), (V, N, kw, "subset")
    g = synth.gen_hwe_genotypes(50, 1003, seed=2)
    exp = np.zeros((50, 1003, 2), dtype=np.int8)
    exp[g == 1] = (0, 1)
    exp[g == 2] = (1, 1)
    exp[g == 3] = (-9, -9)
    with PgenContext(image=synth.fixed_pgen_image(g)) as c:
        assert np.array_equal(c.decode_phased(), exp)


@pytest.mark.parametrize("name", fixture_names())
def test_decode_i8_valid_child_vector_layout(name, golden):
    """pd_decode_i8_valid: DuckDB child data + validity (src/pgen_reader.cpp:1008-1046) for every genotype
    filter mask, tiny tiles (batches start at any bit of the bitmap), sub-ranges and a sample subset."""
    from oracle import port

    key = name + "/i8"
    if key not in golden:
        pytest.skip("no int8 golden")
    exp = golden[key]
    V = exp.shape[0]
    rng = np.random.default_rng(11)
    with _ctx(name, tile_bytes=900, n_staging=2) as c:
        for mask in (7, 0, 1, 2, 4, 6, 3, 5):
            allowed = tuple(bool(mask >> k & 1) for k in range(3))
            d, v = c.decode_i8_valid(call_mask=mask)
            ed, ev = port.child_vector_layout(exp, allowed)
            assert np.array_equal(d, ed), mask
            assert np.array_equal(v, ev), mask
        for _ in range(3):
            a = int(rng.integers(0, V))
            b = int(rng.integers(a, V + 1))
            d, v = c.decode_i8_valid(a, b, call_mask=6)
            ed, ev = port.child_vector_layout(exp[a:b], (False, True, True))
            assert np.array_equal(d, ed) and np.array_equal(v, ev), (a, b)
        if name + "/subset_i8" in golden:
            c.set_subset(golden[name + "/subset"])
            d, v = c.decode_i8_valid(call_mask=5)
            ed, ev = port.child_vector_layout(golden[name + "/subset_i8"], (True, False, True))
            assert np.array_equal(d, ed) and np.array_equal(v, ev)


def test_decode_i8_valid_many_batches(tmp_path):
    """Rows of 70001 samples (odd: every row starts at a new bit offset) over enough variants that the
    128 MiB output batches and the tiles both split the call; resident and streamed agree with pgenlib."""
    import synth
    from oracle import port, ref

    if not ref.available():
        pytest.skip("oracle/_ref not built")
    V, N = 4200, 70001
    g = synth.gen_mixed_genotypes(V, N, seed=5)
    path = str(tmp_path / "big.pgen")
    ref.write_pgen(path, synth.to_genovec_words(g), N)
    with ref.RefPgen(path) as r:
        exp = r.decode_i8()
    ed, ev = port.child_vector_layout(exp, (True, True, False))
    from plinking_duck_b200.pgen import PgenContext

    with PgenContext(path) as c:
        d, v = c.decode_i8_valid(call_mask=3)
        assert np.array_equal(d, ed) and np.array_equal(v, ev)
        c.load_resident()
        d, v = c.decode_i8_valid(call_mask=3)
        assert np.array_equal(d, ed) and np.array_equal(v, ev)
        # page-locked destinations: both streams are DMA-ed straight into the caller's arrays
        import ctypes as C

        import torch

        pd_ = torch.full((V, N), 0x55, dtype=torch.int8).pin_memory().numpy()
        pv = torch.full((ev.size,), -1, dtype=torch.int64).pin_memory().numpy().view(np.uint64)
        c._check(c._lib.pd_decode_i8_valid(c._h, 0, V, 3, pd_.ctypes.data_as(C.c_void_p), pv.ctypes.data_as(C.c_void_p)))
        assert np.array_equal(pd_, ed) and np.array_equal(pv, ev)


def test_c4_compressed_fileset_at_scale(tmp_path):
    """BASELINE.json configs[3] shape at test scale (24 000 variants x 50 000 samples, tests/c4.py): a pgenlib-written
    file that really holds LD-compressed (2 / 3), difflist (4 / 6), one-bit and raw records; genotype counts,
    per-sample missingness and per-sample het / hom-alt / missing counts bit-exact against the reference's own
    pgenlib on every variant and sample -- streamed in default and tiny tiles, and resident."""
    import os
    import tempfile

    import torch

    import c4
    from oracle import ref
    from plinking_duck_b200.pgen import PgenContext

    if not ref.available():
        pytest.skip("oracle/_ref not built: no pgenlib writer")
    V, N = 24_000, 50_000
    shm = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    with tempfile.TemporaryDirectory(dir=shm or str(tmp_path)) as d:
        path = os.path.join(d, "c4.pgen")
        c4.make_c4_file(path, V, N, 777, torch.device("cuda", 0))
        hist = c4.vrtype_histogram(path)
        assert hist.get(2, 0) > 500 and hist.get(3, 0) > 100, hist   # LD-compressed records at scale
        assert hist.get(4, 0) > 5000 and hist.get(6, 0) > 500 and hist.get(1, 0) > 500, hist
        assert os.path.getsize(path) < 0.35 * V * ((N + 3) // 4)
        with ref.RefPgen(path) as r:
            exp_counts = r.counts()
            _, exp_ms = r.missingness()
            eh, ea, em = r.sample_counts(sparse=True)
            a, b = 4000, 4200
            exp_i8 = r.decode_i8(a, b)
        for kw in (dict(), dict(tile_bytes=1 << 16, n_staging=2)):
            with PgenContext(path, **kw) as c:
                for resident in (False, True):
                    if resident:
                        c.load_resident()
                    assert np.array_equal(c.counts(), exp_counts)
                    assert np.array_equal(c.missing_by_sample(), exp_ms)
                    h, al, m = c.sample_counts()
                    assert np.array_equal(h, eh) and np.array_equal(al, ea) and np.array_equal(m, em)
                assert np.array_equal(c.decode_i8(a, b), exp_i8)
                lo, hi = 7001, 19003   # a range that starts and ends inside LD blocks
                assert np.array_equal(c.counts(lo, hi), exp_counts[lo:hi])
                counts, p = c.hardy(lo, hi)
                assert np.array_equal(counts, exp_counts[lo:hi]) and np.all((p >= 0) & (p <= 1))


def _tidy_expected(i8, sample_ids, emit_mask, v0=0):
    """The rows PfileTidyScan emits (src/pfile_reader.cpp:3043-3075 filter): variant-major, ascending sample."""
    cat = np.where(i8 == -9, 3, i8).astype(np.int64)
    keep = ((emit_mask >> cat) & 1).astype(bool)
    vv, ss = np.nonzero(keep)
    return (vv + v0).astype(np.uint32), np.asarray(sample_ids, dtype=np.uint32)[ss], i8[vv, ss]


@pytest.mark.parametrize("name", ["pgen_example", "pca_example", "syn_n1000", "streaming_example", "syn_n70000"])
def test_tidy_rows_match_reference_emission_order(name, golden):
    from plinking_duck_b200 import ops

    if name + "/i8" not in golden:
        pytest.skip("no int8 golden")
    i8 = golden[name + "/i8"]
    V, N = i8.shape
    for kw in (dict(), dict(tile_bytes=2048, n_staging=2), dict(devices=[0])):
        with _ctx(name, **kw) as c:
            for mask in (15, 7, 6, 2, 8, 9):
                var, smp, geno = c.tidy(emit_mask=mask)
                ev, es, eg = _tidy_expected(i8, np.arange(N), mask)
                assert np.array_equal(var, ev) and np.array_equal(smp, es) and np.array_equal(geno, eg), (name, mask)
            a, b = V // 4, V - V // 5
            var, smp, geno = c.tidy(a, b, emit_mask=14)
            ev, es, eg = _tidy_expected(i8[a:b], np.arange(N), 14, v0=a)
            assert np.array_equal(var, ev) and np.array_equal(smp, es) and np.array_equal(geno, eg)
            if N >= 8:
                sub = np.arange(1, N, 3)
                c.set_subset(sub)
                var, smp, geno = c.tidy(emit_mask=11)
                ev, es, eg = _tidy_expected(i8[:, sub], sub, 11)   # sample column = FILE index of the subset sample
                assert np.array_equal(var, ev) and np.array_equal(smp, es) and np.array_equal(geno, eg)
                c.set_subset(None)
    le = ops.list_entries(10, 3, 7)
    assert le.tolist() == [[10, 7], [17, 7], [24, 7]]


def test_tidy_many_batches_fixed_width():
    import synth
    from plinking_duck_b200.pgen import PgenContext

    V, N = 900, 70_001   # 24 Mi rows per batch -> several batches with mask 15
    g = synth.gen_hwe_genotypes(V, N, seed=2, missing=0.02)
    i8 = np.where(g == 3, -9, g).astype(np.int8)
    with PgenContext(image=synth.fixed_pgen_image(g)) as c:
        var, smp, geno = c.tidy(emit_mask=15)
        assert var.size == V * N
        assert np.array_equal(geno.reshape(V, N), i8)
        assert np.array_equal(smp.reshape(V, N)[5], np.arange(N)) and np.array_equal(var.reshape(V, N)[:, 3], np.arange(V))
        var, smp, geno = c.tidy(emit_mask=4 | 8)
        ev, es, eg = _tidy_expected(i8, np.arange(N), 12)
        assert np.array_equal(var, ev) and np.array_equal(smp, es) and np.array_equal(geno, eg)
