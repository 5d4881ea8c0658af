"""pd_fileset: variant-sharded file lists (read_pfile's multi-file sources, src/pfile_reader.cpp:325-400,2133-2151).
A fileset over k files that hold consecutive variant ranges of one dataset must return exactly what a single context
over the unsplit file returns (and therefore what pgenlib returns: tests/golden): per-variant outputs concatenated,
per-sample outputs summed over the files."""
import ctypes as C
import os

import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu


def _open(paths, **kw):
    from plinking_duck_b200 import _native as nat

    lib = nat.load()
    arr = (C.c_char_p * len(paths))(*[os.fsencode(p) for p in paths])
    opts, keep = nat.open_opts(**kw)
    h = C.c_void_p()
    err = C.create_string_buffer(512)
    rc = lib.pd_fileset_open(arr, len(paths), C.byref(opts), C.byref(h), err, len(err))
    return lib, nat, rc, h, err.value.decode()


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.mark.parametrize("kw", [dict(), dict(tile_bytes=1 << 16, n_staging=2), dict(devices=[0]), dict(devices=[0, 1])])
def test_fileset_equals_the_unsplit_file(tmp_path, kw):
    from conftest import need_devices

    need_devices(kw)
    from plinking_duck_b200.pgen import PgenContext

    V, N, K = 5000, 4099, 3
    g = synth.gen_hwe_genotypes(V, N, seed=12, missing=0.02)
    whole = str(tmp_path / "whole.pgen")
    synth.write_fixed_pgen(whole, g)
    cuts = [0, 1700, 1701, 3900, V]  # a one-variant file among them
    paths = []
    for i in range(len(cuts) - 1):
        p = str(tmp_path / f"part{i}.pgen")
        synth.write_fixed_pgen(p, g[cuts[i]:cuts[i + 1]])
        paths.append(p)
    rng = np.random.default_rng(4)
    vid = np.sort(rng.choice(V, 1200, replace=False)).astype(np.uint32)
    W = rng.standard_normal((vid.size, K))
    flip = (rng.random(vid.size) < 0.3).astype(np.uint8)
    with PgenContext(whole) as c:
        exp_counts, exp_p = c.hardy()
        exp_i8 = c.decode_i8(1500, 4000)
        exp_sc = c.sample_counts(100, 4500)
        exp_ms = c.missing_by_sample()
        exp_score = c.score(vid, flip, W, 1)
        sub = np.arange(3, N, 5)
        c.set_subset(sub)
        exp_sub_counts = c.counts(1000, 4200)
        exp_sub_ms = c.missing_by_sample(1000, 4200)
    lib, nat, rc, fs, msg = _open(paths, **kw)
    assert rc == nat.PD_OK, msg
    try:
        assert lib.pd_fileset_file_count(fs) == 4 and lib.pd_fileset_variant_ct(fs) == V and lib.pd_fileset_sample_ct(fs) == N
        off = np.zeros(5, dtype=np.uint32)
        assert lib.pd_fileset_offsets(fs, _p(off)) == nat.PD_OK and off.tolist() == cuts
        counts = np.zeros((V, 4), dtype=np.uint32)
        p = np.zeros(V)
        assert lib.pd_fileset_hardy(fs, 0, V, 0, _p(counts), _p(p)) == nat.PD_OK
        assert np.array_equal(counts, exp_counts) and np.array_equal(p, exp_p)
        c2 = np.zeros((2400, 4), dtype=np.uint32)
        assert lib.pd_fileset_counts(fs, 1600, 4000, _p(c2)) == nat.PD_OK and np.array_equal(c2, exp_counts[1600:4000])
        i8 = np.zeros((2500, N), dtype=np.int8)
        assert lib.pd_fileset_decode_i8(fs, 1500, 4000, _p(i8)) == nat.PD_OK and np.array_equal(i8, exp_i8)
        h, a, m = (np.zeros(N, dtype=np.uint32) for _ in range(3))
        assert lib.pd_fileset_sample_counts(fs, 100, 4500, _p(h), _p(a), _p(m)) == nat.PD_OK
        assert np.array_equal(h, exp_sc[0]) and np.array_equal(a, exp_sc[1]) and np.array_equal(m, exp_sc[2])
        ms = np.zeros(N, dtype=np.uint32)
        assert lib.pd_fileset_missing_by_sample(fs, 0, V, _p(ms)) == nat.PD_OK and np.array_equal(ms, exp_ms)
        ss, ds, ac = np.zeros((N, K)), np.zeros(N), np.zeros(N, dtype=np.uint32)
        rc = lib.pd_fileset_score(fs, _p(vid), _p(flip), _p(W), vid.size, K, 1, _p(ss), _p(ds), _p(ac))
        assert rc == nat.PD_OK, lib.pd_fileset_last_error(fs)
        tol = 1e-12 * (2 * np.abs(W).sum() + 1)  # the files are added in a different order than one batch adds
        assert np.abs(ss - exp_score[0]).max() <= tol and np.abs(ds - exp_score[1]).max() <= tol
        assert np.array_equal(ac, exp_score[2])
        # subset: applied to every file
        bits = np.zeros((N + 63) // 64, dtype=np.uint64)
        np.bitwise_or.at(bits, sub >> 6, np.uint64(1) << (sub & 63).astype(np.uint64))
        assert lib.pd_fileset_set_sample_subset(fs, _p(bits), sub.size) == nat.PD_OK
        c3 = np.zeros((3200, 4), dtype=np.uint32)
        assert lib.pd_fileset_counts(fs, 1000, 4200, _p(c3)) == nat.PD_OK and np.array_equal(c3, exp_sub_counts)
        ms2 = np.zeros(sub.size, dtype=np.uint32)
        assert lib.pd_fileset_missing_by_sample(fs, 1000, 4200, _p(ms2)) == nat.PD_OK and np.array_equal(ms2, exp_sub_ms)
        # errors: a range past the end; the per-file context is reachable for everything else
        assert lib.pd_fileset_counts(fs, 0, V + 1, _p(counts)) == nat.PD_ERR_INVALID_ARG
        assert b"out of bounds" in lib.pd_fileset_last_error(fs)
        assert lib.pd_fileset_ctx(fs, 3) and not lib.pd_fileset_ctx(fs, 4)
    finally:
        lib.pd_fileset_close(fs)


def test_fileset_rejects_mismatched_sample_counts_and_missing_files(tmp_path):
    g = synth.gen_hwe_genotypes(50, 100, seed=1)
    a, b = str(tmp_path / "a.pgen"), str(tmp_path / "b.pgen")
    synth.write_fixed_pgen(a, g)
    synth.write_fixed_pgen(b, g[:, :99])
    lib, nat, rc, fs, msg = _open([a, b])
    assert rc == nat.PD_ERR_INVALID_ARG and "99 samples" in msg and not fs
    lib, nat, rc, fs, msg = _open([a, str(tmp_path / "nope.pgen")])
    assert rc == nat.PD_ERR_OPEN and "nope.pgen" in msg


def test_fileset_of_compressed_fixtures_vs_golden(golden):
    """A compressed fixture of the reference listed twice: the per-variant rows repeat, the per-sample sums double."""
    from conftest import fixture_path

    name = "streaming_example"
    path = fixture_path(name)
    lib, nat, rc, fs, msg = _open([path, path], tile_bytes=4096)
    assert rc == nat.PD_OK, msg
    try:
        V = golden[name + "/counts"].shape[0]
        N = lib.pd_fileset_sample_ct(fs)
        counts = np.zeros((2 * V, 4), dtype=np.uint32)
        assert lib.pd_fileset_counts(fs, 0, 2 * V, _p(counts)) == nat.PD_OK
        assert np.array_equal(counts[:V], golden[name + "/counts"]) and np.array_equal(counts[V:], golden[name + "/counts"])
        ms = np.zeros(N, dtype=np.uint32)
        assert lib.pd_fileset_missing_by_sample(fs, 0, 2 * V, _p(ms)) == nat.PD_OK
        assert np.array_equal(ms, 2 * golden[name + "/missing_by_sample"])
        ms1 = np.zeros(N, dtype=np.uint32)  # a range inside the second file only
        assert lib.pd_fileset_missing_by_sample(fs, V, 2 * V, _p(ms1)) == nat.PD_OK
        assert np.array_equal(ms1, golden[name + "/missing_by_sample"])
    finally:
        lib.pd_fileset_close(fs)
