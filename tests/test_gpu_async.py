"""Asynchronous jobs behind the scan callbacks: pd_counts_start, pd_hardy_start and pd_decode_start
jobs fetched by 16 threads that claim 128-variant batches from a shared counter, the way DuckDB's scan threads do
(src/plink_freq.cpp:434-441, src/plink_hardy.cpp:481-510, src/pfile_reader.cpp:2790-2835).  Every row must come out
once, bit-exact (p-values identical to the blocking call), on single- and multi-GPU contexts."""
import itertools
import threading

import numpy as np
import pytest

from conftest import fixture_path

pytestmark = pytest.mark.gpu


def _claim_loop(V, fetch, n_threads=16, batch=128):
    """n_threads workers claim [a, a + batch) from a shared counter and call fetch(a, b); returns {a: result}."""
    counter = itertools.count()
    lock = threading.Lock()
    out, errors = {}, []

    def work():
        try:
            while True:
                with lock:
                    a = next(counter) * batch
                if a >= V:
                    return
                b = min(V, a + batch)
                r = fetch(a, b)
                with lock:
                    out[a] = r
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    ts = [threading.Thread(target=work) for _ in range(n_threads)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errors, errors
    return out


def _devices_or_skip(k):
    import torch

    if k and torch.cuda.device_count() < k:
        pytest.skip(f"needs {k} GPUs")
    return list(range(k)) if k else None


@pytest.mark.parametrize("k", [0, 1, 2])
@pytest.mark.parametrize("name", ["pca_example", "syn_n1000", "streaming_example"])
def test_hardy_job_16_threads(name, k, golden):
    from plinking_duck_b200.pgen import PgenContext

    exp = golden[name + "/counts"]
    V = exp.shape[0]
    with PgenContext(fixture_path(name), devices=_devices_or_skip(k), tile_bytes=4096) as c:
        _, p_ref = c.hardy(midp=True)
        job = c.hardy_async(0, V, midp=True)
        got = _claim_loop(V, job.fetch, batch=16)
        job.close()
        cj = c.counts_async(0, V)
        gc = _claim_loop(V, cj.fetch, batch=16)
        cj.close()
    assert sorted(got) == list(range(0, V, 16))
    counts = np.concatenate([got[a][0] for a in sorted(got)])
    p = np.concatenate([got[a][1] for a in sorted(got)])
    assert np.array_equal(counts, exp) and np.array_equal(p, p_ref)
    assert np.array_equal(np.concatenate([gc[a] for a in sorted(gc)]), exp)


@pytest.mark.parametrize("k", [0, 2])
def test_hardy_job_fixed_width_fused_kernel(k, tmp_path):
    import synth
    from plinking_duck_b200.pgen import PgenContext

    V, N = 5000, 7001
    g = synth.gen_hwe_genotypes(V, N, seed=3, missing=0.02)
    img = synth.fixed_pgen_image(g)
    exp = synth.counts_numpy(g)
    with PgenContext(image=img, devices=_devices_or_skip(k), tile_bytes=1 << 20) as c:
        _, p_ref = c.hardy()
        for resident in (False, True):
            if resident:
                c.load_resident()
            job = c.hardy_async(100, V - 50)
            got = _claim_loop(V - 150, lambda a, b: job.fetch(100 + a, 100 + b))
            job.close()
            counts = np.concatenate([got[a][0] for a in sorted(got)])
            p = np.concatenate([got[a][1] for a in sorted(got)])
            assert np.array_equal(counts, exp[100:V - 50]) and np.array_equal(p, p_ref[100:V - 50])


@pytest.mark.parametrize("k", [0, 1, 2])
def test_decode_job_16_threads(k, tmp_path):
    """Many chunks (the ring of four chunk buffers is recycled several times), ragged last chunk, a sample subset."""
    import synth
    from plinking_duck_b200.pgen import PgenContext

    V, N = 3000, 150_001  # 150 KB of int8 per row -> 384-row chunks -> 8 chunks
    g = synth.gen_hwe_genotypes(V, N, seed=5, missing=0.03)
    exp = np.where(g == 3, -9, g).astype(np.int8)
    with PgenContext(image=synth.fixed_pgen_image(g), devices=_devices_or_skip(k)) as c:
        job = c.decode_async(7, V)
        got = _claim_loop(V - 7, lambda a, b: job.fetch(7 + a, 7 + b))
        job.close()
        out = np.concatenate([got[a] for a in sorted(got)])
        assert np.array_equal(out, exp[7:])
        sub = np.arange(1, N, 5)
        c.set_subset(sub)
        job = c.decode_async(0, 500)
        got = _claim_loop(500, job.fetch, n_threads=4, batch=64)
        job.close()
        assert np.array_equal(np.concatenate([got[a] for a in sorted(got)]), exp[:500][:, sub])
        # a job that is dropped half-way must not hang
        job = c.decode_async(0, V)
        job.fetch(0, 128)
        job.close()


def test_decode_job_compressed_fixture(golden):
    from plinking_duck_b200.pgen import PgenContext

    name = "syn_n1000"
    exp = golden[name + "/i8"]
    V = exp.shape[0]
    with PgenContext(fixture_path(name)) as c:
        job = c.decode_async(0, V)
        got = _claim_loop(V, job.fetch, batch=32)
        job.close()
    assert np.array_equal(np.concatenate([got[a] for a in sorted(got)]), exp)
