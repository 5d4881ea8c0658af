"""pd_hwe_from_counts_noalloc: the allocation-free form of pd_hwe_from_counts / pd_hwe_lnp_from_counts (scratch in a
static device array; include/plinkgpu.h).  It has no counterpart in the reference,
so the yardstick is the existing entry point: same kernels, so the results must agree -- checked to the
1e-12 relative bound the HWE tests state (tests/hwe_check.py::EXACT_ABS), and in fact bit for bit.

Collected last on purpose; it must pass alone in a fresh process and after every
other test.  Inputs and outputs are device buffers: the TEST allocates them (torch), the entry point may not."""
import ctypes as C

import numpy as np
import pytest

import hwe_check as hc

pytestmark = pytest.mark.gpu

REL_TOL = hc.EXACT_ABS  # 1e-12


def _lib():
    from plinking_duck_b200 import _native as nat

    return nat.load(), nat


def _existing(counts, midp, want_lnp):
    from conftest import fixture_path
    from plinking_duck_b200.pgen import PgenContext

    with PgenContext(fixture_path("pgen_example")) as c:
        return c.hwe_lnp_from_counts(counts, midp=bool(midp)) if want_lnp else c.hwe_from_counts(counts, midp=bool(midp))


def _noalloc(counts, midp, want_lnp, stream=None, poison=None):
    """One call through the C ABI on device buffers owned by the test; returns the results as a host array."""
    import torch

    lib, nat = _lib()
    n = counts.shape[0]
    d_counts = torch.from_numpy(np.ascontiguousarray(counts, dtype=np.uint32).view(np.int32).copy()).cuda()
    d_out = torch.full((max(n, 1),), float("nan") if poison is None else poison, dtype=torch.float64, device="cuda")
    err = C.create_string_buffer(256)
    s = stream if stream is not None else torch.cuda.current_stream()
    with torch.cuda.stream(s):
        rc = lib.pd_hwe_from_counts_noalloc(d_counts.data_ptr(), n, int(midp), int(want_lnp), d_out.data_ptr(),
                                            C.c_void_p(s.cuda_stream), err, len(err))
    assert rc == nat.PD_OK, err.value.decode()
    s.synchronize()
    return d_out[:n].cpu().numpy()


def _agree(got, want, want_lnp):
    assert got.shape == want.shape
    assert not np.isnan(got).any()
    if want_lnp:
        fin = np.isfinite(want)
        assert np.array_equal(np.isfinite(got), fin) and np.array_equal(got[~fin], want[~fin])
        assert np.all(np.abs(got[fin] - want[fin]) <= REL_TOL + 4 * np.spacing(np.abs(want[fin])))
    else:
        assert np.all(np.abs(got - want) <= REL_TOL * np.abs(want))
    assert np.array_equal(got, want)  # the same kernels on the same tables: identical bits


def _tables():
    # tables from equilibrium to ln p ~ -350 000 at three sample sizes: the ones at N >= 100003 overflow the fp64 walk
    # and go through the scratch list (k_hwe_scaled)
    return np.concatenate([hc.departure_grid(1000, 5), hc.departure_grid(100003, 6), hc.departure_grid(500000, 7)])


@pytest.mark.parametrize("want_lnp", [0, 1])
@pytest.mark.parametrize("midp", [0, 1])
def test_noalloc_matches_existing_entry_point_twice_in_a_row(midp, want_lnp):
    counts = _tables()
    want = _existing(counts, midp, want_lnp)
    # two calls in a row on different inputs of different lengths: the shared scratch keeps the first call's
    # contents and the second must not see them
    first = _noalloc(counts, midp, want_lnp)
    _agree(first, want, want_lnp)
    sub = counts[::-1][: counts.shape[0] // 3].copy()
    second = _noalloc(sub, midp, want_lnp)
    _agree(second, _existing(sub, midp, want_lnp), want_lnp)
    third = _noalloc(counts, midp, want_lnp)
    _agree(third, want, want_lnp)


def test_noalloc_on_a_side_stream_and_at_a_larger_size():
    import torch

    rng = np.random.default_rng(11)
    base = _tables()
    counts = base[rng.integers(0, base.shape[0], 200_000)]
    want = _existing(counts, 0, 0)
    s = torch.cuda.Stream()
    _agree(_noalloc(counts, 0, 0, stream=s), want, 0)
    _agree(_noalloc(counts, 0, 0, stream=s), want, 0)


def test_noalloc_rejects_sizes_beyond_the_static_scratch_and_bad_pointers():
    import torch

    lib, nat = _lib()
    limit = lib.pd_hwe_noalloc_max_variants()
    assert limit == 1 << 20
    d = torch.zeros(16, dtype=torch.float64, device="cuda")
    err = C.create_string_buffer(256)
    rc = lib.pd_hwe_from_counts_noalloc(d.data_ptr(), limit + 1, 0, 0, d.data_ptr(), None, err, len(err))
    assert rc == nat.PD_ERR_INVALID_ARG
    msg = err.value.decode()
    assert "PD_NOALLOC_HWE_MAX_VARIANTS" in msg and str(limit) in msg
    # host memory is not accepted (no staging buffers exist to copy it through)
    host = np.zeros(64, dtype=np.uint32)
    rc = lib.pd_hwe_from_counts_noalloc(host.ctypes.data, 4, 0, 0, d.data_ptr(), None, err, len(err))
    assert rc == nat.PD_ERR_INVALID_ARG and "device memory" in err.value.decode()
    assert lib.pd_hwe_from_counts_noalloc(None, 0, 0, 0, None, None, None, 0) == nat.PD_OK  # n = 0: nothing to do


def test_noalloc_call_allocates_no_device_memory():
    """cudaMemGetInfo before and after a call (buffers made beforehand, library already initialised)."""
    import torch

    counts = _tables()
    _noalloc(counts, 0, 0)  # first use on this device: symbol address, module load
    lib, nat = _lib()
    n = counts.shape[0]
    d_counts = torch.from_numpy(counts.view(np.int32).copy()).cuda()
    d_out = torch.empty(n, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        rc = lib.pd_hwe_from_counts_noalloc(d_counts.data_ptr(), n, 1, 1, d_out.data_ptr(),
                                            C.c_void_p(torch.cuda.current_stream().cuda_stream), None, 0)
        assert rc == nat.PD_OK
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
