"""GPU robustness: corrupted variable-width records must come back as PD_ERR_MALFORMED (or as some result when
the corruption still parses), never as a hang, a crash or an out-of-bounds access.  The reference's negative
tests (test/sql/*_negative.test) stop at header-level damage; pgenlib itself guards every difflist component
against the record end (ParseDifflistHeader, pgenlib_read.cc:2184-2218) -- the device parsers mirror that
(every component is bounds-checked against the record's end before it is read; staged tiles and the resident
copy carry 64 bytes of slack for the aligned vector loads)."""
import numpy as np
import pytest

from conftest import fixture_path

pytestmark = pytest.mark.gpu


def _corrupt(raw: np.ndarray, data_start: int, rng, n_flips: int) -> np.ndarray:
    out = raw.copy()
    pos = rng.integers(data_start, raw.size, size=n_flips)
    out[pos] = rng.integers(0, 256, size=n_flips).astype(np.uint8)
    return out


@pytest.mark.parametrize("name", ["rare_small", "syn_n1000", "pca_example", "syn_dosage"])
def test_corrupted_records_fail_cleanly(name):
    from plinking_duck_b200._native import PlinkGpuError as PgenError
    from plinking_duck_b200.pgen import PgenContext

    raw = np.fromfile(fixture_path(name), dtype=np.uint8)
    with PgenContext(image=raw) as c:
        _, fpos = c.index()
        V, N = c.variant_ct, c.raw_sample_ct
        good = c.counts()
    data_start = int(fpos[0])
    rng = np.random.default_rng(1234)
    outcomes = {"ok": 0, "error": 0}
    for trial in range(24):
        bad = _corrupt(raw, data_start, rng, n_flips=1 + trial % 6)
        with PgenContext(image=bad, tile_bytes=4096 if trial % 2 else 0) as c:
            if trial % 3 == 2:
                c.load_resident()
            for op in (lambda: c.counts(), lambda: c.missing_by_sample(), lambda: c.sample_counts(),
                       lambda: c.decode_i8(), lambda: c.decode_phased(), lambda: c.hardy(),
                       lambda: c.dosage_counts(),
                       lambda: c.score(np.arange(V, dtype=np.uint32), None, np.ones(V), 0)):
                try:
                    op()
                    outcomes["ok"] += 1
                except PgenError:
                    outcomes["error"] += 1
    assert outcomes["ok"] + outcomes["error"] == 24 * 8
    # the context (and the GPU) must still be healthy afterwards
    with PgenContext(image=raw) as c:
        assert np.array_equal(c.counts(), good)
