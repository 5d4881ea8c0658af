"""CPU: the variant-range plan of a multi-GPU context (pd_plan_partition -> PartitionByBytes, csrc/api_group.cu):
contiguous, covering, byte-balanced, pieces start on a non-LD record; a multi-GPU open without a device fails
loudly like a single-GPU one."""
import numpy as np
import pytest

from conftest import fixture_names, fixture_path


@pytest.mark.parametrize("name", fixture_names())
@pytest.mark.parametrize("k", [1, 2, 3, 8])
def test_partition_plan_properties(name, k, golden, built):
    from plinking_duck_b200.pgen import plan_partition

    vt = golden[name + "/vrtypes"]
    fpos = golden[name + "/fpos"].astype(np.int64)
    V = vt.size
    cuts = plan_partition(fixture_path(name), k).astype(np.int64)
    assert cuts[0] == 0 and cuts[-1] == V and np.all(np.diff(cuts) >= 0)
    for c in cuts[1:-1]:
        if c < V:
            assert (int(vt[c]) & 6) != 2, "a piece must not start on an LD-compressed record"
    # balance: no piece exceeds its fair share by more than the records an LD snap-back can move
    total = fpos[V] - fpos[0]
    widths = np.diff(fpos)
    ld_runs = 0
    run = 0
    for v in range(V):
        run = run + widths[v] if (int(vt[v]) & 6) == 2 else 0
        ld_runs = max(ld_runs, run)
    bytes_per = fpos[cuts[1:]] - fpos[cuts[:-1]]
    assert bytes_per.sum() == total
    slack = widths.max() + ld_runs + 1
    assert np.all(bytes_per <= total // k + 2 * slack), (bytes_per.tolist(), total, k)
    # sub-ranges
    a, b = V // 5, V - V // 7
    sub = plan_partition(fixture_path(name), k, a, b).astype(np.int64)
    assert sub[0] == a and sub[-1] == b and np.all(np.diff(sub) >= 0)


def test_partition_plan_fixed_width(tmp_path, built):
    import synth
    from plinking_duck_b200.pgen import plan_partition

    g = synth.gen_hwe_genotypes(1001, 37, seed=3)
    path = str(tmp_path / "fx.pgen")
    synth.write_fixed_pgen(path, g)
    cuts = plan_partition(path, 8)
    assert cuts[0] == 0 and cuts[-1] == 1001
    assert np.diff(cuts.astype(np.int64)).max() - np.diff(cuts.astype(np.int64)).min() <= 1
    assert plan_partition(path, 4, 10, 10).tolist() == [10] * 5


def test_multi_gpu_open_has_no_cpu_fallback(built):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from plinking_duck_b200 import _native as nat
    from plinking_duck_b200.pgen import PgenContext

    with pytest.raises(nat.PlinkGpuError) as e:
        PgenContext(fixture_path("pgen_example"), devices=[0, 1])
    assert e.value.code == nat.PD_ERR_CUDA
