"""GPU: the sharded front-end on real devices.  With one visible GPU this exercises the device-side
result path (pd_device_result + in-place reduction) at world_size 1; tests/multi_rank_check.py is the
same check under torchrun with one rank per GPU (scripts/gpu_multi.sh)."""
import numpy as np
import pytest

from conftest import fixture_path

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["pca_example", "syn_n1000", "rare_small", "syn_dosage"])
def test_sharded_world1_device_results(name, golden):
    import torch

    from plinking_duck_b200.multi import ShardedPgen
    from plinking_duck_b200.pgen import PgenContext

    with PgenContext(fixture_path(name), device=0) as c:
        sp = ShardedPgen(c, rank=0, world=1, device=torch.device("cuda", 0))
        assert np.array_equal(sp.counts(), golden[name + "/counts"])
        assert np.array_equal(sp.missing_by_sample(), golden[name + "/missing_by_sample"])
        het, hom, mis = sp.sample_counts()
        sc = golden[name + "/sample_counts"]
        assert np.array_equal(het, sc[0]) and np.array_equal(hom, sc[1]) and np.array_equal(mis, sc[2])
        w, fl = golden[name + "/score_w"], golden[name + "/score_flip"]
        s, d, a = sp.score(np.arange(w.size), fl, w, 0)
        assert np.array_equal(a, golden[name + "/score_m0_ct"])
        assert np.max(np.abs(s[:, 0] - golden[name + "/score_m0_sum"])) <= 1e-12 * (2 * np.abs(w).sum() + 1)


def test_device_resident_image_is_used_in_place():
    import torch
    import synth
    from plinking_duck_b200.pgen import PgenContext

    g = synth.gen_hwe_genotypes(3000, 4099, seed=77, missing=0.03)
    img = synth.fixed_pgen_image(g)
    dev = torch.zeros(img.size + 256, dtype=torch.uint8, device="cuda:0")
    dev[: img.size].copy_(torch.from_numpy(img))
    with PgenContext.from_device_pointer(dev.data_ptr(), img.size, device=0) as c:
        assert c.info().resident == 1
        c.reset_stats()
        assert np.array_equal(c.counts(), synth.counts_numpy(g))
        assert c.stats().h2d_bytes == 0  # nothing was staged: the records were already in HBM
        assert np.array_equal(c.missing_by_sample(), (g == 3).sum(0).astype(np.uint32))
        assert np.array_equal(c.decode_i8(10, 40), np.where(g[10:40] == 3, -9, g[10:40]).astype(np.int8))
    # variable-width image in device memory
    raw = np.fromfile(fixture_path("pca_example"), dtype=np.uint8)
    dev2 = torch.zeros(raw.size + 256, dtype=torch.uint8, device="cuda:0")
    dev2[: raw.size].copy_(torch.from_numpy(raw))
    golden = np.load(fixture_path("pca_example").replace("fixtures/pca_example.pgen", "golden.npz"))
    with PgenContext.from_device_pointer(dev2.data_ptr(), raw.size, device=0) as c:
        assert np.array_equal(c.counts(), golden["pca_example/counts"])
        assert np.array_equal(c.decode_i8(), golden["pca_example/i8"])
