"""Run under torchrun, one rank per GPU: every rank scans its variant range with libplinkgpu, per-sample
accumulators are all-reduced over NCCL in place, and every rank checks the result against the goldens.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/multi_rank_check.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist

    from plinking_duck_b200.multi import ShardedPgen
    from plinking_duck_b200.pgen import PgenContext

    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    golden = np.load(os.path.join(ROOT, "tests", "golden", "golden.npz"))
    for name in ("pca_example", "syn_n1000", "streaming_example", "syn_dosage", "rare_small"):
        path = os.path.join(ROOT, "tests", "golden", "fixtures", name + ".pgen")
        with PgenContext(path, device=lr) as c:
            sp = ShardedPgen(c, rank=rank, world=world, device=dev)
            assert np.array_equal(sp.counts(), golden[name + "/counts"]), name
            assert np.array_equal(sp.missing_by_sample(), golden[name + "/missing_by_sample"]), name
            het, hom, mis = sp.sample_counts()
            sc = golden[name + "/sample_counts"]
            assert np.array_equal(het, sc[0]) and np.array_equal(hom, sc[1]) and np.array_equal(mis, sc[2]), name
            if name + "/score_w" in golden:
                w, fl = golden[name + "/score_w"], golden[name + "/score_flip"]
                W = np.stack([w * (k + 1) for k in range(16)], axis=1)
                s, d, a = sp.score(np.arange(w.size), fl, W, 0)
                scale = 2 * np.abs(w).sum() + 1
                for k in (0, 15):
                    assert np.max(np.abs(s[:, k] - (k + 1) * golden[name + "/score_m0_sum"])) <= 1e-12 * scale * (k + 1), name
                assert np.array_equal(a, golden[name + "/score_m0_ct"]), name
                assert np.allclose(d, golden[name + "/score_m0_dsum"], rtol=1e-12), name
        dist.barrier()
    if rank == 0:
        print(f"multi_rank_check ok: world={world}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
