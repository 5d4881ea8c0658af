"""CPU: the multi-GPU host logic (variant-range partition by record bytes + one all-reduce of the
per-sample accumulators) under torch.distributed `gloo`, world_size 2.  The per-rank compute is an
oracle-backed stand-in (the real ranks call libplinkgpu); what is tested here is that the partition
covers every variant exactly once and that the reductions reproduce the single-process goldens."""
import os
import socket

import numpy as np
import pytest

from conftest import GOLDEN_DIR, ROOT, fixture_names, fixture_path


@pytest.mark.parametrize("name", fixture_names())
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_partition_covers_and_balances(name, world, golden):
    from plinking_duck_b200.multi import partition_variants

    vt, fpos = golden[name + "/vrtypes"], golden[name + "/fpos"]
    V = vt.size
    parts = partition_variants(fpos, vt, world)
    assert len(parts) == world and parts[0][0] == 0 and parts[-1][1] == V
    for (a, b), (c, d) in zip(parts[:-1], parts[1:]):
        assert b == c and a <= b
    for a, b in parts:
        if a < b and a > 0:
            assert (int(vt[a]) & 6) != 2 or a == parts[0][0], "range starts on an LD-compressed record"
    total = int(fpos[V] - fpos[0])
    biggest = int((fpos[1:] - fpos[:-1]).max())
    ld_runs = 0
    if ((vt & 6) == 2).any():
        run = 0
        for t in vt:
            run = run + 1 if (int(t) & 6) == 2 else 0
            ld_runs = max(ld_runs, run)
    for a, b in parts:
        assert int(fpos[b] - fpos[a]) <= total // world + (ld_runs + 2) * biggest


def test_partition_list_balances():
    from plinking_duck_b200.multi import partition_list

    fpos = np.arange(0, 1001 * 10, 10, dtype=np.uint64)
    vidx = np.arange(0, 1000, 3)
    parts = partition_list(vidx, fpos, 4)
    assert parts[0][0] == 0 and parts[-1][1] == vidx.size
    sizes = [b - a for a, b in parts]
    assert max(sizes) - min(sizes) <= 1
    assert partition_list(np.array([], dtype=np.int64), fpos, 3) == [(0, 0)] * 3


class OracleLocal:
    """Stand-in for PgenContext on a CPU-only box (test infrastructure: pgenlib through oracle/_ref)."""

    def __init__(self, path):
        from oracle import ref

        self.r = ref.RefPgen(path)
        self.variant_ct, self.sample_ct = self.r.V, self.r.N

    def index(self):
        return self.r.vrtypes(), self.r.fpos()

    def counts(self, a, b):
        return self.r.counts(a, b)

    def missing_by_sample(self, a, b):
        return self.r.missingness(a, b)[1]

    def sample_counts(self, a, b):
        return self.r.sample_counts(a, b)

    def score(self, vidx, flip, W, mode):
        W = np.asarray(W, dtype=np.float64).reshape(len(vidx), -1)
        fl = np.zeros(len(vidx), np.uint8) if flip is None else flip
        cols = [self.r.score(vidx, fl, W[:, k], mode) for k in range(W.shape[1])]
        s = np.stack([c[0] for c in cols], axis=1) if len(vidx) or True else None
        return s, cols[0][1], cols[0][2]


def _worker(rank, world, port, name, out_dir):
    import sys

    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist

    from plinking_duck_b200.multi import ShardedPgen

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sp = ShardedPgen(OracleLocal(fixture_path(name)), rank=rank, world=world)
        counts = sp.counts()
        mis = sp.missing_by_sample()
        het, hom, mis2 = sp.sample_counts()
        g = np.load(os.path.join(GOLDEN_DIR, "golden.npz"))
        w, fl = g[name + "/score_w"], g[name + "/score_flip"]
        s, d, a = sp.score(np.arange(w.size), fl, np.stack([w, 2 * w], axis=1), 0)
        np.savez(os.path.join(out_dir, f"r{rank}.npz"), counts=counts, mis=mis, het=het, hom=hom, mis2=mis2, s=s, d=d,
                 a=a, v0=sp.v0, v1=sp.v1)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["pca_example", "syn_n1000", "rare_small"])
def test_two_rank_gloo_matches_single_process(name, golden, refmod, tmp_path):
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    world = 2
    mp.spawn(_worker, args=(world, port, name, str(tmp_path)), nprocs=world, join=True)
    res = [np.load(str(tmp_path / f"r{r}.npz")) for r in range(world)]
    assert res[0]["v1"] == res[1]["v0"] and res[0]["v0"] == 0
    w = golden[name + "/score_w"]
    for r in res:
        assert np.array_equal(r["counts"], golden[name + "/counts"])
        assert np.array_equal(r["mis"], golden[name + "/missing_by_sample"])
        sc = golden[name + "/sample_counts"]
        assert np.array_equal(r["het"], sc[0]) and np.array_equal(r["hom"], sc[1]) and np.array_equal(r["mis2"], sc[2])
        scale = 2.0 * np.abs(w).sum() + 1.0
        assert np.max(np.abs(r["s"][:, 0] - golden[name + "/score_m0_sum"])) <= 1e-12 * scale
        assert np.max(np.abs(r["s"][:, 1] - 2 * golden[name + "/score_m0_sum"])) <= 2e-12 * scale
        assert np.array_equal(r["a"], golden[name + "/score_m0_ct"])
        assert np.allclose(r["d"], golden[name + "/score_m0_dsum"], rtol=1e-12)


def _worker_shards(rank, world, port, names, out_dir):
    import sys

    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist

    from plinking_duck_b200.multi import ShardedPgen

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        name = names[rank]
        sp = ShardedPgen(OracleLocal(fixture_path(name)), rank=rank, world=world, prepartitioned=True)
        assert (sp.v0, sp.v1) == (0, sp.V)
        mis = sp.missing_by_sample()
        het, hom, mis2 = sp.sample_counts()
        g = np.load(os.path.join(GOLDEN_DIR, "golden.npz"))
        w, fl = g[name + "/score_w"], g[name + "/score_flip"]
        s, d, a = sp.score(np.arange(w.size), fl, w, 0)
        np.savez(os.path.join(out_dir, f"s{rank}.npz"), mis=mis, het=het, hom=hom, mis2=mis2, s=s, d=d, a=a,
                 local_counts=sp.counts_local())
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_prepartitioned_shard_files(golden, refmod, tmp_path):
    """One shard FILE per rank (the reference's shard1 / shard2 fixtures, same 8 samples): per-variant outputs
    stay local, per-sample accumulators and scores are summed over the ranks."""
    import torch.multiprocessing as mp

    names = ["shard1", "shard2"]
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker_shards, args=(2, port, names, str(tmp_path)), nprocs=2, join=True)
    res = [np.load(str(tmp_path / f"s{r}.npz")) for r in range(2)]
    exp_mis = sum(golden[n + "/missing_by_sample"] for n in names)
    exp_sc = sum(golden[n + "/sample_counts"].astype(np.int64) for n in names)
    exp_s = sum(golden[n + "/score_m0_sum"] for n in names)
    exp_a = sum(golden[n + "/score_m0_ct"].astype(np.int64) for n in names)
    scale = sum(2.0 * np.abs(golden[n + "/score_w"]).sum() + 1.0 for n in names)
    for r, n in zip(res, names):
        assert np.array_equal(r["local_counts"], golden[n + "/counts"])
        assert np.array_equal(r["mis"], exp_mis) and np.array_equal(r["mis2"], exp_sc[2])
        assert np.array_equal(r["het"], exp_sc[0]) and np.array_equal(r["hom"], exp_sc[1])
        assert np.max(np.abs(r["s"][:, 0] - exp_s)) <= 1e-12 * scale
        assert np.array_equal(r["a"].astype(np.int64), exp_a)
