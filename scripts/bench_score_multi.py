#!/usr/bin/env python
"""BASELINE.json configs[4] across GPUs (development tool): plink_score with K weight columns, the scored-variant
list cut into contiguous chunks of equal record bytes, one chunk per rank, per-sample partial sums all-reduced
over NCCL in place in HBM (plinking_duck_b200/multi.py).  Weak scaling: V = variants_per_gpu x world.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29531 \
      scripts/bench_score_multi.py

Every rank holds ITS chunk of the fileset resident in HBM (a contiguous variant range, generated from a
per-rank seed) and scores all of it; wall clock between barriers, max over ranks; rank 0 prints one JSON line.
Parity of the partitioned + reduced sums against a single-GPU run is tests/multi_rank_check.py's job."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants-per-gpu", type=int, default=400_000)
    ap.add_argument("--samples", type=int, default=100_000)
    ap.add_argument("--columns", type=int, default=16)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    import bench
    from plinking_duck_b200.multi import ShardedPgen
    from plinking_duck_b200.pgen import PgenContext

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    V, N, K = args.variants_per_gpu, args.samples, args.columns
    W = (N + 3) // 4
    nbytes = 12 + V * W
    img = torch.empty(nbytes + 256, dtype=torch.uint8, device=dev)
    bench.generate_image_gpu(V, N, 99 + rank, dev, img)
    rng = np.random.default_rng(17 + rank)
    vid = np.arange(V, dtype=np.uint32)
    Wt = rng.standard_normal((V, K))
    fl = (rng.random(V) < 0.1).astype(np.uint8)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()

    with PgenContext.from_device_pointer(img.data_ptr(), nbytes, device=local_rank) as c:
        sh = ShardedPgen(c, rank, world, device=dev, prepartitioned=True)
        S, D, A = sh.score(vid, fl, Wt, 0)  # warm-up (NCCL communicator, workspaces)
        ts = []
        for _ in range(args.reps):
            barrier()
            t = time.perf_counter()
            S, D, A = sh.score(vid, fl, Wt, 0)
            barrier()
            ts.append(time.perf_counter() - t)
        local_call_ms = c.stats().call_ms      # this rank's pd_score alone (last repetition)
        local_kernel_ms = c.stats().kernel_ms
        # the three all-reduces alone, on buffers of the same sizes
        bufs = [torch.zeros(N * K, dtype=torch.float64, device=dev), torch.zeros(N, dtype=torch.float64, device=dev),
                torch.zeros(N, dtype=torch.int32, device=dev)]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ar_ms = 0.0
        if world > 1:
            for b in bufs:
                dist.all_reduce(b)
            barrier()
            e0.record()
            for b in bufs:
                dist.all_reduce(b)
            e1.record()
            torch.cuda.synchronize(dev)
            ar_ms = e0.elapsed_time(e1)
        tt = torch.tensor([min(ts)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        best = float(tt.item())
        if rank == 0:
            assert int(A[0]) == 2 * V * world, "allele counts were not summed over all ranks"
            print(json.dumps({
                "op": f"plink_score, {K} weight columns, scored list partitioned over {world} GPU(s), NCCL all-reduce of "
                      f"double[N x K] + double[N] + u32[N] in HBM",
                "n_gpus": world, "variants_total": V * world, "variants_per_gpu": V, "samples": N,
                "s_per_call": best, "Tcalls_per_s": float(V) * world * N / best / 1e12,
                "fp64_equivalent_TFLOPs": 2.0 * V * world * N * K / best / 1e12,
                "rank0_pd_score_call_ms": local_call_ms, "rank0_score_kernel_ms": local_kernel_ms,
                "allreduce_ms": ar_ms, "scaling": "weak"}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
