#!/usr/bin/env python
"""BASELINE.json configs[4]: plink_score with K weight columns (development tool; bench.py is the contract).

A fixed-width hardcall fileset is generated on the GPU (the bench generator), written to tmpfs for the
reference arm, and scored three ways over the same scored-variant list, flips and weights:
  * GPU, records resident in HBM (tcgen05 int8-limb kernel; kernel_ms / call_ms from the library's events)
  * GPU, streamed from the page-locked host image (H2D inside the call)
  * the reference's loop (src/plink_score.cpp:598-653 as restated in oracle/ref_harness.cc: PgrGetD +
    Dosage16ToDoublesMinus9 + K axpy passes, one reader per host thread, thread-local sums merged at the end)
Sums are compared within the stated bound 1e-12 * (2 sum|w| + 1).  One JSON line."""
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
    ap.add_argument("--variants", type=int, default=20_000)
    ap.add_argument("--samples", type=int, default=100_000)
    ap.add_argument("--columns", type=int, default=16)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()

    import torch

    import bench
    from oracle import ref
    from plinking_duck_b200.pgen import PgenContext

    V, N, K = args.variants, args.samples, args.columns
    dev = torch.device("cuda", 0)
    W = (N + 3) // 4
    nbytes = 12 + V * W
    img = torch.empty(nbytes + 256, dtype=torch.uint8, device=dev)
    bench.generate_image_gpu(V, N, 4242, dev, img)
    host = bench.host_image_of(img, V, N)
    rng = np.random.default_rng(9)
    vid = np.arange(V, dtype=np.uint32)
    Wt = rng.standard_normal((V, K))
    fl = (rng.random(V) < 0.1).astype(np.uint8)
    calls = float(V) * N
    flops = 2.0 * V * N * K
    out = {"op": f"plink_score, {K} weight columns, mean imputation", "variants": V, "samples": N}

    with PgenContext.from_device_pointer(img.data_ptr(), nbytes, device=0) as c:
        S, D, A = c.score(vid, fl, Wt, 0)
        ks, cs = [], []
        for _ in range(args.reps):
            c.score(vid, fl, Wt, 0)
            st = c.stats()
            ks.append(st.kernel_ms)
            cs.append(st.call_ms)
        k, cm = float(np.median(ks)), float(np.median(cs))
        out["resident"] = {"kernel_ms": k, "call_ms": cm, "Tcalls_per_s_kernel": calls / (k * 1e-3) / 1e12,
                           "Tcalls_per_s_call": calls / (cm * 1e-3) / 1e12,
                           "fp64_equivalent_TFLOPs_kernel": flops / (k * 1e-3) / 1e12}
    with PgenContext.from_pointer(host.data_ptr(), host.numel(), device=0) as c:
        c.score(vid, fl, Wt, 0)
        ts = []
        for _ in range(args.reps):
            t = time.perf_counter()
            S2, _, _ = c.score(vid, fl, Wt, 0)
            ts.append(time.perf_counter() - t)
        # exact within a batch of scored variants; batches (which follow the staging tiles) are added in fp64
        tol = 1e-12 * (2.0 * np.abs(Wt).sum(axis=0) + 1.0)
        assert np.all(np.abs(S2 - S) <= tol[None, :]), "streamed and resident sums differ"
        out["e2e_from_host_image"] = {"s": min(ts), "Tcalls_per_s": calls / min(ts) / 1e12,
                                      "h2d_bytes": int(V) * W + Wt.nbytes}
    if ref.available():
        d = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
        path = os.path.join(d, f"plinkgpu_score_{V}x{N}.pgen")
        host.numpy().tofile(path)
        try:
            threads = os.cpu_count() or 1
            ref.bench_score(path, threads, vid[:256], fl[:256], Wt[:256])
            best, sums = None, None
            for _ in range(2):
                t, sums, cts = ref.bench_score(path, threads, vid, fl, Wt)
                best = t if best is None else min(best, t)
            bound = 1e-12 * (2.0 * np.abs(Wt).sum(axis=0) + 1.0)
            err = np.abs(S - sums)
            assert np.all(err <= bound[None, :]), float(np.max(err / bound[None, :]))
            assert np.array_equal(A, cts)
            out["cpu_reference"] = {"s": best, "threads": threads, "Tcalls_per_s": calls / best / 1e12,
                                    "max_abs_err_over_bound": float(np.max(err / bound[None, :]))}
        finally:
            if os.path.exists(path):
                os.remove(path)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
