#!/usr/bin/env python
"""BASELINE.json configs[3]: plink_freq counts + plink_missing sample mode on a difflist / LD-compressed
.pgen (development tool; bench.py is the contract).

Builds the SURVEY 8d "S4" fileset on this box: N samples, V variants, MAF model of the reference's
test/data/generate_rare_test_data.sh (95 % log-uniform [1/2N, 1 %], 5 % log-uniform [1 %, 50 %], 1 %
missing), genotypes generated on the GPU and written with pgenlib's own writer (oracle/_ref), so the file
holds whatever mix of vrtypes 0/1/2/3/4/5/6 plink2 itself would emit.  Then times, through the C ABI,
  counts:            streamed from the file (e2e) and HBM-resident (kernel)
  missing_by_sample: resident
against the reference's multi-threaded CPU scan (pgenlib, one reader per thread, 128-variant claims) and
checks the integer results bit for bit against it.  One JSON line per operation."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def make_file(path, V, N, seed, device, plain=False):
    """The C4 fileset of tests/c4.py: rare-variant MAF model + LD blocks + ALT-major + one-bit rows (plain: the MAF
    model alone = the round-1 rare-variant fileset)."""
    import c4

    c4.make_c4_file(path, V, N, seed, device, plain=plain)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants", type=int, default=30_000)
    ap.add_argument("--samples", type=int, default=100_000)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--plain", action="store_true", help="rare-variant MAF model only (vrtypes 0 / 1 / 4), as in round 1")
    args = ap.parse_args()

    import torch

    from oracle import ref
    from plinking_duck_b200.pgen import PgenContext

    if not ref.available():
        raise SystemExit("oracle/_ref/libpgenref.so is needed to write the compressed fileset")
    V, N = args.variants, args.samples
    dev = torch.device("cuda", 0)
    d = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
    path = os.path.join(d, f"plinkgpu_rare_{V}x{N}.pgen")
    t0 = time.time()
    make_file(path, V, N, 777, dev, plain=args.plain)
    size = os.path.getsize(path)
    threads = os.cpu_count() or 1
    try:
        with ref.RefPgen(path) as r:
            vt = np.asarray(r.vrtypes()) & 7
        hist = {int(k): int((vt == k).sum()) for k in np.unique(vt)}
        meta = {"file_bytes": size, "bytes_per_variant": size / V, "dense_bytes_per_variant": (N + 3) // 4,
                "vrtype_histogram": hist, "build_s": round(time.time() - t0, 1)}
        print(json.dumps({"op": "fileset", "variants": V, "samples": N, **meta}), flush=True)

        def cpu(kind):
            ref.bench_scan(path, threads, 0, min(V, 2048), kind)
            return min(ref.bench_scan(path, threads, 0, V, kind)[0] for _ in range(3))

        cpu_counts_s = cpu(0)
        cpu_missing_s = cpu(3)
        with ref.RefPgen(path) as r:
            exp_counts = r.counts()
            _, exp_ms = r.missingness()

        calls = float(V) * N
        with PgenContext(path, device=0) as c:
            c.counts()
            ts = []
            for _ in range(args.reps):
                t = time.perf_counter()
                got = c.counts()
                ts.append(time.perf_counter() - t)
            assert np.array_equal(got, exp_counts), "streamed counts differ from pgenlib"
            e2e_s = min(ts)
            c.load_resident()
            c.counts()
            ks, cs = [], []
            for _ in range(args.reps):
                got = c.counts()
                st = c.stats()
                ks.append(st.kernel_ms)
                cs.append(st.call_ms)
            assert np.array_equal(got, exp_counts), "resident counts differ from pgenlib"
            k, cm = float(np.median(ks)), float(np.median(cs))
            print(json.dumps({
                "op": "counts (plink_freq) on compressed records", "kernel_ms": k, "call_ms": cm,
                "record_GBps_kernel": size / (k * 1e-3) / 1e9, "Tcalls_per_s_kernel": calls / (k * 1e-3) / 1e12,
                "Tcalls_per_s_resident_call": calls / (cm * 1e-3) / 1e12, "Tcalls_per_s_e2e_from_file": calls / e2e_s / 1e12,
                "cpu_reference": {"Tcalls_per_s": calls / cpu_counts_s / 1e12, "threads": threads, "s": cpu_counts_s},
                "bit_exact_vs_pgenlib": True}), flush=True)
            c.missing_by_sample()
            ks, cs = [], []
            for _ in range(args.reps):
                ms = c.missing_by_sample()
                st = c.stats()
                ks.append(st.kernel_ms)
                cs.append(st.call_ms)
            assert np.array_equal(ms, exp_ms), "missing_by_sample differs from pgenlib"
            k, cm = float(np.median(ks)), float(np.median(cs))
            print(json.dumps({
                "op": "missing_by_sample (plink_missing sample mode) on compressed records", "kernel_ms": k, "call_ms": cm,
                "Tcalls_per_s_kernel": calls / (k * 1e-3) / 1e12, "Tcalls_per_s_resident_call": calls / (cm * 1e-3) / 1e12,
                "cpu_reference": {"Tcalls_per_s": calls / cpu_missing_s / 1e12, "threads": threads, "s": cpu_missing_s},
                "bit_exact_vs_pgenlib": True}), flush=True)
            # plink_score over every variant of the compressed file: records are expanded to dense rows
            # (K3b) batch by batch in front of the tensor-core kernel
            K = 16
            rng = np.random.default_rng(3)
            vid = np.arange(V, dtype=np.uint32)
            Wt = rng.standard_normal((V, K))
            c.score(vid, None, Wt, 0)
            ks, cs = [], []
            for _ in range(args.reps):
                c.score(vid, None, Wt, 0)
                st = c.stats()
                ks.append(st.kernel_ms)
                cs.append(st.call_ms)
            k, cm = float(np.median(ks)), float(np.median(cs))
            print(json.dumps({
                "op": f"plink_score K={K} on compressed records (resident)", "kernel_ms": k, "call_ms": cm,
                "Tcalls_per_s_kernel": calls / (k * 1e-3) / 1e12, "Tcalls_per_s_resident_call": calls / (cm * 1e-3) / 1e12}),
                flush=True)
    finally:
        if os.path.exists(path):
            os.remove(path)


if __name__ == "__main__":
    main()
