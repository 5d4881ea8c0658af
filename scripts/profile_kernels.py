#!/usr/bin/env python
"""Per-kernel throughput on device-resident records (development tool; bench.py is the contract).

Generates a fixed-width .pgen image straight into HBM and times every entry point with the library's
own CUDA-event stats: kernel_ms (scan kernels only) and call_ms (whole call incl. result copies).
Prints one JSON line per operation with achieved GB/s of algorithmic bytes against MEASURED_PEAKS.json.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants", type=int, default=200_000)
    ap.add_argument("--samples", type=int, default=100_000)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--ops", default="counts,counts_subset,hardy,missing_by_sample,sample_counts,decode_i8,decode_i8_valid,decode_phased,score1,score16")
    ap.add_argument("--score-variants", type=int, default=20_000)
    ap.add_argument("--decode-variants", type=int, default=20_000)
    ap.add_argument("--cpu-variants", type=int, default=0,
                    help="also time the reference's CPU loops (oracle/_ref, all host threads) for counts / hardy / "
                         "missingness / decode on the first this-many variants of the same data (tmpfs file)")
    args = ap.parse_args()

    import torch

    import bench
    from plinking_duck_b200.pgen import PgenContext

    dev = torch.device("cuda", 0)
    V, N = args.variants, args.samples
    W = (N + 3) // 4
    nbytes = 12 + V * W
    img = torch.empty(nbytes + 256, dtype=torch.uint8, device=dev)
    bench.generate_image_gpu(V, N, 1234, dev, img)
    peak, _ = bench.peaks()
    ctx = PgenContext.from_device_pointer(img.data_ptr(), nbytes, device=0)
    ops = args.ops.split(",")
    rng = np.random.default_rng(5)

    def report(name, fn, alg_bytes, calls, extra=None):
        fn()  # warm-up
        ks, cs = [], []
        for _ in range(args.reps):
            fn()
            st = ctx.stats()
            ks.append(st.kernel_ms)
            cs.append(st.call_ms)
        k, c = float(np.median(ks)), float(np.median(cs))
        line = {"op": name, "kernel_ms": k, "call_ms": c, "alg_GBps_kernel": alg_bytes / (k * 1e-3) / 1e9 if k else None,
                "frac_of_peak": alg_bytes / (k * 1e-3) / 1e9 / peak if k else None,
                "Tcalls_per_s_kernel": calls / (k * 1e-3) / 1e12 if k else None,
                "Tcalls_per_s_call": calls / (c * 1e-3) / 1e12 if c else None}
        if extra:
            line.update(extra)
        print(json.dumps(line), flush=True)

    if "counts" in ops:
        report("counts", lambda: ctx.counts(), V * W, V * N)
    if "counts_subset" in ops:
        ctx.set_subset(np.arange(0, N, 2))
        report("counts_subset50", lambda: ctx.counts(), V * W, V * N)
        ctx.set_subset(None)
    if "hardy" in ops:
        report("hardy", lambda: ctx.hardy(), V * W, V * N)
    if "hwe_only" in ops:
        cnts = ctx.counts()
        def hw():
            ctx.hwe_from_counts(cnts)
        hw()
        a = []
        for _ in range(args.reps):
            hw()
            a.append(ctx.stats().aux_kernel_ms)
        print(json.dumps({"op": "hwe_only", "variants": V, "samples": N, "kernel_ms": float(np.median(a))}), flush=True)
    if "missing_by_sample" in ops:
        report("missing_by_sample", lambda: ctx.missing_by_sample(), V * W, V * N)
    if "sample_counts" in ops:
        report("sample_counts", lambda: ctx.sample_counts(), V * W, V * N)
    if "decode_i8" in ops:
        Vd = min(V, args.decode_variants)
        out = torch.empty((Vd, N), dtype=torch.int8, pin_memory=True).numpy()
        import ctypes as C

        def dec():
            ctx._check(ctx._lib.pd_decode_i8(ctx._h, 0, Vd, out.ctypes.data_as(C.c_void_p)))

        report("decode_i8", dec, Vd * (W + N), Vd * N, {"note": "kernel_ms = expand kernels; call_ms includes D2H of N bytes/variant"})
    if "decode_i8_valid" in ops:
        Vd = min(V, args.decode_variants)
        outv = torch.empty((Vd, N), dtype=torch.int8, pin_memory=True).numpy()
        val = torch.empty(((Vd * N + 63) // 64,), dtype=torch.int64, pin_memory=True).numpy().view(np.uint64)
        import ctypes as C

        def decv():
            ctx._check(ctx._lib.pd_decode_i8_valid(ctx._h, 0, Vd, 7, outv.ctypes.data_as(C.c_void_p), val.ctypes.data_as(C.c_void_p)))

        report("decode_i8_valid", decv, Vd * (W + N + N // 8), Vd * N,
               {"note": "kernel_ms = expand kernels (calls + validity bits); call_ms includes D2H of 1.125 N bytes/variant"})
        plain = ctx.decode_i8(0, min(Vd, 2000))
        bits = np.unpackbits(val.view(np.uint8), bitorder="little")[: min(Vd, 2000) * N].reshape(-1, N)
        assert np.array_equal(bits.astype(bool), plain != -9) and np.array_equal(outv[: plain.shape[0]], np.where(plain == -9, 0, plain))
    if "decode_phased" in ops:
        Vd = min(V, args.decode_variants // 2)
        outp = torch.empty((Vd, N, 2), dtype=torch.int8, pin_memory=True).numpy()
        import ctypes as C

        def decp():
            ctx._check(ctx._lib.pd_decode_phased(ctx._h, 0, Vd, outp.ctypes.data_as(C.c_void_p)))

        report("decode_phased", decp, Vd * (W + 2 * N), Vd * N,
               {"note": "kernel_ms = expand kernels; call_ms includes D2H of 2N bytes/variant"})
    for K in (1, 16):
        if f"score{K}" in ops:
            Vs = min(V, args.score_variants)
            vid = np.arange(Vs, dtype=np.uint32)
            Wt = rng.standard_normal((Vs, K))
            fl = (rng.random(Vs) < 0.1).astype(np.uint8)
            report(f"score_K{K}", lambda: ctx.score(vid, fl, Wt, 0), Vs * W, Vs * N,
                   {"flops": 2.0 * Vs * N * K, "note": "TFLOP/s = flops / kernel_ms"})
    ctx.close()
    if args.cpu_variants:
        # the reference's loops beside the kernels: pgenlib readers, one per host thread,
        # 128-variant claims (oracle/ref_harness.cc: ref_bench_scan)
        from oracle import ref

        if ref.available():
            Vc = min(V, args.cpu_variants)
            d = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
            path = os.path.join(d, f"plinkgpu_table_{Vc}x{N}.pgen")
            host = img[: 12 + Vc * W].cpu().numpy().copy()
            host[3:7] = np.frombuffer(np.uint32(Vc).tobytes(), dtype=np.uint8)  # variant count of the slice
            host.tofile(path)
            threads = os.cpu_count() or 1
            try:
                for kind, name in ((0, "counts"), (1, "hardy"), (3, "missing_by_sample"), (4, "decode_i8")):
                    ref.bench_scan(path, threads, 0, min(Vc, 2048), kind)
                    t = min(ref.bench_scan(path, threads, 0, Vc, kind)[0] for _ in range(3))
                    print(json.dumps({"op": name, "cpu_reference": "pgenlib, %d threads, %d variants on tmpfs" % (threads, Vc),
                                      "s": t, "Tcalls_per_s": Vc * N / t / 1e12}), flush=True)
            finally:
                os.unlink(path)


if __name__ == "__main__":
    main()
