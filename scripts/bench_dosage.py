#!/usr/bin/env python
"""Dosage path at scale (development tool): plink_freq(dosage := true) = PgrGetDCounts, and
read_pgen(dosages := true) = PgrGetD + Dosage16ToDoublesMinus9, on a pgenlib-written fileset in which a
fraction of the samples of every variant carries an explicit dosage (list / bitarray / unconditional tracks
as the writer chooses).  GPU (records resident) vs the reference's pgenlib on ONE host thread (the harness
has no threaded scan for these two); every integer result and IMP_R2 compared bit for bit.  One JSON line
per operation."""
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
    ap.add_argument("--variants", type=int, default=4000)
    ap.add_argument("--samples", type=int, default=100_000)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()

    import synth
    from oracle import ref
    from plinking_duck_b200.pgen import PgenContext

    if not ref.available():
        raise SystemExit("oracle/_ref/libpgenref.so is needed to write the dosage fileset")
    V, N = args.variants, args.samples
    d = "/dev/shm" if os.path.isdir("/dev/shm") else "/tmp"
    path = os.path.join(d, f"plinkgpu_dosage_{V}x{N}.pgen")
    t0 = time.time()
    g, dos = synth.gen_dosage_variants(V, N, seed=31)
    ref.write_pgen(path, synth.to_genovec_words(g), N, dosage=dos)
    size = os.path.getsize(path)
    try:
        print(json.dumps({"op": "fileset", "variants": V, "samples": N, "file_bytes": size,
                          "bytes_per_variant": size / V, "build_s": round(time.time() - t0, 1)}), flush=True)
        with ref.RefPgen(path) as r:
            t = time.perf_counter()
            ec, ead, er2 = r.dcounts()
            cpu_counts_s = time.perf_counter() - t
            pick = np.arange(0, V, max(1, V // 200), dtype=np.uint32)
            t = time.perf_counter()
            ed = np.stack([r.dosage(int(v)) for v in pick])
            cpu_dosage_s = time.perf_counter() - t
        calls = float(V) * N
        with PgenContext(path, device=0) as c:
            c.load_resident()
            c.dosage_counts()
            ks, cs = [], []
            for _ in range(args.reps):
                gc, gad, gr2 = c.dosage_counts()
                st = c.stats()
                ks.append(st.kernel_ms + st.aux_kernel_ms)
                cs.append(st.call_ms)
            assert np.array_equal(gc, ec) and np.array_equal(gad, ead)
            ok = ~np.isnan(er2)
            assert np.array_equal(gr2[ok], er2[ok]) and np.array_equal(np.isnan(gr2), np.isnan(er2))
            k, cm = float(np.median(ks)), float(np.median(cs))
            print(json.dumps({
                "op": "dosage_counts (plink_freq dosage := true, PgrGetDCounts)", "kernel_ms": k, "call_ms": cm,
                "record_GBps_call": size / (cm * 1e-3) / 1e9, "Tcalls_per_s_call": calls / (cm * 1e-3) / 1e12,
                "cpu_reference": {"Tcalls_per_s": calls / cpu_counts_s / 1e12, "threads": 1, "s": cpu_counts_s},
                "bit_exact_vs_pgenlib": True}), flush=True)
            c.dosage_f64(pick)
            ts = []
            for _ in range(args.reps):
                t = time.perf_counter()
                gd = c.dosage_f64(pick)
                ts.append(time.perf_counter() - t)
            assert np.array_equal(gd, ed)
            pc = float(pick.size) * N
            print(json.dumps({
                "op": "dosage_f64 (read_pgen dosages := true, PgrGetD + Dosage16ToDoublesMinus9)", "variants": int(pick.size),
                "s_per_call": min(ts), "Tcalls_per_s_incl_D2H_of_8B_per_call": pc / min(ts) / 1e12,
                "cpu_reference": {"Tcalls_per_s": pc / cpu_dosage_s / 1e12, "threads": 1, "s": cpu_dosage_s},
                "bit_exact_vs_pgenlib": True}), flush=True)
    finally:
        if os.path.exists(path):
            os.remove(path)


if __name__ == "__main__":
    main()
