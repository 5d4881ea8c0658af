#!/usr/bin/env python
"""Regenerates tests/golden/ (run in the build container, where /root/reference exists).

1. Copies the reference's small binary .pgen test FIXTURES (test data, not source) from
   /root/reference/test/data into tests/golden/fixtures/ so that parity tests can run on the
   GPU box, where /root/reference does not exist.
2. Writes synthetic compressed fixtures with pgenlib's own writer (oracle/_ref) so that
   vrtypes 1/2/3/4/6 at sample counts around the 1/2/3-byte sample-id boundaries are covered.
3. Runs the reference's own pgenlib (oracle/_ref/libpgenref.so) over every fixture and stores
   the outputs (counts, int8 genotypes, missingness, dosages, scores, HWE ln p) in golden.npz.
4. Stores the known-answer values quoted from the reference's sqllogictests in kat.json.
"""
import json
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref  # noqa: E402
import synth  # noqa: E402

REF_DATA = "/root/reference/test/data"
FIX = os.path.join(HERE, "fixtures")
COPIED = [
    "pgen_example", "all_missing", "large_example", "streaming_example", "shard1", "shard2", "shard3",
    "rare_small", "pca_example", "dosage_example", "phased_example", "sexchr_example", "pgen_split",
]


def main():
    os.makedirs(FIX, exist_ok=True)
    for name in COPIED:
        for ext in (".pgen", ".pgen.pgi", ".psam", ".pvar"):
            src = os.path.join(REF_DATA, name + ext)
            if os.path.exists(src) and (ext.startswith(".pgen") or os.path.getsize(src) < 20000):
                shutil.copyfile(src, os.path.join(FIX, name + ext))
    # synthetic compressed fixtures (pgenlib's writer chooses the record types)
    synth_specs = {
        "syn_n255": dict(V=300, N=255, seed=11),
        "syn_n256": dict(V=300, N=256, seed=12),
        "syn_n257": dict(V=300, N=257, seed=13),
        "syn_n1000": dict(V=600, N=1000, seed=14),
        "syn_n70000": dict(V=40, N=70000, seed=15),
    }
    for name, sp in synth_specs.items():
        g = synth.gen_mixed_genotypes(sp["V"], sp["N"], sp["seed"])
        ref.write_pgen(os.path.join(FIX, name + ".pgen"), synth.to_genovec_words(g), sp["N"])
    # dosage fixture
    gd, dos = synth.gen_dosage_variants(40, 300, seed=21)
    ref.write_pgen(os.path.join(FIX, "syn_dosage.pgen"), synth.to_genovec_words(gd), 300, dosage=dos)

    out = {}
    names = sorted(f[:-5] for f in os.listdir(FIX) if f.endswith(".pgen"))
    rng = np.random.default_rng(20261104)
    meta = {}
    for name in names:
        path = os.path.join(FIX, name + ".pgen")
        with ref.RefPgen(path) as r:
            V, N = r.V, r.N
            vt = r.vrtypes()
            out[name + "/vrtypes"] = vt
            out[name + "/fpos"] = r.fpos()
            out[name + "/counts"] = r.counts()
            if V * N <= 4_000_000:
                out[name + "/i8"] = r.decode_i8()
            pv, ps = r.missingness()
            out[name + "/missing_by_variant"] = pv
            out[name + "/missing_by_sample"] = ps
            het, hom_alt, miss = r.sample_counts(sparse=False)
            h2, a2, m2 = r.sample_counts(sparse=True)
            assert (het == h2).all() and (hom_alt == a2).all() and (miss == m2).all()
            out[name + "/sample_counts"] = np.stack([het, hom_alt, miss])
            c = out[name + "/counts"]
            out[name + "/hwe_lnp"] = ref.hwe_lnp_batch(c, False)
            out[name + "/hwe_lnp_midp"] = ref.hwe_lnp_batch(c, True)
            has_unsupported = bool(((vt & 0x88) != 0).any())
            if not has_unsupported:
                dc, ad, r2 = r.dcounts()
                out[name + "/dcounts"] = dc
                out[name + "/all_dosages"] = ad
                out[name + "/imp_r2"] = r2
                nd = min(V, 64)
                out[name + "/dosage_first"] = np.stack([r.dosage(v) for v in range(nd)])
                # score: all variants, random weights, 10% flips, three modes
                w = rng.standard_normal(V)
                w[rng.random(V) < 0.05] = 0.0
                fl = (rng.random(V) < 0.1).astype(np.uint8)
                vid = np.arange(V, dtype=np.uint32)
                out[name + "/score_w"] = w
                out[name + "/score_flip"] = fl
                for mode in (0, 1, 2):
                    s, d, a = r.score(vid, fl, w, mode)
                    out[name + f"/score_m{mode}_sum"] = s
                    out[name + f"/score_m{mode}_dsum"] = d
                    out[name + f"/score_m{mode}_ct"] = a
            # subset goldens: deterministic ~40% subset (and a single-sample subset)
            if N >= 4:
                sub = np.sort(rng.choice(N, size=max(1, (N * 2) // 5), replace=False)).astype(np.uint32)
                r.set_subset(sub)
                out[name + "/subset"] = sub
                out[name + "/subset_counts"] = r.counts()
                if V * sub.size <= 2_000_000:
                    out[name + "/subset_i8"] = r.decode_i8()
                pv, ps = r.missingness()
                out[name + "/subset_missing_by_sample"] = ps
                r.set_subset(None)
            meta[name] = dict(V=int(V), N=int(N), vrtypes={int(k): int(v) for k, v in zip(*np.unique(vt, return_counts=True))})
    np.savez_compressed(os.path.join(HERE, "golden.npz"), **out)
    with open(os.path.join(HERE, "fixtures.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print(json.dumps(meta, indent=1, sort_keys=True))


if __name__ == "__main__":
    main()
