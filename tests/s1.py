"""S1 = BASELINE.json configs[0] fileset: V = 10 000 variants x N = 1 000 samples, hardcalls,
p_v ~ U(0.05, 0.5), g ~ Binomial(2, p_v), 1 % missing, seed 20261104; chromosomes 1, 2, X (PAR and non-PAR
positions), Y and MT; a SEX column with males, females and a few unknowns.  Written both as a fixed-width mode 0x02
file and through pgenlib's own writer (mode 0x10, mixed record types).  Test infrastructure.

Also the numpy restatement of the extension-side arithmetic that is not in pgenlib: ComputeSexAwareCounts
(src/plink_common.cpp:1996-2108) and the plink_freq derivations (src/plink_freq.cpp:487-544)."""
from __future__ import annotations

import os

import numpy as np

import synth

V, N, SEED = 10_000, 1_000, 20261104


def layout():
    """(chrom[V], pos[V], sex[N]): 1 -> male, 2 -> female, 0 -> unknown."""
    chrom = np.array(["1"] * 5000 + ["2"] * 4400 + ["X"] * 400 + ["Y"] * 100 + ["MT"] * 100)
    pos = np.arange(1, V + 1, dtype=np.int64) * 1000
    # chrX: the first 60 inside GRCh38 PAR1 (10 001 - 2 781 479), the rest non-PAR, the last 20 inside PAR2
    x0 = 9400
    pos[x0:x0 + 60] = 20_000 + 40_000 * np.arange(60)
    pos[x0 + 60:x0 + 380] = 3_000_000 + 400_000 * np.arange(320)
    pos[x0 + 380:x0 + 400] = 155_701_383 + 10_000 * np.arange(20)
    rng = np.random.default_rng(SEED + 1)
    sex = rng.choice([1, 2, 0], size=N, p=[0.47, 0.48, 0.05]).astype(np.uint8)
    return chrom, pos, sex


def genotypes() -> np.ndarray:
    g = synth.gen_hwe_genotypes(V, N, seed=SEED, missing=0.01, maf_lo=0.05, maf_hi=0.5)
    g[17] = 3          # an all-missing variant (NULL frequency, test/sql/plink_freq.test:213-229)
    g[9405] = 3        # ... one on chrX
    g[4242] = 0        # monomorphic
    return g


def write(dirpath: str, g: np.ndarray, with_pgenlib: bool):
    """Returns {'fixed': path[, 'pgenlib': path]}; also writes .psam / .pvar beside the first (documentation)."""
    chrom, pos, sex = layout()
    out = {"fixed": os.path.join(dirpath, "s1_fixed.pgen")}
    synth.write_fixed_pgen(out["fixed"], g)
    with open(os.path.join(dirpath, "s1_fixed.psam"), "w") as f:
        f.write("#IID\tSEX\n")
        for i, s in enumerate(sex):
            f.write(f"S{i}\t{'NA' if s == 0 else int(s)}\n")
    with open(os.path.join(dirpath, "s1_fixed.pvar"), "w") as f:
        f.write("#CHROM\tPOS\tID\tREF\tALT\n")
        for v in range(V):
            f.write(f"{chrom[v]}\t{int(pos[v])}\tv{v}\tA\tG\n")
    if with_pgenlib:
        from oracle import ref

        out["pgenlib"] = os.path.join(dirpath, "s1_pgenlib.pgen")
        ref.write_pgen(out["pgenlib"], synth.to_genovec_words(g), N)
    return out


def sex_aware_counts(i8: np.ndarray, ploidy: int, sex: np.ndarray):
    """ComputeSexAwareCounts for one variant's int8 calls.  ploidy: 1 chrX, 2 chrY, 3 chrMT (PD_PLOIDY_*).
    Returns dict(obs_allele_ct, alt_allele_ct, hom_ref, het, hom_alt, missing)."""
    g = i8.astype(np.int64)
    male, female = sex == 1, sex == 2
    if ploidy == 3:  # MT: haploid for everyone, het -> missing
        hr, ha = int((g == 0).sum()), int((g == 2).sum())
        return dict(obs=hr + ha, alt=ha, hom_ref=hr, het=0, hom_alt=ha, missing=int(g.size - hr - ha))
    if ploidy == 2:  # Y: males only
        hr, ha = int(((g == 0) & male).sum()), int(((g == 2) & male).sum())
        return dict(obs=hr + ha, alt=ha, hom_ref=hr, het=0, hom_alt=ha, missing=int(g.size - hr - ha))
    # X non-PAR: females diploid, males haploid (het -> missing), unknown sex excluded
    f0, f1, f2 = (int(((g == k) & female).sum()) for k in (0, 1, 2))
    m0, m2 = int(((g == 0) & male).sum()), int(((g == 2) & male).sum())
    return dict(obs=2 * (f0 + f1 + f2) + m0 + m2, alt=f1 + 2 * f2 + m2, hom_ref=f0 + m0, het=f1, hom_alt=f2 + m2,
                missing=int(g.size - f0 - f1 - f2 - m0 - m2))


def expected_freq_rows(counts: np.ndarray, i8: np.ndarray, ploidy: np.ndarray, sex: np.ndarray, dtype):
    """plink_freq rows (src/plink_freq.cpp:487-544) from pgenlib's counts (autosomal / PAR) and int8 calls
    (sex chromosomes)."""
    out = np.zeros(counts.shape[0], dtype=dtype)
    for v in range(counts.shape[0]):
        c = counts[v]
        if ploidy[v] == 0:
            obs = int(c[0]) + int(c[1]) + int(c[2])
            out["HOM_REF_CT"][v], out["HET_CT"][v], out["HOM_ALT_CT"][v], out["MISSING_CT"][v] = c
            if obs == 0:
                out["freq_is_null"][v] = 1
            else:
                out["OBS_CT"][v] = 2 * obs
                out["ALT_FREQ"][v] = (float(c[1]) + 2.0 * float(c[2])) / (2.0 * float(obs))
        else:
            s = sex_aware_counts(i8[v], int(ploidy[v]), sex)
            out["HOM_REF_CT"][v], out["HET_CT"][v], out["HOM_ALT_CT"][v], out["MISSING_CT"][v] = (
                s["hom_ref"], s["het"], s["hom_alt"], s["missing"])
            if s["obs"] == 0:
                out["freq_is_null"][v] = 1
            else:
                out["OBS_CT"][v] = s["obs"]
                out["ALT_FREQ"][v] = float(s["alt"]) / float(s["obs"])
    return out
