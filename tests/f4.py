"""Fixtures for the record tracks beyond biallelic hardcalls, written with pgenlib's own writer
(oracle/_ref): multiallelic variants (aux1a / aux1b patch sets, pgen_spec.tex:469-564) and dosage-phase tracks
(:632-676).  The extension reads such files through the biallelic calls (PgrGet, PgrGetCounts, PgrGetMissingness,
PgrGetD), so parity is defined by what those return.  Test infrastructure."""
from __future__ import annotations

import numpy as np

import synth


def _bits(mask: np.ndarray) -> np.ndarray:
    n = mask.size
    w = np.zeros((n + 63) // 64 + 1, dtype=np.uint64)
    idx = np.nonzero(mask)[0]
    np.bitwise_or.at(w, idx >> 6, np.uint64(1) << (idx & 63).astype(np.uint64))
    return w


def write_multiallelic(path: str, V: int, N: int, seed: int, rare_every: int = 0):
    """Returns the (V, N) main-track codes that went in (0 REF/REF, 1 REF/ALTx, 2 ALTx/ALTy, 3 missing).
    rare_every = k > 0: every k-th variant has rare ALT2+ alleles, so that pgenlib's writer stores its patch sets as
    sample-id lists (aux1 mode 1) instead of bit arrays (mode 0)."""
    from oracle import ref

    rng = np.random.default_rng(seed)
    allele_cts = rng.choice([2, 2, 3, 4, 5, 7, 18, 40], size=V).astype(np.uint32)
    allele_cts[0] = 2
    allele_cts[1] = 3
    w = ref.RefWriter(path, V, N, 0, allele_cts)
    codes = np.zeros((V, N), dtype=np.int8)
    for v in range(V):
        k = int(allele_cts[v])
        freqs = rng.dirichlet(np.ones(k) * 0.7)
        freqs[0] += 1.0   # REF-major
        if rare_every and v % rare_every == 0 and k > 2:
            freqs[1] += 1.0
            freqs[2:] *= 0.01  # a handful of ALT2+ carriers among many REF/ALT1 and ALT1/ALT1 calls
        freqs /= freqs.sum()
        a = rng.choice(k, size=N, p=freqs)
        b = rng.choice(k, size=N, p=freqs)
        lo, hi = np.minimum(a, b), np.maximum(a, b)
        miss = rng.random(N) < 0.03
        code = np.where(hi == 0, 0, np.where(lo == 0, 1, 2)).astype(np.int8)
        code[miss] = 3
        codes[v] = code
        gv = synth.to_genovec_words(code[None, :])[0]
        if k == 2:
            w.append(gv)
            continue
        s01 = (code == 1) & (hi != 1)
        s10 = (code == 2) & ~((lo == 1) & (hi == 1))
        vals10 = np.stack([lo[s10], hi[s10]], axis=1)
        w.append_multiallelic(gv, _bits(s01), hi[s01], _bits(s10), vals10, k)
    w.close()
    return codes, allele_cts


def write_multiallelic_phased(path: str, V: int, N: int, seed: int):
    """Multiallelic AND biallelic variants, most of them with a hardcall-phase track (vrtype bits 3 + 4 together): the
    phase track counts the ALTx/ALTy (x != y) calls among its hets.  Returns (allele pairs (V, N, 2) with 255 = missing
    in file order lo <= hi, phasepresent, phaseinfo, allele counts)."""
    from oracle import ref

    rng = np.random.default_rng(seed)
    allele_cts = rng.choice([2, 3, 3, 4, 5, 7, 18, 40], size=V).astype(np.uint32)
    allele_cts[0] = 2
    allele_cts[1] = 3
    w = ref.RefWriter(path, V, N, 2, allele_cts)
    pairs = np.zeros((V, N, 2), dtype=np.uint8)
    pps = np.zeros((V, N), dtype=bool)
    pis = np.zeros((V, N), dtype=bool)
    for v in range(V):
        k = int(allele_cts[v])
        freqs = rng.dirichlet(np.ones(k) * 0.7)
        freqs[0] += 0.6
        if v % 3 == 0 and k > 2:
            freqs[1] += 1.0
            freqs[2:] *= 0.02   # patch sets as id lists
        freqs /= freqs.sum()
        a = rng.choice(k, size=N, p=freqs)
        b = rng.choice(k, size=N, p=freqs)
        lo, hi = np.minimum(a, b), np.maximum(a, b)
        miss = rng.random(N) < 0.03
        code = np.where(hi == 0, 0, np.where(lo == 0, 1, 2)).astype(np.int8)
        code[miss] = 3
        pairs[v, :, 0] = np.where(miss, 255, lo)
        pairs[v, :, 1] = np.where(miss, 255, hi)
        het = ~miss & (lo != hi)
        mode = v % 4
        if mode == 0:      # no phase track at all
            pp = np.zeros(N, dtype=bool)
        elif mode == 1:    # every het phased (aux2 without an explicit phasepresent array)
            pp = het.copy()
        else:
            pp = het & (rng.random(N) < 0.6)
        pi = pp & (rng.random(N) < 0.5)
        pps[v], pis[v] = pp, pi
        gv = synth.to_genovec_words(code[None, :])[0]
        s01 = (code == 1) & (hi != 1)
        s10 = (code == 2) & ~((lo == 1) & (hi == 1))
        vals10 = np.stack([lo[s10], hi[s10]], axis=1)
        if k == 2:
            if pp.any():
                w.append_hphase(gv, _bits(pp), _bits(pi))
            else:
                w.append(gv)
        elif pp.any():
            w.append_multiallelic_hphase(gv, _bits(s01), hi[s01], _bits(s10), vals10, _bits(pp), _bits(pi), k)
        else:
            w.append_multiallelic(gv, _bits(s01), hi[s01], _bits(s10), vals10, k)
    w.close()
    return pairs, pps, pis, allele_cts


def write_dphase(path: str, V: int, N: int, seed: int):
    """Hardcalls + dosages, a third of the variants with an explicit dosage-phase track.  Returns the hardcall
    codes."""
    from oracle import ref

    rng = np.random.default_rng(seed)
    g = synth.gen_hwe_genotypes(V, N, seed=seed, missing=0.03)
    w = ref.RefWriter(path, V, N, 4, None)
    for v in range(V):
        code = g[v]
        gv = synth.to_genovec_words(code[None, :])[0]
        kind = v % 3
        if kind == 0:
            w.append(gv)
            continue
        het = code == 1
        phasepresent = het & (rng.random(N) < 0.6)
        phaseinfo = phasepresent & (rng.random(N) < 0.5)
        dosage_present = (code != 3) & (rng.random(N) < 0.2)
        base = np.array([0, 16384, 32768, 0], dtype=np.int64)[code]
        dos = np.clip(base + rng.integers(-5000, 5000, size=N), 0, 32768)
        dm = dos[dosage_present].astype(np.uint16)
        if kind == 1:
            dphase_present = np.zeros(N, dtype=bool)
            dd = np.zeros(0, dtype=np.int16)
        else:
            dphase_present = dosage_present & (rng.random(N) < 0.5)
            dv = dos[dphase_present]
            lim = np.minimum(dv, 32768 - dv)      # |delta| <= min(d, 2 - d)
            dd = (rng.integers(-1, 2, size=dv.size) * (lim // 2)).astype(np.int16)
            dd[dd == 0] = 1
            dd = np.where(np.abs(dd) > lim, 0, dd).astype(np.int16)
            keep = dd != 0
            idx = np.nonzero(dphase_present)[0]
            dphase_present[idx[~keep]] = False
            dd = dd[keep]
        w.append_dphase(gv, _bits(phasepresent), _bits(phaseinfo), _bits(dosage_present), _bits(dphase_present), dm, dd)
    w.close()
    return g
