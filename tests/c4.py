"""C4 = BASELINE.json configs[3] fileset: a difflist / LD-compressed .pgen as plink2 itself writes
it.  Genotypes come from the MAF model of the reference's rare-variant fixture (test/data/generate_rare_test_data.sh:
32-46: 95 % log-uniform [1/2N, 1 %], 5 % log-uniform [1 %, 50 %], 1 % missing), generated on the GPU; on top of it,
as in real data, runs of variants in strong LD (near-copies of the previous variant, some with the alleles swapped),
ALT-major variants and a few two-category rows -- so pgenlib's writer (oracle/_ref, SpgwAppendBiallelicGenovec)
emits every main-track record type: 0 raw, 1 one-bit, 2 / 3 LD-compressed, 4 / 6 / 7 difflists.  Test infrastructure."""
from __future__ import annotations

import numpy as np


def make_c4_file(path: str, V: int, N: int, seed: int, device, plain: bool = False) -> np.ndarray:
    """Writes the file; returns the genovec words (V, ceil(N/32)) uint64 that went in.  plain = True: the MAF model
    alone (no LD runs, ALT-major or two-category rows): the round-1 rare-variant fileset, vrtypes 0 / 1 / 4."""
    import torch

    from oracle import ref

    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    W64 = (N + 31) // 32
    rows = np.zeros((V, W64), dtype=np.uint64)
    chunk = max(16, ((192 << 20) // N) // 16 * 16)
    for v0 in range(0, V, chunk):
        vt = min(chunk, V - v0)
        common = torch.rand((vt, 1), device=device, generator=gen) >= 0.95
        u = torch.rand((vt, 1), device=device, generator=gen)
        lo = torch.where(common, torch.full_like(u, np.log(0.01)), torch.full_like(u, np.log(1.0 / (2 * N))))
        hi = torch.where(common, torch.full_like(u, np.log(0.5)), torch.full_like(u, np.log(0.01)))
        maf = torch.exp(lo + u * (hi - lo))
        x = torch.rand((vt, W64 * 32), device=device, generator=gen)
        p2 = maf * maf
        code = (x < p2).to(torch.int64) * 2 + ((x >= p2) & (x < p2 + 2 * maf * (1 - maf))).to(torch.int64)
        miss = torch.rand((vt, W64 * 32), device=device, generator=gen) < 0.01
        code = torch.where(miss, torch.full_like(code, 3), code)
        r = torch.arange(vt, device=device)
        # common variants in LD blocks: rows 16k+1 .. 16k+4 copy their predecessor with ~N/1500 calls changed;
        # row 16k+4 additionally swaps the alleles (0 <-> 2), which pgenlib stores as an inverted LD record (type 3)
        base = r[(r % 16 == 0)] if not plain else r[:0]
        if base.numel():
            pb = 0.2 + 0.3 * torch.rand((base.numel(), 1), device=device, generator=gen)
            xb = torch.rand((base.numel(), W64 * 32), device=device, generator=gen)
            cb = (xb < pb * pb).to(torch.int64) * 2 + ((xb >= pb * pb) & (xb < pb * pb + 2 * pb * (1 - pb))).to(torch.int64)
            code[base] = torch.where(miss[base], torch.full_like(cb, 3), cb)
        for k in (1, 2, 3, 4):
            idx = r[(r % 16 == k)] if not plain else r[:0]
            if not idx.numel():
                continue
            prev = code[idx - 1]
            if k == 4:
                prev = torch.where(prev == 0, torch.full_like(prev, 2), torch.where(prev == 2, torch.zeros_like(prev), prev))
            flip = torch.rand(prev.shape, device=device, generator=gen) < (1.0 / 1500.0)
            newv = torch.randint(0, 4, prev.shape, device=device, generator=gen)
            code[idx] = torch.where(flip, newv, prev)
        # ALT-major rare variants (difflist against all hom-alt, type 6)
        idx = r[(r % 16 == 9)] if not plain else r[:0]
        if idx.numel():
            c9 = code[idx]
            code[idx] = torch.where(c9 == 3, c9, 2 - c9)
        # two dominant categories (one-bit records, type 1)
        idx = r[(r % 16 == 12)] if not plain else r[:0]
        if idx.numel():
            two = (torch.rand((idx.numel(), W64 * 32), device=device, generator=gen) < 0.4).to(torch.int64)
            code[idx] = torch.where(miss[idx], torch.full_like(two, 3), two)
        code[:, N:] = 0
        sh = (torch.arange(32, device=device, dtype=torch.int64) * 2).view(1, 1, 32)
        words = (code.view(vt, W64, 32) << sh).sum(dim=2)  # < 2^64 as two's complement int64
        rows[v0:v0 + vt] = words.cpu().numpy().view(np.uint64)
        del x, code, miss, words
    ref.write_pgen(path, rows, N)
    return rows


def vrtype_histogram(path: str) -> dict:
    from oracle import ref

    with ref.RefPgen(path) as r:
        vt = np.asarray(r.vrtypes()) & 7
    return {int(k): int((vt == k).sum()) for k in np.unique(vt)}
