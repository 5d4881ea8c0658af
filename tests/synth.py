"""Seeded synthetic genotype data and .pgen writers used by tests and bench (test infrastructure).

Genotype codes follow the .pgen 2-bit encoding: 0 hom-ref, 1 het, 2 hom-alt, 3 missing.
"""
from __future__ import annotations

import struct

import numpy as np


def gen_hwe_genotypes(V: int, N: int, seed: int, missing: float = 0.01, maf_lo=0.05, maf_hi=0.5) -> np.ndarray:
    """SURVEY 8d model: p_v ~ U(maf_lo, maf_hi), g ~ Binomial(2, p_v), `missing` fraction set to 3."""
    rng = np.random.default_rng(seed)
    p = rng.uniform(maf_lo, maf_hi, size=(V, 1))
    u = rng.random((V, N))
    g = (u < p * p).astype(np.int8) * 2
    g[(u >= p * p) & (u < p * p + 2 * p * (1 - p))] = 1
    if missing > 0:
        g[rng.random((V, N)) < missing] = 3
    return g


def gen_mixed_genotypes(V: int, N: int, seed: int) -> np.ndarray:
    """A mixture that makes pgenlib's writer emit every main-track record type:
    dense (0), two-category dominated (1-bit), near-copies of the previous variant (LD, 2),
    inverted near-copies (3), rare variants (4), ALT-major (6), mostly-missing (7), monomorphic."""
    rng = np.random.default_rng(seed)
    g = np.zeros((V, N), dtype=np.int8)
    prev = None
    for v in range(V):
        kind = rng.integers(0, 10)
        if kind <= 1 or prev is None:
            p = rng.uniform(0.1, 0.5)
            u = rng.random(N)
            row = np.where(u < p * p, 2, np.where(u < p * p + 2 * p * (1 - p), 1, 0)).astype(np.int8)
            row[rng.random(N) < 0.02] = 3
        elif kind == 2:  # two dominant categories + a few others -> 1-bit
            a, b = rng.choice(4, size=2, replace=False)
            row = np.where(rng.random(N) < rng.uniform(0.2, 0.5), a, b).astype(np.int8)
            k = rng.integers(0, max(1, N // 40) + 1)
            if k:
                idx = rng.choice(N, size=k, replace=False)
                row[idx] = rng.integers(0, 4, size=k)
        elif kind == 3 or kind == 4:  # LD: previous variant with a few changes
            row = prev.copy()
            k = rng.integers(0, max(1, N // 30) + 1)
            if k:
                idx = rng.choice(N, size=k, replace=False)
                row[idx] = rng.integers(0, 4, size=k)
        elif kind == 5:  # inverted LD
            row = prev.copy()
            swap0, swap2 = row == 0, row == 2
            row[swap0] = 2
            row[swap2] = 0
            k = rng.integers(0, max(1, N // 50) + 1)
            if k:
                idx = rng.choice(N, size=k, replace=False)
                row[idx] = rng.integers(0, 4, size=k)
        elif kind == 6 or kind == 7:  # rare variant
            row = np.zeros(N, dtype=np.int8)
            k = rng.integers(0, max(1, N // 20) + 1)
            if k:
                idx = rng.choice(N, size=k, replace=False)
                row[idx] = rng.choice([1, 1, 1, 2, 3], size=k)
        elif kind == 8:  # ALT-major
            row = np.full(N, 2, dtype=np.int8)
            k = rng.integers(0, max(1, N // 20) + 1)
            if k:
                idx = rng.choice(N, size=k, replace=False)
                row[idx] = rng.choice([0, 1, 3], size=k)
        else:  # mostly missing
            row = np.full(N, 3, dtype=np.int8)
            k = rng.integers(0, max(1, N // 20) + 1)
            if k:
                idx = rng.choice(N, size=k, replace=False)
                row[idx] = rng.choice([0, 1, 2], size=k)
        g[v] = row
        prev = row
    return g


def gen_rare_genotypes(V: int, N: int, seed: int) -> np.ndarray:
    """MAF model of the reference's rare-variant fixture (test/data/generate_rare_test_data.sh:32-46):
    95 % log-uniform [1/2N, 1 %], 5 % log-uniform [1 %, 50 %], 1 % missing."""
    rng = np.random.default_rng(seed)
    g = np.zeros((V, N), dtype=np.int8)
    for v in range(V):
        if rng.random() < 0.95:
            maf = np.exp(rng.uniform(np.log(1.0 / (2 * N)), np.log(0.01)))
        else:
            maf = np.exp(rng.uniform(np.log(0.01), np.log(0.5)))
        u = rng.random(N)
        row = np.where(u < maf * maf, 2, np.where(u < maf * maf + 2 * maf * (1 - maf), 1, 0)).astype(np.int8)
        row[rng.random(N) < 0.01] = 3
        g[v] = row
    return g


def pack_2bit_bytes(g: np.ndarray) -> np.ndarray:
    """(V, N) codes -> (V, ceil(N/4)) uint8, sample 0 in the low bits, trailing bits zero."""
    V, N = g.shape
    nb = (N + 3) // 4
    pad = np.zeros((V, nb * 4), dtype=np.uint8)
    pad[:, :N] = g.astype(np.uint8) & 3
    q = pad.reshape(V, nb, 4)
    return (q[:, :, 0] | (q[:, :, 1] << 2) | (q[:, :, 2] << 4) | (q[:, :, 3] << 6)).astype(np.uint8)


def to_genovec_words(g: np.ndarray) -> np.ndarray:
    """(V, N) codes -> (V, ceil(N/32)) uint64 genovecs."""
    V, N = g.shape
    b = pack_2bit_bytes(g)
    nw = (N + 31) // 32
    buf = np.zeros((V, nw * 8), dtype=np.uint8)
    buf[:, : b.shape[1]] = b
    return buf.view(np.uint64).reshape(V, nw)


def fixed_header(V: int, N: int) -> bytes:
    """Mode 0x02 header: magic, mode, V, N (u32 LE), header_ctrl 0x40 (pgenlib_read.cc:877-911)."""
    return bytes([0x6C, 0x1B, 0x02]) + int(V).to_bytes(4, "little") + int(N).to_bytes(4, "little") + bytes([0x40])


def fixed_pgen_image(g: np.ndarray) -> np.ndarray:
    V, N = g.shape
    body = pack_2bit_bytes(g).reshape(-1)
    return np.concatenate([np.frombuffer(fixed_header(V, N), dtype=np.uint8), body])


def write_fixed_pgen(path: str, g: np.ndarray):
    fixed_pgen_image(g).tofile(path)


def counts_numpy(g: np.ndarray) -> np.ndarray:
    return np.stack([(g == k).sum(axis=1) for k in range(4)], axis=1).astype(np.uint32)


def gen_dosage_variants(V: int, N: int, seed: int):
    """Hardcalls + per-variant dosage lists in the form oracle.ref.write_pgen takes.
    Dosages are consistent with the hardcalls (|dosage - hardcall| < 0.5)."""
    rng = np.random.default_rng(seed)
    g = np.zeros((V, N), dtype=np.int8)
    dos = []
    for v in range(V):
        frac = [0.0, 0.02, 0.5, 1.0][v % 4]
        has = rng.random(N) < frac if frac < 1.0 else np.ones(N, dtype=bool)
        d16 = rng.integers(0, 32769, size=N)
        # keep away from the rounding boundaries 8192 / 24576
        d16 = np.where(np.abs(d16 - 8192) < 40, 8000, d16)
        d16 = np.where(np.abs(d16 - 24576) < 40, 24700, d16)
        hard = np.where(d16 < 8192, 0, np.where(d16 < 24576, 1, 2)).astype(np.int8)
        p = rng.uniform(0.1, 0.5)
        u = rng.random(N)
        plain = np.where(u < p * p, 2, np.where(u < p * p + 2 * p * (1 - p), 1, 0)).astype(np.int8)
        row = np.where(has, hard, plain)
        miss = (rng.random(N) < 0.03) & ~has
        row[miss] = 3
        g[v] = row
        if has.any():
            words = np.zeros((N + 63) // 64, dtype=np.uint64)
            idx = np.nonzero(has)[0]
            np.bitwise_or.at(words, idx >> 6, np.uint64(1) << (idx & 63).astype(np.uint64))
            dos.append((words, d16[has].astype(np.uint16)))
        else:
            dos.append(None)
    return g, dos


def rewrite_with_all_ref_records(src: str, dst: str) -> int:
    """Re-emit a variable-width .pgen (mode 0x10) so that every all-ref record -- which pgenlib's writer
    stores as a type-4 record holding an empty difflist -- becomes a zero-length vrtype-5 record (the writer
    only emits those under FUTURE_ENCODER, pgenlib_write.cc:1009-1014, but every reader accepts them).
    Records that serve as an LD base are left alone.  Index layout: 8-bit vrtypes, 4-byte record lengths
    (header_ctrl 0x47).  Returns the number of records converted."""
    from oracle.port import PortPgen

    with PortPgen(src) as p:
        vt, fpos = p.index()
        V, N = p.variant_ct, p.sample_ct
    vt = np.asarray(vt).astype(np.uint8)
    fpos = np.asarray(fpos).astype(np.int64)
    raw = np.fromfile(src, dtype=np.uint8)
    recs, types, converted = [], [], 0
    for v in range(V):
        rec = raw[fpos[v]:fpos[v + 1]].tobytes()
        t = int(vt[v])
        next_is_ld = v + 1 < V and (int(vt[v + 1]) & 7) in (2, 3)
        if (t & 7) == 4 and t == 4 and rec == b"\x00" and not next_is_ld:
            rec, t = b"", 5
            converted += 1
        recs.append(rec)
        types.append(t)
    nblocks = (V + 65535) // 65536
    hdr_len = 12 + 8 * nblocks + V * 5
    out = bytearray(b"\x6c\x1b\x10" + struct.pack("<II", V, N) + b"\x47")
    lens = np.array([len(r) for r in recs], dtype=np.int64)
    starts = hdr_len + np.concatenate([[0], np.cumsum(lens)])
    for b in range(nblocks):
        out += struct.pack("<Q", int(starts[b * 65536]))
    for b in range(nblocks):
        lo, hi = b * 65536, min(V, (b + 1) * 65536)
        out += bytes(types[lo:hi])
        out += lens[lo:hi].astype("<u4").tobytes()
    assert len(out) == hdr_len
    for r in recs:
        out += r
    with open(dst, "wb") as f:
        f.write(bytes(out))
    return converted


def gen_phase(g: np.ndarray, seed: int):
    """Per-variant hardcall phase for oracle.ref.write_pgen(phase=...): a third of the variants carry no
    phase track (None), a third have every het phased (pgenlib stores the implicit form), a third have a
    random subset of their hets phased (explicit phasepresent form).  Returns (list, expected_pairs) with
    expected_pairs (V, N, 2) int8 following UnpackPhasedGenotypes."""
    rng = np.random.default_rng(seed)
    V, N = g.shape
    nw = (N + 63) // 64
    phase = []
    exp = np.zeros((V, N, 2), dtype=np.int8)
    exp[g == 1] = (0, 1)
    exp[g == 2] = (1, 1)
    exp[g == 3] = (-9, -9)

    def pack(bits):
        b = np.zeros(nw * 64, dtype=np.uint8)
        b[:N] = bits
        return np.packbits(b, bitorder="little").view(np.uint64).copy()

    for v in range(V):
        het = g[v] == 1
        kind = v % 3
        if kind == 0 or not het.any():
            phase.append(None)
            continue
        present = het.copy() if kind == 1 else (het & (rng.random(N) < 0.6))
        if not present.any():
            phase.append(None)
            continue
        info = present & (rng.random(N) < 0.5)
        phase.append((pack(present), pack(info)))
        exp[v, info] = (1, 0)
    return phase, exp


# header_ctrl modes 8-14 (pgenlib_read.cc:928-957): code -> (vrtype, bytes beyond / instead of the main track).
# n4 = ceil(N / 4) is added where the record has a main 2-bit track.
_COMPACT = {
    8: (1, {0: (5, None), 1: (0, 0)}),
    9: (2, {0: (5, None), 1: (0, 0), 2: (8, 2), 3: (8, 3)}),
    10: (2, {0: (5, None), 1: (0, 0), 2: (16, 1), 3: (16, 2)}),
    11: (4, dict([(0, (5, None)), (1, (0, 0)), (9, (16, 1))] + [(i, (8, i)) for i in range(2, 8)]
                 + [(i + 8, (24, 1 + i)) for i in range(2, 8)])),
    12: (2, {0: (5, None), 1: (0, 0), 2: (0x45, -2), 3: (0x40, -3)}),
    14: (4, {0: (5, None), 1: (0, 0), 3: (0x10, -2), 4: (0x45, -2), 5: (0x40, -3), 12: (0xc5, -4), 13: (0xc0, -5),
             15: (0xe0, -6)}),
}


def compact_pgen_image(storage: int, g: np.ndarray, seed: int = 0):
    """A .pgen in one of the compact header_ctrl modes 8-14, built by hand (pgenlib's writer never emits them):
    g is (V, N) genotypes in {0,1,2,3}; a variant whose calls are all hom-ref may get code 0 (vrtype 5, empty
    record), the others a random code of the mode whose record starts with the 2-bit main track, followed by
    random bytes where the code says the record carries more (those tracks are not parsed by counts / PgrGet).
    Single-sample modes 12 / 14: codes without a main track (vrtype 0x45 / 0xc5, 2-4 bytes) stand for all-ref
    variants too.  Returns (image, vrtypes, fpos)."""
    bits, table = _COMPACT[storage]
    V, N = g.shape
    n4 = (N + 3) // 4
    rng = np.random.default_rng(seed)
    packed = pack_2bit_bytes(g)
    with_main = [c for c, (vt, extra) in table.items() if (vt & 7) != 5 and (extra is None or extra >= 0 or -extra > n4)]
    no_main = [c for c, (vt, extra) in table.items() if (vt & 7) == 5]
    codes = np.zeros(V, dtype=np.uint8)
    recs, vrtypes, lens = [], np.zeros(V, np.uint8), np.zeros(V, np.uint64)
    for v in range(V):
        if not g[v].any() and rng.random() < 0.5:
            c = int(rng.choice(no_main))
        else:
            c = int(rng.choice(with_main))
        vt, extra = table[c]
        if (vt & 7) == 5:
            rec = b"" if extra is None else bytes(rng.integers(0, 256, -extra, dtype=np.uint8))
        else:
            total = n4 + extra if extra >= 0 else -extra
            rec = packed[v].tobytes() + bytes(rng.integers(0, 256, total - n4, dtype=np.uint8))
        if storage == 12 and (vt & 0x60) == 0x40:  # single sample, unconditional dosage: one valid u16 (65535 = missing)
            d16 = 65535 if rng.random() < 0.1 else int(rng.integers(0, 32769))
            rec = rec[:-2] + int(d16).to_bytes(2, "little")
        codes[v], vrtypes[v], lens[v] = c, vt, len(rec)
        recs.append(rec)
    nblk = (V + 65535) // 65536
    hdr = bytearray(b"\x6c\x1b\x10") + np.array([V, N], "<u4").tobytes() + bytes([0x40 | storage])
    blocks = bytearray()
    for b in range(nblk):
        cb = codes[b * 65536:(b + 1) * 65536]
        out = np.zeros((len(cb) * bits + 7) // 8, dtype=np.uint8)
        for i, c in enumerate(cb):
            out[(i * bits) >> 3] |= int(c) << ((i * bits) & 7)
        blocks += out.tobytes()
    first = len(hdr) + 8 * nblk + len(blocks)
    fpos = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64) + np.uint64(first)
    offs = np.array([fpos[b * 65536] for b in range(nblk)], "<u8").tobytes()
    image = np.frombuffer(bytes(hdr) + offs + bytes(blocks) + b"".join(recs), dtype=np.uint8).copy()
    return image, vrtypes, fpos


def genotypes_from_counts(counts: np.ndarray, seed: int) -> np.ndarray:
    """(V, N) int8 genotype matrix whose per-variant tallies are exactly counts[v] = (hom_ref, het, hom_alt,
    missing), samples shuffled: lets a test dictate the table an exact test sees."""
    counts = np.asarray(counts, dtype=np.int64)
    N = int(counts[0].sum())
    rng = np.random.default_rng(seed)
    g = np.empty((counts.shape[0], N), dtype=np.int8)
    for v, c in enumerate(counts):
        assert int(c.sum()) == N
        row = np.repeat(np.arange(4, dtype=np.int8), c)
        rng.shuffle(row)
        g[v] = row
    return g


def write_bed(path: str, g: np.ndarray):
    """PLINK 1 variant-major .bed (mode 0x01): magic 6c 1b 01, then ceil(N/4) bytes per variant with the PLINK 1
    codes 3 hom-ref, 2 het, 0 hom-alt, 1 missing (pgen_spec.tex:428-432); pad codes are 0."""
    to_bed = np.array([3, 2, 0, 1], dtype=np.uint8)
    V, N = g.shape
    nb = (N + 3) // 4
    pad = np.zeros((V, nb * 4), dtype=np.uint8)
    pad[:, :N] = to_bed[g.astype(np.uint8) & 3]
    q = pad.reshape(V, nb, 4)
    body = (q[:, :, 0] | (q[:, :, 1] << 2) | (q[:, :, 2] << 4) | (q[:, :, 3] << 6)).astype(np.uint8)
    with open(path, "wb") as f:
        f.write(bytes([0x6C, 0x1B, 0x01]))
        f.write(body.tobytes())
