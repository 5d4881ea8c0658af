"""ctypes binding of the operator layer (include/plinkgpu_ops.h): the reference's table functions
as numpy record arrays.  Argument names follow the SQL surface (midp, dosage, samples, region is a
variant range here because .pvar parsing stays in the extension)."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _native as nat
from .pgen import PgenContext, _p

PLOIDY_AUTOSOMAL, PLOIDY_CHR_X, PLOIDY_CHR_Y, PLOIDY_CHR_MT = 0, 1, 2, 3

FREQ_DTYPE = np.dtype([("ALT_FREQ", "<f8"), ("IMP_R2", "<f8"), ("OBS_CT", "<u4"), ("HOM_REF_CT", "<i4"), ("HET_CT", "<i4"),
                       ("HOM_ALT_CT", "<i4"), ("MISSING_CT", "<i4"), ("freq_is_null", "u1"), ("counts_are_null", "u1"),
                       ("pad", "u1", (2,))], align=True)
HARDY_DTYPE = np.dtype([("O_HET", "<f8"), ("E_HET", "<f8"), ("P_HWE", "<f8"), ("HOM_REF_CT", "<i4"), ("HET_CT", "<i4"),
                        ("HOM_ALT_CT", "<i4"), ("stats_are_null", "u1"), ("counts_are_null", "u1"), ("pad", "u1", (2,))],
                       align=True)
MISSING_DTYPE = np.dtype([("MISSING_CT", "<u4"), ("OBS_CT", "<u4"), ("F_MISS", "<f8")], align=True)
GENO_STATS_DTYPE = np.dtype([("hom_ref", "<u4"), ("het", "<u4"), ("hom_alt", "<u4"), ("missing", "<u4"), ("n", "<u4"),
                             ("carrier_count", "<u4"), ("af", "<f8"), ("maf", "<f8"), ("missing_rate", "<f8"),
                             ("het_rate", "<f8")], align=True)
SCORE_DTYPE = np.dtype([("ALLELE_CT", "<u4"), ("DENOM", "<u4"), ("NAMED_ALLELE_DOSAGE_SUM", "<f8"), ("SCORE_SUM", "<f8"),
                        ("SCORE_AVG", "<f8")], align=True)

_vp, _u32 = C.c_void_p, C.c_uint32
_OPS = {
    "pdop_classify_ploidy": (C.c_int, [C.c_char_p, C.c_int32, C.c_char_p]),
    "pdop_freq": (C.c_int, [_vp, _u32, _u32, _vp, _vp, _u32, _vp]),
    "pdop_hardy": (C.c_int, [_vp, _u32, _u32, _vp, _vp, _u32, _vp]),
    "pdop_missing_variant": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pdop_missing_sample": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pdop_score": (C.c_int, [_vp, _vp, _vp, _vp, _u32, _u32, _vp]),
    "pdop_counts_pass": (C.c_int, [_vp, C.c_double, C.c_double, C.c_int64, C.c_int64]),
    "pdop_genotype_range": (None, [_vp, _u32, _u32, _vp, _vp]),
    "pdop_predecomp_filter": (C.c_int, [_vp, _u32, _u32, _vp, _vp, _vp]),
    "pdop_genotype_column": (C.c_int, [_vp, _u32, _u32, _vp, _vp, _vp, _vp, _vp]),
    "pdop_ld_stats": (None, [_vp, _u32, _vp]),
    "pdop_list_entries": (None, [C.c_uint64, _u32, _u32, _vp]),
    "pdop_genotype_stats_variant": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pdop_genotype_stats_sample": (C.c_int, [_vp, _u32, _u32, _vp]),
}


class PredecompFilter(C.Structure):
    """pd_predecomp_filter of include/plinkgpu_ops.h."""
    _fields_ = [("use_count_filter", C.c_uint32), ("af_min", C.c_double), ("af_max", C.c_double),
                ("ac_min", C.c_int64), ("ac_max", C.c_int64), ("use_genotype_filter", C.c_uint32),
                ("allowed_mask", C.c_uint32), ("include_missing", C.c_uint32)]
_bound = False


def _lib():
    global _bound
    lib = nat.load()
    if not _bound:
        for name, (res, args) in _OPS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _bound = True
    return lib


def classify_ploidy(chrom: str, pos: int, build: str = "GRCh38") -> int:
    r = _lib().pdop_classify_ploidy(chrom.encode(), int(pos), build.encode())
    if r < 0:
        raise ValueError(f"unrecognized build '{build}' (expected 'GRCh38'/'hg38', 'GRCh37'/'hg19', or 'none')")
    return r


def _u8(a, n, what):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=np.uint8)
    if a.size != n:
        raise ValueError(f"{what} must have {n} entries")
    return a


def _range(ctx, v0, v1):
    return int(v0), int(ctx.variant_ct if v1 is None else v1)


def plink_freq(ctx: PgenContext, v0=0, v1=None, *, ploidy=None, sex=None, dosage=False) -> np.ndarray:
    v0, v1 = _range(ctx, v0, v1)
    out = np.zeros(max(v1 - v0, 0), dtype=FREQ_DTYPE)
    pl, sx = _u8(ploidy, ctx.variant_ct, "ploidy"), _u8(sex, ctx.raw_sample_ct, "sex")
    ctx._check(_lib().pdop_freq(ctx._h, v0, v1, _p(pl), _p(sx), int(bool(dosage)), _p(out)))
    return out


def plink_hardy(ctx: PgenContext, v0=0, v1=None, *, ploidy=None, sex=None, midp=False) -> np.ndarray:
    v0, v1 = _range(ctx, v0, v1)
    out = np.zeros(max(v1 - v0, 0), dtype=HARDY_DTYPE)
    pl, sx = _u8(ploidy, ctx.variant_ct, "ploidy"), _u8(sex, ctx.raw_sample_ct, "sex")
    ctx._check(_lib().pdop_hardy(ctx._h, v0, v1, _p(pl), _p(sx), int(bool(midp)), _p(out)))
    return out


def plink_missing(ctx: PgenContext, v0=0, v1=None, *, mode="variant") -> np.ndarray:
    v0, v1 = _range(ctx, v0, v1)
    if mode == "variant":
        out = np.zeros(max(v1 - v0, 0), dtype=MISSING_DTYPE)
        ctx._check(_lib().pdop_missing_variant(ctx._h, v0, v1, _p(out)))
    elif mode == "sample":
        out = np.zeros(ctx.sample_ct, dtype=MISSING_DTYPE)
        ctx._check(_lib().pdop_missing_sample(ctx._h, v0, v1, _p(out)))
    else:
        raise ValueError("plink_missing: mode must be 'variant' or 'sample'")
    return out


def plink_score(ctx: PgenContext, vidx: Sequence[int], weights, *, flip=None, no_mean_imputation=False,
                center=False) -> np.ndarray:
    vidx = np.ascontiguousarray(vidx, dtype=np.uint32)
    w = np.ascontiguousarray(weights, dtype=np.float64)
    if w.shape != vidx.shape:
        raise ValueError("plink_score: weights list length must match the number of variants")
    fl = None if flip is None else np.ascontiguousarray(flip, dtype=np.uint8)
    mode = 2 if center else (1 if no_mean_imputation else 0)
    out = np.zeros(ctx.sample_ct, dtype=SCORE_DTYPE)
    ctx._check(_lib().pdop_score(ctx._h, _p(vidx), _p(fl), _p(w), vidx.size, mode, _p(out)))
    return out


def counts_pass(counts, af_min=-1.0, af_max=-1.0, ac_min=-1, ac_max=-1) -> bool:
    c = np.ascontiguousarray(counts, dtype=np.uint32)
    return bool(_lib().pdop_counts_pass(_p(c), float(af_min), float(af_max), int(ac_min), int(ac_max)))


def genotype_range(counts, allowed=(), include_missing=False):
    """CheckGenotypeRange: allowed = iterable of categories 0 hom_ref / 1 het / 2 hom_alt -> (any_pass, all_pass)."""
    c = np.ascontiguousarray(counts, dtype=np.uint32)
    mask = sum(1 << int(g) for g in allowed)
    a, b = C.c_uint32(0), C.c_uint32(0)
    _lib().pdop_genotype_range(_p(c), mask, int(bool(include_missing)), C.addressof(a), C.addressof(b))
    return bool(a.value), bool(b.value)


def predecomp_filter(ctx, v0=0, v1=None, af=None, ac=None, allowed=None, include_missing=False):
    """CheckPreDecompFilters over [v0, v1) on GPU counts -> (skip, all_pass) uint8 arrays.
    af / ac: (min, max) with None for an open end; allowed: iterable of genotype categories or None."""
    v0, v1 = ctx._range(v0, v1)
    n = max(v1 - v0, 0)
    f = PredecompFilter()
    f.use_count_filter = int(af is not None or ac is not None)
    f.af_min, f.af_max = [(-1.0 if x is None else float(x)) for x in (af or (None, None))]
    f.ac_min, f.ac_max = [(-1 if x is None else int(x)) for x in (ac or (None, None))]
    f.use_genotype_filter = int(allowed is not None)
    f.allowed_mask = sum(1 << int(g) for g in (allowed or ()))
    f.include_missing = int(bool(include_missing))
    skip = np.zeros(n, dtype=np.uint8)
    allp = np.zeros(n, dtype=np.uint8)
    ctx._check(_lib().pdop_predecomp_filter(ctx._h, v0, v1, C.addressof(f), _p(skip), _p(allp)))
    return skip, allp


def _filter_struct(af, ac, allowed, include_missing):
    f = PredecompFilter()
    f.use_count_filter = int(af is not None or ac is not None)
    f.af_min, f.af_max = [(-1.0 if x is None else float(x)) for x in (af or (None, None))]
    f.ac_min, f.ac_max = [(-1 if x is None else int(x)) for x in (ac or (None, None))]
    f.use_genotype_filter = int(allowed is not None)
    f.allowed_mask = sum(1 << int(g) for g in (allowed or ()))
    f.include_missing = int(bool(include_missing))
    return f


def genotype_column(ctx, v0=0, v1=None, af=None, ac=None, allowed=None, include_missing=False):
    """read_pgen's genotype column over [v0, v1): (kept variant indices, int8 child data, uint64 validity
    words) for the variants that survive CheckPreDecompFilters, rows compacted as the scan emits them."""
    v0, v1 = ctx._range(v0, v1)
    n = max(v1 - v0, 0)
    f = _filter_struct(af, ac, allowed, include_missing)
    out = np.full((n, ctx.sample_ct), 0x55, dtype=np.int8)
    validity = np.full((n * ctx.sample_ct + 63) // 64, 0xAAAAAAAAAAAAAAAA, dtype=np.uint64)
    kept = np.zeros(n, dtype=np.uint32)
    nk = C.c_uint32(0)
    ctx._check(_lib().pdop_genotype_column(ctx._h, v0, v1, C.addressof(f), _p(out), _p(validity), _p(kept), C.byref(nk)))
    m = nk.value
    return kept[:m], out[:m], validity[: (m * ctx.sample_ct + 63) // 64]


def genotype_stats(ctx: PgenContext, v0=0, v1=None, *, orient="variant") -> np.ndarray:
    """read_pfile(..., genotypes := 'stats' [, orient := 'sample']): the STRUCT column as a record array
    (genotypes := 'counts' is its first four fields)."""
    v0, v1 = _range(ctx, v0, v1)
    if orient == "variant":
        out = np.zeros(max(v1 - v0, 0), dtype=GENO_STATS_DTYPE)
        ctx._check(_lib().pdop_genotype_stats_variant(ctx._h, v0, v1, _p(out)))
    elif orient == "sample":
        out = np.zeros(ctx.sample_ct, dtype=GENO_STATS_DTYPE)
        ctx._check(_lib().pdop_genotype_stats_sample(ctx._h, v0, v1, _p(out)))
    else:
        raise ValueError("genotype_stats: orient must be 'variant' or 'sample'")
    return out


def list_entries(base_offset: int, rows: int, n: int) -> np.ndarray:
    """(rows, 2) uint64 {offset, length} pairs of a LIST(TINYINT) genotype column."""
    out = np.zeros((rows, 2), dtype=np.uint64)
    _lib().pdop_list_entries(int(base_offset), int(rows), int(n), _p(out))
    return out


LD_DTYPE = np.dtype([("R2", "<f8"), ("D_PRIME", "<f8"), ("OBS_CT", "<u4"), ("is_valid", "<u4")], align=True)


def plink_ld(ctx: PgenContext, va, vb) -> np.ndarray:
    """R2 / D_PRIME / OBS_CT rows of plink_ld for the given variant pairs (pd_ld_pairs + pdop_ld_stats)."""
    sums = ctx.ld_pairs(va, vb)
    out = np.zeros(sums.shape[0], dtype=LD_DTYPE)
    _lib().pdop_ld_stats(_p(sums), sums.shape[0], _p(out))
    return out
