"""PgenContext: one opened .pgen on one B200, a thin object over the C ABI."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

from . import _native as nat


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def include_bits(indices, n: int) -> np.ndarray:
    """sample_include bitmask (uint64 words, LSB first) from sample indices."""
    words = np.zeros((n + 63) // 64 + 1, dtype=np.uint64)
    idx = np.unique(np.asarray(indices, dtype=np.int64))
    if idx.size and (idx[0] < 0 or idx[-1] >= n):
        raise ValueError("sample index out of range")
    np.bitwise_or.at(words, idx >> 6, np.uint64(1) << (idx & 63).astype(np.uint64))
    return words


def probe(path: str):
    """Header-only parse (no GPU): (Info, vrtypes, fpos)."""
    lib = nat.load()
    info = nat.Info()
    err = C.create_string_buffer(512)
    rc = lib.pd_probe(os.fsencode(path), C.byref(info), None, None, 0, err, 512)
    if rc:
        raise nat.PlinkGpuError(rc, err.value.decode(errors="replace"))
    vt = np.zeros(info.variant_ct, dtype=np.uint8)
    fpos = np.zeros(info.variant_ct + 1, dtype=np.uint64)
    rc = lib.pd_probe(os.fsencode(path), C.byref(info), _p(vt), _p(fpos), info.variant_ct, err, 512)
    if rc:
        raise nat.PlinkGpuError(rc, err.value.decode(errors="replace"))
    return info, vt, fpos


def plan_partition(path: str, k: int, v0: int = 0, v1: Optional[int] = None) -> np.ndarray:
    """cuts[k + 1]: the variant ranges a k-GPU context gives its GPUs (pd_plan_partition; no device needed)."""
    lib = nat.load()
    if v1 is None:
        v1 = probe(path)[0].variant_ct
    cuts = np.zeros(k + 1, dtype=np.uint32)
    err = C.create_string_buffer(512)
    rc = lib.pd_plan_partition(os.fsencode(path), int(v0), int(v1), int(k), _p(cuts), err, 512)
    if rc:
        raise nat.PlinkGpuError(rc, err.value.decode(errors="replace"))
    return cuts


class PgenContext:
    """Owns a pd_ctx.  All heavy lifting happens in libplinkgpu.so on the GPU."""

    def __init__(self, path: Optional[str] = None, *, image: Optional[np.ndarray] = None, pgi: Optional[np.ndarray] = None,
                 device: int = -1, tile_bytes: int = 0, n_staging: int = 0, devices=None, pin_file: bool = False,
                 raw_counts=(0, 0)):
        """devices: list of CUDA ordinals for ONE context over several GPUs of this box (pd_open_opts.n_devices):
        variant ranges of equal record bytes per GPU, per-sample sums all-reduced with NCCL inside the library.
        raw_counts: (variant_ct, sample_ct) for a PLINK 1 .bed file (mode 0x01), which stores neither.
        pin_file: PD_OPEN_PIN_FILE -- page-lock the file's mapping at open so scans DMA straight from the page cache."""
        self._lib = nat.load()
        self._h = C.c_void_p()
        opts, dev_keep = nat.open_opts(device, tile_bytes, n_staging, devices, nat.PD_OPEN_PIN_FILE if pin_file else 0, raw_counts)
        self._keep = (image, pgi, dev_keep)
        err = C.create_string_buffer(512)
        if image is not None:
            pg = pgi.ctypes.data_as(C.c_void_p) if pgi is not None else None
            rc = self._lib.pd_open_mem(image.ctypes.data_as(C.c_void_p), image.nbytes, pg,
                                       pgi.nbytes if pgi is not None else 0, C.byref(opts), C.byref(self._h), err, 512)
        else:
            rc = self._lib.pd_open(os.fsencode(path), C.byref(opts), C.byref(self._h), err, 512)
        if rc:
            raise nat.PlinkGpuError(rc, err.value.decode(errors="replace"))
        info = self.info()
        self.variant_ct = info.variant_ct
        self.raw_sample_ct = info.sample_ct
        self.sample_ct = info.sample_ct  # effective (after subset)

    @classmethod
    def from_pointer(cls, ptr: int, nbytes: int, **kw):
        """Open a .pgen image living at a raw host address (e.g. a pinned torch tensor)."""
        arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(nbytes,))
        return cls(image=arr, **kw)

    @classmethod
    def from_device_pointer(cls, ptr: int, nbytes: int, device: int = -1, **kw):
        """Open a .pgen image that already lives in this GPU's memory (used in place, always resident).
        The allocation must extend >= 64 bytes past nbytes."""
        self = cls.__new__(cls)
        self._lib = nat.load()
        self._h = C.c_void_p()
        self._keep = None
        opts, _ = nat.open_opts(device, kw.get("tile_bytes", 0), kw.get("n_staging", 0))
        err = C.create_string_buffer(512)
        rc = self._lib.pd_open_mem(C.c_void_p(ptr), nbytes, None, 0, C.byref(opts), C.byref(self._h), err, 512)
        if rc:
            raise nat.PlinkGpuError(rc, err.value.decode(errors="replace"))
        info = self.info()
        self.variant_ct = info.variant_ct
        self.raw_sample_ct = info.sample_ct
        self.sample_ct = info.sample_ct
        return self

    # -- lifetime --
    def close(self):
        if self._h:
            self._lib.pd_close(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int):
        if rc:
            raise nat.PlinkGpuError(rc, self._lib.pd_last_error(self._h).decode(errors="replace"))

    def _range(self, v0, v1):
        v1 = self.variant_ct if v1 is None else v1
        return int(v0), int(v1)

    # -- metadata --
    def info(self) -> nat.Info:
        out = nat.Info()
        self._check(self._lib.pd_info(self._h, C.byref(out)))
        return out

    def index(self):
        vt = np.zeros(self.variant_ct, dtype=np.uint8)
        fpos = np.zeros(self.variant_ct + 1, dtype=np.uint64)
        self._check(self._lib.pd_index(self._h, _p(vt), _p(fpos)))
        return vt, fpos

    def stats(self) -> nat.Stats:
        out = nat.Stats()
        self._check(self._lib.pd_get_stats(self._h, C.byref(out)))
        return out

    def reset_stats(self):
        self._check(self._lib.pd_reset_stats(self._h))

    # -- residency / subset --
    def load_resident(self, v0=0, v1=None):
        v0, v1 = self._range(v0, v1)
        self._check(self._lib.pd_load_resident(self._h, v0, v1))

    def unload(self):
        self._check(self._lib.pd_unload(self._h))

    def set_subset(self, indices):
        if indices is None:
            self._check(self._lib.pd_set_sample_subset(self._h, None, 0))
            self.sample_ct = self.raw_sample_ct
            return
        bits = include_bits(indices, self.raw_sample_ct)
        cnt = int(np.unique(np.asarray(indices)).size)
        self._check(self._lib.pd_set_sample_subset(self._h, _p(bits), cnt))
        self.sample_ct = cnt

    # -- per-variant reductions --
    def counts(self, v0=0, v1=None, out=None) -> np.ndarray:
        v0, v1 = self._range(v0, v1)
        out = np.zeros((max(v1 - v0, 0), 4), dtype=np.uint32) if out is None else out
        self._check(self._lib.pd_counts(self._h, v0, v1, _p(out)))
        return out

    def counts_masked(self, stratum_indices, v0=0, v1=None) -> np.ndarray:
        v0, v1 = self._range(v0, v1)
        bits = include_bits(stratum_indices, self.raw_sample_ct)
        out = np.zeros((max(v1 - v0, 0), 4), dtype=np.uint32)
        self._check(self._lib.pd_counts_masked(self._h, v0, v1, _p(bits), _p(out)))
        return out

    def counts_async(self, v0=0, v1=None) -> "CountsJob":
        v0, v1 = self._range(v0, v1)
        return CountsJob(self, v0, v1)

    def hardy_async(self, v0=0, v1=None, midp=False) -> "HardyJob":
        v0, v1 = self._range(v0, v1)
        return HardyJob(self, v0, v1, midp)

    def hardy(self, v0=0, v1=None, midp=False, out_counts=None, out_p=None):
        """out_counts / out_p: optional preallocated (e.g. page-locked) result arrays."""
        v0, v1 = self._range(v0, v1)
        n = max(v1 - v0, 0)
        counts = np.zeros((n, 4), dtype=np.uint32) if out_counts is None else out_counts
        p = np.zeros(n, dtype=np.float64) if out_p is None else out_p
        assert counts.size >= 4 * n and p.size >= n
        self._check(self._lib.pd_hardy(self._h, v0, v1, int(bool(midp)), _p(counts), _p(p)))
        return counts, p

    def hwe_from_counts(self, counts, midp=False) -> np.ndarray:
        counts = np.ascontiguousarray(counts, dtype=np.uint32).reshape(-1, 4)
        p = np.zeros(counts.shape[0], dtype=np.float64)
        self._check(self._lib.pd_hwe_from_counts(self._h, _p(counts), counts.shape[0], int(bool(midp)), _p(p)))
        return p

    def hwe_lnp_from_counts(self, counts, midp=False) -> np.ndarray:
        """ln p of the exact test (plink2::HweLnP's own return value) for (n, 4) counts."""
        counts = np.ascontiguousarray(counts, dtype=np.uint32).reshape(-1, 4)
        out = np.zeros(counts.shape[0], dtype=np.float64)
        self._check(self._lib.pd_hwe_lnp_from_counts(self._h, _p(counts), counts.shape[0], int(bool(midp)), _p(out)))
        return out

    def set_allele_counts(self, allele_cts) -> None:
        """Per-variant allele counts from the .pvar (None clears them); needed by decode_alleles."""
        if allele_cts is None:
            self._check(self._lib.pd_set_allele_counts(self._h, None, 0))
            return
        a = np.ascontiguousarray(allele_cts, dtype=np.uint32)
        self._check(self._lib.pd_set_allele_counts(self._h, _p(a), a.size))

    def decode_alleles(self, v0=0, v1=None) -> np.ndarray:
        """(rows, n_out, 2) uint8 allele codes (PgrGetM + PglMultiallelicSparseToDenseMiss), 255 = missing."""
        info = self.info()
        v1 = info.variant_ct if v1 is None else v1
        out = np.zeros((max(v1 - v0, 0), info.effective_sample_ct, 2), dtype=np.uint8)
        self._check(self._lib.pd_decode_alleles(self._h, v0, v1, _p(out)))
        return out

    def hwe_xchr_from_counts(self, c5, midp=False) -> np.ndarray:
        """c5: (n, 5) female hom_ref / het / hom_alt, male ref / alt."""
        c5 = np.ascontiguousarray(c5, dtype=np.uint32).reshape(-1, 5)
        p = np.zeros(c5.shape[0], dtype=np.float64)
        self._check(self._lib.pd_hwe_xchr_from_counts(self._h, _p(c5), c5.shape[0], int(bool(midp)), _p(p)))
        return p

    # -- per-sample accumulations --
    def missing_by_sample(self, v0=0, v1=None) -> np.ndarray:
        v0, v1 = self._range(v0, v1)
        out = np.zeros(self.sample_ct, dtype=np.uint32)
        self._check(self._lib.pd_missing_by_sample(self._h, v0, v1, _p(out)))
        return out

    def sample_counts(self, v0=0, v1=None):
        v0, v1 = self._range(v0, v1)
        het = np.zeros(self.sample_ct, dtype=np.uint32)
        hom_alt = np.zeros(self.sample_ct, dtype=np.uint32)
        missing = np.zeros(self.sample_ct, dtype=np.uint32)
        self._check(self._lib.pd_sample_counts(self._h, v0, v1, _p(het), _p(hom_alt), _p(missing)))
        return het, hom_alt, missing

    # -- decode --
    def decode_i8(self, v0=0, v1=None) -> np.ndarray:
        v0, v1 = self._range(v0, v1)
        out = np.zeros((max(v1 - v0, 0), self.sample_ct), dtype=np.int8)
        self._check(self._lib.pd_decode_i8(self._h, v0, v1, _p(out)))
        return out

    def decode_async(self, v0=0, v1=None) -> "DecodeJob":
        v0, v1 = self._range(v0, v1)
        return DecodeJob(self, v0, v1)

    def decode_i8_valid(self, v0=0, v1=None, call_mask=7):
        """Genotype column in DuckDB's child-vector layout: (int8 calls with 0 where invalid, uint64 validity
        words over the flattened rows x samples calls).  call_mask bit g: category g in {0,1,2} is selected."""
        v0, v1 = self._range(v0, v1)
        rows = max(v1 - v0, 0)
        out = np.full((rows, self.sample_ct), 0x55, dtype=np.int8)
        validity = np.full((rows * self.sample_ct + 63) // 64, 0xAAAAAAAAAAAAAAAA, dtype=np.uint64)
        self._check(self._lib.pd_decode_i8_valid(self._h, v0, v1, call_mask, _p(out), _p(validity)))
        return out, validity

    def tidy(self, v0=0, v1=None, emit_mask=15):
        """Genotype-orient rows (variant, sample, call) of the calls whose category is in emit_mask."""
        v0, v1 = self._range(v0, v1)
        rows = C.c_uint64()
        self._check(self._lib.pd_tidy_i8(self._h, v0, v1, emit_mask, 0, None, None, None, C.byref(rows)))
        n = rows.value
        var = np.zeros(n, dtype=np.uint32)
        smp = np.zeros(n, dtype=np.uint32)
        geno = np.zeros(n, dtype=np.int8)
        if n:
            self._check(self._lib.pd_tidy_i8(self._h, v0, v1, emit_mask, n, _p(var), _p(smp), _p(geno), C.byref(rows)))
            assert rows.value == n
        return var, smp, geno

    def decode_phased(self, v0=0, v1=None) -> np.ndarray:
        """(v1 - v0, sample_ct, 2) int8 allele pairs (read_pgen / read_pfile with phased := true)."""
        v0, v1 = self._range(v0, v1)
        out = np.zeros((max(v1 - v0, 0), self.sample_ct, 2), dtype=np.int8)
        self._check(self._lib.pd_decode_phased(self._h, v0, v1, _p(out)))
        return out

    def dosage_f64(self, vidx) -> np.ndarray:
        vidx = np.ascontiguousarray(vidx, dtype=np.uint32)
        out = np.zeros((vidx.size, self.sample_ct), dtype=np.float64)
        self._check(self._lib.pd_dosage_f64(self._h, _p(vidx), vidx.size, _p(out)))
        return out

    def ld_pairs(self, va, vb) -> np.ndarray:
        """(n, 6) uint32 sums of plink_ld's estimator for the variant pairs (va[i], vb[i])."""
        va = np.ascontiguousarray(va, dtype=np.uint32)
        vb = np.ascontiguousarray(vb, dtype=np.uint32)
        out = np.zeros((va.size, 6), dtype=np.uint32)
        self._check(self._lib.pd_ld_pairs(self._h, _p(va), _p(vb), va.size, _p(out)))
        return out

    def standardized(self, vidx):
        """(rows (n, sample_ct) float64, skip (n,) uint8): NormalizeGenotypes of the listed variants."""
        vidx = np.ascontiguousarray(vidx, dtype=np.uint32)
        out = np.zeros((vidx.size, self.sample_ct), dtype=np.float64)
        skip = np.zeros(vidx.size, dtype=np.uint8)
        self._check(self._lib.pd_standardized_f64(self._h, _p(vidx), vidx.size, _p(out), _p(skip)))
        return out, skip

    def dosage_counts(self, v0=0, v1=None):
        v0, v1 = self._range(v0, v1)
        n = max(v1 - v0, 0)
        counts = np.zeros((n, 4), dtype=np.uint32)
        ad = np.zeros((n, 2), dtype=np.uint64)
        r2 = np.zeros(n, dtype=np.float64)
        self._check(self._lib.pd_dosage_counts(self._h, v0, v1, _p(counts), _p(ad), _p(r2)))
        return counts, ad, r2

    def score(self, vidx, flip, weights, mode=0):
        vidx = np.ascontiguousarray(vidx, dtype=np.uint32)
        weights = np.ascontiguousarray(weights, dtype=np.float64)
        if weights.ndim == 1:
            weights = weights.reshape(-1, 1)
        n, K = weights.shape
        if n != vidx.size:
            raise ValueError("weights rows must match vidx")
        fl = None if flip is None else np.ascontiguousarray(flip, dtype=np.uint8)
        sums = np.zeros((self.sample_ct, K), dtype=np.float64)
        dsum = np.zeros(self.sample_ct, dtype=np.float64)
        act = np.zeros(self.sample_ct, dtype=np.uint32)
        self._check(self._lib.pd_score(self._h, _p(vidx), _p(fl), _p(weights), n, K, int(mode), _p(sums), _p(dsum), _p(act)))
        return sums, dsum, act


class CountsJob:
    """pd_counts_start / pd_counts_fetch: what a DuckDB scan's claim loop would hold."""

    def __init__(self, ctx: PgenContext, v0: int, v1: int):
        self._ctx = ctx
        self._j = C.c_void_p()
        self.v0, self.v1 = v0, v1
        ctx._check(ctx._lib.pd_counts_start(ctx._h, v0, v1, C.byref(self._j)))

    def fetch(self, a: int, b: int) -> np.ndarray:
        out = np.zeros((max(b - a, 0), 4), dtype=np.uint32)
        rc = self._ctx._lib.pd_counts_fetch(self._j, a, b, _p(out))
        if rc:
            raise nat.PlinkGpuError(rc, self._ctx._lib.pd_last_error(self._ctx._h).decode(errors="replace"))
        return out

    def close(self):
        if self._j:
            self._ctx._lib.pd_job_free(self._j)
            self._j = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class HardyJob(CountsJob):
    """pd_hardy_start / pd_hardy_fetch: counts + exact-test p-values, fetched in claimed row ranges."""

    def __init__(self, ctx: PgenContext, v0: int, v1: int, midp: bool = False):
        self._ctx = ctx
        self._j = C.c_void_p()
        self.v0, self.v1 = v0, v1
        ctx._check(ctx._lib.pd_hardy_start(ctx._h, v0, v1, int(bool(midp)), C.byref(self._j)))

    def fetch(self, a: int, b: int):
        counts = np.zeros((max(b - a, 0), 4), dtype=np.uint32)
        p = np.zeros(max(b - a, 0), dtype=np.float64)
        rc = self._ctx._lib.pd_hardy_fetch(self._j, a, b, _p(counts), _p(p))
        if rc:
            raise nat.PlinkGpuError(rc, self._ctx._lib.pd_last_error(self._ctx._h).decode(errors="replace"))
        return counts, p


class DecodeJob(CountsJob):
    """pd_decode_start / pd_decode_fetch: int8 genotype rows decoded ahead of the scan threads that fetch them."""

    def __init__(self, ctx: PgenContext, v0: int, v1: int):
        self._ctx = ctx
        self._j = C.c_void_p()
        self.v0, self.v1 = v0, v1
        self.n_out = ctx.sample_ct
        ctx._check(ctx._lib.pd_decode_start(ctx._h, v0, v1, C.byref(self._j)))

    def fetch(self, a: int, b: int) -> np.ndarray:
        out = np.zeros((max(b - a, 0), self.n_out), dtype=np.int8)
        rc = self._ctx._lib.pd_decode_fetch(self._j, a, b, _p(out))
        if rc:
            raise nat.PlinkGpuError(rc, self._ctx._lib.pd_last_error(self._ctx._h).decode(errors="replace"))
        return out
