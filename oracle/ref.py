"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/_ref/libpgenref.so.

libpgenref.so is the reference's own pgenlib (compiled by oracle/Makefile from the sources
under /root/reference, never copied) plus oracle/ref_harness.cc.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libpgenref.so")
_SO_AVX2 = os.path.join(_HERE, "_ref", "avx2", "libpgenref.so")  # `make ref_avx2`: -march=x86-64-v3 build

_lib = None
_lib_avx2 = None


def available() -> bool:
    return os.path.exists(_SO)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError(f"{_SO} missing: run `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(_SO)
        vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int32
        L.ref_open.argtypes = [C.c_char_p, C.POINTER(vp), C.c_char_p, C.c_size_t]
        L.ref_open.restype = C.c_int
        L.ref_open_counts.argtypes = [C.c_char_p, u32, u32, C.POINTER(vp), C.c_char_p, C.c_size_t]
        L.ref_open_counts.restype = C.c_int
        L.ref_open_alleles.argtypes = [C.c_char_p, vp, C.POINTER(vp), C.c_char_p, C.c_size_t]
        L.ref_open_alleles.restype = C.c_int
        L.ref_decode_alleles.argtypes = [vp, u32, u32, vp]
        L.ref_decode_alleles.restype = C.c_int
        L.ref_close.argtypes = [vp]
        L.ref_info.argtypes = [vp, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)]
        L.ref_vrtype.argtypes = [vp, u32]
        L.ref_vrtype.restype = u32
        L.ref_fpos.argtypes = [vp, u32]
        L.ref_fpos.restype = u64
        L.ref_set_subset.argtypes = [vp, vp, u32]
        L.ref_counts.argtypes = [vp, u32, u32, vp]
        L.ref_dcounts.argtypes = [vp, u32, u32, vp, vp, vp]
        L.ref_decode_i8.argtypes = [vp, u32, u32, vp]
        L.ref_missingness.argtypes = [vp, u32, u32, vp, vp]
        L.ref_dosage.argtypes = [vp, u32, vp, C.POINTER(u32)]
        L.ref_score.argtypes = [vp, vp, vp, vp, u32, u32, vp, vp, vp]
        L.ref_sample_counts.argtypes = [vp, u32, u32, u32, vp, vp, vp]
        L.ref_hwe_lnp.argtypes = [i32, i32, i32, u32]
        L.ref_hwe_lnp.restype = C.c_double
        L.ref_hwe_xchr_lnp.argtypes = [i32, i32, i32, i32, i32, u32]
        L.ref_hwe_xchr_lnp.restype = C.c_double
        L.ref_hwe_lnp_batch.argtypes = [vp, u32, u32, vp]
        L.ref_bench_scan.argtypes = [C.c_char_p, u32, u32, u32, u32, C.POINTER(u64)]
        L.ref_bench_scan.restype = C.c_double
        L.ref_bench_score.argtypes = [C.c_char_p, u32, vp, vp, vp, u32, u32, vp, vp]
        L.ref_bench_score.restype = C.c_double
        L.ref_writer_open.argtypes = [C.c_char_p, u32, u32, u32, C.POINTER(vp)]
        L.ref_writer_append.argtypes = [vp, vp]
        L.ref_writer_append_dosage.argtypes = [vp, vp, vp, vp, u32]
        L.ref_writer_append_phased.argtypes = [vp, vp, vp, vp]
        L.ref_decode_phased.argtypes = [vp, u32, u32, vp]
        L.ref_ld_pair.argtypes = [vp, u32, u32, vp, C.POINTER(u32), C.POINTER(u32)]
        L.ref_writer_open_ex.argtypes = [C.c_char_p, u32, u32, u32, vp, C.POINTER(vp)]
        L.ref_writer_append_multiallelic.argtypes = [vp, vp, vp, vp, u32, vp, vp, u32, u32]
        L.ref_writer_append_multiallelic_hphase.argtypes = [vp, vp, vp, vp, u32, vp, vp, u32, vp, vp, u32]
        L.ref_writer_append_multiallelic_hphase.restype = C.c_int
        L.ref_writer_append_dphase.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, u32, u32]
        L.ref_writer_close.argtypes = [vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class RefPgen:
    """One pgenlib reader (PgenFileInfo + PgenReader), as each DuckDB scan thread owns."""

    def __init__(self, path: str, raw_counts=None, allele_cts=None):
        """raw_counts = (V, N) for PLINK 1 .bed files, which store neither.  allele_cts = per-variant allele counts
        (from the .pvar) for the multiallelic view (decode_alleles)."""
        self._h = C.c_void_p()
        err = C.create_string_buffer(512)
        if allele_cts is not None:
            ac = np.ascontiguousarray(allele_cts, dtype=np.uint32)
            rc = lib().ref_open_alleles(os.fsencode(path), _p(ac), C.byref(self._h), err, 512)
        elif raw_counts is None:
            rc = lib().ref_open(os.fsencode(path), C.byref(self._h), err, 512)
        else:
            rc = lib().ref_open_counts(os.fsencode(path), int(raw_counts[0]), int(raw_counts[1]), C.byref(self._h), err, 512)
        if rc:
            raise IOError(f"ref_open({path}) failed rc={rc}: {err.value.decode(errors='replace')}")
        self._refresh()

    def _refresh(self):
        V, N, S = C.c_uint32(), C.c_uint32(), C.c_uint32()
        lib().ref_info(self._h, C.byref(V), C.byref(N), C.byref(S))
        self.V, self.N, self.sample_ct = V.value, N.value, S.value

    def decode_alleles(self, v0=0, v1=None) -> np.ndarray:
        """PgrGetM + PglMultiallelicSparseToDenseMiss: (rows, sample_ct, 2) uint8 allele codes, 255 = missing."""
        v1 = self.V if v1 is None else v1
        out = np.zeros((v1 - v0, self.sample_ct, 2), dtype=np.uint8)
        rc = lib().ref_decode_alleles(self._h, v0, v1, _p(out))
        if rc:
            raise IOError(f"ref_decode_alleles failed rc={rc}")
        return out

    def close(self):
        if self._h:
            lib().ref_close(self._h)
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

    def vrtypes(self) -> np.ndarray:
        return np.array([lib().ref_vrtype(self._h, v) for v in range(self.V)], dtype=np.uint8)

    def fpos(self) -> np.ndarray:
        return np.array([lib().ref_fpos(self._h, v) for v in range(self.V + 1)], dtype=np.uint64)

    def set_subset(self, idx):
        if idx is None:
            lib().ref_set_subset(self._h, None, 0)
        else:
            a = np.ascontiguousarray(np.sort(np.asarray(idx, dtype=np.uint32)))
            lib().ref_set_subset(self._h, _p(a), a.size)
        self._refresh()

    def _rng(self, v0, v1):
        v1 = self.V if v1 is None else v1
        return v0, v1

    def counts(self, v0=0, v1=None) -> np.ndarray:
        v0, v1 = self._rng(v0, v1)
        out = np.zeros((v1 - v0, 4), dtype=np.uint32)
        rc = lib().ref_counts(self._h, v0, v1, _p(out))
        if rc:
            raise IOError(f"PgrGetCounts rc={rc}")
        return out

    def dcounts(self, v0=0, v1=None):
        v0, v1 = self._rng(v0, v1)
        n = v1 - v0
        counts = np.zeros((n, 4), dtype=np.uint32)
        ad = np.zeros((n, 2), dtype=np.uint64)
        r2 = np.zeros(n, dtype=np.float64)
        rc = lib().ref_dcounts(self._h, v0, v1, _p(counts), _p(ad), _p(r2))
        if rc:
            raise IOError(f"PgrGetDCounts rc={rc}")
        return counts, ad, r2

    def decode_i8(self, v0=0, v1=None) -> np.ndarray:
        v0, v1 = self._rng(v0, v1)
        out = np.zeros((v1 - v0, self.sample_ct), dtype=np.int8)
        rc = lib().ref_decode_i8(self._h, v0, v1, _p(out))
        if rc:
            raise IOError(f"PgrGet rc={rc}")
        return out

    def decode_phased(self, v0=0, v1=None) -> np.ndarray:
        """(v1-v0, sample_ct, 2) int8 allele pairs: PgrGetP + UnpackPhasedGenotypes."""
        v0, v1 = self._rng(v0, v1)
        out = np.zeros((v1 - v0, self.sample_ct, 2), dtype=np.int8)
        rc = lib().ref_decode_phased(self._h, v0, v1, _p(out))
        if rc:
            raise IOError(f"PgrGetP rc={rc}")
        return out

    def missingness(self, v0=0, v1=None):
        v0, v1 = self._rng(v0, v1)
        pv = np.zeros(v1 - v0, dtype=np.uint32)
        ps = np.zeros(self.sample_ct, dtype=np.uint32)
        rc = lib().ref_missingness(self._h, v0, v1, _p(pv), _p(ps))
        if rc:
            raise IOError(f"PgrGetMissingness rc={rc}")
        return pv, ps

    def dosage(self, vidx: int) -> np.ndarray:
        out = np.zeros(self.sample_ct, dtype=np.float64)
        ct = C.c_uint32()
        rc = lib().ref_dosage(self._h, vidx, _p(out), C.byref(ct))
        if rc:
            raise IOError(f"PgrGetD rc={rc}")
        return out

    def score(self, vidx, flip, weight, mode=0):
        vidx = np.ascontiguousarray(vidx, dtype=np.uint32)
        flip = np.ascontiguousarray(flip, dtype=np.uint8)
        weight = np.ascontiguousarray(weight, dtype=np.float64)
        s = np.zeros(self.sample_ct, dtype=np.float64)
        d = np.zeros(self.sample_ct, dtype=np.float64)
        a = np.zeros(self.sample_ct, dtype=np.uint32)
        rc = lib().ref_score(self._h, _p(vidx), _p(flip), _p(weight), vidx.size, mode, _p(s), _p(d), _p(a))
        if rc:
            raise IOError(f"ref_score rc={rc}")
        return s, d, a

    def ld_pair(self, va: int, vb: int):
        """(r2, d_prime, obs_ct, is_valid) of plink_ld's ComputeLdStats for one variant pair."""
        out = np.zeros(2, dtype=np.float64)
        obs, valid = C.c_uint32(), C.c_uint32()
        rc = lib().ref_ld_pair(self._h, int(va), int(vb), _p(out), C.byref(obs), C.byref(valid))
        if rc:
            raise IOError(f"PgrGet rc={rc}")
        return float(out[0]), float(out[1]), obs.value, valid.value

    def sample_counts(self, v0=0, v1=None, sparse=False):
        v0, v1 = self._rng(v0, v1)
        het = np.zeros(self.sample_ct, dtype=np.uint32)
        hom_alt = np.zeros(self.sample_ct, dtype=np.uint32)
        missing = np.zeros(self.sample_ct, dtype=np.uint32)
        rc = lib().ref_sample_counts(self._h, v0, v1, int(sparse), _p(het), _p(hom_alt), _p(missing))
        if rc:
            raise IOError(f"ref_sample_counts rc={rc}")
        return het, hom_alt, missing


def hwe_lnp(hets, hom1, hom2, midp=False) -> float:
    return lib().ref_hwe_lnp(int(hets), int(hom1), int(hom2), int(bool(midp)))


def hwe_xchr_lnp(fhets, fhom1, fhom2, male1, male2, midp=False) -> float:
    return lib().ref_hwe_xchr_lnp(int(fhets), int(fhom1), int(fhom2), int(male1), int(male2), int(bool(midp)))


def hwe_lnp_batch(counts: np.ndarray, midp=False) -> np.ndarray:
    counts = np.ascontiguousarray(counts, dtype=np.uint32)
    out = np.zeros(counts.shape[0], dtype=np.float64)
    lib().ref_hwe_lnp_batch(_p(counts), counts.shape[0], int(bool(midp)), _p(out))
    return out


def lnp_to_pvalue(lnp):
    """LnPToPvalue, src/plink_hardy.cpp:52-64."""
    lnp = np.asarray(lnp, dtype=np.float64)
    p = np.exp(lnp)
    p = np.where(np.isnan(lnp), 1.0, np.clip(p, 0.0, 1.0))
    return p


def avx2_available() -> bool:
    return os.path.exists(_SO_AVX2)


def _bench_lib(avx2: bool):
    """The library a timed scan runs in: the reference-flags (SSE2) build, or the AVX2 build (timing only)."""
    global _lib_avx2
    if not avx2:
        return lib()
    if _lib_avx2 is None:
        L = C.CDLL(_SO_AVX2)
        L.ref_bench_scan.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64)]
        L.ref_bench_scan.restype = C.c_double
        _lib_avx2 = L
    return _lib_avx2


def bench_scan(path, threads, v0, v1, kind, avx2: bool = False):
    cs = C.c_uint64()
    t = _bench_lib(avx2).ref_bench_scan(os.fsencode(path), threads, v0, v1, kind, C.byref(cs))
    if t < 0:
        raise IOError("ref_bench_scan: open failed")
    return t, cs.value


def bench_score(path, threads, vidx, flip, weights):
    vidx = np.ascontiguousarray(vidx, dtype=np.uint32)
    flip = np.ascontiguousarray(flip, dtype=np.uint8)
    weights = np.ascontiguousarray(weights, dtype=np.float64)
    n, K = weights.shape
    with RefPgen(path) as r:
        N = r.N
    sums = np.zeros((N, K), dtype=np.float64)
    cts = np.zeros(N, dtype=np.uint32)
    t = lib().ref_bench_score(os.fsencode(path), threads, _p(vidx), _p(flip), _p(weights), n, K, _p(sums), _p(cts))
    if t < 0:
        raise IOError("ref_bench_score: open failed")
    return t, sums, cts


def write_pgen(path: str, genovecs: np.ndarray, N: int, dosage=None, phase=None):
    """Write a variable-width (mode 0x10) .pgen with pgenlib's own writer.

    genovecs: (V, ceil(N/32)) uint64, 2-bit codes, trailing bits zero.
    dosage:   optional list of (present_bits uint64[ceil(N/64)], u16 values) per variant or None.
    phase:    optional list of (phasepresent uint64[ceil(N/64)], phaseinfo uint64[ceil(N/64)]) per variant or
              None (hardcall phase, SpgwAppendBiallelicGenovecHphase); not combined with dosage here.
    """
    genovecs = np.ascontiguousarray(genovecs, dtype=np.uint64)
    V = genovecs.shape[0]
    h = C.c_void_p()
    assert dosage is None or phase is None
    flags = (1 if dosage is not None else 0) | (2 if phase is not None else 0)
    rc = lib().ref_writer_open(os.fsencode(path), V, N, flags, C.byref(h))
    if rc:
        raise IOError(f"SpgwInitPhase1 rc={rc}")
    for v in range(V):
        row = np.ascontiguousarray(genovecs[v])
        if dosage is not None and dosage[v] is not None:
            pres, vals = dosage[v]
            pres = np.ascontiguousarray(pres, dtype=np.uint64)
            vals = np.ascontiguousarray(vals, dtype=np.uint16)
            rc = lib().ref_writer_append_dosage(h, _p(row), _p(pres), _p(vals), vals.size)
        elif phase is not None and phase[v] is not None:
            pp = np.ascontiguousarray(phase[v][0], dtype=np.uint64)
            pi = np.ascontiguousarray(phase[v][1], dtype=np.uint64)
            rc = lib().ref_writer_append_phased(h, _p(row), _p(pp), _p(pi))
        else:
            rc = lib().ref_writer_append(h, _p(row))
        if rc:
            raise IOError(f"SpgwAppend rc={rc} at variant {v}")
    rc = lib().ref_writer_close(h)
    if rc:
        raise IOError(f"SpgwFinish rc={rc}")


class RefWriter:
    """pgenlib's own writer, record by record (fixtures with multiallelic / dosage-phase tracks).
    flags: 1 dosage, 2 hardcall phase, 4 dosage phase; allele_cts: per-variant allele counts or None."""

    def __init__(self, path: str, V: int, N: int, flags: int = 0, allele_cts=None):
        self._h = C.c_void_p()
        self.N = N
        ac = None if allele_cts is None else np.ascontiguousarray(allele_cts, dtype=np.uint32)
        rc = lib().ref_writer_open_ex(os.fsencode(path), V, N, flags, _p(ac), C.byref(self._h))
        if rc:
            raise IOError(f"SpgwInitPhase1 rc={rc}")

    def _bits(self, a):
        return np.ascontiguousarray(a, dtype=np.uint64)

    def append(self, genovec):
        rc = lib().ref_writer_append(self._h, _p(np.ascontiguousarray(genovec, dtype=np.uint64)))
        if rc:
            raise IOError(f"SpgwAppendBiallelicGenovec rc={rc}")

    def append_hphase(self, genovec, phasepresent, phaseinfo):
        rc = lib().ref_writer_append_phased(self._h, _p(np.ascontiguousarray(genovec, dtype=np.uint64)), _p(self._bits(phasepresent)),
                                            _p(self._bits(phaseinfo)))
        if rc:
            raise IOError(f"SpgwAppendBiallelicGenovecHphase rc={rc}")

    def append_multiallelic(self, genovec, set01, vals01, set10, vals10, allele_ct):
        v01 = np.ascontiguousarray(vals01, dtype=np.uint8)
        v10 = np.ascontiguousarray(vals10, dtype=np.uint8).reshape(-1)
        rc = lib().ref_writer_append_multiallelic(self._h, _p(np.ascontiguousarray(genovec, dtype=np.uint64)), _p(self._bits(set01)),
                                                  _p(v01), v01.size, _p(self._bits(set10)), _p(v10), v10.size // 2, int(allele_ct))
        if rc:
            raise IOError(f"SpgwAppendMultiallelicSparse rc={rc}")

    def append_multiallelic_hphase(self, genovec, set01, vals01, set10, vals10, phasepresent, phaseinfo, allele_ct):
        v01 = np.ascontiguousarray(vals01, dtype=np.uint8)
        v10 = np.ascontiguousarray(vals10, dtype=np.uint8).reshape(-1)
        rc = lib().ref_writer_append_multiallelic_hphase(
            self._h, _p(np.ascontiguousarray(genovec, dtype=np.uint64)), _p(self._bits(set01)), _p(v01), v01.size,
            _p(self._bits(set10)), _p(v10), v10.size // 2, _p(self._bits(phasepresent)), _p(self._bits(phaseinfo)), int(allele_ct))
        if rc:
            raise IOError(f"SpgwAppendMultiallelicGenovecHphase rc={rc}")

    def append_dphase(self, genovec, phasepresent, phaseinfo, dosage_present, dphase_present, dosage_main, dphase_delta):
        dm = np.ascontiguousarray(dosage_main, dtype=np.uint16)
        dd = np.ascontiguousarray(dphase_delta, dtype=np.int16)
        rc = lib().ref_writer_append_dphase(self._h, _p(np.ascontiguousarray(genovec, dtype=np.uint64)), _p(self._bits(phasepresent)),
                                            _p(self._bits(phaseinfo)), _p(self._bits(dosage_present)), _p(self._bits(dphase_present)),
                                            _p(dm), _p(dd), dm.size, dd.size)
        if rc:
            raise IOError(f"SpgwAppendBiallelicGenovecDphase16 rc={rc}")

    def close(self):
        if self._h:
            rc = lib().ref_writer_close(self._h)
            self._h = C.c_void_p()
            if rc:
                raise IOError(f"SpgwFinish rc={rc}")
