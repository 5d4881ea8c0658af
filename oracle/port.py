"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/libpgen_oracle.so (pgen_oracle.c, the plain-C
restatement of the reference's genotype path).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline leg may import this."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libpgen_oracle.so")
_lib = None


def available() -> bool:
    return os.path.exists(_SO)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError(f"{_SO} missing: run `make -C oracle port`")
        L = C.CDLL(_SO)
        vp, u32 = C.c_void_p, C.c_uint32
        L.po_open.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(vp), C.c_char_p, C.c_size_t]
        L.po_close.argtypes = [vp]
        L.po_info.argtypes = [vp, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)]
        L.po_vrtype.argtypes = [vp, u32]
        L.po_vrtype.restype = u32
        L.po_fpos.argtypes = [vp, u32]
        L.po_fpos.restype = C.c_uint64
        L.po_read_codes.argtypes = [vp, u32, vp]
        L.po_counts.argtypes = [vp, u32, u32, vp, vp]
        L.po_decode_i8.argtypes = [vp, u32, u32, vp]
        L.po_decode_phased.argtypes = [vp, u32, u32, vp]
        L.po_missing_by_sample.argtypes = [vp, u32, u32, vp]
        L.po_hwe_exact.argtypes = [u32, u32, u32, C.c_int]
        L.po_hwe_exact.restype = C.c_double
        L.po_hwe_lnp_exact.argtypes = [u32, u32, u32, C.c_int]
        L.po_hwe_lnp_exact.restype = C.c_double
        L.po_score.argtypes = [vp, vp, vp, vp, u32, C.c_int, vp, vp, vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def hwe_exact(hets: int, hom1: int, hom2: int, midp: bool = False) -> float:
    return float(lib().po_hwe_exact(hets, hom1, hom2, 1 if midp else 0))


def hwe_lnp_exact(hets: int, hom1: int, hom2: int, midp: bool = False) -> float:
    """ln p in x87 long double with an unbounded exponent (po_hwe_lnp_exact): the yardstick for HWE parity."""
    return float(lib().po_hwe_lnp_exact(int(hets), int(hom1), int(hom2), 1 if midp else 0))


class PortPgen:
    def __init__(self, path: str):
        L = lib()
        self._h = C.c_void_p()
        err = C.create_string_buffer(256)
        pgi = path + ".pgi"
        rc = L.po_open(path.encode(), pgi.encode() if os.path.exists(pgi) else None, C.byref(self._h), err, 256)
        if rc:
            raise RuntimeError(f"po_open({path}): {err.value.decode()} (rc={rc})")
        v, n, m = C.c_uint32(), C.c_uint32(), C.c_uint32()
        L.po_info(self._h, C.byref(v), C.byref(n), C.byref(m))
        self.variant_ct, self.sample_ct, self.mode = v.value, n.value, m.value

    def close(self):
        if self._h:
            lib().po_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _chk(self, rc):
        if rc:
            raise RuntimeError(f"pgen_oracle error {rc}")

    def index(self):
        L = lib()
        vt = np.array([L.po_vrtype(self._h, v) for v in range(self.variant_ct)], dtype=np.uint8)
        fp = np.array([L.po_fpos(self._h, v) for v in range(self.variant_ct + 1)], dtype=np.uint64)
        return vt, fp

    def counts(self, v0=0, v1=None, subset=None) -> np.ndarray:
        v1 = self.variant_ct if v1 is None else v1
        out = np.zeros((v1 - v0, 4), dtype=np.uint32)
        inc = None
        if subset is not None:
            bits = np.zeros(self.sample_ct, dtype=np.uint8)
            bits[np.asarray(subset, dtype=np.int64)] = 1
            inc = np.packbits(bits, bitorder="little")
        self._chk(lib().po_counts(self._h, v0, v1, _p(inc) if inc is not None else None, _p(out)))
        return out

    def decode_i8(self, v0=0, v1=None) -> np.ndarray:
        v1 = self.variant_ct if v1 is None else v1
        out = np.zeros((v1 - v0, self.sample_ct), dtype=np.int8)
        self._chk(lib().po_decode_i8(self._h, v0, v1, _p(out)))
        return out

    def decode_phased(self, v0=0, v1=None) -> np.ndarray:
        v1 = self.variant_ct if v1 is None else v1
        out = np.zeros((v1 - v0, self.sample_ct, 2), dtype=np.int8)
        self._chk(lib().po_decode_phased(self._h, v0, v1, _p(out)))
        return out

    def missing_by_sample(self, v0=0, v1=None) -> np.ndarray:
        v1 = self.variant_ct if v1 is None else v1
        out = np.zeros(self.sample_ct, dtype=np.uint32)
        self._chk(lib().po_missing_by_sample(self._h, v0, v1, _p(out)))
        return out

    def score(self, vidx, flip, weights, mode=0):
        vidx = np.ascontiguousarray(vidx, dtype=np.uint32)
        w = np.ascontiguousarray(weights, dtype=np.float64).reshape(-1)
        fl = None if flip is None else np.ascontiguousarray(flip, dtype=np.uint8)
        s = np.zeros(self.sample_ct)
        d = np.zeros(self.sample_ct)
        a = np.zeros(self.sample_ct, dtype=np.uint32)
        self._chk(lib().po_score(self._h, _p(vidx), _p(fl) if fl is not None else None, _p(w), vidx.size, mode,
                                 _p(s), _p(d), _p(a)))
        return s, d, a


def child_vector_layout(i8: np.ndarray, allowed=(True, True, True)):
    """The unphased genotype-column fill loop of the reference (src/pgen_reader.cpp:1008-1025 for
    ARRAY(TINYINT, N), :1026-1046 for LIST(TINYINT)), restated on a decoded (rows x samples) int8 matrix:
    a call is invalid when it is -9 or when GenotypeFilter::AllowsCall (src/include/plink_common.hpp:471-477)
    rejects its category; child_data is 0 there and the call otherwise; child_validity bit (row * N + s)
    is set where valid (duckdb::ValidityMask: 64-bit words, LSB first).  Returns (data, validity words)."""
    g = np.asarray(i8, dtype=np.int8)
    ok = np.zeros(g.shape, dtype=bool)
    for cat in (0, 1, 2):
        if allowed[cat]:
            ok |= g == cat
    data = np.where(ok, g, 0).astype(np.int8)
    bits = np.packbits(ok.reshape(-1), bitorder="little")
    words = np.zeros(((ok.size + 63) // 64) * 8, dtype=np.uint8)
    words[: bits.size] = bits
    return data, words.view(np.uint64)


def child_vector_layout_loop(i8, allowed=(True, True, True)):
    """Element-by-element form of the same loop (small cases only): pins child_vector_layout."""
    rows, n = len(i8), (len(i8[0]) if len(i8) else 0)
    data = [[0] * n for _ in range(rows)]
    words = [0] * ((rows * n + 63) // 64)
    for r in range(rows):
        base = r * n
        for s in range(n):
            geno = int(i8[r][s])
            if geno == -9 or not (0 <= geno <= 2 and allowed[geno]):
                data[r][s] = 0
            else:
                data[r][s] = geno
                words[(base + s) >> 6] |= 1 << ((base + s) & 63)
    return np.array(data, dtype=np.int8).reshape(rows, n), np.array(words, dtype=np.uint64)
