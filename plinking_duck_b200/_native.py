"""ctypes binding of libplinkgpu.so (include/plinkgpu.h).

This is the same binding a reference maintainer would write against the C ABI; there is no
Python or CPU fallback behind it: a missing library or a missing GPU raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libplinkgpu.so")

PD_OK = 0
PD_ERR_OPEN, PD_ERR_READ, PD_ERR_MALFORMED, PD_ERR_NOT_SUPPORTED = 1, 2, 3, 4
PD_ERR_NOMEM, PD_ERR_INVALID_ARG, PD_ERR_CUDA = 5, 6, 7
PD_RESULT_SCORE, PD_RESULT_DOSAGE_SUM, PD_RESULT_ALLELE_CT, PD_RESULT_SAMPLE_COUNTS = 0, 1, 2, 3
ERR_NAMES = {
    1: "PD_ERR_OPEN", 2: "PD_ERR_READ", 3: "PD_ERR_MALFORMED", 4: "PD_ERR_NOT_SUPPORTED",
    5: "PD_ERR_NOMEM", 6: "PD_ERR_INVALID_ARG", 7: "PD_ERR_CUDA",
}


class OpenOpts(C.Structure):
    _fields_ = [("device", C.c_int32), ("tile_bytes", C.c_uint64), ("n_staging", C.c_uint32), ("n_devices", C.c_uint32),
                ("devices", C.POINTER(C.c_int32)), ("flags", C.c_uint32), ("raw_variant_ct", C.c_uint32),
                ("raw_sample_ct", C.c_uint32), ("reserved", C.c_uint32)]


PD_OPEN_PIN_FILE = 1


def open_opts(device: int = -1, tile_bytes: int = 0, n_staging: int = 0, devices=None, flags: int = 0, raw_counts=(0, 0)):
    """pd_open_opts; `devices` (a list of CUDA ordinals) asks for ONE context over several GPUs.  Returns
    (opts, keepalive) -- the device array must outlive the pd_open call."""
    if devices is None:
        return OpenOpts(device, tile_bytes, n_staging, 0, None, flags, int(raw_counts[0]), int(raw_counts[1]), 0), None
    arr = (C.c_int32 * len(devices))(*[int(d) for d in devices])
    return OpenOpts(-1, tile_bytes, n_staging, len(devices), C.cast(arr, C.POINTER(C.c_int32)), flags, int(raw_counts[0]),
                    int(raw_counts[1]), 0), arr


class Info(C.Structure):
    _fields_ = [
        ("variant_ct", C.c_uint32), ("sample_ct", C.c_uint32), ("mode", C.c_uint32), ("fixed_width", C.c_uint32),
        ("has_ld", C.c_uint32), ("has_difflist", C.c_uint32), ("has_dosage", C.c_uint32), ("has_phase", C.c_uint32),
        ("has_multiallelic", C.c_uint32), ("resident", C.c_uint32), ("data_bytes", C.c_uint64),
        ("resident_bytes", C.c_uint64), ("effective_sample_ct", C.c_uint32), ("n_devices", C.c_uint32),
        ("pinned_file", C.c_uint32), ("reserved", C.c_uint32),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("launches", C.c_uint64), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("record_bytes", C.c_uint64),
        ("kernel_ms", C.c_double), ("aux_kernel_ms", C.c_double), ("total_ms", C.c_double),
        ("scan_launches", C.c_uint64), ("call_ms", C.c_double), ("host_stage_ms", C.c_double),
        ("host_wait_ms", C.c_double), ("allreduce_ms", C.c_double), ("allreduce_bytes", C.c_uint64),
        ("n_devices", C.c_uint32), ("reserved", C.c_uint32),
    ]


# Every symbol include/plinkgpu.h declares: (name, restype, argtypes)
_vp, _u32, _u64 = C.c_void_p, C.c_uint32, C.c_uint64
SYMBOLS = {
    "pd_open": (C.c_int, [C.c_char_p, C.POINTER(OpenOpts), C.POINTER(_vp), C.c_char_p, C.c_size_t]),
    "pd_open_mem": (C.c_int, [_vp, _u64, _vp, _u64, C.POINTER(OpenOpts), C.POINTER(_vp), C.c_char_p, C.c_size_t]),
    "pd_probe": (C.c_int, [C.c_char_p, C.POINTER(Info), _vp, _vp, _u32, C.c_char_p, C.c_size_t]),
    "pd_plan_partition": (C.c_int, [C.c_char_p, _u32, _u32, _u32, _vp, C.c_char_p, C.c_size_t]),
    "pd_close": (None, [_vp]),
    "pd_last_error": (C.c_char_p, [_vp]),
    "pd_info": (C.c_int, [_vp, C.POINTER(Info)]),
    "pd_index": (C.c_int, [_vp, _vp, _vp]),
    "pd_load_resident": (C.c_int, [_vp, _u32, _u32]),
    "pd_unload": (C.c_int, [_vp]),
    "pd_set_sample_subset": (C.c_int, [_vp, _vp, _u32]),
    "pd_counts": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pd_counts_masked": (C.c_int, [_vp, _u32, _u32, _vp, _vp]),
    "pd_counts_start": (C.c_int, [_vp, _u32, _u32, C.POINTER(_vp)]),
    "pd_counts_fetch": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pd_job_free": (None, [_vp]),
    "pd_hardy": (C.c_int, [_vp, _u32, _u32, _u32, _vp, _vp]),
    "pd_hardy_start": (C.c_int, [_vp, _u32, _u32, _u32, C.POINTER(_vp)]),
    "pd_hardy_fetch": (C.c_int, [_vp, _u32, _u32, _vp, _vp]),
    "pd_hwe_from_counts": (C.c_int, [_vp, _vp, _u32, _u32, _vp]),
    "pd_hwe_lnp_from_counts": (C.c_int, [_vp, _vp, _u32, _u32, _vp]),
    "pd_hwe_xchr_from_counts": (C.c_int, [_vp, _vp, _u32, _u32, _vp]),
    "pd_hwe_noalloc_max_variants": (_u32, []),
    "pd_hwe_from_counts_noalloc": (C.c_int, [_vp, _u32, _u32, _u32, _vp, _vp, C.c_char_p, C.c_size_t]),
    "pd_missing_by_sample": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pd_sample_counts": (C.c_int, [_vp, _u32, _u32, _vp, _vp, _vp]),
    "pd_decode_i8": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pd_decode_alleles": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pd_set_allele_counts": (C.c_int, [_vp, _vp, _u32]),
    "pd_decode_start": (C.c_int, [_vp, _u32, _u32, C.POINTER(_vp)]),
    "pd_decode_fetch": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pd_decode_i8_valid": (C.c_int, [_vp, _u32, _u32, _u32, _vp, _vp]),
    "pd_decode_i8_valid_list": (C.c_int, [_vp, _vp, _u32, _u32, _vp, _vp]),
    "pd_tidy_i8": (C.c_int, [_vp, _u32, _u32, _u32, _u64, _vp, _vp, _vp, C.POINTER(_u64)]),
    "pd_decode_phased": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pd_dosage_f64": (C.c_int, [_vp, _vp, _u32, _vp]),
    "pd_dosage_counts": (C.c_int, [_vp, _u32, _u32, _vp, _vp, _vp]),
    "pd_score": (C.c_int, [_vp, _vp, _vp, _vp, _u32, _u32, _u32, _vp, _vp, _vp]),
    "pd_ld_pairs": (C.c_int, [_vp, _vp, _vp, _u32, _vp]),
    "pd_standardized_f64": (C.c_int, [_vp, _vp, _u32, _vp, _vp]),
    "pd_device_result": (C.c_int, [_vp, _u32, C.POINTER(_vp), C.POINTER(_u64)]),
    "pd_fileset_open": (C.c_int, [C.POINTER(C.c_char_p), _u32, C.POINTER(OpenOpts), C.POINTER(_vp), C.c_char_p, C.c_size_t]),
    "pd_fileset_close": (None, [_vp]),
    "pd_fileset_last_error": (C.c_char_p, [_vp]),
    "pd_fileset_file_count": (_u32, [_vp]),
    "pd_fileset_variant_ct": (_u32, [_vp]),
    "pd_fileset_sample_ct": (_u32, [_vp]),
    "pd_fileset_offsets": (C.c_int, [_vp, _vp]),
    "pd_fileset_ctx": (_vp, [_vp, _u32]),
    "pd_fileset_set_sample_subset": (C.c_int, [_vp, _vp, _u32]),
    "pd_fileset_counts": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pd_fileset_hardy": (C.c_int, [_vp, _u32, _u32, _u32, _vp, _vp]),
    "pd_fileset_decode_i8": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pd_fileset_missing_by_sample": (C.c_int, [_vp, _u32, _u32, _vp]),
    "pd_fileset_sample_counts": (C.c_int, [_vp, _u32, _u32, _vp, _vp, _vp]),
    "pd_fileset_score": (C.c_int, [_vp, _vp, _vp, _vp, _u32, _u32, _u32, _vp, _vp, _vp]),
    "pd_get_stats": (C.c_int, [_vp, C.POINTER(Stats)]),
    "pd_reset_stats": (C.c_int, [_vp]),
    "pd_version": (C.c_char_p, []),
}

_lib = None


def load():
    """Load libplinkgpu.so; raises if it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: the CUDA extension is required (no CPU fallback). "
                "Build it with `make -C plinking_duck_b200/csrc`."
            )
        # PLINKGPU_LIB: a differently built copy of the same library (A/B runs of development scripts only)
        lib = C.CDLL(os.environ.get("PLINKGPU_LIB") or LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError if the library does not export it
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class PlinkGpuError(IOError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"{ERR_NAMES.get(code, code)}: {msg}")
        self.code = code
