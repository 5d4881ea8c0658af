"""CPU: the C-ABI library loads, exports every symbol include/plinkgpu.h declares, and its host-only
entry points (pd_probe: header / index parse) agree with pgenlib.  No compute calls are made."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, fixture_names, fixture_path


def _declared_symbols():
    out = []
    for h in sorted(os.listdir(os.path.join(ROOT, "include"))):
        if h.endswith(".h"):
            src = open(os.path.join(ROOT, "include", h)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            out += re.findall(r"\b(pd[a-z]*_[a-z0-9_]+)\s*\(", src)
    return sorted(set(out))


def test_header_symbols_are_bound_and_exported(built):
    from plinking_duck_b200 import _native as nat
    from plinking_duck_b200 import ops

    declared = _declared_symbols()
    assert len(declared) >= 30
    lib = C.CDLL(nat.LIB_PATH)
    bound = set(nat.SYMBOLS) | set(ops._OPS)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/*.h but not exported"
        assert name in bound, f"{name} has no ctypes binding"
    assert bound <= set(declared)
    assert nat.load().pd_version().decode().endswith("sm_100a")


def test_no_cpu_fallback(built):
    """Without a CUDA device pd_open must fail loudly (PD_ERR_CUDA), never compute on the host."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from plinking_duck_b200 import _native as nat
    from plinking_duck_b200.pgen import PgenContext

    with pytest.raises(nat.PlinkGpuError) as e:
        PgenContext(fixture_path("pgen_example"))
    assert e.value.code == nat.PD_ERR_CUDA
    assert "no CPU fallback" in str(e.value)


def test_fileset_and_noalloc_entry_points_fail_loudly_without_a_gpu(built):
    """The layers added on top of pd_open keep its contract: no device, no result -- and argument errors that need
    no device are still reported as such."""
    import ctypes as C
    import os

    import torch

    from plinking_duck_b200 import _native as nat

    lib = nat.load()
    err = C.create_string_buffer(256)
    # sizes beyond the static scratch are rejected before any device is touched
    limit = lib.pd_hwe_noalloc_max_variants()
    assert limit == 1 << 20
    rc = lib.pd_hwe_from_counts_noalloc(None, limit + 1, 0, 0, None, None, err, len(err))
    assert rc == nat.PD_ERR_INVALID_ARG and b"PD_NOALLOC_HWE_MAX_VARIANTS" in err.value
    assert lib.pd_hwe_from_counts_noalloc(None, 0, 0, 0, None, None, None, 0) == nat.PD_OK
    if torch.cuda.is_available():
        return
    path = os.fsencode(fixture_path("pgen_example"))
    arr = (C.c_char_p * 2)(path, path)
    fs = C.c_void_p()
    rc = lib.pd_fileset_open(arr, 2, None, C.byref(fs), err, len(err))
    assert rc == nat.PD_ERR_CUDA and not fs and b"no CPU fallback" in err.value
    host = (C.c_uint32 * 16)()
    out = (C.c_double * 4)()
    rc = lib.pd_hwe_from_counts_noalloc(host, 4, 0, 0, out, None, err, len(err))
    assert rc != nat.PD_OK and err.value


@pytest.mark.parametrize("name", fixture_names())
def test_probe_index_matches_pgenlib(name, golden, built, fixture_meta):
    from plinking_duck_b200.pgen import probe

    info, vt, fpos = probe(fixture_path(name))
    assert info.variant_ct == fixture_meta[name]["V"] and info.sample_ct == fixture_meta[name]["N"]
    assert np.array_equal(vt, golden[name + "/vrtypes"])
    assert np.array_equal(fpos, golden[name + "/fpos"])
    assert bool(info.has_ld) == bool(((vt & 6) == 2).any())
    assert bool(info.has_dosage) == bool((vt & 0x60).any())


def test_probe_rejects_bad_files(tmp_path, built):
    import synth
    from plinking_duck_b200 import _native as nat
    from plinking_duck_b200.pgen import probe

    with pytest.raises(nat.PlinkGpuError) as e:
        probe(str(tmp_path / "nope.pgen"))
    assert e.value.code == nat.PD_ERR_OPEN
    bad = tmp_path / "bad.pgen"
    bad.write_bytes(bytes(range(13)))
    with pytest.raises(nat.PlinkGpuError) as e:
        probe(str(bad))
    assert e.value.code == nat.PD_ERR_MALFORMED
    g = synth.gen_hwe_genotypes(10, 10, 1)
    trunc = tmp_path / "trunc.pgen"
    synth.fixed_pgen_image(g)[:-1].tofile(str(trunc))
    with pytest.raises(nat.PlinkGpuError) as e:
        probe(str(trunc))
    assert e.value.code == nat.PD_ERR_MALFORMED
    ok = tmp_path / "ok.pgen"
    synth.write_fixed_pgen(str(ok), g)
    info, vt, fpos = probe(str(ok))
    assert (info.variant_ct, info.sample_ct, info.mode, info.fixed_width) == (10, 10, 2, 3)
    assert np.array_equal(fpos, 12 + 3 * np.arange(11))


def test_ploidy_classification_matches_reference_rules(built):
    """ClassifyChromPloidy / ResolveParBounds (src/plink_common.cpp:1926-1979)."""
    from plinking_duck_b200 import ops

    assert ops.classify_ploidy("1", 100) == ops.PLOIDY_AUTOSOMAL
    assert ops.classify_ploidy("chrX", 50_000_000) == ops.PLOIDY_CHR_X
    assert ops.classify_ploidy("23", 50_000_000) == ops.PLOIDY_CHR_X
    assert ops.classify_ploidy("X", 1_000_000) == ops.PLOIDY_AUTOSOMAL          # GRCh38 PAR1
    assert ops.classify_ploidy("X", 2_781_480) == ops.PLOIDY_CHR_X
    assert ops.classify_ploidy("X", 2_700_000, "hg19") == ops.PLOIDY_CHR_X     # outside GRCh37 PAR1
    assert ops.classify_ploidy("X", 1_000_000, "none") == ops.PLOIDY_CHR_X
    assert ops.classify_ploidy("PAR1", 5) == ops.PLOIDY_AUTOSOMAL
    assert ops.classify_ploidy("chrY", 5) == ops.PLOIDY_CHR_Y and ops.classify_ploidy("24", 5) == ops.PLOIDY_CHR_Y
    for mt in ("MT", "chrM", "M", "26"):
        assert ops.classify_ploidy(mt, 5) == ops.PLOIDY_CHR_MT
    with pytest.raises(ValueError):
        ops.classify_ploidy("X", 5, "GRCh99")
    assert ops.counts_pass([1, 1, 1, 1], af_min=0.4, af_max=0.6) and not ops.counts_pass([1, 1, 1, 1], af_min=0.6)
    assert not ops.counts_pass([0, 0, 0, 4]) and ops.counts_pass([2, 1, 1, 0], ac_min=3, ac_max=3)
    # CheckGenotypeRange (src/plink_common.cpp:1465-1493)
    assert ops.genotype_range([5, 0, 0, 2], allowed=(0,)) == (True, True)
    assert ops.genotype_range([5, 1, 0, 0], allowed=(0,)) == (True, False)
    assert ops.genotype_range([5, 0, 0, 0], allowed=(1, 2)) == (False, False)
    assert ops.genotype_range([0, 0, 0, 3], allowed=(1,)) == (False, True)
    assert ops.genotype_range([0, 0, 0, 3], allowed=(1,), include_missing=True) == (True, True)
