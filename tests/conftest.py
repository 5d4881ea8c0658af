import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
FIX_DIR = os.path.join(GOLDEN_DIR, "fixtures")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(GOLDEN_DIR, "golden.npz"))


@pytest.fixture(scope="session")
def fixture_meta():
    with open(os.path.join(GOLDEN_DIR, "fixtures.json")) as f:
        return json.load(f)


def fixture_names():
    with open(os.path.join(GOLDEN_DIR, "fixtures.json")) as f:
        return sorted(json.load(f).keys())


def fixture_path(name):
    return os.path.join(FIX_DIR, name + ".pgen")


def ensure_built():
    """Build libplinkgpu.so (and, where the reference tree exists, oracle/_ref) if absent.
    nvcc cross-compiles for sm_100a without a GPU, so this also works on the CPU-only box."""
    import subprocess

    lib = os.path.join(ROOT, "plinking_duck_b200", "libplinkgpu.so")
    if not os.path.exists(lib):
        subprocess.run(["make", "-j8"], cwd=os.path.join(ROOT, "plinking_duck_b200", "csrc"), check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.STDOUT)
    ref_so = os.path.join(ROOT, "oracle", "_ref", "libpgenref.so")
    if not os.path.exists(ref_so) and os.path.isdir("/root/reference/third_party/plink-ng/2.0/include"):
        subprocess.run(["make", "-j8", "ref"], cwd=os.path.join(ROOT, "oracle"), check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.STDOUT)


@pytest.fixture(scope="session")
def built():
    ensure_built()
    return True


@pytest.fixture(scope="session")
def refmod(built):
    """oracle.ref, or skip when oracle/_ref/libpgenref.so could not be built (no reference tree)."""
    from oracle import ref

    if not ref.available():
        pytest.skip("oracle/_ref/libpgenref.so not available")
    return ref


def need_devices(kw):
    """Skips a GPU test whose context options ask for more devices than the box has (devices=[0, 1] on one GPU)."""
    devs = (kw or {}).get("devices")
    if devs:
        import torch

        if torch.cuda.device_count() <= max(devs):
            pytest.skip(f"needs {max(devs) + 1} GPUs")
