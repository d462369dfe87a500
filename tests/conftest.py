"""pytest plumbing: the `gpu` marker, the package loader (directory name has a hyphen), the
oracle and the CPU-model builds, seeded input generators shared by the parity tests."""
import ctypes
import importlib.util
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_pkg():
    name = "gnark_fork_b200"
    if name in sys.modules:
        return sys.modules[name]
    d = os.path.join(ROOT, "gnark-fork_b200")
    spec = importlib.util.spec_from_file_location(name, os.path.join(d, "__init__.py"), submodule_search_locations=[d])
    m = importlib.util.module_from_spec(spec)
    sys.modules[name] = m
    spec.loader.exec_module(m)
    return m


@pytest.fixture(scope="session")
def pkg():
    return load_pkg()


@pytest.fixture(scope="session")
def cref():
    from oracle import cref as c
    c.lib()
    return c


@pytest.fixture(scope="session")
def model():
    d = os.path.join(ROOT, "tests", "cpu_model")
    so = os.path.join(d, "libmodel.so")
    src = os.path.join(d, "model.cpp")
    hdrs = [os.path.join(ROOT, "gnark-fork_b200", "csrc", h) for h in ("ntt.cuh", "msm.cuh", "field.cuh", "ec.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(p) > os.path.getmtime(so) for p in [src] + hdrs):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-I/usr/local/cuda/include", "-o", so, src])
    return ctypes.CDLL(so)


@pytest.fixture(scope="session")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()


TOP_LIMB = 0x30644E72E131A029  # top u64 limb of both BN254 moduli


def rand_field(rng: np.random.Generator, n: int) -> np.ndarray:
    """n pseudo-random reduced field elements (valid for Fp and Fr), taken as Montgomery limbs."""
    a = rng.integers(0, 1 << 63, size=(n, 4), dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=(n, 4), dtype=np.uint64)
    a[:, 3] = rng.integers(0, TOP_LIMB, size=n, dtype=np.uint64)
    return a
