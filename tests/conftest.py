"""Shared fixtures.  `-m "not gpu"`: oracle vs golden vectors / brute force / closed forms, host logic, ABI export
check (runs on CPU).  `-m gpu`: parity of the CUDA path (through the C ABI) against the oracle on a B200."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    from oracle import pyoracle
    pyoracle.lib()
    return pyoracle


@pytest.fixture(scope="session")
def S():
    from raytracerfacility_b200 import structs
    return structs


@pytest.fixture(scope="session")
def scenes():
    from raytracerfacility_b200 import scenes as sc
    return sc


@pytest.fixture(scope="session")
def mod():
    from raytracerfacility_b200 import module
    return module


@pytest.fixture()
def rt(mod):
    """A fresh context on cuda:0.  Raises (test fails loudly) when the CUDA library or the GPU is missing."""
    r = mod.RayTracer(0)
    yield r
    r.destroy()


def bits(a):
    """float32 array -> uint32 view for bit-exact comparisons (NaN-safe, -0 != +0)."""
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def assert_bit_equal(a, b, what=""):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    if a.dtype.kind == "f":
        a, b = a.view(np.uint32), b.view(np.uint32)
    if a.dtype.names:
        a, b = a.view(np.uint8), b.view(np.uint8)
    bad = np.nonzero(a != b)[0]
    assert bad.size == 0, f"{what}: {bad.size} mismatches, first at {bad[:5]}"
