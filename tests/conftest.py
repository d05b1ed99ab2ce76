import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "fluid-simulation_b200"))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_py
    oracle_py.build()
    return oracle_py.Oracle()


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference build; present in the build container and (as a
    prebuilt .so that travels with the snapshot) on the GPU box."""
    import oracle_py
    if not os.path.exists(oracle_py.REF_SO):
        pytest.skip("oracle/_ref/libfluid_ref.so not built (needs /root/reference)")
    return oracle_py.Ref()


@pytest.fixture(scope="session")
def golden_hashes():
    return json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))["frames"]


@pytest.fixture(scope="session")
def golden_bordered():
    """FSType::ORIGINAL_BORDERED frames of the unmodified reference (make_golden.py: bordered_frames)."""
    return json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))["bordered_frames"]


@pytest.fixture(scope="session")
def golden_mf():
    """FSType::ORIGINAL_BORDERED_MF frames of the unmodified reference, two fluids (make_golden.py: mf_frames)."""
    return json.load(open(os.path.join(GOLDEN, "golden_hashes.json")))["mf_frames"]


@pytest.fixture(scope="session")
def golden_fields():
    return np.load(os.path.join(GOLDEN, "golden_fields.npz"))


@pytest.fixture(scope="session")
def fl():
    """The product library through its ctypes view.  GPU tests must FAIL, not
    skip, when the CUDA library is missing or no device is usable."""
    import fluid_b200
    fluid_b200.lib()  # raises FluidError if libfluid_cuda.so is not built
    return fluid_b200


def rand_field(rng, N, scale=1.0):
    return (rng.uniform(-1, 1, (N + 2) ** 2) * scale).astype(np.float32)


def assert_bits(a, b, what=""):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    if a.tobytes() != b.tobytes():
        ai, bi = a.view(np.uint32).ravel(), b.view(np.uint32).ravel()
        bad = np.nonzero(ai != bi)[0]
        with np.errstate(invalid="ignore"):
            err = np.nanmax(np.abs(a.ravel()[bad].astype(np.float64) - b.ravel()[bad]))
        raise AssertionError("%s: %d of %d values differ bitwise, first at flat index %d (%r vs %r), max abs diff %g"
                             % (what, bad.size, ai.size, bad[0], a.ravel()[bad[0]], b.ravel()[bad[0]], err))
