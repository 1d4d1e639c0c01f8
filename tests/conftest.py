"""Shared fixtures.  GPU tests are marked @pytest.mark.gpu; everything else runs on CPU.

The oracle (oracle/) is used here as the CHECKER only; the product path under test is the CUDA
library reached through the C ABI (simplepathtracer_b200.PathTracer -> include/spt_cuda.h).
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import Oracle, build_oracle

    build_oracle(ref=True)
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    from oracle import Reference, ref_available

    if not ref_available():
        pytest.skip("oracle/_ref/libspt_ref.so not built (needs /root/reference at build time)")
    return Reference()


@pytest.fixture(scope="session")
def reference_wide():
    from oracle import Reference, ref_available

    if not ref_available(wide=True):
        pytest.skip("oracle/_ref/libspt_ref_wide.so not built")
    return Reference(wide=True)


@pytest.fixture(scope="session")
def tracer():
    import simplepathtracer_b200 as spt

    t = spt.PathTracer(0)  # raises loudly if the CUDA library or a device is missing
    yield t
    t.close()


@pytest.fixture(scope="session")
def cam(oracle):
    return oracle.camera()  # the reference camera (Globals.hpp:22-24)


@pytest.fixture(scope="session")
def scene_init(oracle):
    return oracle.init_spheres(1)


@pytest.fixture(scope="session")
def scene_random(oracle):
    return oracle.generate_spheres(1)


def golden_path(name):
    return os.path.join(GOLDEN, name)
