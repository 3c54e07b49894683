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
def oracle():
    from oracle import orc
    orc.build()
    return orc.Oracle("port")


@pytest.fixture(scope="session")
def ref_oracle():
    """The reference's own CPU code (only where oracle/_ref/libhako_ref.so exists)."""
    from oracle import orc
    if not orc.ref_available():
        pytest.skip("oracle/_ref/libhako_ref.so not built (no /root/reference on this machine)")
    return orc.Oracle("reference")


@pytest.fixture(scope="session")
def golden():
    import json
    with open(os.path.join(GOLDEN, "golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def small_cases():
    return np.load(os.path.join(GOLDEN, "small_cases.npz"))


@pytest.fixture(scope="session")
def bunny():
    """(tris[T,3,3], colours, emissions) -- the bunny fixture with the deterministic vertex attributes
    make_golden.py used."""
    from scenes import bunny_scene
    return bunny_scene()
