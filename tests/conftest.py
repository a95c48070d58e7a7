import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    """The product package (its directory name carries the reference's hyphen, hence importlib)."""
    import __graft_entry__ as g
    if not os.path.exists(os.path.join(ROOT, "mitsuba3-amvpt_b200", "libamvpt_b200.so")):
        g.build()
    return importlib.import_module("mitsuba3-amvpt_b200")


@pytest.fixture(scope="session")
def scenes(pkg):
    return importlib.import_module("mitsuba3-amvpt_b200.scenes")


@pytest.fixture(scope="session")
def ob():
    import oracle_binding
    oracle_binding.build_oracle()
    return oracle_binding


@pytest.fixture(scope="session")
def gpu(pkg):
    if pkg.device_count() < 1:
        pytest.fail("the gpu-marked tests need a CUDA device; the product has no CPU fallback")
    return 0
