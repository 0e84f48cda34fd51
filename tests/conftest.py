import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


@pytest.fixture(scope="session")
def cpu():
    """The CPU checker: the real reference build when present, else the C port."""
    import cpu_libs
    return cpu_libs.best_cpu()


@pytest.fixture(scope="session")
def model():
    """(rates[8], T[4]) of the reference's test/test.param, rates scaled."""
    import cpu_libs
    from synth import TEST_PARAMS
    rates, T, _, _ = cpu_libs.oracle().model_from_params(*TEST_PARAMS)
    return rates, T


@pytest.fixture()
def ctx():
    from epievo_b200.device import Context
    c = Context(0)
    yield c
    c.close()
