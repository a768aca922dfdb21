import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "ref: needs oracle/_ref (the compiled reference); skipped when absent")


@pytest.fixture(scope="session")
def orc():
    import oracle
    return oracle.lib()


@pytest.fixture(scope="session")
def refimpl():
    import oracle
    r = oracle.ref()
    if r is None:
        pytest.skip("oracle/_ref/libppo_ref.so not built (needs /root/reference); "
                    "golden fixtures in tests/golden cover the same cases")
    return r


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    path = os.path.join(ROOT, "tests", "golden")

    def load(name):
        return np.load(os.path.join(path, name + ".npz"))
    return load
