import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from tests import oracle_binding
    return oracle_binding.load_oracle()


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference compiled into oracle/_ref (None when it has not been built)."""
    from tests import oracle_binding
    return oracle_binding.load_ref()


@pytest.fixture(scope="session")
def ctx():
    import hashmergejoin_b200 as hmj
    c = hmj.Context(0, use_torch_stream=True)
    yield c
    c.close()
