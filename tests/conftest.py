import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Native artefacts are built in-tree by __graft_entry__.build(); build them if a fresh checkout lacks them."""
    from cpp_volume_rendering_b200 import _capi, inputs
    from oracle import pyoracle
    if not (os.path.exists(_capi.LIB_PATH) and os.path.exists(inputs.HOSTLIB_PATH) and os.path.exists(pyoracle.LIB_PATH)):
        import __graft_entry__
        __graft_entry__.build()


@pytest.fixture(scope="session")
def bonsai_table():
    from cpp_volume_rendering_b200 import inputs
    return inputs.tf_build_linear(inputs.bonsai_tf())


@pytest.fixture(scope="session")
def ctx():
    from cpp_volume_rendering_b200 import _capi
    c = _capi.Context(0)
    yield c
    c.close()
