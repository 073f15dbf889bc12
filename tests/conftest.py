import importlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    """The product package (hyphenated directory name -> importlib)."""
    return importlib.import_module("nvidia-optix-ray-tracer_b200")


@pytest.fixture(scope="session")
def synth():
    return importlib.import_module("nvidia-optix-ray-tracer_b200.synth")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure).  Built on demand; prebuilt on the GPU box."""
    import oracle_lib
    return oracle_lib.load()


@pytest.fixture(scope="session")
def ref():
    """oracle/_ref: pieces of the unmodified reference compiled as host code (may be absent)."""
    import oracle_lib
    return oracle_lib.load_ref()


@pytest.fixture(scope="session")
def built_lib(pkg):
    """librtb200.so, built if missing (nvcc cross-compiles without a GPU)."""
    if not os.path.exists(pkg.LIB_PATH):
        pkg.build()
    return pkg.lib()
