import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    import cpu_libs
    return cpu_libs.load_oracle()


@pytest.fixture(scope="session")
def ref():
    import cpu_libs
    lib = cpu_libs.load_ref()
    if lib is None:
        pytest.skip("oracle/_ref/libref_lora_phy.so not built (needs /root/reference)")
    return lib
