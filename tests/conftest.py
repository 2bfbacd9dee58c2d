import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def nrb():
    import importlib
    return importlib.import_module("neural-renderer_b200")


@pytest.fixture(scope="session")
def oracle_lib():
    """The CPU oracle: the reference's own headers compiled for the host when that build is present
    (oracle/_ref, built where /root/reference exists and shipped as a binary), else our restatement."""
    from oracle import loader
    if not (loader.have("ref") or loader.have("port")):
        import __graft_entry__ as g
        g.build_oracle()
    return loader.oracle_lib("auto")
