import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parents[1]
for p in (str(REPO), str(REPO / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _native_built():
    # The CUDA library is built in-tree (nvcc cross-compiles without a GPU); tests never run
    # against a missing library.
    from ugu_b200 import build
    build.build_native()
    from oracle import build_port
    build_port()
