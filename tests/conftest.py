"""pytest configuration: the `gpu` marker and the repo root on sys.path."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with `-m gpu` on a B200)")


@pytest.fixture(scope="session", autouse=True)
def _built_libraries():
    """The shared objects are built in-tree by `make` / __graft_entry__.build(); build
    them on demand when a fresh checkout runs the CPU suite."""
    from simple_raytracer_b200 import abi
    need = [os.path.join(abi.LIB_DIR, "librt_host.so"), os.path.join(abi.LIB_DIR, "librt_b200.so"),
            os.path.join(ROOT, "oracle", "liboracle.so")]
    if not all(os.path.exists(p) for p in need):
        import __graft_entry__
        __graft_entry__.build()
