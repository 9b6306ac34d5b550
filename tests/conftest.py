import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_gpu() -> bool:
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import columnar

    columnar.build()
    return columnar


@pytest.fixture(scope="session")
def ctx():
    """Process-wide igpu context; fails loudly when the CUDA library is missing."""
    from headlamp_intel_gpu_plugin_b200 import native

    return native.default_context()


@pytest.fixture
def switch(monkeypatch, ctx):
    """Sets an IGPU_* switch for one test: the library reads those when a context is created, so the shared context is told to
    read them again now and once more after the test, when monkeypatch has restored the environment."""
    def set_(name, value="1"):
        monkeypatch.setenv(name, value)
        ctx.reload_env()

    yield set_
    monkeypatch.undo()
    ctx.reload_env()
