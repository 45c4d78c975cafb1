"""pytest configuration: the `gpu` marker and shared fixtures.

`-m "not gpu"`  : oracle vs golden vectors, host logic, C-ABI load/export checks (no CUDA device needed)
`-m gpu`        : parity tests proper -- the CUDA path, called through the C ABI, against the oracle
"""
from __future__ import annotations

import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

os.environ.setdefault("MASTER_ADDR", "127.0.0.1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _cuda_ok() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _cuda_ok():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def lib():
    """The built C-ABI library (built on demand: nvcc cross-compiles without a GPU)."""
    from sglang_custom_b200 import _lib, build
    if not _lib.LIB_PATH.exists():
        build.build()
    return _lib.load()


@pytest.fixture(scope="session")
def device():
    import torch
    return torch.device("cuda:0")
