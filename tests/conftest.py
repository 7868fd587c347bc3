"""Shared fixtures.  GPU tests are marked `gpu`; everything else runs on CPU.

oracle/ is test infrastructure: these tests (and only tests/, smoke() and the
bench's CPU-baseline legs) may import it.
"""
import importlib
import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "oracle"))

REFERENCE_PRESENT = os.path.isdir("/root/reference/oni-ml/oni-lda-c")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pkg():
    """The product package (its directory name has hyphens, so import by string)."""
    return importlib.import_module("open-network-insight_b200")


@pytest.fixture(scope="session")
def oni():
    m = pkg()
    if not m.LIB_PATH.exists() or not m.LDA_BIN.exists():
        m.build()            # a fresh checkout: nvcc cross-compiles without a GPU (same as __graft_entry__.build())
    return m


@pytest.fixture(scope="session")
def oracle():
    import oracle_py
    oracle_py.build()
    return oracle_py


@pytest.fixture(scope="session")
def corpusgen():
    return importlib.import_module("open-network-insight_b200.corpusgen")


needs_reference = pytest.mark.skipif(not REFERENCE_PRESENT, reason="/root/reference not present on this box")
