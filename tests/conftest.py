"""pytest configuration: the `gpu` marker, and session fixtures that build the TEST-ONLY checkers (oracle/, tests/sim)."""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


def _have_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container (run with gpurun)")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def lib():
    """libspeq_cuda.so, built in-tree by nvcc (cross-compiles without a GPU)."""
    from speq_b200 import build, capi
    build.build_lib()
    return capi.lib()


@pytest.fixture(scope="session")
def oracle_bin() -> Path:
    import helpers
    return helpers.oracle_binary()


@pytest.fixture(scope="session")
def ref_bin():
    """oracle/_ref/speq_ref: the same glue over the REAL vendored sdsl-lite objects (needs /root/reference to build;
    prebuilt copies travel to the GPU box).  None when neither is available."""
    import helpers
    return helpers.ref_binary()


@pytest.fixture(scope="session")
def sim():
    import helpers
    return helpers.sim_library()
