import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_sessionstart(session):
    """The library is a build artefact (git-ignored): a fresh checkout has none.  Build it once (nvcc cross-compiles
    without a GPU, ~1 min) instead of failing the first test that loads it; __graft_entry__.build() is the same call
    the driver makes and also places oracle/_ref when the reference tree is present."""
    lib = os.path.join(ROOT, "video_style_transfer_b200", "libvst_b200.so")
    if not os.path.exists(lib):
        import __graft_entry__
        __graft_entry__.build()


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
