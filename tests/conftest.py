import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting_cc-comments_b200"), os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


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


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Everything the tests load is built in-tree first (no-op when up to date)."""
    import __graft_entry__ as G
    lib = os.path.join(G.PKG, "libgsr_b200.so")
    ext = os.path.join(G.PKG, "diff_gaussian_rasterization", "_C.so")
    if not (os.path.exists(lib) and os.path.exists(ext)):
        G.build()
    from oracle import gs_oracle, gs_next_oracle
    gs_oracle.build()
    gs_next_oracle.build()
    yield
