import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(autouse=True)
def _default_hooks(request):
    """GPU tests start and end on the product configuration: tensor-core path, all stages, default lazy-max threshold
    (the hooks are thread-local in the library; pytest runs the tests on one thread)."""
    if "gpu" not in request.keywords:
        yield
        return
    import haet_b200 as hb
    l = hb.lib()

    def reset():
        l.haet_set_path(0)
        l.haet_set_stage_mask(15)
        l.haet_set_raise_threshold(0.0)
    reset()
    yield
    reset()
