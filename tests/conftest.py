import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "reference: needs the reference checkout at /root/reference (build container only)")


def pytest_collection_modifyitems(config, items):
    import torch
    has_gpu = torch.cuda.is_available()
    has_ref = os.path.isfile(os.path.join(os.environ.get("AMYLOID_REFERENCE_ROOT", "/root/reference"), "models.py"))
    for item in items:
        if "gpu" in item.keywords and not has_gpu:
            item.add_marker(pytest.mark.skip(reason="no CUDA device"))
        if "reference" in item.keywords and not has_ref:
            item.add_marker(pytest.mark.skip(reason="reference checkout not mounted"))


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    def load(name):
        path = os.path.join(ROOT, "tests", "golden", name)
        if name.endswith(".json"):
            import json
            return json.load(open(path))
        return np.load(path)
    return load


@pytest.fixture(scope="session", autouse=True)
def built_library():
    """Make sure the in-tree CUDA library exists (nvcc cross-compiles without a GPU)."""
    from amyloid_yolo_paper_b200 import _lib
    if os.path.exists("/usr/local/cuda/bin/nvcc"):
        _lib.build()
    return _lib.LIB_PATH
