import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the checkers (oracle/) and the product library once per session; both are no-ops when fresh.
    oracle/_ref is only (re)built where /root/reference exists; elsewhere the prebuilt .so is used."""
    import oracle
    oracle.build()
    from raytracing_engine_b200 import build
    build.build_library()
    yield


@pytest.fixture(scope="session")
def golden():
    import json
    import numpy as np
    g = os.path.join(ROOT, "tests", "golden")
    with open(os.path.join(g, "golden.json")) as fh:
        meta = json.load(fh)
    imgs = np.load(os.path.join(g, "images.npz"))
    return meta, imgs


@pytest.fixture(scope="session")
def case_table():
    from cases import cases
    return {c[0]: c for c in cases()}


@pytest.fixture(scope="session")
def shade_golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "shade_golden.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="session")
def shade_case_table():
    from shade_cases import shade_cases
    return shade_cases()
