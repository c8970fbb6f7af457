"""pytest configuration: the `gpu` marker and session fixtures that make sure the native libraries exist."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def native_libs():
    """Builds libptscene.so / libpt.so / libptoracle.so in-tree if they are missing or stale (incremental)."""
    from opengl_ray_tracing_b200 import build as B
    B.build_all()
    from oracle import pyoracle
    pyoracle.build()
    return True


@pytest.fixture(scope="session")
def golden():
    import json
    import numpy as np
    g = os.path.join(ROOT, "tests", "golden")
    with open(os.path.join(g, "kats.json")) as f:
        kats = json.load(f)
    return kats, np.load(os.path.join(g, "images.npz"))


def have_ref():
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_host.so"))


def have_reference_tree():
    return os.path.isdir(os.environ.get("PT_REFERENCE_DIR", "/root/reference"))
