import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def built_lib():
    """Build (if stale) and load the C-ABI library; CPU-only machines can still dlopen it."""
    from neuralstyletransferv1_b200.build import build_lib
    build_lib()
    from neuralstyletransferv1_b200 import _lib
    return _lib.load()


def golden_names():
    """fixtures of the three generator families (the post-network fixtures have their own tests)"""
    return sorted(f[:-4] for f in os.listdir(GOLDEN) if f.endswith(".npz") and f.split("_")[0] in ("johnson", "nst", "reconet"))
