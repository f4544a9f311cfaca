import os
import sys
from pathlib import Path

# the library honours its developer knobs (FFS_CAND_CAP, FFS_FORCE_BRUTE, ... used by a few tests to
# force rare paths) only when this is set; it is read once, at the first call into the library
os.environ.setdefault("FFS_DEV_KNOBS", "1")

import pytest

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "fast-feedback-service_b200"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _have_gpu() -> bool:
    try:
        from ffs_b200 import lib
        return lib.load().ffs_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # `-m gpu` on a box without a GPU must fail loudly, not skip: only skip gpu tests
    # when they were not asked for explicitly.
    if "gpu" in (config.getoption("-m") or ""):
        return
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
