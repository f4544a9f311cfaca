"""CPU-only: the C-ABI library loads and exports every symbol include/ffs_b200.h declares;
host-side logic that needs no GPU.  No compute call is made here."""
import re
from pathlib import Path

import pytest

from ffs_b200 import lib
from ffs_b200.spotfinder import parse_algorithm

ROOT = Path(__file__).resolve().parents[1]


def declared_symbols():
    text = (ROOT / "include" / "ffs_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = set(re.findall(r"\b((?:ffs|spotfinder)_[a-z0-9_]+)\s*\(", text))
    return sorted(names)


def test_every_declared_symbol_is_exported_and_bound():
    L = lib.load()
    names = declared_symbols()
    assert len(names) >= 35
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/ffs_b200.h but not exported"
        assert n in lib.SYMBOLS, f"{n} has no ctypes prototype in ffs_b200/lib.py"


def test_no_gpu_means_loud_failure_not_fallback():
    L = lib.load()
    if L.ffs_device_count() > 0:
        pytest.skip("a GPU is present")
    from ffs_b200.spotfinder import Spotfinder
    with pytest.raises(lib.FfsError) as e:
        Spotfinder(64, 64)
    assert e.value.status == -3  # FFS_ERR_NO_DEVICE


def test_algorithm_names():
    assert parse_algorithm("dispersion") == 0
    assert parse_algorithm("Dispersion_Extended") == 1
    with pytest.raises(ValueError):
        parse_algorithm("sobel")


def test_product_never_imports_oracle():
    pkg = ROOT / "fast-feedback-service_b200"
    for f in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu*")) + list(pkg.rglob("*.inl")) + list(pkg.rglob("*.cc")):
        txt = f.read_text(errors="ignore")
        assert "oracle" not in txt.replace("the oracle", "").replace("CPU oracle", "").replace("oracle's", "") \
            or f.name == "synth.py", f
