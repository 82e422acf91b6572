"""CPU-side checks of the drop-in boundary: libimdcu.so loads, exports every symbol that
include/imdcu.h declares, and refuses to work without a device (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "imdcu.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(imdcu_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from imd_b200 import capi
    L = capi.load()
    names = header_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), "missing export %s" % n
    # and the binding declares all of them
    assert set(names) == set(L._declared), set(names) ^ set(L._declared)
    assert L.imdcu_abi_version() == 3


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from imd_b200 import capi
    with pytest.raises(capi.ImdcuError) as e:
        capi.Context(0)
    assert e.value.code == -3
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_touch_oracle():
    """The product -- the library package, the C ABI header and the IMD-side drop-in with its build recipe -- must
    never import, link or call anything under oracle/ (comments may name oracle/_ref as the sibling output directory)."""
    for top in ("imd_b200", "integration", "include", "shmpi"):
        for dirpath, dirs, files in os.walk(os.path.join(ROOT, top)):
            dirs[:] = [d for d in dirs if d not in ("_bin", "__pycache__")]
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".c", "Makefile")):
                    src = open(os.path.join(dirpath, f), errors="ignore").read()
                    assert "oracle" not in src.replace("oracle/_ref", ""), os.path.join(dirpath, f)
