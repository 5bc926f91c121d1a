"""CPU: the C-ABI library loads and exports every symbol include/*.h declares;
without a GPU it fails loudly instead of falling back."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(dc_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(pkg):
    L = pkg.lib()
    names = declared_symbols("depthconv_b200.h")
    assert len(names) >= 15
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/depthconv_b200.h but not exported"
    assert sorted(pkg.EXPORTED_SYMBOLS) == names
    assert b"sm_100a" in L.dc_version()


def test_no_cpu_fallback_without_gpu(pkg):
    L = pkg.lib()
    if L.dc_device_count() > 0:
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = L.dc_create(0, C.byref(h))
    assert rc == pkg.DC_ERR_NO_DEVICE and not h
    assert b"no CPU fallback" in L.dc_last_error(None)
    with pytest.raises(pkg.DcError):
        pkg.Context(0)


def test_null_arguments_are_rejected(pkg):
    L = pkg.lib()
    assert L.dc_create(0, None) == pkg.DC_ERR_INVALID
    assert L.dc_set_pinhole(None, 1, 1, 1.0, 1.0, 1.0, 1.0) == pkg.DC_ERR_INVALID
    assert L.dc_convert_host(None, None) == pkg.DC_ERR_INVALID
    assert L.dc_width(None) == 0
    L.dc_destroy(None)
