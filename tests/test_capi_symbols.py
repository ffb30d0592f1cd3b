"""CPU: liblazac_b200.so loads and exports every symbol include/lazac_b200.h declares.
No compute call is made (there is no GPU here); the no-GPU behaviour of lzg_init is checked:
it must fail loudly, never fall back."""
import ctypes as C
import os
import re

import pytest

from lazac_copy_number_b200 import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    with open(os.path.join(ROOT, "include", "lazac_b200.h")) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lzg_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_table_agree():
    assert declared_symbols() == sorted(name for name, _, _ in _capi.SYMBOLS)


def test_library_exports_every_declared_symbol():
    if not os.path.exists(_capi.LIB_PATH):
        _capi.build()
    lib = C.CDLL(_capi.LIB_PATH)
    for name in declared_symbols():
        assert getattr(lib, name) is not None
    _capi.load()
    assert b"sm_100a" in _capi.load().lzg_version()


def test_status_strings():
    lib = _capi.load()
    assert lib.lzg_status_string(0) == b"ok"
    # the reference's logic_error text, src/copy_number.cxx:242
    assert lib.lzg_status_string(2) == b"every child must have exactly two children"


def test_no_gpu_means_error_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from lazac_copy_number_b200.api import Context
    from lazac_copy_number_b200._capi import LazacGpuError
    with pytest.raises(LazacGpuError) as e:
        Context(0)
    assert e.value.status == 5  # LZG_ERR_NO_DEVICE


def test_product_package_does_not_touch_the_oracle():
    pkg = os.path.join(ROOT, "lazac_copy_number_b200")
    for dp, _, files in os.walk(pkg):
        if os.sep + "build" in dp:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp", "Makefile")):
                with open(os.path.join(dp, f)) as fh:
                    text = fh.read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "zcnt_oracle" not in text and "libref_shim" not in text, f
