"""CPU-side checks of the C ABI: the library builds/loads without a GPU, exports every symbol
include/nichecompass_b200.h declares, and the ctypes signature table mirrors the header."""
import ctypes
import os
import re

import pytest

from nichecompass_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "nichecompass_b200.h")


def parse_header():
    return _lib._parse_header(HEADER)


def test_header_symbols_exported_and_signatures_match():
    decls = parse_header()
    assert len(decls) >= 8
    lib = _lib.lib()
    for name, kinds in decls.items():
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
        if name in _lib.VALUE_FUNCS:
            continue
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"
        if name in _lib.HOST_FUNCS:          # host-memory management: synchronous, no stream
            continue
        assert kinds.endswith("s"), f"{name}: last argument must be the stream"
    for name in _lib.SIGNATURES:
        assert name in decls, f"{name} bound in _lib.py but not declared in the header"


def test_version_and_error_string():
    lib = _lib.lib()
    assert lib.nc_version() >= 100
    assert isinstance(lib.nc_last_error(), bytes)


def test_no_cpu_fallback():
    import torch
    with pytest.raises(_lib.NicheCompassCudaError):
        _lib.ptr(torch.zeros(4))
