"""CPU-side checks of the drop-in boundary: the C-ABI library builds/loads and exports every symbol that
include/dl4jb200.h declares, and the shape arithmetic entry point (which needs no GPU) matches the oracle."""
import os
import re

import pytest

from conftest import ROOT


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "dl4jb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dl4jb200_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from deeplearning4j_b200 import build
    build.build()
    from deeplearning4j_b200 import _lib
    return _lib


def test_header_symbols_exported_and_bound(lib):
    syms = _declared_symbols()
    assert len(syms) >= 35
    dll = lib.cdll()
    for s in syms:
        assert hasattr(dll, s), f"{s} declared in include/dl4jb200.h but not exported"
    # every declared entry point has a ctypes prototype (and nothing is bound that the header lacks)
    assert sorted(lib.exported_names()) == syms


def test_out_size_matches_oracle_and_reference_errors(lib, oracle):
    for mode in (0, 1, 2):
        for inp in (1, 3, 4, 7, 9, 10, 11, 28, 224, 230):
            for k in (1, 2, 3, 5, 7, 11):
                for s in (1, 2, 3, 4, 7):
                    for p in (0, 1, 2, 3):
                        for d in (1, 2):
                            try:
                                exp = oracle.out_size(inp, k, s, p, d, mode)
                            except oracle.DL4JInvalidInputException:
                                with pytest.raises(lib.DL4JInvalidInputException):
                                    lib.out_size(inp, k, s, p, d, mode)
                                continue
                            except oracle.DL4JInvalidConfigException:
                                with pytest.raises(lib.DL4JInvalidConfigException):
                                    lib.out_size(inp, k, s, p, d, mode)
                                continue
                            assert lib.out_size(inp, k, s, p, d, mode) == exp


def test_missing_library_fails_loudly(lib, monkeypatch):
    monkeypatch.setattr(lib, "_cdll", None)
    monkeypatch.setattr(lib, "LIB_PATH", "/nonexistent/libdl4jb200.so")
    with pytest.raises(ImportError):
        lib.cdll()


def test_no_gpu_means_error_not_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from deeplearning4j_b200 import helpers
    with pytest.raises(Exception):
        helpers.Context(0)


def test_java_binding_names_exist_in_the_header():
    """java/ cannot be compiled in this image (no JDK); at least every native method Dl4jB200.java declares must be an entry
    point include/dl4jb200.h declares, so the reference-side glue cannot drift from the C ABI unnoticed."""
    java = open(os.path.join(ROOT, "java", "org", "deeplearning4j", "b200", "Dl4jB200.java")).read()
    natives = set(re.findall(r"native\s+[^;(]*?\b(dl4jb200_[a-z0-9_]+)\s*\(", java))
    assert len(natives) >= 45
    missing = sorted(natives - set(_declared_symbols()))
    assert not missing, missing
