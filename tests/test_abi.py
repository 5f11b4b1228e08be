"""The C-ABI library: it loads here (no GPU), exports every function include/hmj.h declares,
and refuses to create a context without a device (no CPU fallback).  CPU only."""
import ctypes as C
import os
import re

import pytest

import hashmergejoin_b200 as hmj
from hashmergejoin_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "hmj.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hmj_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_what_binding_lists():
    assert declared_functions() == sorted(_lib.SIGNATURES)


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(hmj.LIB_PATH)
    for name in declared_functions():
        assert hasattr(lib, name), name


def test_version_and_host_helpers_without_gpu():
    lib = hmj.load()
    assert b"sm_100a" in lib.hmj_version()
    # optimal_partition is a host scalar (radix_hash.h:38-57): 1M->10, 64M->13, 1B->10
    assert hmj.optimal_partition(10 ** 6) == 10
    assert hmj.optimal_partition(64 * 10 ** 6) == 13
    assert hmj.optimal_partition(10 ** 9) == 10
    assert hmj.optimal_partition(5) == 6


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = hmj.load()
    h = C.c_void_p()
    assert lib.hmj_create(C.byref(h), 0) == _lib.HMJ_ENODEVICE
    with pytest.raises(hmj.HmjError):
        hmj.Context(0)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "hashmergejoin_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_binding" not in text and "hmj_oracle" not in text and "libhmj_ref" not in text, f
