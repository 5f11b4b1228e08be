"""The C++ facade (hashmergejoin_b200/facade/*.h: the reference's header names, template
parameters and argument order over the C ABI) against the reference's own unit-test cases,
restated in tests/cpp/facade_test.cc."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "build", "facade_test")


def test_facade_headers_compile_against_the_c_abi():
    # host compiler only; links libhmj.so.  (the GPU box runs the prebuilt binary)
    subprocess.run(["make", "-C", ROOT, "lib", "tests"], check=True, capture_output=True)
    assert os.path.exists(BIN)


def test_facade_without_a_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = subprocess.run([BIN], capture_output=True, text=True)
    assert r.returncode != 0 and "hmj_create failed" in r.stdout, "no CPU fallback may exist"


@pytest.mark.gpu
def test_facade_reference_cases_on_gpu():
    assert os.path.exists(BIN), "build/facade_test missing: run `make tests` (graft build())"
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failures" in r.stdout
