"""Runs the C++ acceptance program built against the drop-in header include/vo.h (tests/cpp)."""
import os
import subprocess

import pytest

from conftest import ROOT, has_gpu

EXE = os.path.join(ROOT, "tests", "cpp", "test_vo_dropin")


def build():
    subprocess.run(["make", "-C", os.path.join(ROOT, "tests", "cpp")], check=True, capture_output=True)


def test_dropin_header_compiles_without_opencv():
    build()
    assert os.path.exists(EXE)


@pytest.mark.skipif(has_gpu(), reason="only meaningful without a GPU")
def test_dropin_fails_loudly_without_gpu():
    build()
    r = subprocess.run([EXE], capture_output=True, text=True)
    assert r.returncode != 0 and "ALL TESTS PASS" not in r.stdout


@pytest.mark.gpu
def test_reference_known_answers_through_cpp_header():
    build()
    r = subprocess.run([EXE], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ALL TESTS PASS" in r.stdout
