"""Runs the C++ acceptance program built against the drop-in header include/vo.h (tests/cpp)."""
import os
import subprocess

import pytest

from conftest import ROOT, has_gpu

EXE = os.path.join(ROOT, "tests", "cpp", "test_vo_dropin")


def build():
    subprocess.run(["make", "-C", os.path.join(ROOT, "tests", "cpp"), "all"], check=True, capture_output=True)


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


STREAM_EXE = os.path.join(ROOT, "tests", "cpp", "test_vo_stream")


def test_streaming_wrapper_host_parts(tmp_path):
    """include/vo_stream.h (the stand-in for the reference's ROS node, src/stereo_vo.cpp, and its debug overlays):
    approximate-time pairing, MONO8 conversion, both calibration key styles, cv::circle overlay -- host only."""
    build()
    r = subprocess.run([STREAM_EXE, "host", str(tmp_path)], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "ALL TESTS PASS" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_streaming_wrapper_node_on_gpu(tmp_path):
    build()
    r = subprocess.run([STREAM_EXE, "gpu", str(tmp_path)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "TEST NODE" in r.stdout and "ALL TESTS PASS" in r.stdout, r.stdout + r.stderr
