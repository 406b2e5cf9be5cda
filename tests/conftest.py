import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def has_gpu() -> bool:
    try:
        from stereo_visual_odometry_b200 import _lib
        return _lib.lib().svo_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def hostcores():
    """Host-compiled arithmetic cores (tests/host_cores.cpp) as a ctypes library."""
    import ctypes
    out = os.path.join(ROOT, "tests", "_build", "libsvo_hostcores.so")
    src = os.path.join(ROOT, "tests", "host_cores.cpp")
    hdr = os.path.join(ROOT, "stereo_visual_odometry_b200", "csrc", "svo_core.cuh")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wno-unknown-pragmas", "-o", out, src],
                       check=True)
    return ctypes.CDLL(out)


@pytest.fixture(scope="session")
def run1():
    """The reference's run1 sequence as committed fixtures (tests/golden/make_golden.py): dict(bgr=[(L,R)] and
    xyz=golden rows for frames 0..7 as the CLI reads them, gray=[(L,R)] for frames 0..47 as 8UC1)."""
    import cv2
    z = np.load(os.path.join(ROOT, "tests", "golden", "run1_first8.npz"))
    bgr, gray = [], []
    for i in range(8):
        L = cv2.imdecode(z[f"left_{i}"], cv2.IMREAD_COLOR)
        R = cv2.imdecode(z[f"right_{i}"], cv2.IMREAD_COLOR)
        bgr.append((L, R))
        gray.append((cv2.cvtColor(L, cv2.COLOR_BGR2GRAY), cv2.cvtColor(R, cv2.COLOR_BGR2GRAY)))
    z2 = np.load(os.path.join(ROOT, "tests", "golden", "run1_gray_8_47.npz"))
    for i in range(8, 48):
        gray.append((cv2.imdecode(z2[f"left_{i}"], cv2.IMREAD_GRAYSCALE), cv2.imdecode(z2[f"right_{i}"], cv2.IMREAD_GRAYSCALE)))
    return dict(bgr=bgr, gray=gray, xyz=z["result_xyz"])


def add_triangle(img, x, y, r):
    """addTriangle, src/main.cpp:89-100 of the reference."""
    for i in range(-r, r + 1):
        for j in range(0, r + 1):
            if abs(i) <= r - abs(j):
                img[i + y, j + x] = 120
    img[y, x] = 0
