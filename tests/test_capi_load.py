"""The C-ABI library loads and exports every symbol include/svo.h declares; without a CUDA device
every compute entry point fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, has_gpu
from stereo_visual_odometry_b200 import _lib


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "svo.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(svo_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_exported():
    L = _lib.lib()
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/svo.h but not exported"
    assert sorted(_lib.SYMBOLS) == names


def test_struct_layout_matches_header():
    p = _lib.default_params(1241, 376)
    assert (p.width, p.height, p.buckets_along_height, p.buckets_along_width) == (1241, 376, 92, 160)
    assert (p.bucket_start_row, p.features_per_bucket, p.features_threshold) == (4, 1, 15)
    assert (p.fast_threshold, p.age_threshold, p.win_size, p.max_level, p.lk_max_iters) == (20, 20, 10, 3, 30)
    assert abs(p.circular_threshold - 0.15) < 1e-7 and p.ransac_iterations == 100
    assert p.max_translation_norm == 0.1 and p.max_rotation_norm == 0.5
    assert C.sizeof(_lib.FrameResult) == 8 * 4 + 28 * 8


def test_inverse_transform_host():
    from stereo_visual_odometry_b200 import getInverseTransform
    import cv2
    R, _ = cv2.Rodrigues(np.array([0.1, -0.2, 0.05]))
    t = np.array([0.3, -0.1, 0.7])
    T = getInverseTransform(R, t)
    M = np.eye(4)
    M[:3, :3], M[:3, 3] = R, t
    assert np.abs(T - np.linalg.inv(M)).max() < 1e-14


@pytest.mark.skipif(has_gpu(), reason="only meaningful without a GPU")
def test_no_cpu_fallback():
    from stereo_visual_odometry_b200 import SvoError, VisualOdometry, featureDetectionFast
    img = np.zeros((64, 64), np.uint8)
    with pytest.raises(SvoError):
        featureDetectionFast(img, 20)
    with pytest.raises(SvoError):
        VisualOdometry(64, 64)


def test_invalid_parameters_are_rejected_before_any_device_work():
    """svo_create validates its parameters first: these fail with SVO_ERR_INVALID messages with or without a GPU."""
    import ctypes as C
    from stereo_visual_odometry_b200._lib import default_params, lib
    L = lib()
    for kw, needle in [(dict(width=8), "image size"), (dict(n_seq=0), "n_seq"), (dict(channels=2), "channels"),
                       (dict(win_size=21), "winSize"), (dict(max_level=7), "max_level"), (dict(fast_threshold=0), "fast_threshold"),
                       (dict(ransac_hypotheses=0), "ransac_hypotheses"), (dict(buckets_along_width=0), "bucket")]:
        p = default_params(640, 480)
        for k, v in kw.items():
            setattr(p, k, v)
        ctx = L.svo_create(C.byref(p))
        assert not ctx, kw
        assert needle in L.svo_last_error().decode(), (kw, L.svo_last_error())
