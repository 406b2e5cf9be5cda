"""Pins the CPU oracle (oracle/ref_cv2.py) against every golden vector the reference holds for
the hot path: run1/result.csv and the known answers of the eight assert-tests in
src/main.cpp:50-264."""
import os

import numpy as np
import pytest

from conftest import add_triangle
from oracle import ref_cv2 as O

REF_RUN1 = "/root/reference/run1"


def test_run1_result_csv_first8(run1):
    """run1/result.csv rows 0..7 (CLI-faithful 8UC3 mode, identity initial pose: SURVEY F3/F4)."""
    xyz, oks, _ = O.run_sequence(iter(run1["bgr"]), O.CLI_P_LEFT, O.CLI_P_RIGHT, mode="cli_bgr", pose0=np.eye(4))
    assert np.abs(xyz - run1["xyz"]).max() < 2e-5   # 6 significant digits in the CSV


@pytest.mark.skipif(not os.path.isdir(REF_RUN1), reason="reference checkout not present (GPU box)")
def test_run1_result_csv_full():
    import cv2

    def frames():
        for i in range(128):   # right frame 128 is missing in the reference
            yield (cv2.imread(f"{REF_RUN1}/left/frame{i:06d}.png"), cv2.imread(f"{REF_RUN1}/right/frame{i:06d}.png"))

    xyz, oks, _ = O.run_sequence(frames(), O.CLI_P_LEFT, O.CLI_P_RIGHT, mode="cli_bgr", pose0=np.eye(4))
    gold = np.loadtxt(f"{REF_RUN1}/result.csv", delimiter=",", skiprows=1)[:128, :3]
    assert np.abs(xyz - gold).max() < 2e-5
    # today's CLI starts from a 26 degree pitch (main.cpp:368-373): same trajectory rotated
    xyz26, _, _ = O.run_sequence(frames(), O.CLI_P_LEFT, O.CLI_P_RIGHT, mode="cli_bgr", pose0=O.initial_pose("pitch26"))
    P0 = O.initial_pose("pitch26")[:3, :3]
    assert np.abs(xyz26 - gold @ P0.T).max() < 2e-5


def test_bucket_empty():   # main.cpp:50-60
    b = O.Bucket(0)
    for a, s in zip([6, 2, 3, 4, 5], [60, 20, 30, 40, 50]):
        b.add_feature((1, 1), a, s)
    assert b.max_size == 0 and b.size() == 0


def test_bucket_nonempty():   # main.cpp:62-78
    b = O.Bucket(3)
    for a, s in zip([6, 2, 3, 4, 5], [60, 20, 30, 40, 50]):
        b.add_feature((1, 1), a, s)
    assert b.size() == 3 and b.ages == [6, 4, 5] and b.strengths == [60, 40, 50]


def featureset_image():
    img = np.zeros((300, 200), np.uint8)
    for i in range(11):
        add_triangle(img, 20, (i + 1) * 20, 8)
    return img


def test_featureset():   # main.cpp:102-127
    img = featureset_image()
    pts, resp = O.fast_detect(img, 1)
    p, a, s = O.bucket_filter(pts, np.zeros(len(pts), np.int32), resp.astype(np.int32), 300, 200, 92, 160, 4, 1)
    assert len(p) == 11 and (a == 0).all() and (s <= 128).all()
    p2, _, _ = O.bucket_filter(p, a, s, 300, 200, 1, 1, 0, 7)
    assert len(p2) == 7


def filter_fixture():
    pts = [(299.0, 299.0)] * 15 + [(299.0, 0.0)] * 10 + [(0.0, 150.0)] * 5
    return np.array(pts, np.float32), np.zeros(30, np.int32), np.full(30, 40, np.int32)


def test_featureset_filter():   # main.cpp:128-159
    p, a, s = filter_fixture()
    p1, a1, s1 = O.bucket_filter(p, a, s, 300, 300, 2, 2, 0, 11)
    assert len(p1) == 26
    assert len(O.bucket_filter(p1, a1, s1, 300, 300, 2, 1, 0, 11)[0]) == 21
    assert len(O.bucket_filter(p1, a1, s1, 300, 300, 1, 2, 0, 11)[0]) == 16


def test_find_unmoved_points():   # main.cpp:161-172
    i = np.arange(35)
    p1 = np.stack([i, i], 1).astype(np.float32)
    p2 = np.stack([i + (i % 5 == 0), i + (i % 7 == 0)], 1).astype(np.float32)
    ok = O.find_close_points(p1, p2, 0.5)
    assert (ok == ((i % 5 != 0) & (i % 7 != 0))).all()


def circular_images():
    imgs = [np.zeros((600, 600), np.uint8) for _ in range(4)]   # L0, R0, L1, R1
    for i in range(11):
        for j in range(11):
            add_triangle(imgs[0], (j + 1) * 40, (i + 1) * 40, 8)
            add_triangle(imgs[2], (j + 1) * 40 + 1, (i + 1) * 40, 8)
            add_triangle(imgs[1], (j + 1) * 40, (i + 1) * 40 + 1, 8)
            add_triangle(imgs[3], (j + 1) * 40 + 1, (i + 1) * 40 + 1, 8)
    return imgs


def test_circular_matching():   # main.cpp:174-209
    L0, R0, L1, R1 = circular_images()
    pts, resp = O.fast_detect(L0, O.FAST_THRESHOLD)
    p, a, s = O.bucket_filter(pts, np.zeros(len(pts), np.int32), resp.astype(np.int32), 600, 600, 92, 160, 4, 1)
    assert len(p) == 121
    pL1, s0 = O.lk_leg(L0, L1, p)
    pR1, s1 = O.lk_leg(L1, R1, pL1)
    pR0, s2 = O.lk_leg(R1, R0, pR1)
    pC, s3 = O.lk_leg(R0, L0, pR0)
    keep = O.find_close_points(p, pC, O.CIRCULAR_MATCHING_SUCCESS_THRESHOLD) & (s0 & s1 & s2 & s3).astype(bool)
    assert keep.sum() == 121


def camera_to_world_fixture():
    wp, cp = [], []
    for i in (-1, 0, 1):
        for j in (-1, 0, 1):
            for k in (5, 6, 7):
                wp.append((-j, i, k))
                cp.append((i / (k + 1), j / (k + 1)))
    return np.array(wp, np.float32), np.array(cp, np.float32)


def test_camera_to_world():   # main.cpp:211-264
    wp, cp = camera_to_world_fixture()
    inl, ok, R, t = O.camera_to_world(np.eye(3, dtype=np.float32), cp, wp, np.eye(3), np.zeros((3, 1)))
    assert ok and len(inl) == 27
    assert abs(t[0, 0]) < 1e-6 and abs(t[1, 0]) < 1e-6 and abs(t[2, 0] - 1) < 1e-6
    assert np.abs(R - np.array([[0, 1, 0], [-1, 0, 0], [0, 0, 1]], float)).max() < 1e-8


def test_oracle_pose_moves_when_one_feature_is_dropped():
    """Why the pipeline pose bar is stated per matched set (tests/test_gpu_pipeline.py::check): cv2.solvePnPRansac
    itself, as the reference calls it (vo.cpp:301-304), is not stable to 1e-3 when ONE of ~2000 matched features is
    removed -- the fixed-seed sample stream indexes the array, so every later 5-point sample changes.  Measured on
    the oracle alone: a large share of the single-feature drops moves the translation by more than 1e-3 relative
    (up to ~2e-2), while 1e-4 px noise on the SAME set moves it by < 1e-4."""
    import cv2
    import synthdata as synth
    from oracle import ref_cv2 as O
    cv2.setNumThreads(0)
    seq = synth.make_sequence(n_frames=5, seed=0)
    vo = O.RefVO(seq.Pl, seq.Pr)
    rng = np.random.default_rng(0)
    drop, jitter = [], []
    for L, R in seq.frames():
        Rg, tg = vo.R.copy(), vo.t.copy()
        vo.stereo_callback(L, R)
        tp = vo.taps
        if tp.X3 is None:
            continue
        for _ in range(8):
            m = np.ones(len(tp.pL1), bool)
            m[rng.integers(len(m))] = False
            _, ok, _, t1 = O.camera_to_world(vo.K, tp.pL1[m], tp.X3[m], Rg, tg)
            drop.append(np.linalg.norm(t1 - tp.t) / np.linalg.norm(tp.t))
            noisy = (tp.pL1 + rng.normal(0, 1e-4, tp.pL1.shape)).astype(np.float32)
            _, ok, _, t2 = O.camera_to_world(vo.K, noisy, tp.X3, Rg, tg)
            jitter.append(np.linalg.norm(t2 - tp.t) / np.linalg.norm(tp.t))
    drop, jitter = np.array(drop), np.array(jitter)
    print(f"oracle: one feature dropped -> dt > 1e-3 in {(drop > 1e-3).mean():.0%} of the trials, max {drop.max():.1e}; "
          f"same set + 1e-4 px noise -> max {jitter.max():.1e}")
    assert jitter.max() < 1e-3
    assert (drop > 1e-3).mean() > 0.05 and drop.max() < 5e-2
