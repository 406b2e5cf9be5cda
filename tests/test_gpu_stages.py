"""-m gpu parity tests, stage by stage, through the C ABI (stereo_visual_odometry_b200.vo ->
libsvo_b200.so) against the cv2 oracle.  Bit-exact for integer work (FAST, bucketing, pyramid,
compaction); stated tolerances for LK / triangulation / PnP."""
import numpy as np
import pytest

import stereo_visual_odometry_b200 as svo
from stereo_visual_odometry_b200 import vo as V
from conftest import add_triangle
from oracle import ref_cv2 as O
import synthdata as synth
from test_oracle_golden import (camera_to_world_fixture, circular_images, featureset_image, filter_fixture)

pytestmark = pytest.mark.gpu


def images(run1):
    rng = np.random.default_rng(0)
    seq = synth.make_sequence(n_frames=2, seed=1)
    L, R = seq.render(1)
    return [run1["gray"][0][0], run1["gray"][5][1], L, R, rng.integers(0, 256, (47, 156), dtype=np.uint8),
            rng.integers(0, 256, (131, 257), dtype=np.uint8), np.zeros((40, 40), np.uint8)]


def test_fast_bit_exact(run1):
    for img in images(run1):
        for thr in (1, 5, 20, 60):
            pts, resp = V.featureDetectionFast(img, thr)
            rp, rr = O.fast_detect(np.ascontiguousarray(img), thr)
            assert len(pts) == len(rp), (img.shape, thr)
            assert np.array_equal(pts, rp) and np.array_equal(resp, rr)


def test_fast_strided_view(run1):
    big = np.zeros((300, 700), np.uint8)
    big[:288, 100:612] = run1["gray"][2][0]
    view = big[:288, 100:612]
    pts, resp = V.featureDetectionFast(view, 20)
    rp, rr = O.fast_detect(np.ascontiguousarray(view), 20)
    assert np.array_equal(pts, rp) and np.array_equal(resp, rr)


def test_pyramid_bit_exact(run1):
    import cv2
    for img in images(run1)[:6]:
        lv = V.buildOpticalFlowPyramid(img, 3, 10)
        ref = np.ascontiguousarray(img)
        for l in range(1, len(lv)):
            ref = cv2.pyrDown(ref)
            assert np.array_equal(lv[l], ref), (img.shape, l)
        h, w = img.shape
        expect = 1
        while expect <= 3 and (w + 1) // 2 > 10 and (h + 1) // 2 > 10:
            w, h, expect = (w + 1) // 2, (h + 1) // 2, expect + 1
        assert len(lv) == expect


def test_bucket_generic_matches_reference_policy():
    rng = np.random.default_rng(5)
    for per_bucket, grid, start in [(1, (92, 160), 4), (3, (10, 12), 1), (7, (1, 1), 0), (2, (5, 3), 0), (0, (4, 4), 0)]:
        n = 3000
        H, W = 376, 1241
        pts = np.stack([rng.uniform(0, W, n), rng.uniform(0, H, n)], 1).astype(np.float32)
        pts[:500] = np.floor(pts[:500])
        ages = rng.integers(0, 25, n).astype(np.int32)
        strs = rng.integers(1, 255, n).astype(np.int32)
        fs = V.FeatureSet()
        fs.points, fs.ages, fs.strengths = pts, ages, strs
        fs.filterByBucketLocationInternal(np.zeros((H, W), np.uint8), grid[0], grid[1], start, per_bucket)
        rp, ra, rs = O.bucket_filter(pts, ages, strs, H, W, grid[0], grid[1], start, per_bucket)
        if per_bucket == 0:
            assert fs.size() == 0
            continue
        assert fs.size() == len(rp)
        assert np.array_equal(fs.points, rp) and np.array_equal(fs.ages, ra) and np.array_equal(fs.strengths, rs)


def test_reference_test_bucket_nonempty():   # main.cpp:62-78 through the generic kernel (1x1 grid, capacity 3)
    fs = V.FeatureSet()
    fs.points = np.ones((5, 2), np.float32)
    fs.ages = np.array([6, 2, 3, 4, 5], np.int32)
    fs.strengths = np.array([60, 20, 30, 40, 50], np.int32)
    fs.filterByBucketLocationInternal(np.zeros((10, 10), np.uint8), 1, 1, 0, 3)
    assert fs.ages.tolist() == [6, 4, 5] and fs.strengths.tolist() == [60, 40, 50]
    fs.filterByBucketLocationInternal(np.zeros((10, 10), np.uint8), 1, 1, 0, 0)   # Bucket(0), main.cpp:50-60
    assert fs.size() == 0


def test_reference_test_featureset():   # main.cpp:102-127
    img = featureset_image()
    fs = V.FeatureSet()
    pts, resp = V.featureDetectionFast(img, 1)
    fs.points, fs.ages, fs.strengths = pts, np.zeros(len(pts), np.int32), resp.astype(np.int32)
    fs.filterByBucketLocation(img)
    assert fs.size() == 11 and (fs.ages == 0).all() and (fs.strengths <= 128).all()
    fs.filterByBucketLocationInternal(img, 1, 1, 0, 7)
    assert fs.size() == 7


def test_reference_test_featureset_filter():   # main.cpp:128-159
    p, a, s = filter_fixture()
    img = np.zeros((300, 300), np.uint8)
    fs = V.FeatureSet()
    fs.points, fs.ages, fs.strengths = p, a, s
    fs.filterByBucketLocationInternal(img, 2, 2, 0, 11)
    assert fs.size() == 26
    cp = V.FeatureSet()
    cp.points, cp.ages, cp.strengths = fs.points.copy(), fs.ages.copy(), fs.strengths.copy()
    fs.filterByBucketLocationInternal(img, 2, 1, 0, 11)
    assert fs.size() == 21
    cp.filterByBucketLocationInternal(img, 1, 2, 0, 11)
    assert cp.size() == 16


def test_reference_test_find_unmoved_points():   # main.cpp:161-172
    i = np.arange(35)
    p1 = np.stack([i, i], 1).astype(np.float32)
    p2 = np.stack([i + (i % 5 == 0), i + (i % 7 == 0)], 1).astype(np.float32)
    assert (V.findClosePoints(p1, p2, 0.5) == ((i % 5 != 0) & (i % 7 != 0))).all()
    assert len(V.findClosePoints(np.zeros((0, 2)), np.zeros((0, 2)), 0.5)) == 0
    nan = np.array([[np.nan, 1.0]], np.float32)
    assert V.findClosePoints(nan, nan, 0.5).tolist() == O.find_close_points(nan, nan, 0.5).tolist()


def test_compaction_stable():
    rng = np.random.default_rng(6)
    for n in (0, 1, 31, 1024, 5000):
        p = rng.uniform(0, 100, (n, 2)).astype(np.float32)
        k = rng.integers(0, 2, n).astype(bool)
        out = V.deletePointsWithFailureStatus(p, k)
        assert np.array_equal(out, p[k])


def test_lk_leg_matches_opencv(run1):
    g = run1["gray"]
    rng = np.random.default_rng(2)
    border = np.array([[1.2, 1.7], [510.5, 286.9], [0, 0], [511, 287], [3.3, 150.2], [508.8, 10.1], [-3, 5], [600, 100]],
                      np.float32)
    for (I, J) in [(g[0][0], g[1][0]), (g[3][0], g[3][1]), (g[5][1], g[4][1]), (g[6][1], g[6][0])]:
        pts, _ = O.fast_detect(np.ascontiguousarray(I), 5)
        pts = np.concatenate([pts + rng.uniform(-0.5, 0.5, pts.shape).astype(np.float32), border])
        o, s = V.calcOpticalFlowPyrLK(I, J, pts)
        ro, rs = O.lk_leg(I, J, pts)
        # the window sums are formed in OpenCV's float32 order (svo_core.cuh cv_lk_sum_*): identical bits
        assert np.array_equal(s, rs)
        assert np.array_equal(o, ro), (np.abs(o - ro).max(), (o != ro).any(1).mean())


def test_lk_leg_three_channel_matches_opencv(run1):
    """calcOpticalFlowPyrLK on interleaved 8UC3 images (what the reference CLI passes)."""
    import cv2
    b = run1["bgr"]
    rng = np.random.default_rng(3)
    border = np.array([[1.2, 1.7], [510.5, 286.9], [0, 0], [511, 287], [3.3, 150.2], [508.8, 10.1], [-3, 5], [600, 100]],
                      np.float32)
    for (I, J) in [(b[0][0], b[1][0]), (b[3][0], b[3][1]), (b[5][1], b[4][1])]:
        pts, _ = O.fast_detect(cv2.cvtColor(I, cv2.COLOR_BGR2GRAY), 5)
        pts = np.concatenate([pts + rng.uniform(-0.5, 0.5, pts.shape).astype(np.float32), border])
        o, s = V.calcOpticalFlowPyrLK(I, J, pts)
        ro, rs = O.lk_leg(I, J, pts)
        assert (s == rs).mean() >= 0.99
        assert (np.abs(o - ro).max(1) <= 0.05).mean() >= 0.99


def test_lk_kitti_shape():
    seq = synth.make_sequence(n_frames=2, seed=2)
    (L0, R0), (L1, R1) = seq.render(0), seq.render(1)
    pts, _ = O.fast_detect(L0, 20)
    for I, J in [(L0, L1), (L1, R1), (R1, R0)]:
        o, s = V.calcOpticalFlowPyrLK(I, J, pts)
        ro, rs = O.lk_leg(I, J, pts)
        assert np.array_equal(s, rs)
        assert np.array_equal(o, ro), (np.abs(o - ro).max(), (o != ro).any(1).mean())


def test_lk_small_image_fewer_levels():
    rng = np.random.default_rng(7)
    import cv2
    I = cv2.GaussianBlur(rng.integers(0, 256, (60, 90), dtype=np.uint8), (0, 0), 1.5)
    J = np.roll(I, 1, axis=1)
    pts = np.stack([rng.uniform(5, 85, 50), rng.uniform(5, 55, 50)], 1).astype(np.float32)
    o, s = V.calcOpticalFlowPyrLK(I, J, pts)
    ro, rs = O.lk_leg(I, J, pts)
    assert (s == rs).mean() >= 0.95
    ok = (s != 0) & (rs != 0)
    assert (np.abs(o - ro).max(1)[ok] <= 0.05).mean() >= 0.95


def test_reference_test_circular_matching():   # main.cpp:174-209
    L0, R0, L1, R1 = circular_images()
    fs = V.FeatureSet()
    fs.appendFeaturesFromImage(L0, V.FAST_THRESHOLD)
    assert fs.size() == 121
    vo = V.VisualOdometry(600, 600)
    vo.initalize_projection_matricies(*synth.projection_matrices("cli"))
    vo.stereo_callback(L0, R0)
    # boundary conditions: empty and single-element inputs must not crash (main.cpp:194-198)
    out = vo.circularMatching(L1, R1, np.zeros((0, 2), np.float32))
    assert all(len(o) == 0 for o in out)
    out = vo.circularMatching(L1, R1, fs.points[:1])
    assert len(out[0]) == len(out[1]) == len(out[2]) == len(out[3])
    pl0, pr0, pl1, pr1 = vo.circularMatching(L1, R1, fs.points, fs)
    assert len(pl0) == len(pr0) == len(pl1) == len(pr1) == fs.size() == 121
    assert np.abs(pl1 - pl0 - [1, 0]).max() < 0.05 and np.abs(pr0 - pl0 - [0, 1]).max() < 0.05


def test_triangulate_matches_opencv():
    rng = np.random.default_rng(3)
    for calib in ("kitti00", "cli", "zed"):
        Pl, Pr = synth.projection_matrices(calib)
        n = 2000
        X = np.stack([rng.uniform(-10, 10, n), rng.uniform(-2, 2, n), rng.uniform(3, 60, n)], 1)
        Xh = np.concatenate([X, np.ones((n, 1))], 1)
        pl, pr = Xh @ Pl.astype(np.float64).T, Xh @ Pr.astype(np.float64).T
        pL = (pl[:, :2] / pl[:, 2:] + rng.normal(0, 0.1, (n, 2))).astype(np.float32)
        pR = (pr[:, :2] / pr[:, 2:] + rng.normal(0, 0.1, (n, 2))).astype(np.float32)
        X3 = V.triangulatePoints(Pl, Pr, pL, pR)
        ref = O.triangulate(Pl, Pr, pL, pR)
        assert np.allclose(X3, ref, rtol=1e-5, atol=1e-6)
    assert V.triangulatePoints(Pl, Pr, np.zeros((0, 2)), np.zeros((0, 2))).shape == (0, 3)


def test_reference_test_camera_to_world():   # main.cpp:211-264
    wp, cp = camera_to_world_fixture()
    inl, ok, R, t = V.cameraToWorld(np.eye(3, dtype=np.float32), cp, wp, np.eye(3), np.zeros((3, 1)))
    assert ok and len(inl) == 27 and inl.tolist() == list(range(27))
    assert abs(t[0, 0]) < 1e-6 and abs(t[1, 0]) < 1e-6 and abs(t[2, 0] - 1) < 1e-6
    assert np.abs(R - np.array([[0, 1, 0], [-1, 0, 0], [0, 0, 1]], float)).max() < 1e-8


def test_pnp_ransac_matches_opencv():
    """cameraToWorld on the device vs cv2.solvePnPRansac on the same inputs (12 % gross outliers, 0.2 px noise):
    identical ascending inlier list, pose far inside north_star's 1e-3 rad / 1e-3 relative translation."""
    import cv2
    rng = np.random.default_rng(4)
    Pl, _ = synth.projection_matrices("cli")
    K = Pl[:, :3]
    n = 1500
    X3 = np.stack([rng.uniform(-3, 3, n), rng.uniform(-1, 1, n), rng.uniform(1.5, 20, n)], 1).astype(np.float32)
    dist = []
    for trial in range(12):
        rv, tt = rng.normal(0, 0.01, 3), rng.normal(0, 0.03, 3)
        Rm, _ = cv2.Rodrigues(rv)
        Xc = (Rm @ X3.astype(np.float64).T).T + tt
        uv = (Xc[:, :2] / Xc[:, 2:]) * np.array([K[0, 0], K[1, 1]]) + np.array([K[0, 2], K[1, 2]])
        uv += rng.normal(0, 0.2, uv.shape)
        oi = rng.choice(n, n // 8, replace=False)
        uv[oi] += rng.uniform(-40, 40, (len(oi), 2))
        uv = uv.astype(np.float32)
        inl, ok, R, t = V.cameraToWorld(K, uv, X3, np.eye(3), np.zeros((3, 1)))
        ri, rok, rR, rt = O.camera_to_world(K, uv, X3, np.eye(3), np.zeros((3, 1)))
        assert ok and rok
        dR = float(np.linalg.norm(cv2.Rodrigues(R @ rR.T)[0]))
        dt = float(np.linalg.norm(t - rt) / np.linalg.norm(rt))
        dist.append((len(set(inl.tolist()) ^ set(ri.tolist())), dR, dt))
    print("pnp vs opencv (inlier symmetric difference, dR rad, dt relative):", dist)
    assert all(sd == 0 for sd, _, _ in dist), dist
    assert all(dR < 1e-6 and dt < 1e-6 for _, dR, dt in dist), dist


def test_pnp_failure_restores_translation():
    rng = np.random.default_rng(8)
    n = 40
    X3 = rng.uniform(-1, 1, (n, 3)).astype(np.float32) + [0, 0, 5]
    uv = rng.uniform(0, 600, (n, 2)).astype(np.float32)   # garbage correspondences
    t0 = np.array([[0.01], [0.02], [0.03]])
    inl, ok, R, t = V.cameraToWorld(synth.projection_matrices("cli")[0], uv, X3, np.eye(3), t0)
    rinl, rok, rR, rt = O.camera_to_world(synth.projection_matrices("cli")[0][:, :3], uv, X3, np.eye(3), t0)
    assert ok == rok
    if not ok:
        assert np.array_equal(t, t0) and np.array_equal(R, np.eye(3))


def test_matching_features_boundary_matches_oracle(run1):
    """VisualOdometry::matchingFeatures (include/vo.h:354-362, src/vo.cpp:315-366) through the boundary function
    (svo_matching_features via vo.py): all five containers against the oracle's taps, with an empty feature set
    (frame 1: FAST + bucketing only) and with the tracked set the reference hands over on the next frame."""
    from scipy.spatial import cKDTree
    cases = [("run1", [run1["gray"][i] for i in range(4)], O.CLI_P_LEFT, O.CLI_P_RIGHT)]
    seq = synth.make_sequence(width=640, height=240, n_frames=4, seed=3)
    cases.append(("synthetic", list(seq.frames()), seq.Pl, seq.Pr))
    for name, frames, Pl, Pr in cases:
        ref = O.RefVO(Pl, Pr)
        vo = V.VisualOdometry()
        ref.stereo_callback(*frames[0])
        for i in range(1, 4):
            fs = V.FeatureSet()
            fs.points, fs.ages, fs.strengths = ref.pts.copy(), ref.ages.copy(), ref.strengths.copy()   # state before the call
            (L0, R0), (L1, R1) = frames[i - 1], frames[i]
            ref.stereo_callback(L1, R1)
            tp = ref.taps
            pL0, pR0, pL1, pR1 = vo.matchingFeatures(L0, R0, L1, R1, fs)
            n = len(pL0)
            assert len(pR0) == len(pL1) == len(pR1) == fs.size() == n
            assert abs(n - len(tp.pL0)) <= 0.01 * len(tp.pL0) + 1, (name, i, n, len(tp.pL0))
            assert np.array_equal(fs.points, pL0)                       # currentVOFeatures holds t0 coordinates (vo.cpp:360)
            d, j = cKDTree(tp.pL0).query(pL0, distance_upper_bound=1e-3)
            hit = np.isfinite(d)
            assert hit.mean() >= 0.99 and hit.sum() >= 0.99 * len(tp.pL0), (name, i, hit.mean())
            g, r = np.nonzero(hit)[0], j[hit]
            assert (np.diff(r) > 0).all()                               # same (bucket row-major) order
            keep = tp.keep
            ages = tp.bucketed[1][keep][r]                              # before stereo_callback's ages++ (vo.cpp:70-72)
            strengths = tp.bucketed[2][keep][r]
            assert np.array_equal(fs.ages[g], ages) and np.array_equal(fs.strengths[g], strengths)
            for mine, theirs in ((pL1, tp.pL1), (pR0, tp.pR0), (pR1, tp.pR1)):
                assert (np.abs(mine[g] - theirs[r]).max(1) <= 0.05).mean() >= 0.99
            if i == 1:   # identical integer inputs: the whole set is identical, positions to LK float noise
                assert n == len(tp.pL0) and np.array_equal(pL0, tp.pL0)
