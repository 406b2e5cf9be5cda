"""CPU checks of the kernels' arithmetic cores (svo_core.cuh compiled for the host with LANES=1,
tests/host_cores.cpp) against the cv2 oracle.  The same inline functions are what the CUDA kernels
execute; the -m gpu tests repeat these comparisons through the C ABI on the device."""
import ctypes as C

import numpy as np

from oracle import ref_cv2 as O
import synthdata as synth

u8p, fp, dp, ip = C.POINTER(C.c_uint8), C.POINTER(C.c_float), C.POINTER(C.c_double), C.POINTER(C.c_int)


def P(a, t):
    return a.ctypes.data_as(t)


def fast_map(hc, img, thr):
    h, w = img.shape
    m = np.zeros_like(img)
    n = hc.hc_fast_map(P(img, u8p), w, h, img.strides[0], thr, P(m, u8p))
    ys, xs = np.nonzero(m)
    return n, np.stack([xs, ys], 1).astype(np.float32), m[ys, xs].astype(np.float32)


def test_fast_bit_exact(hostcores, run1):
    rng = np.random.default_rng(0)
    imgs = [run1["gray"][0][0], run1["gray"][5][1], rng.integers(0, 256, (47, 156), dtype=np.uint8),
            rng.integers(0, 256, (64, 64), dtype=np.uint8) // 8 * 8]
    for img in imgs:
        img = np.ascontiguousarray(img)
        for thr in (1, 5, 20):
            n, pts, resp = fast_map(hostcores, img, thr)
            rp, rr = O.fast_detect(img, thr)
            assert n == len(rp) and np.array_equal(pts, rp) and np.array_equal(resp, rr)


def test_fast_packed_pair_arithmetic(hostcores, run1):
    """svo_fast_packed.cuh (what fast_kernel executes per pixel pair) == the scalar arc score, compass bound and
    response byte on every quad, incl. saturated and flat images."""
    rng = np.random.default_rng(1)
    imgs = [run1["gray"][0][0], rng.integers(0, 256, (40, 90), dtype=np.uint8),
            rng.integers(0, 2, (40, 90), dtype=np.uint8) * 255, np.full((20, 40), 255, np.uint8),
            np.zeros((20, 40), np.uint8), rng.integers(0, 256, (30, 64), dtype=np.uint8) // 64 * 85]
    for img in imgs:
        img = np.ascontiguousarray(img)
        h, w = img.shape
        for thr in (1, 5, 20, 60, 200):
            assert hostcores.hc_fast_packed_check(P(img, u8p), w, h, img.strides[0], thr) == 0


def test_pyrdown_bit_exact(hostcores, run1):
    import cv2
    rng = np.random.default_rng(1)
    for img in [run1["gray"][0][0], rng.integers(0, 256, (47, 156), dtype=np.uint8),
                rng.integers(0, 256, (376, 1241), dtype=np.uint8), rng.integers(0, 256, (37, 50), dtype=np.uint8)]:
        img = np.ascontiguousarray(img)
        h, w = img.shape
        d = np.zeros(((h + 1) // 2, (w + 1) // 2), np.uint8)
        hostcores.hc_pyrdown(P(img, u8p), w, h, w, P(d, u8p), d.shape[1])
        assert np.array_equal(d, cv2.pyrDown(img))


def host_lk(hc, I, J, pts):
    hc.hc_lk.argtypes = [u8p, u8p] + [C.c_int] * 5 + [C.c_double, C.c_float, fp, C.c_int, fp, u8p]
    pts = np.ascontiguousarray(pts, np.float32)
    n = len(pts)
    out = np.zeros((n, 2), np.float32)
    st = np.zeros(n, np.uint8)
    h, w = I.shape
    hc.hc_lk(P(I, u8p), P(J, u8p), w, h, w, 3, 30, 1e-4, 1e-3, P(pts, fp), n, P(out, fp), P(st, u8p))
    return out, st


def test_lk_matches_opencv(hostcores, run1):
    """north_star tolerance: <= 0.05 px on >= 99 % of features, identical status on >= 99 %."""
    g = run1["gray"]
    rng = np.random.default_rng(2)
    border = np.array([[1.2, 1.7], [510.5, 286.9], [0, 0], [511, 287], [3.3, 150.2], [508.8, 10.1], [-3, 5], [600, 100]],
                      np.float32)
    for (I, J) in [(g[0][0], g[1][0]), (g[3][0], g[3][1]), (g[5][1], g[4][1]), (g[6][1], g[6][0])]:
        I, J = np.ascontiguousarray(I), np.ascontiguousarray(J)
        pts, _ = O.fast_detect(I, 5)
        pts = np.concatenate([pts + rng.uniform(-0.5, 0.5, pts.shape).astype(np.float32), border])
        o, s = host_lk(hostcores, I, J, pts)
        ro, rs = O.lk_leg(I, J, pts)
        # north_star's bar is 0.05 px on >= 99 %; the window sums follow OpenCV's float32 order, so: identical bits
        assert np.array_equal(s, rs)
        assert np.array_equal(o, ro), (np.abs(o - ro).max(), (o != ro).any(1).mean())


def host_lk_cn(hc, I, J, pts):
    hc.hc_lk_cn.argtypes = [u8p, u8p] + [C.c_int] * 6 + [C.c_double, C.c_float, fp, C.c_int, fp, u8p]
    pts = np.ascontiguousarray(pts, np.float32)
    n = len(pts)
    out = np.zeros((n, 2), np.float32)
    st = np.zeros(n, np.uint8)
    h, w, cn = I.shape
    hc.hc_lk_cn(P(I, u8p), P(J, u8p), w, h, w * cn, cn, 3, 30, 1e-4, 1e-3, P(pts, fp), n, P(out, fp), P(st, u8p))
    return out, st


def test_lk_three_channel_matches_opencv(hostcores, run1):
    """The CLI-faithful mode: calcOpticalFlowPyrLK on the BGR images cv::imread returns."""
    import cv2
    b = run1["bgr"]
    rng = np.random.default_rng(3)
    border = np.array([[1.2, 1.7], [510.5, 286.9], [0, 0], [511, 287], [3.3, 150.2], [508.8, 10.1], [-3, 5], [600, 100]],
                      np.float32)
    for (I, J) in [(b[0][0], b[1][0]), (b[3][0], b[3][1]), (b[5][1], b[4][1])]:
        I, J = np.ascontiguousarray(I), np.ascontiguousarray(J)
        pts, _ = O.fast_detect(cv2.cvtColor(I, cv2.COLOR_BGR2GRAY), 5)
        pts = np.concatenate([pts + rng.uniform(-0.5, 0.5, pts.shape).astype(np.float32), border])
        o, s = host_lk_cn(hostcores, I, J, pts)
        ro, rs = O.lk_leg(I, J, pts)
        assert (s == rs).mean() >= 0.99
        d = np.abs(o - ro).max(1)
        assert (d <= 0.05).mean() >= 0.99
        # three-channel pyrDown is per channel
    d = np.zeros((144, 256, 3), np.uint8)
    I = np.ascontiguousarray(b[2][0])
    hostcores.hc_pyrdown_cn(P(I, u8p), 512, 288, 512 * 3, 3, P(d, u8p), 256 * 3)
    assert np.array_equal(d, cv2.pyrDown(I))


def test_triangulate_matches_opencv(hostcores):
    rng = np.random.default_rng(3)
    Pl, Pr = synth.projection_matrices("kitti00")
    n = 500
    X = np.stack([rng.uniform(-10, 10, n), rng.uniform(-2, 2, n), rng.uniform(3, 60, n)], 1)
    Xh = np.concatenate([X, np.ones((n, 1))], 1)
    pl = Xh @ Pl.astype(np.float64).T
    pr = Xh @ Pr.astype(np.float64).T
    pL = (pl[:, :2] / pl[:, 2:] + rng.normal(0, 0.1, (n, 2))).astype(np.float32)
    pR = (pr[:, :2] / pr[:, 2:] + rng.normal(0, 0.1, (n, 2))).astype(np.float32)
    X3 = np.zeros((n, 3), np.float32)
    hostcores.hc_triangulate(P(np.ascontiguousarray(Pl), fp), P(np.ascontiguousarray(Pr), fp), P(pL, fp), P(pR, fp), n,
                             P(X3, fp))
    ref = O.triangulate(Pl, Pr, pL, pR)
    assert np.allclose(X3, ref, rtol=1e-5, atol=1e-6)


def host_pnp(hc, K, uv, X3, R, t):
    hc.hc_pnp_ransac.argtypes = [fp, fp, fp, C.c_int, dp, dp, C.c_int, C.c_int, C.c_double, C.c_float, ip, ip]
    R = np.array(R, np.float64).copy()
    t = np.array(t, np.float64).reshape(3).copy()
    inl = np.zeros(len(uv), np.int32)
    ok = C.c_int(0)
    K = np.ascontiguousarray(K, np.float32)
    uv = np.ascontiguousarray(uv, np.float32)
    X3 = np.ascontiguousarray(X3, np.float32)
    n = hc.hc_pnp_ransac(P(K, fp), P(uv, fp), P(X3, fp), len(uv), P(R, dp), P(t, dp), 100, 100,
                         float(np.float32(0.98)), 64.0, P(inl, ip), C.byref(ok))
    return inl[:n], ok.value, R, t


def test_pnp_reference_known_answer(hostcores):   # main.cpp:211-264
    from test_oracle_golden import camera_to_world_fixture
    wp, cp = camera_to_world_fixture()
    inl, ok, R, t = host_pnp(hostcores, np.eye(3), cp, wp, np.eye(3), np.zeros(3))
    assert ok and len(inl) == 27
    assert abs(t[0]) < 1e-6 and abs(t[1]) < 1e-6 and abs(t[2] - 1) < 1e-6
    assert np.abs(R - np.array([[0, 1, 0], [-1, 0, 0], [0, 0, 1]], float)).max() < 1e-8


def test_epnp_hypothesis_is_bit_identical_to_opencv(hostcores):
    """One RANSAC hypothesis = cv::solvePnP(SOLVEPNP_EPNP) on 5 points.  svo_epnp.cuh restates OpenCV's one-sided
    Jacobi SVD and epnp.cpp operation for operation, so even the samples whose 2-D null-space basis is decided by
    rounding noise must give the same pose: translation within 1e-12 (measured: identical doubles)."""
    import cv2
    hostcores.hc_epnp5.argtypes = [fp, fp, fp, dp, dp]
    rng = np.random.default_rng(11)
    Pl, _ = synth.projection_matrices("cli")
    K = np.ascontiguousarray(Pl[:, :3], np.float32)
    worst = 0.0
    for trial in range(300):
        X = np.stack([rng.uniform(-3, 3, 5), rng.uniform(-1, 1, 5), rng.uniform(1.5, 20, 5)], 1).astype(np.float32)
        Rm, _ = cv2.Rodrigues(rng.normal(0, 0.02, 3))
        Xc = (Rm @ X.astype(np.float64).T).T + rng.normal(0, 0.05, 3)
        uv = (Xc[:, :2] / Xc[:, 2:]) * np.array([K[0, 0], K[1, 1]]) + np.array([K[0, 2], K[1, 2]])
        uv = (uv + rng.normal(0, [0.0, 0.3, 3.0][trial % 3], uv.shape)).astype(np.float32)   # exact / noisy / gross
        ok, rv, tv = cv2.solvePnP(X.reshape(-1, 1, 3), uv.reshape(-1, 1, 2), K, np.zeros((4, 1), np.float32),
                                  flags=cv2.SOLVEPNP_EPNP)
        R, t = np.zeros(9), np.zeros(3)
        assert hostcores.hc_epnp5(P(K, fp), P(uv, fp), P(X, fp), P(R, dp), P(t, dp))
        Rc, _ = cv2.Rodrigues(rv)
        worst = max(worst, np.abs(t - tv.ravel()).max(), np.abs(R.reshape(3, 3) - Rc).max())
    assert worst < 1e-12, worst


def test_pnp_ransac_matches_opencv(hostcores):
    """cameraToWorld (vo.cpp:282-313): OpenCV's sample stream, EPnP hypotheses, adaptive stop and LM refit are all
    replayed, so on the same inputs the inlier set is IDENTICAL and the pose agrees far inside north_star's
    1e-3 rad / 1e-3 relative translation -- with 12 % gross outliers and 0.2 px noise in the data."""
    import cv2
    rng = np.random.default_rng(4)
    Pl, _ = synth.projection_matrices("cli")
    K = Pl[:, :3]
    n = 1500
    X3 = np.stack([rng.uniform(-3, 3, n), rng.uniform(-1, 1, n), rng.uniform(1.5, 20, n)], 1).astype(np.float32)
    for trial in range(12):
        rv, tt = rng.normal(0, 0.01, 3), rng.normal(0, 0.03, 3)
        Rm, _ = cv2.Rodrigues(rv)
        Xc = (Rm @ X3.astype(np.float64).T).T + tt
        uv = (Xc[:, :2] / Xc[:, 2:]) * np.array([K[0, 0], K[1, 1]]) + np.array([K[0, 2], K[1, 2]])
        uv += rng.normal(0, 0.2, uv.shape)
        oi = rng.choice(n, n // 8, replace=False)
        uv[oi] += rng.uniform(-40, 40, (len(oi), 2))
        uv = uv.astype(np.float32)
        inl, ok, R, t = host_pnp(hostcores, K, uv, X3, np.eye(3), np.zeros(3))
        ri, rok, rR, rt = O.camera_to_world(K, uv, X3, np.eye(3), np.zeros((3, 1)))
        assert ok and rok
        assert np.array_equal(inl, ri), (trial, len(set(inl) ^ set(ri)))
        dR = np.linalg.norm(cv2.Rodrigues(R @ rR.T)[0])
        dt = np.linalg.norm(t - rt.ravel()) / np.linalg.norm(rt)
        assert dR < 1e-6 and dt < 1e-6, (trial, dR, dt)


def test_bucket_key_order(hostcores):
    assert hostcores.hc_bucket_key_roundtrip(7, 12345, 254)
    assert hostcores.hc_bucket_key_roundtrip(-3, (1 << 24) + 466615, 0)
    assert hostcores.hc_bucket_key_less(3, 5, 4, 900)       # higher score wins
    assert hostcores.hc_bucket_key_less(3, 9, 3, 2)         # tie: earlier order wins
    for a, s in [(0, 5), (0, 19), (0, 20), (0, 39), (0, 40), (3, 254), (19, 1)]:
        assert hostcores.hc_bucket_score(a, s, 20) == O.compute_score(a, s)
