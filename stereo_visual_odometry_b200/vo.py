"""Host-side mirror of the reference's `include/vo.h` interface (namespace
visual_odometry) over the C ABI.  Same names, argument meaning and failure
behaviour as the reference; every call runs the sm_100a kernels -- there is no
CPU path here.  (The C++ drop-in header is include/vo.h of this repo; this module
is the same surface for Python callers, tests and bench.py.)
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import FrameResult, Params, SvoError, check, default_params, lib, ptr

# include/vo.h:53-127 of the reference
BUCKET_START_ROW = 4
BUCKETS_ALONG_HEIGHT = 92
BUCKETS_ALONG_WIDTH = 160
FEATURES_PER_BUCKET = 1
FEATURES_THRESHOLD = 15
PRE_MATCHING_FEATURE_THRESHOLD = 100
AGE_THRESHOLD = 20
FAST_THRESHOLD = 20
RANSAC_REPROJECTION_ERROR = 8.0
RANSAC_ITERATIONS = 100
OPTICAL_FLOW_MIN_EIG_THRESHOLD = 0.001
CIRCULAR_MATCHING_SUCCESS_THRESHOLD = 0.15
MAX_TRANSLATION_NORM = 0.1
MAX_ROTATION_NORM = 0.5


def _img(a) -> np.ndarray:
    a = np.asarray(a)
    if a.dtype == np.uint8 and a.ndim == 3 and a.shape[2] == 3:   # CV_8UC3 as the reference CLI feeds it (F4)
        if a.strides[2] != 1 or a.strides[1] != 3:
            a = np.ascontiguousarray(a)
        return a
    if a.dtype != np.uint8 or a.ndim != 2:
        raise SvoError("images must be 2-D uint8 (CV_8UC1) or HxWx3 uint8 (CV_8UC3, CLI-faithful mode)")
    if a.strides[1] != 1:
        a = np.ascontiguousarray(a)
    return a


def _pts(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, np.float32).reshape(-1, 2))


# --------------------------------------------------------------------------- free functions
def featureDetectionFast(image, fast_threshold: int, cap: int | None = None):
    """vo.h:393-395 / feature_set.cpp:55-68.  Returns (points [K,2] f32, responses [K] f32)."""
    im = _img(image)
    h, w = im.shape
    cap = cap or max(1024, (w * h) // 4)
    xy = np.zeros((cap, 2), np.float32)
    resp = np.zeros((cap,), np.float32)
    n = check(lib().svo_fast_detect(ptr(im), w, h, im.strides[0], int(fast_threshold), ptr(xy, C.c_float),
                                    ptr(resp, C.c_float), cap))
    if n > cap:
        return featureDetectionFast(image, fast_threshold, cap=n)
    return xy[:n].copy(), resp[:n].copy()


def findClosePoints(points_1, points_2, threshold: float) -> np.ndarray:
    """vo.h:430-432 / vo.cpp:265-280."""
    a, b = _pts(points_1), _pts(points_2)
    ok = np.zeros((len(a),), np.uint8)
    check(lib().svo_find_close_points(ptr(a, C.c_float), ptr(b, C.c_float), len(a), float(threshold), ptr(ok)))
    return ok.astype(bool)


def deletePointsWithFailureStatus(point_vector, isok) -> np.ndarray:
    """vo.h:406-408 / vo.cpp:144-153: stable compaction; returns the kept points."""
    p = _pts(point_vector).copy()
    k = np.ascontiguousarray(np.asarray(isok).astype(np.uint8))
    n = check(lib().svo_compact(ptr(p), 8, len(p), ptr(k)))
    return p[:n].copy()


def cameraToWorld(cameraProjection, cameraPoints, worldPoints, rotation, translation, params: Params | None = None):
    """vo.h:452-456 / vo.cpp:282-313.  Returns (inliers int32[N_in], success, rotation, translation);
    rotation / translation follow the reference's in/out rule (R updated on success, t restored on failure)."""
    K = np.ascontiguousarray(np.asarray(cameraProjection, np.float32)[:3, :3])
    cam = _pts(cameraPoints)
    world = np.ascontiguousarray(np.asarray(worldPoints, np.float32).reshape(-1, 3))
    R = np.ascontiguousarray(np.asarray(rotation, np.float64).reshape(3, 3)).copy()
    t = np.ascontiguousarray(np.asarray(translation, np.float64).reshape(3)).copy()
    p = params or default_params(64, 64)
    inl = np.zeros((len(cam),), np.int32)
    ok = C.c_int(0)
    n = check(lib().svo_camera_to_world(C.byref(p), ptr(K, C.c_float), ptr(cam, C.c_float), ptr(world, C.c_float),
                                        len(cam), ptr(R, C.c_double), ptr(t, C.c_double), ptr(inl, C.c_int),
                                        C.byref(ok)))
    return inl[:n].copy(), bool(ok.value), R, t.reshape(3, 1)


def getInverseTransform(rotation, translation_stereo) -> np.ndarray:
    """vo.h:469-470 / vo.cpp:246-258 (host, FP64)."""
    R = np.ascontiguousarray(np.asarray(rotation, np.float64).reshape(3, 3))
    t = np.ascontiguousarray(np.asarray(translation_stereo, np.float64).reshape(3))
    T = np.zeros((4, 4), np.float64)
    check(lib().svo_get_inverse_transform(ptr(R, C.c_double), ptr(t, C.c_double), ptr(T, C.c_double)))
    return T


def triangulatePoints(Pl, Pr, pL, pR) -> np.ndarray:
    """cv::triangulatePoints + convertPointsFromHomogeneous as at vo.cpp:89-94 -> [N,3] f32."""
    Pl = np.ascontiguousarray(np.asarray(Pl, np.float32).reshape(12))
    Pr = np.ascontiguousarray(np.asarray(Pr, np.float32).reshape(12))
    a, b = _pts(pL), _pts(pR)
    X = np.zeros((len(a), 3), np.float32)
    check(lib().svo_triangulate(ptr(Pl, C.c_float), ptr(Pr, C.c_float), ptr(a, C.c_float), ptr(b, C.c_float), len(a),
                                ptr(X, C.c_float)))
    return X


def buildOpticalFlowPyramid(image, max_level: int = 3, win_size: int = 10):
    """Image levels of cv::buildOpticalFlowPyramid (vo.cpp:50-53, 200-201) as a list of arrays."""
    im = _img(image)
    h, w = im.shape
    out = np.zeros((w * h,), np.uint8)
    n = check(lib().svo_build_pyramid(ptr(im), w, h, im.strides[0], max_level, win_size, ptr(out), out.size))
    levels, o = [np.ascontiguousarray(im)], 0
    lw, lh = w, h
    for _ in range(1, n):
        lw, lh = (lw + 1) // 2, (lh + 1) // 2
        levels.append(out[o:o + lw * lh].reshape(lh, lw).copy())
        o += lw * lh
    return levels


def calcOpticalFlowPyrLK(prev_img, next_img, prev_pts, params: Params | None = None):
    """One LK leg as at vo.cpp:203-215.  Returns (next_pts [N,2], status [N] u8)."""
    a, b = _img(prev_img), _img(next_img)
    if a.shape != b.shape or a.strides[0] != b.strides[0]:
        a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    h, w = a.shape[:2]
    p = params or default_params(w, h, channels=3 if a.ndim == 3 else 1)
    pts = _pts(prev_pts)
    out = np.zeros_like(pts)
    st = np.zeros((len(pts),), np.uint8)
    check(lib().svo_lk_track(C.byref(p), ptr(a), ptr(b), a.strides[0], ptr(pts, C.c_float), len(pts),
                             ptr(out, C.c_float), ptr(st)))
    return out, st


# --------------------------------------------------------------------------- FeatureSet
class FeatureSet:
    """vo.h:132-189: parallel arrays points / ages / strengths."""

    def __init__(self):
        self.clear()

    def clear(self):
        self.points = np.zeros((0, 2), np.float32)
        self.ages = np.zeros((0,), np.int32)
        self.strengths = np.zeros((0,), np.int32)

    def size(self) -> int:
        return len(self.points)

    def appendFeaturesFromImage(self, image, fast_threshold: int):
        """feature_set.cpp:75-89."""
        pts, resp = featureDetectionFast(image, fast_threshold)
        self.points = np.concatenate([self.points, pts], 0)
        self.ages = np.concatenate([self.ages, np.zeros(len(pts), np.int32)])
        self.strengths = np.concatenate([self.strengths, resp.astype(np.int32)])
        self.filterByBucketLocation(image)

    def filterByBucketLocation(self, image):
        self.filterByBucketLocationInternal(image, BUCKETS_ALONG_HEIGHT, BUCKETS_ALONG_WIDTH, BUCKET_START_ROW,
                                            FEATURES_PER_BUCKET)

    def filterByBucketLocationInternal(self, image, buckets_along_height, buckets_along_width, bucket_start_row,
                                       features_per_bucket):
        """vo.h:167-174 / feature_set.cpp:95-147."""
        h, w = np.asarray(image).shape[:2]
        xy = np.ascontiguousarray(self.points, np.float32).copy()
        ages = np.ascontiguousarray(self.ages, np.int32).copy()
        strs = np.ascontiguousarray(self.strengths, np.int32).copy()
        n = check(lib().svo_bucket_filter(ptr(xy, C.c_float), ptr(ages, C.c_int), ptr(strs, C.c_int), len(ages), h, w,
                                          buckets_along_height, buckets_along_width, bucket_start_row,
                                          features_per_bucket, AGE_THRESHOLD, FAST_THRESHOLD))
        self.points, self.ages, self.strengths = xy[:n].copy(), ages[:n].copy(), strs[:n].copy()


def deleteFeaturesWithFailureStatus(fs: FeatureSet, isok):
    """vo.h:415-417 / vo.cpp:154-168."""
    k = np.asarray(isok).astype(bool)
    fs.points = deletePointsWithFailureStatus(fs.points, k)
    fs.ages = fs.ages[k].copy()
    fs.strengths = fs.strengths[k].copy()


# --------------------------------------------------------------------------- VisualOdometry
class VisualOdometry:
    """vo.h:231-380.  All cross-frame state (t0 pyramids, currentVOFeatures, rotation /
    translation guess, last_transform) lives on the GPU inside an svo_ctx."""

    def __init__(self, width: int | None = None, height: int | None = None, n_seq: int = 1, **param_overrides):
        self._ctx = None
        self._size = (width, height)
        self._n_seq = n_seq
        self._overrides = param_overrides
        self._proj = None
        self.leftCameraProjection_ = None
        self.rightCameraProjection_ = None
        self._last_images = None
        if width is not None and height is not None:
            self._create(width, height)

    def _create(self, width, height, channels=1):
        kw = dict(self._overrides)
        kw.setdefault("channels", channels)
        self.params = default_params(width, height, n_seq=self._n_seq, **kw)
        ctx = lib().svo_create(C.byref(self.params))
        if not ctx:
            raise SvoError("svo_create failed: " + lib().svo_last_error().decode())
        self._ctx = ctx
        self._size = (width, height)
        if self._proj is not None:
            self.initalize_projection_matricies(*self._proj)

    def close(self):
        if self._ctx:
            lib().svo_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def initalize_projection_matricies(self, leftCameraProjection, rightCameraProjection):
        """vo.h:307-309 / vo.cpp:8-26 (sic)."""
        Pl = np.ascontiguousarray(np.asarray(leftCameraProjection, np.float32).reshape(3, 4)).copy()
        Pr = np.ascontiguousarray(np.asarray(rightCameraProjection, np.float32).reshape(3, 4)).copy()
        self._proj = (Pl, Pr)
        self.leftCameraProjection_, self.rightCameraProjection_ = Pl, Pr
        if self._ctx:
            check(lib().svo_set_projection(self._ctx, ptr(Pl, C.c_float), ptr(Pr, C.c_float)))

    def stereo_callback(self, image_left, image_right):
        """vo.h:333-334 / vo.cpp:41-137.  Returns (success, 4x4 float64)."""
        L, R = _img(image_left), _img(image_right)
        if L.shape != R.shape:
            raise SvoError("left / right image sizes differ")
        if L.strides[0] != R.strides[0]:
            L, R = np.ascontiguousarray(L), np.ascontiguousarray(R)
        if self._ctx is None:
            self._create(L.shape[1], L.shape[0], 3 if L.ndim == 3 else 1)
        if (L.shape[1], L.shape[0]) != self._size or (3 if L.ndim == 3 else 1) != self.params.channels:
            raise SvoError("image size changed")
        T = np.zeros((4, 4), np.float64)
        ok = check(lib().svo_process_frame(self._ctx, ptr(L), ptr(R), L.strides[0], ptr(T, C.c_double)))
        self._last_images = (L, R)
        return bool(ok), T

    def submit(self, image_left, image_right):
        """Streaming form of stereo_callback (svo_submit_frame): enqueue one frame and return at once.  The
        arrays must stay alive (ideally pinned) until the matching collect()."""
        L, R = _img(image_left), _img(image_right)
        if L.shape != R.shape or L.strides[0] != R.strides[0]:
            raise SvoError("left / right image layouts differ")
        if self._ctx is None:
            self._create(L.shape[1], L.shape[0], 3 if L.ndim == 3 else 1)
        check(lib().svo_submit_frame(self._ctx, ptr(L), ptr(R), L.strides[0]))
        self._inflight = getattr(self, "_inflight", [])
        self._inflight.append((L, R))

    def submit_frames(self, pairs):
        """Look-ahead form of submit (svo_submit_frames): `pairs` = up to `lookahead` (left, right) frames that are
        already there; pyramids + FAST of all of them run in one launch set, results are collected one by one."""
        imgs = [(_img(L), _img(R)) for (L, R) in pairs]
        L0 = imgs[0][0]
        if any(a.shape != L0.shape or b.shape != L0.shape or a.strides[0] != L0.strides[0] or b.strides[0] != L0.strides[0]
               for a, b in imgs):
            raise SvoError("all frames of a look-ahead batch must share one layout")
        if self._ctx is None:
            self._create(L0.shape[1], L0.shape[0], 1)
        n = len(imgs)
        lp = (C.c_void_p * n)(*[a.ctypes.data for a, _ in imgs])
        rp = (C.c_void_p * n)(*[b.ctypes.data for _, b in imgs])
        check(lib().svo_submit_frames(self._ctx, n, lp, rp, L0.strides[0]))
        self._inflight = getattr(self, "_inflight", [])
        self._inflight += imgs

    def collect(self):
        """Result of the oldest submitted frame: (success, 4x4 float64), identical to stereo_callback's."""
        r = FrameResult()
        check(lib().svo_collect(self._ctx, C.byref(r)))
        if getattr(self, "_inflight", None):
            self._last_images = self._inflight.pop(0)
        return bool(r.success), np.array(r.transform, np.float64).reshape(4, 4)

    def stereo_callback_batch(self, left, right):
        """n_seq sequences per call: left/right are [n_seq, H, W] uint8.  Returns list[FrameResult]."""
        L, R = np.ascontiguousarray(left), np.ascontiguousarray(right)
        S, H, W = L.shape
        res = (FrameResult * S)()
        check(lib().svo_process_frame_batch(self._ctx, ptr(L), ptr(R), W, W * H, res))
        return list(res)

    def last_result(self, seq: int = 0) -> FrameResult:
        r = FrameResult()
        check(lib().svo_get_result(self._ctx, seq, C.byref(r)))
        return r

    def tap(self, which: int, seq: int = 0):
        """Parity taps (svo_get_tap)."""
        cap = self.params.buckets_along_height * self.params.buckets_along_width
        nf = {0: 2, 3: 2, 1: 8, 2: 11}[which]
        f = np.zeros((nf * cap,), np.float32)
        i = np.zeros((2 * cap,), np.int32)
        u = np.zeros((cap,), np.uint8)
        n = check(lib().svo_get_tap(self._ctx, seq, which, cap, ptr(f, C.c_float), ptr(i, C.c_int), ptr(u)))
        if which in (0, 3):
            return dict(points=f[:2 * n].reshape(n, 2).copy(), ages=i[:n].copy(), strengths=i[cap:cap + n].copy())
        g = lambda k: f[2 * cap * k:2 * cap * k + 2 * n].reshape(n, 2).copy()
        if which == 1:
            return dict(pL1=g(0), pR1=g(1), pR0=g(2), pC=g(3), status=u[:n].copy())
        r = self.last_result(seq)
        return dict(pL0=g(0), pL1=g(1), pR0=g(2), pR1=g(3), X3=f[8 * cap:8 * cap + 3 * n].reshape(n, 3).copy(),
                    inliers=i[:r.n_inliers].copy())

    def circularMatching(self, imgLeftT1, imgRightT1, pointsLeftT0, current_features: FeatureSet | None = None):
        """vo.h:374-379 / vo.cpp:169-240 against the images of the previous stereo_callback.
        Returns (pL0, pR0, pL1, pR1) compacted; `current_features` is compacted in place."""
        if self._ctx is None or self._last_images is None:
            raise SvoError("circularMatching needs a prior stereo_callback (t0 pyramids)")
        # the images of the cached t0 pyramids (vo.h:257-258) come from the context: after an empty frame they are
        # still the OLD ones (vo.cpp:179-181), exactly as in the reference
        L0 = np.empty_like(np.ascontiguousarray(self._last_images[0]))
        R0 = np.empty_like(L0)
        check(lib().svo_get_frame_images(self._ctx, 0, ptr(L0), ptr(R0), L0.strides[0]))
        return circularMatching(L0, R0, imgLeftT1, imgRightT1, pointsLeftT0, current_features, self.params)

    def matchingFeatures(self, imageLeftT0, imageRightT0, imageLeftT1, imageRightT1, currentVOFeatures: FeatureSet):
        """vo.h:354-362 / vo.cpp:315-366.  Returns (pL0, pR0, pL1, pR1); the feature set is updated in place."""
        imgs = [np.ascontiguousarray(_img(a)) for a in (imageLeftT0, imageRightT0, imageLeftT1, imageRightT1)]
        h, w = imgs[0].shape
        p = default_params(w, h)
        cap = p.buckets_along_height * p.buckets_along_width + currentVOFeatures.size()
        xy = np.zeros((cap, 2), np.float32)
        age = np.zeros((cap,), np.int32)
        st = np.zeros((cap,), np.int32)
        n0 = currentVOFeatures.size()
        xy[:n0], age[:n0], st[:n0] = currentVOFeatures.points, currentVOFeatures.ages, currentVOFeatures.strengths
        n = C.c_int(n0)
        out = [np.zeros((cap, 2), np.float32) for _ in range(4)]
        m = check(lib().svo_matching_features(C.byref(p), *[ptr(a) for a in imgs], w, ptr(xy, C.c_float),
                                              ptr(age, C.c_int), ptr(st, C.c_int), C.byref(n), cap,
                                              *[ptr(o, C.c_float) for o in out]))
        currentVOFeatures.points, currentVOFeatures.ages, currentVOFeatures.strengths = \
            xy[:m].copy(), age[:m].copy(), st[:m].copy()
        return tuple(o[:m].copy() for o in out)


def circularMatching(imgLeftT0, imgRightT0, imgLeftT1, imgRightT1, pointsLeftT0, current_features=None,
                     params: Params | None = None):
    """Stateless VisualOdometry::circularMatching (vo.cpp:169-240).  Returns (pL0, pR0, pL1, pR1)."""
    imgs = [np.ascontiguousarray(_img(a)) for a in (imgLeftT0, imgRightT0, imgLeftT1, imgRightT1)]
    h, w = imgs[0].shape
    p = params or default_params(w, h)
    p0 = _pts(pointsLeftT0).copy()
    n = len(p0)
    if n == 0:   # vo.cpp:179-181
        z = np.zeros((0, 2), np.float32)
        return z, z.copy(), z.copy(), z.copy()
    pR0, pL1, pR1 = (np.zeros((n, 2), np.float32) for _ in range(3))
    keep = np.zeros((n,), np.uint8)
    m = check(lib().svo_circular_match(C.byref(p), *[ptr(a) for a in imgs], w, ptr(p0, C.c_float), n,
                                       ptr(pR0, C.c_float), ptr(pL1, C.c_float), ptr(pR1, C.c_float), ptr(keep)))
    if current_features is not None and current_features.size() == n:
        deleteFeaturesWithFailureStatus(current_features, keep)
    return p0[:m].copy(), pR0[:m].copy(), pL1[:m].copy(), pR1[:m].copy()
