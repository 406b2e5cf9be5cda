"""CPU ORACLE (test infrastructure, NOT product code).

Restatement of the reference's per-frame stereo-VO front end
(``VisualOdometry::stereo_callback`` and everything below it) on top of
``cv2`` -- the same OpenCV C++ kernels the reference links (the reference
pins OpenCV only implicitly: soname ``libopencv_*.so.405`` => 4.5.x; this
image has opencv-python-headless 4.13).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product path never
does.

Pinned against (see tests/test_oracle_golden.py):
  * ``run1/result.csv`` (the reference's only valid end-to-end golden; mode
    ``cli_bgr`` + identity initial pose),
  * the known answers in the reference's own tests, src/main.cpp:50-264.

Reference citations are relative to /root/reference.
"""
from __future__ import annotations

import math
import time
from dataclasses import dataclass, field

import cv2
import numpy as np

# --- include/vo.h:53-127 -----------------------------------------------------
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
WIN_SIZE = (10, 10)          # include/vo.h:251
MAX_LEVEL = 3                # include/vo.h:252
TERMCRIT = (cv2.TERM_CRITERIA_COUNT + cv2.TERM_CRITERIA_EPS, 30, 0.0001)  # src/vo.cpp:183-184


def _trunc_div(a: int, b: int) -> int:
    """C integer division (truncation toward zero)."""
    q = abs(a) // abs(b)
    return q if (a >= 0) == (b > 0) else -q


def compute_score(age: int, strength: int) -> int:
    """Bucket::compute_score, src/feature_set.cpp:16-18."""
    return age + _trunc_div(strength - FAST_THRESHOLD, 20)


class Bucket:
    """src/feature_set.cpp:4-53."""

    def __init__(self, max_size: int):
        self.max_size = max_size
        self.points: list = []
        self.ages: list = []
        self.strengths: list = []

    def size(self) -> int:
        return len(self.points)

    def add_feature(self, point, age: int, strength: int) -> None:
        if not self.max_size:
            return
        if age < AGE_THRESHOLD:
            if self.size() < self.max_size:
                self.points.append(point)
                self.ages.append(age)
                self.strengths.append(strength)
            else:
                score = compute_score(age, strength)
                score_min = compute_score(self.ages[0], self.strengths[0])
                idx = 0
                for i in range(1, self.size()):
                    s = compute_score(self.ages[i], self.strengths[i])
                    if s < score_min:
                        score_min, idx = s, i
                if score > score_min:
                    self.points[idx] = point
                    self.ages[idx] = age
                    self.strengths[idx] = strength


def fast_detect(img_u8: np.ndarray, threshold: int):
    """featureDetectionFast, src/feature_set.cpp:55-68 -> cv::FAST(img, kps, thr, true).

    ``img_u8`` must be single-channel here (cv2's FastFeatureDetector would
    gray-convert a 3-channel image whereas the C++ free function scans raw
    bytes, SURVEY F4 -- the caller reinterprets the bytes itself).
    Returns (points float32 [K,2], responses float32 [K]) in row-major order.
    """
    det = cv2.FastFeatureDetector_create(threshold=int(threshold), nonmaxSuppression=True,
                                         type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    kps = det.detect(np.ascontiguousarray(img_u8), None)
    if not kps:
        return np.zeros((0, 2), np.float32), np.zeros((0,), np.float32)
    pts = np.array([k.pt for k in kps], np.float32)
    resp = np.array([k.response for k in kps], np.float32)
    return pts, resp


def bucket_filter(points, ages, strengths, rows, cols, n_bh, n_bw, start_row, per_bucket):
    """FeatureSet::filterByBucketLocationInternal, src/feature_set.cpp:95-147.

    Inputs are arrays in feature order; returns the new (points, ages,
    strengths) in bucket-row-major order.  The fast path for per_bucket==1 is
    semantically identical to the Bucket loop (earliest element with maximum
    score among age < 20); the general path runs the Bucket objects.
    """
    points = np.asarray(points, np.float32).reshape(-1, 2)
    ages = np.asarray(ages, np.int32).reshape(-1)
    strengths = np.asarray(strengths, np.int32).reshape(-1)
    bh = (rows + n_bh - 1) // n_bh
    bw = (cols + n_bw - 1) // n_bw
    n = len(ages)
    if n == 0:
        return points.copy(), ages.copy(), strengths.copy()
    # (int)(float / int): float32 division then truncation
    r = (points[:, 1] / np.float32(bh)).astype(np.int32)
    c = (points[:, 0] / np.float32(bw)).astype(np.int32)
    bidx = r * n_bw + c
    if per_bucket == 1:
        q = np.abs(strengths - FAST_THRESHOLD) // 20
        q = np.where(strengths - FAST_THRESHOLD >= 0, q, -q)
        score = ages + q
        ok = (r >= start_row) & (ages < AGE_THRESHOLD)
        idx = np.nonzero(ok)[0]
        # earliest element with max score per bucket: sort by (bucket, -score, index)
        order = np.lexsort((idx, -score[idx], bidx[idx]))
        sidx = idx[order]
        sb = bidx[sidx]
        first = np.ones(len(sidx), bool)
        first[1:] = sb[1:] != sb[:-1]
        win = sidx[first]           # already ascending in bucket index
        return points[win].copy(), ages[win].copy(), strengths[win].copy()
    buckets = {}
    for i in range(n):
        b = int(bidx[i])
        if b not in buckets:
            buckets[b] = Bucket(per_bucket if int(r[i]) >= start_row else 0)
        buckets[b].add_feature(points[i], int(ages[i]), int(strengths[i]))
    op, oa, os_ = [], [], []
    for b in sorted(buckets):
        op += buckets[b].points
        oa += buckets[b].ages
        os_ += buckets[b].strengths
    return (np.array(op, np.float32).reshape(-1, 2), np.array(oa, np.int32), np.array(os_, np.int32))


def find_close_points(p1, p2, threshold):
    """findClosePoints, src/vo.cpp:265-280 (float32 arithmetic; NaN => kept)."""
    p1 = np.asarray(p1, np.float32).reshape(-1, 2)
    p2 = np.asarray(p2, np.float32).reshape(-1, 2)
    off = np.maximum(np.abs(p1[:, 0] - p2[:, 0]), np.abs(p1[:, 1] - p2[:, 1]))
    return ~(off > np.float32(threshold))


def lk_leg(prev_img, next_img, pts):
    """One calcOpticalFlowPyrLK call as at src/vo.cpp:203-215."""
    p = np.ascontiguousarray(np.asarray(pts, np.float32).reshape(-1, 1, 2))
    nxt, st, _err = cv2.calcOpticalFlowPyrLK(
        prev_img, next_img, p, None, winSize=WIN_SIZE, maxLevel=MAX_LEVEL, criteria=TERMCRIT,
        flags=0, minEigThreshold=OPTICAL_FLOW_MIN_EIG_THRESHOLD)
    return nxt.reshape(-1, 2), st.reshape(-1).astype(np.uint8)


def triangulate(Pl, Pr, pL, pR):
    """src/vo.cpp:89-94: triangulatePoints + convertPointsFromHomogeneous (f32 out)."""
    X4 = cv2.triangulatePoints(np.asarray(Pl, np.float32), np.asarray(Pr, np.float32),
                               np.ascontiguousarray(np.asarray(pL, np.float32).reshape(-1, 2).T),
                               np.ascontiguousarray(np.asarray(pR, np.float32).reshape(-1, 2).T))
    X3 = cv2.convertPointsFromHomogeneous(np.ascontiguousarray(X4.T))
    return X3.reshape(-1, 3).astype(np.float32)


def camera_to_world(K, cam_pts, world_pts, rotation, translation):
    """cameraToWorld, src/vo.cpp:282-313.  Returns (inliers[int32], success, R, t)."""
    rvec, _ = cv2.Rodrigues(np.asarray(rotation, np.float64))
    tvec = np.asarray(translation, np.float64).reshape(3, 1).copy()
    old_t = tvec.copy()
    ok, rvec_o, tvec_o, inl = cv2.solvePnPRansac(
        np.ascontiguousarray(np.asarray(world_pts, np.float32).reshape(-1, 1, 3)),
        np.ascontiguousarray(np.asarray(cam_pts, np.float32).reshape(-1, 1, 2)),
        np.asarray(K, np.float32), np.zeros((4, 1), np.float32),
        rvec=rvec.copy(), tvec=tvec, useExtrinsicGuess=True,
        iterationsCount=RANSAC_ITERATIONS, reprojectionError=RANSAC_REPROJECTION_ERROR,
        confidence=0.98, flags=cv2.SOLVEPNP_ITERATIVE)
    inl = np.zeros((0,), np.int32) if inl is None else np.asarray(inl, np.int32).reshape(-1)
    if ok:
        R, _ = cv2.Rodrigues(rvec_o)
        return inl, True, R, np.asarray(tvec_o, np.float64).reshape(3, 1)
    return inl, False, np.asarray(rotation, np.float64).copy(), old_t


def get_inverse_transform(R, t):
    """getInverseTransform, src/vo.cpp:246-258."""
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = np.asarray(t).reshape(3)
    return np.linalg.inv(T)


@dataclass
class FrameTaps:
    """Per-stage outputs of one stereo_callback (parity probes)."""
    fast_pts: np.ndarray | None = None
    fast_resp: np.ndarray | None = None
    retry: bool = False
    bucketed: tuple | None = None          # (pts, ages, strengths) fed to circular matching
    lk: dict = field(default_factory=dict)  # pL1,pR1,pR0,pC,s0..s3
    keep: np.ndarray | None = None         # mask over the bucketed set after circular + bounds
    pL0: np.ndarray | None = None
    pL1: np.ndarray | None = None
    pR0: np.ndarray | None = None
    pR1: np.ndarray | None = None
    X3: np.ndarray | None = None
    inliers: np.ndarray | None = None
    pnp_ok: bool = False
    R: np.ndarray | None = None
    t: np.ndarray | None = None
    success: bool = False
    reason: str = ""
    times: dict = field(default_factory=dict)


class RefVO:
    """VisualOdometry, include/vo.h:231-380 + src/vo.cpp.

    mode 'gray': images are HxW uint8 (the 8UC1 contract).
    mode 'cli_bgr': images are HxWx3 uint8 as the CLI feeds them (F4): FAST
    scans the first W bytes of every 3W-byte row, LK runs on 3 channels.
    """

    def __init__(self, Pl, Pr, mode: str = "gray", params: dict | None = None):
        self.Pl = np.asarray(Pl, np.float32).reshape(3, 4).copy()   # vo.cpp:8-26
        self.Pr = np.asarray(Pr, np.float32).reshape(3, 4).copy()
        self.K = self.Pl[:, :3].copy()
        self.mode = mode
        self.frame_id = 0
        self.L0 = self.R0 = None
        self.pyr_valid = False          # lastLeft/RightPyramid follow L0/R0 unless circularMatching bailed early
        self.lastL = self.lastR = None  # images the cached pyramids were built from
        self.pts = np.zeros((0, 2), np.float32)
        self.ages = np.zeros((0,), np.int32)
        self.strengths = np.zeros((0,), np.int32)
        self.R = np.eye(3)
        self.t = np.zeros((3, 1))
        self.last_transform = np.eye(4)
        p = dict(n_bh=BUCKETS_ALONG_HEIGHT, n_bw=BUCKETS_ALONG_WIDTH, start_row=BUCKET_START_ROW,
                 per_bucket=FEATURES_PER_BUCKET, fast_threshold=FAST_THRESHOLD)
        if params:
            p.update(params)
        self.p = p
        self.taps: FrameTaps | None = None

    # -- FeatureSet::appendFeaturesFromImage, feature_set.cpp:75-89
    def _fast_view(self, img):
        if self.mode == "cli_bgr":
            h, w = img.shape[:2]
            return np.ascontiguousarray(img.reshape(h, 3 * w)[:, :w])
        return img

    def _append(self, img, thr, taps):
        t0 = time.perf_counter()
        pts, resp = fast_detect(self._fast_view(img), thr)
        t1 = time.perf_counter()
        if taps.fast_pts is None:
            taps.fast_pts, taps.fast_resp = pts, resp
        self.pts = np.concatenate([self.pts, pts], 0)
        self.ages = np.concatenate([self.ages, np.zeros(len(pts), np.int32)])
        self.strengths = np.concatenate([self.strengths, resp.astype(np.int32)])
        h, w = img.shape[:2]
        self.pts, self.ages, self.strengths = bucket_filter(
            self.pts, self.ages, self.strengths, h, w, self.p["n_bh"], self.p["n_bw"],
            self.p["start_row"], self.p["per_bucket"])
        t2 = time.perf_counter()
        taps.times["fast"] = taps.times.get("fast", 0.0) + (t1 - t0)
        taps.times["bucket"] = taps.times.get("bucket", 0.0) + (t2 - t1)

    def stereo_callback(self, L, R):
        """vo.cpp:41-137.  Returns (success, 4x4 float64)."""
        taps = FrameTaps()
        self.taps = taps
        fail = (False, self.last_transform.copy())
        if self.frame_id == 0:                      # vo.cpp:47-56
            self.L0, self.R0 = L, R
            self.lastL, self.lastR = L, R
            self.frame_id += 1
            taps.reason = "first"
            return fail
        self.frame_id += 1
        H, W = L.shape[:2]
        # ---- matchingFeatures, vo.cpp:315-366
        self._append(self.L0, self.p["fast_threshold"], taps)
        if len(self.ages) < PRE_MATCHING_FEATURE_THRESHOLD:
            taps.retry = True
            self._append(self.L0, self.p["fast_threshold"] // 4, taps)
        taps.bucketed = (self.pts.copy(), self.ages.copy(), self.strengths.copy())
        pL0 = self.pts.copy()
        n0 = len(pL0)
        if n0 == 0:                                 # vo.cpp:179-181 (cached pyramids NOT refreshed)
            pL1 = pR1 = pR0 = np.zeros((0, 2), np.float32)
            keep = np.zeros((0,), bool)
        else:
            t0 = time.perf_counter()
            pL1, s0 = lk_leg(self.lastL, L, pL0)            # vo.cpp:203-205
            pR1, s1 = lk_leg(L, R, pL1)                     # vo.cpp:206-208
            pR0, s2 = lk_leg(R, self.lastR, pR1)            # vo.cpp:209-211
            pC, s3 = lk_leg(self.lastR, self.lastL, pR0)    # vo.cpp:212-215
            t1 = time.perf_counter()
            taps.times["lk"] = t1 - t0
            taps.lk = dict(pL1=pL1, pR1=pR1, pR0=pR0, pC=pC, s0=s0, s1=s1, s2=s2, s3=s3)
            keep = find_close_points(pL0, pC, CIRCULAR_MATCHING_SUCCESS_THRESHOLD)
            keep &= (s0 != 0) & (s1 != 0) & (s2 != 0) & (s3 != 0)   # vo.cpp:227-230
            self.lastL, self.lastR = L, R                   # vo.cpp:231-232

            def inside(p):                                  # vo.cpp:341-359
                return (p[:, 0] >= 0) & (p[:, 1] >= 0) & (p[:, 1] < H) & (p[:, 0] < W)
            keep &= inside(pL0) & inside(pL1) & inside(pR0) & inside(pR1)
            taps.times["select"] = time.perf_counter() - t1
        taps.keep = keep
        pL0, pL1, pR0, pR1 = pL0[keep], pL1[keep], pR0[keep], pR1[keep]
        self.pts, self.ages, self.strengths = self.pts[keep], self.ages[keep], self.strengths[keep]
        taps.pL0, taps.pL1, taps.pR0, taps.pR1 = pL0, pL1, pR0, pR1
        self.ages = self.ages + 1                           # vo.cpp:70-72
        self.L0, self.R0 = L, R                             # vo.cpp:74-75
        if len(pL0) <= max(4, FEATURES_THRESHOLD):          # vo.cpp:82-84
            taps.reason = "few_features"
            return fail
        t0 = time.perf_counter()
        X3 = triangulate(self.Pl, self.Pr, pL0, pR0)        # vo.cpp:89-94
        t1 = time.perf_counter()
        inl, ok, Rn, tn = camera_to_world(self.K, pL1, X3, self.R, self.t)   # vo.cpp:101-102
        t2 = time.perf_counter()
        taps.times["triangulate"] = t1 - t0
        taps.times["pnp"] = t2 - t1
        self.R, self.t = Rn, tn
        taps.X3, taps.inliers, taps.pnp_ok, taps.R, taps.t = X3, inl, ok, Rn.copy(), tn.copy()
        if len(inl) < FEATURES_THRESHOLD or not ok:         # vo.cpp:106-113
            taps.reason = "pnp_failed"
            return fail
        self.pts = pL1[inl].copy()                          # vo.cpp:115-121
        self.ages, self.strengths = self.ages[inl], self.strengths[inl]
        tnorm = float(np.linalg.norm(self.t))
        rv, _ = cv2.Rodrigues(self.R)
        angle = float(np.linalg.norm(rv))
        if tnorm > MAX_TRANSLATION_NORM or angle > MAX_ROTATION_NORM:   # vo.cpp:123-132
            taps.reason = "too_big"
            return fail
        T = get_inverse_transform(self.R, self.t)           # vo.cpp:133-136
        self.last_transform = T
        taps.success = True
        return True, T.copy()


def initial_pose(kind: str = "pitch26") -> np.ndarray:
    """src/main.cpp:368-373 ('identity' is what run1/result.csv was produced with, F3)."""
    if kind == "identity":
        return np.eye(4)
    th = np.float32((26.0 / 360) * 2 * math.pi)   # `float theta`
    c, s = math.cos(float(th)), math.sin(float(th))
    return np.array([[1, 0, 0, 0], [0, c, s, 0], [0, -s, c, 0], [0, 0, 0, 1]], np.float64)


CLI_P_LEFT = np.array([[322.11376, 0.0, 327.47336, 0.0],
                       [0.0, 322.11376, 176.33722, 0.0],
                       [0.0, 0.0, 1.0, 0.0]], np.float32)       # src/main.cpp:357-362
CLI_P_RIGHT = np.array([[322.11376, 0.0, 327.47336, -22.5428],
                        [0.0, 322.11376, 176.33722, 0.0],
                        [0.0, 0.0, 1.0, 0.0]], np.float32)


def run_sequence(frames, Pl, Pr, mode="gray", pose0=None, params=None, collect=None):
    """The CLI loop, src/main.cpp:374-401.  ``frames`` yields (L, R).

    Returns (positions [n,3], success flags [n], transforms [n,4,4])."""
    vo = RefVO(Pl, Pr, mode=mode, params=params)
    pose = np.eye(4) if pose0 is None else np.array(pose0, np.float64)
    xyz, oks, Ts = [], [], []
    for L, R in frames:
        ok, T = vo.stereo_callback(L, R)
        pose = pose @ T                                  # main.cpp:396 (also on failure)
        xyz.append(pose[:3, 3].copy())
        oks.append(ok)
        Ts.append(T)
        if collect is not None:
            collect(vo)
    return np.array(xyz), np.array(oks, bool), np.array(Ts)
