"""Deterministic synthetic stereo sequences (TEST / BENCH INFRASTRUCTURE, not product code).

SURVEY.md 8(d): a static world made of a textured ground plane (1.65 below the camera) and a
fronto-parallel far wall, seen by a rectified stereo pair moving forward with small yaw / pitch
changes.  Per-frame translation stays <= 0.08 (< MAX_TRANSLATION_NORM = 0.1, include/vo.h of the
reference), otherwise every frame would be rejected.  Images are uint8 single channel
(the 8UC1 contract).  Ground-truth poses are returned so accuracy can be judged independently of
the oracle.  Imported by bench.py (input generation, outside every timed region), tools/, tests/ and smoke().
"""
from __future__ import annotations

import functools
from dataclasses import dataclass

import cv2
import numpy as np

ATLAS = 4096

CALIB = {   # calibration/*.yaml of the reference: (fx, fy, cx, cy, bf)
    "kitti00": (718.856, 718.856, 607.1928, 185.2157, -386.1448),
    "zed": (684.3679, 684.3679, 689.8894, 406.8742, -82.1242),
    "powerranger": (665.92872, 665.95872, 275.01847, 224.42522, -50.70819),
    "cli": (322.11376, 322.11376, 327.47336, 176.33722, -22.5428),   # src/main.cpp:357-362
}


def projection_matrices(calib: str):
    fx, fy, cx, cy, bf = CALIB[calib]
    Pl = np.array([[fx, 0, cx, 0], [0, fy, cy, 0], [0, 0, 1, 0]], np.float32)
    Pr = Pl.copy()
    Pr[0, 3] = bf   # src/stereo_vo.cpp:46-47
    return Pl, Pr


@functools.lru_cache(maxsize=4)
def texture_atlas(seed: int = 1234, density: float = 1.0) -> np.ndarray:
    """4096x4096 u8: base 96 + blurred noise, stamped triangles / rectangles with a 1-px hole
    (the reference's own synthetic fixture idea, src/main.cpp:89-100)."""
    rng = np.random.default_rng(seed)
    noise = rng.normal(0.0, 1.0, (ATLAS, ATLAS)).astype(np.float32)
    noise = cv2.GaussianBlur(noise, (0, 0), 2.0)
    noise *= 24.0 / max(1e-6, float(np.abs(noise).max()))
    img = np.clip(96.0 + noise, 0, 255).astype(np.uint8)
    n_shapes = int(density * ATLAS * ATLAS / (14 * 14))
    xs = rng.integers(12, ATLAS - 12, n_shapes)
    ys = rng.integers(12, ATLAS - 12, n_shapes)
    sides = rng.integers(6, 21, n_shapes)
    vals = rng.integers(0, 256, n_shapes)
    kinds = rng.integers(0, 2, n_shapes)
    for x, y, s, v, k in zip(xs.tolist(), ys.tolist(), sides.tolist(), vals.tolist(), kinds.tolist()):
        h = s // 2
        if k == 0:
            cv2.rectangle(img, (x - h, y - h), (x + h, y + h), int(v), -1)
        else:
            cv2.fillConvexPoly(img, np.array([[x, y - h], [x + h, y + h], [x - h, y + h]], np.int32), int(v))
        img[y, x] = 255 - v if abs(v - 128) > 40 else 0
    return img


def _rot_xyz(rx, ry, rz):
    Rx, _ = cv2.Rodrigues(np.array([rx, 0, 0], np.float64))
    Ry, _ = cv2.Rodrigues(np.array([0, ry, 0], np.float64))
    Rz, _ = cv2.Rodrigues(np.array([0, 0, rz], np.float64))
    return Rz @ Ry @ Rx


@dataclass
class Sequence:
    width: int
    height: int
    Pl: np.ndarray
    Pr: np.ndarray
    R_wc: np.ndarray        # [n,3,3] camera-to-world rotation of the left camera
    C: np.ndarray           # [n,3] left camera centre in the world
    baseline: float
    seed: int
    noise: float
    density: float
    wall_z: float
    ground_y: float = 1.65
    s_ground: float = 32.0  # atlas px per world unit
    s_wall: float = 10.0

    def __len__(self):
        return len(self.C)

    def _grid(self):
        if not hasattr(self, "_uv"):
            u, v = np.meshgrid(np.arange(self.width, dtype=np.float32), np.arange(self.height, dtype=np.float32))
            self._uv = (u, v)
        return self._uv

    def _render_cam(self, R_wc, C, frame, cam):
        """Both planes are projective maps of the atlas: two inverse-map warps + a half-plane mask."""
        atlas = texture_atlas(1234, self.density)
        K = self.Pl[:, :3].astype(np.float64)
        M = R_wc @ np.linalg.inv(K)                      # pixel -> world ray direction
        m0, m1, m2 = M[0], M[1], M[2]
        g, q = self.ground_y - C[1], self.wall_z - C[2]
        sg, sw = self.s_ground, self.s_wall
        ug0, vg0 = ATLAS / 2.0, 200.0                    # ground: atlas u = sg*X + ug0, v = sg*Z + vg0
        uw0, vw0 = ATLAS / 2.0 + 300.0, ATLAS / 2.0      # wall:   atlas u = sw*X + uw0, v = sw*Y + vw0
        Hg = np.stack([sg * (C[0] * m1 + g * m0) + ug0 * m1, sg * (C[2] * m1 + g * m2) + vg0 * m1, m1])
        Hw = np.stack([sw * (C[0] * m2 + q * m0) + uw0 * m2, sw * (C[1] * m2 + q * m1) + vw0 * m2, m2])
        fl = cv2.INTER_LINEAR | cv2.WARP_INVERSE_MAP
        size = (self.width, self.height)
        ig = cv2.warpPerspective(atlas, Hg, size, flags=fl, borderMode=cv2.BORDER_REFLECT_101)
        iw = cv2.warpPerspective(atlas, Hw, size, flags=fl, borderMode=cv2.BORDER_REFLECT_101)
        u, v = self._grid()
        dy = (m1[0] * u + m1[1] * v + m1[2]).astype(np.float32)
        dz = (m2[0] * u + m2[1] * v + m2[2]).astype(np.float32)
        ground = (dy > 0) & (np.float32(g) * dz < np.float32(q) * dy)   # ground hit nearer than the wall
        img = np.where(ground, ig, iw)
        if self.noise > 0:
            rng = np.random.default_rng(1000 * self.seed + 2 * frame + cam)
            img = np.clip(img.astype(np.float32) + rng.normal(0, self.noise, img.shape), 0, 255).astype(np.uint8)
        return np.ascontiguousarray(img)

    def render(self, i: int):
        """(left, right) uint8 [H,W] of frame i."""
        R_wc, C = self.R_wc[i], self.C[i]
        Cr = C + R_wc @ np.array([self.baseline, 0.0, 0.0])
        return self._render_cam(R_wc, C, i, 0), self._render_cam(R_wc, Cr, i, 1)

    def frames(self):
        for i in range(len(self)):
            yield self.render(i)

    def gt_transform(self, i: int) -> np.ndarray:
        """Ground-truth stereo_callback transform for frame i (pose of camera i in camera i-1)."""
        T0 = np.eye(4)
        T0[:3, :3], T0[:3, 3] = self.R_wc[i - 1], self.C[i - 1]
        T1 = np.eye(4)
        T1[:3, :3], T1[:3, 3] = self.R_wc[i], self.C[i]
        return np.linalg.inv(T0) @ T1


def make_sequence(width=1241, height=376, calib="kitti00", n_frames=1000, seed=0, noise=0.0, density=1.0,
                  step=(0.03, 0.08)) -> Sequence:
    Pl, Pr = projection_matrices(calib)
    if calib == "kitti00" and (width, height) != (1241, 376):   # keep the principal point inside smaller test images
        Pl = Pl.copy(); Pr = Pr.copy()
        Pl[0, 2] = Pr[0, 2] = width / 2.0 - 0.5
        Pl[1, 2] = Pr[1, 2] = height / 2.0 - 0.5
    baseline = float(-Pr[0, 3] / Pr[0, 0])
    rng = np.random.default_rng(42 + seed)
    steps = rng.uniform(step[0], step[1], n_frames)
    dyaw = rng.normal(0, 0.002, n_frames)
    dpitch = rng.normal(0, 0.002, n_frames)
    k = np.ones(5) / 5.0   # smooth
    dyaw = np.convolve(dyaw, k, mode="same")
    dpitch = np.convolve(dpitch, k, mode="same")
    yaw = np.cumsum(dyaw)
    pitch = np.clip(np.cumsum(dpitch), -0.05, 0.05)
    R_wc = np.stack([_rot_xyz(pitch[i], yaw[i], 0.0) for i in range(n_frames)])
    C = np.zeros((n_frames, 3))
    for i in range(1, n_frames):
        C[i] = C[i - 1] + R_wc[i - 1] @ np.array([0.0, 0.0, steps[i]])
    C[:, 1] = 0.0   # stay at constant height above the ground
    wall_z = float(C[:, 2].max() + 35.0)
    return Sequence(width, height, Pl, Pr, R_wc, C, baseline, seed, noise, density, wall_z)
