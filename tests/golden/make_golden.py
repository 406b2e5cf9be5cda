"""Generates tests/golden/run1_first8.npz and run1_gray_8_47.npz from the reference's bundled run1 sequence.

Run in the build container (needs /root/reference):  python tests/golden/make_golden.py
Contents: the PNG files of frames 0..7 (left/right, byte-for-byte, so cv2.imdecode gives the
very pixels the reference CLI reads at src/main.cpp:37-38) and the first 8 rows of the
reference's own committed output run1/result.csv (x,y,z) -- the golden trajectory.
run1_gray_8_47.npz: frames 8..47 as single-channel PNGs (cv2.cvtColor(BGR2GRAY) of the decoded files, lossless),
so the 8UC1 pipeline can be run against the oracle on the GPU box (no /root/reference there) for 48 frames --
long enough to cross AGE_THRESHOLD = 20 (include/vo.h:84, src/feature_set.cpp:24) on real images.
"""
import os

import numpy as np

REF = "/root/reference/run1"
N = 8
out = {}
for i in range(N):
    for cam in ("left", "right"):
        with open(f"{REF}/{cam}/frame{i:06d}.png", "rb") as f:
            out[f"{cam}_{i}"] = np.frombuffer(f.read(), np.uint8)
out["result_xyz"] = np.loadtxt(f"{REF}/result.csv", delimiter=",", skiprows=1)[:N, :3]
np.savez(os.path.join(os.path.dirname(os.path.abspath(__file__)), "run1_first8.npz"), **out)
print("wrote run1_first8.npz")

import cv2

N2 = 48
out = {}
for i in range(N, N2):
    for cam in ("left", "right"):
        img = cv2.imread(f"{REF}/{cam}/frame{i:06d}.png", cv2.IMREAD_COLOR)
        ok, enc = cv2.imencode(".png", cv2.cvtColor(img, cv2.COLOR_BGR2GRAY), [cv2.IMWRITE_PNG_COMPRESSION, 9])
        assert ok
        out[f"{cam}_{i}"] = enc.ravel()
np.savez(os.path.join(os.path.dirname(os.path.abspath(__file__)), "run1_gray_8_47.npz"), **out)
print("wrote run1_gray_8_47.npz")
