"""Generates tests/golden/run1_first8.npz from the reference's bundled run1 sequence.

Run in the build container (needs /root/reference):  python tests/golden/make_golden.py
Contents: the PNG files of frames 0..7 (left/right, byte-for-byte, so cv2.imdecode gives the
very pixels the reference CLI reads at src/main.cpp:37-38) and the first 8 rows of the
reference's own committed output run1/result.csv (x,y,z) -- the golden trajectory.
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
