"""Debug tool: per-kernel timeline of the synchronous svo_process_frame path (SVO_TRACE=<csv>)."""
import csv
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, '.')
TR = os.environ.get("SVO_TRACE", "")
import torch  # noqa: E402

import synthdata as synth  # noqa: E402
from stereo_visual_odometry_b200._lib import check, default_params, lib, ptr  # noqa: E402

W, H = 1241, 376
N = 30
seq = synth.make_sequence(width=W, height=H, calib="kitti00", n_frames=N, seed=0)
Ls, Rs = zip(*[seq.render(i) for i in range(N)])
Lp = torch.from_numpy(np.stack(Ls)).pin_memory()
Rp = torch.from_numpy(np.stack(Rs)).pin_memory()
L_ = lib()
p = default_params(W, H)
ctx = L_.svo_create(C.byref(p))
Pl, Pr = np.ascontiguousarray(seq.Pl), np.ascontiguousarray(seq.Pr)
check(L_.svo_set_projection(ctx, ptr(Pl, C.c_float), ptr(Pr, C.c_float)))
T = np.zeros(16)


def pp(i, n=N):
    q = i % (2 * n - 2)
    return q if q < n else 2 * n - 2 - q


ts = []
for k in range(200):
    j = pp(k)
    t0 = time.perf_counter()
    check(L_.svo_process_frame(ctx, Lp[j].data_ptr(), Rp[j].data_ptr(), W, ptr(T, C.c_double)))
    ts.append(time.perf_counter() - t0)
print("sync us/frame median %.1f  mean %.1f" % (1e6 * np.median(ts[20:]), 1e6 * np.mean(ts[20:])))
L_.svo_destroy(ctx)
if TR:
    rows = list(csv.DictReader(open(TR)))
    t0 = min(float(r["start_us"]) for r in rows if int(r["frame"]) == 36)
    for r in rows:
        if 36 <= int(r["frame"]) <= 37:
            a, b = float(r["start_us"]) - t0, float(r["end_us"]) - t0
            print(f'{r["name"]:10s} f{r["frame"]} s{r["stream"]} {a:8.1f} -> {b:8.1f}  ({b - a:6.1f})')
