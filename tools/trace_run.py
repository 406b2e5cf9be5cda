"""Debug tool: per-kernel timeline of the streaming pipeline (the library reads SVO_TRACE=<csv>)."""
import csv
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, '.')
os.environ.setdefault("SVO_TRACE", "gpurun_out/trace.csv")
import torch  # noqa: E402

import synthdata as synth  # noqa: E402
from stereo_visual_odometry_b200._lib import FrameResult, check, default_params, lib, ptr  # noqa: E402

W, H = 1241, 376
N = 30
OFF = int(os.environ.get("TRACE_OFFSET", "0"))   # first frame of the 30-frame window (the bench's 1000-frame sequence gets richer in features)
seq = synth.make_sequence(width=W, height=H, calib="kitti00", n_frames=max(N + OFF, int(os.environ.get("TRACE_TOTAL", "0"))), seed=0)
Ls, Rs = zip(*[seq.render(OFF + i) for i in range(N)])
Lp = torch.from_numpy(np.stack(Ls)).pin_memory()
Rp = torch.from_numpy(np.stack(Rs)).pin_memory()
L_ = lib()
p = default_params(W, H)
ctx = L_.svo_create(C.byref(p))
Pl, Pr = np.ascontiguousarray(seq.Pl), np.ascontiguousarray(seq.Pr)
check(L_.svo_set_projection(ctx, ptr(Pl, C.c_float), ptr(Pr, C.c_float)))
fr = FrameResult()


def pp(i, n=N):
    q = i % (2 * n - 2)
    return q if q < n else 2 * n - 2 - q


for k in range(60):
    j = pp(k)
    check(L_.svo_submit_frame(ctx, Lp[j].data_ptr(), Rp[j].data_ptr(), W))
    if L_.svo_pending(ctx) > 3:
        check(L_.svo_collect(ctx, C.byref(fr)))
while L_.svo_pending(ctx) > 0:
    check(L_.svo_collect(ctx, C.byref(fr)))
L_.svo_destroy(ctx)
rows = list(csv.DictReader(open(os.environ["SVO_TRACE"])))
t0 = min(float(r["start_us"]) for r in rows if int(r["frame"]) == 36)
for r in rows:
    if 36 <= int(r["frame"]) <= 39:
        a, b = float(r["start_us"]) - t0, float(r["end_us"]) - t0
        print(f'{r["name"]:10s} f{r["frame"]} s{r["stream"]} {a:8.1f} -> {b:8.1f}  ({b - a:6.1f})')
