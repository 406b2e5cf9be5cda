"""Debug tool: where the synchronous svo_process_frame call spends its time on the host (submit vs collect)."""
import ctypes as C
import sys
import time

import numpy as np

sys.path.insert(0, '.')
import torch  # noqa: E402

import synthdata as synth  # noqa: E402
from stereo_visual_odometry_b200._lib import FrameResult, check, default_params, lib, ptr  # noqa: E402

W, H, N = 1241, 376, 30
seq = synth.make_sequence(width=W, height=H, calib="kitti00", n_frames=N, seed=0)
Ls, Rs = zip(*[seq.render(i) for i in range(N)])
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


ts, tc, tt = [], [], []
for k in range(300):
    j = pp(k)
    t0 = time.perf_counter()
    check(L_.svo_submit_frame(ctx, Lp[j].data_ptr(), Rp[j].data_ptr(), W))
    t1 = time.perf_counter()
    check(L_.svo_collect(ctx, C.byref(fr)))
    t2 = time.perf_counter()
    ts.append(t1 - t0)
    tc.append(t2 - t1)
    tt.append(t2 - t0)
us = lambda a: 1e6 * float(np.median(a[50:]))
print("submit %.1f us  collect %.1f us  total %.1f us per frame (medians, one frame in flight)" % (us(ts), us(tc), us(tt)))
# the same with the Python loop overhead measured alone
t0 = time.perf_counter()
for k in range(300):
    j = pp(k)
    _ = (Lp[j].data_ptr(), Rp[j].data_ptr())
print("python loop + data_ptr: %.1f us per frame" % (1e6 * (time.perf_counter() - t0) / 300))
L_.svo_destroy(ctx)
