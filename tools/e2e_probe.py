"""Probe: end-to-end streaming rate vs frames in flight, and where the host thread spends its time."""
import ctypes as C
import sys
import time

import numpy as np

sys.path.insert(0, '.')
import torch  # noqa: E402

import synthdata as synth  # noqa: E402
from stereo_visual_odometry_b200._lib import FrameResult, check, default_params, lib, ptr  # noqa: E402

W, H, N = 1241, 376, 60
seq = synth.make_sequence(width=W, height=H, calib="kitti00", n_frames=N, seed=0)
Ls, Rs = zip(*[seq.render(i) for i in range(N)])
Lp = torch.from_numpy(np.stack(Ls)).pin_memory()
Rp = torch.from_numpy(np.stack(Rs)).pin_memory()
L_ = lib()


def pp(i, n=N):
    q = i % (2 * n - 2)
    return q if q < n else 2 * n - 2 - q


for depth in (int(a) for a in sys.argv[1:]):
    ctx = L_.svo_create(C.byref(default_params(W, H)))
    Pl, Pr = np.ascontiguousarray(seq.Pl), np.ascontiguousarray(seq.Pr)
    check(L_.svo_set_projection(ctx, ptr(Pl, C.c_float), ptr(Pr, C.c_float)))
    fr = FrameResult()
    k = 0
    for phase, n in (("warm", 100), ("timed", 1500)):
        t_sub = t_col = 0.0
        t0 = time.perf_counter()
        for _ in range(n):
            j = pp(k)
            a = time.perf_counter()
            check(L_.svo_submit_frame(ctx, Lp[j].data_ptr(), Rp[j].data_ptr(), W))
            b = time.perf_counter()
            t_sub += b - a
            k += 1
            if L_.svo_pending(ctx) > depth:
                check(L_.svo_collect(ctx, C.byref(fr)))
                t_col += time.perf_counter() - b
        while L_.svo_pending(ctx) > 0:
            check(L_.svo_collect(ctx, C.byref(fr)))
        dt = time.perf_counter() - t0
    print(f"depth {depth}: {n / dt:7.1f} frames/s  submit {1e6 * t_sub / n:6.1f} us/frame  collect-wait {1e6 * t_col / n:6.1f} us/frame",
          flush=True)
    L_.svo_destroy(ctx)
