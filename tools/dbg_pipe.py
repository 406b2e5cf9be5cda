"""Debug tool: counters of the streaming pipeline in steady state (cache misses, FAST candidates tracked / skipped)."""
import ctypes as C
import sys

import numpy as np

sys.path.insert(0, '.')
import torch  # noqa: E402

import synthdata as synth  # noqa: E402
from stereo_visual_odometry_b200._lib import FrameResult, check, default_params, lib, ptr  # noqa: E402

W, H, N = 1241, 376, 40
seq = synth.make_sequence(width=W, height=H, calib="kitti00", n_frames=N, seed=0)
Ls, Rs = zip(*[seq.render(i) for i in range(N)])
Lp = torch.from_numpy(np.stack(Ls)).pin_memory()
Rp = torch.from_numpy(np.stack(Rs)).pin_memory()
L_ = lib()
fr = FrameResult()
info = (C.c_int * 8)()
for stop in (25, 30, 35, 40):   # stream `stop` frames with 3 in flight, then look at the counters of the last one
    ctx = L_.svo_create(C.byref(default_params(W, H)))
    Pl, Pr = np.ascontiguousarray(seq.Pl), np.ascontiguousarray(seq.Pr)
    check(L_.svo_set_projection(ctx, ptr(Pl, C.c_float), ptr(Pr, C.c_float)))
    for k in range(stop):
        check(L_.svo_submit_frame(ctx, Lp[k].data_ptr(), Rp[k].data_ptr(), W))
        if L_.svo_pending(ctx) > 3:
            check(L_.svo_collect(ctx, C.byref(fr)))
    while L_.svo_pending(ctx) > 0:
        check(L_.svo_collect(ctx, C.byref(fr)))
    check(L_.svo_pipeline_info(ctx, info))
    print(f"frame {stop - 1}: bucketed {fr.n_bucketed} matched {fr.n_matched} inliers {fr.n_inliers} cache misses {info[4]} "
          f"FAST candidates tracked {info[7] & 0xFFFF} skipped {info[7] >> 16}")
    L_.svo_destroy(ctx)
