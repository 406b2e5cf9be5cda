"""A/B helper: CUDA-event time of the batched ingest stage (64 sequences per launch set) + per-frame stage times."""
import ctypes as C
import sys

import numpy as np

sys.path.insert(0, '.')
import torch  # noqa: E402

import synthdata as synth  # noqa: E402
from stereo_visual_odometry_b200._lib import check, default_params, lib, ptr  # noqa: E402

W, H, S, P, NF = 1241, 376, 64, 1280, 8
seq = synth.make_sequence(width=W, height=H, calib="kitti00", n_frames=NF, seed=0)
fr = [seq.render(i) for i in range(NF)]
L_ = lib()
p = default_params(W, H, n_seq=S)
ctx = L_.svo_create(C.byref(p))
Pl, Pr = np.ascontiguousarray(seq.Pl), np.ascontiguousarray(seq.Pr)
check(L_.svo_set_projection(ctx, ptr(Pl, C.c_float), ptr(Pr, C.c_float)))
Lb = torch.zeros((NF, S, H, P), dtype=torch.uint8, device="cuda")
Rb = torch.zeros((NF, S, H, P), dtype=torch.uint8, device="cuda")
for i in range(NF):
    Lb[i, :, :, :W].copy_(torch.from_numpy(np.stack([fr[(i + s) % NF][0] for s in range(S)])))
    Rb[i, :, :, :W].copy_(torch.from_numpy(np.stack([fr[(i + s) % NF][1] for s in range(S)])))
torch.cuda.synchronize()
for i in range(4):
    check(L_.svo_process_frame_batch_dev(ctx, Lb[i].data_ptr(), Rb[i].data_ptr(), P, P * H, None))
check(L_.svo_synchronize(ctx))
check(L_.svo_enable_timers(ctx, 1))
check(L_.svo_reset_timers(ctx))
n = 12
for k in range(n):
    i = (4 + k) % NF
    check(L_.svo_process_frame_batch_dev(ctx, Lb[i].data_ptr(), Rb[i].data_ptr(), P, P * H, None))
check(L_.svo_synchronize(ctx))
out = []
for st in range(6):
    ms, ln = C.c_double(0), C.c_longlong(0)
    L_.svo_get_stage_time(ctx, st, C.byref(ms), C.byref(ln))
    out.append(1e3 * ms.value / n)
print("batched stage us per launch set [ingest bucket klt select hyp refit]:", " ".join("%.1f" % v for v in out))
L_.svo_destroy(ctx)
