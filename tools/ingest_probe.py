"""Profiling helper: a few frames of a 64-sequence context (the batched ingest launches), for ncu."""
import ctypes as C
import sys

import numpy as np

sys.path.insert(0, '.')
import torch  # noqa: E402

import synthdata as synth  # noqa: E402
from stereo_visual_odometry_b200._lib import check, default_params, lib, ptr  # noqa: E402

W, H, S, P = 1241, 376, 64, 1280
seq = synth.make_sequence(width=W, height=H, calib="kitti00", n_frames=6, seed=0)
fr = [seq.render(i) for i in range(6)]
L_ = lib()
p = default_params(W, H, n_seq=S)
ctx = L_.svo_create(C.byref(p))
Pl, Pr = np.ascontiguousarray(seq.Pl), np.ascontiguousarray(seq.Pr)
check(L_.svo_set_projection(ctx, ptr(Pl, C.c_float), ptr(Pr, C.c_float)))
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    Lb = torch.zeros((S, H, P), dtype=torch.uint8, device="cuda")   # 128-byte aligned row pitch (TMA-capable planes)
    Rb = torch.zeros((S, H, P), dtype=torch.uint8, device="cuda")
    Lb[..., :W].copy_(torch.from_numpy(np.stack([fr[(i + s) % 6][0] for s in range(S)])))
    Rb[..., :W].copy_(torch.from_numpy(np.stack([fr[(i + s) % 6][1] for s in range(S)])))
    torch.cuda.synchronize()
    check(L_.svo_process_frame_batch_dev(ctx, Lb.data_ptr(), Rb.data_ptr(), P, P * H, None))
    check(L_.svo_synchronize(ctx))
L_.svo_destroy(ctx)
print("done")
